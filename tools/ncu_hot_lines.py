"""Aggregate the per-instruction stall samples of an ncu report by CUDA source line.

    python tools/ncu_hot_lines.py gpurun_out/prof.ncu-rep vit_colmap_b200/build/match_fused2.o match_fused2_kernel [top] [mangled]

(`mangled`: substring of the kernel's section name in the cubin when it differs from the demangled name in the
report, e.g. "vit_gemm2_kernel<(int)0>" and "vit_gemm2_kernelILi0".)

ncu's `--page source --csv` lists SASS instructions with their sample counts but no source line; the
line table comes from `nvdisasm -g` on the cubin of the object file built with -lineinfo.  The n-th
instruction of the kernel in both listings is the same instruction."""
import csv
import io
import re
import subprocess
import sys
import tempfile
from collections import defaultdict
from pathlib import Path


def sass_lines(obj: Path, kernel: str):
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", str(obj.resolve())], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
        cubin = next(Path(tmp).glob("*.cubin"))
        text = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], check=True, capture_output=True, text=True).stdout
    out, line, inside, inline_depth = [], None, False, 0
    for ln in text.splitlines():
        if ln.startswith("\t.section") or ln.startswith(".section"):
            inside = kernel in ln and ".text." in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            # keep the outermost (non-inlined) location when nvdisasm prints an inline chain
            if "inlined at" in m.group(3):
                m2 = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
                line = (Path(m2[-1][0]).name, int(m2[-1][1]), f"{Path(m.group(1)).name}:{m.group(2)}")
            else:
                line = (Path(m.group(1)).name, int(m.group(2)), "")
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
            out.append((line, ln.split("*/", 1)[1].strip().rstrip(";")))
    return out


def main():
    rep, obj, kernel = sys.argv[1], Path(sys.argv[2]), sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
    mangled = sys.argv[5] if len(sys.argv) > 5 else kernel
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    start = next(i for i, r in enumerate(rows) if r and r[0] == "Kernel Name" and kernel in r[1])
    hdr = rows[start + 1]
    body = []
    for r in rows[start + 2:]:
        if not r or r[0] == "Kernel Name":
            break
        body.append(r)
    table = sass_lines(obj, mangled)
    if len(table) != len(body):
        print(f"warning: {len(body)} instructions in the report, {len(table)} in the cubin", file=sys.stderr)
    i_s, i_ns, i_ex = hdr.index("# Samples"), hdr.index("Warp Stall Sampling (Not-issued Samples)"), hdr.index("Instructions Executed")
    stall_cols = [(k, h) for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    by_line = defaultdict(lambda: [0, 0, 0, defaultdict(int)])
    total = 0
    for k, r in enumerate(body):
        loc = table[k][0] if k < len(table) else None
        key = (loc[0], loc[1]) if loc else ("?", 0)
        n = int(r[i_s] or 0)
        total += n
        e = by_line[key]
        e[0] += n
        e[1] += int(r[i_ns] or 0)
        e[2] += int(r[i_ex] or 0)
        for c, h in stall_cols:
            v = int(r[c] or 0)
            if v:
                e[3][h] += v
    print(f"{total} samples; top {top} source lines")
    src_cache = {}
    for (f, ln), (n, ns, ex, st) in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:top]:
        path = next(iter(Path("vit_colmap_b200/csrc").glob(f)), None) if f != "?" else None
        if path and path not in src_cache:
            src_cache[path] = path.read_text().splitlines()
        text = src_cache[path][ln - 1].strip()[:90] if path and 0 < ln <= len(src_cache[path]) else ""
        stalls = ", ".join(f"{h[6:]} {v}" for h, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{100.0 * n / max(total, 1):5.1f}%  {f}:{ln:<4d} exec {ex:>11d}  [{stalls}]  {text}")


if __name__ == "__main__":
    main()
