"""In-tree build of the CUDA extension (C-ABI shared library) for sm_100a.

    python -m vit_colmap_b200.build [--force]

Produces vit_colmap_b200/_vc_b200.so with nvcc (cross-compiles without a GPU).  The .so is
git-ignored but travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OUT = PKG / "_vc_b200.so"
SOURCES = ["match_kernel.cu", "match_fused.cu", "match_fused2.cu", "vc_api.cu", "score_kernels.cu", "select_kernels.cu",
           "sample_kernels.cu", "vc_api_extract.cu", "vc_api_noalloc.cu", "vc_api_sampler.cu", "vit_gemm.cu", "vit_gemm2.cu", "vit_attention.cu",
           "vit_pointwise.cu", "vc_api_vit.cu", "conv_gemm.cu", "conv_gemm2.cu", "head_pointwise.cu", "vc_api_head.cu"]
EXTRA_SOURCES_GLOB = "*_kernels.cu"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found")


def sources() -> list[Path]:
    srcs = [CSRC / s for s in SOURCES]
    srcs += sorted(p for p in CSRC.glob(EXTRA_SOURCES_GLOB) if p not in srcs)
    return srcs


def needs_build() -> bool:
    if not OUT.exists():
        return True
    t = OUT.stat().st_mtime
    deps = list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "vc_b200.h"]
    return any(p.stat().st_mtime > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return OUT
    objdir = PKG / "build"
    objdir.mkdir(exist_ok=True)
    nvcc = _nvcc()
    headers = list(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "vc_b200.h"]
    newest_header = max(h.stat().st_mtime for h in headers)

    def compile_one(src: Path) -> str:
        obj = objdir / (src.stem + ".o")
        if not force and obj.exists() and obj.stat().st_mtime > max(src.stat().st_mtime, newest_header):
            return str(obj)  # object is newer than its source and every header
        cmd = [nvcc, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src.name}")
        (objdir / (src.stem + ".ptxas.log")).write_text(res.stdout + res.stderr)
        return str(obj)

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        objs = list(pool.map(compile_one, sources()))
    cmd = [nvcc, "-shared", "-o", str(OUT), *objs, "-gencode", "arch=compute_100a,code=sm_100a",
           "-cudart", "static"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
