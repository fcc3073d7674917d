"""Times the bandwidth-bound extraction stages (score map, selection, sample/pack) on a
batch of images with CUDA events on the launching stream and prints achieved GB/s (algorithmic
bytes / time) against the measured HBM peak.

    python tools/time_post.py [batch=256] [model_dim=768] [gh=37] [gw=37]
"""
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


ONCE = "--once" in sys.argv   # one launch per stage, for an ncu capture
if ONCE:
    sys.argv.remove("--once")


def timed(fn, iters=20, warm=3, flush=None):
    if ONCE:
        iters, warm = 1, 0
    for _ in range(warm):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    C = int(sys.argv[2]) if len(sys.argv) > 2 else 768
    gh = int(sys.argv[3]) if len(sys.argv) > 3 else 37
    gw = int(sys.argv[4]) if len(sys.argv) > 4 else 37
    from vit_colmap_b200.extraction import PostProcessor
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = None
    for k, v in peaks.items():
        if "hbm" in k.lower() and isinstance(v, (int, float)):
            hbm = float(v)
            break
    pp = PostProcessor(0)
    g = torch.Generator(device="cuda").manual_seed(0)
    tokens = torch.randn((B, gh * gw, C), device="cuda", generator=g)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    out = {"batch": B, "C": C, "grid": [gh, gw], "hbm_peak_GBps": hbm, "l2": "256 MB flush between iterations",
           "stages": {}}

    def report(name, ms, nbytes):
        gbps = nbytes / ms / 1e6
        out["stages"][name] = {"ms": round(ms, 4), "us_per_image": round(1e3 * ms / B, 3),
                               "algorithmic_MB": round(nbytes / 1e6, 2), "GBps": round(gbps, 1),
                               "frac_of_hbm_peak": round(gbps / hbm, 3) if hbm else None}

    for method in ("harris", "combined"):
        ms = timed(lambda: pp.score_map(tokens, (gh, gw), method), flush=flush)
        report(f"score_map[{method}] f32 tokens", ms, B * gh * gw * (C * 4 + 4))
    tok16 = tokens.to(torch.bfloat16)
    ms = timed(lambda: pp.score_map(tok16, (gh, gw), "harris"), flush=flush)
    report("score_map[harris] bf16 tokens", ms, B * gh * gw * (C * 2 + 4))
    score = pp.score_map(tokens, (gh, gw), "harris")
    for K in (2048, 8192):
        ms = timed(lambda: pp.select_keypoints(score, K), flush=flush)
        report(f"select_keypoints[K={K}]", ms, B * gh * gw * 4)
    # descriptor sampling at K keypoints per image (random cells: the matcher-sized case)
    for K, d in ((8192, 768), (8192, 128), (2048, 128)):
        if d > C:
            continue
        cap = K  # random cells (repeats allowed): the gather is what is timed
        coords = torch.stack([torch.randint(0, gh, (B, cap), device="cuda", generator=g),
                              torch.randint(0, gw, (B, cap), device="cuda", generator=g)], dim=2).int().contiguous()
        counts = torch.full((B,), cap, dtype=torch.int32, device="cuda")
        proj = None if d == C else torch.linalg.qr(torch.randn(C, d, device="cuda", generator=g))[0].contiguous()
        ms = timed(lambda: pp.sample_pack(tokens, (gh, gw), coords, counts, (518, 518), (518, 518), projection=proj),
                   iters=10, flush=flush)
        nbytes = B * cap * (C * 4 + 3 * d + 8)   # one token row in, bf16 + u8 rows + keypoint out
        report(f"sample_pack[K={cap}, d={d}] f32 tokens", ms, nbytes)
        nbytes = B * cap * (4 * C * 4 + 3 * d + 8)  # bilinear: four token rows per keypoint
        ms = timed(lambda: pp.sample_pack(tokens, (gh, gw), coords.float(), counts, (518, 518), (518, 518),
                                          projection=proj, rootsift=True), iters=10, flush=flush)
        report(f"sample_pack[K={cap}, d={d}] rootsift float coords", ms, nbytes)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
