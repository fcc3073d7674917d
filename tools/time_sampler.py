"""Times the training-time sampler kernels with CUDA events on the launching stream,
L2 flushed between iterations, and the CPU oracle (the reference's own torch calls, all host cores) on a
bounded sample beside them.

    python tools/time_sampler.py [batch=16] [C=768] [gh=85] [gw=114] [K=512]
"""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def timed(fn, flush, iters=20, warm=3):
    for _ in range(warm):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev]))


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    C = int(sys.argv[2]) if len(sys.argv) > 2 else 768
    gh = int(sys.argv[3]) if len(sys.argv) > 3 else 85
    gw = int(sys.argv[4]) if len(sys.argv) > 4 else 114
    K = int(sys.argv[5]) if len(sys.argv) > 5 else 512
    from oracle import sampler_oracle as so   # CPU baseline leg + input recipe only
    from vit_colmap_b200.dataloader._device import SamplerDevice, TokenMap
    from vit_colmap_b200.dataloader.homography_utils import inverse_homography
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = next((float(v) for k, v in peaks.items() if "hbm" in k.lower() and isinstance(v, (int, float))), None)
    dev = SamplerDevice(0)
    HW = gh * gw
    g = torch.Generator(device="cuda").manual_seed(0)
    tm1 = TokenMap(torch.randn((B, HW, C), device="cuda", generator=g), (gh, gw))
    tm2 = TokenMap(torch.randn((B, HW, C), device="cuda", generator=g), (gh, gw))
    _, _, H, image_size = so.synthetic_pair(0, B, 4, gh, gw)
    h_inv = inverse_homography(H).cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {"batch": B, "C": C, "grid": [gh, gw], "K": K, "hbm_peak_GBps": hbm, "l2": "256 MB flush between iterations",
           "stages": {}}

    ms = timed(lambda: dev.warp_similarity(tm1, tm2, h_inv, 14, image_size), flush)
    nbytes = 2 * B * HW * C * 4 + B * HW * 4
    out["stages"]["warp_similarity"] = {"ms": round(ms, 4), "algorithmic_GB": round(nbytes / 1e9, 4),
                                        "GBps": round(nbytes / ms / 1e6, 1),
                                        "frac_of_hbm_peak": round(nbytes / ms / 1e6 / hbm, 3) if hbm else None}
    sim = dev.warp_similarity(tm1, tm2, h_inv, 14, image_size).reshape(B, HW)
    ms = timed(lambda: dev.topk_rows(sim, min(K, HW), 0.0), flush)
    out["stages"]["topk_rows"] = {"ms": round(ms, 4), "us_per_row": round(1e3 * ms / B, 2), "n": HW, "k": min(K, HW)}
    order, _, _ = dev.topk_rows(sim, min(K, HW))
    xy = torch.stack([(order % gw).float(), (order // gw).float()], dim=-1)
    ms = timed(lambda: dev.sample_bilinear(tm2, xy), flush)
    nb = B * xy.shape[1] * C * 4 * 2
    out["stages"]["sample_bilinear"] = {"ms": round(ms, 4), "GBps_rows_in_plus_out": round(nb / ms / 1e6, 1)}
    pos = dev.sample_bilinear(tm2, xy)
    ms = timed(lambda: dev.hard_negatives(tm2, pos, xy, 5, 16.0), flush)
    fl = 2.0 * B * xy.shape[1] * HW * C
    out["stages"]["hard_negatives"] = {"ms": round(ms, 4), "GFLOP": round(fl / 1e9, 2),
                                       "TFLOPs_fp32": round(fl / ms / 1e9, 2),
                                       "scores_per_s": round(B * xy.shape[1] * HW / ms * 1e3, 0)}

    # CPU oracle beside it (bounded: 2 images)
    nb_cpu = min(B, 2)
    f1 = tm1.tokens[:nb_cpu].cpu().reshape(nb_cpu, gh, gw, C).permute(0, 3, 1, 2).contiguous()
    f2 = tm2.tokens[:nb_cpu].cpu().reshape(nb_cpu, gh, gw, C).permute(0, 3, 1, 2).contiguous()
    t0 = time.perf_counter()
    so.similarity_map(f1, f2, H[:nb_cpu], 14, image_size)
    t_sim = (time.perf_counter() - t0) / nb_cpu
    kk = min(64, xy.shape[1])
    t0 = time.perf_counter()
    so.hard_negative_indices(f2, pos[:nb_cpu, :kk].cpu(), xy[:nb_cpu, :kk].cpu(), 5, 16.0)
    t_hn = (time.perf_counter() - t0) / nb_cpu * (xy.shape[1] / kk)
    out["cpu_oracle"] = {"threads": torch.get_num_threads(), "similarity_ms_per_pair": round(1e3 * t_sim, 2),
                         "hard_negatives_ms_per_image": round(1e3 * t_hn, 2),
                         "sample": f"{nb_cpu} images; hard negatives on {kk} of {xy.shape[1]} positives, scaled"}
    out["gpu_ms_per_pair"] = {"similarity": round(out["stages"]["warp_similarity"]["ms"] / B, 4),
                              "hard_negatives": round(out["stages"]["hard_negatives"]["ms"] / B, 4)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
