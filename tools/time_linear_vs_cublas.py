"""Times vc_vit_linear (bias epilogue, bf16 out) against torch.matmul (cuBLAS, no epilogue) on the ViT-B GEMM
shapes, CUDA events, L2 flushed between iterations.  cuBLAS is the yardstick only; it is not on the product path.

    python tools/time_linear_vs_cublas.py"""
import ctypes
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vit_colmap_b200 import _lib  # noqa: E402

lib = _lib.load()
h = ctypes.c_void_p()
_lib.check(lib.vc_create(0, ctypes.byref(h)))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3


M = 43840
for name, n, k, epi in (("qkv", 2304, 768, 0), ("proj", 768, 768, 2), ("fc1", 3072, 768, 1), ("fc2", 768, 3072, 2)):
    a = torch.randn(M, k, device="cuda").to(torch.bfloat16)
    w = (torch.randn(n, k, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn(n, device="cuda")
    gamma = torch.randn(n, device="cuda")
    out = torch.zeros(M, n, dtype=torch.float32 if epi == 2 else torch.bfloat16, device="cuda")
    s = int(torch.cuda.current_stream().cuda_stream)
    ours = timed(lambda: lib.vc_vit_linear(h, a.data_ptr(), w.data_ptr(), bias.data_ptr(), gamma.data_ptr(), M, n, k, epi,
                                           out.data_ptr(), s))
    wt = w.t()
    ref = timed(lambda: torch.matmul(a, wt))
    fl = 2.0 * M * n * k
    print(f"{name}: M={M} N={n} K={k}: ours (epilogue {epi}) {ours:7.1f} us = {fl / ours / 1e6:7.1f} TFLOP/s | "
          f"cuBLAS matmul (no epilogue) {ref:7.1f} us = {fl / ref / 1e6:7.1f} TFLOP/s")
