"""Times vc_vit_attention alone (CUDA events): python tools/time_attn.py [B] [T] [heads]"""
import ctypes
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vit_colmap_b200 import _lib  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1370
heads = int(sys.argv[3]) if len(sys.argv) > 3 else 12
lib = _lib.load()
h = ctypes.c_void_p()
_lib.check(lib.vc_create(0, ctypes.byref(h)))
C = heads * 64
qkv = (torch.randn((B, T, 3 * C), device="cuda") * 1.5).to(torch.bfloat16)
out = torch.zeros((B, T, C), dtype=torch.bfloat16, device="cuda")
st = int(torch.cuda.current_stream().cuda_stream)
run = lambda: _lib.check(lib.vc_vit_attention(h, qkv.data_ptr(), B, T, heads, out.data_ptr(), st))  # noqa: E731
for _ in range(5):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
n = 20
e0.record()
for _ in range(n):
    run()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / n * 1e3
fl = 4.0 * B * heads * T * T * 64
print(f"attention B={B} T={T} heads={heads}: {us:.1f} us, {fl / us / 1e6:.1f} TFLOP/s, "
      f"{B * heads * T * T / us / 1e3:.1f} Gexp/s")
