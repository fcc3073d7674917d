"""Times the fused matcher (top-2 stage only, CUDA events) over descriptor dimensions:
python tools/time_match_dims.py [n_images=16] [kpts=8192] [d ...]   (VC_MATCH_2CTA=1 selects the two-SM kernel;
VC_MATCH_NO_PRUNE=1 switches the pruning bound off; VC_BENCH_NONNEG=1 uses |unit rows|, cosine ~0.64 between unrelated rows)"""
import ctypes
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vit_colmap_b200 import _lib  # noqa: E402
from vit_colmap_b200.matching import Matcher  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 16
K = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
m = Matcher(0)
pairs = torch.tensor([(i, j) for i in range(n_img) for j in range(i + 1, n_img)], dtype=torch.int32, device="cuda")
counts = torch.full((n_img,), K, dtype=torch.int32, device="cuda")
opt = m._opts(0.8, 0.7, True, 1.0)
dims = [int(a) for a in sys.argv[3:]] or [64, 128, 256, 384, 768, 1024]
for d in dims:
    g = torch.Generator(device="cuda").manual_seed(d)
    x = torch.randn((n_img, K, d), device="cuda", generator=g)
    if os.environ.get("VC_BENCH_NONNEG"):
        x = x.abs()
    x = (x / x.norm(dim=2, keepdim=True)).to(torch.bfloat16)
    need = int(m.lib.vc_match_workspace_bytes(len(pairs), K))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")

    def run():
        _lib.check(m.lib.vc_match_pairs(m.handle, x.data_ptr(), counts.data_ptr(), n_img, K, d, pairs.data_ptr(),
                                        len(pairs), ctypes.byref(opt), None, None, ws.data_ptr(), need, m._stream()))
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    tf = 2.0 * K * K * d * len(pairs) / ms / 1e9
    print(f"d={d:5d}: {ms:8.3f} ms / {len(pairs)} pairs = {ms / len(pairs) * 1e3:7.1f} us/pair, {tf:7.1f} TFLOP/s, "
          f"{K * K * len(pairs) / ms / 1e9:6.2f} Tscores/s")
