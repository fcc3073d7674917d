"""Times guided matching (vc_match_pairs_guided: geometric gate in the GEMM epilogue) next to the
unguided matcher on the same descriptors:  python tools/time_guided.py [n_images=8] [kpts=8192] [d=128] [signed]
Rates are per pair (both directions, cross check, emit) with CUDA events on the launching stream."""
import ctypes
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vit_colmap_b200 import _lib  # noqa: E402
from vit_colmap_b200.matching import Matcher, PairGeometry  # noqa: E402

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 8
K = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
d = int(sys.argv[3]) if len(sys.argv) > 3 else 128
signed = len(sys.argv) > 4 and sys.argv[4] == "signed"   # default: non-negative rows (uint8-descriptor-like, cosine ~0.64)
m = Matcher(0)
# scene: image i > 0 shows image 0's points moved by a homography (H test) or along their epipolar
# lines (F test), in permuted row order, descriptors perturbed; pairs (0, i) carry that model
pairs_l = [(0, j) for j in range(1, n_img)]
pairs = torch.tensor(pairs_l, dtype=torch.int32, device="cuda")
counts = torch.full((n_img,), K, dtype=torch.int32, device="cuda")
g = torch.Generator(device="cuda").manual_seed(1)
H = np.array([[1.01, 0.02, 5.0], [-0.01, 0.99, -3.0], [1e-5, 2e-5, 1.0]])
F = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
Ht = torch.tensor(H, dtype=torch.float64, device="cuda")


def scene(kind):
    fold = (lambda t: t) if signed else torch.abs
    base = fold(torch.randn((K, d), device="cuda", generator=g))
    kp0 = torch.rand((K, 2), device="cuda", generator=g, dtype=torch.float64) * torch.tensor([1600.0, 1200.0], device="cuda")
    xs, ks = [base], [kp0]
    for _ in range(1, n_img):
        perm = torch.randperm(K, device="cuda", generator=g)
        xs.append(fold(base + 0.15 * torch.randn((K, d), device="cuda", generator=g))[perm])
        if kind == "H":
            q = torch.cat([kp0, torch.ones((K, 1), device="cuda", dtype=torch.float64)], 1) @ Ht.T
            moved = q[:, :2] / q[:, 2:]
        else:
            moved = kp0.clone()
            moved[:, 0] += (torch.rand((K,), device="cuda", generator=g, dtype=torch.float64) - 0.5) * 100
        ks.append((moved + 0.5 * torch.randn((K, 2), device="cuda", generator=g, dtype=torch.float64))[perm])
    x = torch.stack(xs)
    return (x / x.norm(dim=2, keepdim=True)).to(torch.bfloat16).contiguous(), torch.stack(ks).float().contiguous()


x, kp = scene("H")
need = int(m.lib.vc_match_workspace_bytes(len(pairs_l), K))
ws = torch.empty(need, dtype=torch.uint8, device="cuda")
matches = torch.empty((len(pairs_l), K, 2), dtype=torch.int32, device="cuda")
n_matches = torch.empty((len(pairs_l),), dtype=torch.int32, device="cuda")
opt = m._opts(0.8, 0.7, True, 1.0)


def timed(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(3):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 3


def plain():
    _lib.check(m.lib.vc_match_pairs(m.handle, x.data_ptr(), counts.data_ptr(), n_img, K, d, pairs.data_ptr(), len(pairs_l),
                                    ctypes.byref(opt), matches.data_ptr(), n_matches.data_ptr(), ws.data_ptr(), need,
                                    m._stream()))


ms = timed(plain)
print(f"unguided (fused single-pass kernel): {ms / len(pairs_l) * 1e3:8.1f} us/pair")
for name, kind, geo in (("H (SiftGPU rule)", "H", PairGeometry(4, None, H, 4.0)), ("F (SiftGPU rule)", "F", PairGeometry(3, F, None, 4.0))):
    x, kp = scene(kind)
    arr = (_lib.PairGeometry * len(pairs_l))(*[geo.to_c(0) for _ in pairs_l])
    geo_dev = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).cuda()

    def guided():
        _lib.check(m.lib.vc_match_pairs_guided(m.handle, x.data_ptr(), kp.data_ptr(), counts.data_ptr(), n_img, K, d,
                                               pairs.data_ptr(), geo_dev.data_ptr(), len(pairs_l), ctypes.byref(opt),
                                               matches.data_ptr(), n_matches.data_ptr(), ws.data_ptr(), need, m._stream()))
    ms = timed(guided)
    print(f"guided {name}: {ms / len(pairs_l) * 1e3:8.1f} us/pair, {2 * K * K * len(pairs_l) / ms / 1e9:6.2f} T gated candidates/s, "
          f"mean matches {n_matches.float().mean().item():.1f}")
