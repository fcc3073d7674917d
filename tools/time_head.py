"""Times the ViTFeatureModel conv head layer by layer (CUDA events) on synthetic activations:
    python tools/time_head.py [C] [batch] [grid] [d]"""
import ctypes
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from vit_colmap_b200 import _lib  # noqa: E402
from vit_colmap_b200.head import ConvHead, synthetic_head_state_dict  # noqa: E402

C = int(sys.argv[1]) if len(sys.argv) > 1 else 768
B = int(sys.argv[2]) if len(sys.argv) > 2 else 16
G = int(sys.argv[3]) if len(sys.argv) > 3 else 37
d = int(sys.argv[4]) if len(sys.argv) > 4 else 128
head = ConvHead(synthetic_head_state_dict(C, d, 0), C, d, 0)
lib, h = head.lib, head.handle
st = int(torch.cuda.current_stream().cuda_stream)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def conv(name, H, W, cin, cout, kind, epi, wkey, skey, tkey):
    x = torch.randn((B, H, W, cin), device="cuda").to(torch.bfloat16)
    s = 2 if kind == 1 else 1
    out = torch.empty((B, H * s, W * s, cout), dtype=torch.float32 if epi == 2 else torch.bfloat16, device="cuda")
    k = head._k
    fn = lambda: _lib.check(lib.vc_conv2d(h, x.data_ptr(), B, H, W, cin, k[wkey].data_ptr(), cout, kind, epi,  # noqa: E731
                                          k[skey].data_ptr() if skey else None, k[tkey].data_ptr(), out.data_ptr(), st))
    ms = timed(fn)
    taps = {0: 9, 1: 4, 2: 1}[kind]
    fl = 2.0 * B * H * W * s * s * taps * cin * cout
    print(f"{name:28s} {H:4d}x{W:<4d} {cin:4d}->{cout:<4d} {ms:8.3f} ms  {fl / ms / 1e9:8.1f} TFLOP/s")
    return ms


oh, ow = head.target_size(G, G)
total = 0.0
total += conv("upsampler.0.deconv", G, G, C, 512, 1, 0, "deconv_w0", None, "deconv_b0")
total += conv("upsampler.0.conv+bn+gelu", 2 * G, 2 * G, 512, 512, 0, 1, "conv_w0", "conv_s0", "conv_t0")
total += conv("upsampler.1.deconv", 2 * G, 2 * G, 512, 512, 1, 0, "deconv_w1", None, "deconv_b1")
total += conv("upsampler.1.conv+bn+gelu", 4 * G, 4 * G, 512, 512, 0, 1, "conv_w1", "conv_s1", "conv_t1")
total += conv("trunk conv+bn+gelu", oh, ow, 512, 256, 0, 1, "trunk_w", "trunk_s", "trunk_t")
total += conv("heads conv+bn+gelu (64+128)", oh, ow, 256, 192, 0, 1, "heads_w", "heads_s", "heads_t")
total += conv("descriptor 1x1 (dense)", oh, ow, 128, d, 2, 2, "desc_w", None, "desc_b")
print(f"sum of conv layers: {total:.3f} ms")

tok = torch.randn((B, G * G, C), device="cuda")
for dense in (True, False):
    ms = timed(lambda: head.forward(tok, (G, G), dense_descriptors=dense))
    fl = head.flops_per_image(G, G) * B
    print(f"head.forward dense={dense}: {ms:.3f} ms/batch, {B / ms * 1e3:.1f} img/s, {fl / ms / 1e9:.1f} TFLOP/s")
