"""Times the ViT forward (CUDA events) on synthetic images: python tools/time_vit.py [name] [batch] [size]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from oracle.vit_oracle import synthetic_state_dict  # noqa: E402  (weights recipe only)
from vit_colmap_b200.vit import VitBackbone  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "dinov2_vitb14"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
size = int(sys.argv[3]) if len(sys.argv) > 3 else 518
net = VitBackbone(synthetic_state_dict(name, 0), name, 0)
img = torch.randint(0, 256, (B, size, size, 3), dtype=torch.uint8, device="cuda")
for _ in range(3):
    net.forward_patch_tokens(img)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
n = 10
e0.record()
for _ in range(n):
    net.forward_patch_tokens(img)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
g = size // 14
fl = net.flops_per_image(g, g) * B
print(f"{name} B={B} {size}px: {ms:.3f} ms/batch, {B / ms * 1e3:.1f} img/s, {fl / ms / 1e9:.1f} TFLOP/s")
