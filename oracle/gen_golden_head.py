"""Generates tests/golden/head_*.npz by running the REFERENCE's own ViTFeatureModel
(vit_colmap/model/vit_feature_model.py:32-295) in eval mode on seeded backbone features with the
seeded head weights of oracle.head_oracle.synthetic_head_state_dict.  torch.hub.load is replaced by
a stub backbone (the DINOv2 hub model cannot be downloaded here); the head code that runs is the
reference's, byte for byte.  Build container only (needs /root/reference).

    python -m oracle.gen_golden_head
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")

CASES = [
    # name, backbone, C, descriptor_dim, grid (H, W), seed
    ("s4x5_d128", "dinov2_vits14", 384, 128, (4, 5), 11),
    ("b3x3_d64", "dinov2_vitb14", 768, 64, (3, 3), 12),
    ("s9x6_d32", "dinov2_vits14", 384, 32, (9, 6), 13),
]


class _StubBackbone(torch.nn.Module):
    def __init__(self, C):
        super().__init__()
        self.embed_dim = C
        self.feat = None  # [1, HW, C], set per run

    def forward_features(self, x):
        return {"x_norm_patchtokens": self.feat}


def synthetic_backbone_features(seed: int, C: int, H: int, W: int) -> torch.Tensor:
    rng = np.random.default_rng(seed)
    return torch.from_numpy(rng.standard_normal((1, C, H, W)).astype(np.float32))


def run_reference(backbone, C, d, grid, seed):
    sys.path.insert(0, str(REF))
    from oracle.head_oracle import synthetic_head_state_dict
    stub = _StubBackbone(C)
    real_load = torch.hub.load
    torch.hub.load = lambda *a, **k: stub
    try:
        from vit_colmap.model.vit_feature_model import ViTFeatureModel
        model = ViTFeatureModel(backbone_name=backbone, descriptor_dim=d, freeze_backbone=True)
    finally:
        torch.hub.load = real_load
    sd = synthetic_head_state_dict(C, d, seed)
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected and all(k.startswith("backbone.") for k in missing), (missing, unexpected)
    model.eval()
    H, W = grid
    feat = synthetic_backbone_features(seed + 1000, C, H, W)
    with torch.inference_mode():
        a = model.forward_from_backbone_features(feat)
        # the image path: forward() pulls the same features through the (stub) backbone
        stub.feat = feat.permute(0, 2, 3, 1).reshape(1, H * W, C)
        b = model(torch.zeros(1, 3, 14 * H, 14 * W))
    for k in a:
        assert torch.equal(a[k], b[k]), k
    return {k: v.numpy() for k, v in a.items()}


def main():
    out_dir = ROOT / "tests" / "golden"
    for name, backbone, C, d, grid, seed in CASES:
        out = run_reference(backbone, C, d, grid, seed)
        np.savez_compressed(out_dir / f"head_{name}.npz", meta=np.array([C, d, grid[0], grid[1], seed], np.int64),
                            keypoints=out["keypoints"], descriptors=out["descriptors"].astype(np.float16),
                            features_sample=out["features"][:, ::16])
        print(name, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
