"""Generates tests/golden/hybrid_*.npz by running the REFERENCE's own hybrid extractor code
(vit_colmap/features/hybrid_extractor.py:224-294, _extract_descriptors_at_keypoints).

Runs only in the build container (needs /root/reference).  The extractor object is created with
object.__new__ (no torch.hub, no cv2 detector); keypoints are seeded sub-pixel positions, feature
maps come from oracle.extract_oracle.synthetic_feature_map, so the files hold keypoints + outputs.

    python -m oracle.gen_golden_hybrid
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")

CASES = [
    # name, seed, C, H, W, n_keypoints, descriptor_dim, orig (w,h), feature (w,h)
    ("h37_direct", 11, 64, 37, 37, 300, 64, (518, 518), (518, 518)),
    ("h42x57_pca", 12, 96, 42, 57, 500, 32, (800, 600), (798, 588)),
    ("h85x114_pca", 13, 48, 85, 114, 900, 16, (1600, 1200), (1596, 1190)),
    ("h9x20_random_proj", 14, 32, 9, 20, 12, 16, (280, 126), (280, 126)),
]


def reference_class():
    import sys
    sys.path.insert(0, str(REF))
    from vit_colmap.features import hybrid_extractor  # the reference module
    return hybrid_extractor.ViTExtractor


def main():
    from oracle.extract_oracle import synthetic_feature_map
    cls = reference_class()
    out_dir = ROOT / "tests" / "golden"
    for name, seed, C, H, W, n, d, orig, feat in CASES:
        fm = synthetic_feature_map(seed, C, H, W)
        rng = np.random.default_rng(seed)
        kps = np.stack([rng.uniform(0, orig[0] - 1, n), rng.uniform(0, orig[1] - 1, n)], axis=1).astype(np.float32)
        kps[0] = (0.0, 0.0)                       # border clamps
        kps[1] = (orig[0] - 1, orig[1] - 1)
        ex = object.__new__(cls)
        ex.device = torch.device("cpu")
        ex.descriptor_dim = d
        ex.descriptor_projection = None
        torch.manual_seed(seed)                   # the m <= d branch draws torch.randn
        desc_u8 = ex._extract_descriptors_at_keypoints(fm, kps, orig, feat)
        proj = ex.descriptor_projection
        np.savez_compressed(out_dir / f"hybrid_{name}.npz",
                            meta=np.array([seed, C, H, W, n, d, orig[0], orig[1], feat[0], feat[1]], np.int64),
                            keypoints=kps, desc_u8=desc_u8,
                            projection=(proj.numpy() if proj is not None else np.zeros((0, 0), np.float32)))
        print(f"{name}: {desc_u8.shape}, projection {None if proj is None else tuple(proj.shape)}")


if __name__ == "__main__":
    main()
