"""Generates tests/golden/extract_*.npz by running the REFERENCE's own code.

Runs only in the build container (needs /root/reference; pycolmap / torch.hub are not touched:
the extractor object is created with object.__new__ and only its pure post-processing methods are
called).  Inputs are regenerated from seeds by oracle.extract_oracle.synthetic_feature_map, so the
files hold outputs only.

    python -m oracle.gen_golden_extract
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")

CASES = [
    # name, seed, C, H, W, method, num_keypoints, descriptor_dim, orig (w,h), resized (w,h)
    ("s37_harris", 1, 64, 37, 37, "harris", 2048, 64, (518, 518), (518, 518)),
    ("s37_dog", 2, 48, 37, 37, "dog", 2048, 48, (518, 518), (518, 518)),
    ("s37_combined", 3, 48, 37, 37, "combined", 8192, 48, (518, 518), (518, 518)),
    ("hp42x57_pca", 4, 96, 42, 57, "harris", 2048, 32, (800, 600), (798, 588)),
    ("dtu85x114", 5, 32, 85, 114, "harris", 8192, 32, (1600, 1200), (1596, 1190)),
    ("tiny9x20", 6, 16, 9, 20, "harris", 64, 16, (280, 126), (280, 126)),
]


def make_reference_extractor(num_keypoints, descriptor_dim, method):
    sys.path.insert(0, str(REF))
    from vit_colmap.features.vit_extractor import ViTExtractor  # the reference class
    ex = object.__new__(ViTExtractor)
    ex.device = "cpu"
    ex.num_keypoints = num_keypoints
    ex.descriptor_dim = descriptor_dim
    ex.detection_method = method
    ex.descriptor_projection = None
    ex.patch_size = 14
    return ex


def main():
    from oracle.extract_oracle import synthetic_feature_map
    out_dir = ROOT / "tests" / "golden"
    out_dir.mkdir(parents=True, exist_ok=True)
    for name, seed, C, H, W, method, K, d, orig, resized in CASES:
        fm = synthetic_feature_map(seed, C, H, W)
        ex = make_reference_extractor(K, d, method)
        score = ex._compute_distinctiveness(fm, method=method)
        coords, scores = ex._spatial_binning_selection(score, target_keypoints=K, bin_size=16)
        if len(coords) == 0:
            coords, scores = ex._simple_topk_selection(score, K)
        n_cand = len(coords)
        nms_coords, nms_scores = ex._apply_nms(coords, scores, nms_radius=1.5)
        kps, desc_u8 = ex._dense_to_sparse(fm.unsqueeze(0), orig, resized, (H, W))
        proj = ex.descriptor_projection
        np.savez_compressed(
            out_dir / f"extract_{name}.npz",
            meta=np.array([seed, C, H, W, K, d, orig[0], orig[1], resized[0], resized[1]], np.int64),
            method=np.array(method), score=score.numpy().astype(np.float32),
            n_candidates=np.array(n_cand), nms_coords=nms_coords.numpy().astype(np.int64),
            nms_scores=nms_scores.numpy().astype(np.float32), keypoints=kps, desc_u8=desc_u8,
            projection=(proj.numpy() if proj is not None else np.zeros((0, 0), np.float32)))
        print(f"{name}: {n_cand} candidates -> {len(nms_coords)} keypoints, desc {desc_u8.shape}")


if __name__ == "__main__":
    main()
