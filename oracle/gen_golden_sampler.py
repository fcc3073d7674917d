"""Generates tests/golden/sampler_*.npz by running the REFERENCE's own training-sampler code
(vit_colmap/dataloader/training_sampler.py TrainingSampler, vit_colmap/dataloader/homography_utils.py
compute_feature_similarity) on seeded inputs from oracle.sampler_oracle.synthetic_pair.

Runs only in the build container (needs /root/reference); CPU, fp32.  The files hold the seeds and
the reference's outputs; inputs are regenerated from the seed by the tests.

    python -m oracle.gen_golden_sampler
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")

CASES = [
    # name, seed, B, C, Hp, Wp, top_k, threshold, min_pts, max_pts, radius, n_in, n_hard
    ("s37", 21, 2, 64, 37, 37, 128, 0.7, 40, 300, 16, 4, 5),
    ("s42x57", 22, 2, 96, 42, 57, 256, 0.8, 100, 500, 16, 3, 5),
    ("s12x9_small", 23, 3, 32, 12, 9, 64, 0.9, 20, 30, 6, 2, 8),
    ("s20x20_wide_radius", 24, 1, 48, 20, 20, 50, 0.5, 10, 2048, 26, 3, 6),
]


def main():
    # the reference's package __init__ files import pycolmap (absent here): register the two packages as
    # bare namespaces so that its sampler modules load, unmodified, from where they lie
    import types
    for pkg, path in (("vit_colmap", REF / "vit_colmap"), ("vit_colmap.dataloader", REF / "vit_colmap" / "dataloader")):
        m = types.ModuleType(pkg)
        m.__path__ = [str(path)]
        sys.modules[pkg] = m
    from vit_colmap.dataloader.homography_utils import compute_feature_similarity  # the reference
    from vit_colmap.dataloader.training_sampler import TrainingSampler              # the reference
    from oracle.sampler_oracle import synthetic_pair
    out_dir = ROOT / "tests" / "golden"
    for name, seed, B, C, Hp, Wp, top_k, thr, mn, mx, radius, n_in, n_hard in CASES:
        f1, f2, H, image_size = synthetic_pair(seed, B, C, Hp, Wp)
        sim = compute_feature_similarity(f1, f2, H, patch_size=14, input_size=image_size)
        out = {"meta": np.array([seed, B, C, Hp, Wp, top_k, mn, mx, radius, n_in, n_hard], np.int64),
               "threshold": np.float32(thr), "similarity": sim.numpy()}
        for mode in ("top_k", "threshold", "hybrid"):
            s = TrainingSampler(top_k_invariant=top_k, negative_radius=radius, num_in_image_negatives=n_in,
                                num_hard_negatives=n_hard, selection_mode=mode, similarity_threshold=thr,
                                min_invariant_points=mn, max_invariant_points=mx)
            out[f"coords_{mode}"] = s.select_invariant_points(f1, f2, H, image_size).numpy()
        s = TrainingSampler(top_k_invariant=top_k, negative_radius=radius, num_in_image_negatives=n_in,
                            num_hard_negatives=n_hard)
        coords = torch.from_numpy(out["coords_top_k"])[:, :24].contiguous()   # the python loops are slow
        p1, p2 = s.generate_positive_pairs(f1, f2, H, coords, image_size)
        out["pos_coords"] = coords.numpy()
        out["pos_f1"], out["pos_f2"] = p1.numpy(), p2.numpy()
        out["hard"] = s.generate_hard_negatives(f2, p2, coords).numpy()
        torch.manual_seed(seed)
        np.random.seed(seed)
        out["in_image"] = s.generate_in_image_negatives(f2, coords).numpy()
        out["cross_image"] = s.generate_cross_image_negatives(f1, coords).numpy()
        torch.manual_seed(seed + 1)
        np.random.seed(seed + 1)
        out["all_negatives"] = s.generate_all_negatives(f1, f2, p2, coords).numpy()
        np.savez_compressed(out_dir / f"sampler_{name}.npz", **out)
        print(name, {k: v.shape for k, v in out.items() if hasattr(v, "shape") and v.ndim > 0})


if __name__ == "__main__":
    main()
