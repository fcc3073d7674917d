"""Generates tests/golden/trainable_*.npz by running the REFERENCE's own
TrainableViTExtractor._run_inference (vit_colmap/features/trainable_vit_extractor.py:140-269) with
the network replaced by a stub that returns seeded head outputs, so the selection / packing code
that runs is the reference's, byte for byte.  Build container only (needs /root/reference).

    python -m oracle.gen_golden_trainable
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference")

CASES = [
    # name, seed, d, image (h, w), num_keypoints, score_threshold, nms_radius
    ("t518_r1", 1, 128, (518, 518), 20480, 0.4, 1),      # the Pipeline's settings, run_pipeline.py:329-333
    ("t518_r4", 2, 64, (518, 518), 2048, 0.0, 4),        # the class defaults
    ("t480x650", 3, 32, (480, 650), 300, 0.5, 2),        # resized to 476x644, top-K truncation
    ("t56_none", 4, 16, (56, 70), 64, 0.9999, 1),        # threshold removes (almost) everything
]


def run_reference(seed, d, hw, K, thr, r):
    sys.path.insert(0, str(REF))
    import torchvision.transforms as transforms
    from vit_colmap.features.trainable_vit_extractor import TrainableViTExtractor
    from oracle.extract_oracle import synthetic_head_maps
    h, w = hw
    h_new, w_new = (h // 14) * 14, (w // 14) * 14
    h4, w4 = h_new // 4, w_new // 4
    kp, desc = synthetic_head_maps(seed, d, h4, w4)
    ex = object.__new__(TrainableViTExtractor)
    ex.device = torch.device("cpu")
    ex.patch_size = 14
    ex.num_keypoints, ex.score_threshold, ex.nms_radius, ex.descriptor_dim = K, thr, r, d
    ex.transform = transforms.Compose([transforms.ToTensor()])
    ex.model = lambda x: {"keypoints": kp[None], "descriptors": desc[None]}
    img = np.zeros((h, w, 3), np.uint8)
    kps, du8 = ex._run_inference(img)
    return kps, du8, (h4, w4), (w, h), (w_new, h_new)


def main():
    out_dir = ROOT / "tests" / "golden"
    for name, seed, d, hw, K, thr, r in CASES:
        kps, du8, (h4, w4), orig, res = run_reference(seed, d, hw, K, thr, r)
        np.savez_compressed(out_dir / f"trainable_{name}.npz",
                            meta=np.array([seed, d, h4, w4, K, r, orig[0], orig[1], res[0], res[1]], np.int64),
                            thr=np.array(thr, np.float64), keypoints=kps, desc_u8=du8)
        print(f"{name}: grid {h4}x{w4} -> {len(kps)} keypoints")


if __name__ == "__main__":
    main()
