"""Similarity of two patch-token maps after homography warping, on the GPU.

Reference: vit_colmap/dataloader/homography_utils.py ``compute_feature_similarity`` :219-253
(= ``warp_patch_tokens`` :10-92 + F.normalize + channel sum).  One fused kernel
(vc_warp_similarity): the warped map is never materialised."""
from __future__ import annotations

from typing import Tuple

from ._device import get_device


def inverse_homography(H):
    """torch.linalg.inv(H) in fp32 on the host, as the reference computes it (homography_utils.py:38)."""
    import torch
    return torch.linalg.inv(H.detach().to(device="cpu", dtype=torch.float32))


def compute_feature_similarity(features1, features2, H, patch_size: int = 14,
                               input_size: Tuple[int, int] = (1200, 1600), device: int = 0):
    """features1 / features2: (B, C, H_p, W_p) tensors (or TokenMap); H: (B, 3, 3) source -> target.
    Returns the cosine similarity map (B, H_p, W_p) f32 on the GPU."""
    dev = get_device(device)
    tm1, tm2 = dev.token_map(features1), dev.token_map(features2)
    return dev.warp_similarity(tm1, tm2, inverse_homography(H), patch_size, input_size)
