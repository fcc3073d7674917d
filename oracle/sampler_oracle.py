"""CPU oracle of the training-time samplers.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates, vectorised and with a fixed tie order, the reference's
  vit_colmap/dataloader/homography_utils.py  warp_patch_tokens :10-92, compute_feature_similarity :219-253
  vit_colmap/dataloader/training_sampler.py  select_invariant_points :56-181, sample_features_at_coords
                                             :183-211, transform_coords_with_homography :213-265,
                                             generate_positive_pairs :267-316, generate_hard_negatives :452-534
on CPU torch fp32.  F.grid_sample / torch.linalg.inv are the reference's own library calls and are
used directly.  Pinned by tests/golden/sampler_*.npz, which hold the outputs of
the REFERENCE's own functions on seeded inputs (oracle/gen_golden_sampler.py).
Tie order: torch.topk does not define one; here (and in the kernels) equal values come out lower
index first.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np
import torch
import torch.nn.functional as F


def synthetic_pair(seed: int, B: int, C: int, Hp: int, Wp: int, patch: int = 14, noise: float = 0.3):
    """Seeded (features1, features2, H, image_size): smooth-ish maps, image 2 = image 1 seen through a
    mild random homography plus noise, so the similarity map has structure and an invalid border."""
    rng = np.random.default_rng(seed)
    img_h, img_w = Hp * patch, Wp * patch
    base = rng.standard_normal((B, C, Hp, Wp)).astype(np.float32)
    low = rng.standard_normal((B, C, (Hp + 3) // 4, (Wp + 3) // 4)).astype(np.float32)
    low = np.repeat(np.repeat(low, 4, axis=2), 4, axis=3)[:, :, :Hp, :Wp]
    f1 = torch.from_numpy(0.5 * base + low)
    Hs = []
    for _ in range(B):
        a = rng.uniform(-0.08, 0.08)
        s = 1.0 + rng.uniform(-0.05, 0.05)
        tx, ty = rng.uniform(-0.1, 0.1) * img_w, rng.uniform(-0.1, 0.1) * img_h
        p = rng.uniform(-1e-5, 1e-5, 2)
        Hs.append([[s * np.cos(a), -s * np.sin(a), tx], [s * np.sin(a), s * np.cos(a), ty], [p[0], p[1], 1.0]])
    H = torch.tensor(np.array(Hs), dtype=torch.float32)
    f2 = warp(f1, H, patch, (img_h, img_w)) + noise * torch.from_numpy(
        rng.standard_normal((B, C, Hp, Wp)).astype(np.float32))
    return f1, f2, H, (img_h, img_w)


def source_grid(H: torch.Tensor, Hp: int, Wp: int, patch: int, image_size) -> torch.Tensor:
    """Normalised sampling grid (B, Hp, Wp, 2) of warp_patch_tokens (homography_utils.py:36-78)."""
    B = H.shape[0]
    H_inv = torch.linalg.inv(H)
    ys = (torch.arange(Hp, dtype=torch.float32) + 0.5) * patch
    xs = (torch.arange(Wp, dtype=torch.float32) + 0.5) * patch
    yy, xx = torch.meshgrid(ys, xs, indexing="ij")
    pts = torch.stack([xx, yy, torch.ones_like(xx)], dim=-1).reshape(-1, 3)
    src = torch.bmm(H_inv, pts.unsqueeze(0).expand(B, -1, -1).permute(0, 2, 1)).permute(0, 2, 1)
    xy = src[:, :, :2] / src[:, :, 2:3].clamp(min=1e-8)
    xy = xy.reshape(B, Hp, Wp, 2).clone()
    img_h, img_w = image_size
    xy[..., 0] = (xy[..., 0] / img_w) * 2 - 1
    xy[..., 1] = (xy[..., 1] / img_h) * 2 - 1
    return xy


def warp(f1: torch.Tensor, H: torch.Tensor, patch: int, image_size) -> torch.Tensor:
    """warp_patch_tokens :10-92."""
    B, C, Hp, Wp = f1.shape
    return F.grid_sample(f1, source_grid(H, Hp, Wp, patch, image_size), mode="bilinear", padding_mode="zeros",
                         align_corners=False)


def similarity_map(f1, f2, H, patch: int, image_size) -> torch.Tensor:
    """compute_feature_similarity :219-253 / select_invariant_points :83-97 -> (B, Hp, Wp)."""
    w = F.normalize(warp(f1, H, patch, image_size), p=2, dim=1)
    t = F.normalize(f2, p=2, dim=1)
    return (w * t).sum(dim=1)


def _topk_stable(v: torch.Tensor, k: int) -> torch.Tensor:
    """Indices of the k largest values, lower index first among equal values."""
    return torch.sort(v, descending=True, stable=True).indices[:k]


def select_indices(sim_flat: torch.Tensor, mode: str, top_k: int, threshold: float, min_pts: int,
                   max_pts: int) -> List[torch.Tensor]:
    """Per image, the flat indices select_invariant_points picks (:100-151), before padding."""
    out = []
    n = sim_flat.shape[1]
    for b in range(sim_flat.shape[0]):
        v = sim_flat[b]
        if mode == "top_k":
            out.append(_topk_stable(v, min(top_k, n)))
            continue
        above = torch.nonzero(v >= threshold, as_tuple=True)[0]
        if mode == "threshold":
            if len(above) < min_pts:
                idx = _topk_stable(v, min(min_pts, n))
            elif len(above) > max_pts:
                idx = above[_topk_stable(v[above], max_pts)]
            else:
                idx = above
        else:
            if len(above) == 0:
                idx = _topk_stable(v, min(top_k, n))
            elif len(above) > top_k:
                idx = above[_topk_stable(v[above], top_k)]
            else:
                idx = above
        out.append(idx)
    return out


def pad_and_coords(indices: List[torch.Tensor], Wp: int) -> torch.Tensor:
    """:124-134 (pad every row to the longest with its first index) and :168-179 -> (B, K, 2) float (x, y)."""
    m = max(len(i) for i in indices)
    top = torch.zeros(len(indices), m, dtype=torch.long)
    for b, i in enumerate(indices):
        top[b, : len(i)] = i
        if len(i) < m:
            top[b, len(i):] = i[0]
    return torch.stack([top % Wp, top // Wp], dim=-1).float()


def sample_at(features: torch.Tensor, coords: torch.Tensor) -> torch.Tensor:
    """sample_features_at_coords :183-211 -> (B, K, C)."""
    B, C, Hp, Wp = features.shape
    g = coords.clone().float()
    g[..., 0] = (g[..., 0] / (Wp - 1)) * 2 - 1
    g[..., 1] = (g[..., 1] / (Hp - 1)) * 2 - 1
    s = F.grid_sample(features, g.unsqueeze(2), mode="bilinear", align_corners=False)
    return s.squeeze(-1).permute(0, 2, 1)


def transform_coords(coords: torch.Tensor, H: torch.Tensor, patch: int, feature_space: bool = True) -> torch.Tensor:
    """transform_coords_with_homography :213-265."""
    c = coords.clone()
    if feature_space:
        c = (coords + 0.5) * patch
    hom = torch.cat([c, torch.ones_like(c[..., :1])], dim=-1)
    t = torch.bmm(H, hom.permute(0, 2, 1)).permute(0, 2, 1)
    xy = t[..., :2] / t[..., 2:3].clamp(min=1e-8)
    if feature_space:
        xy = xy / patch - 0.5
    return xy


def positive_pairs(f1, f2, H, coords, patch: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """generate_positive_pairs :267-316."""
    return sample_at(f1, transform_coords(coords, torch.linalg.inv(H), patch)), sample_at(f2, coords)


def hard_negative_indices(features, pos_feat, pos_xy, n_hard: int, radius: float):
    """generate_hard_negatives :476-512 without the gather: (idx [B,K,n_hard] int64, -1 padded; counts [B,K];
    masked similarities [B,K,HW])."""
    B, C, Hp, Wp = features.shape
    flat = F.normalize(features.view(B, C, -1).permute(0, 2, 1), p=2, dim=-1)       # (B, HW, C)
    pos = F.normalize(pos_feat, p=2, dim=-1)                                        # (B, K, C)
    sims = torch.bmm(pos, flat.permute(0, 2, 1))                                    # (B, K, HW)
    ys, xs = torch.meshgrid(torch.arange(Hp, dtype=torch.float32), torch.arange(Wp, dtype=torch.float32),
                            indexing="ij")
    allc = torch.stack([xs.flatten(), ys.flatten()], dim=-1)                        # (HW, 2)
    dist = torch.norm(allc[None, None] - pos_xy[:, :, None, :], dim=-1)             # (B, K, HW)
    far = dist >= radius
    masked = sims.masked_fill(~far, -float("inf"))
    K = pos_xy.shape[1]
    idx = torch.full((B, K, n_hard), -1, dtype=torch.long)
    counts = torch.zeros(B, K, dtype=torch.long)
    for b in range(B):
        for k in range(K):
            m = min(n_hard, int(far[b, k].sum()))
            idx[b, k, :m] = _topk_stable(masked[b, k], m)
            counts[b, k] = m
    return idx, counts, masked


def hard_negatives(features, pos_feat, pos_xy, n_hard: int, radius: float) -> torch.Tensor:
    """generate_hard_negatives :452-534 -> (B, K, n_hard, C), zero rows where fewer cells qualify."""
    B, C, Hp, Wp = features.shape
    idx, counts, _ = hard_negative_indices(features, pos_feat, pos_xy, n_hard, radius)
    K = idx.shape[1]
    safe = idx.clamp(min=0).reshape(B, K * n_hard)
    xy = torch.stack([(safe % Wp).float(), (safe // Wp).float()], dim=-1)
    out = sample_at(features, xy).reshape(B, K, n_hard, C).clone()
    out[idx < 0] = 0.0
    return out
