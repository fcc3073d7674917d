"""Positive / negative example generation for self-supervised training, on the GPU.

Mirror of the reference's ``TrainingSampler`` (vit_colmap/dataloader/training_sampler.py:12-568):
same constructor, same method names, argument meaning and return shapes.  The similarity map, the
selections (top-k / threshold / hybrid), the hard-negative search and every bilinear gather are
kernels behind the C ABI (include/vc_b200.h); what stays on the host is O(B*K) bookkeeping: the 3x3
inverses, the coordinate transforms and the reference's own random draws (made with the same torch /
NumPy calls in the same order, so a seeded run draws the same negatives as the reference on CPU).
There is no CPU fallback."""
from __future__ import annotations

from typing import Tuple

import numpy as np

from ._device import TokenMap, get_device
from .homography_utils import inverse_homography


class TrainingSampler:
    """Generate positive and negative examples for self-supervised learning (training_sampler.py:12)."""

    def __init__(self, top_k_invariant: int = 512, negative_radius: int = 16, num_in_image_negatives: int = 10,
                 num_hard_negatives: int = 5, patch_size: int = 14, selection_mode: str = "top_k",
                 similarity_threshold: float = 0.7, min_invariant_points: int = 100,
                 max_invariant_points: int = 2048, device: int = 0):
        assert selection_mode in ["top_k", "threshold", "hybrid"], f"Invalid selection_mode: {selection_mode}"
        self.top_k_invariant = top_k_invariant
        self.negative_radius = negative_radius
        self.num_in_image_negatives = num_in_image_negatives
        self.num_hard_negatives = num_hard_negatives
        self.patch_size = patch_size
        self.selection_mode = selection_mode
        self.similarity_threshold = similarity_threshold
        self.min_invariant_points = min_invariant_points
        self.max_invariant_points = max_invariant_points
        self.device = int(device)

    # ------------------------------------------------------------------ helpers
    @property
    def _dev(self):
        return get_device(self.device)

    @staticmethod
    def _grid(features) -> Tuple[int, int, int, int]:
        return features.shape_nchw if isinstance(features, TokenMap) else tuple(features.shape)

    # ------------------------------------------------------------------ invariant points
    def similarity_map(self, features1, features2, H, image_size):
        """Cosine similarity of features1 warped by H with features2: (B, H_p, W_p) (:83-97)."""
        dev = self._dev
        return dev.warp_similarity(dev.token_map(features1), dev.token_map(features2), inverse_homography(H),
                                   self.patch_size, image_size)

    def select_invariant_points(self, features1, features2, H, image_size):
        """(B, K, 2) float (x, y) feature-map coordinates of the selected points (:56-181)."""
        import torch
        dev = self._dev
        B, C, H_p, W_p = self._grid(features1)
        n = H_p * W_p
        sim = self.similarity_map(features1, features2, H, image_size).reshape(B, n)
        if self.selection_mode == "top_k":
            k = min(self.top_k_invariant, n)
            order, _, _ = dev.topk_rows(sim, k)
            top = order.to(torch.long)
        else:
            if self.selection_mode == "threshold":
                kmax = min(max(self.min_invariant_points, self.max_invariant_points), n)
            else:
                kmax = min(self.top_k_invariant, n)
            order, above, n_above = dev.topk_rows(sim, kmax, self.similarity_threshold)
            counts = n_above.cpu().tolist()
            rows = []
            for b in range(B):
                m = counts[b]
                if self.selection_mode == "threshold":
                    if m < self.min_invariant_points:      # too few: the best min_points overall (:113-116)
                        idx = order[b, : min(self.min_invariant_points, n)]
                    elif m > self.max_invariant_points:    # too many: the best max_points of them (:117-121)
                        idx = order[b, : self.max_invariant_points]
                    else:                                  # all of them, in index order (:108-110)
                        idx = above[b, :m]
                else:
                    if m == 0:                             # none: plain top-k (:143-146)
                        idx = order[b, : min(self.top_k_invariant, n)]
                    elif m > self.top_k_invariant:         # the best k of those above (:147-151)
                        idx = order[b, : self.top_k_invariant]
                    else:
                        idx = above[b, :m]
                rows.append(idx.to(torch.long))
            max_len = max(len(r) for r in rows)
            top = torch.zeros(B, max_len, dtype=torch.long, device=sim.device)
            for b, r in enumerate(rows):                   # padded with the row's first index (:127-134)
                top[b, : len(r)] = r
                if len(r) < max_len:
                    top[b, len(r):] = r[0]
        return torch.stack([top % W_p, top // W_p], dim=-1).float()

    # ------------------------------------------------------------------ sampling
    def sample_features_at_coords(self, features, coords):
        """Bilinear samples (B, K, C) at (B, K, 2) (x, y) feature-map coordinates (:183-211)."""
        dev = self._dev
        return dev.sample_bilinear(dev.token_map(features), coords)

    def transform_coords_with_homography(self, coords, H, from_feature_space: bool = True, image_size=None,
                                         feature_size=None):
        """(B, K, 2) coordinates through H (B, 3, 3); feature space <-> pixels by the patch size (:213-265).
        O(B*K) host arithmetic in fp32."""
        import torch
        c = coords.detach().to("cpu", torch.float32).numpy()
        Hn = H.detach().to("cpu", torch.float32).numpy()
        scale = from_feature_space and feature_size is not None
        ps = np.float32(self.patch_size)
        if scale:
            c = (c + np.float32(0.5)) * ps
        hom = np.concatenate([c, np.ones(c.shape[:2] + (1,), np.float32)], axis=-1)   # (B, K, 3)
        t = np.einsum("bij,bkj->bki", Hn, hom).astype(np.float32)
        w = np.maximum(t[..., 2:3], np.float32(1e-8))
        xy = (t[..., :2] / w).astype(np.float32)
        if scale:
            xy = xy / ps - np.float32(0.5)
        return torch.from_numpy(np.ascontiguousarray(xy)).to(coords.device)

    def generate_positive_pairs(self, features1, features2, H, invariant_coords, image_size):
        """f1, f2: (B, K, C) features of image 1 / image 2 at corresponding locations (:267-316)."""
        B, C, H_p, W_p = self._grid(features1)
        f2 = self.sample_features_at_coords(features2, invariant_coords)
        coords1 = self.transform_coords_with_homography(invariant_coords, inverse_homography(H),
                                                        from_feature_space=True, image_size=image_size,
                                                        feature_size=(H_p, W_p))
        f1 = self.sample_features_at_coords(features1, coords1)
        return f1, f2

    # ------------------------------------------------------------------ negatives
    def _far_counts(self, positive_coords, H_p: int, W_p: int):
        """For every positive: boolean [HW] mask of the cells at least negative_radius away (:349-353)."""
        pc = positive_coords.detach().to("cpu").float().numpy()                      # (B, K, 2) (x, y)
        xs = np.tile(np.arange(W_p, dtype=np.float32), H_p)
        ys = np.repeat(np.arange(H_p, dtype=np.float32), W_p)
        dx = xs[None, None, :] - pc[..., 0:1]
        dy = ys[None, None, :] - pc[..., 1:2]
        return np.sqrt(dx * dx + dy * dy, dtype=np.float32) >= np.float32(self.negative_radius), xs, ys

    def generate_in_image_negatives(self, features, positive_coords):
        """(B, K, N, C): random cells at least negative_radius from each positive (:318-383).  The draws
        are the reference's own (torch.randperm / torch.randint on the CPU generator, same order)."""
        import torch
        B, C, H_p, W_p = self._grid(features)
        K = positive_coords.shape[1]
        N = self.num_in_image_negatives
        far, xs, ys = self._far_counts(positive_coords, H_p, W_p)
        chosen = np.zeros((B, K, N), np.int64)
        for b in range(B):
            for k in range(K):
                valid = np.nonzero(far[b, k])[0]
                if len(valid) >= N:
                    perm = torch.randperm(len(valid))[:N].numpy()
                    idx = valid[perm]
                else:
                    idx = valid
                    if len(idx) < N:
                        pad = torch.randint(0, H_p * W_p, (N - len(idx),)).numpy()
                        idx = np.concatenate([idx, pad])
                chosen[b, k] = idx
        xy = np.stack([xs[chosen], ys[chosen]], axis=-1).reshape(B, K * N, 2)
        out = self.sample_features_at_coords(features, torch.from_numpy(xy))
        return out.view(B, K, N, C)

    def generate_cross_image_negatives(self, features_batch, positive_coords):
        """(B, K, N, C): features at random cells of the other images of the batch (:385-450).  Draws as
        in the reference: np.random.choice for the images, torch.randint for x then y."""
        import torch
        dev = self._dev
        B, C, H_p, W_p = self._grid(features_batch)
        K = positive_coords.shape[1]
        N = self.num_in_image_negatives
        if B < 2:
            return torch.zeros(B, K, N, C, device=dev._dev())
        tm = dev.token_map(features_batch)
        n_sel = min(N, B - 1)
        img = np.zeros((B, K, n_sel), np.int64)
        xy = np.zeros((B, K, n_sel, 2), np.float32)
        for b in range(B):
            others = [i for i in range(B) if i != b]
            for k in range(K):
                sel = np.random.choice(others, size=n_sel, replace=True)
                for j, i in enumerate(sel):
                    rx = torch.randint(0, W_p, (1,)).float()
                    ry = torch.randint(0, H_p, (1,)).float()
                    img[b, k, j] = i
                    xy[b, k, j] = (float(rx), float(ry))
        # one gather per source image: the points drawn from image i, scattered back to their (b, k, j)
        flat_img = img.reshape(-1)
        flat_xy = xy.reshape(-1, 2)
        out = torch.zeros(B * K * n_sel, C, device=dev._dev())
        per = [np.nonzero(flat_img == i)[0] for i in range(B)]
        m = max(len(p) for p in per)
        pts = np.full((B, max(m, 1), 2), np.nan, np.float32)
        for i, p in enumerate(per):
            pts[i, : len(p)] = flat_xy[p]
        got = dev.sample_bilinear(tm, torch.from_numpy(pts))
        for i, p in enumerate(per):
            if len(p):
                out[torch.from_numpy(p).to(out.device)] = got[i, : len(p)]
        return out.view(B, K, n_sel, C)

    def generate_hard_negatives(self, features, positive_features, positive_coords):
        """(B, K, N_hard, C): the cells most similar to each positive that lie at least negative_radius
        away, sampled like every other point; zero rows where fewer qualify (:452-534)."""
        import torch
        dev = self._dev
        tm = dev.token_map(features)
        B, C, H_p, W_p = self._grid(features)
        K = positive_coords.shape[1]
        N = self.num_hard_negatives
        idx, _ = dev.hard_negatives(tm, positive_features, positive_coords, N, float(self.negative_radius))
        idx = idx.reshape(B, K * N)
        xy = torch.stack([(idx % W_p).float(), (idx // W_p).float()], dim=-1)
        xy[idx < 0] = float("nan")      # padding entries come back as zero rows
        return dev.sample_bilinear(tm, xy).view(B, K, N, C)

    def generate_all_negatives(self, features1, features2, positive_features, positive_coords):
        """In-image (image 2), cross-image (image 1) and hard (image 2) negatives, concatenated (:536-568)."""
        import torch
        dev = self._dev
        tm1, tm2 = dev.token_map(features1), dev.token_map(features2)
        a = self.generate_in_image_negatives(tm2, positive_coords)
        b = self.generate_cross_image_negatives(tm1, positive_coords)
        c = self.generate_hard_negatives(tm2, positive_features, positive_coords)
        return torch.cat([a, b, c], dim=2)
