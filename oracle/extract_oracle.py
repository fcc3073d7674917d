"""Oracle for the extraction post-processing (score map, keypoint selection, NMS, descriptor
sampling, PCA projection, normalise + quantise).  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement (torch CPU / numpy, fp32) of ``vit_colmap/features/vit_extractor.py``:
  _harris_response :281-348      _dog_response :350-394        _compute_distinctiveness :254-279
  _get_gaussian_kernel :396-402  _spatial_binning_selection :404-485
  _simple_topk_selection :487-498  _apply_nms :500-543  _extract_descriptors_interp :545-586
  _reduce_descriptor_dim :588-653  _dense_to_sparse :168-252
and of ``trainable_vit_extractor.py:114-138,173-269`` (NMS / selection / packing).

Pinned against the reference itself: oracle/gen_golden_extract.py imports the reference classes
from /root/reference (possible only in the build container) and stores their outputs under
tests/golden/; tests/test_oracle_extract.py checks this restatement against those files.
Tie order of torch.topk / argsort is unspecified in the reference; the restatement (and the CUDA
kernels) break ties by the lower flat index, which is what the golden vectors show.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def gaussian_kernel(kernel_size: int, sigma: float) -> torch.Tensor:
    """vit_extractor.py:396-402."""
    x = torch.arange(kernel_size, dtype=torch.float32) - kernel_size // 2
    g = torch.exp(-(x ** 2) / (2 * sigma ** 2))
    k1 = g / g.sum()
    return (k1.unsqueeze(0) * k1.unsqueeze(1)).unsqueeze(0).unsqueeze(0)


def _minmax(score: torch.Tensor) -> torch.Tensor:
    score = score - score.min()
    if score.max() > 0:
        score = score / score.max()
    return score


def harris_moments(fm: torch.Tensor):
    """Structure-tensor moments before smoothing (vit_extractor.py:298-309). fm: [C,H,W]."""
    gx = F.pad(fm[:, :, 1:] - fm[:, :, :-1], (0, 1, 0, 0))
    gy = F.pad(fm[:, 1:, :] - fm[:, :-1, :], (0, 0, 0, 1))
    return (gx ** 2).mean(dim=0), (gy ** 2).mean(dim=0), (gx * gy).mean(dim=0)


def harris_response(fm: torch.Tensor, normalize: bool = True) -> torch.Tensor:
    """vit_extractor.py:281-348."""
    ixx, iyy, ixy = harris_moments(fm)
    k = gaussian_kernel(3, 1.0)
    sm = lambda t: F.conv2d(t[None, None], k, padding=1).squeeze()  # noqa: E731
    ixx, iyy, ixy = sm(ixx), sm(iyy), sm(ixy)
    det = ixx * iyy - ixy ** 2
    trace = ixx + iyy
    corner = det - 0.04 * (trace ** 2)
    edge = torch.sqrt(ixx + iyy)
    score = 0.7 * corner + 0.3 * edge
    return _minmax(score) if normalize else score


def dog_response(fm: torch.Tensor) -> torch.Tensor:
    """vit_extractor.py:350-394 (kernel sizes int(6*sigma+1) made odd: 7 and 11)."""
    avg = fm.mean(dim=0, keepdim=True).unsqueeze(0)
    out = []
    for sigma in (1.0, 1.6):
        ks = int(6 * sigma + 1)
        ks = ks if ks % 2 == 1 else ks + 1
        out.append(F.conv2d(avg, gaussian_kernel(ks, sigma), padding=ks // 2))
    return _minmax(torch.abs(out[0] - out[1]).squeeze())


def distinctiveness(fm: torch.Tensor, method: str = "harris") -> torch.Tensor:
    """vit_extractor.py:254-279."""
    if method == "harris":
        return harris_response(fm)
    if method == "dog":
        return dog_response(fm)
    if method == "combined":
        h, d = harris_response(fm), dog_response(fm)
        h = (h - h.min()) / (h.max() - h.min() + 1e-8)
        d = (d - d.min()) / (d.max() - d.min() + 1e-8)
        return 0.5 * h + 0.5 * d
    raise ValueError(f"Unknown detection method: {method}")


def _topk_stable(values: np.ndarray, k: int):
    """top-k by value descending, ties by lower index."""
    order = np.lexsort((np.arange(len(values)), -values.astype(np.float64)))[:k]
    return values[order], order


def spatial_binning_selection(score: np.ndarray, target: int, bin_size: int = 16):
    """vit_extractor.py:404-485.  Returns (coords [n,2] int64 (y,x), scores [n])."""
    H, W = score.shape
    nbh, nbw = max(1, H // bin_size), max(1, W // bin_size)
    per_bin = max(1, target // (nbh * nbw))
    coords, scores = [], []
    for i in range(nbh):
        for j in range(nbw):
            h0, h1 = i * bin_size, min((i + 1) * bin_size, H)
            w0, w1 = j * bin_size, min((j + 1) * bin_size, W)
            blk = score[h0:h1, w0:w1]
            if blk.size == 0:
                continue
            flat = blk.reshape(-1)
            k = min(per_bin, flat.size)
            vals, idx = _topk_stable(flat, k)
            bw = w1 - w0
            coords.append(np.stack([idx // bw + h0, idx % bw + w0], axis=1))
            scores.append(vals)
    if not coords:
        return np.zeros((0, 2), np.int64), np.zeros((0,), np.float32)
    coords, scores = np.concatenate(coords).astype(np.int64), np.concatenate(scores)
    if len(coords) > target:
        scores, idx = _topk_stable(scores, target)
        coords = coords[idx]
    return coords, scores


def simple_topk_selection(score: np.ndarray, k: int):
    """vit_extractor.py:487-498."""
    flat = score.reshape(-1)
    vals, idx = _topk_stable(flat, min(k, flat.size))
    W = score.shape[1]
    return np.stack([idx // W, idx % W], axis=1).astype(np.int64), vals


def apply_nms(coords: np.ndarray, scores: np.ndarray, radius: float = 1.5):
    """vit_extractor.py:500-543: greedy in descending score order; suppress 0 < dist < radius."""
    if len(coords) == 0:
        return coords, scores
    order = np.lexsort((np.arange(len(scores)), -scores.astype(np.float64)))
    c, s = coords[order], scores[order]
    keep = np.ones(len(c), bool)
    cf = c.astype(np.float32)
    for i in range(len(c)):
        if not keep[i]:
            continue
        d = np.sqrt(((cf - cf[i]) ** 2).sum(axis=1))
        keep[(d < radius) & (d > 0)] = False
    return c[keep], s[keep]


def sample_descriptors(fm: torch.Tensor, coords: np.ndarray) -> torch.Tensor:
    """vit_extractor.py:545-586 (grid_sample bilinear, border, align_corners=True)."""
    C, H, W = fm.shape
    if len(coords) == 0:
        return torch.empty((0, C))
    cn = torch.from_numpy(coords).float()
    cn[:, 0] = 2.0 * cn[:, 0] / (H - 1) - 1.0
    cn[:, 1] = 2.0 * cn[:, 1] / (W - 1) - 1.0
    grid = cn[:, [1, 0]].unsqueeze(0).unsqueeze(2)
    d = F.grid_sample(fm.unsqueeze(0), grid, mode="bilinear", padding_mode="border", align_corners=True)
    return d.squeeze(0).squeeze(2).t()


def fit_pca(desc: torch.Tensor, d: int, generator=None) -> torch.Tensor:
    """vit_extractor.py:601-639: first-call projection, [C,d]."""
    if desc.shape[0] > d:
        centered = desc - desc.mean(dim=0, keepdim=True)
        _, _, V = torch.svd(centered.cpu())
        return V[:, :d].contiguous()
    proj = torch.randn(desc.shape[1], d, dtype=torch.float32, generator=generator)
    return proj / np.sqrt(desc.shape[1])


def keypoint_xy(coords: np.ndarray, grid_hw, resized_wh, orig_wh) -> np.ndarray:
    """vit_extractor.py:229-236 in fp32 tensor arithmetic with Python-float scales."""
    H, W = grid_hw
    w_res, h_res = resized_wh
    w_orig, h_orig = orig_wh
    c = torch.from_numpy(coords.astype(np.int64))
    x = (c[:, 1] + 0.5) * (w_res / W) * (w_orig / w_res)
    y = (c[:, 0] + 0.5) * (h_res / H) * (h_orig / h_res)
    return torch.stack([x, y], dim=1).numpy().astype(np.float32)


def quantize_vit(desc: torch.Tensor) -> np.ndarray:
    """vit_extractor.py:243-250: L2 normalise (eps 1e-12), trunc(clip(512*v, 0, 255))."""
    d = F.normalize(desc, p=2, dim=1).numpy()
    return (d * 512.0).clip(0, 255).astype(np.uint8)


def dense_to_sparse(fm: torch.Tensor, orig_wh, resized_wh, num_keypoints=2048, descriptor_dim=128,
                    method="harris", projection=None):
    """vit_extractor.py:168-252.  Returns dict with every intermediate the tests compare."""
    C, H, W = fm.shape
    score = distinctiveness(fm, method)
    coords, scores = spatial_binning_selection(score.numpy(), num_keypoints, 16)
    if len(coords) == 0:
        coords, scores = simple_topk_selection(score.numpy(), num_keypoints)
    coords, scores = apply_nms(coords, scores, 1.5)
    desc = sample_descriptors(fm, coords)
    kps = keypoint_xy(coords, (H, W), resized_wh, orig_wh)
    if desc.shape[1] > descriptor_dim:
        if projection is None:
            projection = fit_pca(desc, descriptor_dim)
        desc = desc @ projection
    desc_n = F.normalize(desc, p=2, dim=1)
    return {"score": score.numpy(), "coords": coords, "scores": scores, "keypoints": kps,
            "desc_float": desc_n.numpy(), "desc_u8": quantize_vit(desc), "projection": projection}


def synthetic_feature_map(seed: int, C: int, H: int, W: int) -> torch.Tensor:
    """Seeded smooth-ish feature map [C,H,W] fp32 (numpy Generator: stable across versions)."""
    rng = np.random.default_rng(seed)
    base = rng.standard_normal((C, H, W)).astype(np.float32)
    low = rng.standard_normal((C, (H + 3) // 4, (W + 3) // 4)).astype(np.float32)
    low = np.repeat(np.repeat(low, 4, axis=1), 4, axis=2)[:, :H, :W]
    return torch.from_numpy(0.5 * base + low)


# ------------------------------------------------------------------ hybrid extractor
def hybrid_grid_coords(keypoints_xy: np.ndarray, orig_wh, feat_wh, grid_hw) -> np.ndarray:
    """hybrid_extractor.py:247-255: detector pixels -> feature-map units, f32 array arithmetic with
    Python-float scales.  Returns [n,2] (y, x) float32."""
    H, W = grid_hw
    w_orig, h_orig = orig_wh
    w_feat, h_feat = feat_wh
    kp = np.asarray(keypoints_xy, np.float32).reshape(-1, 2)
    kx = kp[:, 0] * (w_feat / w_orig) * (W / w_feat)
    ky = kp[:, 1] * (h_feat / h_orig) * (H / h_feat)
    return np.stack([ky, kx], axis=1).astype(np.float32)


def hybrid_sample(fm: torch.Tensor, coords_yx: np.ndarray) -> torch.Tensor:
    """hybrid_extractor.py:257-279: grid_sample (bilinear, border, align_corners=True) at float
    feature-map coordinates; [n, C] raw rows."""
    C, H, W = fm.shape
    gx = (2.0 * coords_yx[:, 1] / (W - 1) - 1.0).astype(np.float32)
    gy = (2.0 * coords_yx[:, 0] / (H - 1) - 1.0).astype(np.float32)
    grid = torch.stack([torch.from_numpy(gx), torch.from_numpy(gy)], dim=1).unsqueeze(0).unsqueeze(2)
    d = F.grid_sample(fm.unsqueeze(0), grid, mode="bilinear", padding_mode="border", align_corners=True)
    return d.squeeze(0).squeeze(2).t()


def rootsift(desc: torch.Tensor) -> torch.Tensor:
    """hybrid_extractor.py:286-288: L1 normalise, sqrt(clamp(min=1e-8)), L2 normalise."""
    d = F.normalize(desc, p=1, dim=1)
    d = torch.sqrt(torch.clamp(d, min=1e-8))
    return F.normalize(d, p=2, dim=1)


def hybrid_descriptors(fm: torch.Tensor, keypoints_xy: np.ndarray, orig_wh, feat_wh, descriptor_dim: int,
                       projection=None):
    """hybrid_extractor.py:224-294 (_extract_descriptors_at_keypoints)."""
    C, H, W = fm.shape
    if len(keypoints_xy) == 0:
        return {"desc_float": np.zeros((0, descriptor_dim), np.float32),
                "desc_u8": np.zeros((0, descriptor_dim), np.uint8), "projection": projection}
    coords = hybrid_grid_coords(keypoints_xy, orig_wh, feat_wh, (H, W))
    desc = hybrid_sample(fm, coords)
    if desc.shape[1] > descriptor_dim:
        if projection is None:
            projection = fit_pca(desc, descriptor_dim)
        desc = desc @ projection
    dn = rootsift(desc)
    return {"coords": coords, "desc_float": dn.numpy(), "projection": projection,
            "desc_u8": (dn.numpy() * 512.0).clip(0, 255).astype(np.uint8)}


# ------------------------------------------------------------------ trainable_vit extractor
def simple_nms(scores: torch.Tensor, radius: int) -> torch.Tensor:
    """trainable_vit_extractor.py:114-138: keep == max_pool2d(k=2r+1, s=1, pad=r)."""
    s = scores[None, None]
    mx = F.max_pool2d(s, kernel_size=2 * radius + 1, stride=1, padding=radius)
    return (s == mx).squeeze(0).squeeze(0)


def synthetic_head_maps(seed: int, d: int, h4: int, w4: int):
    """Seeded stand-ins for ViTFeatureModel's outputs: keypoint map [4,h4,w4] (score logit, dx, dy,
    orientation already in [-pi, pi]) and an L2-normalised descriptor map [d,h4,w4]."""
    rng = np.random.default_rng(seed)
    low = rng.standard_normal((1, (h4 + 2) // 3, (w4 + 2) // 3)).astype(np.float32)
    logit = np.repeat(np.repeat(low, 3, axis=1), 3, axis=2)[:, :h4, :w4] + 0.7 * rng.standard_normal((1, h4, w4)).astype(np.float32)
    off = 0.5 * np.tanh(rng.standard_normal((2, h4, w4))).astype(np.float32)
    ori = (np.pi * np.tanh(rng.standard_normal((1, h4, w4)))).astype(np.float32)
    kp = torch.from_numpy(np.concatenate([logit, off, ori], axis=0))
    desc = F.normalize(torch.from_numpy(rng.standard_normal((d, h4, w4)).astype(np.float32)), p=2, dim=0, eps=1e-8)
    return kp, desc


def trainable_select(kp_map: torch.Tensor, desc_map: torch.Tensor, orig_wh, resized_wh, num_keypoints: int,
                     score_threshold: float, nms_radius: int):
    """trainable_vit_extractor.py:173-269 on prepared head outputs.  Ties of torch.topk are resolved
    towards the lower row-major cell index.  Returns (keypoints [K,6] f32, descriptors [K,d] u8,
    coords [K,2] (y,x))."""
    w_orig, h_orig = orig_wh
    w_new, h_new = resized_wh
    scores = torch.sigmoid(kp_map[0])
    valid = (scores > score_threshold) & simple_nms(scores, nms_radius)
    coords = torch.nonzero(valid, as_tuple=False)
    d = desc_map.shape[0]
    if len(coords) == 0:
        return np.zeros((0, 6), np.float32), np.zeros((0, d), np.uint8), np.zeros((0, 2), np.int64)
    vs = scores[coords[:, 0], coords[:, 1]]
    k = min(num_keypoints, len(vs))
    _, order = _topk_stable(vs.numpy(), k)
    order = torch.from_numpy(order.copy())
    sel, sel_scores = coords[order], vs[order]
    sy, sx = sel[:, 0], sel[:, 1]
    kx = ((sx.float() + kp_map[1][sy, sx] + 0.5) * 4.0 * (w_orig / w_new)).clamp(0, w_orig - 1)
    ky = ((sy.float() + kp_map[2][sy, sx] + 0.5) * 4.0 * (h_orig / h_new)).clamp(0, h_orig - 1)
    kps = torch.stack([kx, ky, torch.ones_like(kx), kp_map[3][sy, sx], sel_scores, torch.zeros_like(kx)], dim=1)
    desc = desc_map[:, sy, sx].t().numpy()
    return kps.numpy().astype(np.float32), ((desc + 1.0) * 127.5).clip(0, 255).astype(np.uint8), sel.numpy()
