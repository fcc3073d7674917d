"""Pins oracle/extract_oracle.py against golden vectors produced by the reference's own
ViTExtractor post-processing (oracle/gen_golden_extract.py, run in the build container)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import extract_oracle as eo

GOLDEN = sorted((Path(__file__).parent / "golden").glob("extract_*.npz"))


def canonical_order(scores, coords):
    """score descending, ties by (y, x) ascending."""
    return np.lexsort((coords[:, 1], coords[:, 0], -scores.astype(np.float64)))


def _load(path):
    g = np.load(path)
    seed, C, H, W, K, d, ow, oh, rw, rh = (int(x) for x in g["meta"])
    return g, dict(seed=seed, C=C, H=H, W=W, K=K, d=d, orig=(ow, oh), resized=(rw, rh), method=str(g["method"]))


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_restatement_matches_reference_outputs(path):
    g, m = _load(path)
    fm = eo.synthetic_feature_map(m["seed"], m["C"], m["H"], m["W"])
    proj = torch.from_numpy(g["projection"]) if g["projection"].size else None
    out = eo.dense_to_sparse(fm, m["orig"], m["resized"], m["K"], m["d"], m["method"], projection=proj)
    assert np.abs(out["score"] - g["score"]).max() < 1e-6
    # keypoint indices bit-exact; the only freedom is the order among EXACTLY equal scores
    # (torch.argsort is unstable in the reference, vit_extractor.py:519)
    assert np.array_equal(out["scores"], g["nms_scores"])
    po, pg = canonical_order(out["scores"], out["coords"]), canonical_order(g["nms_scores"], g["nms_coords"])
    assert out["coords"][po].tolist() == g["nms_coords"][pg].tolist()
    assert np.array_equal(out["keypoints"][po], g["keypoints"][pg])
    diff = np.abs(out["desc_u8"][po].astype(int) - g["desc_u8"][pg].astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-3               # fp32 rounding at the trunc edge


def test_pca_fit_matches_reference_projection():
    g, m = _load([p for p in GOLDEN if "pca" in p.stem][0])
    fm = eo.synthetic_feature_map(m["seed"], m["C"], m["H"], m["W"])
    out = eo.dense_to_sparse(fm, m["orig"], m["resized"], m["K"], m["d"], m["method"])
    ref = g["projection"]
    got = out["projection"].numpy()
    # SVD sign is implementation-defined: compare up to per-column sign
    sign = np.sign((ref * got).sum(axis=0))
    assert np.abs(ref - got * sign).max() < 1e-4


def test_binning_ignores_cells_beyond_last_full_bin():
    # 37x37 grids only visit the 32x32 region
    fm = eo.synthetic_feature_map(1, 16, 37, 37)
    score = eo.distinctiveness(fm).numpy()
    coords, _ = eo.spatial_binning_selection(score, 8192, 16)
    assert coords.max() < 32 and len(coords) == 1024


TRAINABLE = sorted((Path(__file__).parent / "golden").glob("trainable_*.npz"))


def load_trainable(path):
    g = np.load(path)
    seed, d, h4, w4, K, r, ow, oh, rw, rh = (int(x) for x in g["meta"])
    return g, dict(seed=seed, d=d, h4=h4, w4=w4, K=K, r=r, thr=float(g["thr"]), orig=(ow, oh), resized=(rw, rh))


@pytest.mark.parametrize("path", TRAINABLE, ids=[p.stem for p in TRAINABLE])
def test_trainable_restatement_matches_reference_outputs(path):
    """oracle.trainable_select vs the reference's own TrainableViTExtractor._run_inference
    (stubbed network, oracle/gen_golden_trainable.py)."""
    g, m = load_trainable(path)
    kp, desc = eo.synthetic_head_maps(m["seed"], m["d"], m["h4"], m["w4"])
    kps, du8, _ = eo.trainable_select(kp, desc, m["orig"], m["resized"], m["K"], m["thr"], m["r"])
    assert kps.shape == g["keypoints"].shape and du8.shape == g["desc_u8"].shape
    if len(kps) == 0:
        return
    assert np.array_equal(kps[:, 4], g["keypoints"][:, 4])        # scores, descending
    # order among exactly equal scores is unspecified in torch.topk: compare as sorted row sets
    key = lambda a: np.lexsort((a[:, 1], a[:, 0], -a[:, 4].astype(np.float64)))  # noqa: E731
    po, pg = key(kps), key(g["keypoints"])
    assert np.array_equal(kps[po], g["keypoints"][pg])
    assert np.array_equal(du8[po], g["desc_u8"][pg])


HYBRID = sorted((Path(__file__).parent / "golden").glob("hybrid_*.npz"))


@pytest.mark.parametrize("path", HYBRID, ids=[p.stem for p in HYBRID])
def test_hybrid_restatement_matches_reference_outputs(path):
    """oracle hybrid_descriptors vs the reference's own _extract_descriptors_at_keypoints
    (hybrid_extractor.py:224-294; oracle/gen_golden_hybrid.py)."""
    g = np.load(path)
    seed, C, H, W, n, d, ow, oh, fw, fh = (int(x) for x in g["meta"])
    fm = eo.synthetic_feature_map(seed, C, H, W)
    proj = torch.from_numpy(g["projection"]) if g["projection"].size else None
    out = eo.hybrid_descriptors(fm, g["keypoints"], (ow, oh), (fw, fh), d, projection=proj)
    assert out["desc_u8"].shape == g["desc_u8"].shape == (n, d)
    assert np.array_equal(out["desc_u8"], g["desc_u8"])


def test_hybrid_empty_keypoints():
    out = eo.hybrid_descriptors(eo.synthetic_feature_map(0, 8, 5, 5), np.zeros((0, 2), np.float32), (70, 70), (70, 70), 8)
    assert out["desc_u8"].shape == (0, 8)
