"""Oracle of the training-time samplers (oracle/sampler_oracle.py) against golden vectors produced by the
REFERENCE's own TrainingSampler / compute_feature_similarity (oracle/gen_golden_sampler.py).  CPU only."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import sampler_oracle as so

GOLDEN = sorted((Path(__file__).parent / "golden").glob("sampler_*.npz"))
TIE = 1e-5  # selections may differ only among similarities closer than this (torch.topk tie order is unspecified)


def load(path):
    g = np.load(path)
    seed, B, C, Hp, Wp, top_k, mn, mx, radius, n_in, n_hard = (int(x) for x in g["meta"])
    m = dict(seed=seed, B=B, C=C, Hp=Hp, Wp=Wp, top_k=top_k, min_pts=mn, max_pts=mx, radius=radius, n_in=n_in,
             n_hard=n_hard, threshold=float(g["threshold"]))
    return g, m, so.synthetic_pair(seed, B, C, Hp, Wp)


def same_selection(got_xy, ref_xy, sim, Wp, tol=TIE):
    """Two (K, 2) coordinate lists pick the same cells up to near-ties: the similarities they pick agree
    position by position (both lists are ordered the same way)."""
    gi = (got_xy[:, 1] * Wp + got_xy[:, 0]).astype(np.int64)
    ri = (ref_xy[:, 1] * Wp + ref_xy[:, 0]).astype(np.int64)
    assert gi.shape == ri.shape
    flat = sim.reshape(-1)
    return np.abs(flat[gi] - flat[ri]).max() <= tol


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_similarity_map_matches_reference(path):
    g, m, (f1, f2, H, image_size) = load(path)
    sim = so.similarity_map(f1, f2, H, 14, image_size).numpy()
    assert np.abs(sim - g["similarity"]).max() < 1e-6
    assert g["similarity"].max() > 0.8 and g["similarity"].min() < 0.5   # real structure, not a constant map


@pytest.mark.parametrize("mode", ["top_k", "threshold", "hybrid"])
@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_selection_matches_reference(path, mode):
    g, m, (f1, f2, H, image_size) = load(path)
    sim = torch.from_numpy(g["similarity"]).reshape(m["B"], -1)
    idx = so.select_indices(sim, mode, m["top_k"], m["threshold"], m["min_pts"], m["max_pts"])
    coords = so.pad_and_coords(idx, m["Wp"]).numpy()
    ref = g[f"coords_{mode}"]
    assert coords.shape == ref.shape
    for b in range(m["B"]):
        assert same_selection(coords[b], ref[b], g["similarity"][b], m["Wp"])


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_positive_pairs_and_hard_negatives_match_reference(path):
    g, m, (f1, f2, H, image_size) = load(path)
    coords = torch.from_numpy(g["pos_coords"])
    p1, p2 = so.positive_pairs(f1, f2, H, coords, 14)
    assert np.abs(p1.numpy() - g["pos_f1"]).max() < 1e-5 and np.abs(p2.numpy() - g["pos_f2"]).max() < 1e-6
    hard = so.hard_negatives(f2, p2, coords, m["n_hard"], float(m["radius"])).numpy()
    ref = g["hard"]
    assert hard.shape == ref.shape
    bad = np.abs(hard - ref).max(axis=-1) > 1e-5
    if bad.any():   # a different cell was picked: only allowed between near-equal similarities
        idx, counts, masked = so.hard_negative_indices(f2, p2, coords, m["n_hard"], float(m["radius"]))
        for b, k, j in zip(*np.nonzero(bad)):
            v = masked[b, k]
            kth = torch.sort(v, descending=True).values[j]
            assert abs(float(v[idx[b, k, j]]) - float(kth)) <= TIE
    if path.stem.endswith("wide_radius"):   # fewer cells than n_hard qualify somewhere: zero-padded rows
        assert (np.abs(ref).max(axis=-1) == 0).any()


def test_host_side_of_the_sampler_mirror():
    """The product's TrainingSampler keeps O(B*K) bookkeeping on the host: the coordinate transform must equal the
    reference arithmetic (oracle), and without a CUDA device every device method must fail loudly (no CPU path)."""
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.dataloader import TrainingSampler
    s = TrainingSampler(patch_size=14)
    rng = np.random.default_rng(3)
    coords = torch.from_numpy(rng.uniform(0, 36, (2, 50, 2)).astype(np.float32))
    _, _, H, _ = so.synthetic_pair(5, 2, 4, 37, 37)
    for feature_space in (True, False):
        got = s.transform_coords_with_homography(coords, H, from_feature_space=feature_space,
                                                 feature_size=(37, 37) if feature_space else None)
        ref = so.transform_coords(coords, H, 14, feature_space)
        assert got.dtype == torch.float32 and np.abs(got.numpy() - ref.numpy()).max() < 1e-4
    if not torch.cuda.is_available():
        with pytest.raises(_lib.VcError):
            s.sample_features_at_coords(torch.zeros(1, 8, 4, 4), torch.zeros(1, 3, 2))
        with pytest.raises(_lib.VcError):
            s.select_invariant_points(torch.zeros(1, 8, 4, 4), torch.zeros(1, 8, 4, 4), torch.eye(3)[None], (56, 56))
