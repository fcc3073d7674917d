"""GPU parity tests of the training-time samplers through the C ABI: against golden
vectors from the REFERENCE's own TrainingSampler / compute_feature_similarity and against the CPU oracle."""
import ctypes
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = sorted((Path(__file__).parent / "golden").glob("sampler_*.npz"))
SIM_TOL = 2e-6    # fp32 summation order; cosines in [-1, 1]
FEAT_TOL = 1e-5   # bilinear samples of O(1) features
TIE = 1e-5        # selections may differ only among similarities closer than this


@pytest.fixture(scope="module")
def dev():
    from vit_colmap_b200.dataloader._device import SamplerDevice
    d = SamplerDevice(0)
    yield d
    d.close()


def load(path):
    from oracle import sampler_oracle as so
    g = np.load(path)
    seed, B, C, Hp, Wp, top_k, mn, mx, radius, n_in, n_hard = (int(x) for x in g["meta"])
    m = dict(seed=seed, B=B, C=C, Hp=Hp, Wp=Wp, top_k=top_k, min_pts=mn, max_pts=mx, radius=radius, n_in=n_in,
             n_hard=n_hard, threshold=float(g["threshold"]))
    return g, m, so.synthetic_pair(seed, B, C, Hp, Wp)


def sampler(m, mode="top_k"):
    from vit_colmap_b200.dataloader import TrainingSampler
    return TrainingSampler(top_k_invariant=m["top_k"], negative_radius=m["radius"], num_in_image_negatives=m["n_in"],
                           num_hard_negatives=m["n_hard"], selection_mode=mode, similarity_threshold=m["threshold"],
                           min_invariant_points=m["min_pts"], max_invariant_points=m["max_pts"])


def same_selection(got_xy, ref_xy, sim, Wp, tol=TIE):
    gi = (got_xy[:, 1] * Wp + got_xy[:, 0]).astype(np.int64)
    ri = (ref_xy[:, 1] * Wp + ref_xy[:, 0]).astype(np.int64)
    flat = sim.reshape(-1)
    return gi.shape == ri.shape and np.abs(flat[gi] - flat[ri]).max() <= tol


# ---------------------------------------------------------------------------------------------- similarity
@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_similarity_map_matches_reference(path):
    from vit_colmap_b200.dataloader import compute_feature_similarity
    g, m, (f1, f2, H, image_size) = load(path)
    sim = compute_feature_similarity(f1, f2, H, patch_size=14, input_size=image_size).cpu().numpy()
    assert np.abs(sim - g["similarity"]).max() < SIM_TOL


@pytest.mark.parametrize("shape", [(2, 768, 85, 114), (3, 384, 37, 37), (1, 1024, 20, 9), (2, 100, 5, 3)])
def test_similarity_map_matches_oracle(dev, shape):
    import torch
    from oracle import sampler_oracle as so
    from vit_colmap_b200.dataloader.homography_utils import inverse_homography
    B, C, Hp, Wp = shape
    f1, f2, H, image_size = so.synthetic_pair(100 + C, B, C, Hp, Wp)
    H = H.clone()
    H[0, 0, 2] += 0.6 * image_size[1]            # a large shift: most of image 0 maps outside (zero vectors)
    if B > 1:
        H[1, 2, 0], H[1, 2, 1] = -2e-3, -1e-3    # strong perspective: the homogeneous w goes through 0 (clamp)
    ref = so.similarity_map(f1, f2, H, 14, image_size).numpy()
    got = dev.warp_similarity(dev.token_map(f1), dev.token_map(f2), inverse_homography(H), 14, image_size)
    got = got.cpu().numpy()
    assert (ref[0] == 0).mean() > 0.3
    # under the strong perspective the source positions reach thousands of pixels and one ulp of them (the 3x3
    # product is summed in another order than torch.bmm's) moves a cosine by a few 1e-6
    assert np.abs(got - ref).max() < 1e-5 and np.abs(got[0] - ref[0]).max() < SIM_TOL


def test_nchw_to_tokens_is_exact(dev):
    import torch
    x = torch.randn(3, 100, 7, 13)
    tm = dev.token_map(x)
    assert tm.grid_hw == (7, 13)
    assert torch.equal(tm.tokens.cpu(), x.permute(0, 2, 3, 1).reshape(3, 91, 100))


# ---------------------------------------------------------------------------------------------- top-k
@pytest.mark.parametrize("case", [(4, 1369, 512, 0), (2, 9690, 2048, 0), (3, 16384, 16384, 0), (5, 100, 7, 8),
                                  (2, 1, 1, 0), (2, 4000, 300, 16)])
def test_topk_rows_matches_oracle(dev, case):
    import torch
    R, n, k, levels = case
    rng = np.random.default_rng(n + k)
    v = rng.standard_normal((R, n)).astype(np.float32)
    if levels:
        v = np.round(v * levels) / levels        # ties everywhere
    v[0, 0] = -np.inf
    thr = 0.25
    order, above, n_above = dev.topk_rows(torch.from_numpy(v), k, thr)
    order, above, n_above = order.cpu().numpy(), above.cpu().numpy(), n_above.cpu().numpy()
    for r in range(R):
        ref = np.lexsort((np.arange(n), -v[r].astype(np.float64)))[:k]     # value desc, index asc
        assert order[r].tolist() == ref.tolist()
        hits = np.nonzero(v[r] >= thr)[0]
        assert n_above[r] == len(hits) and above[r, : len(hits)].tolist() == hits.tolist()


@pytest.mark.parametrize("mode", ["top_k", "threshold", "hybrid"])
@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_select_invariant_points_matches_reference(path, mode):
    g, m, (f1, f2, H, image_size) = load(path)
    coords = sampler(m, mode).select_invariant_points(f1, f2, H, image_size).cpu().numpy()
    ref = g[f"coords_{mode}"]
    assert coords.shape == ref.shape and coords.dtype == np.float32
    for b in range(m["B"]):
        assert same_selection(coords[b], ref[b], g["similarity"][b], m["Wp"])


# ---------------------------------------------------------------------------------------------- sampling / pairs
@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_positive_pairs_match_reference(path):
    import torch
    g, m, (f1, f2, H, image_size) = load(path)
    coords = torch.from_numpy(g["pos_coords"])
    p1, p2 = sampler(m).generate_positive_pairs(f1, f2, H, coords, image_size)
    assert np.abs(p1.cpu().numpy() - g["pos_f1"]).max() < FEAT_TOL
    assert np.abs(p2.cpu().numpy() - g["pos_f2"]).max() < FEAT_TOL


def test_sample_bilinear_edges(dev):
    """Positions on and beyond the border (zero padding), and NaN = padding entry."""
    import torch
    from oracle import sampler_oracle as so
    f = torch.randn(2, 64, 9, 14)
    xy = torch.tensor([[0.0, 0.0], [13.0, 8.0], [-0.4, 3.3], [13.6, 8.2], [-3.0, 2.0], [5.5, 40.0], [6.25, 4.75],
                       [12.99, 0.01]]).repeat(2, 1, 1)
    ref = so.sample_at(f, xy).numpy()
    got = dev.sample_bilinear(dev.token_map(f), xy).cpu().numpy()
    assert np.abs(got - ref).max() < FEAT_TOL
    assert np.abs(ref[:, 4]).max() == 0 and np.abs(ref[:, 1]).max() > 0
    xy[0, 2] = float("nan")
    got = dev.sample_bilinear(dev.token_map(f), xy).cpu().numpy()
    assert np.abs(got[0, 2]).max() == 0 and np.abs(got[1, 2] - ref[1, 2]).max() < FEAT_TOL


# ---------------------------------------------------------------------------------------------- negatives
@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_hard_negatives_match_reference(path):
    import torch
    from oracle import sampler_oracle as so
    g, m, (f1, f2, H, image_size) = load(path)
    coords = torch.from_numpy(g["pos_coords"])
    p2 = torch.from_numpy(g["pos_f2"])
    hard = sampler(m).generate_hard_negatives(f2, p2, coords).cpu().numpy()
    ref = g["hard"]
    assert hard.shape == ref.shape
    bad = np.abs(hard - ref).max(axis=-1) > FEAT_TOL
    if bad.any():     # a different cell: only between near-equal similarities
        idx, counts, masked = so.hard_negative_indices(f2, p2, coords, m["n_hard"], float(m["radius"]))
        flat = f2.reshape(m["B"], m["C"], -1)
        for b, k, j in zip(*np.nonzero(bad)):
            kth = float(torch.sort(masked[b, k], descending=True).values[j])
            cell = int(np.argmin(np.abs(flat[b].numpy().T - hard[b, k, j]).max(axis=1)))
            assert abs(float(masked[b, k, cell]) - kth) <= TIE


@pytest.mark.parametrize("case", [(2, 384, 37, 37, 512, 5, 16.0), (1, 768, 42, 57, 100, 32, 10.0),
                                  (2, 64, 12, 9, 33, 8, 14.5), (1, 32, 6, 6, 5, 4, 100.0)])
def test_hard_negative_indices_match_oracle(dev, case):
    import torch
    from oracle import sampler_oracle as so
    B, C, Hp, Wp, K, n_hard, radius = case
    rng = np.random.default_rng(K + C)
    f = torch.from_numpy(rng.standard_normal((B, C, Hp, Wp)).astype(np.float32))
    f[:, :, 0, 1] = f[:, :, Hp - 1, Wp - 1]      # two identical cells: an exact tie, lower index first
    cells = rng.integers(0, Hp * Wp, (B, K))
    pos_xy = torch.from_numpy(np.stack([cells % Wp, cells // Wp], -1).astype(np.float32))
    pos_xy[:, 0] += 0.37                         # non-integer coordinates too
    pos = so.sample_at(f, pos_xy) + 0.05 * torch.from_numpy(rng.standard_normal((B, K, C)).astype(np.float32))
    ref_idx, ref_cnt, masked = so.hard_negative_indices(f, pos, pos_xy, n_hard, radius)
    idx, cnt = dev.hard_negatives(dev.token_map(f), pos, pos_xy, n_hard, radius)
    idx, cnt = idx.cpu().numpy().astype(np.int64), cnt.cpu().numpy()
    assert np.array_equal(cnt, ref_cnt.numpy())
    ref_idx = ref_idx.numpy()
    diff = idx != ref_idx
    assert not (diff & ((idx < 0) | (ref_idx < 0))).any()      # padding in the same places
    for b, k, j in zip(*np.nonzero(diff)):
        assert abs(float(masked[b, k, idx[b, k, j]]) - float(masked[b, k, ref_idx[b, k, j]])) <= TIE
    assert diff.mean() < 0.01
    if radius > Hp + Wp:
        assert (idx == -1).all() and (cnt == 0).all()


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_random_negatives_draw_like_the_reference(path):
    """In-image / cross-image negatives use the reference's own torch / NumPy draws in the same order, so a
    seeded run reproduces the reference's CPU outputs."""
    import torch
    g, m, (f1, f2, H, image_size) = load(path)
    coords = torch.from_numpy(g["pos_coords"])
    s = sampler(m)
    torch.manual_seed(m["seed"])
    np.random.seed(m["seed"])
    a = s.generate_in_image_negatives(f2, coords).cpu().numpy()
    c = s.generate_cross_image_negatives(f1, coords).cpu().numpy()
    assert a.shape == g["in_image"].shape and np.abs(a - g["in_image"]).max() < FEAT_TOL
    assert c.shape == g["cross_image"].shape and np.abs(c - g["cross_image"]).max() < FEAT_TOL
    torch.manual_seed(m["seed"] + 1)
    np.random.seed(m["seed"] + 1)
    allneg = s.generate_all_negatives(f1, f2, torch.from_numpy(g["pos_f2"]), coords).cpu().numpy()
    ref = g["all_negatives"]
    assert allneg.shape == ref.shape
    n_rand = a.shape[2] + c.shape[2]
    assert np.abs(allneg[:, :, :n_rand] - ref[:, :, :n_rand]).max() < FEAT_TOL
    assert (np.abs(allneg[:, :, n_rand:] - ref[:, :, n_rand:]).max(axis=-1) < FEAT_TOL).mean() > 0.98


# ---------------------------------------------------------------------------------------------- errors
def test_sampler_argument_errors(dev):
    import torch
    from vit_colmap_b200 import _lib
    with pytest.raises(_lib.VcError, match="multiple of 4"):
        dev.token_map(torch.zeros(1, 4, 2, 2)) and dev.sample_bilinear(
            dev.token_map(torch.zeros(1, 6, 2, 2)), torch.zeros(1, 1, 2))
    with pytest.raises(_lib.VcError, match="exceeds"):
        dev.topk_rows(torch.zeros(1, 16385), 4)
    tm = dev.token_map(torch.zeros(1, 8, 4, 4))
    with pytest.raises(_lib.VcError, match="n_hard"):
        dev.hard_negatives(tm, torch.zeros(1, 2, 8), torch.zeros(1, 2, 2), 33, 1.0)
