"""GPU parity tests: extraction post-processing kernels (through the C ABI) against the CPU oracle
and the golden vectors produced by the reference's own ViTExtractor code."""
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = sorted((Path(__file__).parent / "golden").glob("extract_*.npz"))
SCORE_TOL = 2e-6  # fp32 summation order; scores are min-max normalised to [0, 1]


@pytest.fixture(scope="module")
def pp():
    from vit_colmap_b200.extraction import PostProcessor
    p = PostProcessor(0)
    yield p
    p.close()


def _tokens(fm, dtype=None):
    """[C,H,W] torch CPU -> token-major [1, H*W, C] on the GPU."""
    import torch
    C, H, W = fm.shape
    t = fm.permute(1, 2, 0).reshape(1, H * W, C).contiguous().cuda()
    return t.to(dtype or torch.float32)


def _load(path):
    g = np.load(path)
    seed, C, H, W, K, d, ow, oh, rw, rh = (int(x) for x in g["meta"])
    return g, dict(seed=seed, C=C, H=H, W=W, K=K, d=d, orig=(ow, oh), resized=(rw, rh), method=str(g["method"]))


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_score_map_matches_reference(pp, path):
    from oracle import extract_oracle as eo
    g, m = _load(path)
    fm = eo.synthetic_feature_map(m["seed"], m["C"], m["H"], m["W"])
    s = pp.score_map(_tokens(fm), (m["H"], m["W"]), m["method"]).cpu().numpy()[0]
    assert np.abs(s - g["score"]).max() < SCORE_TOL


@pytest.mark.parametrize("method", ["harris", "dog", "combined"])
@pytest.mark.parametrize("shape", [(384, 37, 37), (768, 42, 57), (1024, 20, 9), (100, 5, 3)])
def test_score_map_matches_oracle(pp, method, shape):
    from oracle import extract_oracle as eo
    C, H, W = shape
    fm = eo.synthetic_feature_map(11, C, H, W)
    s = pp.score_map(_tokens(fm), (H, W), method).cpu().numpy()[0]
    ref = eo.distinctiveness(fm, method).numpy()
    assert np.abs(s - ref).max() < SCORE_TOL


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_selection_is_bit_exact_on_reference_scores(pp, path):
    """Same score map in -> identical keypoints, scores and order as the reference's binning + NMS."""
    import torch
    g, m = _load(path)
    score = torch.from_numpy(g["score"]).cuda()[None]
    coords, scores, counts = pp.select_keypoints(score, m["K"])
    n = int(counts[0])
    assert n == len(g["nms_coords"])
    assert np.array_equal(scores[0, :n].cpu().numpy(), g["nms_scores"])
    got = coords[0, :n].cpu().numpy().astype(np.int64)
    # order among exactly equal scores is unspecified in the reference (unstable argsort)
    key = lambda s, c: np.lexsort((c[:, 1], c[:, 0], -s.astype(np.float64)))  # noqa: E731
    assert got[key(g["nms_scores"], got)].tolist() == g["nms_coords"][key(g["nms_scores"], g["nms_coords"])].tolist()


@pytest.mark.parametrize("case", [(37, 37, 2048, 0), (85, 114, 8192, 0), (85, 114, 8192, 16), (64, 64, 10, 0),
                                  (9, 20, 64, 4), (128, 128, 8192, 0), (33, 50, 3, 0), (1, 1, 5, 0)])
def test_selection_matches_oracle_with_ties(pp, case):
    """Random maps, optionally quantised to a few levels so that ties are everywhere."""
    import torch
    from oracle import extract_oracle as eo
    H, W, K, levels = case
    rng = np.random.default_rng(H * 1000 + W + K)
    score = rng.random((3, H, W)).astype(np.float32)
    if levels:
        score = np.floor(score * levels).astype(np.float32) / levels
    coords, scores, counts = pp.select_keypoints(torch.from_numpy(score).cuda(), K)
    for b in range(3):
        c, s = eo.spatial_binning_selection(score[b], K, 16)
        c, s = eo.apply_nms(c, s, 1.5)
        n = int(counts[b])
        assert n == len(c)
        assert coords[b, :n].cpu().numpy().tolist() == c.tolist()   # including the order
        assert np.array_equal(scores[b, :n].cpu().numpy(), s)


@pytest.mark.parametrize("path", GOLDEN, ids=[p.stem for p in GOLDEN])
def test_sample_pack_matches_reference(pp, path):
    import torch
    from oracle import extract_oracle as eo
    g, m = _load(path)
    fm = eo.synthetic_feature_map(m["seed"], m["C"], m["H"], m["W"])
    n = len(g["nms_coords"])
    coords = torch.from_numpy(g["nms_coords"].astype(np.int32)).cuda()[None].contiguous()
    counts = torch.tensor([n], dtype=torch.int32, device="cuda")
    proj = torch.from_numpy(g["projection"]) if g["projection"].size else None
    kps, dbf, du8, df32 = pp.sample_pack(_tokens(fm), (m["H"], m["W"]), coords, counts, m["orig"], m["resized"],
                                         projection=proj, want_f32=True)
    assert np.array_equal(kps[0, :n].cpu().numpy(), g["keypoints"])
    diff = np.abs(du8[0, :n].cpu().numpy().astype(int) - g["desc_u8"].astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-3          # fp32 rounding at the trunc edge
    import torch.nn.functional as F
    ref = eo.sample_descriptors(fm, g["nms_coords"])
    if proj is not None:
        ref = ref @ proj
    ref = F.normalize(ref, p=2, dim=1).numpy()
    got = df32[0, :n].cpu().numpy()
    assert np.abs(got - ref).max() < 1e-5
    bf = dbf[0, :n].float().cpu().numpy()
    assert np.abs(bf - got).max() <= 2 ** -8                       # bf16 rounding of unit rows


def test_bilinear_sampling_at_float_coords(pp):
    """General bilinear path (non-integer grid coordinates, border clamp) vs F.grid_sample."""
    import torch
    import torch.nn.functional as F
    from oracle import extract_oracle as eo
    C, H, W = 96, 21, 30
    fm = eo.synthetic_feature_map(3, C, H, W)
    rng = np.random.default_rng(0)
    n = 500
    yx = np.stack([rng.uniform(-0.5, H - 0.5, n), rng.uniform(-0.5, W - 0.5, n)], 1).astype(np.float32)
    cn = torch.from_numpy(yx.copy())
    cn[:, 0] = 2.0 * cn[:, 0] / (H - 1) - 1.0
    cn[:, 1] = 2.0 * cn[:, 1] / (W - 1) - 1.0
    ref = F.grid_sample(fm[None], cn[:, [1, 0]][None, :, None], mode="bilinear", padding_mode="border",
                        align_corners=True)[0, :, :, 0].t()
    ref = F.normalize(ref, dim=1).numpy()
    coords = torch.from_numpy(yx).cuda()[None].contiguous()
    counts = torch.tensor([n], dtype=torch.int32, device="cuda")
    _, _, _, df32 = pp.sample_pack(_tokens(fm), (H, W), coords, counts, (W * 14, H * 14), (W * 14, H * 14),
                                   want_f32=True)
    assert np.abs(df32[0].cpu().numpy() - ref).max() < 1e-5


def test_batched_bf16_tokens_whole_stage(pp):
    """Whole stage on a batch of bf16 token maps == oracle on the bf16-rounded maps, image by image.
    Keypoint sets must agree except where the deciding score gap is a near-tie (< 1e-5)."""
    import torch
    from oracle import extract_oracle as eo
    C, H, W, K = 384, 37, 37, 2048
    fms = [eo.synthetic_feature_map(20 + i, C, H, W).to(torch.bfloat16) for i in range(5)]
    tok = torch.stack([f.permute(1, 2, 0).reshape(H * W, C) for f in fms]).cuda()
    out = pp.dense_to_sparse(tok, (H, W), (518, 518), (518, 518), K, "harris", want_f32=True)
    for i, f in enumerate(fms):
        ref = eo.dense_to_sparse(f.float(), (518, 518), (518, 518), K, C, "harris")
        n = int(out.counts[i])
        got = {tuple(c) for c in out.coords[i, :n].cpu().numpy().tolist()}
        exp = {tuple(c) for c in ref["coords"].tolist()}
        if got != exp:
            sc = ref["score"]
            odd = got ^ exp
            # every disagreement must sit next to a near-tie of the oracle's score map
            for (y, x) in odd:
                nb = sc[max(0, y - 1):y + 2, max(0, x - 1):x + 2]
                gaps = np.abs(nb - sc[y, x])
                assert np.sort(gaps.ravel())[1] < 1e-5, (i, y, x)
        else:
            pos = {tuple(c): k for k, c in enumerate(ref["coords"].tolist())}
            idx = [pos[tuple(c)] for c in out.coords[i, :n].cpu().numpy().tolist()]
            assert np.abs(out.desc_f32[i, :n].cpu().numpy() - ref["desc_float"][idx]).max() < 1e-5
            assert np.array_equal(out.keypoints[i, :n].cpu().numpy(), ref["keypoints"][idx])


TRAINABLE = sorted((Path(__file__).parent / "golden").glob("trainable_*.npz"))


@pytest.mark.parametrize("path", TRAINABLE, ids=[p.stem for p in TRAINABLE])
def test_trainable_select_pack_matches_reference(pp, path):
    """vc_trainable_select + vc_sample_pack(quant TRAINABLE) vs the reference's own
    TrainableViTExtractor._run_inference outputs (stubbed network, tests/golden/trainable_*.npz).
    Keypoint sets are identical; rows may only swap among exactly equal scores; the sigmoid is
    allowed 1 ulp (CUDA expf vs ATen's vectorised exp)."""
    import torch
    from oracle import extract_oracle as eo
    from vit_colmap_b200.extraction import trainable_select_pack
    g = np.load(path)
    seed, d, h4, w4, K, r, ow, oh, rw, rh = (int(x) for x in g["meta"])
    kp, desc = eo.synthetic_head_maps(seed, d, h4, w4)
    kp_cl = kp.permute(1, 2, 0)[None].contiguous().cuda()
    desc_cl = desc.permute(1, 2, 0)[None].contiguous().cuda()
    kps, du8, dbf, df32, counts = trainable_select_pack(pp, kp_cl, desc_cl, (ow, oh), (rw, rh), K, float(g["thr"]), r,
                                                        apply_tanh=False, want_f32=True)
    n = int(counts[0])
    ref_k, ref_d = g["keypoints"], g["desc_u8"]
    assert n == len(ref_k)
    if n == 0:
        return
    got_k, got_d = kps[0, :n].cpu().numpy(), du8[0, :n].cpu().numpy()
    assert np.all(np.diff(got_k[:, 4]) <= 0)                                   # descending scores
    key = lambda a: np.lexsort((a[:, 1], a[:, 0]))  # noqa: E731  by position: sets must coincide
    po, pg = key(got_k), key(ref_k)
    assert np.array_equal(got_k[po][:, [0, 1, 2, 3, 5]], ref_k[pg][:, [0, 1, 2, 3, 5]])
    assert np.abs(got_k[po][:, 4] - ref_k[pg][:, 4]).max() <= 1.2e-7
    diff = np.abs(got_d[po].astype(int) - ref_d[pg].astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 1e-3      # re-normalising unit rows moves <= 1 ulp
    assert np.abs(dbf[0, :n].float().cpu().numpy() - df32[0, :n].cpu().numpy()).max() <= 2 ** -8


def test_trainable_select_large_map_and_tanh(pp):
    """A 1600x1200-sized score map (399x297 cells, beyond any shared-memory path), batch of 2,
    orientation produced from raw head output (apply_tanh)."""
    import torch
    from oracle import extract_oracle as eo
    from vit_colmap_b200.extraction import trainable_select_pack
    h4, w4, d = 297, 399, 32
    outs = []
    maps = [eo.synthetic_head_maps(30 + i, d, h4, w4) for i in range(2)]
    kp_cl = torch.stack([m[0].permute(1, 2, 0) for m in maps]).contiguous()
    raw = kp_cl.clone()
    raw[..., 3] = torch.atanh((kp_cl[..., 3] / np.pi).clamp(-0.999, 0.999))
    desc_cl = torch.stack([m[1].permute(1, 2, 0) for m in maps]).contiguous()
    kps, du8, _, _, counts = trainable_select_pack(pp, raw.cuda(), desc_cl.cuda(), (1600, 1200), (1596, 1190), 8192,
                                                   0.6, 2, apply_tanh=True)
    for i, (kp, desc) in enumerate(maps):
        kp = kp.clone()
        kp[3] = torch.tanh(raw[i, ..., 3]) * torch.pi
        rk, rd, _ = eo.trainable_select(kp, desc, (1600, 1200), (1596, 1190), 8192, 0.6, 2)
        n = int(counts[i])
        assert n == len(rk) and n > 1000
        gk = kps[i, :n].cpu().numpy()
        po, pg = np.lexsort((gk[:, 1], gk[:, 0])), np.lexsort((rk[:, 1], rk[:, 0]))
        assert np.array_equal(gk[po][:, :3], rk[pg][:, :3])
        assert np.abs(gk[po][:, 3] - rk[pg][:, 3]).max() < 1e-6 and np.abs(gk[po][:, 4] - rk[pg][:, 4]).max() <= 1.2e-7
        assert np.abs(du8[i, :n].cpu().numpy()[po].astype(int) - rd[pg].astype(int)).max() <= 1


HYBRID = sorted((Path(__file__).parent / "golden").glob("hybrid_*.npz"))


@pytest.mark.parametrize("path", HYBRID, ids=[p.stem for p in HYBRID])
def test_hybrid_rootsift_descriptors_match_reference(pp, path):
    """vc_sample_pack (float coordinates, RootSIFT flag) vs the reference's own
    _extract_descriptors_at_keypoints outputs (hybrid_extractor.py:224-294) and the oracle."""
    import torch
    from oracle import extract_oracle as eo
    from vit_colmap_b200.features.hybrid_extractor import HybridViTExtractor
    g = np.load(path)
    seed, C, H, W, n, d, ow, oh, fw, fh = (int(x) for x in g["meta"])
    fm = eo.synthetic_feature_map(seed, C, H, W)
    proj = torch.from_numpy(g["projection"]) if g["projection"].size else None
    yx = HybridViTExtractor.grid_coords(g["keypoints"], (ow, oh), (fw, fh), (H, W))
    ref = eo.hybrid_descriptors(fm, g["keypoints"], (ow, oh), (fw, fh), d, projection=proj)
    assert np.array_equal(yx, ref["coords"])
    coords = torch.from_numpy(yx).cuda()[None].contiguous()
    counts = torch.tensor([n], dtype=torch.int32, device="cuda")
    _, dbf, du8, df32 = pp.sample_pack(_tokens(fm), (H, W), coords, counts, (ow, oh), (fw, fh), projection=proj,
                                       want_f32=True, rootsift=True)
    got = df32[0, :n].cpu().numpy()
    assert np.abs(got - ref["desc_float"]).max() < 1e-5        # tolerance: fp32 summation order
    diff = np.abs(du8[0, :n].cpu().numpy().astype(int) - g["desc_u8"].astype(int))
    assert diff.max() <= 1 and (diff > 0).mean() < 2e-3         # trunc edge of 512 v
    assert np.abs(dbf[0, :n].float().cpu().numpy() - got).max() <= 2 ** -8


def test_rootsift_flag_batched_bf16_ragged(pp):
    """Two images, ragged counts, bf16 tokens, no projection: rows >= count untouched, unit norm."""
    import torch
    from oracle import extract_oracle as eo
    C, H, W = 128, 12, 17
    fms = [eo.synthetic_feature_map(s, C, H, W) for s in (21, 22)]
    tok = torch.cat([_tokens(f, torch.bfloat16) for f in fms])
    rng = np.random.default_rng(5)
    n = [40, 7]
    yx = np.zeros((2, 40, 2), np.float32)
    for i in range(2):
        yx[i, : n[i]] = np.stack([rng.uniform(0, H - 1, n[i]), rng.uniform(0, W - 1, n[i])], 1)
    counts = torch.tensor(n, dtype=torch.int32, device="cuda")
    _, dbf, du8, df32 = pp.sample_pack(tok, (H, W), torch.from_numpy(yx).cuda(), counts, (W * 14, H * 14),
                                       (W * 14, H * 14), want_f32=True, rootsift=True)
    for i in range(2):
        fm16 = fms[i].to(torch.bfloat16).float()
        ref = eo.rootsift(eo.hybrid_sample(fm16, yx[i, : n[i]])).numpy()
        assert np.abs(df32[i, : n[i]].cpu().numpy() - ref).max() < 1e-5
        assert (du8[i, n[i]:] == 0).all() and (df32[i, n[i]:] == 0).all()
