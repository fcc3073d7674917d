"""GPU parity tests: the fused sm_100a matcher (through the C ABI) against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def matcher():
    from vit_colmap_b200.matching import Matcher
    m = Matcher(0)
    yield m
    m.close()


def _unit_rows(rng, n, d):
    x = rng.standard_normal((n, d)).astype(np.float32)
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def _bf16_round(x):
    import torch
    return torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()


def _make_set(rng, n_images, kmax, d, counts):
    """Image j shares half its rows with image j-1 (perturbed) -- SURVEY 8(d) C4 recipe."""
    desc = np.zeros((n_images, kmax, d), np.float32)
    for j in range(n_images):
        n = counts[j]
        x = _unit_rows(rng, n, d)
        if j > 0 and counts[j - 1] > 0 and n > 0:
            m = min(n, counts[j - 1]) // 2
            src = rng.permutation(counts[j - 1])[:m]
            dst = rng.permutation(n)[:m]
            y = desc[j - 1, src] + (0.35 / d ** 0.5) * rng.standard_normal((m, d)).astype(np.float32)
            x[dst] = y / np.linalg.norm(y, axis=1, keepdims=True)
        desc[j, :n] = x
    return _bf16_round(desc)


def test_debug_scores_match_fp32_gemm(matcher):
    """TMA + tcgen05 main loop alone: dense scores vs a float64 host GEMM."""
    import torch
    rng = np.random.default_rng(0)
    kmax, d = 300, 192
    counts = [300, 257]
    desc = _make_set(rng, 2, kmax, d, counts)
    t = torch.from_numpy(desc).cuda().to(torch.bfloat16)
    c = torch.tensor(counts, dtype=torch.int32, device="cuda")
    s = matcher.debug_scores(t, c, 0, 1).cpu().numpy()[: counts[0], : counts[1]]
    ref = desc[0, : counts[0]].astype(np.float64) @ desc[1, : counts[1]].astype(np.float64).T
    assert np.abs(s - ref).max() < 1e-5  # fp32 accumulation of exact bf16 products


def _compare_with_oracle(matcher, desc, counts, pairs, max_ratio, max_distance, cross_check, tol=1e-5):
    """Bit-exact match lists, except where the oracle's own margins are inside the documented
    near-tie tolerance (score gap < 1e-5, north star)."""
    import torch
    from oracle import matching as om
    t = torch.from_numpy(desc).cuda().to(torch.bfloat16)
    c = torch.tensor(counts, dtype=torch.int32, device="cuda")
    p = torch.tensor(pairs, dtype=torch.int32, device="cuda").reshape(-1, 2)
    matches, n = matcher.match_pairs(t, c, p, max_ratio, max_distance, cross_check, 1.0)
    best, second, index = matcher.top2(t, c, p, max_ratio, max_distance, cross_check, 1.0)
    torch.cuda.synchronize()
    matches, n = matches.cpu().numpy().astype(np.uint32), n.cpu().numpy()
    best, second, index = best.cpu().numpy(), second.cpu().numpy(), index.cpu().numpy()
    n_exact = 0
    for k, (i1, i2) in enumerate(pairs):
        a, b = desc[i1, : counts[i1]], desc[i2, : counts[i2]]
        ref, rows, cols = om.match_c(a, b, max_ratio, max_distance, cross_check, return_top2=True)
        got = matches[k, : n[k]]
        for dirn, (top, cnt) in enumerate(((rows, counts[i1]), (cols, counts[i2]))):
            if cnt == 0 or min(counts[i1], counts[i2]) == 0:
                continue
            assert np.abs(best[k, dirn, :cnt] - top[:, 0]).max() < tol
            assert np.abs(second[k, dirn, :cnt] - top[:, 1]).max() < tol
            differ = index[k, dirn, :cnt] != top[:, 2].astype(np.int32)
            # an index may differ only when best and second are within the tolerance
            assert np.all((top[differ, 0] - top[differ, 1]) < tol)
        if got.tolist() == ref.tolist():
            n_exact += 1
            continue
        # differences must be attributable to near-ties in the oracle's margins
        gs, rs = set(map(tuple, got.tolist())), set(map(tuple, ref.tolist()))
        for (i, j) in gs ^ rs:
            t1 = np.arccos(min(rows[i, 0], 1.0)); t2 = np.arccos(min(rows[i, 1], 1.0))
            u1 = np.arccos(min(cols[j, 0], 1.0)); u2 = np.arccos(min(cols[j, 1], 1.0))
            margins = [abs(t1 - max_distance), abs(t1 - max_ratio * t2), abs(u1 - max_distance),
                       abs(u1 - max_ratio * u2), rows[i, 0] - rows[i, 1], cols[j, 0] - cols[j, 1]]
            assert min(margins) < 50 * tol, f"pair {k}: ({i},{j}) differs without a near-tie"
    return n_exact


@pytest.mark.parametrize("kmax,d,counts", [
    (256, 128, [256, 256, 256]),
    (300, 64, [300, 129, 1, 0]),       # ragged, single row, empty image
    (1000, 384, [1000, 777, 513]),
    (2048, 768, [2048, 2047, 1739]),   # north-star d; 1739 = reference keypoint yield at DTU size
    (1100, 1024, [1100, 1025, 640]),   # ViT-L width: the two-SM (cta_group::2) kernel; odd row-tile counts
    (700, 1024, [700, 100, 129]),      # two-SM kernel with a single row tile (second CTA of the pair idles)
])
def test_match_pairs_float_mode(matcher, kmax, d, counts):
    rng = np.random.default_rng(kmax + d)
    desc = _make_set(rng, len(counts), kmax, d, counts)
    pairs = [(i, j) for i in range(len(counts)) for j in range(len(counts)) if i != j]
    for cross in (True, False):
        n_exact = _compare_with_oracle(matcher, desc, counts, pairs, 0.8, 0.7, cross)
        assert n_exact >= 2


def test_ties_zero_rows_and_negative_scores(matcher):
    rng = np.random.default_rng(5)
    kmax, d = 200, 128
    counts = [200, 180]
    desc = _make_set(rng, 2, kmax, d, counts)
    desc[1, 50] = desc[1, 3]      # exact duplicate -> best == second, rejected (sift.cc:778)
    desc[0, 90] = 0.0             # zero row: no score > 0 -> idx -1 (sift.cc:760-762)
    desc[0, 91] = -desc[0, 5]
    _compare_with_oracle(matcher, desc, counts, [(0, 1), (1, 0)], 0.9, 1.2, True)


def test_u8_compat_mode_bit_exact(matcher):
    """COLMAP uint8 descriptors (d=128) through the bf16 path are exact integers."""
    import torch
    from oracle import matching as om
    rng = np.random.default_rng(11)
    kmax, d = 700, 128
    counts = [700, 650, 50]
    f = np.abs(_make_set(rng, 3, kmax, d, counts))
    u8 = np.zeros((3, kmax, d), np.uint8)
    for i, n in enumerate(counts):
        u8[i, :n] = om.to_uint8_colmap(om.l2_normalize(f[i, :n]))
    pairs = [(0, 1), (1, 2), (2, 0)]
    got = matcher.match_pairs_host(u8, np.array(counts, np.int32), np.array(pairs, np.int32), 0.8, 0.7, True)
    for k, (i1, i2) in enumerate(pairs):
        ref = om.match_c(u8[i1, : counts[i1]], u8[i2, : counts[i2]], 0.8, 0.7, True)
        assert got[k].tolist() == ref.tolist()
    # KA2 (sift_test.cc:499-506): 50 ten-hot rows vs reversed -> 50 matches
    d1 = om.ten_hot_descriptors(50, np.random.default_rng(1))
    both = np.stack([d1, d1[::-1].copy()])
    (m,) = matcher.match_pairs_host(both, np.array([50, 50], np.int32), np.array([[0, 1]], np.int32))
    assert len(m) == 50 and all(j == 49 - i for i, j in m.tolist())


def test_full_size_8192_d768_against_oracle_and_properties(matcher):
    """BASELINE.json's headline size (N = M = 8192, d = 768).  (a) one pair against the C oracle
    directly (51 GFLOP on the host cores); (b) size-independent properties: a permuted copy matches
    through the permutation, swapping the pair transposes the match list, an image against itself
    yields the identity, and the ragged count 8191 works."""
    import torch
    rng = np.random.default_rng(8192)
    kmax, d = 8192, 768
    counts = [8192, 8192, 8192, 8191]
    desc = _make_set(rng, 2, kmax, d, counts[:2])
    perm = rng.permutation(kmax)
    desc = np.concatenate([desc, desc[0][perm][None], desc[1][None]], axis=0)
    desc[3, 8191] = 0
    n_exact = _compare_with_oracle(matcher, desc, counts, [(0, 1)], 0.8, 0.7, True)
    assert n_exact == 1
    t = torch.from_numpy(desc).cuda().to(torch.bfloat16)
    c = torch.tensor(counts, dtype=torch.int32, device="cuda")
    pairs = [(0, 1), (1, 0), (0, 2), (0, 0), (0, 3)]
    p = torch.tensor(pairs, dtype=torch.int32, device="cuda")
    m, n = matcher.match_pairs(t, c, p, 0.8, 0.7, True, 1.0)
    torch.cuda.synchronize()
    m, n = m.cpu().numpy(), n.cpu().numpy()
    fwd, bwd = m[0, : n[0]], m[1, : n[1]]
    assert n[0] > 3000 and np.all(np.diff(fwd[:, 0].astype(np.int64)) > 0)      # ascending idx1
    assert sorted(map(tuple, fwd.tolist())) == sorted((j, i) for i, j in bwd.tolist())
    pm = m[2, : n[2]]
    assert n[2] == kmax and np.array_equal(perm[pm[:, 1]], pm[:, 0])            # row perm[k] of image 0 is row k of image 2
    ident = m[3, : n[3]]
    assert n[3] == kmax and np.array_equal(ident[:, 0], ident[:, 1])           # an image against itself: identity
    keep = fwd[fwd[:, 1] != 8191]
    got3 = m[4, : n[4]]
    assert abs(int(n[4]) - len(keep)) <= 2 and len(set(map(tuple, got3.tolist())) ^ set(map(tuple, keep.tolist()))) <= 4


def test_l2_rule_matches_oracle(matcher):
    """rule 'l2' (COLMAP's CPU index semantics, sift.cc:827-915, exact search) through the fused GEMM
    with role-augmented descriptors: bit-exact match lists against the int64 oracle on SIFT-like and
    on uniform-random (dummy extractor) uint8 descriptors, ragged counts, both cross-check settings."""
    from oracle import matching as om
    rng = np.random.default_rng(21)
    kmax, d = 400, 128
    counts = [400, 333, 1, 0]
    f = np.abs(_make_set(rng, 4, kmax, d, counts))
    sift = np.zeros((4, kmax, d), np.uint8)
    for i, n in enumerate(counts):
        sift[i, :n] = om.to_uint8_colmap(om.l2_normalize(f[i, :n])) if n else 0
    dummy = rng.integers(0, 256, (4, kmax, d), dtype=np.uint8)
    dummy[1, :150] = dummy[0, 100:250]           # shared cells
    dummy[1, 200] = dummy[1, 201]                # exact duplicate: best == second -> rejected
    pairs = [(0, 1), (1, 0), (0, 2), (2, 1), (0, 3)]
    for desc, (ratio, dist) in ((sift, (0.8, 0.7)), (dummy, (0.8, 0.7)), (dummy, (0.95, 3.5))):
        for cross in (True, False):
            got = matcher.match_pairs_host(desc, np.array(counts, np.int32), np.array(pairs, np.int32), ratio, dist,
                                           cross, rule="l2")
            for k, (i1, i2) in enumerate(pairs):
                ref = om.match_l2_numpy(desc[i1, : counts[i1]], desc[i2, : counts[i2]], ratio, dist, cross)
                assert got[k].tolist() == ref.tolist(), (k, cross, len(got[k]), len(ref))
    assert len(matcher.match_pairs_host(dummy, np.array(counts, np.int32), np.array([[0, 1]], np.int32), rule="l2")[0]) >= 140


@pytest.mark.parametrize("dtype", ["f32", "bf16", "u8"])
def test_host_entry_chunked_upload_groups(matcher, dtype):
    """vc_match_pairs_host above its chunking threshold (>= 8 images, >= 32 MB): the upload is cut
    into four image chunks, the pairs are re-ordered by the chunk they wait for and the packed match
    lists are scattered back.  Results must come back in the CALLER's pair order and equal the
    device-resident entry (itself held to the oracle above), for shuffled / repeated / swapped /
    self pairs and ragged counts; three pairs are also checked against the C oracle directly."""
    import torch
    from oracle import matching as om
    rng = np.random.default_rng(77)
    n_images, kmax, d = 10, 2304 if dtype == "bf16" else 1536, 2304 if dtype == "u8" else 768
    counts = [kmax, 1500, 0, kmax, 777, kmax, 1, kmax, 1290, kmax]
    desc = _make_set(rng, n_images, kmax, d, counts)
    all_pairs = [(i, j) for i in range(n_images) for j in range(n_images) if i != j]
    pairs = [all_pairs[k] for k in rng.permutation(len(all_pairs))[:40]] + [(3, 3), (9, 0), (9, 0), (0, 9)]
    pairs_np = np.array(pairs, np.int32)
    cnt = np.array(counts, np.int32)
    if dtype == "u8":
        host = np.zeros(desc.shape, np.uint8)
        for i, n in enumerate(counts):
            host[i, :n] = om.to_uint8_colmap(np.abs(desc[i, :n]))   # non-negative, like SIFT rows
        dev = torch.from_numpy(host).cuda().to(torch.bfloat16)
        inv = 1.0 / 512.0 ** 2
    elif dtype == "bf16":
        host = torch.from_numpy(desc).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
        dev = torch.from_numpy(desc).cuda().to(torch.bfloat16)
        inv = 1.0
    else:
        host = desc
        dev = torch.from_numpy(desc).cuda().to(torch.bfloat16)
        inv = 1.0
    assert host.nbytes >= 32 << 20
    got = matcher.match_pairs_host(host, cnt, pairs_np, 0.8, 0.7, True)
    m, n = matcher.match_pairs(dev, torch.from_numpy(cnt).cuda(), torch.from_numpy(pairs_np).cuda(), 0.8, 0.7, True, inv)
    torch.cuda.synchronize()
    total = 0
    for k in range(len(pairs)):
        ref = m[k, : int(n[k])].cpu().numpy().astype(np.uint32)
        assert got[k].tolist() == ref.tolist(), (k, pairs[k])
        total += len(ref)
    assert total > 2000
    for k in (0, 1, len(pairs) - 1) if dtype != "u8" else ():
        i1, i2 = pairs[k]
        ref = om.match_c(desc[i1, : counts[i1]], desc[i2, : counts[i2]], 0.8, 0.7, True)
        assert got[k].tolist() == ref.tolist()
    # a second call with fewer pairs re-uses the staging buffers
    got2 = matcher.match_pairs_host(host, cnt, pairs_np[:5], 0.8, 0.7, True)
    assert [g.tolist() for g in got2] == [g.tolist() for g in got[:5]]
    with pytest.raises(Exception):
        matcher.match_pairs_host(host, cnt, np.array([[0, n_images]], np.int32))


@pytest.mark.parametrize("max_ratio,max_distance", [(0.8, 0.7), (1.0, 0.7), (0.6, 0.9), (0.95, 1.5), (0.5, 0.3),
                                                    (0.8, 1.3), (0.9, 0.2)])
@pytest.mark.parametrize("cross_check", [True, False])
def test_pruning_bound_is_exact_near_the_bound(matcher, max_ratio, max_distance, cross_check):
    """The fused epilogues skip scores with acos(s) > max_distance / max_ratio (they can change no
    accept / reject decision, sift.cc:765-780).  Clustered descriptors put best AND second-best
    cosines on both sides of that bound (0.3 .. 0.99), for several option sets including ones where
    the bound is inactive (max_distance / max_ratio >= pi / 2); match lists must equal the C oracle's
    exactly, and must not depend on the bound (VC_MATCH_NO_PRUNE is covered by the top-2 entry, which
    runs without it and is compared value by value in _compare_with_oracle)."""
    rng = np.random.default_rng(int(max_ratio * 100) * 1000 + int(max_distance * 100))
    kmax, d = 640, 256
    counts = [640, 611, 500]
    centres = _unit_rows(rng, 40, d)
    desc = np.zeros((3, kmax, d), np.float32)
    for i, n in enumerate(counts):
        which = rng.integers(0, len(centres), n)
        spread = rng.uniform(0.05, 1.2, (n, 1)).astype(np.float32)     # cosine to the centre 0.99 .. 0.4
        x = centres[which] + spread / d ** 0.5 * rng.standard_normal((n, d)).astype(np.float32)
        desc[i, :n] = x / np.linalg.norm(x, axis=1, keepdims=True)
    desc = _bf16_round(desc)
    pairs = [(0, 1), (1, 2), (2, 0), (1, 1)]
    n_exact = _compare_with_oracle(matcher, desc, counts, pairs, max_ratio, max_distance, cross_check)
    assert n_exact >= 2


def test_resident_entry_equals_host_entry_and_oracle(matcher):
    """Matcher.upload (ONE upload) + match_resident (vc_match_pairs_resident: only pair lists go up,
    packed match lists come back) against vc_match_pairs_host and the C oracle, for uint8 and float
    descriptors, more pairs than one call takes (pairs_per_call) and ragged counts."""
    from oracle import matching as om
    rng = np.random.default_rng(31)
    n_images, kmax, d = 7, 500, 128
    counts = [500, 499, 0, 257, 1, 500, 333]
    f = _make_set(rng, n_images, kmax, d, counts)
    u8 = np.zeros(f.shape, np.uint8)
    for i, n in enumerate(counts):
        u8[i, :n] = om.to_uint8_colmap(om.l2_normalize(np.abs(f[i, :n]))) if n else 0
    pairs = np.array([(i, j) for i in range(n_images) for j in range(n_images) if i != j] + [(0, 0)], np.int32)
    cnt = np.array(counts, np.int32)
    for host in (u8, f):
        resident = matcher.upload(host, cnt)
        got = matcher.match_resident(resident, pairs, 0.8, 0.7, True, pairs_per_call=16)
        ref = matcher.match_pairs_host(host, cnt, pairs, 0.8, 0.7, True)
        assert [g.tolist() for g in got] == [r.tolist() for r in ref]
        for k in (0, 5, 17, len(pairs) - 1):
            i1, i2 = pairs[k]
            exact = om.match_c(host[i1, : counts[i1]], host[i2, : counts[i2]], 0.8, 0.7, True)
            assert got[k].tolist() == exact.tolist()
    assert sum(len(g) for g in got) > 500
    with pytest.raises(ValueError):
        matcher.match_resident(resident, np.array([[0, n_images]], np.int32))
    with pytest.raises(Exception):  # a host pointer is not a resident descriptor set
        import ctypes
        from vit_colmap_b200 import _lib
        opt = matcher._opts(0.8, 0.7, True, 1.0)
        out, nm = np.zeros((1, kmax, 2), np.uint32), np.zeros(1, np.int32)
        _lib.check(matcher.lib.vc_match_pairs_resident(matcher.handle, f.ctypes.data, cnt.ctypes.data, n_images, kmax,
                                                       d, pairs.ctypes.data, 1, ctypes.byref(opt), out.ctypes.data,
                                                       nm.ctypes.data))


def test_l2_augment_kernel_is_exact(matcher):
    """vc_l2_augment (device) against the integer construction it documents: rows of role 1 dotted
    with rows of role 2 give K - |x - y|^2 exactly in fp32."""
    import ctypes
    import torch
    from vit_colmap_b200 import _lib
    rng = np.random.default_rng(3)
    n_images, kmax, d = 3, 37, 128
    u8 = rng.integers(0, 256, (n_images, kmax, d), dtype=np.uint8)
    u8[1, 5] = 180                                     # 2 |x|^2 + 1 just below 2^23: the largest K allowed
    u8[2, 7] = 0
    src = torch.from_numpy(u8).cuda()
    out = torch.empty((2 * n_images, kmax, d + 8), dtype=torch.bfloat16, device="cuda")
    k = ctypes.c_float(0)
    _lib.check(matcher.lib.vc_l2_augment(matcher.handle, src.data_ptr(), n_images, kmax, d, out.data_ptr(),
                                         ctypes.byref(k), 0))
    torch.cuda.synchronize()
    aug = out.float().cpu().numpy().astype(np.float64)
    K = float(k.value)
    norm2 = (u8.astype(np.int64) ** 2).sum(-1)
    assert K == 2.0 ** np.ceil(np.log2(2 * norm2.max() + 2)) == 2 ** 23
    a, b = aug[:n_images].reshape(-1, d + 8), aug[n_images:].reshape(-1, d + 8)
    x = u8.reshape(-1, d).astype(np.int64)
    dist2 = ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1)
    assert np.array_equal(a @ b.T, K - dist2)
    assert np.abs(aug).max() <= max(K, 65536 * 255) and np.array_equal(aug, np.round(aug))


def test_more_pairs_than_a_grid_dimension(matcher):
    """n_pairs > 65535: the merge kernels index pairs on gridDim.x (a y-dimension launch would fail
    with an invalid configuration after the main kernel has already run)."""
    import torch
    from oracle import matching as om
    rng = np.random.default_rng(9)
    n_images, kmax, d = 6, 24, 32
    counts = [24, 24, 17, 24, 1, 24]
    desc = _make_set(rng, n_images, kmax, d, counts)
    base = [(i, j) for i in range(n_images) for j in range(n_images) if i != j]
    pairs = np.array([base[k % len(base)] for k in range(70001)], np.int32)
    t = torch.from_numpy(desc).cuda().to(torch.bfloat16)
    m, n = matcher.match_pairs(t, torch.tensor(counts, dtype=torch.int32, device="cuda"),
                               torch.from_numpy(pairs).cuda(), 0.9, 1.2, True, 1.0)
    torch.cuda.synchronize()
    m, n = m.cpu().numpy().astype(np.uint32), n.cpu().numpy()
    for k in list(range(len(base))) + [65535, 65536, 70000]:
        i1, i2 = pairs[k]
        ref = om.match_c(desc[i1, : counts[i1]], desc[i2, : counts[i2]], 0.9, 1.2, True)
        assert m[k, : n[k]].tolist() == ref.tolist(), k
    assert np.array_equal(n[: len(base)], n[len(base): 2 * len(base)]) and n.sum() > 70001


def test_two_handles_on_two_devices():
    """A second vc_handle on another GPU of the same process: the > 48 KB shared-memory opt-in is per
    (kernel, device), so every kernel family must work on both.  Needs two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from oracle import matching as om
    from vit_colmap_b200.matching import Matcher
    rng = np.random.default_rng(2)
    kmax, counts = 300, [300, 280]
    for d in (128, 768):  # single-CTA and two-SM kernels
        desc = _make_set(rng, 2, kmax, d, counts)
        ref = om.match_c(desc[0, :300], desc[1, :280], 0.8, 0.7, True)
        for dev in (0, 1, 0):
            m = Matcher(dev)
            with torch.cuda.device(dev):
                got = m.match_pairs_host(desc, np.array(counts, np.int32), np.array([[0, 1]], np.int32))
            assert got[0].tolist() == ref.tolist(), (d, dev)
            m.close()


@pytest.mark.parametrize("kind", ["H", "F"])
@pytest.mark.parametrize("rule", ["angular", "l2"])
def test_guided_matching_against_oracle(matcher, kind, rule):
    """the geometric gate in the GEMM epilogue (vc_match_pairs_guided) against the numpy
    restatement of MultiplyDescriptorG_Kernel + GetBestMatch (angular rule) resp. of
    SiftCPUFeatureMatcher::MatchGuided (l2 rule): match lists bit-exact (every residual operation is a
    separately rounded fp32 op on both sides), both pair orientations, cross check on and off; the gate
    must actually bite (unguided matching returns a different list)."""
    from _scenes import two_view_scene as _two_view_scene
    from oracle import matching as om
    from vit_colmap_b200.matching import PairGeometry
    rng = np.random.default_rng(41 + (kind == "F") + 2 * (rule == "l2"))
    n, kmax, d = 333, 400, 128
    desc, kps, config, G = _two_view_scene(rng, n, kmax, d, kind)
    counts = np.array([n, n], np.int32)
    resident = matcher.upload(desc, counts, rule=rule)
    F = G if kind == "F" else None
    H = G if kind == "H" else None
    H_inv = np.linalg.inv(G) if kind == "H" else None
    geo_fwd = PairGeometry(config, F, H, 2.0)
    geo_bwd = PairGeometry(config, None if F is None else F.T, H_inv, 2.0)
    ratio, dist = (0.9, 1.0) if rule == "angular" else (0.9, 0.9)
    for cross in (True, False):
        got = matcher.match_guided(resident, kps, np.array([[0, 1], [1, 0]], np.int32), [geo_fwd, geo_bwd], ratio, dist,
                                   cross)
        ref_f, filt = om.match_guided_numpy(desc[0, :n], desc[1, :n], kps[0], kps[1], config, F, H, 2.0, rule, ratio,
                                            dist, cross, return_filtered=True)
        ref_b = om.match_guided_numpy(desc[1, :n], desc[0, :n], kps[1], kps[0], config, None if F is None else F.T,
                                      H_inv, 2.0, rule, ratio, dist, cross)
        assert got[0].tolist() == ref_f.tolist(), (cross, len(got[0]), len(ref_f))
        assert got[1].tolist() == ref_b.tolist(), (cross, len(got[1]), len(ref_b))
        plain = matcher.match_resident(resident, np.array([[0, 1]], np.int32), ratio, dist, cross)[0]
        assert plain.tolist() != got[0].tolist()
    assert 0.9 < filt.mean() < 1.0 and len(ref_f) > 100
    with pytest.raises(ValueError):  # DEGENERATE / UNDEFINED carry no model (sift.cc:1122-1124)
        matcher.match_guided(resident, kps, np.array([[0, 1]], np.int32), [PairGeometry(1)], ratio, dist, True)


def test_column_merge_paths_agree(matcher, monkeypatch):
    """The fused epilogue merges column hits either lane by lane (two shared-memory atomics per hit) or,
    when many columns of a chunk are hit at once, through warp-wide maxima with one merging lane per
    column.  Forcing each path for every chunk (VC_MATCH_DENSE_MIN = 33 / 1) and the default must give
    the same top-2 values, indices and match lists -- with the pruning bound off and exact duplicates
    in both images, so that ties and second-best updates go through both paths -- and equal the oracle."""
    import torch
    from oracle import matching as om
    rng = np.random.default_rng(17)
    kmax, counts = 900, [900, 777, 640]
    monkeypatch.setenv("VC_MATCH_NO_PRUNE", "1")
    for d in (128, 768):                                  # single-CTA and two-SM kernels
        desc = np.abs(_make_set(rng, 3, kmax, d, counts))  # non-negative: every score is a candidate
        desc /= np.maximum(np.linalg.norm(desc, axis=2, keepdims=True), 1e-12)
        desc[0, 300:310] = desc[0, 20:30]                 # duplicate rows: column ties, lowest row must win
        desc[1, 400:405] = desc[1, 7:12]                  # duplicate columns: row ties, lowest column must win
        desc = _bf16_round(desc)
        t = torch.from_numpy(desc).cuda().to(torch.bfloat16)
        c = torch.tensor(counts, dtype=torch.int32, device="cuda")
        pairs = [(0, 1), (1, 2), (2, 0)]
        p = torch.tensor(pairs, dtype=torch.int32, device="cuda")
        outs = []
        for dense_min in ("33", "1", None):
            if dense_min is None:
                monkeypatch.delenv("VC_MATCH_DENSE_MIN", raising=False)
            else:
                monkeypatch.setenv("VC_MATCH_DENSE_MIN", dense_min)
            m, n = matcher.match_pairs(t, c, p, 1.05, 3.2, True, 1.0)   # mutual nearest neighbours: indices matter on ties
            top = matcher.top2(t, c, p, 1.05, 3.2, True, 1.0)
            torch.cuda.synchronize()
            outs.append(([m[k, : int(n[k])].cpu().numpy().tolist() for k in range(len(pairs))],
                         [x.cpu().numpy() for x in top]))
        for lists, top in outs[1:]:
            assert lists == outs[0][0]
            for a, b in zip(top, outs[0][1]):
                assert np.array_equal(a, b)
        for k, (i1, i2) in enumerate(pairs):
            ref = om.match_c(desc[i1, : counts[i1]], desc[i2, : counts[i2]], 1.05, 3.2, True)
            got = np.array(outs[0][0][k], np.uint32).reshape(-1, 2)
            differ = set(map(tuple, got.tolist())) ^ set(map(tuple, ref.tolist()))
            assert len(differ) <= 0.01 * len(ref) and len(ref) > 150   # fp32 near-ties only (accumulation order)


def test_descriptor_width_not_a_multiple_of_8(matcher):
    """ViTExtractor(descriptor_dim=100) writes 100-column blobs; match_descriptor_sets pads the rows with
    zero columns (dot products unchanged) -- match lists equal the oracle on the unpadded rows, both rules."""
    from oracle import matching as om
    from vit_colmap_b200.matching import FeatureMatchingOptions, match_descriptor_sets
    rng = np.random.default_rng(100)
    counts, d = [260, 200, 0], 100
    f = np.abs(_make_set(rng, 3, 260, d, counts))
    descs = [om.to_uint8_colmap(om.l2_normalize(f[i, :n])) if n else np.zeros((0, d), np.uint8) for i, n in enumerate(counts)]
    pairs = [(0, 1), (1, 0), (0, 2)]
    for rule in ("angular", "l2"):
        got = match_descriptor_sets(descs, pairs, FeatureMatchingOptions(max_ratio=0.9, max_distance=1.0, rule=rule), matcher)
        for (i, j), m in zip(pairs, got):
            ref = om.match_c(descs[i], descs[j], 0.9, 1.0, True) if rule == "angular" and counts[i] and counts[j] else \
                om.match_l2_numpy(descs[i], descs[j], 0.9, 1.0, True)
            assert m.tolist() == ref.tolist(), (rule, i, j)
    assert len(got[0]) > 50
