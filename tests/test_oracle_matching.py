"""Pins the CPU oracle (oracle/match_oracle.c + numpy restatement) against COLMAP's own
known-answer tests: third_party/colmap/src/colmap/feature/sift_test.cc:377-555 (KA1-KA4),
feature/pairing_test.cc:55-87 (KA5), feature/utils_test.cc:47-77 (KA6)."""
import numpy as np
import pytest

from oracle import matching as om


def _rng(seed=0):
    return np.random.default_rng(seed)


def _all_impls(d1, d2, **kw):
    a = om.match_c(d1, d2, **kw)
    b = om.match_numpy(d1, d2, **kw)
    c = om.match_python(d1, d2, **kw)
    assert a.tolist() == b.tolist() == c.tolist()
    return a


def test_ka1_nominal_two_descriptors_reversed():
    # sift_test.cc:377-428
    d1 = om.ten_hot_descriptors(2, _rng())
    assert set(np.unique(d1)) == {0, 162}  # round(512/sqrt(10))
    d2 = d1[::-1].copy()
    m = _all_impls(d1, d2)
    assert m.tolist() == [[0, 1], [1, 0]]
    empty = np.zeros((0, 128), np.uint8)
    assert len(_all_impls(empty, d2)) == 0
    assert len(_all_impls(d1, empty)) == 0
    assert len(_all_impls(empty, empty)) == 0


def test_ka2_fifty_reversed():
    # sift_test.cc:499-506
    d1 = om.ten_hot_descriptors(50, _rng(1))
    m = _all_impls(d1, d1[::-1].copy())
    assert len(m) == 50
    assert all(j == 49 - i for i, j in m.tolist())


def test_ka3_ratio_test():
    # sift_test.cc:509-535
    for seed in range(20):
        d1 = om.ten_hot_descriptors(50, _rng(seed))
        if d1[0, 0] == 0:  # avoids the uint8 += 100 overflow the C++ test leaves to UB
            break
    d2 = d1.copy()
    assert len(_all_impls(d1, d2)) == 50
    d2[49] = d2[0]
    d2[0, 0] += 50
    d2[0] = om.to_uint8_colmap(om.l2_normalize(d2[0:1].astype(np.float32)))[0]
    d2[49, 0] += 100
    d2[49] = om.to_uint8_colmap(om.l2_normalize(d2[49:50].astype(np.float32)))[0]
    assert len(_all_impls(d1[:49], d2, max_ratio=0.4)) == 48
    assert len(_all_impls(d1, d2, max_ratio=0.6)) == 49


def test_ka4_cross_check():
    # sift_test.cc:538-554
    d1 = om.ten_hot_descriptors(50, _rng(3))
    d2 = d1.copy()
    d1[0] = d1[1]
    assert len(_all_impls(d1, d2, cross_check=False)) == 50
    assert len(_all_impls(d1, d2, cross_check=True)) == 48


def test_ka5_exhaustive_pairs():
    # pairing_test.cc:55-87: 34 images, block 10 -> 16 Next() calls, all 561 pairs once
    ids = np.arange(1, 35, dtype=np.uint32)
    pairs, sizes = om.exhaustive_pairs(ids, block_size=10)
    assert len(sizes) == 16
    assert len(pairs) == 34 * 33 // 2
    uniq = {(min(a, b), max(a, b)) for a, b in pairs.tolist()}
    assert len(uniq) == 34 * 33 // 2
    assert all(a != b for a, b in pairs.tolist())
    # pair id (util/types.h:200-208)
    assert om.pair_id(3, 7) == 2147483647 * 3 + 7 == om.pair_id(7, 3)


def test_ka6_uint8_quantiser():
    # utils_test.cc:65-77: round(512*x) clipped at 255
    x = np.array([[0.0, 0.01, 0.1, 0.3, 0.5, 0.6]], np.float32)
    assert om.to_uint8_colmap(x).tolist() == [[0, 5, 51, 154, 255, 255]]


@pytest.mark.parametrize("n1,n2,d", [(37, 53, 128), (64, 64, 32), (1, 9, 128), (130, 7, 128)])
def test_c_vs_numpy_u8_random(n1, n2, d):
    rng = _rng(n1 * 1000 + n2)
    f1 = rng.standard_normal((n1, d)).astype(np.float32)
    f2 = np.concatenate([f1[: n2 // 2] + 0.05 * rng.standard_normal((n2 // 2, d)).astype(np.float32),
                         rng.standard_normal((n2 - n2 // 2, d)).astype(np.float32)])[:n2]
    d1 = om.to_uint8_colmap(om.l2_normalize(np.abs(f1)))
    d2 = om.to_uint8_colmap(om.l2_normalize(np.abs(f2)))
    for cc in (True, False):
        for ratio in (0.8, 0.95):
            _all_impls(d1, d2, max_ratio=ratio, max_distance=0.7, cross_check=cc)


def test_c_vs_numpy_float_mode_and_ties():
    rng = _rng(7)
    a = om.l2_normalize(rng.standard_normal((96, 64)).astype(np.float32))
    b = np.concatenate([a[:40] + 0.02 * rng.standard_normal((40, 64)).astype(np.float32),
                        rng.standard_normal((30, 64)).astype(np.float32)])
    b = om.l2_normalize(b).astype(np.float32)
    b[50] = b[3]          # exact duplicate column -> best == second for row 3
    a[90] = 0.0           # all-zero row: no score > 0
    a[91] = -a[5]         # all scores flip sign
    m = _all_impls(a, b, max_ratio=0.9, max_distance=1.2)
    assert 3 not in m[:, 0].tolist()
    assert 90 not in m[:, 0].tolist()
    m2, rows, cols = om.match_c(a, b, max_ratio=0.9, max_distance=1.2, return_top2=True)
    assert m2.tolist() == m.tolist()
    assert rows[90, 2] == -1 and rows[3, 0] == rows[3, 1]


def test_dummy_descriptors_saturate_angular_rule():
    # uniform random u8 rows have norm >> 512 -> acos(min(.,1)) = 0 for
    # best and second -> ratio test rejects everything (sift.cc:765-780)
    rng = np.random.RandomState(42)
    d = rng.randint(0, 256, (40, 128)).astype(np.uint8)
    assert len(_all_impls(d, d.copy())) == 0


def test_l2_rule_known_answers():
    """CPU index rule (sift.cc:827-915) done exactly.  KA2-style: 50 ten-hot rows vs reversed -> 50
    matches under both rules; KA8: DummyExtractor descriptors (uniform uint8, norm ~1670 > 512)
    match identical grid cells one-to-one under l2 and not at all under the angular rule."""
    d1 = om.ten_hot_descriptors(50, np.random.default_rng(1))
    m = om.match_l2_numpy(d1, d1[::-1].copy(), 0.8, 0.7, True)
    assert len(m) == 50 and all(j == 49 - i for i, j in m.tolist())
    cells = [np.random.RandomState(42 + 1000 * gx + gy).randint(0, 256, 128, dtype=np.uint8)
             for gy in range(15) for gx in range(20)]  # dummy_extractor.py:100-109
    a = np.stack(cells)
    perm = np.random.default_rng(2).permutation(300)
    b = a[perm]
    m = om.match_l2_numpy(a, b, 0.8, 0.7, True)
    assert len(m) == 300 and np.array_equal(perm[m[:, 1]], m[:, 0])
    assert len(om.match_numpy(a, b, 0.8, 0.7, True)) == 0
    # distance threshold: a near-duplicate farther than 512*max_distance is rejected
    c = a[:1].copy().astype(np.int64)
    c[0, :64] = np.clip(c[0, :64] + 60, 0, 255)
    assert len(om.match_l2_numpy(a[:1], c.astype(np.uint8), 0.8, 0.7, False)) == 0  # |diff| ~ 480 > 358
    assert len(om.match_l2_numpy(a[:1], c.astype(np.uint8), 0.8, 1.0, False)) == 1
    # empty sides
    assert om.match_l2_numpy(a[:0], b).shape == (0, 2)


def test_guided_oracle_structural_answers():
    """Guided matching restatement (sift.cc:1058-1157 / ProgramCU.cu:1521-1663) on a scene whose answer
    is structural: image 2 = permuted image 1 moved by a homography.  With the model every row matches
    through the permutation; moving ten keypoints off the model removes exactly those; a configuration
    without a model returns nothing; the GPU and CPU rules agree when no residual is near the bound."""
    rng = np.random.default_rng(0)
    n = 60
    kp1 = rng.uniform(0, 640, (n, 2)).astype(np.float32)
    H = np.array([[1.01, 0.02, 5.0], [-0.01, 0.99, -3.0], [1e-5, 2e-5, 1.0]])
    p = np.c_[kp1, np.ones(n)] @ H.T
    kp2 = (p[:, :2] / p[:, 2:]).astype(np.float32)
    d1 = om.ten_hot_descriptors(n, rng)
    perm = rng.permutation(n)
    d2, kp2 = d1[perm], kp2[perm]
    for rule in ("angular", "l2"):
        m = om.match_guided_numpy(d1, d2, kp1, kp2, 4, None, H, 4.0, rule)
        assert len(m) == n and all(perm[j] == i for i, j in m)
        moved = kp2.copy()
        moved[:10] += 50
        m2 = om.match_guided_numpy(d1, d2, kp1, moved, 4, None, H, 4.0, rule)
        assert sorted(m2[:, 1].tolist()) == list(range(10, n))
        assert len(om.match_guided_numpy(d1, d2, kp1, kp2, 1, None, H, 4.0, rule)) == 0       # DEGENERATE
        assert len(om.match_guided_numpy(d1[:0], d2, kp1[:0], kp2, 4, None, H, 4.0, rule)) == 0
    # fundamental matrix of a pure x-translation: corresponding points share their y coordinate
    F = np.array([[0.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]])
    kp2f = kp1.copy()
    kp2f[:, 0] += rng.uniform(-80, 80, n).astype(np.float32)
    kp2f = kp2f[perm]
    hd, se = om.guided_residuals(np.eye(3), F, kp1, kp2f)
    assert se.dtype == np.float32 and np.allclose(se[np.arange(n), np.argsort(perm)], 0, atol=1e-6)
    # Sampson-style residual of a point dy pixels off its epipolar line: dy^2 / 2
    off = kp2f.copy()
    off[:, 1] += 3.0
    _, se3 = om.guided_residuals(np.eye(3), F, kp1, off)
    assert np.allclose(se3[np.arange(n), np.argsort(perm)], 4.5, rtol=1e-4)
    m = om.match_guided_numpy(d1, d2, kp1, kp2f, 3, F, None, 1.0, "angular")
    assert len(m) == n
