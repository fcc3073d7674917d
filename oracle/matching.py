"""Oracle for descriptor matching and pair generation.  TEST INFRASTRUCTURE.

Two restatements of COLMAP's brute-force matcher
(third_party/colmap/src/colmap/feature/sift.cc:733-825,922-971):

* ``match_c``      -- ctypes binding of oracle/match_oracle.c (OpenMP, fast enough
                      for N=M=8192 samples; used as the CPU baseline).
* ``match_numpy``  -- independent vectorised numpy restatement, used to cross-check
                      the C code on small cases.
* ``match_python`` -- literal pure-Python loops (tiny cases only).

Parity pin: COLMAP's own known-answer tests (sift_test.cc:377-555,
pairing_test.cc:55-87) are restated in tests/test_oracle_matching.py.
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "libvc_oracle.so"
_lib = None

SQ_SIFT_NORM = 512.0 * 512.0  # feature/sift.cc:63


def build(force: bool = False) -> Path:
    """Compile oracle/match_oracle.c (gcc, OpenMP) into oracle/_build/."""
    src = _HERE / "match_oracle.c"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE)] + (["-B"] if force else []), check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(str(_LIB_PATH))
        lib.vc_oracle_match.restype = ctypes.c_int
        lib.vc_oracle_match.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, ctypes.c_float, ctypes.c_float, ctypes.c_int,
            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.vc_oracle_exhaustive_pairs.restype = ctypes.c_int64
        lib.vc_oracle_exhaustive_pairs.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
            ctypes.c_int, ctypes.c_void_p]
        lib.vc_oracle_pair_id.restype = ctypes.c_int64
        lib.vc_oracle_pair_id.argtypes = [ctypes.c_uint32, ctypes.c_uint32]
        lib.vc_oracle_num_threads.restype = ctypes.c_int
        lib.vc_oracle_set_num_threads.argtypes = [ctypes.c_int]
        lib.vc_oracle_set_num_threads.restype = None
        _lib = lib
    return _lib


def set_num_threads(n: int) -> None:
    _load().vc_oracle_set_num_threads(int(n))


def num_threads() -> int:
    return int(_load().vc_oracle_num_threads())


def match_c(desc1: np.ndarray, desc2: np.ndarray, max_ratio: float = 0.8,
            max_distance: float = 0.7, cross_check: bool = True, fast_float: bool = False,
            return_top2: bool = False):
    """FindBestMatchesBruteForce (sift.cc:789-825) through the C oracle.

    uint8 input -> COLMAP-compatible integer mode (scale 1/512^2);
    float input -> float-d mode (scale 1, rows expected unit-norm).
    Returns matches [n,2] uint32 (and, optionally, row/col (best, second, idx) arrays).
    """
    lib = _load()
    n1, n2 = int(desc1.shape[0]), int(desc2.shape[0])
    d = int(desc1.shape[1]) if desc1.ndim == 2 else int(desc2.shape[1])
    if desc1.dtype == np.uint8:
        a = np.ascontiguousarray(desc1, dtype=np.uint8)
        b = np.ascontiguousarray(desc2, dtype=np.uint8)
        mode = 1
    else:
        a = np.ascontiguousarray(desc1, dtype=np.float32)
        b = np.ascontiguousarray(desc2, dtype=np.float32)
        mode = 2 if fast_float else 0
    out = np.zeros((max(n1, 1), 2), dtype=np.uint32)
    rows = np.zeros((max(n1, 1), 3), dtype=np.float32)
    cols = np.zeros((max(n2, 1), 3), dtype=np.float32)
    n = lib.vc_oracle_match(a.ctypes.data, n1, b.ctypes.data, n2, d, mode,
                            float(max_ratio), float(max_distance), int(bool(cross_check)),
                            out.ctypes.data, rows.ctypes.data if return_top2 else None,
                            cols.ctypes.data if return_top2 else None)
    matches = out[:n].copy()
    if return_top2:
        return matches, rows[:n1], cols[:n2]
    return matches


def _scores(desc1, desc2):
    if desc1.dtype == np.uint8:
        s = desc1.astype(np.int64) @ desc2.astype(np.int64).T  # sift.cc:939-961
        return s.astype(np.float32), np.float32(1.0 / SQ_SIFT_NORM)
    s = desc1.astype(np.float64) @ desc2.astype(np.float64).T
    return s.astype(np.float32), np.float32(1.0)


def _one_way_numpy(s: np.ndarray, inv_norm, max_ratio, max_distance) -> np.ndarray:
    """FindBestMatchesOneWayBruteForce (sift.cc:733-787), vectorised over rows."""
    n1, n2 = s.shape
    out = np.full(n1, -1, dtype=np.int64)
    if n2 == 0:
        return out
    idx = np.argmax(s, axis=1)  # first occurrence of the max == strict '>' scan
    best = s[np.arange(n1), idx]
    masked = s.copy()
    masked[np.arange(n1), idx] = -np.inf
    second = masked.max(axis=1) if n2 > 1 else np.full(n1, -np.inf, dtype=np.float32)
    second = np.maximum(second, 0).astype(np.float32)  # init 0 (sift.cc:743)
    valid = best > 0  # init best = 0, strict '>' (sift.cc:742,747)
    best = np.where(valid, best, 0).astype(np.float32)
    d1 = np.arccos(np.minimum(np.float32(inv_norm) * best, np.float32(1.0))).astype(np.float32)
    d2 = np.arccos(np.minimum(np.float32(inv_norm) * second, np.float32(1.0))).astype(np.float32)
    ok = valid & ~(d1 > np.float32(max_distance)) & ~(d1 >= np.float32(max_ratio) * d2)
    out[ok] = idx[ok]
    return out


def match_numpy(desc1, desc2, max_ratio=0.8, max_distance=0.7, cross_check=True):
    n1, n2 = desc1.shape[0], desc2.shape[0]
    if n1 == 0 or n2 == 0:  # sift.cc:1002-1005
        return np.zeros((0, 2), dtype=np.uint32)
    s, inv = _scores(desc1, desc2)
    m12 = _one_way_numpy(s, inv, max_ratio, max_distance)
    if cross_check:
        m21 = _one_way_numpy(np.ascontiguousarray(s.T), inv, max_ratio, max_distance)
        keep = [(i, j) for i, j in enumerate(m12) if j != -1 and m21[j] == i]
    else:
        keep = [(i, j) for i, j in enumerate(m12) if j != -1]
    return np.asarray(keep, dtype=np.uint32).reshape(-1, 2)


def match_python(desc1, desc2, max_ratio=0.8, max_distance=0.7, cross_check=True):
    """Literal loops of sift.cc:733-825 (tiny inputs only)."""
    n1, n2 = desc1.shape[0], desc2.shape[0]
    if n1 == 0 or n2 == 0:
        return np.zeros((0, 2), dtype=np.uint32)
    s, inv = _scores(desc1, desc2)
    f32 = np.float32

    def one_way(mat):
        res = [-1] * mat.shape[0]
        for i in range(mat.shape[0]):
            best_idx, best, second = -1, f32(0), f32(0)
            for j in range(mat.shape[1]):
                v = mat[i, j]
                if v > best:
                    best_idx, second, best = j, best, v
                elif v > second:
                    second = v
            if best_idx == -1:
                continue
            t1 = f32(math.acos(min(float(f32(inv) * best), 1.0)))
            if t1 > f32(max_distance):
                continue
            t2 = f32(math.acos(min(float(f32(inv) * second), 1.0)))
            if t1 >= f32(max_ratio) * t2:
                continue
            res[i] = best_idx
        return res

    m12 = one_way(s)
    out = []
    if cross_check:
        m21 = one_way(s.T)
        for i, j in enumerate(m12):
            if j != -1 and m21[j] != -1 and m21[j] == i:
                out.append((i, j))
    else:
        out = [(i, j) for i, j in enumerate(m12) if j != -1]
    return np.asarray(out, dtype=np.uint32).reshape(-1, 2)


def match_l2_numpy(desc1, desc2, max_ratio=0.8, max_distance=0.7, cross_check=True, filtered=None):
    """FindBestMatchesIndex / FindBestMatchesOneWayIndex (COLMAP/feature/sift.cc:827-915) with the
    index replaced by an exact search: uint8 descriptors, squared L2 distances (exact in int64),
    first minimum wins (strict '<'), best > 512^2 * max_distance^2 rejected, sqrt(best) >=
    max_ratio * sqrt(second) rejected, second = FLT_MAX when there is only one candidate."""
    a = np.asarray(desc1).astype(np.int64)
    b = np.asarray(desc2).astype(np.int64)
    if a.shape[0] == 0 or b.shape[0] == 0:
        return np.zeros((0, 2), np.uint32)
    d2 = (a * a).sum(1)[:, None] + (b * b).sum(1)[None, :] - 2 * (a @ b.T)
    if filtered is not None:  # guided matching: ComputeSiftDistanceMatrix(L2, guided_filter), sift.cc:945-946
        d2 = np.where(filtered, np.int64(SQ_SIFT_NORM), d2)
    max_l2 = np.float32(SQ_SIFT_NORM) * np.float32(max_distance) * np.float32(max_distance)

    def one_way(mat):
        out = np.full(mat.shape[0], -1, np.int64)
        for i in range(mat.shape[0]):
            row = mat[i]
            j = int(np.argmin(row))  # first minimum
            best = np.float32(row[j])
            second = np.float32(np.partition(row, 1)[1]) if row.shape[0] > 1 else np.float32(np.finfo(np.float32).max)
            if best > max_l2:
                continue
            if np.sqrt(best) >= np.float32(max_ratio) * np.sqrt(second):
                continue
            out[i] = j
        return out

    m12 = one_way(d2)
    if cross_check:
        m21 = one_way(d2.T)
        keep = [(i, j) for i, j in enumerate(m12) if j >= 0 and m21[j] == i]
    else:
        keep = [(i, j) for i, j in enumerate(m12) if j >= 0]
    return np.array(keep, dtype=np.uint32).reshape(-1, 2)


# TwoViewGeometry::ConfigurationType values that carry F resp. H (scene/two_view_geometry.h:40-61)
F_CONFIGS, H_CONFIGS = (2, 9, 3), (4, 5, 6)


def guided_residuals(H, F, kp1, kp2):
    """fp32 residuals of every candidate (i, j), each operation separately rounded (numpy float32):
    hd[i, j] = |hnormalized(H p1_i) - p2_j|^2                      (sift.cc:1116-1121, ProgramCU.cu:1576-1583)
    se[i, j] = (p2_j^T F p1_i)^2 / (Fp1_x^2 + Fp1_y^2 + Ftp2_x^2 + Ftp2_y^2)   (sift.cc:1099-1110, ProgramCU.cu:1586-1597)"""
    f32 = np.float32
    H, F = np.asarray(H, f32).reshape(3, 3), np.asarray(F, f32).reshape(3, 3)
    x1, y1 = np.asarray(kp1, f32)[:, 0][:, None], np.asarray(kp1, f32)[:, 1][:, None]
    x2, y2 = np.asarray(kp2, f32)[:, 0][None, :], np.asarray(kp2, f32)[:, 1][None, :]
    with np.errstate(all="ignore"):
        a = (H[0, 0] * x1 + H[0, 1] * y1) + H[0, 2]
        b = (H[1, 0] * x1 + H[1, 1] * y1) + H[1, 2]
        c = (H[2, 0] * x1 + H[2, 1] * y1) + H[2, 2]
        dx, dy = a / c - x2, b / c - y2
        hd = dx * dx + dy * dy
        fx0 = (F[0, 0] * x1 + F[0, 1] * y1) + F[0, 2]
        fx1 = (F[1, 0] * x1 + F[1, 1] * y1) + F[1, 2]
        fx2 = (F[2, 0] * x1 + F[2, 1] * y1) + F[2, 2]
        ft0 = (F[0, 0] * x2 + F[1, 0] * y2) + F[2, 0]
        ft1 = (F[0, 1] * x2 + F[1, 1] * y2) + F[2, 1]
        e = (x2 * fx0 + y2 * fx1) + fx2
        den = ((fx0 * fx0 + fx1 * fx1) + ft0 * ft0) + ft1 * ft1
        se = (e * e) / den
    assert hd.dtype == f32 and se.dtype == f32
    return hd, se


def match_guided_numpy(desc1, desc2, kp1, kp2, config, F, H, max_error=4.0, rule="angular", max_ratio=0.8,
                       max_distance=0.7, cross_check=True, return_filtered=False):
    """Guided matching.  rule "angular": SiftGPUFeatureMatcher::MatchGuided (sift.cc:1308-1413) ->
    SiftMatchGPU::GetGuidedSiftMatch (SiftMatch.cpp:676-689: the absent model becomes the identity with
    bound 1e20) -> MultiplyDescriptorG_Kernel (ProgramCU.cu:1521-1663: candidate kept iff hd < hmax and
    se < fmax, a dropped one ends <= 0 and never matches) -> GetBestMatch.  rule "l2":
    SiftCPUFeatureMatcher::MatchGuided (sift.cc:1058-1157): one test by `config`, dropped iff residual >
    max_error^2, dropped candidates take part with squared distance 512^2, FindBestMatchesIndex."""
    n1, n2 = desc1.shape[0], desc2.shape[0]
    empty = np.zeros((0, 2), np.uint32)
    use_f = config in F_CONFIGS
    if not use_f and config not in H_CONFIGS:
        return (empty, None) if return_filtered else empty  # sift.cc:1122-1124 / :1385-1387
    if n1 == 0 or n2 == 0:
        return (empty, None) if return_filtered else empty
    eye = np.eye(3, dtype=np.float32)
    hd, se = guided_residuals(eye if use_f else H, F if use_f else eye, kp1[:n1, :2], kp2[:n2, :2])
    if rule == "angular":
        res = np.float32(max_error * max_error)  # sift.cc:1394: the double product, cast to float
        hmax, fmax = (np.float32(1e20), res) if use_f else (res, np.float32(1e20))
        filtered = ~((hd < hmax) & (se < fmax))
        s, inv = _scores(desc1, desc2)
        s = np.where(filtered, np.float32(0), s)
        m12 = _one_way_numpy(s, inv, max_ratio, max_distance)
        if cross_check:
            m21 = _one_way_numpy(np.ascontiguousarray(s.T), inv, max_ratio, max_distance)
            keep = [(i, j) for i, j in enumerate(m12) if j != -1 and m21[j] == i]
        else:
            keep = [(i, j) for i, j in enumerate(m12) if j != -1]
        out = np.asarray(keep, dtype=np.uint32).reshape(-1, 2)
    elif rule == "l2":
        res = np.float32(max_error * max_error)  # sift.cc:1091
        filtered = (se > res) if use_f else (hd > res)
        out = match_l2_numpy(desc1, desc2, max_ratio, max_distance, cross_check, filtered=filtered)
    else:
        raise ValueError(rule)
    return (out, filtered) if return_filtered else out


def exhaustive_pairs(image_ids, block_size: int = 50):
    """ExhaustivePairGenerator (pairing.cc:171-236).  Returns (pairs [n,2] uint32,
    per-Next() block sizes)."""
    lib = _load()
    ids = np.ascontiguousarray(np.asarray(image_ids, dtype=np.uint32))
    n = len(ids)
    out = np.zeros((max(n * (n - 1) // 2, 1), 2), dtype=np.uint32)
    nblk_max = (math.ceil(n / block_size) ** 2) if n else 1
    sizes = np.zeros(max(nblk_max, 1), dtype=np.int32)
    nb = ctypes.c_int(0)
    total = lib.vc_oracle_exhaustive_pairs(ids.ctypes.data, n, int(block_size), out.ctypes.data,
                                           sizes.ctypes.data, len(sizes), ctypes.byref(nb))
    return out[:total].copy(), sizes[:nb.value].copy()


def pair_id(id1: int, id2: int) -> int:
    """ImagePairToPairId (util/types.h:200-208)."""
    return int(_load().vc_oracle_pair_id(int(id1), int(id2)))


def to_uint8_colmap(desc: np.ndarray) -> np.ndarray:
    """FeatureDescriptorsToUnsignedByte (feature/utils.cc:56-71): min(255, round(512*x))."""
    scaled = np.round(512.0 * desc.astype(np.float32))
    return np.clip(scaled, 0, 255).astype(np.uint8)


def l2_normalize(desc: np.ndarray) -> np.ndarray:
    """L2NormalizeFeatureDescriptors (feature/utils.cc:46-48)."""
    n = np.linalg.norm(desc, axis=1, keepdims=True)
    return desc / n


def ten_hot_descriptors(num: int, rng: np.random.Generator, dim: int = 128) -> np.ndarray:
    """CreateRandomFeatureDescriptors (feature/sift_test.cc:300-314): each row has 10
    ones at distinct random positions, L2-normalised, quantised with the COLMAP rule
    (=> non-zeros are round(512/sqrt(10)) = 162).  The C++ mt19937/shuffle stream is
    not reproduced; the known answers are structural for distinct rows."""
    out = np.zeros((num, dim), dtype=np.float32)
    seen = set()
    for i in range(num):
        while True:
            pos = tuple(sorted(rng.choice(dim, size=10, replace=False).tolist()))
            if pos not in seen:
                seen.add(pos)
                break
        out[i, list(pos)] = 1.0
    return to_uint8_colmap(l2_normalize(out))
