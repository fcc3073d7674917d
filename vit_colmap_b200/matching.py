"""Host-side mirror of the reference's matching interface, running on the sm_100a kernels.

Reference boundary: ``pycolmap.match_exhaustive(database_path, matching_options, ...)``
(PYCOLMAP/pipeline/match_features.cc:28-51,210-223) as called by
``vit_colmap/pipeline/run_pipeline.py:351-363``; matcher semantics COLMAP/feature/sift.cc:733-825.
PyTorch is used only for device memory and streams; every computation is a call into the C ABI
(include/vc_b200.h).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import _lib
from .pairing import ExhaustivePairGenerator

SQ_SIFT_NORM = 512.0 * 512.0  # COLMAP/feature/sift.cc:63


@dataclass
class SiftMatchingOptions:
    """COLMAP/feature/sift.h:108-126 (fields used on this path)."""
    max_ratio: float = 0.8
    max_distance: float = 0.7
    cross_check: bool = True


@dataclass
class FeatureMatchingOptions:
    """COLMAP/feature/matcher.h:53-91 (fields used on this path) with the nested SIFT options
    flattened the way vit_colmap/utils/config.py:64-96 fills them."""
    use_gpu: bool = True
    gpu_index: str = "-1"
    num_threads: int = -1
    max_num_matches: int = 32768
    guided_matching: bool = False
    max_ratio: float = 0.8
    max_distance: float = 0.7
    cross_check: bool = True
    # COLMAP/controllers/feature_matching_utils.cc:485-488 blanks pairs with fewer matches
    min_num_matches: int = 15
    # "angular": COLMAP's GPU / brute-force rule (sift.cc:733-787); "l2": its CPU index rule done
    # exactly (sift.cc:827-874) -- what the PyPI pycolmap (no CUDA) applies, and the only one under
    # which uniform-random uint8 descriptors (the `dummy` extractor) produce matches
    rule: str = "angular"


def _ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


class Matcher:
    """One per GPU (COLMAP creates one SiftMatchGPU per device,
    controllers/feature_matching_utils.cc:66-84)."""

    def __init__(self, device: int = 0):
        import torch
        if not torch.cuda.is_available():
            raise _lib.VcError("no CUDA device: vit_colmap_b200 has no CPU fallback")
        self.torch = torch
        self.device = int(device)
        self.lib = _lib.load()
        h = ctypes.c_void_p()
        _lib.check(self.lib.vc_create(self.device, ctypes.byref(h)))
        self.handle = h
        self._ws = None

    def close(self):
        if getattr(self, "handle", None):
            self.lib.vc_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _dev(self):
        return self.torch.device("cuda", self.device)

    def _stream(self) -> int:
        return int(self.torch.cuda.current_stream(self._dev()).cuda_stream)

    def _workspace(self, n_pairs: int, kmax: int):
        need = int(self.lib.vc_match_workspace_bytes(n_pairs, kmax))
        if self._ws is None or self._ws.numel() < need:
            self._ws = self.torch.empty(need, dtype=self.torch.uint8, device=self._dev())
        return self._ws, need

    @staticmethod
    def _opts(max_ratio, max_distance, cross_check, inv_norm, rule: int = 0, l2_offset: float = 0.0) -> _lib.MatchOptions:
        return _lib.MatchOptions(float(max_ratio), float(max_distance), int(bool(cross_check)),
                                 float(inv_norm), int(rule), float(l2_offset))

    def to_bf16(self, desc):
        """uint8 / float32 device tensor -> bf16 device tensor via the library's converters."""
        torch = self.torch
        desc = desc.contiguous()
        out = torch.empty(desc.shape, dtype=torch.bfloat16, device=desc.device)
        if desc.dtype == torch.uint8:
            _lib.check(self.lib.vc_convert_u8_to_bf16(_ptr(desc), _ptr(out), desc.numel(), self._stream()))
        elif desc.dtype == torch.float32:
            _lib.check(self.lib.vc_convert_f32_to_bf16(_ptr(desc), _ptr(out), desc.numel(), self._stream()))
        elif desc.dtype == torch.bfloat16:
            return desc
        else:
            raise TypeError(f"unsupported descriptor dtype {desc.dtype}")
        return out

    # ------------------------------------------------------------------ device-resident path
    def match_pairs(self, desc, counts, pairs, max_ratio=0.8, max_distance=0.7, cross_check=True,
                    inv_norm=1.0, out=None):
        """desc bf16 [n_images,kmax,d], counts int32 [n_images], pairs int32 [n_pairs,2]; all on
        this device.  Returns (matches int32-viewed-uint32 [n_pairs,kmax,2], n_matches int32)."""
        torch = self.torch
        assert desc.dtype == torch.bfloat16 and desc.is_contiguous() and desc.dim() == 3
        n_images, kmax, d = desc.shape
        n_pairs = int(pairs.shape[0])
        if out is None:
            matches = torch.empty((n_pairs, kmax, 2), dtype=torch.int32, device=desc.device)
            n_matches = torch.empty((n_pairs,), dtype=torch.int32, device=desc.device)
        else:
            matches, n_matches = out
        ws, need = self._workspace(n_pairs, kmax)
        opt = self._opts(max_ratio, max_distance, cross_check, inv_norm)
        _lib.check(self.lib.vc_match_pairs(self.handle, _ptr(desc), _ptr(counts), n_images, kmax, d,
                                           _ptr(pairs), n_pairs, ctypes.byref(opt), _ptr(matches),
                                           _ptr(n_matches), _ptr(ws), need, self._stream()))
        return matches, n_matches

    def top2(self, desc, counts, pairs, max_ratio=0.8, max_distance=0.7, cross_check=True, inv_norm=1.0):
        """Diagnostics: (best, second, index) [n_pairs,2,kmax] of the same launch."""
        torch = self.torch
        n_images, kmax, d = desc.shape
        n_pairs = int(pairs.shape[0])
        best = torch.zeros((n_pairs, 2, kmax), dtype=torch.float32, device=desc.device)
        second = torch.zeros_like(best)
        index = torch.full((n_pairs, 2, kmax), -1, dtype=torch.int32, device=desc.device)
        ws, need = self._workspace(n_pairs, kmax)
        opt = self._opts(max_ratio, max_distance, cross_check, inv_norm)
        _lib.check(self.lib.vc_match_pairs_top2(self.handle, _ptr(desc), _ptr(counts), n_images, kmax, d,
                                                _ptr(pairs), n_pairs, ctypes.byref(opt), _ptr(best),
                                                _ptr(second), _ptr(index), _ptr(ws), need,
                                                self._stream()))
        return best, second, index

    def debug_scores(self, desc, counts, image1: int, image2: int):
        torch = self.torch
        n_images, kmax, d = desc.shape
        scores = torch.zeros((kmax, kmax), dtype=torch.float32, device=desc.device)
        _lib.check(self.lib.vc_debug_scores(self.handle, _ptr(desc), _ptr(counts), n_images, kmax, d,
                                            int(image1), int(image2), _ptr(scores), kmax, self._stream()))
        return scores

    # ------------------------------------------------------------------ host-buffer path
    def match_pairs_host(self, desc: np.ndarray, counts: np.ndarray, pairs: np.ndarray, max_ratio=0.8,
                         max_distance=0.7, cross_check=True, inv_norm: Optional[float] = None,
                         rule: str = "angular"):
        """desc: host uint8 or float32 [n_images,kmax,d].  Returns a list of [n,2] uint32 arrays.
        rule "angular": COLMAP's brute-force / GPU semantics (sift.cc:733-787); rule "l2": the exact
        form of its CPU index path (sift.cc:827-874), uint8 descriptors only."""
        desc = np.ascontiguousarray(desc)
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        n_images, kmax, d = desc.shape
        n_pairs = pairs.shape[0]
        if rule == "l2":
            return self._match_pairs_host_l2(desc, counts, pairs, max_ratio, max_distance, cross_check)
        if rule != "angular":
            raise ValueError(f"unknown matching rule {rule!r} (angular | l2)")
        if desc.dtype == np.uint8:
            dtype = _lib.DTYPE_U8
            inv = 1.0 / SQ_SIFT_NORM if inv_norm is None else inv_norm
        elif desc.dtype == np.float32:
            dtype = _lib.DTYPE_F32
            inv = 1.0 if inv_norm is None else inv_norm
        elif desc.dtype == np.uint16:  # raw bf16 bit patterns
            dtype = _lib.DTYPE_BF16
            inv = 1.0 if inv_norm is None else inv_norm
        else:
            raise TypeError("descriptors must be uint8, float32 or uint16 (bf16 bits)")
        matches = np.empty((max(n_pairs, 1), kmax, 2), dtype=np.uint32)
        n_matches = np.zeros((max(n_pairs, 1),), dtype=np.int32)
        opt = self._opts(max_ratio, max_distance, cross_check, inv)
        _lib.check(self.lib.vc_match_pairs_host(self.handle, desc.ctypes.data, dtype, counts.ctypes.data, n_images, kmax,
                                                d, pairs.ctypes.data, n_pairs, ctypes.byref(opt),
                                                matches.ctypes.data, n_matches.ctypes.data))
        return [matches[i, :n_matches[i]].copy() for i in range(n_pairs)]


    def _match_pairs_host_l2(self, desc, counts, pairs, max_ratio, max_distance, cross_check):
        """Squared-L2 nearest neighbours through the SAME fused GEMM kernel: the descriptors are
        augmented so that the dot product of an image-1 row with an image-2 row is exactly
        K - |x - y|^2.  The two roles need different augmentations (K - d^2 is not a Gram matrix), so
        the image axis is doubled: copies [0, n) play image 1, copies [n, 2n) image 2.
            image 1: [2x, D0, D1, D2, 1, 1, 1, 1]             (|x|^2 = D0 + 256 D1 + 65536 D2)
            image 2: [ y, -1, -256, -65536, -E0, -256 E1, -65536 E2, K]   (|y|^2 = E0 + 256 E1 + 65536 E2)
        Every entry is an integer bf16 represents exactly and every partial sum stays below 2^24
        (K = power of two > 2 max|x|^2 >= max d^2, required <= 2^23), so fp32 accumulation is exact
        and ties resolve as in COLMAP (first minimum wins)."""
        if desc.dtype != np.uint8:
            raise TypeError("rule 'l2' is defined for uint8 descriptors (COLMAP FeatureDescriptors)")
        n_images, kmax, d = desc.shape
        x = desc.astype(np.int64)
        norm2 = (x * x).sum(axis=2)
        mask = np.arange(kmax)[None, :] < counts[:, None]
        max_norm2 = int((norm2 * mask).max()) if mask.any() else 0
        K = 1 << max(int(2 * max_norm2 + 1).bit_length(), 1)
        if K > (1 << 23):
            raise ValueError("descriptor norms too large for the exact l2 rule (need 2*max|x|^2 < 2^23)")
        dig = np.stack([norm2 & 255, (norm2 >> 8) & 255, norm2 >> 16], axis=2).astype(np.float32)
        aug = np.zeros((2 * n_images, kmax, d + 8), dtype=np.float32)
        aug[:n_images, :, :d] = 2.0 * desc
        aug[:n_images, :, d:d + 3] = dig
        aug[:n_images, :, d + 3:d + 7] = 1.0
        aug[n_images:, :, :d] = desc
        aug[n_images:, :, d:d + 3] = np.array([-1.0, -256.0, -65536.0], np.float32)
        aug[n_images:, :, d + 3:d + 6] = -dig * np.array([1.0, 256.0, 65536.0], np.float32)
        aug[n_images:, :, d + 6] = float(K)
        import torch
        bits = torch.from_numpy(aug).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
        assert np.array_equal(torch.from_numpy(bits.view(np.int16)).view(torch.bfloat16).float().numpy(), aug)
        counts2 = np.concatenate([counts, counts]).astype(np.int32)
        pairs2 = np.stack([pairs[:, 0], pairs[:, 1] + n_images], axis=1).astype(np.int32)
        n_pairs = pairs2.shape[0]
        matches = np.empty((max(n_pairs, 1), kmax, 2), dtype=np.uint32)
        n_matches = np.zeros((max(n_pairs, 1),), dtype=np.int32)
        opt = self._opts(max_ratio, max_distance, cross_check, 1.0 / SQ_SIFT_NORM, 1, float(K))
        _lib.check(self.lib.vc_match_pairs_host(self.handle, bits.ctypes.data, _lib.DTYPE_BF16, counts2.ctypes.data,
                                                2 * n_images, kmax, d + 8, pairs2.ctypes.data, n_pairs,
                                                ctypes.byref(opt), matches.ctypes.data, n_matches.ctypes.data))
        return [matches[i, :n_matches[i]].copy() for i in range(n_pairs)]

def match_descriptor_sets(descriptors: Sequence[np.ndarray], pairs, options: Optional[FeatureMatchingOptions] = None,
                          matcher: Optional[Matcher] = None, device: int = 0):
    """Match host descriptor arrays (ragged list of [n_i, d] uint8 or float32) for the given index
    pairs.  Pads to a common kmax, calls the host-buffer C entry, returns list of [n,2] uint32."""
    options = options or FeatureMatchingOptions()
    matcher = matcher or Matcher(device)
    n_images = len(descriptors)
    d = max((int(x.shape[1]) for x in descriptors if x.ndim == 2 and x.shape[0] > 0), default=128)
    dtype = np.uint8 if all(x.dtype == np.uint8 for x in descriptors) else np.float32
    kmax = max(max((int(x.shape[0]) for x in descriptors), default=0), 1)
    kmax = min(kmax, options.max_num_matches)  # SIFTGPU/SiftMatchCU.cpp:108 clamps inputs likewise
    buf = np.zeros((n_images, kmax, d), dtype=dtype)
    counts = np.zeros((n_images,), dtype=np.int32)
    for i, x in enumerate(descriptors):
        n = min(int(x.shape[0]), kmax)
        counts[i] = n
        if n:
            buf[i, :n] = x[:n]
    rule = getattr(options, "rule", "angular")
    if rule == "l2" and dtype != np.uint8:
        raise TypeError("rule 'l2' is defined for uint8 descriptors")
    return matcher.match_pairs_host(buf, counts, np.asarray(pairs, dtype=np.int32),
                                    options.max_ratio, options.max_distance, options.cross_check, rule=rule)


def match_exhaustive(database_path, matching_options: Optional[FeatureMatchingOptions] = None,
                     pairing_options=None, verification_options=None, device=None, *,
                     matcher: Optional[Matcher] = None, rank: int = 0, world_size: int = 1,
                     pairs_per_launch: int = 256):
    """Drop-in for ``pycolmap.match_exhaustive(database_path, matching_options, ...)`` restricted to
    the matching half (PYCOLMAP/pipeline/match_features.cc:28-51; geometric verification stays with
    pycolmap and is invoked afterwards when pycolmap is importable).

    Reads ``descriptors`` rows, generates exhaustive pairs block-wise (pairing.cc:202-236), matches
    every block on the GPU and writes ``matches`` rows: pairs with fewer than ``min_num_matches``
    matches are written empty (controllers/feature_matching_utils.cc:485-488), pairs that already
    have a row are skipped (:443-449).  With world_size > 1 each rank matches its share of the pair
    blocks (no collective); only rank 0's caller should write -- use ``return_only=True`` style via
    the returned dict to gather.  Returns {(image_id1, image_id2): matches[n,2] uint32}.
    """
    from .database import Database
    options = matching_options or FeatureMatchingOptions()
    block_size = int(getattr(pairing_options, "block_size", 50)) if pairing_options is not None else 50
    if isinstance(device, int):
        dev_index = device
    else:
        dev_index = 0
    matcher = matcher or Matcher(dev_index)
    results = {}
    with Database.open(str(database_path)) as db:
        images = db.read_all_images()
        ids = [im.image_id for im in images]
        desc = {i: db.read_descriptors(i) for i in ids}
        desc = {i: (v if v.ndim == 2 and v.shape[0] > 0 else np.zeros((0, 128), np.uint8)) for i, v in desc.items()}
        index_of = {i: k for k, i in enumerate(ids)}
        gen = ExhaustivePairGenerator(ids, block_size)
        todo = []
        for pair in gen.all_pairs():
            if db.exists_matches(pair[0], pair[1]):  # resume semantics
                continue
            todo.append(pair)
        lo = len(todo) * rank // world_size
        hi = len(todo) * (rank + 1) // world_size
        mine = todo[lo:hi]
        arrays = [desc[i] for i in ids]
        for s in range(0, len(mine), pairs_per_launch):
            chunk = mine[s:s + pairs_per_launch]
            idx_pairs = np.array([(index_of[a], index_of[b]) for a, b in chunk], dtype=np.int32)
            out = match_descriptor_sets(arrays, idx_pairs, options, matcher)
            for (a, b), m in zip(chunk, out):
                if m.shape[0] < options.min_num_matches:
                    m = np.zeros((0, 2), np.uint32)
                results[(a, b)] = m
        if rank == 0 and world_size == 1:
            with db.transaction():
                for (a, b), m in results.items():
                    db.write_matches(a, b, m)
    return results
