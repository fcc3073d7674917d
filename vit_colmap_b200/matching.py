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
from typing import List, Optional, Sequence

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
        self._host_staging: list = []   # page-locked staging of pad_descriptors, reused across uploads
        self._out_cache = None          # host result slots of match_resident, reused across calls

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
            resident = self.upload(desc, counts, rule="l2")
            return self.match_resident(resident, pairs, max_ratio, max_distance, cross_check)
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

    # ------------------------------------------------------------------ upload once, match many
    def upload(self, desc, counts, rule: str = "angular", inv_norm: Optional[float] = None) -> "ResidentDescriptors":
        """Make a padded descriptor set resident in HBM as the matcher's bf16 operand -- ONCE; every
        later pair chunk only sends its pair list (vc_match_pairs_resident).  `desc`: host ndarray or
        torch tensor (host or this device), uint8 / float32 / bf16, [n_images, kmax, d]; `counts`: host
        int32 [n_images].  COLMAP's GPU matcher re-uploads descriptors per pair instead
        (COLMAP/feature/sift.cc:1270-1284).

        rule "l2" (uint8 only): the rows are augmented on the device (vc_l2_augment) so that the dot
        product of image-1 copy i with image-2 copy n+j is K - |x - y|^2 exactly."""
        torch = self.torch
        if rule not in ("angular", "l2"):
            raise ValueError(f"unknown matching rule {rule!r} (angular | l2)")
        if isinstance(desc, np.ndarray):
            if desc.dtype == np.uint16:  # raw bf16 bit patterns
                t = torch.from_numpy(np.ascontiguousarray(desc).view(np.int16)).view(torch.bfloat16)
            else:
                t = torch.from_numpy(np.ascontiguousarray(desc))
        else:
            t = desc
        if t.dim() != 3:
            raise ValueError("descriptors must be [n_images, kmax, d]")
        if t.dtype not in (torch.uint8, torch.float32, torch.bfloat16):
            raise TypeError("descriptors must be uint8, float32 or bf16")
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        n_images, kmax, d = (int(v) for v in t.shape)
        if counts.shape != (n_images,):
            raise ValueError("counts must be int32 [n_images]")
        is_u8 = t.dtype == torch.uint8
        if t.device.type != "cuda":
            if not t.is_pinned() and t.numel() >= (1 << 20):
                pinned = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
                pinned.copy_(t)
                t = pinned
            t = t.to(self._dev(), non_blocking=True)
        t = t.contiguous()
        if rule == "l2":
            if not is_u8:
                raise TypeError("rule 'l2' is defined for uint8 descriptors (COLMAP FeatureDescriptors)")
            out = torch.empty((2 * n_images, kmax, d + 8), dtype=torch.bfloat16, device=t.device)
            k = ctypes.c_float(0.0)
            _lib.check(self.lib.vc_l2_augment(self.handle, _ptr(t), n_images, kmax, d, _ptr(out), ctypes.byref(k),
                                              self._stream()))
            res = ResidentDescriptors(out, np.concatenate([counts, counts]), n_images, 1.0 / SQ_SIFT_NORM, 1,
                                      float(k.value), n_images)
        else:
            inv = inv_norm if inv_norm is not None else (1.0 / SQ_SIFT_NORM if is_u8 else 1.0)
            res = ResidentDescriptors(self.to_bf16(t), counts, n_images, inv, 0, 0.0, 0)
        torch.cuda.current_stream(self._dev()).synchronize()  # the C entry runs on the handle's own streams
        return res

    def match_resident(self, resident: "ResidentDescriptors", pairs, max_ratio=0.8, max_distance=0.7,
                       cross_check=True, pairs_per_call: int = 1024) -> List[np.ndarray]:
        """Match index pairs [n, 2] (into the uploaded image axis) against resident descriptors.
        Only the pair list goes up and only the packed match lists come back."""
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        if pairs.size and (pairs.min() < 0 or pairs.max() >= resident.n_images):
            raise ValueError("pair index outside the uploaded image range")
        desc = resident.desc
        n_dev, kmax, d = (int(v) for v in desc.shape)
        opt = self._opts(max_ratio, max_distance, cross_check, resident.inv_norm, resident.rule, resident.l2_offset)
        out: List[np.ndarray] = []
        for s in range(0, pairs.shape[0], pairs_per_call):
            chunk = pairs[s:s + pairs_per_call].copy()
            chunk[:, 1] += resident.second_offset
            n = chunk.shape[0]
            if self._out_cache is None or self._out_cache.size < n * kmax * 2:
                self._out_cache = np.empty((n * kmax * 2,), dtype=np.uint32)  # reused: fresh pages fault on first touch
            matches = self._out_cache[: n * kmax * 2].reshape(n, kmax, 2)
            n_matches = np.zeros((n,), dtype=np.int32)
            _lib.check(self.lib.vc_match_pairs_resident(self.handle, _ptr(desc), resident.counts.ctypes.data, n_dev,
                                                        kmax, d, chunk.ctypes.data, n, ctypes.byref(opt),
                                                        matches.ctypes.data, n_matches.ctypes.data))
            out.extend(matches[i, :n_matches[i]].copy() for i in range(n))
        return out


    # ------------------------------------------------------------------ guided matching
    def match_guided(self, resident: "ResidentDescriptors", keypoints, pairs, geometries: Sequence["PairGeometry"],
                     max_ratio=0.8, max_distance=0.7, cross_check=True) -> List[np.ndarray]:
        """Guided matching of index pairs against resident descriptors.

        keypoints: [n_images, kmax, 2] float32 (x, y) per descriptor row (host ndarray or device tensor);
        geometries: one PairGeometry per pair.  Semantics: SiftGPUFeatureMatcher::MatchGuided
        (COLMAP/feature/sift.cc:1308-1413) under the angular rule, SiftCPUFeatureMatcher::MatchGuided
        (:1058-1157) under rule l2 -- see vc_match_pairs_guided in include/vc_b200.h."""
        torch = self.torch
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        n_pairs = pairs.shape[0]
        if len(geometries) != n_pairs:
            raise ValueError("one geometry per pair")
        if n_pairs == 0:
            return []
        if pairs.min() < 0 or pairs.max() >= resident.n_images:
            raise ValueError("pair index outside the uploaded image range")
        desc = resident.desc
        n_dev, kmax, d = (int(v) for v in desc.shape)
        kp = keypoints if not isinstance(keypoints, np.ndarray) else torch.from_numpy(
            np.ascontiguousarray(keypoints, dtype=np.float32))
        kp = kp.to(device=desc.device, dtype=torch.float32).contiguous()
        if tuple(kp.shape) != (resident.n_images, kmax, 2):
            raise ValueError(f"keypoints must be [{resident.n_images}, {kmax}, 2]")
        geo = (_lib.PairGeometry * n_pairs)()
        for k, g in enumerate(geometries):
            geo[k] = g.to_c(resident.rule)
        geo_dev = torch.frombuffer(bytearray(bytes(geo)), dtype=torch.uint8).to(desc.device)
        chunk = pairs.copy()
        chunk[:, 1] += resident.second_offset
        pairs_dev = torch.from_numpy(chunk).to(desc.device)
        counts_dev = torch.from_numpy(resident.counts).to(desc.device)
        matches = torch.empty((n_pairs, kmax, 2), dtype=torch.int32, device=desc.device)
        n_matches = torch.empty((n_pairs,), dtype=torch.int32, device=desc.device)
        ws, need = self._workspace(n_pairs, kmax)
        opt = self._opts(max_ratio, max_distance, cross_check, resident.inv_norm, resident.rule, resident.l2_offset)
        _lib.check(self.lib.vc_match_pairs_guided(self.handle, _ptr(desc), _ptr(kp), _ptr(counts_dev), n_dev, kmax, d,
                                                  _ptr(pairs_dev), _ptr(geo_dev), n_pairs, ctypes.byref(opt),
                                                  _ptr(matches), _ptr(n_matches), _ptr(ws), need, self._stream()))
        n_h = n_matches.cpu().numpy()
        top = int(n_h.max()) if n_pairs else 0
        # one strided device-to-host copy of the filled prefixes [n_pairs, max count, 2] instead of one per pair
        m_h = matches[:, :top].cpu().numpy().view(np.uint32) if top else np.zeros((n_pairs, 0, 2), np.uint32)
        return [m_h[k, : int(n_h[k])].copy() for k in range(n_pairs)]


# TwoViewGeometry::ConfigurationType (COLMAP/scene/two_view_geometry.h:40-61)
CONFIG_UNDEFINED, CONFIG_DEGENERATE, CONFIG_CALIBRATED, CONFIG_UNCALIBRATED = 0, 1, 2, 3
CONFIG_PLANAR, CONFIG_PANORAMIC, CONFIG_PLANAR_OR_PANORAMIC, CONFIG_CALIBRATED_RIG = 4, 5, 6, 9
_F_CONFIGS = (CONFIG_CALIBRATED, CONFIG_CALIBRATED_RIG, CONFIG_UNCALIBRATED)
_H_CONFIGS = (CONFIG_PLANAR, CONFIG_PANORAMIC, CONFIG_PLANAR_OR_PANORAMIC)


@dataclass
class PairGeometry:
    """Two-view geometry of one image pair as guided matching consumes it
    (TwoViewGeometry::{config, F, H}, COLMAP/scene/two_view_geometry.h:63-73).  F / H: 3x3, H maps
    image 1 to image 2, F satisfies p2^T F p1 = 0; max_error in pixels (RANSAC max_error, default 4)."""
    config: int
    F: Optional[np.ndarray] = None
    H: Optional[np.ndarray] = None
    max_error: float = 4.0

    def usable(self) -> bool:
        return self.config in _F_CONFIGS or self.config in _H_CONFIGS

    def to_c(self, rule: int) -> "_lib.PairGeometry":
        if not self.usable():
            raise ValueError(f"two-view configuration {self.config} has no geometry to guide with (sift.cc:1122-1124)")
        use_f = self.config in _F_CONFIGS
        eye = np.eye(3, dtype=np.float32)
        big = 1.0e20  # SIFTGPU/SiftMatch.cpp:684: the absent model becomes identity with an unreachable bound
        m = np.asarray(self.F if use_f else self.H, dtype=np.float64).astype(np.float32).reshape(3, 3)
        res = np.float32(float(self.max_error) * float(self.max_error))  # sift.cc:1091,1394: double product -> float
        g = _lib.PairGeometry()
        g.h[:] = (eye if use_f else m).reshape(-1).tolist()
        g.f[:] = (m if use_f else eye).reshape(-1).tolist()
        if rule == 0:   # GPU path: both tests are evaluated
            g.mode, g.max_h, g.max_f = _lib.GUIDED_SIFTGPU, (big if use_f else float(res)), (float(res) if use_f else big)
        else:           # CPU path: one test, '>' filters
            g.mode, g.max_h, g.max_f = (_lib.GUIDED_F if use_f else _lib.GUIDED_H), float(res), float(res)
        return g


@dataclass
class ResidentDescriptors:
    """Descriptors of a set of images as the matcher reads them: bf16 [n_dev, kmax, d] in HBM."""
    desc: object               # torch bf16 tensor on the matcher's device
    counts: np.ndarray         # host int32 [n_dev]
    n_images: int              # images the caller uploaded (n_dev = 2 * n_images under rule l2)
    inv_norm: float
    rule: int                  # vc_match_options.rule
    l2_offset: float
    second_offset: int         # added to the image-2 index of every pair (rule l2: its role-2 copy)

    @property
    def nbytes(self) -> int:
        return int(self.desc.numel()) * 2


def pad_descriptors(descriptors: Sequence[np.ndarray], max_rows: int, pinned: bool = True, staging=None):
    """Ragged list of [n_i, d] uint8 / float32 arrays -> (padded [n, kmax, d] torch host tensor, counts).
    The rows are written once, directly into page-locked memory, so the upload is one DMA.  `staging`
    (a one-element list holding a flat page-locked uint8 tensor, grown on demand) lets a caller reuse
    the allocation across calls: allocating and first-touching page-locked memory costs more than the
    copy itself.  Only the padding rows are zeroed (the kernels ignore rows >= counts[i]; the l2
    augmentation takes its largest norm over every row, so padding must not be garbage)."""
    import torch
    n_images = len(descriptors)
    d_in = max((int(x.shape[1]) for x in descriptors if x.ndim == 2 and x.shape[0] > 0), default=128)
    d = -(-d_in // 8) * 8  # the kernels take rows of a multiple of 8 columns: zero columns leave every dot product unchanged
    is_u8 = all(x.dtype == np.uint8 for x in descriptors)
    kmax = max(max((int(x.shape[0]) for x in descriptors if x.ndim == 2), default=0), 1)
    kmax = min(kmax, max_rows)  # SIFTGPU/SiftMatchCU.cpp:108 clamps inputs likewise
    dtype = torch.uint8 if is_u8 else torch.float32
    nbytes = n_images * kmax * d * (1 if is_u8 else 4)
    use_pin = pinned and torch.cuda.is_available()
    if staging is not None:
        if not staging or staging[0].numel() < nbytes:
            staging[:] = [torch.empty((max(nbytes, 1 << 20),), dtype=torch.uint8, pin_memory=use_pin)]
        buf = staging[0][:nbytes].view(dtype).view(n_images, kmax, d)
    else:
        buf = torch.empty((n_images, kmax, d), dtype=dtype, pin_memory=use_pin)
    view = buf.numpy()
    counts = np.zeros((n_images,), dtype=np.int32)
    for i, x in enumerate(descriptors):
        n = min(int(x.shape[0]), kmax) if x.ndim == 2 else 0
        counts[i] = n
        if n:
            view[i, :n, : x.shape[1]] = x[:n]
            if x.shape[1] < d:
                view[i, :n, x.shape[1]:] = 0
        if n < kmax:
            view[i, n:] = 0
    return buf, counts


def match_descriptor_sets(descriptors: Sequence[np.ndarray], pairs, options: Optional[FeatureMatchingOptions] = None,
                          matcher: Optional[Matcher] = None, device: int = 0):
    """Match host descriptor arrays (ragged list of [n_i, d] uint8 or float32) for the given index
    pairs: pad + upload ONCE, then match every pair on the resident set.  Returns list of [n,2] uint32."""
    options = options or FeatureMatchingOptions()
    matcher = matcher or Matcher(device)
    buf, counts = pad_descriptors(descriptors, options.max_num_matches, staging=matcher._host_staging)
    rule = getattr(options, "rule", "angular")
    resident = matcher.upload(buf, counts, rule=rule)
    return matcher.match_resident(resident, np.asarray(pairs, dtype=np.int32), options.max_ratio,
                                  options.max_distance, options.cross_check)


def match_exhaustive(database_path, matching_options: Optional[FeatureMatchingOptions] = None,
                     pairing_options=None, verification_options=None, device=None, *,
                     matcher: Optional[Matcher] = None, rank: int = 0, world_size: int = 1,
                     pairs_per_launch: int = 1024, max_resident_bytes: Optional[int] = None,
                     stats: Optional[dict] = None):
    """Drop-in for ``pycolmap.match_exhaustive(database_path, matching_options, ...)`` restricted to
    the matching half (PYCOLMAP/pipeline/match_features.cc:28-51; geometric verification stays with
    pycolmap and is invoked afterwards when pycolmap is importable).

    Reads ``descriptors`` rows, uploads them ONCE (COLMAP's matcher cache + per-pair upload,
    COLMAP/feature/matcher.cc:71-151 and sift.cc:1270-1284, become one resident bf16 tensor), generates
    exhaustive pairs block-wise (pairing.cc:202-236), matches every block on the GPU and writes
    ``matches`` rows: pairs with fewer than ``min_num_matches`` matches are written empty
    (controllers/feature_matching_utils.cc:485-488), pairs that already have a row are skipped
    (:443-449).  If the padded set exceeds ``max_resident_bytes`` (default: 60 % of the free device
    memory) the upload happens per pairing block instead (the <= 2 x block_size images a block names),
    still once per block rather than once per pair chunk.  With world_size > 1 each rank matches its
    share of the pair blocks (no collective); only a single-process call writes.
    ``stats`` (optional dict) receives the wall-clock seconds of the four stages: read_s (sqlite3 reads +
    pair generation), upload_s (pad into page-locked memory, H2D, conversion), match_s, write_s.
    Returns {(image_id1, image_id2): matches[n,2] uint32}.
    """
    import time
    from .database import Database
    clock = {"read_s": 0.0, "upload_s": 0.0, "match_s": 0.0, "write_s": 0.0}
    t_start = time.perf_counter()
    options = matching_options or FeatureMatchingOptions()
    block_size = int(getattr(pairing_options, "block_size", 50)) if pairing_options is not None else 50
    dev_index = device if isinstance(device, int) else 0
    matcher = matcher or Matcher(dev_index)
    rule = getattr(options, "rule", "angular")
    results = {}
    with Database.open(str(database_path)) as db:
        images = db.read_all_images()
        ids = [im.image_id for im in images]
        desc = {i: db.read_descriptors(i) for i in ids}
        desc = {i: (v if v.ndim == 2 and v.shape[0] > 0 else np.zeros((0, 128), np.uint8)) for i, v in desc.items()}
        gen = ExhaustivePairGenerator(ids, block_size)
        blocks = [[p for p in blk if not db.exists_matches(p[0], p[1])] for blk in gen.blocks()]  # resume semantics
        n_todo = sum(len(b) for b in blocks)
        lo, hi = n_todo * rank // world_size, n_todo * (rank + 1) // world_size
        mine, seen = [], 0
        for blk in blocks:  # this rank's contiguous share of the pair list, still grouped by block
            part = blk[max(lo - seen, 0):max(hi - seen, 0)]
            seen += len(blk)
            if part:
                mine.append(part)

        clock["read_s"] = time.perf_counter() - t_start

        def run(block_list, image_ids):
            t0 = time.perf_counter()
            index_of = {i: k for k, i in enumerate(image_ids)}
            buf, counts = pad_descriptors([desc[i] for i in image_ids], options.max_num_matches,
                                          staging=matcher._host_staging)
            resident = matcher.upload(buf, counts, rule=rule)
            clock["upload_s"] += time.perf_counter() - t0
            t0 = time.perf_counter()
            for blk in block_list:
                idx = np.array([(index_of[a], index_of[b]) for a, b in blk], dtype=np.int32)
                out = matcher.match_resident(resident, idx, options.max_ratio, options.max_distance,
                                             options.cross_check, pairs_per_launch)
                for (a, b), m in zip(blk, out):
                    results[(a, b)] = m if m.shape[0] >= options.min_num_matches else np.zeros((0, 2), np.uint32)
            clock["match_s"] += time.perf_counter() - t0

        if mine:
            kmax = min(max(max((int(v.shape[0]) for v in desc.values()), default=1), 1), options.max_num_matches)
            dmax = max((int(v.shape[1]) for v in desc.values()), default=128)
            need = len(ids) * kmax * (dmax + 8) * 2 * (2 if rule == "l2" else 1) + len(ids) * kmax * dmax
            if max_resident_bytes is None:
                free, _total = matcher.torch.cuda.mem_get_info(matcher._dev())
                max_resident_bytes = int(free * 0.6)
            if need <= max_resident_bytes:
                run(mine, ids)
            else:
                for blk in mine:
                    used = sorted({i for pr in blk for i in pr}, key=ids.index)
                    run([blk], used)
        t0 = time.perf_counter()
        if rank == 0 and world_size == 1:
            with db.transaction():
                for (a, b), m in results.items():
                    db.write_matches(a, b, m)
        clock["write_s"] = time.perf_counter() - t0
    if stats is not None:
        stats.update(clock)
    return results


def guided_matching(database_path, matching_options: Optional[FeatureMatchingOptions] = None, max_error: float = 4.0,
                    min_num_inliers: int = 15, device=None, *, matcher: Optional[Matcher] = None,
                    pairs_per_launch: int = 256):
    """Second matching pass with ``guided_matching=True`` (COLMAP/feature/matcher.h:76-77,
    controllers/feature_matching_utils.cc:129-146,300-330): every pair whose ``two_view_geometries`` row
    holds a usable model (written by pycolmap's geometric verification) is matched again with the
    model as a gate on the candidates, and the row's inlier matches are replaced by the result; a pair
    left with fewer than ``min_num_inliers`` gets an empty TwoViewGeometry
    (feature_matching_utils.cc:490-493).  Descriptors and keypoints are uploaded once.
    ``max_error``: TwoViewGeometryOptions.ransac_options.max_error (pixels).
    Returns {(image_id1, image_id2): inlier matches [n, 2] uint32}."""
    from .database import Database, TwoViewGeometry
    options = matching_options or FeatureMatchingOptions()
    dev_index = device if isinstance(device, int) else 0
    matcher = matcher or Matcher(dev_index)
    rule = getattr(options, "rule", "angular")
    out = {}
    with Database.open(str(database_path)) as db:
        todo = [(pr, g) for pr, g in db.read_two_view_geometries()
                if PairGeometry(g.config).usable() and db.exists_descriptors(pr[0]) and db.exists_descriptors(pr[1])]
        if not todo:
            return out
        ids = sorted({i for pr, _ in todo for i in pr})
        index_of = {i: k for k, i in enumerate(ids)}
        descs = [db.read_descriptors(i) for i in ids]
        buf, counts = pad_descriptors(descs, options.max_num_matches, staging=matcher._host_staging)
        kmax = int(buf.shape[1])
        kps = np.zeros((len(ids), kmax, 2), np.float32)
        for k, i in enumerate(ids):
            kp = db.read_keypoints(i)
            n = min(int(counts[k]), kp.shape[0])
            if n != int(counts[k]):
                raise ValueError(f"image {i}: {kp.shape[0]} keypoints for {counts[k]} descriptors (sift.cc:1069-1070)")
            kps[k, :n] = kp[:n, :2]
        resident = matcher.upload(buf, counts, rule=rule)
        kps_dev = matcher.torch.from_numpy(kps).to(matcher._dev())
        for s in range(0, len(todo), pairs_per_launch):
            part = todo[s:s + pairs_per_launch]
            idx = np.array([(index_of[a], index_of[b]) for (a, b), _ in part], dtype=np.int32)
            geos = [PairGeometry(g.config, g.F, g.H, max_error) for _, g in part]
            res = matcher.match_guided(resident, kps_dev, idx, geos, options.max_ratio, options.max_distance,
                                       options.cross_check)
            for (pr, g), m in zip(part, res):
                out[pr] = m
        with db.transaction():
            for pr, g in todo:
                m = out[pr]
                if m.shape[0] < min_num_inliers:
                    new = TwoViewGeometry()
                    out[pr] = new.inlier_matches
                else:
                    new = TwoViewGeometry(g.config, g.F, g.E, g.H, g.qvec, g.tvec, m)
                db.write_two_view_geometry(pr[0], pr[1], new, replace=True)
    return out
