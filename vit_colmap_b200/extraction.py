"""Host-side driver of the extraction post-processing kernels (score map, keypoint selection,
descriptor sampling / projection / packing).

Reference: ``ViTExtractor._dense_to_sparse`` (vit_colmap/features/vit_extractor.py:168-252).
PyTorch only allocates device memory and provides the stream; every computation is a call into
the C ABI (include/vc_b200.h).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

from . import _lib

METHODS = {"harris": 0, "dog": 1, "combined": 2}
NORM_ROOTSIFT = 0x10  # VC_NORM_ROOTSIFT, include/vc_b200.h
QUANT_VIT, QUANT_COLMAP, QUANT_TRAINABLE = 0, 1, 2


def _ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


@dataclass
class SparseFeatures:
    """Batched result; rows >= counts[i] of image i are padding."""
    counts: "torch.Tensor"      # int32 [B]
    coords: "torch.Tensor"      # int32 [B, cap, 2] (y, x) grid cells
    scores: "torch.Tensor"      # f32 [B, cap]
    keypoints: "torch.Tensor"   # f32 [B, cap, 2] (x, y) pixels
    desc_bf16: "torch.Tensor"   # bf16 [B, cap, d]
    desc_u8: "torch.Tensor"     # uint8 [B, cap, d]
    desc_f32: Optional["torch.Tensor"] = None


class PostProcessor:
    """One per GPU; shares the library handle type with the matcher."""

    def __init__(self, device: int = 0, handle=None):
        import torch
        if not torch.cuda.is_available():
            raise _lib.VcError("no CUDA device: vit_colmap_b200 has no CPU fallback")
        self.torch = torch
        self.device = int(device)
        self.lib = _lib.load()
        self._own = handle is None
        if handle is None:
            handle = ctypes.c_void_p()
            _lib.check(self.lib.vc_create(self.device, ctypes.byref(handle)))
        self.handle = handle
        self._ws = None

    def close(self):
        if self._own and getattr(self, "handle", None):
            self.lib.vc_destroy(self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _dev(self):
        return self.torch.device("cuda", self.device)

    def _stream(self) -> int:
        return int(self.torch.cuda.current_stream(self._dev()).cuda_stream)

    @staticmethod
    def _dtype_code(torch, t) -> int:
        if t.dtype == torch.float32:
            return _lib.DTYPE_F32
        if t.dtype == torch.bfloat16:
            return _lib.DTYPE_BF16
        raise TypeError(f"tokens must be float32 or bfloat16, got {t.dtype}")

    # ------------------------------------------------------------------ stages
    def score_map(self, tokens, grid_hw: Tuple[int, int], method: str = "harris"):
        """tokens [B, H*W, C] (token-major, last dim contiguous) -> score f32 [B, H, W]."""
        torch = self.torch
        H, W = grid_hw
        B, T, C = tokens.shape
        assert T == H * W and tokens.stride(2) == 1 and tokens.stride(1) == C
        need = int(self.lib.vc_score_workspace_bytes(B, H, W))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=tokens.device)
        score = torch.empty((B, H, W), dtype=torch.float32, device=tokens.device)
        _lib.check(self.lib.vc_score_map(self.handle, _ptr(tokens), self._dtype_code(torch, tokens), B, H, W, C,
                                         int(tokens.stride(0)), METHODS[method], _ptr(score), _ptr(self._ws),
                                         need, self._stream()))
        return score

    def select_keypoints(self, score, num_keypoints: int, bin_size: int = 16, nms_radius: float = 1.5):
        """score f32 [B, H, W] -> (coords int32 [B,cap,2], scores f32 [B,cap], counts int32 [B])."""
        torch = self.torch
        B, H, W = score.shape
        cap = min(int(num_keypoints), H * W)
        coords = torch.zeros((B, cap, 2), dtype=torch.int32, device=score.device)
        scores = torch.zeros((B, cap), dtype=torch.float32, device=score.device)
        counts = torch.zeros((B,), dtype=torch.int32, device=score.device)
        _lib.check(self.lib.vc_select_keypoints(self.handle, _ptr(score.contiguous()), B, H, W, int(num_keypoints),
                                                int(bin_size), float(nms_radius), _ptr(coords), _ptr(scores),
                                                _ptr(counts), cap, self._stream()))
        return coords, scores, counts

    def sample_pack(self, tokens, grid_hw, coords, counts, orig_wh, resized_wh, projection=None,
                    quant_mode: int = QUANT_VIT, want_f32: bool = False, projection_bias=None,
                    rootsift: bool = False):
        """rootsift: L1-normalise + sqrt before the L2 step (hybrid_extractor.py:286-288)."""
        torch = self.torch
        if rootsift:
            quant_mode = int(quant_mode) | NORM_ROOTSIFT
        H, W = grid_hw
        B, T, C = tokens.shape
        cap = int(coords.shape[1])
        d = C if projection is None else int(projection.shape[1])
        if projection is not None:
            projection = projection.to(device=tokens.device, dtype=torch.float32).contiguous()
        w_res, h_res = resized_wh
        w_orig, h_orig = orig_wh
        scale = (ctypes.c_float * 4)(w_res / W, w_orig / w_res, h_res / H, h_orig / h_res)
        # vc_sample_pack zero-fills the rows beyond counts[i] itself: no memset pass over the outputs
        kps = torch.empty((B, cap, 2), dtype=torch.float32, device=tokens.device)
        dbf = torch.empty((B, cap, d), dtype=torch.bfloat16, device=tokens.device)
        du8 = torch.empty((B, cap, d), dtype=torch.uint8, device=tokens.device)
        df32 = torch.empty((B, cap, d), dtype=torch.float32, device=tokens.device) if want_f32 else None
        is_float = coords.dtype == torch.float32
        _lib.check(self.lib.vc_sample_pack(self.handle, _ptr(tokens), self._dtype_code(torch, tokens), B, H, W, C,
                                           int(tokens.stride(0)), _ptr(coords.contiguous()), int(is_float),
                                           _ptr(counts), cap, _ptr(projection), _ptr(projection_bias), d, int(quant_mode),
                                           scale,
                                           _ptr(kps), _ptr(dbf), _ptr(du8), _ptr(df32), self._stream()))
        return kps, dbf, du8, df32

    # ------------------------------------------------------------------ whole stage
    def dense_to_sparse(self, tokens, grid_hw, orig_wh, resized_wh, num_keypoints=2048, method="harris",
                        projection=None, quant_mode: int = QUANT_VIT, want_f32: bool = False) -> SparseFeatures:
        score = self.score_map(tokens, grid_hw, method)
        coords, scores, counts = self.select_keypoints(score, num_keypoints)
        kps, dbf, du8, df32 = self.sample_pack(tokens, grid_hw, coords, counts, orig_wh, resized_wh,
                                               projection, quant_mode, want_f32)
        return SparseFeatures(counts, coords, scores, kps, dbf, du8, df32)


def trainable_select_pack(pp: PostProcessor, kp_map, desc_map, orig_wh, resized_wh, num_keypoints: int = 2048,
                          score_threshold: float = 0.0, nms_radius: int = 4, apply_tanh: bool = False,
                          want_f32: bool = False, projection=None, projection_bias=None):
    """Selection + packing stage of the `trainable_vit` extractor
    (vit_colmap/features/trainable_vit_extractor.py:173-269) on channel-last head outputs:
    kp_map f32 [B, h4, w4, 4], desc_map f32/bf16 [B, h4, w4, c] (un-normalised is fine: rows are
    L2-normalised with the model's eps 1e-8 as they are packed).  With ``projection`` f32 [c, d]
    (+ ``projection_bias`` [d]) desc_map is the descriptor head's hidden map and its 1x1 convolution
    (vit_feature_model.py:120) is applied only at the selected keypoints.
    Returns (keypoints f32 [B,cap,6], desc_u8 [B,cap,d], desc_bf16, desc_f32|None, counts int32 [B])."""
    torch = pp.torch
    B, h4, w4, _ = kp_map.shape
    c = int(desc_map.shape[3])
    d = c if projection is None else int(projection.shape[1])
    kp_map = kp_map.contiguous()
    desc_map = desc_map.contiguous()
    cap = min(int(num_keypoints), h4 * w4)
    need = int(pp.lib.vc_trainable_workspace_bytes(B, h4, w4))
    ws = torch.empty(need, dtype=torch.uint8, device=kp_map.device)
    coords = torch.zeros((B, cap, 2), dtype=torch.int32, device=kp_map.device)
    kps = torch.zeros((B, cap, 6), dtype=torch.float32, device=kp_map.device)
    counts = torch.zeros((B,), dtype=torch.int32, device=kp_map.device)
    w_orig, h_orig = orig_wh
    w_new, h_new = resized_wh
    scale = (ctypes.c_float * 4)(w_orig / w_new, h_orig / h_new, w_orig - 1, h_orig - 1)
    _lib.check(pp.lib.vc_trainable_select(pp.handle, _ptr(kp_map), B, h4, w4, int(num_keypoints),
                                          float(score_threshold), int(nms_radius), int(bool(apply_tanh)), scale,
                                          _ptr(coords), _ptr(kps), _ptr(counts), cap, _ptr(ws), need, pp._stream()))
    dbf = torch.empty((B, cap, d), dtype=torch.bfloat16, device=kp_map.device)  # tails zeroed by the kernel
    du8 = torch.empty((B, cap, d), dtype=torch.uint8, device=kp_map.device)
    df32 = torch.empty((B, cap, d), dtype=torch.float32, device=kp_map.device) if want_f32 else None
    tok = desc_map.reshape(B, h4 * w4, c)
    if projection is not None:
        projection = projection.to(device=kp_map.device, dtype=torch.float32).contiguous()
        if projection_bias is not None:
            projection_bias = projection_bias.to(device=kp_map.device, dtype=torch.float32).contiguous()
    _lib.check(pp.lib.vc_sample_pack(pp.handle, _ptr(tok), pp._dtype_code(torch, tok), B, h4, w4, c, int(tok.stride(0)),
                                     _ptr(coords), 0, _ptr(counts), cap, _ptr(projection), _ptr(projection_bias), d,
                                     QUANT_TRAINABLE, None, 0, _ptr(dbf), _ptr(du8), _ptr(df32), pp._stream()))
    return kps, du8, dbf, df32, counts
