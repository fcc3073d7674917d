"""Thin driver of the sampler entry points of the C ABI (include/vc_b200.h, "Training-time
similarity and hard-negative mining").  PyTorch allocates device memory and provides the stream;
every computation is a library call.  There is no CPU fallback."""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional

from .. import _lib


def _ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


@dataclass
class TokenMap:
    """A batch of feature maps in the backbone's token-major layout: tokens f32 [B, H*W, C] on the GPU."""
    tokens: "torch.Tensor"
    grid_hw: tuple

    @property
    def shape_nchw(self):
        B, _, C = self.tokens.shape
        return (B, C, self.grid_hw[0], self.grid_hw[1])


class SamplerDevice:
    """One per GPU (shares the library handle type with the matcher and the extractors)."""

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

    def _f32(self, t):
        return t.to(device=self._dev(), dtype=self.torch.float32).contiguous()

    # ------------------------------------------------------------------ layout
    def token_map(self, features) -> TokenMap:
        """(B, C, H, W) tensor (any device) or TokenMap -> TokenMap on this GPU."""
        if isinstance(features, TokenMap):
            return features
        torch = self.torch
        B, C, H, W = features.shape
        src = self._f32(features)
        out = torch.empty((B, H * W, C), dtype=torch.float32, device=self._dev())
        _lib.check(self.lib.vc_nchw_to_tokens(self.handle, _ptr(src), B, C, H, W, _ptr(out), self._stream()))
        return TokenMap(out, (H, W))

    # ------------------------------------------------------------------ kernels
    def warp_similarity(self, tm1: TokenMap, tm2: TokenMap, h_inv, patch_size: int, image_size):
        """-> similarity f32 [B, H, W]; h_inv = torch.linalg.inv(H) as f32 [B, 3, 3]."""
        torch = self.torch
        B, _, C = tm1.tokens.shape
        H, W = tm1.grid_hw
        assert tm2.tokens.shape == tm1.tokens.shape and tm2.grid_hw == tm1.grid_hw
        hi = self._f32(h_inv.reshape(B, 9))
        sim = torch.empty((B, H, W), dtype=torch.float32, device=self._dev())
        img_h, img_w = image_size
        _lib.check(self.lib.vc_warp_similarity(self.handle, _ptr(tm1.tokens), _ptr(tm2.tokens), B, H, W, C, _ptr(hi),
                                               float(patch_size), float(img_h), float(img_w), _ptr(sim),
                                               self._stream()))
        return sim

    def topk_rows(self, values, k: int, threshold: Optional[float] = None):
        """values f32 [R, n] -> (order int32 [R, k], above int32 [R, n] | None, n_above int32 [R] | None)."""
        torch = self.torch
        values = self._f32(values)
        R, n = values.shape
        k = int(k)
        order = torch.empty((R, max(k, 1)), dtype=torch.int32, device=self._dev())
        above = n_above = None
        if threshold is not None:
            above = torch.empty((R, n), dtype=torch.int32, device=self._dev())
            n_above = torch.empty((R,), dtype=torch.int32, device=self._dev())
        _lib.check(self.lib.vc_topk_rows(self.handle, _ptr(values), R, n, k,
                                         float(threshold if threshold is not None else 0.0), _ptr(order),
                                         _ptr(above), _ptr(n_above), self._stream()))
        return order[:, :k], above, n_above

    def hard_negatives(self, tm: TokenMap, pos_feat, pos_xy, n_hard: int, radius: float):
        """-> (idx int32 [B, K, n_hard] (-1 padded), counts int32 [B, K])."""
        torch = self.torch
        B, _, C = tm.tokens.shape
        H, W = tm.grid_hw
        pos_feat = self._f32(pos_feat)
        pos_xy = self._f32(pos_xy)
        K = int(pos_xy.shape[1])
        idx = torch.empty((B, K, int(n_hard)), dtype=torch.int32, device=self._dev())
        counts = torch.empty((B, K), dtype=torch.int32, device=self._dev())
        _lib.check(self.lib.vc_hard_negatives(self.handle, _ptr(tm.tokens), _ptr(pos_feat), _ptr(pos_xy), B, H, W, C,
                                              K, int(n_hard), float(radius), _ptr(idx), _ptr(counts),
                                              self._stream()))
        return idx, counts

    def sample_bilinear(self, tm: TokenMap, xy):
        """xy f32 [B, n, 2] (x, y) feature-map units (NaN -> zero row) -> f32 [B, n, C]."""
        torch = self.torch
        B, _, C = tm.tokens.shape
        H, W = tm.grid_hw
        xy = self._f32(xy)
        n = int(xy.shape[1])
        out = torch.empty((B, n, C), dtype=torch.float32, device=self._dev())
        if n == 0:
            return out
        _lib.check(self.lib.vc_sample_bilinear(self.handle, _ptr(tm.tokens), B, H, W, C, _ptr(xy), n, _ptr(out),
                                               self._stream()))
        return out


_devices = {}


def get_device(device: int = 0) -> SamplerDevice:
    """Process-wide SamplerDevice of a GPU (created on first use)."""
    d = _devices.get(int(device))
    if d is None or d.handle is None:
        d = _devices[int(device)] = SamplerDevice(int(device))
    return d
