"""Convolutional head of ViTFeatureModel on the sm_100a kernels (host-side weight container).

Replaces the upsampler / trunk / keypoint head / descriptor head modules of
vit_colmap/model/vit_feature_model.py:90-121 and the head half of ``forward`` (:193-232).  The
state dict uses the reference module's own keys (``upsampler.0.deconv.weight`` ...), so the
``model_state_dict`` of a checkpoint written by scripts/train_vit_features.py loads directly.
Weights are repacked once on the host into the GEMM layouts of include/vc_b200.h (vc_head_weights);
all computation is vc_conv_head_forward.  Eval mode only (BatchNorm running statistics); no CPU
fallback.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Tuple

from . import _lib

BN_EPS = 1e-5
PATCH = 14


def _ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


def pack_conv3x3(w):
    """[Cout, Cin, 3, 3] -> [Cout, 9*Cin], column (ky*3+kx)*Cin + ci."""
    co, ci = w.shape[:2]
    return w.permute(0, 2, 3, 1).reshape(co, 9 * ci)


def pack_deconv4x4s2(w):
    """ConvTranspose2d weight [Cin, Cout, 4, 4] -> four sub-pixel phases [4*Cout, 4*Cin]:
    row (2*py+px)*Cout + co, column (2*ty+tx)*Cin + ci.  oy = 2*iy - 1 + ky, so phase 0 uses
    (offset -1: k=3, offset 0: k=1) and phase 1 uses (offset 0: k=2, offset +1: k=0)."""
    import torch
    ci, co = w.shape[:2]
    k_of = {0: (3, 1), 1: (2, 0)}
    rows = []
    for py in (0, 1):
        for px in (0, 1):
            taps = [w[:, :, k_of[py][ty], k_of[px][tx]].t() for ty in (0, 1) for tx in (0, 1)]  # each [Cout, Cin]
            rows.append(torch.cat(taps, dim=1))
    return torch.cat(rows, dim=0)


def bn_scale_shift(sd, prefix: str, conv_bias):
    s = sd[prefix + ".weight"].float() / (sd[prefix + ".running_var"].float() + BN_EPS).sqrt()
    t = (conv_bias.float() - sd[prefix + ".running_mean"].float()) * s + sd[prefix + ".bias"].float()
    return s, t


class ConvHead:
    """Head weights resident on one GPU + the forward call."""

    def __init__(self, state_dict: Dict[str, "torch.Tensor"], backbone_dim: int, descriptor_dim: int = 128,
                 device: int = 0, handle=None):
        import torch
        if not torch.cuda.is_available():
            raise _lib.VcError("no CUDA device: vit_colmap_b200 has no CPU fallback")
        if backbone_dim % 64 or descriptor_dim % 32:
            raise ValueError("backbone_dim must be a multiple of 64 and descriptor_dim a multiple of 32")
        self.torch = torch
        self.backbone_dim, self.descriptor_dim = int(backbone_dim), int(descriptor_dim)
        self.device = int(device)
        self.lib = _lib.load()
        self._own = handle is None
        if handle is None:
            handle = ctypes.c_void_p()
            _lib.check(self.lib.vc_create(self.device, ctypes.byref(handle)))
        self.handle = handle
        dev = torch.device("cuda", self.device)
        sd = {k: v.detach().cpu() for k, v in state_dict.items()}
        f32 = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()  # noqa: E731
        bf16 = lambda t: t.float().to(device=dev, dtype=torch.bfloat16).contiguous()  # noqa: E731
        if tuple(sd["upsampler.0.deconv.weight"].shape[:2]) != (self.backbone_dim, 512):
            raise ValueError("upsampler.0.deconv.weight does not match backbone_dim")
        if sd["descriptor_head.3.weight"].shape[0] != self.descriptor_dim:
            raise ValueError("descriptor_head.3.weight does not match descriptor_dim")
        k = {}
        for i in range(2):
            p = f"upsampler.{i}."
            k[f"deconv_w{i}"] = bf16(pack_deconv4x4s2(sd[p + "deconv.weight"].float()))
            k[f"deconv_b{i}"] = f32(sd[p + "deconv.bias"])
            k[f"conv_w{i}"] = bf16(pack_conv3x3(sd[p + "conv.weight"].float()))
            s, t = bn_scale_shift(sd, p + "bn", sd[p + "conv.bias"])
            k[f"conv_s{i}"], k[f"conv_t{i}"] = f32(s), f32(t)
        k["trunk_w"] = bf16(pack_conv3x3(sd["trunk.0.weight"].float()))
        s, t = bn_scale_shift(sd, "trunk.1", sd["trunk.0.bias"])
        k["trunk_s"], k["trunk_t"] = f32(s), f32(t)
        ks, kt = bn_scale_shift(sd, "keypoint_head.1", sd["keypoint_head.0.bias"])
        ds, dt = bn_scale_shift(sd, "descriptor_head.1", sd["descriptor_head.0.bias"])
        k["heads_w"] = bf16(torch.cat([pack_conv3x3(sd["keypoint_head.0.weight"].float()),
                                       pack_conv3x3(sd["descriptor_head.0.weight"].float())], dim=0))
        k["heads_s"], k["heads_t"] = f32(torch.cat([ks, ds])), f32(torch.cat([kt, dt]))
        k["kp_w"] = f32(sd["keypoint_head.3.weight"].reshape(4, 64))
        k["kp_b"] = f32(sd["keypoint_head.3.bias"])
        dw = sd["descriptor_head.3.weight"].reshape(self.descriptor_dim, 128).float()
        k["desc_w"] = bf16(dw)
        k["desc_b"] = f32(sd["descriptor_head.3.bias"])
        # the sparse path projects only the selected keypoints: P = W^T [128, d] fp32 (bf16-rounded
        # values, so the dense tensor-core path and the sparse path see the same weights)
        k["desc_proj"] = k["desc_w"].float().t().contiguous()
        self._k = k
        two = ctypes.c_void_p * 2
        self._w = _lib.HeadWeights(
            self.backbone_dim, self.descriptor_dim,
            two(_ptr(k["deconv_w0"]), _ptr(k["deconv_w1"])), two(_ptr(k["deconv_b0"]), _ptr(k["deconv_b1"])),
            two(_ptr(k["conv_w0"]), _ptr(k["conv_w1"])), two(_ptr(k["conv_s0"]), _ptr(k["conv_s1"])),
            two(_ptr(k["conv_t0"]), _ptr(k["conv_t1"])),
            _ptr(k["trunk_w"]), _ptr(k["trunk_s"]), _ptr(k["trunk_t"]),
            _ptr(k["heads_w"]), _ptr(k["heads_s"]), _ptr(k["heads_t"]),
            _ptr(k["kp_w"]), _ptr(k["kp_b"]), _ptr(k["desc_w"]), _ptr(k["desc_b"]))
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

    @staticmethod
    def target_size(gh: int, gw: int) -> Tuple[int, int]:
        """vit_feature_model.py:195-197 / :257-262: a quarter of the (multiple-of-14) image size."""
        return (gh * PATCH) // 4, (gw * PATCH) // 4

    def forward(self, tokens, grid_hw: Tuple[int, int], target_size: Optional[Tuple[int, int]] = None,
                dense_descriptors: bool = True, want_features: bool = False):
        """tokens: device f32/bf16 [B, gh*gw, C] patch tokens.  Returns a dict of channels-last maps:
        ``keypoints`` f32 [B,h4,w4,4]; ``descriptors`` f32 [B,h4,w4,d] (L2-normalised) when
        dense_descriptors, else ``desc_hidden`` bf16 [B,h4,w4,128] for the sparse projection;
        ``features`` bf16 [B,h4,w4,256] when want_features."""
        torch = self.torch
        gh, gw = grid_hw
        B, T, C = tokens.shape
        if T != gh * gw or C != self.backbone_dim:
            raise ValueError(f"tokens {tuple(tokens.shape)} do not match grid {grid_hw} x C={self.backbone_dim}")
        tokens = tokens.contiguous()
        oh, ow = target_size if target_size is not None else self.target_size(gh, gw)
        dev = tokens.device
        need = int(self.lib.vc_head_workspace_bytes(B, gh, gw, C, oh, ow))
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        kp = torch.empty((B, oh, ow, 4), dtype=torch.float32, device=dev)
        desc = torch.empty((B, oh, ow, self.descriptor_dim), dtype=torch.float32, device=dev) if dense_descriptors else None
        hid = None if dense_descriptors else torch.empty((B, oh, ow, 128), dtype=torch.bfloat16, device=dev)
        feat = torch.empty((B, oh, ow, 256), dtype=torch.bfloat16, device=dev) if want_features else None
        dtype = _lib.DTYPE_BF16 if tokens.dtype == torch.bfloat16 else _lib.DTYPE_F32
        if tokens.dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("tokens must be float32 or bfloat16")
        stream = int(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(self.lib.vc_conv_head_forward(self.handle, ctypes.byref(self._w), _ptr(tokens), dtype, B, gh, gw,
                                                 oh, ow, _ptr(kp), _ptr(desc), _ptr(feat), _ptr(hid), _ptr(self._ws),
                                                 need, stream))
        out = {"keypoints": kp}
        if desc is not None:
            out["descriptors"] = desc
        if hid is not None:
            out["desc_hidden"] = hid
        if feat is not None:
            out["features"] = feat
        return out

    @property
    def desc_projection(self):
        """(P f32 [128, d], bias f32 [d]) for vc_sample_pack's projecting path."""
        return self._k["desc_proj"], self._k["desc_b"]

    def flops_per_image(self, gh: int, gw: int, target_size: Optional[Tuple[int, int]] = None) -> float:
        """2*MACs of the head: deconv1, conv1, deconv2, conv2, trunk, heads, 1x1s."""
        oh, ow = target_size if target_size is not None else self.target_size(gh, gw)
        C, d = self.backbone_dim, self.descriptor_dim
        p1, p2, p4 = gh * gw * 4, gh * gw * 16, oh * ow
        macs = (p1 * 4 * C * 512 + p1 * 9 * 512 * 512 + p2 * 4 * 512 * 512 + p2 * 9 * 512 * 512
                + p4 * 9 * 512 * 256 + p4 * 9 * 256 * 192 + p4 * (64 * 4 + 128 * d))
        return 2.0 * macs


def synthetic_head_state_dict(backbone_dim: int, descriptor_dim: int, seed: int = 0):
    """Seeded random-init head weights in the reference module's key layout (no trained checkpoint
    ships with the reference).  Fan-in scaled; throughput does not depend on the values."""
    import math
    import numpy as np
    import torch
    rng = np.random.default_rng(seed)
    n = lambda *shape, std=1.0: torch.from_numpy((rng.standard_normal(shape) * std).astype(np.float32))  # noqa: E731
    sd = {}

    def bn(prefix, c):
        sd[prefix + ".weight"] = 1 + n(c, std=0.1)
        sd[prefix + ".bias"] = n(c, std=0.1)
        sd[prefix + ".running_mean"] = n(c, std=0.1)
        sd[prefix + ".running_var"] = torch.from_numpy(rng.uniform(0.5, 1.5, c).astype(np.float32))

    for i, cin in enumerate((backbone_dim, 512)):
        p = f"upsampler.{i}."
        sd[p + "deconv.weight"] = n(cin, 512, 4, 4, std=1.0 / math.sqrt(4 * cin))
        sd[p + "deconv.bias"] = n(512, std=0.1)
        sd[p + "conv.weight"] = n(512, 512, 3, 3, std=1.0 / math.sqrt(9 * 512))
        sd[p + "conv.bias"] = n(512, std=0.1)
        bn(p + "bn", 512)
    sd["trunk.0.weight"] = n(256, 512, 3, 3, std=1.5 / math.sqrt(9 * 512))
    sd["trunk.0.bias"] = n(256, std=0.1)
    bn("trunk.1", 256)
    sd["keypoint_head.0.weight"] = n(64, 256, 3, 3, std=1.5 / math.sqrt(9 * 256))
    sd["keypoint_head.0.bias"] = n(64, std=0.1)
    bn("keypoint_head.1", 64)
    sd["keypoint_head.3.weight"] = n(4, 64, 1, 1, std=2.0 / math.sqrt(64))
    sd["keypoint_head.3.bias"] = n(4, std=0.1)
    sd["descriptor_head.0.weight"] = n(128, 256, 3, 3, std=1.5 / math.sqrt(9 * 256))
    sd["descriptor_head.0.bias"] = n(128, std=0.1)
    bn("descriptor_head.1", 128)
    sd["descriptor_head.3.weight"] = n(descriptor_dim, 128, 1, 1, std=1.0 / math.sqrt(128))
    sd["descriptor_head.3.bias"] = n(descriptor_dim, std=0.05)
    return sd
