"""Oracle for the convolutional head of ViTFeatureModel.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Restates vit_colmap/model/vit_feature_model.py: UpsampleBlock :12-29 (ConvTranspose2d k4 s2 p1 ->
Conv3x3 -> BatchNorm -> GELU), the two-block upsampler :90-95, the bilinear resize to
(14H//4, 14W//4) :199-212 / :263-279, the trunk :98-102, keypoint head :108-113, descriptor head
:116-121, tanh(.)*pi on channel 3 :222-223 and F.normalize(dim=1, eps=1e-8) :226, all in eval mode
(BatchNorm uses running statistics).  Written op by op (explicit sub-pixel phases for the transposed
convolution, explicit bilinear source index arithmetic) so that it documents what the kernels
compute; tests/test_oracle_head.py pins it to outputs of the reference's own ViTFeatureModel
(tests/golden/head_*.npz, made by oracle/gen_golden_head.py).

State-dict keys are the reference module's own (``upsampler.0.deconv.weight``,
``keypoint_head.3.bias`` ...), so a checkpoint written by scripts/train_vit_features.py drops in.

``emulate_bf16=True`` rounds every tensor-core operand (activations and conv weights) and every
stored activation to bf16 where the CUDA path does, with fp32 accumulation, so the GPU parity test
can hold a tight tolerance; the default is the reference's fp32 arithmetic.
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm2d default, vit_feature_model.py:21


def synthetic_head_state_dict(backbone_dim: int, descriptor_dim: int, seed: int = 0) -> Dict[str, torch.Tensor]:
    """Seeded head weights (numpy Generator: stable across versions).  Fan-in scaled so activations
    stay O(1) through the stack; BatchNorm statistics and affine terms are non-trivial so every
    epilogue term is observable."""
    rng = np.random.default_rng(seed)

    def n(*shape, std=1.0):
        return torch.from_numpy((rng.standard_normal(shape) * std).astype(np.float32))

    def bn(prefix, c, sd):
        sd[prefix + ".weight"] = 1 + n(c, std=0.1)
        sd[prefix + ".bias"] = n(c, std=0.1)
        sd[prefix + ".running_mean"] = n(c, std=0.1)
        sd[prefix + ".running_var"] = torch.from_numpy(rng.uniform(0.5, 1.5, c).astype(np.float32))
        sd[prefix + ".num_batches_tracked"] = torch.tensor(1, dtype=torch.long)

    sd: Dict[str, torch.Tensor] = {}
    for i, cin in enumerate((backbone_dim, 512)):
        p = f"upsampler.{i}."
        sd[p + "deconv.weight"] = n(cin, 512, 4, 4, std=1.0 / math.sqrt(4 * cin))
        sd[p + "deconv.bias"] = n(512, std=0.1)
        sd[p + "conv.weight"] = n(512, 512, 3, 3, std=1.0 / math.sqrt(9 * 512))
        sd[p + "conv.bias"] = n(512, std=0.1)
        bn(p + "bn", 512, sd)
    sd["trunk.0.weight"] = n(256, 512, 3, 3, std=1.5 / math.sqrt(9 * 512))
    sd["trunk.0.bias"] = n(256, std=0.1)
    bn("trunk.1", 256, sd)
    sd["keypoint_head.0.weight"] = n(64, 256, 3, 3, std=1.5 / math.sqrt(9 * 256))
    sd["keypoint_head.0.bias"] = n(64, std=0.1)
    bn("keypoint_head.1", 64, sd)
    sd["keypoint_head.3.weight"] = n(4, 64, 1, 1, std=2.0 / math.sqrt(64))
    sd["keypoint_head.3.bias"] = n(4, std=0.1)
    sd["descriptor_head.0.weight"] = n(128, 256, 3, 3, std=1.5 / math.sqrt(9 * 256))
    sd["descriptor_head.0.bias"] = n(128, std=0.1)
    bn("descriptor_head.1", 128, sd)
    sd["descriptor_head.3.weight"] = n(descriptor_dim, 128, 1, 1, std=1.0 / math.sqrt(128))
    sd["descriptor_head.3.bias"] = n(descriptor_dim, std=0.05)
    return sd


def _r(t: torch.Tensor, on: bool) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32) if on else t


def bn_scale_shift(sd, prefix: str, conv_bias: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Eval-mode BatchNorm after a convolution as y = conv_nobias * s + t."""
    s = sd[prefix + ".weight"].float() / torch.sqrt(sd[prefix + ".running_var"].float() + BN_EPS)
    t = (conv_bias.float() - sd[prefix + ".running_mean"].float()) * s + sd[prefix + ".bias"].float()
    return s, t


def deconv4x4s2(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """ConvTranspose2d(k=4, s=2, p=1) written as four sub-pixel phases (the form the kernel uses):
    out[2y+py, 2x+px] = sum over 2x2 input taps.  oy = 2*iy - 1 + ky, so phase 0 reads rows
    (y-1: ky=3, y: ky=1) and phase 1 reads (y: ky=2, y+1: ky=0); columns likewise."""
    B, Cin, H, W = x.shape
    Cout = w.shape[1]
    out = x.new_zeros((B, Cout, 2 * H, 2 * W))
    xp = F.pad(x, (1, 1, 1, 1))
    taps = {0: ((-1, 3), (0, 1)), 1: ((0, 2), (1, 0))}  # phase -> ((d, k), (d, k))
    for py in (0, 1):
        for px in (0, 1):
            acc = x.new_zeros((B, Cout, H, W))
            for dy, ky in taps[py]:
                for dx, kx in taps[px]:
                    win = xp[:, :, 1 + dy:1 + dy + H, 1 + dx:1 + dx + W]
                    acc = acc + torch.einsum("bihw,io->bohw", win, w[:, :, ky, kx])
            out[:, :, py::2, px::2] = acc
    return out + b.view(1, -1, 1, 1)


def bilinear_resize(x: torch.Tensor, size: Tuple[int, int]) -> torch.Tensor:
    """F.interpolate(mode="bilinear", align_corners=False) restated: src = (dst + 0.5) * in/out - 0.5
    clamped at 0, i1 = min(i0 + 1, in - 1), weights (1 - l, l), fp32 throughout."""
    B, C, H, W = x.shape
    oh, ow = size

    def axis(n_in, n_out):
        scale = np.float32(n_in) / np.float32(n_out)
        dst = np.arange(n_out, dtype=np.float32)
        src = np.maximum(scale * (dst + np.float32(0.5)) - np.float32(0.5), np.float32(0)).astype(np.float32)
        i0 = np.minimum(src.astype(np.int64), n_in - 1)
        i1 = np.minimum(i0 + 1, n_in - 1)
        l1 = (src - i0.astype(np.float32)).astype(np.float32)
        return torch.from_numpy(i0), torch.from_numpy(i1), torch.from_numpy(l1), torch.from_numpy(np.float32(1) - l1)

    y0, y1, ly1, ly0 = axis(H, oh)
    x0, x1, lx1, lx0 = axis(W, ow)
    top = x[:, :, y0][:, :, :, x0] * lx0 + x[:, :, y0][:, :, :, x1] * lx1
    bot = x[:, :, y1][:, :, :, x0] * lx0 + x[:, :, y1][:, :, :, x1] * lx1
    return top * ly0.view(1, 1, -1, 1) + bot * ly1.view(1, 1, -1, 1)


def head_forward(sd: Dict[str, torch.Tensor], feat: torch.Tensor, target_size: Optional[Tuple[int, int]] = None,
                 emulate_bf16: bool = False, return_intermediates: bool = False) -> Dict[str, torch.Tensor]:
    """forward_from_backbone_features (vit_feature_model.py:234-295) on feat [B, C, H, W] fp32.
    Returns the reference's dict: keypoints [B,4,h4,w4], descriptors [B,d,h4,w4], features [B,256,h4,w4]."""
    e = emulate_bf16
    inter = {}
    x = _r(feat.float(), e)
    for i in range(2):
        p = f"upsampler.{i}."
        x = _r(deconv4x4s2(x, _r(sd[p + "deconv.weight"].float(), e), sd[p + "deconv.bias"].float()), e)
        inter[f"deconv{i}"] = x
        s, t = bn_scale_shift(sd, p + "bn", sd[p + "conv.bias"])
        y = F.conv2d(x, _r(sd[p + "conv.weight"].float(), e), None, padding=1)
        x = _r(F.gelu(y * s.view(1, -1, 1, 1) + t.view(1, -1, 1, 1)), e)
        inter[f"up{i}"] = x
    H, W = feat.shape[2:]
    th, tw = target_size if target_size is not None else ((H * 14) // 4, (W * 14) // 4)
    if tuple(x.shape[2:]) != (th, tw):
        x = _r(bilinear_resize(x, (th, tw)), e)
    inter["resized"] = x
    s, t = bn_scale_shift(sd, "trunk.1", sd["trunk.0.bias"])
    trunk = _r(F.gelu(F.conv2d(x, _r(sd["trunk.0.weight"].float(), e), None, padding=1) * s.view(1, -1, 1, 1)
                      + t.view(1, -1, 1, 1)), e)
    s, t = bn_scale_shift(sd, "keypoint_head.1", sd["keypoint_head.0.bias"])
    kh = _r(F.gelu(F.conv2d(trunk, _r(sd["keypoint_head.0.weight"].float(), e), None, padding=1)
                   * s.view(1, -1, 1, 1) + t.view(1, -1, 1, 1)), e)
    kp = F.conv2d(kh, sd["keypoint_head.3.weight"].float(), sd["keypoint_head.3.bias"].float())  # fp32 SIMT on the GPU
    kp = kp.clone()
    kp[:, 3] = torch.tanh(kp[:, 3]) * math.pi
    s, t = bn_scale_shift(sd, "descriptor_head.1", sd["descriptor_head.0.bias"])
    dh = _r(F.gelu(F.conv2d(trunk, _r(sd["descriptor_head.0.weight"].float(), e), None, padding=1)
                   * s.view(1, -1, 1, 1) + t.view(1, -1, 1, 1)), e)
    desc = F.conv2d(dh, _r(sd["descriptor_head.3.weight"].float(), e), sd["descriptor_head.3.bias"].float())
    desc = F.normalize(desc, p=2, dim=1, eps=1e-8)
    out = {"keypoints": kp, "descriptors": desc, "features": trunk}
    if return_intermediates:
        out.update(inter)
        out["kp_hidden"], out["desc_hidden"] = kh, dh
    return out
