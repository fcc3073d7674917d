"""DINOv2 ViT-S/B/L/14 backbone on the sm_100a kernels (host-side weight container and driver).

Replaces the ``torch.hub.load("facebookresearch/dinov2", name)`` model object the reference holds
(vit_colmap/features/vit_extractor.py:95-98) for the one call it makes:
``model.forward_features(x)["x_norm_patchtokens"]`` (:135-146), with the preprocessing of :117-132
(ToTensor + ImageNet normalisation) fused into the first kernel.  PyTorch holds the device memory;
all computation is vc_vit_forward (include/vc_b200.h).  No CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Tuple

from . import _lib

CONFIGS = {  # name -> (embed_dim, depth, heads); facebookresearch/dinov2 hub names
    "dinov2_vits14": (384, 12, 6),
    "dinov2_vitb14": (768, 12, 12),
    "dinov2_vitl14": (1024, 24, 16),
}
PATCH = 14
BASE_GRID = 37
PATCH_KPAD = 640  # 3*14*14 = 588 padded to a multiple of the 64-element GEMM k-block


def _ptr(t) -> int:
    return 0 if t is None else int(t.data_ptr())


class VitBackbone:
    """Weights resident on one GPU + the forward call.  ``state_dict`` uses the hub model's keys."""

    def __init__(self, state_dict: Dict[str, "torch.Tensor"], name: str = "dinov2_vitb14", device: int = 0,
                 handle=None, interpolate_offset: float = 0.1):
        import torch
        if not torch.cuda.is_available():
            raise _lib.VcError("no CUDA device: vit_colmap_b200 has no CPU fallback")
        self.torch = torch
        self.name = name
        # 0.1 = the torch.hub dinov2_vit*14 models the reference loads (vit_extractor.py:95-98) are built
        # with interpolate_offset=0.1 (position table resized by scale_factor (g + 0.1) / 37); 0 = the
        # size-based resize of HF transformers.  Same table at 37x37.  See DESIGN.md section 4.
        self.interpolate_offset = float(interpolate_offset)
        self.embed_dim, depth_default, self.num_heads = CONFIGS[name]
        self.depth = 1 + max(int(k.split(".")[1]) for k in state_dict if k.startswith("blocks."))
        self.device = int(device)
        self.lib = _lib.load()
        self._own = handle is None
        if handle is None:
            handle = ctypes.c_void_p()
            _lib.check(self.lib.vc_create(self.device, ctypes.byref(handle)))
        self.handle = handle
        dev = torch.device("cuda", self.device)
        C = self.embed_dim
        f32 = lambda t: t.detach().to(device=dev, dtype=torch.float32).contiguous()  # noqa: E731
        bf16 = lambda t: t.detach().to(device=dev, dtype=torch.bfloat16).contiguous()  # noqa: E731
        sd = state_dict
        self._keep = []  # device tensors referenced by raw pointers
        pw = torch.zeros((C, PATCH_KPAD), dtype=torch.float32)
        pw[:, :3 * PATCH * PATCH] = sd["patch_embed.proj.weight"].detach().float().reshape(C, -1).cpu()
        self.patch_w = bf16(pw)
        self.patch_b = f32(sd["patch_embed.proj.bias"])
        self.cls_token = f32(sd["cls_token"].reshape(C))
        self.pos_embed_base = sd["pos_embed"].detach().float().cpu()
        self.norm_w, self.norm_b = f32(sd["norm.weight"]), f32(sd["norm.bias"])
        self._blocks = (_lib.VitBlock * self.depth)()
        for i in range(self.depth):
            p = f"blocks.{i}."
            t = dict(ln1_w=f32(sd[p + "norm1.weight"]), ln1_b=f32(sd[p + "norm1.bias"]),
                     qkv_w=bf16(sd[p + "attn.qkv.weight"]), qkv_b=f32(sd[p + "attn.qkv.bias"]),
                     proj_w=bf16(sd[p + "attn.proj.weight"]), proj_b=f32(sd[p + "attn.proj.bias"]),
                     ls1=f32(sd[p + "ls1.gamma"]),
                     ln2_w=f32(sd[p + "norm2.weight"]), ln2_b=f32(sd[p + "norm2.bias"]),
                     fc1_w=bf16(sd[p + "mlp.fc1.weight"]), fc1_b=f32(sd[p + "mlp.fc1.bias"]),
                     fc2_w=bf16(sd[p + "mlp.fc2.weight"]), fc2_b=f32(sd[p + "mlp.fc2.bias"]),
                     ls2=f32(sd[p + "ls2.gamma"]))
            # LayerNorm folded into the next GEMM (VC_VIT_LN=fold, vit_kernels.cuh): W' = bf16(W * ln_w),
            # colsum over the ROUNDED W' (what the tensor core multiplies), bias' = bias + W ln_b
            for nm, wk, bk, lw, lb in (("qkv", "attn.qkv", "attn.qkv", "norm1", "norm1"), ("fc1", "mlp.fc1", "mlp.fc1", "norm2", "norm2")):
                W = sd[p + wk + ".weight"].detach().float().cpu()
                wf = (W * sd[p + lw + ".weight"].detach().float().cpu()[None, :]).to(torch.bfloat16)
                t[nm + "_wf"] = wf.to(dev).contiguous()
                t[nm + "_cs"] = f32(wf.float().sum(dim=1))
                t[nm + "_bf"] = f32(sd[p + bk + ".bias"].detach().float().cpu() + W @ sd[p + lb + ".bias"].detach().float().cpu())
            self._keep.append(t)
            for k, v in t.items():
                setattr(self._blocks[i], k, _ptr(v))
        self._pos = {}
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

    # ------------------------------------------------------------------ host-side one-offs
    def pos_embed(self, gh: int, gw: int):
        """[1+gh*gw, C] f32 on the device; bicubic resize (align_corners=False, no antialias) of the 37x37
        table when the grid differs: by scale_factor (g + interpolate_offset) / 37 as DINOv2's
        interpolate_pos_encoding does for the hub models, or by size when the offset is 0.  A one-off
        host-side step per grid (cached); the result is an input of vc_vit_forward."""
        key = (gh, gw)
        if key not in self._pos:
            torch = self.torch
            pos = self.pos_embed_base[0]
            if key != (BASE_GRID, BASE_GRID):
                C = pos.shape[1]
                grid = pos[1:].reshape(1, BASE_GRID, BASE_GRID, C).permute(0, 3, 1, 2)
                if self.interpolate_offset:
                    scale = (float(gh + self.interpolate_offset) / BASE_GRID, float(gw + self.interpolate_offset) / BASE_GRID)
                    grid = torch.nn.functional.interpolate(grid, scale_factor=scale, mode="bicubic", align_corners=False)
                else:
                    grid = torch.nn.functional.interpolate(grid, size=key, mode="bicubic", align_corners=False)
                assert tuple(grid.shape[-2:]) == key
                pos = torch.cat([pos[:1], grid.permute(0, 2, 3, 1).reshape(gh * gw, C)], dim=0)
            self._pos[key] = pos.to(device=torch.device("cuda", self.device), dtype=torch.float32).contiguous()
        return self._pos[key]

    # ------------------------------------------------------------------ forward
    def forward_patch_tokens(self, images_u8, bgr: bool = False):
        """images_u8: device uint8 [B, 14*gh, 14*gw, 3] -> x_norm_patchtokens f32 [B, gh*gw, C]."""
        torch = self.torch
        assert images_u8.dtype == torch.uint8 and images_u8.dim() == 4 and images_u8.shape[3] == 3
        images_u8 = images_u8.contiguous()
        B, Hp, Wp, _ = images_u8.shape
        if Hp % PATCH or Wp % PATCH:
            raise ValueError("image size must be a multiple of 14 (resize first, vit_extractor.py:118-129)")
        gh, gw = Hp // PATCH, Wp // PATCH
        C = self.embed_dim
        need = int(self.lib.vc_vit_workspace_bytes(B, gh, gw, C))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=images_u8.device)
        pos = self.pos_embed(gh, gw)
        w = _lib.VitWeights(C, self.depth, self.num_heads, PATCH_KPAD, _ptr(self.patch_w), _ptr(self.patch_b),
                            _ptr(self.cls_token), _ptr(pos), _ptr(self.norm_w), _ptr(self.norm_b),
                            ctypes.cast(self._blocks, ctypes.POINTER(_lib.VitBlock)))
        tokens = torch.empty((B, gh * gw, C), dtype=torch.float32, device=images_u8.device)
        stream = int(torch.cuda.current_stream(images_u8.device).cuda_stream)
        _lib.check(self.lib.vc_vit_forward(self.handle, ctypes.byref(w), _ptr(images_u8), int(bool(bgr)), B, gh, gw,
                                           _ptr(tokens), _ptr(self._ws), need, stream))
        return tokens

    def forward_features(self, images_u8, bgr: bool = False):
        """Same key the reference reads from the hub model's output dict."""
        return {"x_norm_patchtokens": self.forward_patch_tokens(images_u8, bgr)}

    def flops_per_image(self, gh: int, gw: int) -> float:
        """L*(24 T C^2 + 4 T^2 C) + 2 HW 588 C."""
        C, L, T = self.embed_dim, self.depth, gh * gw + 1
        return L * (24.0 * T * C * C + 4.0 * T * T * C) + 2.0 * gh * gw * 588 * C


def synthetic_state_dict(name: str, seed: int = 0, depth: Optional[int] = None):
    """Seeded random-init weights in the hub model's key layout (no checkpoint can be downloaded in
    the build / benchmark environment): N(0, 0.02^2) linears, position and cls embeddings, LayerNorm
    and LayerScale scales near 1.  Throughput does not depend on the values."""
    import torch
    C, L, _ = CONFIGS[name]
    L = depth or L
    g = torch.Generator().manual_seed(seed)
    n = lambda *s, std=0.02: torch.randn(*s, generator=g) * std  # noqa: E731
    sd = {"cls_token": n(1, 1, C), "pos_embed": n(1, 1 + BASE_GRID * BASE_GRID, C),
          "patch_embed.proj.weight": n(C, 3, PATCH, PATCH), "patch_embed.proj.bias": n(C),
          "norm.weight": 1 + n(C, std=0.1), "norm.bias": n(C, std=0.1)}
    for i in range(L):
        p = f"blocks.{i}."
        sd.update({p + "norm1.weight": 1 + n(C, std=0.1), p + "norm1.bias": n(C, std=0.1),
                   p + "attn.qkv.weight": n(3 * C, C, std=0.05), p + "attn.qkv.bias": n(3 * C),
                   p + "attn.proj.weight": n(C, C), p + "attn.proj.bias": n(C), p + "ls1.gamma": 1 + n(C, std=0.1),
                   p + "norm2.weight": 1 + n(C, std=0.1), p + "norm2.bias": n(C, std=0.1),
                   p + "mlp.fc1.weight": n(4 * C, C), p + "mlp.fc1.bias": n(4 * C),
                   p + "mlp.fc2.weight": n(C, 4 * C), p + "mlp.fc2.bias": n(C), p + "ls2.gamma": 1 + n(C, std=0.1)})
    return sd
