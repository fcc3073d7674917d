"""Oracle for the DINOv2 ViT-S/B/L/14 backbone.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

The backbone is EXTERNAL to the reference: vit_colmap/features/vit_extractor.py:95-98 pulls
``torch.hub.load("facebookresearch/dinov2", model_name)`` (un-pinned, not vendored, no network here)
and uses ``forward_features(x)["x_norm_patchtokens"]`` (:135-146).  PARITY UNPINNED against real
DINOv2 weights; the restatement of the published architecture below is
cross-checked against the independent implementation in HuggingFace transformers
(models/dinov2, random init) by tests/test_oracle_vit.py.

State-dict keys follow the hub model (``patch_embed.proj.weight``, ``blocks.{i}.attn.qkv.weight``,
``blocks.{i}.ls1.gamma`` ...), so a real checkpoint drops in.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn.functional as F

CONFIGS = {  # name -> (embed_dim, depth, heads)
    "dinov2_vits14": (384, 12, 6),
    "dinov2_vitb14": (768, 12, 12),
    "dinov2_vitl14": (1024, 24, 16),
}
PATCH = 14
BASE_GRID = 37  # pos_embed is trained at 518 / 14
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


def synthetic_state_dict(name: str, seed: int = 0, depth: int | None = None) -> Dict[str, torch.Tensor]:
    """SURVEY 8(d) C1 recipe: N(0, 0.02^2) linears / pos / cls, LayerNorm and LayerScale near 1.
    Small random biases and LN offsets keep every epilogue term observable in the parity tests."""
    C, L, _ = CONFIGS[name]
    L = depth or L
    g = torch.Generator().manual_seed(seed)
    n = lambda *s, std=0.02: torch.randn(*s, generator=g) * std  # noqa: E731
    sd = {
        "cls_token": n(1, 1, C), "pos_embed": n(1, 1 + BASE_GRID * BASE_GRID, C),
        "patch_embed.proj.weight": n(C, 3, PATCH, PATCH), "patch_embed.proj.bias": n(C),
        "norm.weight": 1 + n(C, std=0.1), "norm.bias": n(C, std=0.1),
    }
    for i in range(L):
        p = f"blocks.{i}."
        sd.update({
            p + "norm1.weight": 1 + n(C, std=0.1), p + "norm1.bias": n(C, std=0.1),
            p + "attn.qkv.weight": n(3 * C, C, std=0.05), p + "attn.qkv.bias": n(3 * C),
            p + "attn.proj.weight": n(C, C), p + "attn.proj.bias": n(C),
            p + "ls1.gamma": 1 + n(C, std=0.1),
            p + "norm2.weight": 1 + n(C, std=0.1), p + "norm2.bias": n(C, std=0.1),
            p + "mlp.fc1.weight": n(4 * C, C), p + "mlp.fc1.bias": n(4 * C),
            p + "mlp.fc2.weight": n(C, 4 * C), p + "mlp.fc2.bias": n(C),
            p + "ls2.gamma": 1 + n(C, std=0.1),
        })
    return sd


def depth_of(sd) -> int:
    return 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("blocks."))


# torch.hub's dinov2_vit{s,b,l}14 are built with interpolate_offset=0.1, interpolate_antialias=False
# (facebookresearch/dinov2 hub/backbones.py, restated from the public source; the reference does not pin
# or vendor it, vit_extractor.py:95-98, so this stays "parity unpinned").
HUB_INTERPOLATE_OFFSET = 0.1


def interpolate_pos_embed(pos_embed: torch.Tensor, gh: int, gw: int,
                          interpolate_offset: float = HUB_INTERPOLATE_OFFSET) -> torch.Tensor:
    """[1, 1+37*37, C] -> [1+gh*gw, C], bicubic, align_corners=False, no antialias.
    interpolate_offset > 0 (hub default 0.1): DINOv2's interpolate_pos_encoding passes
    scale_factor = ((gh + 0.1) / 37, (gw + 0.1) / 37), so source coordinates are
    (dst + 0.5) * 37 / (g + 0.1) - 0.5 -- slightly different samples from the size-based resize.
    interpolate_offset == 0: size=(gh, gw) (the HF transformers convention, used to cross-check the
    rest of the restatement against transformers.Dinov2Model).  Identical at 37x37 (no resize)."""
    pos = pos_embed[0].float()
    if gh == BASE_GRID and gw == BASE_GRID:
        return pos
    C = pos.shape[1]
    grid = pos[1:].reshape(1, BASE_GRID, BASE_GRID, C).permute(0, 3, 1, 2)
    if interpolate_offset:
        scale = (float(gh + interpolate_offset) / BASE_GRID, float(gw + interpolate_offset) / BASE_GRID)
        grid = F.interpolate(grid, scale_factor=scale, mode="bicubic", align_corners=False)
    else:
        grid = F.interpolate(grid, size=(gh, gw), mode="bicubic", align_corners=False)
    assert tuple(grid.shape[-2:]) == (gh, gw)
    return torch.cat([pos[:1], grid.permute(0, 2, 3, 1).reshape(gh * gw, C)], dim=0)


def preprocess(images_u8: torch.Tensor) -> torch.Tensor:
    """uint8 RGB [B,H,W,3] -> ImageNet-normalised f32 [B,3,H,W] (ToTensor + Normalize,
    vit_extractor.py:76-83,132)."""
    x = images_u8.permute(0, 3, 1, 2).float() / 255.0
    mean = torch.tensor(IMAGENET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(IMAGENET_STD).view(1, 3, 1, 1)
    return (x - mean) / std


def _r(t: torch.Tensor, emulate: bool) -> torch.Tensor:
    return t.to(torch.bfloat16).float() if emulate else t


def _folded_linear(t: torch.Tensor, W: torch.Tensor, b: torch.Tensor, ln_w: torch.Tensor, ln_b: torch.Tensor):
    """LayerNorm (eps 1e-6) + Linear with the rounding points of the kernels' VC_VIT_LN=fold mode: the GEMM runs
    on bf16(t) with weights bf16(W * ln_w); mean and 1/std of the fp32 rows (from sum x and sum x^2) are applied
    after it: rstd * (acc - mean * colsum) + (b + W ln_b)."""
    C = t.shape[-1]
    wf = (W * ln_w[None, :]).to(torch.bfloat16).float()
    acc = t.to(torch.bfloat16).float() @ wf.t()
    mean = t.sum(dim=-1, keepdim=True) / C
    var = ((t * t).sum(dim=-1, keepdim=True) / C - mean * mean).clamp(min=0)
    rstd = torch.rsqrt(var + 1e-6)
    return rstd * (acc - mean * wf.sum(dim=1)) + (b + W @ ln_b)


def forward_features(sd: Dict[str, torch.Tensor], x: torch.Tensor, heads: int, emulate_bf16=False,
                     interpolate_offset: float = HUB_INTERPOLATE_OFFSET):
    """x: normalised f32 [B,3,14*gh,14*gw].  emulate_bf16 rounds every GEMM / attention operand to
    bf16 at the points where the sm_100a kernels do (fp32 accumulation everywhere); emulate_bf16="fold" does
    so for the kernels' folded-LayerNorm mode (every LayerNorm but the first and the last: _folded_linear)."""
    fold = emulate_bf16 == "fold"
    e = bool(emulate_bf16)
    B, _, Hp, Wp = x.shape
    gh, gw = Hp // PATCH, Wp // PATCH
    C = sd["cls_token"].shape[-1]
    L = depth_of(sd)
    # patch embedding == conv 14x14 / stride 14
    patches = F.unfold(x, kernel_size=PATCH, stride=PATCH).transpose(1, 2)          # [B, HW, 588]
    wpe = sd["patch_embed.proj.weight"].reshape(C, -1)
    tok = _r(patches, e) @ _r(wpe, e).t() + sd["patch_embed.proj.bias"]
    pos = interpolate_pos_embed(sd["pos_embed"], gh, gw, interpolate_offset)
    tok = tok + pos[1:]
    cls = (sd["cls_token"].reshape(1, 1, C) + pos[:1]).expand(B, 1, C)
    t = torch.cat([cls, tok], dim=1)                                                # [B, T, C]
    hd = C // heads
    for i in range(L):
        p = f"blocks.{i}."
        if fold and i > 0:
            qkv = _r(_folded_linear(t, sd[p + "attn.qkv.weight"], sd[p + "attn.qkv.bias"], sd[p + "norm1.weight"],
                                    sd[p + "norm1.bias"]), e)
        else:
            y = F.layer_norm(t, (C,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], eps=1e-6)
            qkv = _r(_r(y, e) @ _r(sd[p + "attn.qkv.weight"], e).t() + sd[p + "attn.qkv.bias"], e)
        q, k, v = qkv.reshape(B, -1, 3, heads, hd).permute(2, 0, 3, 1, 4)           # [B, h, T, hd]
        s = (q @ k.transpose(-1, -2)) / math.sqrt(hd)
        if e:  # the kernel rounds un-normalised probabilities to bf16 and divides by the fp32 sum
            pr = torch.exp(s - s.amax(dim=-1, keepdim=True))
            a = (_r(pr, True) @ v) / pr.sum(dim=-1, keepdim=True)
        else:
            a = torch.softmax(s, dim=-1) @ v
        a = _r(a.transpose(1, 2).reshape(B, -1, C), e)
        y = a @ _r(sd[p + "attn.proj.weight"], e).t() + sd[p + "attn.proj.bias"]
        t = t + sd[p + "ls1.gamma"] * y
        if fold:
            hdn = F.gelu(_folded_linear(t, sd[p + "mlp.fc1.weight"], sd[p + "mlp.fc1.bias"], sd[p + "norm2.weight"],
                                        sd[p + "norm2.bias"]))
        else:
            y = F.layer_norm(t, (C,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], eps=1e-6)
            hdn = F.gelu(_r(y, e) @ _r(sd[p + "mlp.fc1.weight"], e).t() + sd[p + "mlp.fc1.bias"])
        y = _r(hdn, e) @ _r(sd[p + "mlp.fc2.weight"], e).t() + sd[p + "mlp.fc2.bias"]
        t = t + sd[p + "ls2.gamma"] * y
    out = F.layer_norm(t, (C,), sd["norm.weight"], sd["norm.bias"], eps=1e-6)
    return {"x_norm_clstoken": out[:, 0], "x_norm_patchtokens": out[:, 1:], "x_prenorm": t}


class OracleDinoV2(torch.nn.Module):
    """Drop-in for the hub model object the reference holds (``.forward_features``, ``.embed_dim``)."""

    def __init__(self, name: str = "dinov2_vitb14", seed: int = 0, state_dict=None):
        super().__init__()
        self.embed_dim, _, self.num_heads = CONFIGS[name]
        self.sd = state_dict if state_dict is not None else synthetic_state_dict(name, seed)

    @torch.no_grad()
    def forward_features(self, x):
        return forward_features(self.sd, x.float().cpu(), self.num_heads)
