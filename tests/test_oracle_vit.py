"""Pins oracle/vit_oracle.py (restatement of the public DINOv2 architecture) against the independent
implementation shipped in HuggingFace transformers (random init, same weights copied across)."""
import numpy as np
import pytest
import torch

from oracle import vit_oracle as vo


def _hf_model(sd, name, depth):
    from transformers import Dinov2Config, Dinov2Model
    C, _, heads = vo.CONFIGS[name]
    cfg = Dinov2Config(hidden_size=C, num_hidden_layers=depth, num_attention_heads=heads, mlp_ratio=4,
                       image_size=518, patch_size=14, layer_norm_eps=1e-6, qkv_bias=True,
                       layerscale_value=1.0, hidden_act="gelu", use_swiglu_ffn=False,
                       attention_probs_dropout_prob=0.0, hidden_dropout_prob=0.0, drop_path_rate=0.0)
    m = Dinov2Model(cfg).eval()
    hf = {"embeddings.cls_token": sd["cls_token"], "embeddings.position_embeddings": sd["pos_embed"],
          "embeddings.mask_token": torch.zeros(1, C),
          "embeddings.patch_embeddings.projection.weight": sd["patch_embed.proj.weight"],
          "embeddings.patch_embeddings.projection.bias": sd["patch_embed.proj.bias"],
          "layernorm.weight": sd["norm.weight"], "layernorm.bias": sd["norm.bias"]}
    for i in range(depth):
        p, q = f"blocks.{i}.", f"encoder.layer.{i}."
        w, b = sd[p + "attn.qkv.weight"], sd[p + "attn.qkv.bias"]
        for j, nm in enumerate(("query", "key", "value")):
            hf[q + f"attention.attention.{nm}.weight"] = w[j * C:(j + 1) * C]
            hf[q + f"attention.attention.{nm}.bias"] = b[j * C:(j + 1) * C]
        hf[q + "attention.output.dense.weight"] = sd[p + "attn.proj.weight"]
        hf[q + "attention.output.dense.bias"] = sd[p + "attn.proj.bias"]
        hf[q + "layer_scale1.lambda1"] = sd[p + "ls1.gamma"]
        hf[q + "layer_scale2.lambda1"] = sd[p + "ls2.gamma"]
        for nm in ("norm1", "norm2"):
            hf[q + nm + ".weight"] = sd[p + nm + ".weight"]
            hf[q + nm + ".bias"] = sd[p + nm + ".bias"]
        for nm in ("fc1", "fc2"):
            hf[q + f"mlp.{nm}.weight"] = sd[p + f"mlp.{nm}.weight"]
            hf[q + f"mlp.{nm}.bias"] = sd[p + f"mlp.{nm}.bias"]
    missing, unexpected = m.load_state_dict(hf, strict=False)
    assert not unexpected and all("pooler" in k for k in missing), (missing, unexpected)
    return m


@pytest.mark.parametrize("name,size", [("dinov2_vits14", (518, 518)), ("dinov2_vitb14", (140, 196))])
def test_restatement_matches_transformers_dinov2(name, size):
    depth = 2
    sd = vo.synthetic_state_dict(name, seed=3, depth=depth)
    rng = np.random.default_rng(0)
    img = torch.from_numpy(rng.integers(0, 256, (1, size[0], size[1], 3), dtype=np.uint8))
    x = vo.preprocess(img)
    ours = vo.forward_features(sd, x, vo.CONFIGS[name][2], interpolate_offset=0.0)  # HF resizes the table by size
    with torch.no_grad():
        ref = _hf_model(sd, name, depth)(pixel_values=x, interpolate_pos_encoding=True).last_hidden_state
    assert np.abs(ours["x_norm_patchtokens"].numpy() - ref[:, 1:].numpy()).max() < 2e-4
    assert np.abs(ours["x_norm_clstoken"].numpy() - ref[:, 0].numpy()).max() < 2e-4


def test_bf16_emulation_stays_close_to_fp32():
    sd = vo.synthetic_state_dict("dinov2_vits14", seed=1, depth=3)
    rng = np.random.default_rng(1)
    x = vo.preprocess(torch.from_numpy(rng.integers(0, 256, (1, 140, 140, 3), dtype=np.uint8)))
    a = vo.forward_features(sd, x, 6)["x_norm_patchtokens"]
    b = vo.forward_features(sd, x, 6, emulate_bf16=True)["x_norm_patchtokens"]
    assert (a - b).abs().max() < 0.1 and (a - b).abs().mean() < 0.01


def test_oracle_module_mirrors_hub_interface():
    m = vo.OracleDinoV2("dinov2_vits14", seed=0, state_dict=vo.synthetic_state_dict("dinov2_vits14", 0, depth=1))
    out = m.forward_features(torch.zeros(1, 3, 28, 42))
    assert m.embed_dim == 384 and out["x_norm_patchtokens"].shape == (1, 6, 384)


def _bicubic_1d_weights(dst: int, src: int, scale: float):
    """Index / weight table of one axis of ATen's upsample_bicubic2d (align_corners=False, A = -0.75):
    source coordinate (o + 0.5) / scale - 0.5, four taps with clamped indices."""
    A = -0.75
    idx = np.zeros((dst, 4), np.int64)
    w = np.zeros((dst, 4), np.float64)
    for o in range(dst):
        x = (o + 0.5) / scale - 0.5
        x0 = int(np.floor(x))
        t = x - x0
        cc = [((A * (t + 1) - 5 * A) * (t + 1) + 8 * A) * (t + 1) - 4 * A,
              ((A + 2) * t - (A + 3)) * t * t + 1,
              ((A + 2) * (1 - t) - (A + 3)) * (1 - t) * (1 - t) + 1,
              ((A * (2 - t) - 5 * A) * (2 - t) + 8 * A) * (2 - t) - 4 * A]
        for k in range(4):
            idx[o, k] = min(max(x0 - 1 + k, 0), src - 1)
            w[o, k] = cc[k]
    return idx, w


@pytest.mark.parametrize("grid", [(10, 14), (42, 57), (85, 114)])
def test_hub_position_table_resize(grid):
    """The torch.hub DINOv2 models resize the 37x37 position table with scale_factor (g + 0.1) / 37
    (interpolate_offset = 0.1), which samples other source coordinates than a size-based resize
    (ADVICE r01).  Checked against a hand-written separable bicubic with exactly those coordinates;
    the two variants agree at 37x37 (nothing is resized) and differ elsewhere."""
    gh, gw = grid
    g = torch.Generator().manual_seed(0)
    pos = torch.randn(1, 1 + 37 * 37, 8, generator=g)
    hub = vo.interpolate_pos_embed(pos, gh, gw).numpy()
    size_based = vo.interpolate_pos_embed(pos, gh, gw, interpolate_offset=0.0).numpy()
    iy, wy = _bicubic_1d_weights(gh, 37, (gh + 0.1) / 37)
    ix, wx = _bicubic_1d_weights(gw, 37, (gw + 0.1) / 37)
    table = pos[0, 1:].numpy().astype(np.float64).reshape(37, 37, 8)
    rows = (table[iy] * wy[:, :, None, None]).sum(1)                  # [gh, 37, C]
    want = (rows[:, ix] * wx[None, :, :, None]).sum(2).reshape(gh * gw, 8)
    assert np.abs(hub[1:] - want).max() < 5e-5  # ATen evaluates the source coordinates in fp32
    assert np.array_equal(hub[0], pos[0, 0].numpy())
    assert np.abs(hub - size_based).max() > 1e-3
    same = vo.interpolate_pos_embed(pos, 37, 37).numpy()
    assert np.array_equal(same, vo.interpolate_pos_embed(pos, 37, 37, interpolate_offset=0.0).numpy())
