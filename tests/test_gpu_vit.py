"""GPU parity tests of the ViT forward kernels (through the C ABI).  Per-op tests compare against a
plain PyTorch fp32 reference of the same op on the same bf16-rounded operands (tolerances written in
each test); the whole forward is compared with oracle/vit_oracle.py."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib_handle():
    from vit_colmap_b200 import _lib
    lib = _lib.load()
    h = ctypes.c_void_p()
    _lib.check(lib.vc_create(0, ctypes.byref(h)))
    yield lib, h
    lib.vc_destroy(h)


def _stream():
    import torch
    return int(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("shape", [(1370, 2304, 768), (300, 384, 384), (2740, 768, 3072), (129, 1152, 384),
                                   (64, 3072, 768)])
@pytest.mark.parametrize("epilogue", [0, 1, 2])
def test_linear_matches_fp32_reference(lib_handle, shape, epilogue):
    import torch
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    m, n, k = shape
    g = torch.Generator(device="cuda").manual_seed(m + n + k + epilogue)
    a = torch.randn((m, k), generator=g, device="cuda").to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn((n,), generator=g, device="cuda")
    gamma = torch.randn((n,), generator=g, device="cuda")
    acc = a.float() @ w.float().t() + bias
    if epilogue == 2:
        out = torch.randn((m, n), generator=g, device="cuda")
        ref = out + gamma * acc
    else:
        out = torch.zeros((m, n), dtype=torch.bfloat16, device="cuda")
        ref = acc if epilogue == 0 else torch.nn.functional.gelu(acc)
    _lib.check(lib.vc_vit_linear(h, a.data_ptr(), w.data_ptr(), bias.data_ptr(), gamma.data_ptr(), m, n, k,
                                 epilogue, out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    err = (out.float() - ref).abs()
    scale = ref.abs().max().item()
    # fp32 accumulation of exact bf16 products; bf16 outputs carry one rounding (2^-9 relative)
    tol = 1e-4 * scale if epilogue == 2 else 2 ** -8 * scale
    assert err.max().item() <= tol, (err.max().item(), scale)


@pytest.mark.parametrize("shape", [(1370, 2304, 768), (129, 768, 768), (257, 384, 768), (1, 3072, 768),
                                   (5000, 768, 3072), (640, 1152, 1024)])
@pytest.mark.parametrize("epilogue", [0, 1, 2])
def test_two_sm_linear_equals_one_sm(lib_handle, monkeypatch, shape, epilogue):
    """The cta_group::2 GEMM (vit_gemm2.cu, default for K >= 768) and the one-SM GEMM (VC_VIT_2CTA=0) share the
    per-tile epilogue and the order of accumulation: identical outputs, bit for bit -- also when the second CTA
    of a pair has one valid row (M = 129), none (M = 1), or a half-empty column tile (N = 384, 1152)."""
    import torch
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    m, n, k = shape
    g = torch.Generator(device="cuda").manual_seed(m + n + k + epilogue)
    a = torch.randn((m, k), generator=g, device="cuda").to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn((n,), generator=g, device="cuda")
    gamma = torch.randn((n,), generator=g, device="cuda")
    x0 = torch.randn((m, n), generator=g, device="cuda")
    outs = []
    for env in ("1", "0"):
        monkeypatch.setenv("VC_VIT_2CTA", env)
        out = x0.clone() if epilogue == 2 else torch.full((m, n), float("nan"), dtype=torch.bfloat16, device="cuda")
        _lib.check(lib.vc_vit_linear(h, a.data_ptr(), w.data_ptr(), bias.data_ptr(), gamma.data_ptr(), m, n, k,
                                     epilogue, out.data_ptr(), _stream()))
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.isfinite(outs[0].float()).all() and torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("rows,C", [(1370, 768), (77, 384), (513, 1024)])
def test_layernorm_matches_torch(lib_handle, rows, C):
    import torch
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    g = torch.Generator(device="cuda").manual_seed(rows)
    x = torch.randn((rows, C), generator=g, device="cuda") * 3 + 0.5
    w = torch.randn((C,), generator=g, device="cuda")
    b = torch.randn((C,), generator=g, device="cuda")
    out = torch.zeros((rows, C), dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.vc_vit_layernorm(h, x.data_ptr(), w.data_ptr(), b.data_ptr(), rows, C, out.data_ptr(), _stream()))
    ref = torch.nn.functional.layer_norm(x, (C,), w, b, eps=1e-6)
    assert (out.float() - ref).abs().max().item() <= 2 ** -8 * ref.abs().max().item()


@pytest.mark.parametrize("B,T,heads", [(2, 1370, 6), (1, 128, 2), (3, 257, 12), (1, 100, 1), (1, 2395, 2)])
def test_attention_matches_fp32_reference(lib_handle, B, T, heads):
    import torch
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    C = heads * 64
    g = torch.Generator(device="cuda").manual_seed(T + heads)
    qkv = (torch.randn((B, T, 3 * C), generator=g, device="cuda") * 1.5).to(torch.bfloat16)
    out = torch.zeros((B, T, C), dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.vc_vit_attention(h, qkv.data_ptr(), B, T, heads, out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    q, k, v = qkv.float().reshape(B, T, 3, heads, 64).permute(2, 0, 3, 1, 4)
    ref = (torch.softmax(q @ k.transpose(-1, -2) / 8.0, dim=-1) @ v).transpose(1, 2).reshape(B, T, C)
    err = (out.float() - ref).abs().max().item()
    # probabilities and outputs are rounded to bf16 (2^-9 relative each)
    assert err <= 1.5e-2 * ref.abs().max().item(), err


@pytest.mark.parametrize("name,size,depth,B", [("dinov2_vits14", (518, 518), 1, 2), ("dinov2_vits14", (518, 518), 12, 2),
                                                ("dinov2_vitb14", (518, 518), 3, 3), ("dinov2_vitb14", (588, 798), 2, 1),
                                                ("dinov2_vitl14", (224, 224), 2, 1)])
def test_forward_matches_oracle(name, size, depth, B):
    """Whole backbone vs the restatement with bf16 rounding at the kernels' rounding points.
    Tolerance on the LayerNorm-ed patch tokens (|x| ~ 1): 1e-3 relative (north star) for one block.
    Two correct bf16 pipelines diverge by rounding-boundary flips (1 bf16 ulp = 3.9e-3) that
    accumulate over depth, so deeper stacks are held to the oracle's own bf16-vs-fp32 gap: the mean
    error against the bf16 oracle must stay below 0.75x that gap, the max error against the fp32
    oracle below 2x its max, and every token's cosine similarity above 1 - 1e-4."""
    import torch
    from oracle import vit_oracle as vo
    from vit_colmap_b200.vit import VitBackbone
    sd = vo.synthetic_state_dict(name, seed=5, depth=depth)
    rng = np.random.default_rng(7)
    img = torch.from_numpy(rng.integers(0, 256, (B, size[0], size[1], 3), dtype=np.uint8))
    x = vo.preprocess(img)
    heads = vo.CONFIGS[name][2]
    ref = vo.forward_features(sd, x, heads, emulate_bf16=True)["x_norm_patchtokens"]
    ref32 = vo.forward_features(sd, x, heads)["x_norm_patchtokens"]
    net = VitBackbone(sd, name, 0)
    got = net.forward_patch_tokens(img.cuda()).cpu()
    net.close()
    err = (got - ref).abs()
    scale = ref.abs().max().item()
    bf16_gap = (ref - ref32).abs()          # what bf16 operands cost the oracle itself
    cos = torch.nn.functional.cosine_similarity(got, ref, dim=-1)
    if depth == 1:
        assert err.max().item() < 1e-3 * scale, err.max().item() / scale
    assert err.mean().item() < 0.75 * bf16_gap.mean().item(), (err.mean().item(), bf16_gap.mean().item())
    assert (got - ref32).abs().max().item() < 2.0 * bf16_gap.max().item()
    assert cos.min().item() > 1 - 1e-4, cos.min().item()


@pytest.mark.parametrize("name,size,depth,B", [("dinov2_vits14", (518, 518), 12, 2), ("dinov2_vitb14", (518, 518), 3, 3),
                                                ("dinov2_vitb14", (588, 798), 2, 1), ("dinov2_vitl14", (224, 224), 3, 1)])
def test_forward_with_folded_layernorm(monkeypatch, name, size, depth, B):
    """VC_VIT_LN=fold: every LayerNorm between the first and the last is folded into the GEMMs around it (the residual
    GEMM stores bf16(x) and per-slice row sums; the next GEMM multiplies by W * ln_w and applies mean / rstd in its
    epilogue).  The rounding points differ from the default mode's, so the yardstick is the oracle emulating THIS
    mode; the criteria are those of test_forward_matches_oracle, and the distance to the fp32 oracle must not be
    worse than the default mode's own."""
    import torch
    from oracle import vit_oracle as vo
    from vit_colmap_b200.vit import VitBackbone
    sd = vo.synthetic_state_dict(name, seed=5, depth=depth)
    rng = np.random.default_rng(7)
    img = torch.from_numpy(rng.integers(0, 256, (B, size[0], size[1], 3), dtype=np.uint8))
    x = vo.preprocess(img)
    heads = vo.CONFIGS[name][2]
    ref = vo.forward_features(sd, x, heads, emulate_bf16="fold")["x_norm_patchtokens"]
    ref32 = vo.forward_features(sd, x, heads)["x_norm_patchtokens"]
    net = VitBackbone(sd, name, 0)
    default = net.forward_patch_tokens(img.cuda()).cpu()
    monkeypatch.setenv("VC_VIT_LN", "fold")
    got = net.forward_patch_tokens(img.cuda()).cpu()
    again = net.forward_patch_tokens(img.cuda()).cpu()
    net.close()
    assert torch.equal(got, again)                      # deterministic: no atomics in the row statistics
    assert not torch.equal(got, default)                # the mode is really on
    bf16_gap = (ref - ref32).abs()
    err = (got - ref).abs()
    cos = torch.nn.functional.cosine_similarity(got, ref, dim=-1)
    assert err.mean().item() < 0.75 * bf16_gap.mean().item(), (err.mean().item(), bf16_gap.mean().item())
    assert (got - ref32).abs().max().item() < 2.0 * bf16_gap.max().item()
    assert cos.min().item() > 1 - 1e-4, cos.min().item()
    assert (got - ref32).abs().mean().item() < 1.1 * (default - ref32).abs().mean().item()


def test_forward_bgr_flag_swaps_channels():
    import torch
    from oracle import vit_oracle as vo
    from vit_colmap_b200.vit import VitBackbone
    sd = vo.synthetic_state_dict("dinov2_vits14", seed=2, depth=1)
    rng = np.random.default_rng(3)
    img = torch.from_numpy(rng.integers(0, 256, (1, 56, 70, 3), dtype=np.uint8)).cuda()
    net = VitBackbone(sd, "dinov2_vits14", 0)
    a = net.forward_patch_tokens(img, bgr=False)
    b = net.forward_patch_tokens(img.flip(-1).contiguous(), bgr=True)
    net.close()
    assert torch.equal(a, b)


def test_forward_dtu_grid_85x114():
    """Config C3's grid (1600x1200 resized to 1596x1190 -> 85 x 114 patches, T = 9691): bicubic
    position-embedding resize, 76 key tiles per query tile in attention, one ViT-B block."""
    import torch
    from oracle import vit_oracle as vo
    from vit_colmap_b200.vit import VitBackbone
    sd = vo.synthetic_state_dict("dinov2_vitb14", seed=9, depth=1)
    rng = np.random.default_rng(11)
    img = torch.from_numpy(rng.integers(0, 256, (1, 1190, 1596, 3), dtype=np.uint8))
    ref = vo.forward_features(sd, vo.preprocess(img), 12, emulate_bf16=True)["x_norm_patchtokens"]
    net = VitBackbone(sd, "dinov2_vitb14", 0)
    got = net.forward_patch_tokens(img.cuda()).cpu()
    net.close()
    assert got.shape == (1, 85 * 114, 768)
    err = (got - ref).abs().flatten() / ref.abs().max().item()
    # north-star tolerance for one block (1e-3) on all but a 1e-5 fraction of the 7.4 M values; the
    # extreme tail is a bf16 rounding flip of an intermediate (1 ulp = 3.9e-3 of that value)
    assert torch.quantile(err[::7], 0.9999).item() < 1e-3 and err.max().item() < 2.5e-3, err.max().item()


@pytest.mark.parametrize("shape", [(1370, 768, 768), (43840 // 8, 768, 3072), (300, 384, 384), (129, 1024, 1024),
                                   (5000, 384, 1536)])
def test_linear_with_fused_layernorm(lib_handle, shape):
    """vc_vit_linear_ln: the residual GEMM whose epilogue also normalises every 128-row block it has
    completed (per-block counter, the completing CTA reads the rows back from L2).  The fp32 stream must
    equal the plain residual epilogue bit for bit, and the bf16 LayerNorm rows must equal the stand-alone
    LayerNorm kernel applied to that stream bit for bit (same arithmetic), for row counts that leave a
    partial last block, several column tiles per row (C = 768 / 1024: 3 / 4 tiles whose completion order is
    not fixed) and a single one (C = 384 / 256-wide tile); twice in a row (the counters must be left zero)."""
    import torch
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    m, n, k = shape
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    a = torch.randn((m, k), generator=g, device="cuda").to(torch.bfloat16)
    w = (torch.randn((n, k), generator=g, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn((n,), generator=g, device="cuda")
    gamma = torch.randn((n,), generator=g, device="cuda")
    ln_w = torch.randn((n,), generator=g, device="cuda")
    ln_b = torch.randn((n,), generator=g, device="cuda")
    x0 = torch.randn((m, n), generator=g, device="cuda") * 2 + 0.3
    for _ in range(2):
        x_ref = x0.clone()
        _lib.check(lib.vc_vit_linear(h, a.data_ptr(), w.data_ptr(), bias.data_ptr(), gamma.data_ptr(), m, n, k, 2,
                                     x_ref.data_ptr(), _stream()))
        ln_ref = torch.zeros((m, n), dtype=torch.bfloat16, device="cuda")
        _lib.check(lib.vc_vit_layernorm(h, x_ref.data_ptr(), ln_w.data_ptr(), ln_b.data_ptr(), m, n, ln_ref.data_ptr(),
                                        _stream()))
        x = x0.clone()
        ln = torch.full((m, n), float("nan"), dtype=torch.bfloat16, device="cuda")
        _lib.check(lib.vc_vit_linear_ln(h, a.data_ptr(), w.data_ptr(), bias.data_ptr(), gamma.data_ptr(),
                                        ln_w.data_ptr(), ln_b.data_ptr(), m, n, k, x.data_ptr(), ln.data_ptr(), _stream()))
        torch.cuda.synchronize()
        assert torch.equal(x, x_ref)
        assert torch.equal(ln.view(torch.int16), ln_ref.view(torch.int16))
    ref = torch.nn.functional.layer_norm(x0 + gamma * (a.float() @ w.float().t() + bias), (n,), ln_w, ln_b, eps=1e-6)
    assert (ln.float() - ref).abs().max().item() <= 2 ** -7 * ref.abs().max().item()
    with pytest.raises(Exception):   # the normalised rows may not overwrite the operand other CTAs still read
        _lib.check(lib.vc_vit_linear_ln(h, a.data_ptr(), w.data_ptr(), bias.data_ptr(), gamma.data_ptr(),
                                        ln_w.data_ptr(), ln_b.data_ptr(), m, n, k, x.data_ptr(), a.data_ptr(), _stream()))


def test_forward_fused_layernorm_equals_separate_launches(monkeypatch):
    """vc_vit_forward with one launch per LayerNorm (default) and with the LayerNorms in the GEMM epilogues
    (VC_VIT_FUSED_LN=1, kept as an option: measured slower): identical tokens, bit for bit, at a grid with a
    partial last row block."""
    import torch
    from oracle import vit_oracle as vo
    from vit_colmap_b200.vit import VitBackbone
    sd = vo.synthetic_state_dict("dinov2_vitb14", seed=5, depth=3)
    rng = np.random.default_rng(5)
    img = torch.from_numpy(rng.integers(0, 256, (3, 518, 518, 3), dtype=np.uint8)).cuda()
    net = VitBackbone(sd, "dinov2_vitb14", 0)
    separate = net.forward_patch_tokens(img).clone()
    monkeypatch.setenv("VC_VIT_FUSED_LN", "1")
    fused = net.forward_patch_tokens(img).clone()
    # third placement: a LayerNorm kernel that follows the residual GEMM's row blocks while the GEMM still runs
    # (VC_VIT_LN=follow, programmatic dependent launch + per-block counters; also measured slower)
    monkeypatch.delenv("VC_VIT_FUSED_LN")
    monkeypatch.setenv("VC_VIT_LN", "follow")
    follow = [net.forward_patch_tokens(img).clone() for _ in range(2)]   # twice: the counters must come back to zero
    torch.cuda.synchronize()
    net.close()
    assert torch.equal(fused, separate) and torch.isfinite(fused).all()
    assert torch.equal(follow[0], separate) and torch.equal(follow[1], separate)
