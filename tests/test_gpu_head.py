"""GPU parity tests of the ViTFeatureModel convolutional head (through the C ABI).
Per-op tests compare each implicit-GEMM convolution with a plain PyTorch fp32 reference of the same
op on the same bf16-rounded operands; the whole head is compared with oracle/head_oracle.py (which
tests/test_oracle_head.py pins to the reference's own ViTFeatureModel) and with the reference's
golden outputs directly.  Tolerances are written in each test."""
import ctypes
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN = sorted((Path(__file__).parent / "golden").glob("head_*.npz"))


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


def _nhwc(t):  # [B,C,H,W] -> channels-last contiguous
    return t.permute(0, 2, 3, 1).contiguous()


# (B, H, W, Cin, Cout): tile multiples, ragged edges, tiny maps, N tiles of 256/192/64/32
CONV_SHAPES = [(2, 8, 16, 64, 256), (1, 37, 37, 384, 512), (3, 5, 7, 128, 192), (1, 74, 74, 512, 512),
               (2, 13, 31, 256, 64), (1, 1, 1, 64, 32), (1, 129, 129, 256, 192)]


@pytest.mark.parametrize("shape", CONV_SHAPES, ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("epilogue", [0, 1, 2])
def test_conv3x3_matches_fp32_reference(lib_handle, shape, epilogue):
    import torch
    import torch.nn.functional as F
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.head import pack_conv3x3
    lib, h = lib_handle
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device="cuda").manual_seed(sum(shape) + epilogue)
    x = torch.randn((B, Cin, H, W), generator=g, device="cuda").to(torch.bfloat16)
    w = (torch.randn((Cout, Cin, 3, 3), generator=g, device="cuda") / (3 * Cin ** 0.5)).to(torch.bfloat16)
    scale = 1 + 0.2 * torch.randn((Cout,), generator=g, device="cuda")
    shift = torch.randn((Cout,), generator=g, device="cuda")
    acc = F.conv2d(x.float(), w.float(), None, padding=1)
    if epilogue == 1:
        ref = F.gelu(acc * scale.view(1, -1, 1, 1) + shift.view(1, -1, 1, 1))
    else:
        ref = acc + shift.view(1, -1, 1, 1)
    out = torch.full((B, H, W, Cout), float("nan"), dtype=torch.float32 if epilogue == 2 else torch.bfloat16, device="cuda")
    wp = pack_conv3x3(w).contiguous()
    _lib.check(lib.vc_conv2d(h, _nhwc(x).data_ptr(), B, H, W, Cin, wp.data_ptr(), Cout, 0, epilogue,
                             scale.data_ptr(), shift.data_ptr(), out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    err = (out.float() - _nhwc(ref)).abs().max().item()
    s = ref.abs().max().item()
    # exact bf16 products, fp32 accumulation; bf16 outputs carry one rounding (2^-9 relative)
    tol = 2e-5 * s if epilogue == 2 else 2 ** -8 * s
    assert err <= tol, (err, s)


@pytest.mark.parametrize("shape", [(1, 37, 37, 384, 512), (3, 5, 7, 128, 192), (1, 1, 1, 64, 32), (2, 13, 31, 256, 64),
                                   (1, 17, 16, 64, 256)], ids=lambda s: "x".join(map(str, s)))
@pytest.mark.parametrize("kind,epilogue", [(0, 1), (0, 2), (1, 0)])
def test_two_sm_conv_equals_one_sm(lib_handle, monkeypatch, shape, kind, epilogue):
    """The cta_group::2 convolution (conv_gemm2.cu, default) and the one-SM kernel (VC_CONV_2CTA=0) share the per-tile
    epilogue and the accumulation order: identical outputs, bit for bit -- also with an odd number of 8 x 16 patches
    (the last pair's second CTA works on nothing) and channel blocks of 192 / 64 / 32."""
    import torch
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.head import pack_conv3x3, pack_deconv4x4s2
    lib, h = lib_handle
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device="cuda").manual_seed(sum(shape) + kind)
    x = _nhwc(torch.randn((B, Cin, H, W), generator=g, device="cuda").to(torch.bfloat16))
    if kind == 0:
        wp = pack_conv3x3((torch.randn((Cout, Cin, 3, 3), generator=g, device="cuda") / (3 * Cin ** 0.5)).to(torch.bfloat16))
        oh, ow = H, W
    else:
        wp = pack_deconv4x4s2((torch.randn((Cin, Cout, 4, 4), generator=g, device="cuda") / (2 * Cin ** 0.5)).to(torch.bfloat16))
        oh, ow = 2 * H, 2 * W
    wp = wp.contiguous()
    scale = 1 + 0.2 * torch.randn((Cout,), generator=g, device="cuda")
    shift = torch.randn((Cout,), generator=g, device="cuda")
    outs = []
    for env in ("1", "0"):
        monkeypatch.setenv("VC_CONV_2CTA", env)
        out = torch.full((B, oh, ow, Cout), float("nan"), dtype=torch.float32 if epilogue == 2 else torch.bfloat16, device="cuda")
        _lib.check(lib.vc_conv2d(h, x.data_ptr(), B, H, W, Cin, wp.data_ptr(), Cout, kind, epilogue,
                                 scale.data_ptr() if epilogue == 1 else None, shift.data_ptr(), out.data_ptr(), _stream()))
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.isfinite(outs[0].float()).all() and torch.equal(outs[0], outs[1])


@pytest.mark.parametrize("shape", [(2, 8, 16, 64, 256), (1, 37, 37, 768, 512), (3, 5, 7, 128, 64), (1, 74, 74, 512, 512),
                                   (1, 3, 3, 384, 512)], ids=lambda s: "x".join(map(str, s)))
def test_deconv_matches_conv_transpose2d(lib_handle, shape):
    import torch
    import torch.nn.functional as F
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.head import pack_deconv4x4s2
    lib, h = lib_handle
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device="cuda").manual_seed(sum(shape))
    x = torch.randn((B, Cin, H, W), generator=g, device="cuda").to(torch.bfloat16)
    w = (torch.randn((Cin, Cout, 4, 4), generator=g, device="cuda") / (2 * Cin ** 0.5)).to(torch.bfloat16)
    bias = torch.randn((Cout,), generator=g, device="cuda")
    ref = F.conv_transpose2d(x.float(), w.float(), bias, stride=2, padding=1)
    out = torch.full((B, 2 * H, 2 * W, Cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    wp = pack_deconv4x4s2(w).contiguous()
    _lib.check(lib.vc_conv2d(h, _nhwc(x).data_ptr(), B, H, W, Cin, wp.data_ptr(), Cout, 1, 0, None,
                             bias.data_ptr(), out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    err = (out.float() - _nhwc(ref)).abs().max().item()
    s = ref.abs().max().item()
    assert err <= 2 ** -8 * s, (err, s)


@pytest.mark.parametrize("shape", [(2, 14, 17, 128, 128), (1, 129, 129, 128, 768), (1, 9, 9, 128, 384), (1, 4, 4, 64, 32)],
                         ids=lambda s: "x".join(map(str, s)))
def test_conv1x1_f32_matches_reference(lib_handle, shape):
    import torch
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    B, H, W, Cin, Cout = shape
    g = torch.Generator(device="cuda").manual_seed(sum(shape))
    x = torch.randn((B, H, W, Cin), generator=g, device="cuda").to(torch.bfloat16)
    w = (torch.randn((Cout, Cin), generator=g, device="cuda") / Cin ** 0.5).to(torch.bfloat16)
    bias = torch.randn((Cout,), generator=g, device="cuda")
    ref = x.float() @ w.float().t() + bias
    out = torch.full((B, H, W, Cout), float("nan"), dtype=torch.float32, device="cuda")
    _lib.check(lib.vc_conv2d(h, x.data_ptr(), B, H, W, Cin, w.data_ptr(), Cout, 2, 2, None, bias.data_ptr(),
                             out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    assert (out - ref).abs().max().item() <= 2e-5 * ref.abs().max().item()


@pytest.mark.parametrize("size", [((8, 8), (7, 7)), ((148, 148), (129, 129)), ((16, 20), (14, 17)), ((36, 24), (31, 21)),
                                  ((5, 5), (9, 11))])
def test_resize_matches_interpolate(lib_handle, size):
    import torch
    import torch.nn.functional as F
    from vit_colmap_b200 import _lib
    lib, h = lib_handle
    (hi, wi), (ho, wo) = size
    x = torch.randn((2, 64, hi, wi), generator=torch.Generator(device="cuda").manual_seed(hi), device="cuda").to(torch.bfloat16)
    ref = F.interpolate(x.float(), size=(ho, wo), mode="bilinear", align_corners=False)
    out = torch.empty((2, ho, wo, 64), dtype=torch.bfloat16, device="cuda")
    _lib.check(lib.vc_resize_bilinear(h, _nhwc(x).data_ptr(), 2, hi, wi, ho, wo, 64, out.data_ptr(), _stream()))
    torch.cuda.synchronize()
    # fp32 interpolation of bf16 inputs, one bf16 rounding on the way out
    assert (out.float() - _nhwc(ref)).abs().max().item() <= 2 ** -8 * ref.abs().max().item()


def _head_case(C, d, H, W, seed, B=1):
    import torch
    from oracle import head_oracle as ho
    sd = ho.synthetic_head_state_dict(C, d, seed)
    rng = np.random.default_rng(seed + 1000)
    feat = torch.from_numpy(rng.standard_normal((B, C, H, W)).astype(np.float32))
    return sd, feat


@pytest.mark.parametrize("case", [(384, 128, 4, 5, 11, 1), (768, 64, 3, 3, 12, 1), (384, 32, 9, 6, 13, 2),
                                  (768, 128, 37, 37, 14, 1), (1024, 384, 8, 11, 15, 2)],
                         ids=lambda c: "C{}_d{}_{}x{}".format(*c[:4]))
def test_head_matches_oracle(case):
    """Whole head vs the oracle with bf16 rounding emulated at the same points: what remains is
    fp32 summation order plus the occasional bf16 rounding flip it causes, so maps agree to a few
    bf16 ulps of their range."""
    import torch
    from oracle import head_oracle as ho
    from vit_colmap_b200.head import ConvHead
    C, d, H, W, seed, B = case
    sd, feat = _head_case(C, d, H, W, seed, B)
    ref = ho.head_forward(sd, feat, emulate_bf16=True)
    head = ConvHead(sd, C, d, 0)
    tokens = feat.permute(0, 2, 3, 1).reshape(B, H * W, C).cuda()
    out = head.forward(tokens, (H, W), want_features=True)
    torch.cuda.synchronize()
    kp = out["keypoints"].permute(0, 3, 1, 2).cpu()
    desc = out["descriptors"].permute(0, 3, 1, 2).cpu()
    ft = out["features"].float().permute(0, 3, 1, 2).cpu()
    assert kp.shape == ref["keypoints"].shape and desc.shape == ref["descriptors"].shape
    assert (ft - ref["features"]).abs().max().item() <= 2 ** -6 * ref["features"].abs().max().item()
    assert (kp - ref["keypoints"]).abs().max().item() <= 2e-2 * ref["keypoints"].abs().max().item()
    assert (desc - ref["descriptors"]).abs().max().item() <= 6e-3  # unit-norm rows; a flipped bf16 rounding of a hidden value moves a component by ~2^-8
    assert torch.allclose(desc.norm(dim=1), torch.ones_like(desc[:, 0]), atol=1e-5)
    # bf16 tokens in, sparse outputs: same keypoint map, hidden map instead of dense descriptors
    out2 = head.forward(tokens.to(torch.bfloat16), (H, W), dense_descriptors=False)
    torch.cuda.synchronize()
    assert torch.equal(out2["keypoints"], out["keypoints"])
    assert out2["desc_hidden"].shape == (B, (14 * H) // 4, (14 * W) // 4, 128)
    head.close()


@pytest.mark.parametrize("path", GOLDEN, ids=lambda p: p.stem)
def test_head_vs_reference_golden(path):
    """Against the reference's own fp32 ViTFeatureModel outputs: the gap is the bf16 operand rounding
    of a 7-convolution stack (the oracle's own bf16-vs-fp32 gap, tests/test_oracle_head.py)."""
    import torch
    from vit_colmap_b200.head import ConvHead
    g = np.load(path)
    C, d, H, W, seed = (int(v) for v in g["meta"])
    sd, feat = _head_case(C, d, H, W, seed)
    head = ConvHead(sd, C, d, 0)
    out = head.forward(feat.permute(0, 2, 3, 1).reshape(1, H * W, C).cuda(), (H, W))
    torch.cuda.synchronize()
    kp = out["keypoints"].permute(0, 3, 1, 2).cpu().numpy()
    desc = out["descriptors"].permute(0, 3, 1, 2).cpu().numpy()
    assert np.abs(kp - g["keypoints"]).max() < 0.1
    cos = (desc * g["descriptors"].astype(np.float32)).sum(axis=1)
    assert cos.min() > 0.999
    head.close()


def test_head_target_size_equal_skips_resize():
    """target_size == 4x grid: the resize branch of vit_feature_model.py:205 is not taken."""
    import torch
    from oracle import head_oracle as ho
    from vit_colmap_b200.head import ConvHead
    sd, feat = _head_case(384, 32, 4, 4, 21)
    ref = ho.head_forward(sd, feat, target_size=(16, 16), emulate_bf16=True)
    head = ConvHead(sd, 384, 32, 0)
    out = head.forward(feat.permute(0, 2, 3, 1).reshape(1, 16, 384).cuda(), (4, 4), target_size=(16, 16))
    torch.cuda.synchronize()
    kp = out["keypoints"].permute(0, 3, 1, 2).cpu()
    assert (kp - ref["keypoints"]).abs().max().item() <= 2e-2 * ref["keypoints"].abs().max().item()
    head.close()


def _oracle_trainable(sd_b, sd_h, name, img_rgb_u8, K, thr, r, d, orig_wh, resized_wh):
    """Oracle pipeline for one image: ViT oracle (bf16 emulation) -> head oracle (bf16 emulation) ->
    the reference's selection / packing restatement."""
    import torch
    from oracle import extract_oracle as eo
    from oracle import head_oracle as ho
    from oracle import vit_oracle as vo
    heads = vo.CONFIGS[name][2]
    x = vo.preprocess(torch.from_numpy(img_rgb_u8[None]))
    tok = vo.forward_features(sd_b, x, heads, emulate_bf16=True)["x_norm_patchtokens"]
    gh, gw = img_rgb_u8.shape[0] // 14, img_rgb_u8.shape[1] // 14
    feat = tok.reshape(1, gh, gw, -1).permute(0, 3, 1, 2)
    out = ho.head_forward(sd_h, feat, emulate_bf16=True)
    return eo.trainable_select(out["keypoints"][0], out["descriptors"][0], orig_wh, resized_wh, K, thr, r), out


def test_trainable_extractor_matches_oracle_pipeline():
    """TrainableViTExtractor._run_inference (backbone + head + selection + sparse descriptor
    projection) vs the chained oracles on one 112x140 image.  Both pipelines are bf16, so score maps
    differ by rounding noise: keypoints are compared as sets of cells where the oracle's score clears
    the NMS/threshold decisions by a margin, and descriptors of common cells within 2 u8 steps."""
    import torch
    from oracle import head_oracle as ho
    from oracle import vit_oracle as vo
    from vit_colmap_b200.features.trainable_vit_extractor import TrainableViTExtractor
    name, d, K, thr, r = "dinov2_vits14", 64, 200, 0.3, 2
    sd_b = vo.synthetic_state_dict(name, seed=3, depth=2)
    sd_h = ho.synthetic_head_state_dict(384, d, 7)
    rng = np.random.default_rng(5)
    img_bgr = rng.integers(0, 256, (112, 140, 3), dtype=np.uint8)
    ex = TrainableViTExtractor(model_name=name, num_keypoints=K, descriptor_dim=d, score_threshold=thr,
                               nms_radius=r, state_dict=sd_b, head_state_dict=sd_h)
    kps, du8 = ex._run_inference(img_bgr)
    assert kps.dtype == np.float32 and kps.shape[1] == 6 and du8.dtype == np.uint8 and du8.shape == (len(kps), d)
    assert 0 < len(kps) <= K
    assert np.all(kps[:, 2] == 1.0) and np.all(kps[:, 5] == 0.0)
    assert np.all(np.diff(kps[:, 4]) <= 0)  # descending score
    (rk, rd, rc), out = _oracle_trainable(sd_b, sd_h, name, img_bgr[:, :, ::-1].copy(), K, thr, r, d, (140, 112), (140, 112))
    # the two bf16 pipelines agree on the dense maps ...
    gh, gw = 8, 10
    tokens = ex.backbone.forward_patch_tokens(torch.from_numpy(img_bgr[None]).cuda(), bgr=True)
    maps = ex.head.forward(tokens, (gh, gw))
    torch.cuda.synchronize()
    kp_map = maps["keypoints"][0].permute(2, 0, 1).cpu()
    assert (kp_map - out["keypoints"][0]).abs().max().item() < 0.05 * out["keypoints"][0].abs().max().item()
    cos = (maps["descriptors"][0].permute(2, 0, 1).cpu() * out["descriptors"][0]).sum(dim=0)
    assert cos.min().item() > 0.999
    # ... and therefore on most keypoints; descriptors of the common ones match to quantisation noise
    # (cells are paired through their sub-pixel coordinates: same cell <=> (x, y) within 0.25 px)
    dist = np.abs(kps[:, None, :2] - rk[None, :, :2]).max(axis=2)
    pairs = [(i, int(dist[i].argmin())) for i in range(len(kps)) if dist[i].min() < 0.25]
    assert len(pairs) >= 0.8 * min(len(kps), len(rk)), (len(pairs), len(kps), len(rk))
    diffs = [np.abs(du8[i].astype(int) - rd[j].astype(int)).max() for i, j in pairs]
    assert np.percentile(diffs, 90) <= 2
    assert np.abs(kps[[i for i, _ in pairs], 3:5] - rk[[j for _, j in pairs], 3:5]).max() < 0.05
    ex.backbone.close()


def test_trainable_sparse_projection_equals_dense_path():
    """Projecting only the selected keypoints (vc_sample_pack with the 1x1 weights) gives the same
    descriptors as gathering from the dense normalised map: same keypoints, u8 within one step."""
    import torch
    from oracle import head_oracle as ho
    from vit_colmap_b200.extraction import PostProcessor, trainable_select_pack
    from vit_colmap_b200.head import ConvHead
    sd, feat = _head_case(384, 128, 6, 7, 31, B=2)
    head = ConvHead(sd, 384, 128, 0)
    pp = PostProcessor(0, handle=head.handle)
    tokens = feat.permute(0, 2, 3, 1).reshape(2, 42, 384).cuda()
    dense = head.forward(tokens, (6, 7))
    sparse = head.forward(tokens, (6, 7), dense_descriptors=False)
    P, b = head.desc_projection
    a = trainable_select_pack(pp, dense["keypoints"], dense["descriptors"], (98, 84), (98, 84), 150, 0.2, 1, want_f32=True)
    s = trainable_select_pack(pp, sparse["keypoints"], sparse["desc_hidden"], (98, 84), (98, 84), 150, 0.2, 1,
                              want_f32=True, projection=P, projection_bias=b)
    torch.cuda.synchronize()
    assert torch.equal(a[4], s[4]) and torch.equal(a[0], s[0])
    n = a[4].cpu().numpy()
    assert n.min() > 10
    for i in range(2):
        assert (a[3][i, :n[i]] - s[3][i, :n[i]]).abs().max().item() < 2e-5
        assert (a[1][i, :n[i]].int() - s[1][i, :n[i]].int()).abs().max().item() <= 1
    head.close()
