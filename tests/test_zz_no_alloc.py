"""GPU test of the allocation-free score-map entry point (vc_score_map_static, include/vc_b200.h).

Collected last on purpose.  The entry point runs the same two
kernels as vc_score_map with its scratch in a file-scope ``static __device__`` array, so the results
must be identical bit for bit; two calls in a row share that array and must not see each other's
left-overs.  Passes alone in a fresh process and after the other tests.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 0.0  # same kernels, same order of operations: bit-exact


@pytest.fixture(scope="module")
def env():
    import torch
    from vit_colmap_b200 import _lib
    lib = _lib.load()
    h = ctypes.c_void_p()
    _lib.check(lib.vc_create(0, ctypes.byref(h)))
    yield torch, _lib, lib, h
    lib.vc_destroy(h)


def _stream(torch):
    return int(torch.cuda.current_stream().cuda_stream)


def _with_workspace(env, tokens, H, W, method):
    torch, _lib, lib, h = env
    B, _, C = tokens.shape
    score = torch.full((B, H, W), -7.0, dtype=torch.float32, device="cuda")
    nbytes = int(lib.vc_score_workspace_bytes(B, H, W))
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    dt = _lib.DTYPE_BF16 if tokens.dtype == torch.bfloat16 else _lib.DTYPE_F32
    _lib.check(lib.vc_score_map(h, tokens.data_ptr(), dt, B, H, W, C, H * W * C, method, score.data_ptr(),
                                ws.data_ptr(), nbytes, _stream(torch)))
    return score


def _static(env, tokens, H, W, method, score):
    torch, _lib, lib, h = env
    B, _, C = tokens.shape
    dt = _lib.DTYPE_BF16 if tokens.dtype == torch.bfloat16 else _lib.DTYPE_F32
    return lib.vc_score_map_static(h, tokens.data_ptr(), dt, B, H, W, C, H * W * C, method, score.data_ptr(),
                                   _stream(torch))


def _tokens(torch, seed, B, H, W, C, dtype):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, H * W, C, generator=g).to(dtype).cuda()


def test_static_scratch_matches_workspace_entry_point_on_two_calls_in_a_row(env):
    torch, _lib, lib, h = env
    # call 1: a large "combined" job (uses all five scratch planes); call 2: a smaller harris job whose
    # scratch region overlaps what call 1 left behind; call 3: dog, bf16 tokens
    cases = [(2, 85, 114, 768, torch.float32, 2), (3, 37, 37, 384, torch.float32, 0),
             (5, 42, 57, 768, torch.bfloat16, 1)]
    inputs = [(_tokens(torch, 100 + i, B, H, W, C, dt), H, W, m) for i, (B, H, W, C, dt, m) in enumerate(cases)]
    expect = [_with_workspace(env, *a) for a in inputs]
    outs = [torch.full_like(e, -7.0) for e in expect]
    torch.cuda.synchronize()
    for rep in range(2):  # the whole sequence twice: every call follows a different job's left-overs
        for a, out in zip(inputs, outs):
            out.fill_(-7.0)
            _lib.check(_static(env, *a, out))
        torch.cuda.synchronize()
        for e, out in zip(expect, outs):
            assert np.abs(out.cpu().numpy() - e.cpu().numpy()).max() <= TOL
            assert torch.equal(out, e)
    # against the CPU oracle as well, within the tolerance of tests/test_gpu_extract.py
    from oracle import extract_oracle as eo
    tok, H, W, m = inputs[1]
    fm = tok[0].cpu().reshape(H, W, -1).permute(2, 0, 1).contiguous()
    ref = eo.distinctiveness(fm, "harris").numpy()
    assert np.abs(outs[1][0].cpu().numpy() - ref).max() < 2e-6


def test_static_call_allocates_no_device_memory(env):
    torch, _lib, lib, h = env
    tok = _tokens(torch, 7, 4, 37, 37, 384, torch.float32)
    out = torch.empty(4, 37, 37, dtype=torch.float32, device="cuda")
    _lib.check(_static(env, tok, 37, 37, 2, out))  # first use loads the module and looks the symbol up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        _lib.check(_static(env, tok, 37, 37, 2, out))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_static_rejects_sizes_beyond_the_array(env):
    torch, _lib, lib, h = env
    B, H, W, C = 22, 85, 114, 384  # 22 * 5 * 9690 floats > 2^20; 21 images still fit
    tok = torch.zeros(B, H * W, C, dtype=torch.bfloat16, device="cuda")
    out = torch.full((B, H, W), -7.0, dtype=torch.float32, device="cuda")
    rc = _static(env, tok, H, W, 0, out)
    assert rc != 0
    msg = lib.vc_last_error().decode()
    assert "static array holds 4194304 B" in msg and "vc_score_map" in msg
    torch.cuda.synchronize()
    assert bool((out == -7.0).all())  # nothing was launched
    ok = _static(env, tok[:21], H, W, 0, out[:21])
    assert ok == 0, lib.vc_last_error()
    torch.cuda.synchronize()
