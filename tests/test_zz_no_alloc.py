"""sf_dcn_forward_static_f32: the DCN forward with its scratch in a static device array (no allocation inside the
call; include/stereoformer_b200.h, csrc/dcn_static.cu).  No counterpart in the reference: it is checked against the
existing entry point sf_dcn_forward_ws_f32 (caller-provided scratch) through the C ABI, on two calls in a row that
use DIFFERENT weights and shapes, so a stale or misaddressed scratch array cannot pass.

Collected last on purpose."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 3e-5      # relative to max |out|: the fast-path DCN tolerance of tests/test_gpu_precision.py (same kernels: 0 expected)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _case(seed, B, Cin, Cout, H, W, groups, dg, with_mask):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, Cin, H, W, generator=g).cuda()
    w = (torch.randn(Cout, Cin // groups, 3, 3, generator=g) * 0.1).cuda()
    b = torch.randn(Cout, generator=g).cuda()
    off = (torch.randn(B, dg * 18, H, W, generator=g) * 2).cuda()
    msk = torch.rand(B, dg * 9, H, W, generator=g).cuda() if with_mask else None
    return x, w, b, off, msk, groups, dg


def _shape_args(x, w, groups, dg):
    B, Cin, H, W = x.shape
    return (B, Cin, H, W, w.shape[0], 3, 3, 1, 1, 1, 1, 1, 1, groups, dg)


def _run_ws(L, ops, case):
    x, w, b, off, msk, groups, dg = case
    out = torch.full((x.shape[0], w.shape[0], x.shape[2], x.shape[3]), float("nan"), device="cuda")
    n = int(L.lib().sf_dcn_workspace_bytes(x.shape[1], w.shape[0], 3, 3, groups))
    assert n > 0, "the case must take the tensor-core path (the one that uses scratch)"
    ws = torch.empty(n, device="cuda", dtype=torch.uint8)
    L.call("sf_dcn_forward_ws_f32", _p(x), _p(w), _p(b), _p(off), _p(msk), _p(out), *_shape_args(x, w, groups, dg),
           _p(ws), ctypes.c_size_t(n), ops.SF_ALGO_TC, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out


def _run_static(L, ops, case, out):
    x, w, b, off, msk, groups, dg = case
    L.call("sf_dcn_forward_static_f32", _p(x), _p(w), _p(b), _p(off), _p(msk), _p(out),
           *_shape_args(x, w, groups, dg), ops.SF_ALGO_TC,
           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    from stereoformer_b200 import _lib as L
    from stereoformer_b200 import ops
    assert int(L.lib().sf_dcn_static_scratch_bytes()) == 4 * 1024 * 1024
    cases = [_case(11, 2, 128, 128, 40, 56, 1, 1, True),        # 1.1 MiB of scratch
             _case(12, 1, 64, 128, 33, 47, 2, 2, False),        # other weights, groups, DCNv1, ragged tile: 0.3 MiB
             _case(11, 2, 128, 128, 40, 56, 1, 1, True)]        # the first one again after the array held other weights
    wants = [_run_ws(L, ops, c) for c in cases]
    outs = [torch.full_like(w, float("nan")) for w in wants]
    torch.cuda.synchronize()
    free_before = None
    for i, (c, want, out) in enumerate(zip(cases, wants, outs)):
        if i == 1:                                              # kernels are loaded by now: the call must not allocate
            free_before = torch.cuda.mem_get_info()[0]
        got = _run_static(L, ops, c, out)
        if i == 1:
            assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated inside the call"
        assert torch.isfinite(got).all()
        scale = want.abs().max().item()
        err = (got - want).abs().max().item()
        print("static-scratch DCN call %d: max |diff| %.3e (max |out| %.2f)" % (i, err, scale))
        assert err <= TOL * scale


def test_static_scratch_rejects_oversized_shapes():
    from stereoformer_b200 import _lib as L
    from stereoformer_b200 import ops
    Cin, Cout = 512, 256                                        # 2 * 256 * 4608 * 2 B = 4.5 MiB > 4 MiB
    assert int(L.lib().sf_dcn_workspace_bytes(Cin, Cout, 3, 3, 1)) > int(L.lib().sf_dcn_static_scratch_bytes())
    x, w, b, off, msk, groups, dg = _case(13, 1, Cin, Cout, 8, 8, 1, 1, True)
    out = torch.zeros(1, Cout, 8, 8, device="cuda")
    with pytest.raises(ValueError, match="SF_DCN_STATIC_SCRATCH_BYTES"):
        L.call("sf_dcn_forward_static_f32", _p(x), _p(w), _p(b), _p(off), _p(msk), _p(out),
               *_shape_args(x, w, groups, dg), ops.SF_ALGO_AUTO,
               ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert float(out.abs().max()) == 0.0                        # nothing was launched
