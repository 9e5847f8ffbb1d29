"""a10 pinned on the REFERENCE'S OWN compiled extension (oracle/_ref/native/DCN.so, built from the unmodified
deform/src sources for sm_100a): forward and backward of the drop-in against it, random offsets and masks."""
import pytest
import torch

from oracle import refenv

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refenv.available()["dcn"],
                                                  reason="reference DCN extension not built")]


def _ref():
    refenv.activate(dcn="reference")
    import sys
    import stereoformer_b200.DCN as ours
    mod = sys.modules["DCN"]
    if mod is ours:                       # an earlier test installed the drop-in under the name
        import importlib.util
        import os
        so = os.path.join(refenv.NATIVE_DST, "DCN.so")
        spec = importlib.util.spec_from_file_location("DCN", so)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    return mod


CASES = [  # B, Cin, Cout, H, W, groups, dg, stride, pad, dil
    (2, 4, 4, 4, 4, 2, 1, 1, 1, 1),            # deform/test.py:15-18 shape
    (2, 16, 16, 34, 60, 1, 1, 1, 1, 1),        # the shipped 16->16 layer (dcn_small_kernel)
    (1, 64, 128, 40, 56, 1, 2, 1, 1, 1),       # deform/test.py:526 (tcgen05 path)
    (1, 128, 128, 31, 45, 1, 1, 1, 1, 1),      # disp_residual.py:213 (tcgen05 path)
    (2, 32, 48, 23, 37, 1, 4, 2, 1, 1),        # stride 2
    (1, 24, 24, 20, 33, 3, 3, 1, 2, 2),        # groups, dilation
    # channels-last gather kernel (dcn_tc2.cu: channels per group and per deformable group multiples of 64)
    (2, 128, 64, 19, 27, 2, 1, 1, 1, 1),       # two groups of 64 channels, ragged 8x16 tiles, batch 2
    (1, 128, 128, 17, 33, 1, 2, 2, 2, 2),      # two deformable groups, stride 2, dilation 2
    (1, 64, 256, 9, 50, 1, 1, 1, 1, 1),        # 256 output channels (full accumulator width)
    (1, 192, 96, 16, 16, 1, 3, 1, 0, 1),       # three 64-channel blocks / deformable groups, no padding, ragged N
    (2, 64, 32, 21, 30, 2, 2, 1, 1, 1),        # 32 channels per group: two taps per chunk, the ninth tap shares one with padding
]
TC2 = CASES[-5:]


@pytest.mark.parametrize("case", CASES)
def test_forward_matches_reference_extension(case):
    import stereoformer_b200.DCN as ours
    ref = _ref()
    B, Cin, Cout, H, W, G, dg, s, p, d = case
    torch.manual_seed(3)
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin // G, 3, 3, device="cuda") * 0.1
    b = torch.randn(Cout, device="cuda")
    Ho = (H + 2 * p - (d * 2 + 1)) // s + 1
    Wo = (W + 2 * p - (d * 2 + 1)) // s + 1
    off = torch.randn(B, dg * 18, Ho, Wo, device="cuda") * 2
    msk = torch.sigmoid(torch.randn(B, dg * 9, Ho, Wo, device="cuda"))
    args = (3, 3, s, s, p, p, d, d, G, dg, 64)
    want = ref.modulated_deform_conv_forward(x, w, b, off, msk, *args)
    got = ours.modulated_deform_conv_forward(x, w, b, off, msk, *args)
    assert got.shape == want.shape
    scale = want.abs().max()
    assert (got - want).abs().max() <= 1e-4 * scale, float((got - want).abs().max() / scale)
    want1 = ref.deform_conv_forward(x, w, b, off, *args)
    got1 = ours.deform_conv_forward(x, w, b, off, *args)
    assert (got1 - want1).abs().max() <= 1e-4 * want1.abs().max()


def test_channels_last_cases_take_the_channels_last_kernel():
    """The workspace query is what routes a call: the cases meant for dcn_tc2 must be eligible for it."""
    from stereoformer_b200 import _lib
    for B, Cin, Cout, H, W, G, dg, s, p, d in TC2 + [CASES[2], CASES[3]]:
        small = int(_lib.lib().sf_dcn_workspace_bytes(Cin, Cout, 3, 3, G))
        big = int(_lib.lib().sf_dcn_workspace_bytes_ex(B, Cin, H, W, Cout, 3, 3, G, dg))
        assert big >= small + B * Cin * H * W * 4, (Cin, Cout, G, dg)
    assert int(_lib.lib().sf_dcn_workspace_bytes_ex(2, 16, 34, 60, 16, 3, 3, 1, 1)) == \
        int(_lib.lib().sf_dcn_workspace_bytes(16, 16, 3, 3, 1))            # 16 channels per group: the specialised FFMA kernel


BWD_TC = [  # shapes that take the channels-last backward kernels (output width a multiple of 4)
    (1, 128, 128, 24, 36, 1, 1, 1, 1, 1),      # one 128-row block of output channels, ragged 8x16 tiles
    (2, 64, 64, 16, 20, 1, 1, 1, 1, 1),        # batch 2, a single 64-channel block
    (1, 128, 64, 18, 28, 2, 2, 1, 1, 1),       # two convolution groups = two deformable groups (offset / mask sums per group)
    (1, 64, 256, 9, 52, 1, 1, 1, 1, 1),        # two 128-row blocks of output channels
    (1, 128, 128, 33, 47, 1, 1, 2, 1, 1),      # stride 2
    (1, 32, 32, 12, 16, 1, 1, 1, 2, 2),        # 32 channels per group (two taps per chunk), dilation 2
]


@pytest.mark.parametrize("case", CASES[:3] + CASES[4:6] + BWD_TC)
def test_backward_matches_reference_extension(case):
    import stereoformer_b200.DCN as ours
    ref = _ref()
    B, Cin, Cout, H, W, G, dg, s, p, d = case
    torch.manual_seed(4)
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin // G, 3, 3, device="cuda") * 0.1
    b = torch.randn(Cout, device="cuda")
    Ho = (H + 2 * p - (d * 2 + 1)) // s + 1
    Wo = (W + 2 * p - (d * 2 + 1)) // s + 1
    off = torch.randn(B, dg * 18, Ho, Wo, device="cuda") * 2
    msk = torch.sigmoid(torch.randn(B, dg * 9, Ho, Wo, device="cuda"))
    go = torch.randn(B, Cout, Ho, Wo, device="cuda")
    args = (3, 3, s, s, p, p, d, d, G, dg, 64)
    want = ref.modulated_deform_conv_backward(x, w, b, off, msk, go, *args)
    got = ours.modulated_deform_conv_backward(x, w, b, off, msk, go, *args)
    for name, g, r in zip(("input", "offset", "mask", "weight", "bias"), got, want):
        scale = r.abs().max().clamp_min(1e-6)
        assert (g - r).abs().max() <= 2e-4 * scale, (name, float((g - r).abs().max() / scale))


def test_backward_v1_and_partial_needs_on_the_channels_last_kernels():
    """DCNv1 (no mask) backward against the reference extension at a shape the channels-last kernels take, and requests
    for a subset of the gradients must return exactly what the full call returns for them."""
    import stereoformer_b200.DCN as ours
    from stereoformer_b200 import ops
    ref = _ref()
    B, Cin, Cout, H, W = 1, 64, 64, 14, 24
    torch.manual_seed(6)
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.1
    b = torch.randn(Cout, device="cuda")
    off = torch.randn(B, 18, H, W, device="cuda") * 2
    msk = torch.sigmoid(torch.randn(B, 9, H, W, device="cuda"))
    go = torch.randn(B, Cout, H, W, device="cuda")
    args = (3, 3, 1, 1, 1, 1, 1, 1, 1, 1, 64)
    want = ref.deform_conv_backward(x, w, b, off, go, *args)
    got = ours.deform_conv_backward(x, w, b, off, go, *args)
    for name, g, r in zip(("input", "offset", "weight", "bias"), got, want):
        scale = r.abs().max().clamp_min(1e-6)
        assert (g - r).abs().max() <= 2e-4 * scale, (name, float((g - r).abs().max() / scale))
    full = ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), 1, 1)
    for needs in ((True, False, False, False, False), (False, True, True, False, False), (False, False, False, True, True)):
        part = ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), 1, 1, needs=needs)
        for want_it, p_, f_ in zip(needs, part, full):
            assert (p_ is not None) == want_it
            if want_it:                                   # reductions run in a different order from call to call
                assert (p_ - f_).abs().max() <= 1e-5 * f_.abs().max()
