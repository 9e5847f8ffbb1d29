"""The precision switch: `fast` = tcgen05 kernels with a bf16 hi + mid operand split where the channel
count allows (the default), `exact` = fp32 FFMA kernels everywhere.  The reference's own analytic DCN checks
(deform/test.py:69-110: zero offsets + unit mask == nn.Conv2d within 1e-5) are run on BOTH at channel counts that the
default dispatch sends to the tensor cores, and the tolerance the fast path really meets is written down here."""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu

FAST_DCN_TOL = 3e-5      # relative to max |out|: bf16 hi+mid operands carry 16 mantissa bits (2^-17 per product, fp32 sums)
FAST_CV_TOL = 2e-5       # relative to max |volume| (north_star allows 1e-4)


@pytest.mark.parametrize("inC,outC,groups", [(64, 64, 1), (128, 128, 1), (64, 128, 2)])
def test_reference_zero_offset_check_on_both_precisions(inC, outC, groups):
    import stereoformer_b200 as sf
    from stereoformer_b200.dcn_modules import ModulatedDeformConv
    torch.manual_seed(3)                                    # deform/test.py:21
    N, H, W = 2, 20, 24
    conv = nn.Conv2d(inC, outC, 3, stride=1, padding=1, groups=groups, bias=True).cuda()
    md = ModulatedDeformConv(inC, outC, 3, 1, 1, 1, groups=groups, deformable_groups=1, im2col_step=1).cuda()
    md.weight.data.copy_(conv.weight.data)
    md.bias.data.copy_(conv.bias.data)
    x = torch.randn(N, inC, H, W).cuda()
    off = torch.zeros(N, 18, H, W).cuda()
    msk = torch.ones(N, 9, H, W).cuda()
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            want = conv(x)
            with sf.precision("exact"):
                err_exact = (md(x, off, msk) - want).abs().max().item()
            assert sf.get_precision() == "fast"
            err_fast = (md(x, off, msk) - want).abs().max().item()
    finally:
        torch.backends.cudnn.allow_tf32 = old
    scale = want.abs().max().item()
    print("zero-offset DCN %d->%d: exact %.2e, fast %.2e (max |out| %.2f)" % (inC, outC, err_exact, err_fast, scale))
    assert err_exact <= 1e-5 * max(1.0, scale)               # the reference's own bound (its test has |out| ~ 1, K = 36)
    assert err_fast <= FAST_DCN_TOL * scale


def test_cost_volume_precisions():
    import stereoformer_b200 as sf
    from oracle import stereo_oracle as O
    torch.manual_seed(5)
    fl, fr = torch.randn(2, 256, 17, 120), torch.randn(2, 256, 17, 120)
    want = O.corr_volume(fl, fr, 24)
    cvm = sf.CostVolume(24)
    with sf.precision("exact"):
        e_exact = (cvm(fl.cuda(), fr.cuda()).cpu() - want).abs().max().item()
    e_fast = (cvm(fl.cuda(), fr.cuda()).cpu() - want).abs().max().item()
    scale = want.abs().max().item()
    print("cost volume C=256: exact %.2e, fast %.2e of max |cv|" % (e_exact / scale, e_fast / scale))
    assert e_exact <= 2e-6 * scale                           # fp32 sums in a different order
    assert e_fast <= FAST_CV_TOL * scale


def test_explicit_algo_is_not_overridden():
    import stereoformer_b200 as sf
    from stereoformer_b200 import ops
    fl, fr = torch.randn(1, 64, 4, 64, device="cuda"), torch.randn(1, 64, 4, 64, device="cuda")
    with sf.precision("exact"):
        a = ops.corr_volume(fl, fr, 8, algo=ops.SF_ALGO_TC)
        b = ops.corr_volume(fl, fr, 8)
    c = ops.corr_volume(fl, fr, 8)
    d = ops.corr_volume(fl, fr, 8, algo=ops.SF_ALGO_EXACT)
    assert torch.equal(a, c) and torch.equal(b, d)
    with pytest.raises(ValueError):
        sf.set_precision("fp8")
