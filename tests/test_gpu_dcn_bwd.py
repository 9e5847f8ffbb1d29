"""GPU parity for the DCN backward: sf_dcn_backward_f32 through the pybind-shaped DCN module and
the autograd Functions, against the C oracle (the reference's col2im / col2im_coord restated and pinned on autograd
by tests/test_oracle_dcn.py) and with the reference's own gradient tests (deform/test.py:304-349, 405-434)."""
import pytest
import torch

from oracle import dcn_c

pytestmark = pytest.mark.gpu

CFGS = [dict(Cin=8, Cout=6, G=2, dg=2, s=1, p=1, d=1, H=7, W=9, k=3, B=2),
        dict(Cin=4, Cout=4, G=1, dg=1, s=2, p=2, d=2, H=9, W=8, k=3, B=1),
        dict(Cin=6, Cout=3, G=3, dg=1, s=1, p=0, d=1, H=6, W=6, k=3, B=2),
        dict(Cin=4, Cout=8, G=1, dg=2, s=1, p=0, d=1, H=5, W=7, k=1, B=2),
        dict(Cin=16, Cout=16, G=1, dg=1, s=1, p=1, d=1, H=20, W=37, k=3, B=3),      # shipped-model channel count
        dict(Cin=64, Cout=128, G=1, dg=2, s=1, p=1, d=1, H=16, W=24, k=3, B=2),     # deform/test.py:526 shape class
        dict(Cin=40, Cout=300, G=1, dg=1, s=1, p=1, d=1, H=9, W=11, k=3, B=1),      # > 256 outputs: two o-blocks
        dict(Cin=6, Cout=4, G=2, dg=3, s=1, p=2, d=1, H=8, W=8, k=5, B=1)]          # 25 taps


def _inputs(c, v2, seed=11):
    torch.manual_seed(seed)
    k = c["k"]
    Ho = (c["H"] + 2 * c["p"] - (c["d"] * (k - 1) + 1)) // c["s"] + 1
    Wo = (c["W"] + 2 * c["p"] - (c["d"] * (k - 1) + 1)) // c["s"] + 1
    x = torch.randn(c["B"], c["Cin"], c["H"], c["W"])
    w = torch.randn(c["Cout"], c["Cin"] // c["G"], k, k) * 0.3
    b = torch.randn(c["Cout"])
    off = torch.randn(c["B"], c["dg"] * 2 * k * k, Ho, Wo) * 2
    msk = torch.sigmoid(torch.randn(c["B"], c["dg"] * k * k, Ho, Wo)) if v2 else None
    go = torch.randn(c["B"], c["Cout"], Ho, Wo)
    return x, w, b, off, msk, go


def _close(a, r, name, tol=1e-4):
    a, r = a.cpu().double(), r.double()
    scale = max(1.0, r.abs().max().item())
    assert (a - r).abs().max().item() < tol * scale, (name, (a - r).abs().max().item(), scale)


@pytest.mark.parametrize("cfg", CFGS)
@pytest.mark.parametrize("v2", [True, False])
def test_dcn_backward_vs_oracle(cfg, v2):
    from stereoformer_b200 import DCN
    x, w, b, off, msk, go = _inputs(cfg, v2)
    sp = ((cfg["s"],) * 2, (cfg["p"],) * 2, (cfg["d"],) * 2, cfg["G"], cfg["dg"])
    ref = dcn_c.dcn_backward(x, w, off, msk, go, *sp)
    k, s, p, d = cfg["k"], cfg["s"], cfg["p"], cfg["d"]
    if v2:
        got = DCN.modulated_deform_conv_backward(x.cuda(), w.cuda(), b.cuda(), off.cuda(), msk.cuda(), go.cuda(), k, k,
                                                 s, s, p, p, d, d, cfg["G"], cfg["dg"], 64)
        names = ("input", "offset", "mask", "weight", "bias")
        refs = ref
    else:
        got = DCN.deform_conv_backward(x.cuda(), w.cuda(), b.cuda(), off.cuda(), go.cuda(), k, k, s, s, p, p, d, d,
                                       cfg["G"], cfg["dg"], 64)
        names = ("input", "offset", "weight", "bias")
        refs = (ref[0], ref[1], ref[3], ref[4])
    assert len(got) == len(names)
    for n, a, r in zip(names, got, refs):
        assert a.shape == r.shape, n
        _close(a, r, n)


def test_dcn_backward_partial_and_errors():
    import stereoformer_b200.ops as ops
    from stereoformer_b200 import DCN
    c = CFGS[0]
    x, w, b, off, msk, go = (t.cuda() for t in _inputs(c, True))
    full = ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), 2, 2)
    part = ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), 2, 2, needs=(False, True, False, False, True))
    assert part[0] is None and part[2] is None and part[3] is None
    assert torch.equal(part[1], full[1]) and torch.equal(part[4], full[4])
    with pytest.raises(RuntimeError, match="grad_out shape"):
        DCN.modulated_deform_conv_backward(x, w, b, off, msk, go[:, :, :-1], 3, 3, 1, 1, 1, 1, 1, 1, 2, 2, 64)
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        DCN.modulated_deform_conv_backward(x.cpu(), w, b, off, msk, go, 3, 3, 1, 1, 1, 1, 1, 1, 2, 2, 64)
    with pytest.raises(ValueError, match="36 taps"):
        k7 = torch.randn(4, 4, 7, 7, device="cuda")
        ops.dcn_backward(torch.randn(1, 4, 9, 9, device="cuda"), k7, torch.zeros(1, 98, 3, 3, device="cuda"), None,
                         torch.zeros(1, 4, 3, 3, device="cuda"), (1, 1), (0, 0), (1, 1), 1, 1)


def test_autograd_modules_match_dense_autograd():
    """ModulatedDeformConvPack / DeformConv under autograd: every parameter and input gradient against autograd
    through the float64 dense restatement of the forward (oracle)."""
    import stereoformer_b200 as sf
    from oracle import stereo_oracle as O
    torch.manual_seed(3)
    m = sf.ModulatedDeformConv(8, 6, 3, stride=1, padding=1, dilation=1, groups=2, deformable_groups=2).cuda()
    x = torch.randn(2, 8, 7, 9, device="cuda", requires_grad=True)
    off = (torch.randn(2, 36, 7, 9, device="cuda") * 2).requires_grad_(True)
    msk = torch.sigmoid(torch.randn(2, 18, 7, 9, device="cuda")).requires_grad_(True)
    out = m(x, off, msk)
    go = torch.randn_like(out)
    got = torch.autograd.grad(out, [x, off, msk, m.weight, m.bias], go)
    leaves = [t.detach().cpu().double().requires_grad_(True) for t in (x, off, msk, m.weight, m.bias)]
    ref_out = O.dcn_forward(leaves[0], leaves[3], leaves[4], leaves[1], leaves[2], (1, 1), (1, 1), (1, 1), 2, 2)
    ref = torch.autograd.grad(ref_out, leaves, go.cpu().double())
    for n, a, r in zip(("input", "offset", "mask", "weight", "bias"), got, ref):
        _close(a, r, n)
    # DCNv1 module
    m1 = sf.DeformConv(8, 6, 3, stride=1, padding=1, dilation=1, groups=2, deformable_groups=2).cuda()
    out1 = m1(x, off)
    got1 = torch.autograd.grad(out1, [x, off, m1.weight, m1.bias], go)
    l1 = [t.detach().cpu().double().requires_grad_(True) for t in (x, off, m1.weight, m1.bias)]
    ref1 = torch.autograd.grad(O.dcn_forward(l1[0], l1[2], l1[3], l1[1], None, (1, 1), (1, 1), (1, 1), 2, 2), l1,
                               go.cpu().double())
    for n, a, r in zip(("input", "offset", "weight", "bias"), got1, ref1):
        _close(a, r, n)


def test_pack_module_trains_one_step():
    """deform/test.py:405-434 style: the Pack module (offsets and mask predicted by its own conv) runs
    forward + backward and every parameter receives a finite gradient."""
    import stereoformer_b200 as sf
    torch.manual_seed(4)
    m = sf.ModulatedDeformConvPack(16, 16, 3, stride=1, padding=1, deformable_groups=2).cuda()
    torch.nn.init.normal_(m.conv_offset_mask.weight, std=0.05)          # the reference zero-initialises it
    x = torch.randn(2, 16, 12, 20, device="cuda", requires_grad=True)
    loss = (m(x) ** 2).mean()
    loss.backward()
    for n, p_ in m.named_parameters():
        assert p_.grad is not None and torch.isfinite(p_.grad).all() and p_.grad.abs().sum() > 0, n
    assert x.grad is not None and torch.isfinite(x.grad).all()


def test_gradcheck_like_reference():
    """The float32 kernels under the same kind of check as deform/test.py:304-349: a central-difference probe of a few
    offset / mask / input entries with float32 tolerances (the double-precision gradcheck itself is
    test_reference_gradcheck_double below)."""
    import stereoformer_b200.ops as ops
    c = CFGS[0]
    x, w, b, off, msk, go = (t.cuda() for t in _inputs(c, True, seed=5))
    sp = ((1, 1), (1, 1), (1, 1), 2, 2)
    g_in, g_off, g_m, g_w, g_b = ops.dcn_backward(x, w, off, msk, go, *sp)

    def loss(x_, off_, msk_, w_):
        return (ops.dcn_forward(x_, w_, b, off_, msk_, *sp, algo=ops.SF_ALGO_GENERIC).double() * go.double()).sum().item()

    gen = torch.Generator().manual_seed(0)
    eps = 1e-2
    for which, t, g in (("x", x, g_in), ("mask", msk, g_m), ("w", w, g_w), ("off", off, g_off)):
        for _ in range(5):
            idx = tuple(int(torch.randint(0, n, (1,), generator=gen)) for n in t.shape)
            tp, tm = t.clone(), t.clone()
            tp[idx] += eps
            tm[idx] -= eps
            args_p = [tp if which == n else v for n, v in (("x", x), ("off", off), ("mask", msk), ("w", w))]
            args_m = [tm if which == n else v for n, v in (("x", x), ("off", off), ("mask", msk), ("w", w))]
            num = (loss(*args_p) - loss(*args_m)) / (2 * eps)
            # offsets: the bilinear surface is piecewise linear, a kink inside +-eps breaks the difference quotient
            tol = 0.15 if which == "off" else 2e-2
            assert abs(num - g[idx].item()) < tol * max(1.0, abs(num)), (which, idx, num, g[idx].item())


# ------------------------------------------------------------------------------------------ double precision
@pytest.mark.parametrize("cfg", CFGS[:4] + CFGS[7:])
@pytest.mark.parametrize("v2", [True, False])
def test_dcn_f64_forward_backward_vs_oracle(cfg, v2):
    """The reference dispatches float and double (modulated_deform_conv_cuda.cu:93): sf_dcn_forward_f64 /
    sf_dcn_backward_f64 against autograd through the oracle's float64 dense restatement (1e-10)."""
    from oracle import stereo_oracle as O
    from stereoformer_b200 import DCN
    x, w, b, off, msk, go = (t.double() if t is not None else None for t in _inputs(cfg, v2))
    k, s, p, d = cfg["k"], cfg["s"], cfg["p"], cfg["d"]
    leaves = [t.clone().requires_grad_(True) for t in (x, off, w, b)] + ([msk.clone().requires_grad_(True)] if v2 else [])
    ref_out = O.dcn_forward(leaves[0], leaves[2], leaves[3], leaves[1], leaves[4] if v2 else None, (s, s), (p, p), (d, d),
                            cfg["G"], cfg["dg"])
    ref = torch.autograd.grad(ref_out, leaves, go)
    cu = lambda t: t.cuda()
    if v2:
        out = DCN.modulated_deform_conv_forward(cu(x), cu(w), cu(b), cu(off), cu(msk), k, k, s, s, p, p, d, d, cfg["G"],
                                                cfg["dg"], 64)
        g_in, g_off, g_m, g_w, g_b = DCN.modulated_deform_conv_backward(cu(x), cu(w), cu(b), cu(off), cu(msk), cu(go), k,
                                                                       k, s, s, p, p, d, d, cfg["G"], cfg["dg"], 64)
        assert (g_m.cpu() - ref[4]).abs().max() < 1e-10
    else:
        out = DCN.deform_conv_forward(cu(x), cu(w), cu(b), cu(off), k, k, s, s, p, p, d, d, cfg["G"], cfg["dg"], 64)
        g_in, g_off, g_w, g_b = DCN.deform_conv_backward(cu(x), cu(w), cu(b), cu(off), cu(go), k, k, s, s, p, p, d, d,
                                                         cfg["G"], cfg["dg"], 64)
    assert out.dtype == torch.float64 and (out.cpu() - ref_out.detach()).abs().max() < 1e-10
    for n, a, r in zip(("input", "offset", "weight", "bias"), (g_in, g_off, g_w, g_b), ref):
        assert a.dtype == torch.float64 and (a.cpu() - r).abs().max() < 1e-9, n


def test_reference_gradcheck_double():
    """deform/test.py:376-403 (check_gradient_dconv) and :405-434 (check_gradient_mdconv) as the reference runs them:
    torch.autograd.gradcheck on the autograd Functions with the reference's shapes (N,inC,inH,inW = 2,4,4,4; outC 4;
    3x3; groups 2; seed 3; deform/test.py:15-21) and tolerances (eps 1e-3, atol 1e-3, rtol 1e-2), in double."""
    from torch.autograd import gradcheck
    from stereoformer_b200.dcn_modules import DeformConvFunction, ModulatedDeformConvFunction
    torch.manual_seed(3)
    N, inC, inH, inW, outC, kH, kW, groups, dgs = 2, 4, 4, 4, 4, 3, 3, 2, 1
    inp = (torch.rand(N, inC, inH, inW).double().cuda() * 0.01).requires_grad_(True)
    offset = (torch.randn(N, dgs * 2 * kW * kH, inH, inW).double().cuda() * 2).requires_grad_(True)
    weight = torch.randn(outC, inC // groups, kH, kW).double().cuda().requires_grad_(True)
    bias = torch.rand(outC).double().cuda().requires_grad_(True)
    assert gradcheck(DeformConvFunction.apply, (inp, offset, weight, bias, 1, 1, 1, groups, dgs, 1),
                     eps=1e-3, atol=1e-3, rtol=1e-2, raise_exception=True, nondet_tol=1e-12)
    mask = torch.sigmoid(torch.rand(N, dgs * kW * kH, inH, inW).double().cuda()).requires_grad_(True)
    assert gradcheck(ModulatedDeformConvFunction.apply, (inp, offset, mask, weight, bias, 1, 1, 1, groups, dgs, 1),
                     eps=1e-3, atol=1e-3, rtol=1e-2, raise_exception=True, nondet_tol=1e-12)
