"""GPU parity: warp, guidance, CorrBlock1D, DCN, epipolar attention, convex upsampling and the drop-in
modules -- CUDA kernels (through the C ABI) vs the oracle / golden vectors."""
import pytest
import torch
import torch.nn as nn

from conftest import load_golden
from oracle import dcn_c
from oracle import stereo_oracle as O

pytestmark = pytest.mark.gpu


def dev(t):
    return t.cuda()


def close_rel(a, b, rtol=1e-4):
    """|a-b| <= rtol * max(|b|, rms(b)): the 1e-4 relative bound of north_star taken against the volume's
    magnitude (entries of a correlation volume are sums with cancellation)."""
    a, b = a.float().cpu(), b.float().cpu()
    s = b.pow(2).mean().sqrt().item()
    return bool(((a - b).abs() <= rtol * torch.clamp(b.abs(), min=s)).all())


@pytest.fixture(scope="module")
def ops():
    import stereoformer_b200.ops as ops
    return ops


# ------------------------------------------------------------------------------------------ a7 warp
def test_disp_warp_golden(ops):
    g = load_golden("disp_warp")
    w, m, dbg = ops.disp_warp(dev(g["img"]), dev(g["disp"]), "border", div_mode=ops.SF_DIV_TRUE, debug=True)
    assert torch.equal(dbg["grid"].cpu(), g["grid"])                     # normalised sampling grid: bit-exact
    H, W = g["img"].shape[-2:]
    ix, iy = O.warp_source_coords(g["grid"], H, W, border=True)
    assert torch.equal(dbg["src"].cpu()[..., 0], ix) and torch.equal(dbg["src"].cpu()[..., 1], iy)
    assert torch.allclose(w.cpu(), g["warped_border"], rtol=1e-4, atol=1e-5)
    assert torch.equal(m.cpu(), g["mask_border"])
    wz, mz = ops.disp_warp(dev(g["img"]), dev(g["disp"]), "zeros", div_mode=ops.SF_DIV_TRUE)
    assert torch.allclose(wz.cpu(), g["warped_zeros"], rtol=1e-4, atol=1e-5)
    assert torch.equal(mz.cpu(), g["mask_zeros"])


@pytest.mark.parametrize("shape", [(8, 3, 68, 120), (1, 3, 248, 360), (2, 5, 7, 33)])
def test_disp_warp_oracle(ops, shape):
    torch.manual_seed(23)
    B, C, H, W = shape
    img, disp = torch.randn(B, C, H, W), torch.rand(B, 1, H, W) * W * 0.4
    rw, rm = O.disp_warp(img, disp)
    w, m = ops.disp_warp(dev(img), dev(disp), div_mode=ops.SF_DIV_TRUE)
    assert torch.allclose(w.cpu(), rw, rtol=1e-4, atol=1e-4)
    assert (m.cpu() != rm).float().mean() < 1e-4        # threshold 0.9999 may flip on a last-bit coordinate
    # zero disparity is NOT the identity (align_corners quirk, SURVEY 8c iv)
    w0, _ = ops.disp_warp(dev(img), torch.zeros(B, 1, H, W, device="cuda"), div_mode=ops.SF_DIV_TRUE)
    r0, _ = O.disp_warp(img, torch.zeros(B, 1, H, W))
    assert torch.allclose(w0.cpu(), r0, rtol=1e-4, atol=1e-4)
    assert not torch.allclose(w0.cpu(), img, atol=1e-3)


def test_disp_warp_matches_pytorch_cuda(ops):
    """Against PyTorch's own CUDA kernels (what the reference executes on a GPU), reciprocal division."""
    torch.manual_seed(29)
    img, disp = torch.randn(2, 3, 20, 120, device="cuda"), torch.rand(2, 1, 20, 120, device="cuda") * 30
    grid = O.warp_grid(disp, 20, 120)
    _, _, dbg = ops.disp_warp(img, disp, div_mode=ops.SF_DIV_RECIP, debug=True)
    assert torch.equal(dbg["grid"], grid)
    rw, rm = O.disp_warp(img, disp)
    w, m = ops.disp_warp(img, disp, div_mode=ops.SF_DIV_RECIP)
    assert torch.allclose(w, rw, rtol=1e-4, atol=1e-4)
    assert (m != rm).float().mean() < 1e-4


def test_disp_warp_negative_raises(ops):
    img, disp = torch.randn(1, 3, 4, 16, device="cuda"), torch.rand(1, 1, 4, 16, device="cuda")
    disp[0, 0, 1, 3] = -0.5
    with pytest.raises(AssertionError):
        ops.disp_warp(img, disp)
    ops.disp_warp(img, disp, check=False)      # sync-free path does not raise


def test_guidance_oracle(ops):
    g = load_golden("uncertainty")
    torch.manual_seed(31)
    B, D, H, W = g["cv"].shape
    left, right = torch.randn(B, 3, H, W), torch.randn(B, 3, H, W)
    e_ref, u_ref = O.guidance(g["cv"], g["cur"], left, right)
    err, unc = ops.guidance(dev(g["cv"]), dev(g["cur"]), dev(left), dev(right), div_mode=ops.SF_DIV_TRUE)
    assert torch.allclose(unc.cpu(), g["unc"], rtol=1e-4, atol=1e-5)
    assert torch.allclose(unc.cpu(), u_ref, rtol=1e-4, atol=1e-5)
    assert torch.allclose(err.cpu(), e_ref, rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------------------------------ a8 / a9 CorrBlock1D
def test_corr1d_golden(ops):
    g = load_golden("corr1d")
    pyr = ops.corr1d_pyramid(dev(g["f1"]), dev(g["f2"]), g["num_levels"])
    assert len(pyr) == g["num_levels"] + 1
    for i, p in enumerate(pyr):
        assert p.shape == g["pyr%d" % i].shape
        assert close_rel(p, g["pyr%d" % i])
    for i, p in enumerate(ops.corr1d_pyramid(dev(g["f1"]), dev(g["f2"]), g["num_levels"], algo=ops.SF_ALGO_GENERIC)):
        assert torch.allclose(p.cpu(), g["pyr%d" % i], rtol=1e-4, atol=1e-5)       # fp32 FFMA kernel
    # lookup on the golden pyramid isolates the lookup kernel
    gp = [dev(g["pyr%d" % i]) for i in range(g["num_levels"])]
    out, dbg = ops.corr1d_lookup(gp, dev(g["coords"]), g["radius"], g["num_levels"], div_mode=ops.SF_DIV_TRUE,
                                 debug=True)
    assert torch.allclose(out.cpu(), g["lookup"], rtol=1e-4, atol=1e-5)
    B, _, H, W = g["coords"].shape
    taps = 2 * g["radius"] + 1
    for l in range(g["num_levels"]):
        x0, _ = O.corr1d_lookup_coords(g["coords"][:, 0].reshape(-1), l, g["radius"], gp[l].shape[-1])
        got = dbg["x0"][:, l * taps:(l + 1) * taps].permute(0, 2, 3, 1).reshape(-1, taps)
        assert torch.equal(got.cpu(), x0)                               # lookup coordinates: bit-exact


@pytest.mark.parametrize("shape", [(1, 256, 8, 240), (2, 32, 3, 100), (1, 20, 2, 37)])
def test_corr1d_oracle(ops, shape):
    torch.manual_seed(37)
    B, C, H, W = shape
    f1, f2 = torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    ref = O.corr1d_pyramid(f1, f2, 4)
    pyr = ops.corr1d_pyramid(dev(f1), dev(f2), 4)
    for p, r in zip(pyr, ref):
        assert p.shape == r.shape
        assert close_rel(p, r)
    xs = torch.arange(W).float().view(1, 1, 1, W).repeat(B, 1, H, 1)
    coords = torch.cat((xs - torch.rand(B, 1, H, W) * W / 4, torch.zeros(B, 1, H, W)), 1)
    lref = O.corr1d_lookup(ref, coords, 4, 4)
    out = ops.corr1d_lookup([dev(r) for r in ref], dev(coords), 4, 4, div_mode=ops.SF_DIV_TRUE)
    assert torch.allclose(out.cpu(), lref, rtol=1e-4, atol=2e-5)
    # KAT (SURVEY 8c v): integer coordinates at level 0 return V[x1, x1 + dx]
    ci = torch.cat((xs, torch.zeros(B, 1, H, W)), 1)
    o0 = ops.corr1d_lookup([dev(r) for r in ref], dev(ci), 4, 4, div_mode=ops.SF_DIV_TRUE).cpu()
    v0 = ref[0].view(B, H, W, W)
    for t, dx in enumerate(range(-4, 5)):
        x2 = torch.arange(W) + dx
        ok = (x2 >= 0) & (x2 < W)
        want = torch.where(ok, v0[:, :, torch.arange(W), x2.clamp(0, W - 1)], torch.zeros(()))
        assert torch.allclose(o0[:, t], want, rtol=1e-4, atol=2e-5)


# ------------------------------------------------------------------------------------------ a10 DCN
def test_dcn_reference_analytic_checks():
    """deform/test.py:69-110, 142-181, 219-260 translated 1:1 to the drop-in DCN module."""
    from stereoformer_b200 import DCN
    from stereoformer_b200.dcn_modules import ModulatedDeformConv, DeformConv, ModulatedDeformConvPack
    torch.manual_seed(3)
    N, inC, inH, inW, outC = 2, 4, 4, 4, 4
    conv = nn.Conv2d(inC, outC, 3, stride=1, padding=1, groups=2, bias=True).cuda()
    md = ModulatedDeformConv(inC, outC, 3, 1, 1, 1, groups=2, deformable_groups=1, im2col_step=1).cuda()
    md.weight.data.copy_(conv.weight.data)
    md.bias.data.copy_(conv.bias.data)
    x = torch.randn(N, inC, inH, inW).cuda()
    off = torch.zeros(N, 18, inH, inW).cuda()
    msk = torch.ones(N, 9, inH, inW).cuda()
    with torch.no_grad():
        assert (md(x, off, msk) - conv(x)).abs().max() < 1e-5          # zero offset == conv2d
        d1 = DeformConv(inC, outC, 3, 1, 1, 1, groups=2, deformable_groups=1, im2col_step=1).cuda()
        d1.weight.data.copy_(conv.weight.data)
        d1.bias.data.copy_(conv.bias.data)
        assert (d1(x, off) - conv(x)).abs().max() < 1e-5
        # identity weights, mask 0.5 -> 2 * out == in
        mi = ModulatedDeformConv(inC, inC, 3, 1, 1, 1, groups=inC, deformable_groups=1, im2col_step=1, bias=False).cuda()
        mi.weight.data.zero_()
        mi.weight.data[:, 0, 1, 1] = 1.0
        mi.bias.data.zero_()
        assert (2 * mi(x, off, msk * 0.5) - x).abs().max() < 1e-6
        # im2col_step invariance + the divisibility error
        offr = torch.randn(N, 18, inH, inW).cuda() * 2
        mr = torch.sigmoid(torch.randn(N, 9, inH, inW).cuda())
        md2 = ModulatedDeformConv(inC, outC, 3, 1, 1, 1, groups=2, deformable_groups=1, im2col_step=2).cuda()
        md2.load_state_dict(md.state_dict())
        assert torch.equal(md(x, offr, mr), md2(x, offr, mr))
        with pytest.raises(RuntimeError, match="must divide im2col_step"):
            DCN.modulated_deform_conv_forward(torch.cat([x, x[:1]]), md.weight, md.bias, torch.cat([offr, offr[:1]]),
                                              torch.cat([mr, mr[:1]]), 3, 3, 1, 1, 1, 1, 1, 1, 2, 1, 2)
        # a fresh Pack is 0.5 * conv2d + bias (zero-initialised offsets, mask = sigmoid(0))
        pk = ModulatedDeformConvPack(inC, outC, 3, 1, 1).cuda()
        ref = 0.5 * nn.functional.conv2d(x, pk.weight, None, 1, 1) + pk.bias.view(1, -1, 1, 1)
        assert (pk(x) - ref).abs().max() < 1e-5
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        md.cpu()(x.cpu(), off.cpu(), msk.cpu())


@pytest.mark.parametrize("cfg", [dict(B=2, Cin=16, Cout=16, H=17, W=30, G=1, dg=1, s=1, p=1, d=1),     # shipped shape
                                 dict(B=2, Cin=64, Cout=128, H=12, W=20, G=1, dg=2, s=1, p=1, d=1),    # test.py:526
                                 dict(B=1, Cin=24, Cout=40, H=9, W=11, G=2, dg=3, s=2, p=2, d=2),
                                 dict(B=1, Cin=6, Cout=3, H=6, W=6, G=3, dg=3, s=1, p=0, d=1),
                                 dict(B=1, Cin=128, Cout=128, H=8, W=10, G=1, dg=1, s=1, p=1, d=1)])
def test_dcn_vs_c_oracle(ops, cfg):
    torch.manual_seed(41)
    B, Cin, Cout, H, W, G, dg, s, p, d = (cfg[k] for k in ("B", "Cin", "Cout", "H", "W", "G", "dg", "s", "p", "d"))
    x = torch.randn(B, Cin, H, W)
    w = torch.randn(Cout, Cin // G, 3, 3) / (3 * (Cin // G) ** 0.5)
    b = torch.randn(Cout)
    Ho = (H + 2 * p - (d * 2 + 1)) // s + 1
    Wo = (W + 2 * p - (d * 2 + 1)) // s + 1
    off = torch.randn(B, dg * 18, Ho, Wo) * 2
    msk = torch.sigmoid(torch.randn(B, dg * 9, Ho, Wo))
    ref = dcn_c.dcn_forward(x, w, b, off, msk, (s, s), (p, p), (d, d), G, dg)
    out = ops.dcn_forward(dev(x), dev(w), dev(b), dev(off), dev(msk), (s, s), (p, p), (d, d), G, dg)
    assert out.shape == ref.shape
    assert torch.allclose(out.cpu(), ref, rtol=1e-4, atol=1e-4)
    ref1 = dcn_c.dcn_forward(x, w, b, off, None, (s, s), (p, p), (d, d), G, dg)
    out1 = ops.dcn_forward(dev(x), dev(w), dev(b), dev(off), None, (s, s), (p, p), (d, d), G, dg)
    assert torch.allclose(out1.cpu(), ref1, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("cfg", [dict(B=2, Cin=64, Cout=128, H=12, W=20, G=1, dg=2, s=1, p=1, d=1),    # test.py:526
                                 dict(B=1, Cin=48, Cout=80, H=13, W=21, G=2, dg=3, s=2, p=2, d=2),     # K, N padding
                                 dict(B=1, Cin=128, Cout=128, H=19, W=23, G=1, dg=1, s=1, p=1, d=1),   # 437 pixels
                                 dict(B=1, Cin=32, Cout=256, H=9, W=15, G=1, dg=4, s=1, p=0, d=1),     # N = 256
                                 dict(B=3, Cin=8, Cout=8, H=6, W=40, G=1, dg=1, s=1, p=1, d=1)])       # tiny K (forced)
def test_dcn_tensor_core(ops, cfg):
    """tcgen05 implicit-GEMM path (bf16 hi/mid split, A operand in tensor memory) against the C oracle and the
    exact FFMA kernel; tolerance 1e-4 of the output magnitude."""
    torch.manual_seed(43)
    B, Cin, Cout, H, W, G, dg, s, p, d = (cfg[k] for k in ("B", "Cin", "Cout", "H", "W", "G", "dg", "s", "p", "d"))
    x = torch.randn(B, Cin, H, W)
    w = torch.randn(Cout, Cin // G, 3, 3) / (3 * (Cin // G) ** 0.5)
    b = torch.randn(Cout)
    Ho = (H + 2 * p - (d * 2 + 1)) // s + 1
    Wo = (W + 2 * p - (d * 2 + 1)) // s + 1
    off = torch.randn(B, dg * 18, Ho, Wo) * 2
    msk = torch.sigmoid(torch.randn(B, dg * 9, Ho, Wo))
    for m in (msk, None):
        ref = dcn_c.dcn_forward(x, w, b, off, m, (s, s), (p, p), (d, d), G, dg)
        args = (dev(x), dev(w), dev(b), dev(off), dev(m) if m is not None else None, (s, s), (p, p), (d, d), G, dg)
        tc = ops.dcn_forward(*args, algo=ops.SF_ALGO_TC)
        ex = ops.dcn_forward(*args, algo=ops.SF_ALGO_GENERIC)
        scale = ref.abs().max().item()
        assert (tc.cpu() - ref).abs().max().item() <= 1e-4 * scale
        assert (tc - ex).abs().max().item() <= 1e-4 * scale
    # zero offsets + unit mask == plain convolution (deform/test.py:69-110) on the tensor-core path too
    conv = torch.nn.functional.conv2d(x, w, b, stride=s, padding=p, dilation=d, groups=G)
    z = ops.dcn_forward(dev(x), dev(w), dev(b), dev(torch.zeros_like(off)), dev(torch.ones_like(msk)), (s, s), (p, p),
                        (d, d), G, dg, algo=ops.SF_ALGO_TC)
    assert (z.cpu() - conv).abs().max().item() <= 1e-4 * conv.abs().max().item()


def test_dcn_tensor_core_limits(ops):
    x = torch.randn(1, 8, 5, 5, device="cuda")
    w = torch.randn(4, 8, 3, 3, device="cuda")          # 4 output channels: no tensor-core shape
    off = torch.zeros(1, 18, 5, 5, device="cuda")
    with pytest.raises(ValueError):
        ops.dcn_forward(x, w, torch.zeros(4, device="cuda"), off, None, (1, 1), (1, 1), (1, 1), 1, 1, algo=ops.SF_ALGO_TC)


# ------------------------------------------------------------------------------------------ a11 attention (spec-derived)
@pytest.mark.parametrize("shape,band", [((1, 128, 6, 312), 48), ((2, 32, 3, 40), 8), ((1, 256, 2, 100), 64),
                                        ((1, 16, 2, 10), 24)])
def test_epipolar_attention_oracle(ops, shape, band):
    torch.manual_seed(43)
    q, k, v = (torch.randn(*shape).bfloat16() for _ in range(3))
    o_ref, s_ref = O.epipolar_attention(q, k, v, band)
    out, sc = ops.epipolar_attention(dev(q), dev(k), dev(v), band, return_scores=True)
    assert torch.allclose(out.float().cpu(), o_ref, rtol=2e-2, atol=2e-2)      # bf16 output
    fin = torch.isfinite(s_ref)
    assert torch.equal(torch.isfinite(sc.cpu()), fin)
    assert torch.allclose(sc.cpu()[fin], s_ref[fin], rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------------------------------ f1 upsampling
def test_upsample_convex_golden(ops):
    g = load_golden("upsample_convex8")
    out = ops.upsample_convex(dev(g["disp"]), dev(g["mask"]), 8)
    assert torch.allclose(out.cpu(), g["out"], rtol=1e-4, atol=1e-4)
    torch.manual_seed(47)
    d, m = torch.rand(2, 1, 9, 13) * 20, torch.randn(2, 9 * 16, 9, 13)
    assert torch.allclose(ops.upsample_convex(dev(d), dev(m), 4).cpu(), O.upsample_convex(d, m, 4), rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("shape", [(2, 1, 5, 9), (1, 1, 68, 120), (1, 1, 2, 3)])
def test_upsample_simple8(ops, shape):
    """upsample_simple8 (models/utils/upsample.py:40-43): golden vector of the reference, the oracle, and PyTorch's
    own CUDA kernel on the same device (1e-4 px; values reach 8*disp)."""
    import torch.nn.functional as F
    import stereoformer_b200 as sf
    g = load_golden("upsample_simple8")
    out = sf.upsample_simple8(dev(g["disp"].detach()))
    assert out.shape == g["out"].shape and (out.cpu() - g["out"]).abs().max() < 1e-4
    dg = dev(g["disp"].detach()).requires_grad_(True)
    gd = torch.autograd.grad(sf.upsample_simple8(dg), dg, dev(g["grad_out"]))[0]
    assert torch.allclose(gd.cpu(), g["grad_disp"], rtol=1e-5, atol=1e-4)
    torch.manual_seed(48)
    d = torch.rand(*shape) * 24
    got = ops.upsample_simple(dev(d), 8)
    assert (got.cpu() - O.upsample_simple(d, 8)).abs().max() < 1e-4
    Hh, Ww = shape[2] * 8, shape[3] * 8
    assert (got - 8 * F.interpolate(dev(d), size=(Hh, Ww), mode="bilinear", align_corners=True)).abs().max() < 1e-4
    go = torch.randn(shape[0], 1, Hh, Ww)
    d2 = d.clone().requires_grad_(True)
    ref_g = torch.autograd.grad(O.upsample_simple(d2, 8), d2, go)[0]
    assert torch.allclose(ops.upsample_simple_backward(dev(go), 8).cpu(), ref_g, rtol=1e-5, atol=1e-4)
    with pytest.raises(NotImplementedError):
        sf.upsample_simple8(dev(d), mode="nearest")


def test_corr1d_alt(ops):
    """PytorchAlternateCorrBlock1D: golden vector of the reference class, the oracle on a second shape (bit-exact
    sampling coordinates, values <= 1e-4 of the magnitude), and -- pooling and correlation commute -- agreement with the
    volume form (CorrBlock1D) at the config-3 row shape."""
    import stereoformer_b200 as sf
    g = load_golden("corr1d_alt")
    r, L = int(g["radius"]), int(g["num_levels"])
    out, dbg = ops.corr1d_alt(dev(g["f1"]), dev(g["f2"]), dev(g["coords"]), r, L, div_mode=ops.SF_DIV_TRUE, debug=True)
    assert out.shape == g["out"].shape and close_rel(out, g["out"])
    # un-normalised sampling coordinates against the oracle's round trip, bit for bit
    B, _, H, W = g["coords"].shape
    cx = g["coords"][:, :1].permute(0, 2, 3, 1).reshape(-1)
    for lvl in range(L):
        Wl = W >> lvl
        _, xg = O.corr1d_lookup_coords(cx, lvl, r, Wl)
        ix = ((xg + 1) * 0.5) * (Wl - 1)
        got = dbg["src"][:, lvl * (2 * r + 1):(lvl + 1) * (2 * r + 1), :, :, 0].permute(0, 2, 3, 1).reshape(-1, 2 * r + 1)
        assert torch.equal(got.cpu(), ix)
    blk = sf.PytorchAlternateCorrBlock1D(dev(g["f1"]), dev(g["f2"]), num_levels=L, radius=r)
    blk.div_mode = 0
    assert torch.equal(blk(dev(g["coords"])), out)
    torch.manual_seed(49)
    f1, f2 = torch.randn(1, 40, 6, 52), torch.randn(1, 40, 6, 52)
    coords = torch.cat((torch.arange(52.).view(1, 1, 1, 52).repeat(1, 1, 6, 1) - torch.rand(1, 1, 6, 52) * 20,
                        torch.arange(6.).view(1, 1, 6, 1).repeat(1, 1, 1, 52) + torch.rand(1, 1, 6, 52) * 0.5), 1)
    assert close_rel(ops.corr1d_alt(dev(f1), dev(f2), dev(coords), 4, 4, div_mode=ops.SF_DIV_TRUE),
                     O.corr1d_alt(f1, f2, coords, 4, 4))           # fractional y: two-row bilinear
    f1, f2 = torch.randn(2, 256, 8, 240, device="cuda"), torch.randn(2, 256, 8, 240, device="cuda")
    coords = torch.cat((torch.arange(240., device="cuda").view(1, 1, 1, 240).repeat(2, 1, 8, 1) -
                        torch.rand(2, 1, 8, 240, device="cuda") * 60,
                        torch.arange(8., device="cuda").view(1, 1, 8, 1).repeat(2, 1, 1, 240)), 1)
    reg = sf.CorrBlock1D(f1, f2, num_levels=4, radius=4)
    reg.div_mode = 0
    assert close_rel(ops.corr1d_alt(f1, f2, coords, 4, 4, div_mode=ops.SF_DIV_TRUE), reg(coords), rtol=2e-4)


# ------------------------------------------------------------------------------------------ drop-in modules
def test_dropin_modules_chain():
    """The LowCNN LC hot-path chain (baseline.py:114-127) through the drop-in nn.Modules."""
    import stereoformer_b200 as sf
    torch.manual_seed(53)
    fl, fr = torch.randn(1, 256, 32, 64), torch.randn(1, 256, 32, 64)      # BASELINE config 1
    cv_ref = O.corr_volume(fl, fr, 24)
    d_ref = O.soft_argmin(cv_ref)
    cv = sf.CostVolume(24, "correlation")(dev(fl), dev(fr))
    disp = sf.DisparityEstimation(24, True)(cv)
    assert disp.shape == d_ref.shape and (disp.cpu() - d_ref).abs().max() < 1e-3
    lc = sf.LocalCostVolume(radius=2, sample_points=20)
    lc.div_mode = 0
    out = lc(cv, disp.unsqueeze(1), True)
    ref = O.local_cost_volume(cv.cpu(), disp.unsqueeze(1).cpu(), 2, 20, True)
    assert out.shape == ref.shape and (out.cpu() - ref).abs().max() < 1e-3
    blk = sf.CorrBlock1D(dev(fl[:, :, :4]), dev(fr[:, :, :4]), num_levels=4, radius=4)
    assert len(blk.corr_pyramid) == 5 and blk.corr_pyramid[0].shape == (4 * 64, 1, 1, 64)
    coords = torch.rand(1, 2, 4, 64, device="cuda") * 60
    assert blk(coords).shape == (1, 36, 4, 64)
    x = dev(fl).requires_grad_(True)
    y = sf.CostVolume(24)(x, dev(fr))
    y.sum().backward()                                  # backward kernels: tests/test_gpu_autograd.py
    assert x.grad is not None and torch.isfinite(x.grad).all()
    blk2 = sf.CorrBlock1D(dev(fl[:, :, :4]).requires_grad_(True), dev(fr[:, :, :4]), num_levels=4, radius=4)
    with pytest.raises(NotImplementedError):            # CorrBlock1D stays forward-only
        blk2(coords).sum().backward()


@pytest.mark.parametrize("shape,band", [((1, 128, 6, 312), 48), ((2, 64, 3, 40), 8), ((1, 256, 2, 104), 64),
                                        ((1, 192, 3, 264), 17), ((3, 128, 5, 128), 1), ((2, 128, 96, 312), 48)])
def test_epipolar_attention_tcgen05(ops, shape, band):
    """The tcgen05/TMEM kernel (forced with SF_ALGO_TMA) against the oracle and against the FFMA kernel."""
    torch.manual_seed(59)
    q, k, v = (torch.randn(*shape).bfloat16() for _ in range(3))
    out = ops.epipolar_attention(dev(q), dev(k), dev(v), band, algo=ops.SF_ALGO_TMA)
    gen = ops.epipolar_attention(dev(q), dev(k), dev(v), band, algo=ops.SF_ALGO_GENERIC)
    assert torch.allclose(out.float(), gen.float(), rtol=2e-2, atol=2e-2)
    # against the (spec-derived) oracle at EVERY size, the C4-sized case included: the oracle's tensor program runs on
    # the device (same fp32 arithmetic as on the CPU up to summation order, far below the bf16 output rounding)
    o_ref, _ = O.epipolar_attention(dev(q), dev(k), dev(v), band)
    assert torch.allclose(out.float(), o_ref, rtol=2e-2, atol=2e-2)
    if shape[2] <= 8:
        o_cpu, _ = O.epipolar_attention(q, k, v, band)
        assert torch.allclose(out.float().cpu(), o_cpu, rtol=2e-2, atol=2e-2)
    with pytest.raises(ValueError):
        ops.epipolar_attention(dev(q[:, :24]), dev(k[:, :24]), dev(v[:, :24]), band, algo=ops.SF_ALGO_TMA)


@pytest.mark.parametrize("shape", [(1, 256, 8, 240), (2, 32, 3, 100), (1, 40, 2, 36), (1, 64, 2, 520), (1, 256, 136, 240)])
def test_corr1d_tcgen05(ops, shape):
    """bf16 hi/mid split on tcgen05 (forced with SF_ALGO_TMA) vs the fp32 FFMA kernel and the oracle:
    <= 1e-4 relative to the volume's magnitude; pooled levels are exact pairwise means of the level below."""
    torch.manual_seed(61)
    B, C, H, W = shape
    f1, f2 = torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    tc = ops.corr1d_pyramid(dev(f1), dev(f2), 4, algo=ops.SF_ALGO_TMA)
    ff = ops.corr1d_pyramid(dev(f1), dev(f2), 4, algo=ops.SF_ALGO_GENERIC)
    assert len(tc) == len(ff) == 5
    for a, b in zip(tc, ff):
        assert a.shape == b.shape
        scale = b.pow(2).mean().sqrt().item()
        assert ((a - b).abs() <= 1e-4 * torch.clamp(b.abs(), min=scale)).all()
    for l in range(1, 5):
        src = tc[l - 1]
        n = src.shape[-1] // 2
        assert torch.equal(tc[l], (src[..., 0:2 * n:2] + src[..., 1:2 * n:2]) * 0.5)
    if H <= 8:
        ref = O.corr1d_pyramid(f1, f2, 4)
        for a, r in zip(tc, ref):
            scale = r.pow(2).mean().sqrt().item()
            assert ((a.cpu() - r).abs() <= 1e-4 * torch.clamp(r.abs(), min=scale)).all()


def test_guidance_negative_disparity_check_is_opt_in():
    """the DLC path dropped the reference's `assert disp.min() >= 0`; it is available as an opt-in."""
    import stereoformer_b200 as sf
    cv = torch.randn(1, 8, 6, 16, device="cuda")
    disp = torch.rand(1, 1, 6, 16, device="cuda") * 7
    left, right = torch.randn(1, 3, 6, 16, device="cuda"), torch.randn(1, 3, 6, 16, device="cuda")
    a = sf.guidance_inputs(cv, disp, left, right)
    b = sf.guidance_inputs(cv, disp, left, right, check=True)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    disp[0, 0, 2, 3] = -0.5
    sf.guidance_inputs(cv, disp, left, right)                      # default: sync-free, no assertion
    with pytest.raises(AssertionError, match="disp.min"):
        sf.guidance_inputs(cv, disp, left, right, check=True)
