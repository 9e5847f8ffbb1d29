"""GPU parity for the backward kernels of the pure-PyTorch operators: gradients through the drop-in
modules against PyTorch autograd through the oracle's restatement of the reference modules (same op graph as the
reference: straight-through ceil/floor ending in .long(), torch.clamp, comparison masks as constants)."""
import pytest
import torch

from conftest import load_golden
from oracle import stereo_oracle as O

pytestmark = pytest.mark.gpu


def _close(a, r, name, tol=1e-4):
    a, r = a.detach().cpu().double(), r.detach().double()
    scale = max(1e-3, r.abs().max().item())
    err = (a - r).abs().max().item()
    assert err < tol * max(scale, 1.0) or err < tol * scale * 10, (name, err, scale)


def _leaf(t):
    return t.clone().requires_grad_(True)


@pytest.mark.parametrize("shape,D,groups", [((2, 16, 5, 40), 8, 1), ((1, 12, 3, 37), 6, 1), ((2, 32, 4, 24), 12, 4),
                                            ((1, 256, 6, 120), 24, 1), ((1, 8, 2, 10), 16, 2)])
def test_corr_volume_backward(shape, D, groups):
    import stereoformer_b200 as sf
    torch.manual_seed(21)
    L, R = torch.randn(*shape), torch.randn(*shape)
    lc, rc = _leaf(L), _leaf(R)
    ref = O.corr_volume(lc, rc, D) if groups == 1 else O.gwc_volume(lc, rc, D, groups)
    go = torch.randn_like(ref)
    rl, rr = torch.autograd.grad(ref, [lc, rc], go)
    lg, rg = _leaf(L.cuda()), _leaf(R.cuda())
    out = sf.CostVolume(D)(lg, rg) if groups == 1 else sf.build_gwc_volume(lg, rg, D, groups)
    gl, gr = torch.autograd.grad(out, [lg, rg], go.cuda())
    _close(gl, rl, "grad_left")
    _close(gr, rr, "grad_right")
    # only one side requires grad
    out2 = sf.CostVolume(D)(lg, R.cuda()) if groups == 1 else sf.build_gwc_volume(lg, R.cuda(), D, groups)
    (gl2,) = torch.autograd.grad(out2, [lg], go.cuda())
    assert torch.equal(gl2, gl)


@pytest.mark.parametrize("shape,D,groups", [((2, 256, 5, 120), 24, 1),      # the bench grid's row width: one ragged segment
                                            ((1, 128, 3, 360), 48, 1),      # C5 row width: 3 segments, 2 K blocks per band
                                            ((1, 64, 2, 132), 8, 1),        # 2 segments, the second almost empty
                                            ((1, 96, 2, 64), 65, 1),        # widest halo, ragged channel block (3 quarters)
                                            ((1, 320, 2, 200), 33, 1),      # 3 channel blocks, the last with 2 quarters
                                            ((2, 256, 3, 48), 12, 4)])      # group-wise: 64 channels per group
def test_corr_volume_backward_tensor_core(shape, D, groups):
    """The tcgen05 band-product backward (csrc/corr_volume_bwd_tc.cu) against the fp32 FFMA kernel and against autograd
    through the oracle; one-sided calls must give the same bits as the two-sided one."""
    from stereoformer_b200 import ops
    torch.manual_seed(22)
    L, R = torch.randn(*shape), torch.randn(*shape)
    lc, rc = _leaf(L), _leaf(R)
    ref = O.corr_volume(lc, rc, D) if groups == 1 else O.gwc_volume(lc, rc, D, groups)
    go = torch.randn_like(ref)
    rl, rr = torch.autograd.grad(ref, [lc, rc], go)
    Lc, Rc, G = L.cuda(), R.cuda(), go.cuda().contiguous()
    gl, gr = ops.corr_volume_backward(Lc, Rc, G, D, groups, algo=ops.SF_ALGO_TC)
    el, er = ops.corr_volume_backward(Lc, Rc, G, D, groups, algo=ops.SF_ALGO_GENERIC)
    for got, exact, want, name in ((gl, el, rl, "grad_left"), (gr, er, rr, "grad_right")):
        scale = want.abs().max().item()
        e_tc = (got.cpu() - want).abs().max().item()
        e_ex = (exact.cpu() - want).abs().max().item()
        print("%s %s D=%d: tcgen05 %.2e, FFMA %.2e of max |grad| %.3f" % (name, shape, D, e_tc / scale, e_ex / scale, scale))
        assert e_ex <= 1e-5 * scale
        assert e_tc <= 1e-4 * scale
    gl1, none = ops.corr_volume_backward(Lc, Rc, G, D, groups, needs=(True, False), algo=ops.SF_ALGO_TC)
    none2, gr1 = ops.corr_volume_backward(Lc, Rc, G, D, groups, needs=(False, True), algo=ops.SF_ALGO_TC)
    assert none is None and none2 is None
    assert torch.equal(gl1, gl) and torch.equal(gr1, gr)


def test_corr_volume_backward_precision_switch_and_limits():
    import stereoformer_b200 as sf
    from stereoformer_b200 import ops
    torch.manual_seed(23)
    L, R, G = torch.randn(1, 128, 2, 64).cuda(), torch.randn(1, 128, 2, 64).cuda(), torch.randn(1, 16, 2, 64).cuda()
    tc = ops.corr_volume_backward(L, R, G, 16, algo=ops.SF_ALGO_TC)
    fast = ops.corr_volume_backward(L, R, G, 16)
    with sf.precision("exact"):
        exact = ops.corr_volume_backward(L, R, G, 16)
    ffma = ops.corr_volume_backward(L, R, G, 16, algo=ops.SF_ALGO_GENERIC)
    assert torch.equal(fast[0], tc[0]) and torch.equal(fast[1], tc[1])
    assert torch.equal(exact[0], ffma[0]) and torch.equal(exact[1], ffma[1])
    with pytest.raises(ValueError, match="does not fit"):                     # 24 channels per group
        ops.corr_volume_backward(L[:, :24].contiguous(), R[:, :24].contiguous(), G, 16, algo=ops.SF_ALGO_TC)


@pytest.mark.parametrize("sim", [True, False])
def test_soft_argmin_backward(sim):
    import stereoformer_b200 as sf
    g = load_golden("soft_argmin")
    cv = g["cv"]
    c = _leaf(cv)
    ref = O.soft_argmin(c, match_similarity=sim)
    go = torch.randn_like(ref)
    (rg,) = torch.autograd.grad(ref, [c], go)
    cg = _leaf(cv.cuda())
    out = sf.DisparityEstimation(cv.shape[1], sim)(cg)
    (gg,) = torch.autograd.grad(out, [cg], go.cuda())
    _close(gg, rg, "grad_cv")


@pytest.mark.parametrize("valid", [0, 1])
def test_fixed_window_backward(valid):
    import stereoformer_b200 as sf
    g = load_golden("local_window_valid%d" % valid)
    cv, cur = g["cv"], g["cur"]
    c, d = _leaf(cv), _leaf(cur)
    ref = O.local_cost_volume(c, d, g["radius"], g["n_points"], bool(valid))
    go = torch.randn_like(ref)
    rc, rd = torch.autograd.grad(ref, [c, d], go)
    m = sf.LocalCostVolume(g["radius"], g["n_points"])
    m.div_mode = 0
    cg, dg = _leaf(cv.cuda()), _leaf(cur.cuda())
    out = m(cg, dg, bool(valid))
    gc, gd = torch.autograd.grad(out, [cg, dg], go.cuda())
    _close(gc, rc, "grad_cv")
    _close(gd, rd, "grad_disp")


@pytest.mark.parametrize("valid", [0, 1])
def test_variance_window_backward(valid):
    """a5: Variance_Based_Local_Cost_Volume under autograd against the reference module's own gradients (golden) and
    against autograd through the oracle on a second, larger input."""
    import stereoformer_b200 as sf
    g = load_golden("variance_window_valid%d" % valid)
    gg = load_golden("gradients_variance")
    m = sf.Variance_Based_Local_Cost_Volume(float(g["gamma"]), int(g["n_points"]))
    m.div_mode = 0
    cg, dg = _leaf(g["cv"].cuda()), _leaf(g["cur"].cuda())
    gc, gd = torch.autograd.grad(m(cg, dg, bool(valid)), [cg, dg], gg["var_valid%d_0" % valid].cuda())
    _close(gc, gg["var_valid%d_1" % valid], "grad_cv (reference)")
    _close(gd, gg["var_valid%d_2" % valid], "grad_disp (reference)")
    torch.manual_seed(33)
    cv = (torch.randn(2, 24, 17, 40) * 3 - 2).abs()
    cur = torch.rand(2, 1, 17, 40) * 23
    c, d = _leaf(cv), _leaf(cur)
    ref = O.variance_local_cost_volume(c, d, 0.7, 10, bool(valid))
    go = torch.randn_like(ref)
    rc, rd = torch.autograd.grad(ref, [c, d], go)
    m2 = sf.Variance_Based_Local_Cost_Volume(0.7, 10)
    m2.div_mode = 0
    cg, dg = _leaf(cv.cuda()), _leaf(cur.cuda())
    gc, gd = torch.autograd.grad(m2(cg, dg, bool(valid)), [cg, dg], go.cuda())
    _close(gc, rc, "grad_cv")
    _close(gd, rd, "grad_disp")


@pytest.mark.parametrize("valid", [0, 1])
@pytest.mark.parametrize("mode", ["offsets", "absolute", "head"])
def test_dlc_window_backward(valid, mode):
    import stereoformer_b200 as sf
    import stereoformer_b200.modules as M
    from stereoformer_b200 import update as U
    g = load_golden("dlc_window_valid%d" % valid)
    cv, cur, lo, up = g["cv"], g["cur"], g["lo"], g["up"]
    N = g["n_points"]
    torch.manual_seed(31)
    old = M.DEFAULT_DIV_MODE
    M.DEFAULT_DIV_MODE = 0
    try:
        if mode == "head":
            pre = torch.randn(cv.shape[0], 2, cv.shape[2], cv.shape[3]) * 3
            c, d, p_ = _leaf(cv), _leaf(cur), _leaf(pre)
            ref = O.head_window(c, d, p_, N, bool(valid))
            go = torch.randn_like(ref)
            refs = torch.autograd.grad(ref, [c, d, p_], go)
            leaves = [_leaf(t.cuda()) for t in (cv, cur, pre)]
            out = U._head_window(leaves[0], leaves[1], leaves[2], N, bool(valid))
            names = ("grad_cv", "grad_disp", "grad_head")
        else:
            a_lo, a_up = (cur - lo, cur + up) if mode == "absolute" else (lo, up)
            c, d, l_, u_ = _leaf(cv), _leaf(cur), _leaf(a_lo), _leaf(a_up)
            if mode == "absolute":
                D = cv.shape[1]
                cand, _ = O.window_candidates(l_, u_, d, D, N, bool(valid), "fixed")
                ref = O.window_expectation(c, cand)
            else:
                ref = O.dlc_window(c, d, l_, u_, N, bool(valid))
            go = torch.randn_like(ref)
            refs = torch.autograd.grad(ref, [c, d, l_, u_], go, allow_unused=True)
            leaves = [_leaf(t.cuda()) for t in (cv, cur, a_lo, a_up)]
            out = sf.dlc_window(*leaves, N, bool(valid), absolute=(mode == "absolute"), div_mode=0)
            names = ("grad_cv", "grad_disp", "grad_lo", "grad_up")
        got = torch.autograd.grad(out, leaves, go.cuda(), allow_unused=True)
    finally:
        M.DEFAULT_DIV_MODE = old
    for n, a, r in zip(names, got, refs):
        if r is None:
            assert a is None or float(a.abs().max()) == 0.0, n
            continue
        _close(a, r, n)


def test_sample_volume_backward():
    import stereoformer_b200 as sf
    g = load_golden("sample_volume")
    c, k = _leaf(g["cv"]), _leaf(g["cand"])
    ref = O.sample_volume(c, k)
    go = torch.randn_like(ref)
    rc, rk = torch.autograd.grad(ref, [c, k], go)
    cg, kg = _leaf(g["cv"].cuda()), _leaf(g["cand"].cuda())
    out = sf.build_cost_volume_from_volume(cg, kg)
    gc, gk = torch.autograd.grad(out, [cg, kg], go.cuda())
    _close(gc, rc, "grad_volume")
    _close(gk, rk, "grad_candidates")


@pytest.mark.parametrize("padding", ["border", "zeros"])
def test_disp_warp_backward(padding):
    import stereoformer_b200 as sf
    import stereoformer_b200.modules as M
    g = load_golden("disp_warp")
    img, disp = g["img"], g["disp"]
    i_, d_ = _leaf(img), _leaf(disp)
    ref, _ = O.disp_warp(i_, d_, padding)
    go = torch.randn_like(ref)
    ri, rd = torch.autograd.grad(ref, [i_, d_], go)
    old = M.DEFAULT_DIV_MODE
    M.DEFAULT_DIV_MODE = 0
    try:
        ig, dg = _leaf(img.cuda()), _leaf(disp.cuda())
        out, mask = sf.disp_warp(ig, dg, padding)
        assert not mask.requires_grad
        gi, gd = torch.autograd.grad(out, [ig, dg], go.cuda())
    finally:
        M.DEFAULT_DIV_MODE = old
    _close(gi, ri, "grad_img")
    _close(gd, rd, "grad_disp")


def test_lc_chain_trains():
    """LowCNN LC chain (baseline.py:114-127) under autograd: feature gradients arrive through window, soft-argmin and
    cost volume kernels and match autograd through the oracle chain."""
    import stereoformer_b200 as sf
    torch.manual_seed(37)
    fl, fr = torch.randn(1, 32, 6, 40), torch.randn(1, 32, 6, 40)
    a, b = _leaf(fl), _leaf(fr)
    cv = O.corr_volume(a, b, 8)
    d0 = O.soft_argmin(cv).unsqueeze(1)
    ref = O.local_cost_volume(cv, d0, 2, 10, True)
    ra, rb = torch.autograd.grad(ref.sum(), [a, b])
    ag, bg = _leaf(fl.cuda()), _leaf(fr.cuda())
    cvg = sf.CostVolume(8)(ag, bg)
    d0g = sf.DisparityEstimation(8)(cvg).unsqueeze(1)
    lc = sf.LocalCostVolume(2, 10)
    lc.div_mode = 0
    ga, gb = torch.autograd.grad(lc(cvg, d0g, True).sum(), [ag, bg])
    _close(ga, ra, "grad_left", tol=5e-4)
    _close(gb, rb, "grad_right", tol=5e-4)


def test_forward_only_ops_still_raise():
    """Operators without a backward kernel refuse to be differentiated instead of silently dropping gradients."""
    import stereoformer_b200 as sf
    q = torch.randn(1, 64, 2, 16, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    out = sf.EpipolarCrossAttention(8)(q, q.detach(), q.detach())
    with pytest.raises(NotImplementedError):
        out.float().sum().backward()


# ------------------------------------------------------------------------------------------ DLC-GRU step, trainable
def test_gru_gate_backward():
    import stereoformer_b200.ops as ops
    torch.manual_seed(41)
    B, C, H, W = 2, 8, 5, 12
    zb, gp, h0, x = torch.randn(B, 2 * C, H, W), torch.randn(B, C, H, W), torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    for softsign in (False, True):
        lz, lg, lh, lx = (_leaf(t) for t in (zb, gp, h0, x))
        s_ref = O.gru_reset(lz[:, C:], lh, lx)
        h_ref = O.gru_blend(lz[:, :C], lg, lh, softsign)
        gs, gh = torch.randn_like(s_ref), torch.randn_like(h_ref)
        r_b, r_h1, r_x = torch.autograd.grad(s_ref, [lz, lh, lx], gs)
        r_z, r_g, r_h2 = torch.autograd.grad(h_ref, [lz, lg, lh], gh)
        zbd = zb.cuda()
        gb, gh1, gx = ops.gru_reset_backward(zbd[:, C:], h0.cuda(), x.shape, gs.cuda())
        gz, gg, gh2 = ops.gru_blend_backward(zbd[:, :C], gp.cuda(), h0.cuda(), gh.cuda(), softsign)
        _close(gb, r_b[:, C:], "g_b_pre"); _close(gh1, r_h1, "g_prev_h(reset)"); _close(gx, r_x, "g_input")
        _close(gz, r_z[:, :C], "g_z_pre"); _close(gg, r_g, "g_g_pre"); _close(gh2, r_h2, "g_prev_h(blend)")


def test_upsample_convex_backward():
    import stereoformer_b200 as sf
    g = load_golden("upsample_convex8")
    d, m = _leaf(g["disp"]), _leaf(g["mask"])
    ref = O.upsample_convex(d, m, 8)
    go = torch.randn_like(ref)
    rd, rm = torch.autograd.grad(ref, [d, m], go)
    dg, mg = _leaf(g["disp"].cuda()), _leaf(g["mask"].cuda())
    gd, gm = torch.autograd.grad(sf.upsample_convex8(dg, mg), [dg, mg], go.cuda())
    _close(gd, rd, "grad_disp")
    _close(gm, rm, "grad_mask")


def test_update_step_trains():
    """One DisparityUpdateDLCWithMask step + convex upsampling under autograd: gradients w.r.t. the cost volume, the
    current disparity, the hidden state and every parameter against autograd through the oracle's step."""
    import stereoformer_b200 as sf
    import stereoformer_b200.modules as M
    from test_gpu_update import _load_step, _sd
    g = load_golden("dlc_update_step")
    old_tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    old = M.DEFAULT_DIV_MODE
    M.DEFAULT_DIV_MODE = 0
    try:
        step = _load_step(g)                             # eval(): BatchNorm uses running statistics, like the golden
        params = dict(step.named_parameters())
        cvg, dg, hg = _leaf(g["cv"].cuda()), _leaf(g["cur"].cuda()), _leaf(g["hidden0"].cuda())
        disp, hid, mask = step(cvg, dg, g["left"].cuda(), g["right"].cuda(), hg, True)
        full = sf.upsample_convex8(disp, mask)
        gen = torch.Generator().manual_seed(9)
        go_f, go_h = torch.randn(full.shape, generator=gen), torch.randn(hid.shape, generator=gen)
        loss = (full * go_f.cuda()).sum() + (hid * go_h.cuda()).sum()
        names = [n for n in params if not n.startswith(("gru.conv_zz", "gru.conv_bb", "gru.conv_gg"))]
        got = torch.autograd.grad(loss, [cvg, dg, hg] + [params[n] for n in names], allow_unused=True)
        # oracle on CPU
        sd = {k: v.detach().cpu().clone().requires_grad_(v.dtype.is_floating_point and "running" not in k)
              for k, v in step.state_dict().items()}
        c, d, h = _leaf(g["cv"]), _leaf(g["cur"]), _leaf(g["hidden0"])
        d1, h1, m1 = O.dlc_update_step(sd, c, d, g["left"], g["right"], h, 20, True)
        loss_ref = (O.upsample_convex(d1, m1, 8) * go_f).sum() + (h1 * go_h).sum()
        ref = torch.autograd.grad(loss_ref, [c, d, h] + [sd[n] for n in names], allow_unused=True)
    finally:
        M.DEFAULT_DIV_MODE = old
        torch.backends.cudnn.allow_tf32 = old_tf32
    for n, a, r in zip(["cv", "disp", "hidden"] + names, got, ref):
        assert (a is None) == (r is None), n
        if a is not None:
            _close(a, r, n, tol=2e-3)
