"""Pins oracle/stereo_oracle.py to outputs of the unmodified reference (tests/golden/*.npz,
made by oracle/gen_golden.py).  CPU only."""
import pytest
import torch

from oracle import stereo_oracle as O
from conftest import load_golden


def eq(a, b):
    return torch.equal(a, b)


@pytest.mark.parametrize("tag", ["small", "odd", "wide_d"])
def test_cost_volumes(tag):
    g = load_golden("corr_volume_" + tag)
    L, R, D = g["left"], g["right"], g["D"]
    assert eq(O.corr_volume(L, R, D), g["corr"])
    assert eq(O.diff_volume(L, R, D), g["diff"])
    assert eq(O.concat_volume(L, R, D), g["concat"])
    assert eq(O.gwc_volume(L, R, D, 4), g["gwc"])


def test_soft_argmin():
    g = load_golden("soft_argmin")
    assert eq(O.soft_argmin(g["cv"]), g["disp_sim"])
    assert eq(O.soft_argmin(g["cv"], match_similarity=False), g["disp_cost"])
    d, p = O.soft_argmin(g["cv"], with_prob=True)
    assert eq(p, g["prob"])


@pytest.mark.parametrize("valid", [0, 1])
def test_fixed_window(valid):
    g = load_golden("local_window_valid%d" % valid)
    cv, cur, D = g["cv"], g["cur"], g["cv"].shape[1]
    cand, _ = O.window_candidates(cur - g["radius"], cur + g["radius"], cur, D, g["n_points"], bool(valid))
    assert eq(cand, g["cand"])                       # sampling coordinates: bit-exact
    jf, ic = O.window_indices(cand, D)
    assert eq(jf, g["idx_floor"]) and eq(ic, g["idx_ceil"])   # window indices: bit-exact
    assert eq(O.sample_volume(cv, cand), g["local"])
    assert eq(O.local_cost_volume(cv, cur, g["radius"], g["n_points"], bool(valid)), g["out"])


def test_sample_volume_arbitrary_candidates():
    """Out-of-range and integer candidates through build_cost_volume_from_volume; the
    straight-through ceil of fixed_local_cost_volume.py:5 is evaluated as written."""
    g = load_golden("sample_volume")
    jf, ic = O.window_indices(g["cand"], g["cv"].shape[1])
    assert eq(jf, g["idx_floor"]) and eq(ic, g["idx_ceil"])
    assert eq(O.sample_volume(g["cv"], g["cand"]), g["local"])


@pytest.mark.parametrize("valid", [0, 1])
def test_variance_window(valid):
    g = load_golden("variance_window_valid%d" % valid)
    assert eq(O.disparity_sigma(torch.softmax(g["cv"], 1), g["cur"]), g["sigma"])
    assert eq(O.variance_local_cost_volume(g["cv"], g["cur"], g["gamma"], g["n_points"], bool(valid)), g["out"])
    # window indices / candidates of the variance window (north_star: bit-exact), from the reference's own bounds
    cv, cur, D, W = g["cv"], g["cur"], g["cv"].shape[1], g["cv"].shape[3]
    sigma = O.disparity_sigma(torch.softmax(cv, 1), cur)
    lower, upper = cur - g["gamma"] * sigma, cur + g["gamma"] * sigma
    assert eq(lower, g["lo"]) and eq(upper, g["up"])
    extra = (upper > torch.arange(W, dtype=cur.dtype).view(1, 1, 1, W)).float() if valid else None
    cand, _ = O.window_candidates(lower, upper, cur, D, g["n_points"], bool(valid), "variance", extra)
    assert eq(cand, g["cand"])
    jf, ic = O.window_indices(cand, D)
    assert eq(jf, g["idx_floor"]) and eq(ic, g["idx_ceil"])


@pytest.mark.parametrize("valid", [0, 1])
def test_dlc_window(valid):
    g = load_golden("dlc_window_valid%d" % valid)
    cv, cur, D = g["cv"], g["cur"], g["cv"].shape[1]
    cand, _ = O.window_candidates(cur - g["lo"], cur + g["up"], cur, D, g["n_points"], bool(valid))
    assert eq(cand, g["cand"])
    jf, ic = O.window_indices(cand, D)
    assert eq(jf, g["idx_floor"]) and eq(ic, g["idx_ceil"])
    assert eq(O.dlc_window(cv, cur, g["lo"], g["up"], g["n_points"], bool(valid)), g["out"])


def test_uncertainty():
    g = load_golden("uncertainty")
    assert eq(O.uncertainty_volume(g["cv"], g["cur"]), g["unc"])


def test_disp_warp():
    g = load_golden("disp_warp")
    H, W = g["img"].shape[-2:]
    assert eq(O.warp_grid(g["disp"], H, W), g["grid"])           # normalised coordinates: bit-exact
    wb, mb = O.disp_warp(g["img"], g["disp"])
    assert eq(wb, g["warped_border"]) and eq(mb, g["mask_border"])
    wz, mz = O.disp_warp(g["img"], g["disp"], "zeros")
    assert eq(wz, g["warped_zeros"]) and eq(mz, g["mask_zeros"])


def test_corr1d():
    g = load_golden("corr1d")
    pyr = O.corr1d_pyramid(g["f1"], g["f2"], g["num_levels"])
    for i, p in enumerate(pyr):
        assert eq(p, g["pyr%d" % i])
    assert eq(O.corr1d_lookup(pyr, g["coords"], g["radius"], g["num_levels"]), g["lookup"])


def test_upsample_simple8():
    g = load_golden("upsample_simple8")
    out = O.upsample_simple(g["disp"], 8)
    assert (out - g["out"]).abs().max() < 1e-4                     # px, values up to 160 (2 ulp)
    d = g["disp"].clone().requires_grad_(True)
    gd = torch.autograd.grad(O.upsample_simple(d, 8), d, g["grad_out"])[0]
    assert torch.allclose(gd, g["grad_disp"], rtol=1e-5, atol=1e-4)


def test_upsample_convex8():
    g = load_golden("upsample_convex8")
    assert eq(O.upsample_convex(g["disp"], g["mask"], 8), g["out"])


def test_attention_is_spec_derived_unpinned():
    """No reference exists (empty file); only internal consistency can be checked."""
    torch.manual_seed(0)
    q, k, v = (torch.randn(1, 8, 2, 12) for _ in range(3))
    o, s = O.epipolar_attention(q, k, v, band=4)
    # band scores -> softmax -> explicit PV must reproduce O
    p = torch.softmax(s, dim=1)
    ref = torch.zeros_like(o)
    for d in range(4):
        ref[..., d:] += p[:, d:d + 1, :, d:] * v[..., : 12 - d]
    assert torch.allclose(o, ref, atol=1e-5)


def _sd(g):
    return {k[3:]: v for k, v in g.items() if k.startswith("sd_")}


def test_conv_gru_gates():
    """f2: the gate arithmetic around the three convolutions, pinned on the reference ConvGRU's own intermediate
    tensors (forward hooks on conv_z / conv_b / conv_g)."""
    g = load_golden("conv_gru")
    assert eq(O.gru_reset(g["b_pre"], g["prev_h"], g["input"]), g["s"])
    assert eq(O.gru_blend(g["z_pre"], g["g_pre"], g["prev_h"]), g["h"])
    assert eq(O.conv_gru(_sd(g), "", g["input"], g["prev_h"]), g["h"])
    assert eq(O.conv_gru(_sd(g), "", g["input"], g["prev_h"], softsign=True), load_golden("conv_gru_softsign")["h"])


def test_dlc_update_step():
    """f2: two chained DisparityUpdateDLCWithMask steps (consider_valid True then False) from the state dict."""
    g = load_golden("dlc_update_step")
    sd = _sd(g)
    d1, h1, m1 = O.dlc_update_step(sd, g["cv"], g["cur"], g["left"], g["right"], g["hidden0"], 20, True)
    assert eq(h1, g["hidden1"]) and eq(m1, g["mask1_head"]) and eq(d1, g["disp1"])
    d2, h2, m2 = O.dlc_update_step(sd, g["cv"], d1, g["left"], g["right"], h1, 20, False)
    assert eq(h2, g["hidden2"]) and eq(m2, g["mask2_head"]) and eq(d2, g["disp2"])


# ------------------------------------------------------------------------------------------ f3 gradients
def _grads(fn, g, name, *ts):
    leaves = [t.clone().requires_grad_(True) for t in ts]
    out = fn(*leaves)
    out = out[0] if isinstance(out, tuple) else out
    got = torch.autograd.grad(out, leaves, g[name + "_0"], allow_unused=True)
    for k, a in enumerate(got):
        ref = g.get("%s_%d" % (name, k + 1))
        assert (a is None) == (ref is None), (name, k)
        if a is not None:
            assert torch.allclose(a, ref, rtol=1e-5, atol=1e-6), (name, k, (a - ref).abs().max())


def test_oracle_autograd_matches_reference_autograd():
    """The oracle restates the reference with the same differentiable torch ops; its autograd graph (straight-through
    ceil/floor, clamps, constant masks, grid_sample) must give the gradients of the unmodified reference modules."""
    g = load_golden("gradients")
    cvs = load_golden("corr_volume_small")
    L, R, D = cvs["left"], cvs["right"], cvs["D"]
    _grads(lambda l, r: O.corr_volume(l, r, D), g, "corr", L, R)
    _grads(lambda l, r: O.gwc_volume(l, r, D, 4), g, "gwc", L, R)
    cv = load_golden("soft_argmin")["cv"]
    _grads(lambda c: O.soft_argmin(c, True), g, "argmin_sim", cv)
    _grads(lambda c: O.soft_argmin(c, False), g, "argmin_cost", cv)
    for valid in (0, 1):
        w = load_golden("local_window_valid%d" % valid)
        _grads(lambda c, d: O.local_cost_volume(c, d, w["radius"], w["n_points"], bool(valid)), g,
               "win_valid%d" % valid, w["cv"], w["cur"])
    s = load_golden("sample_volume")
    _grads(O.sample_volume, g, "sample", s["cv"], s["cand"])
    wp = load_golden("disp_warp")
    _grads(lambda i, d: O.disp_warp(i, d, "border"), g, "warp_border", wp["img"], wp["disp"])
    _grads(lambda i, d: O.disp_warp(i, d, "zeros"), g, "warp_zeros", wp["img"], wp["disp"])


def test_oracle_variance_window_gradients_match_reference():
    """a5 under autograd: gradients of the unmodified Variance_Based_Local_Cost_Volume (through sigma = sqrt(sum p (d -
    disp)^2), the [0, D-1] clamps / the three invalid masks and the sampler) against the oracle's graph."""
    g = load_golden("gradients_variance")
    for valid in (0, 1):
        w = load_golden("variance_window_valid%d" % valid)
        _grads(lambda c, d: O.variance_local_cost_volume(c, d, float(w["gamma"]), int(w["n_points"]), bool(valid)), g,
               "var_valid%d" % valid, w["cv"], w["cur"])


def test_corr1d_alt():
    """a9 alternative (PytorchAlternateCorrBlock1D, corr/corr.py:8-51): the oracle reproduces the reference bit for bit."""
    g = load_golden("corr1d_alt")
    assert eq(O.corr1d_alt(g["f1"], g["f2"], g["coords"], int(g["radius"]), int(g["num_levels"])), g["out"])
