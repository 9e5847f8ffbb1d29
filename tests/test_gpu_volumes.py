"""GPU parity: cost volumes, soft-argmin, window sampling -- CUDA kernels (through the C ABI) vs the oracle
and the committed golden vectors of the unmodified reference.

Tolerances (BASELINE.json north_star): bit-exact window indices and sampling coordinates; <= 1e-4 relative on
cost volumes; <= 1e-3 px on disparities."""
import pytest
import torch

from conftest import load_golden
from oracle import stereo_oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-4


def dev(t):
    return t.cuda()


def close_rel(a, b, rtol=RTOL, scale=None):
    """|a-b| <= rtol * max(|b|, scale) elementwise; scale defaults to the rms of b (cost volumes are sums of
    products with cancellation, so the relative bound is taken against the volume's magnitude)."""
    a, b = a.float().cpu(), b.float().cpu()
    s = b.pow(2).mean().sqrt().item() if scale is None else scale
    return bool(((a - b).abs() <= rtol * torch.clamp(b.abs(), min=s)).all())


@pytest.fixture(scope="module")
def ops():
    import stereoformer_b200.ops as ops
    return ops


# ------------------------------------------------------------------------------------------ cost volumes
@pytest.mark.parametrize("tag", ["small", "odd", "wide_d"])
def test_corr_volume_golden(ops, tag):
    g = load_golden("corr_volume_" + tag)
    L, R, D = dev(g["left"]), dev(g["right"]), g["D"]
    out = ops.corr_volume(L, R, D)
    assert out.shape == g["corr"].shape
    assert close_rel(out, g["corr"])
    gen = ops.corr_volume(L, R, D, algo=ops.SF_ALGO_GENERIC)
    assert close_rel(gen, g["corr"])
    assert close_rel(ops.gwc_volume(L, R, D, 4), g["gwc"])
    assert torch.equal(ops.shift_volume(L, R, D, False).cpu(), g["diff"])      # bit-exact
    assert torch.equal(ops.shift_volume(L, R, D, True).cpu(), g["concat"])


@pytest.mark.parametrize("shape", [(1, 256, 32, 64, 24),      # C1
                                   (2, 256, 68, 120, 24),     # C2 (two pairs)
                                   (1, 64, 9, 360, 48),       # C5 row shape, D=48 (two disparity blocks)
                                   (1, 32, 5, 312, 8), (1, 40, 3, 264, 16), (2, 10, 4, 48, 32), (1, 24, 3, 96, 64),
                                   (1, 16, 2, 200, 100),      # D > 96: disparity tiles
                                   (1, 7, 3, 20, 24), (1, 3, 1, 8, 30)])  # ragged channel chunk, D > W
def test_corr_volume_oracle(ops, shape):
    B, C, H, W, D = shape
    torch.manual_seed(1024)
    L, R = torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    ref = O.corr_volume(L, R, D)
    out = ops.corr_volume(dev(L), dev(R), D, algo=ops.SF_ALGO_TMA)
    assert close_rel(out, ref)
    assert (out.cpu()[ref == 0] == 0).all()          # the x < d triangle is exactly zero
    gen = ops.corr_volume(dev(L), dev(R), D, algo=ops.SF_ALGO_GENERIC)
    assert close_rel(gen, ref)


@pytest.mark.parametrize("shape", [(1, 256, 32, 64, 24),      # C1
                                   (2, 256, 68, 120, 24),     # C2 (two pairs)
                                   (1, 64, 9, 360, 48),       # C5 row shape: three x segments, D=48
                                   (1, 32, 5, 312, 8), (1, 40, 3, 264, 16), (2, 10, 4, 48, 32), (1, 24, 3, 96, 64),
                                   (1, 48, 2, 200, 65),       # widest halo (64) + a clipped second segment
                                   (1, 7, 3, 20, 24), (1, 3, 1, 8, 30),   # ragged channel chunk, D > W
                                   (2, 72, 2, 128, 40), (3, 33, 7, 132, 1)])
def test_corr_volume_tensor_core(ops, shape):
    """tcgen05 path (bf16 hi/mid split, fp32 TMEM accumulation) against the fp32 oracle at the same 1e-4 bound."""
    B, C, H, W, D = shape
    torch.manual_seed(1024)
    L, R = torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    ref = O.corr_volume(L, R, D)
    out = ops.corr_volume(dev(L), dev(R), D, algo=ops.SF_ALGO_TC)
    assert close_rel(out, ref)
    assert (out.cpu()[ref == 0] == 0).all()          # the x < d triangle is exactly zero
    # large dynamic range: the split must track fp32, not bf16, accuracy
    L2 = L * torch.logspace(-3, 3, C).view(1, C, 1, 1)
    ref2 = O.corr_volume(L2, R, D)
    assert close_rel(ops.corr_volume(dev(L2), dev(R), D, algo=ops.SF_ALGO_TC), ref2)


def test_corr_volume_tensor_core_limits(ops):
    x = torch.randn(1, 32, 2, 64, device="cuda")
    with pytest.raises(ValueError):
        ops.corr_volume(x, x, 70, algo=ops.SF_ALGO_TC)         # halo > 64 pixels
    assert close_rel(ops.corr_volume(x, x, 70), O.corr_volume(x.cpu(), x.cpu(), 70))   # AUTO falls back


def test_corr_volume_groups(ops):
    torch.manual_seed(3)
    L, R = torch.randn(2, 48, 6, 64), torch.randn(2, 48, 6, 64)
    for G in (1, 4, 8, 48):
        ref = O.gwc_volume(L, R, 12, G)
        assert close_rel(ops.gwc_volume(dev(L), dev(R), 12, G), ref)


def test_corr_volume_shift_property_full_size(ops):
    """C2-sized known answer: right = left shifted by k pixels  =>  argmax_d cv = k for x >= k, and the
    volume is linear in the left features."""
    torch.manual_seed(7)
    B, C, H, W, D, k = 8, 256, 68, 120, 24, 5
    L = torch.randn(B, C, H, W, device="cuda")
    R = torch.zeros_like(L)
    R[..., : W - k] = L[..., k:]
    cv = ops.corr_volume(L, R, D)
    assert (cv[..., k:].argmax(dim=1) == k).all()
    cv2 = ops.corr_volume(2 * L, R, D)
    assert torch.allclose(cv2, 2 * cv, rtol=1e-6, atol=1e-6)
    # unaligned base pointer falls back to the generic kernel and must agree
    buf = torch.empty(L.numel() + 1, device="cuda")
    Lu = buf[1:].view_as(L).copy_(L)
    assert torch.allclose(ops.corr_volume(Lu, R, D), cv, rtol=1e-5, atol=1e-5)


def test_argument_errors(ops):
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        ops.corr_volume(torch.randn(1, 4, 2, 8), torch.randn(1, 4, 2, 8), 4)
    x = torch.randn(1, 6, 2, 8, device="cuda")
    with pytest.raises(ValueError):
        ops.corr_volume(x, x, 4, groups=4)
    with pytest.raises(ValueError):
        ops.corr_volume(x[..., :7].contiguous(), x[..., :7].contiguous(), 4, algo=ops.SF_ALGO_TMA)
    with pytest.raises(TypeError):
        ops.corr_volume(x.double(), x.double(), 4)


# ------------------------------------------------------------------------------------------ soft-argmin
def test_soft_argmin_golden(ops):
    g = load_golden("soft_argmin")
    cv = dev(g["cv"])
    assert (ops.soft_argmin(cv).cpu() - g["disp_sim"]).abs().max() < 1e-3
    assert (ops.soft_argmin(cv, match_similarity=False).cpu() - g["disp_cost"]).abs().max() < 1e-3
    d, p = ops.soft_argmin(cv, with_prob=True)
    assert torch.allclose(p.cpu(), g["prob"], rtol=1e-4, atol=1e-6)
    assert (d.cpu() - g["disp_sim"]).abs().max() < 1e-3


@pytest.mark.parametrize("shape", [(8, 24, 68, 120), (1, 48, 31, 77), (2, 7, 5, 33), (1, 192, 8, 40)])
def test_soft_argmin_oracle(ops, shape):
    torch.manual_seed(5)
    cv = torch.randn(*shape) * 4
    ref = O.soft_argmin(cv)
    assert (ops.soft_argmin(dev(cv)).cpu() - ref).abs().max() < 1e-3


# ------------------------------------------------------------------------------------------ window sampling
def _check_window(out, dbg, g, exact_out=False):
    assert torch.equal(dbg["cand"].cpu(), g["cand"])                           # sampling coordinates: bit-exact
    assert torch.equal(dbg["idx_floor"].cpu().long(), g["idx_floor"])          # window indices: bit-exact
    assert torch.equal(dbg["idx_ceil"].cpu().long(), g["idx_ceil"])
    if "local" in g:
        assert torch.equal(dbg["local"].cpu(), g["local"])                     # interpolated samples: bit-exact
    assert (out.cpu() - g["out"]).abs().max() < 1e-3
    assert torch.allclose(out.cpu(), g["out"], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("valid", [0, 1])
def test_fixed_window_golden(ops, valid):
    g = load_golden("local_window_valid%d" % valid)
    out, dbg = ops.local_window(dev(g["cv"]), dev(g["cur"]), mode=ops.SF_WIN_FIXED, radius=g["radius"],
                                n_points=g["n_points"], consider_valid=bool(valid), div_mode=ops.SF_DIV_TRUE,
                                debug=True)
    _check_window(out, dbg, g)


@pytest.mark.parametrize("valid", [0, 1])
def test_dlc_window_golden(ops, valid):
    g = load_golden("dlc_window_valid%d" % valid)
    out, dbg = ops.local_window(dev(g["cv"]), dev(g["cur"]), dev(g["lo"]), dev(g["up"]), mode=ops.SF_WIN_OFFSETS,
                                n_points=g["n_points"], consider_valid=bool(valid), div_mode=ops.SF_DIV_TRUE,
                                debug=True)
    _check_window(out, dbg, g)
    # absolute-bounds mode (pure_learning_method.py:112) must give the same window
    out2 = ops.local_window(dev(g["cv"]), dev(g["cur"]), dev(g["cur"] - g["lo"]), dev(g["cur"] + g["up"]),
                            mode=ops.SF_WIN_ABSOLUTE, n_points=g["n_points"], consider_valid=bool(valid),
                            div_mode=ops.SF_DIV_TRUE)
    assert torch.equal(out2, out)


@pytest.mark.parametrize("valid", [0, 1])
def test_variance_window_indices_bit_exact_given_reference_bounds(ops, valid):
    """a5 (VERDICT r1 missing #4): with the reference's own bounds disp -/+ gamma * sigma fed in, the variance window's
    candidates, floor / ceil indices and interpolated samples are bit-exact -- the validity rule `upper > x`, the
    [0, D-1] clamps and the index arithmetic of variance_local_cost_volume.py:44-66 (the only part of a5 that differs
    from a4) carry no tolerance."""
    g = load_golden("variance_window_valid%d" % valid)
    out, dbg = ops.local_window(dev(g["cv"]), dev(g["cur"]), dev(g["lo"]), dev(g["up"]),
                                mode=ops.SF_WIN_VARIANCE_BOUNDS, n_points=g["n_points"], consider_valid=bool(valid),
                                div_mode=ops.SF_DIV_TRUE, debug=True)
    _check_window(out, dbg, g)


@pytest.mark.parametrize("valid", [0, 1])
def test_variance_window_golden(ops, valid):
    """a5 end to end (the kernel derives sigma itself: exp() and the sum order differ from ATen in the last bits, so
    the bounds are not bit-comparable).  Every pixel whose indices differ from the reference's must be explained by a
    bound or candidate sitting within float rounding of a decision threshold; all other pixels must match to 1e-3 px
    with identical indices."""
    g = load_golden("variance_window_valid%d" % valid)
    out, dbg = ops.local_window(dev(g["cv"]), dev(g["cur"]), mode=ops.SF_WIN_VARIANCE, gamma=g["gamma"],
                                n_points=g["n_points"], consider_valid=bool(valid), div_mode=ops.SF_DIV_TRUE,
                                debug=True)
    D, W = g["cv"].shape[1], g["cv"].shape[3]
    same = (dbg["idx_floor"].cpu().long() == g["idx_floor"]).all(1) & (dbg["idx_ceil"].cpu().long() == g["idx_ceil"]).all(1)
    err = (out.cpu() - g["out"]).abs().squeeze(1)
    assert err[same].max() < 1e-3
    assert (dbg["cand"].cpu() - g["cand"])[same.unsqueeze(1).expand_as(g["cand"])].abs().max() < 1e-4
    # pixels with different indices: a candidate within 1e-4 of an integer (ceil / floor flips), or a bound within
    # 1e-4 of a validity threshold (0, D-1, x)
    cand, lo, up = g["cand"], g["lo"].squeeze(1), g["up"].squeeze(1)
    near_int = ((cand - cand.round()).abs() < 1e-4).any(1)
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, W)
    near_thr = (lo.abs() < 1e-4) | ((up - (D - 1)).abs() < 1e-4) | ((up - xs).abs() < 1e-4)
    unexplained = ~same & ~near_int & ~near_thr
    print("variance window valid=%d: %d of %d pixels differ in an index, %d unexplained; max err %.3g"
          % (valid, int((~same).sum()), same.numel(), int(unexplained.sum()), float(err.max())))
    assert int(unexplained.sum()) == 0
    assert float((~same).float().mean()) < 0.02


def test_sample_volume_golden(ops):
    g = load_golden("sample_volume")
    out, jf, ic = ops.sample_volume(dev(g["cv"]), dev(g["cand"]), return_indices=True)
    assert torch.equal(out.cpu(), g["local"])
    assert torch.equal(jf.cpu().long(), g["idx_floor"]) and torch.equal(ic.cpu().long(), g["idx_ceil"])


@pytest.mark.parametrize("shape", [(8, 24, 68, 120), (1, 48, 20, 360), (2, 24, 7, 37), (1, 300, 3, 16)])
@pytest.mark.parametrize("recip", [False, True])
def test_window_oracle_sizes(ops, shape, recip):
    """Larger / ragged shapes (TMA-staged and fallback staging) against the oracle, both division roundings."""
    torch.manual_seed(11)
    B, D, H, W = shape
    cv = torch.abs(torch.randn(B, D, H, W) * 3 - 2)
    cur = torch.rand(B, 1, H, W) * (D - 1)
    for valid in (True, False):
        cand, _ = O.window_candidates(cur - 2, cur + 2, cur, D, 20, valid, recip=recip)
        jf, ic = O.window_indices(cand, D)
        ref = O.window_expectation(cv, cand)
        out, dbg = ops.local_window(dev(cv), dev(cur), mode=ops.SF_WIN_FIXED, radius=2, n_points=20,
                                    consider_valid=valid, div_mode=ops.SF_DIV_RECIP if recip else ops.SF_DIV_TRUE,
                                    debug=True)
        assert torch.equal(dbg["cand"].cpu(), cand)
        assert torch.equal(dbg["idx_floor"].cpu().long(), jf) and torch.equal(dbg["idx_ceil"].cpu().long(), ic)
        assert (out.cpu() - ref).abs().max() < 1e-3


def test_window_matches_pytorch_cuda_division(ops):
    """The reference run on a GPU divides by Python scalars through PyTorch's CUDA kernels (x * (1/s)).
    Running the oracle's tensor program on the device must reproduce the kernel's SF_DIV_RECIP candidates."""
    torch.manual_seed(13)
    cv = torch.abs(torch.randn(2, 24, 16, 120, device="cuda") * 3 - 2)
    cur = torch.rand(2, 1, 16, 120, device="cuda") * 23
    cand, _ = O.window_candidates(cur - 2, cur + 2, cur, 24, 20, True)           # torch CUDA ops, true '/'
    _, dbg = ops.local_window(cv, cur, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                              div_mode=ops.SF_DIV_RECIP, debug=True)
    assert torch.equal(dbg["cand"], cand)


def test_fused_softargmin_window(ops):
    """disp=None: the current disparity is the soft-argmin of the same volume (LowCNN baseline.py:118-127)."""
    torch.manual_seed(17)
    cv = torch.randn(2, 24, 9, 120) * 2
    d0 = O.soft_argmin(cv).unsqueeze(1)
    ref = O.local_cost_volume(cv, d0, 2, 20, True)
    out, got0 = ops.local_window(dev(cv), None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                                 return_disp0=True)
    assert (got0.cpu() - d0).abs().max() < 1e-3
    # the window is placed from the kernel's own soft-argmin; compare against the oracle fed the same d0
    ref2 = O.local_cost_volume(cv, got0.cpu(), 2, 20, True)
    assert (out.cpu() - ref2).abs().max() < 1e-3
    assert (out.cpu() - ref).abs().median() < 1e-4


def test_invalid_window_returns_current_disparity(ops):
    """KAT (SURVEY 8c iii): invalid-mask pixels => all candidates == disp => output == disp exactly."""
    cv = torch.randn(1, 24, 4, 40, device="cuda")
    cur = torch.full((1, 1, 4, 40), 0.5, device="cuda")     # lower bound < 0 -> invalid everywhere
    out = ops.local_window(cv, cur, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True)
    assert torch.equal(out, cur)


def test_integer_candidates_return_volume_entries(ops):
    """KAT (SURVEY 8c ii): integer candidates => build_cost_volume_from_volume returns exact volume entries."""
    torch.manual_seed(19)
    cv = torch.randn(2, 24, 5, 40, device="cuda")
    idx = torch.randint(0, 24, (2, 6, 5, 40), device="cuda")
    out = ops.sample_volume(cv, idx.float())
    assert torch.equal(out, torch.gather(cv, 1, idx))


def test_window_fused_peer_gather_single_rank(ops):
    """sf_local_window_gather_f32: the result is also stored at peer_offset of every buffer in the peer list
    (here two local buffers stand in for the NVLink peer mappings of the multi-GPU bench)."""
    torch.manual_seed(5)
    B, D, H, W = 2, 24, 7, 40
    cv = (torch.randn(B, D, H, W, device="cuda") * 3 - 2).abs()
    ref = ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True)
    n = B * H * W
    bufs = [torch.zeros(3 * n, device="cuda") for _ in range(2)]
    out = ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                           peers=[b.data_ptr() for b in bufs], peer_offset=n)
    assert torch.equal(out, ref)
    for b in bufs:
        assert torch.equal(b[n:2 * n].view_as(ref), ref)
        assert b[:n].abs().sum() == 0 and b[2 * n:].abs().sum() == 0
    with pytest.raises(ValueError):
        ops.local_window(cv, None, n_points=20, peers=[bufs[0].data_ptr()] * 9)


@pytest.mark.parametrize("name,shape", [("C2", (8, 256, 68, 120, 24)), ("C5", (1, 256, 248, 360, 48))])
def test_whole_chain_epe_at_baseline_shapes(ops, name, shape):
    """the default (tensor-core) cost volume -> soft-argmin -> local window chain against the fp32
    oracle chain at the BASELINE shapes, asserted on the MEAN (EPE) and the MAX of |disp - ref|, not on a median.
    north_star: <= 1e-3 px EPE on the output disparity (1/8-scale px here; x8 at full resolution is asserted too)."""
    torch.manual_seed(1024)
    B, C, H, W, D = shape
    fl, fr = torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    # features with structure: the right map is the left map shifted by a smooth disparity, plus noise, so that the
    # soft-argmin is peaked and windows sit all over [0, D-1] (incl. the validity edges)
    disp = torch.rand(B, 1, H, 1) * (D - 2)
    xs = torch.arange(W).view(1, 1, 1, W).float()
    idx = (xs - disp.round()).long().clamp(0, W - 1).expand(B, C, H, W)
    fr = 0.7 * torch.gather(fl, 3, idx) + 0.3 * fr
    fl, fr = fl * 4, fr * 4
    cv = ops.corr_volume(dev(fl), dev(fr), D)
    out = ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                           div_mode=ops.SF_DIV_RECIP)
    cv_ref = O.corr_volume(dev(fl), dev(fr), D)                 # oracle tensor program, fp32, evaluated on the device
    ref = O.local_cost_volume(cv_ref, O.soft_argmin(cv_ref).unsqueeze(1), 2, 20, True, recip=True)
    err = (out - ref).abs()
    rel = float((cv - cv_ref).abs().max() / cv_ref.abs().max())
    print("%s chain: cost volume %.2e of max, disparity EPE %.3e px, max %.3e px, pixels over 1e-3: %d of %d"
          % (name, rel, float(err.mean()), float(err.max()), int((err > 1e-3).sum()), err.numel()))
    assert rel <= 1e-4
    assert float(err.mean()) * 8 <= 1e-3                        # EPE at full resolution
    # the maximum is bounded too: a pixel may only be far off where a validity decision (lower < 0, upper >= D-1) of the
    # reference sits within float rounding of its threshold
    d0 = O.soft_argmin(cv_ref).unsqueeze(1)
    near = ((d0 - 2).abs() < 1e-3) | ((d0 + 2 - (D - 1)).abs() < 1e-3)
    assert float(err[~near].max()) <= 1e-2


def test_bound_tensors_are_size_checked(ops):
    """A wrongly shaped lower / upper bound must be an error, not an out-of-bounds device read."""
    cv = torch.randn(1, 8, 4, 16, device="cuda")
    disp = torch.rand(1, 1, 4, 16, device="cuda") * 7
    good = torch.rand(1, 1, 4, 16, device="cuda")
    bad = (ValueError, RuntimeError)            # the ctypes binding raises ValueError, the torch-extension front end RuntimeError
    with pytest.raises(bad):
        ops.local_window(cv, disp, lo=good[:, :, :2].contiguous(), up=good, mode=ops.SF_WIN_OFFSETS, n_points=10)
    with pytest.raises(bad):
        ops.local_window(cv, disp, lo=good, up=None, mode=ops.SF_WIN_ABSOLUTE, n_points=10)
    with pytest.raises(bad):
        ops.local_window(cv, disp, lo=good, mode=ops.SF_WIN_HEAD, n_points=10)
    with pytest.raises(ValueError, match="elements"):          # debug outputs force the ctypes binding
        ops.local_window(cv, disp, lo=good[:, :, :2].contiguous(), up=good, mode=ops.SF_WIN_OFFSETS, n_points=10, debug=True)
    with pytest.raises(ValueError, match="elements"):
        ops.local_window_backward(cv, disp, good[:, :, :2].contiguous(), good, good, ops.SF_WIN_OFFSETS, 0.0, 10, True)
    ops.local_window(cv, disp, lo=good, up=good + 1, mode=ops.SF_WIN_OFFSETS, n_points=10)
