"""GPU parity for the DLC-GRU refinement step: ConvGRU gate kernels, the head-fused window
mode, the mask-scale-fused upsampling, the drop-in update modules and the graph-captured loop -- against the
oracle (pinned on the reference by tests/test_oracle_golden.py) and the golden vectors."""
import pytest
import torch

from conftest import load_golden
from oracle import stereo_oracle as O

pytestmark = pytest.mark.gpu


def dev(t):
    return t.cuda()


@pytest.fixture(scope="module")
def ops():
    import stereoformer_b200.ops as ops
    return ops


@pytest.fixture(autouse=True)
def fp32_convs():
    """The golden vectors are fp32 CPU runs: keep cuDNN from using TF32 for the convolutions around the kernels."""
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32 = old


def _sd(g):
    return {k[3:]: v for k, v in g.items() if k.startswith("sd_")}


# ------------------------------------------------------------------------------------------ gate kernels
def test_gru_gates_golden(ops):
    g = load_golden("conv_gru")
    s = ops.gru_reset(dev(g["b_pre"]), dev(g["prev_h"]), dev(g["input"]))
    assert torch.allclose(s.cpu(), g["s"], rtol=1e-6, atol=1e-6)
    C = g["b_pre"].shape[1]
    assert torch.equal(s.cpu()[:, C:], g["input"])                       # the concatenated half is a copy
    h = ops.gru_blend(dev(g["z_pre"]), dev(g["g_pre"]), dev(g["prev_h"]))
    assert torch.allclose(h.cpu(), g["h"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("shape", [(2, 8, 5, 12), (1, 3, 7, 9), (3, 64, 17, 30), (8, 64, 68, 120)])
@pytest.mark.parametrize("softsign", [False, True])
def test_gru_gates_random(ops, shape, softsign):
    """Vector and scalar paths (C*H*W % 4 != 0), gates as channel slices of one [B,2C,H,W] tensor (batch stride),
    prev_h = None (update.py:38-39: zeros)."""
    B, C, H, W = shape
    torch.manual_seed(B * 1000 + C)
    zb = torch.randn(B, 2 * C, H, W) * 3
    g_pre, h0, x = torch.randn(B, C, H, W) * 3, torch.randn(B, C, H, W), torch.randn(B, C, H, W)
    zbd = dev(zb)
    s = ops.gru_reset(zbd[:, C:], dev(h0), dev(x))
    assert torch.allclose(s.cpu(), O.gru_reset(zb[:, C:], h0, x), rtol=1e-6, atol=1e-6)
    h = ops.gru_blend(zbd[:, :C], dev(g_pre), dev(h0), softsign)
    assert torch.allclose(h.cpu(), O.gru_blend(zb[:, :C], g_pre, h0, softsign), rtol=1e-6, atol=2e-6)
    s0 = ops.gru_reset(zbd[:, C:], None, dev(x))
    assert torch.equal(s0.cpu(), O.gru_reset(zb[:, C:], None, x))
    hz = ops.gru_blend(zbd[:, :C], dev(g_pre), None, softsign)
    assert torch.allclose(hz.cpu(), O.gru_blend(zb[:, :C], g_pre, None, softsign), rtol=1e-6, atol=2e-6)


def test_gru_gate_errors(ops):
    x = torch.randn(1, 4, 3, 5, device="cuda")
    with pytest.raises(ValueError):
        ops.gru_reset(torch.randn(1, 4, 3, 6, device="cuda"), None, x)
    with pytest.raises(RuntimeError):
        ops.gru_reset(torch.randn(1, 4, 3, 5), None, x)                  # CPU tensor: no CPU path
    with pytest.raises(ValueError):
        ops.gru_blend(x, x, torch.randn(1, 5, 3, 5, device="cuda"))


# ------------------------------------------------------------------------------------------ fused head / mask scale
@pytest.mark.parametrize("valid", [0, 1])
def test_head_window_equals_offsets_mode(ops, valid):
    """SF_WIN_HEAD on the pre-ReLU head output == SF_WIN_OFFSETS on relu + chunk of it, bit for bit."""
    g = load_golden("dlc_window_valid%d" % valid)
    torch.manual_seed(5 + valid)
    cv, cur = dev(g["cv"]), dev(g["cur"])
    pre = torch.randn(cv.shape[0], 2, cv.shape[2], cv.shape[3], device="cuda") * 3
    lo, up = torch.chunk(torch.relu(pre), 2, dim=1)
    a = ops.local_window(cv, cur, lo=pre, mode=ops.SF_WIN_HEAD, n_points=20, consider_valid=bool(valid),
                         div_mode=ops.SF_DIV_TRUE)
    b = ops.local_window(cv, cur, lo.contiguous(), up.contiguous(), mode=ops.SF_WIN_OFFSETS, n_points=20,
                         consider_valid=bool(valid), div_mode=ops.SF_DIV_TRUE)
    assert torch.equal(a, b)
    ref = O.head_window(g["cv"], g["cur"], pre.cpu(), 20, bool(valid))
    assert (a.cpu() - ref).abs().max() < 1e-3
    with pytest.raises(ValueError):
        ops.local_window(cv, cur, mode=ops.SF_WIN_HEAD, n_points=20)     # head tensor missing


def test_upsample_mask_scale(ops):
    g = load_golden("upsample_convex8")
    d, m = dev(g["disp"]), dev(g["mask"])
    a = ops.upsample_convex(d, m, 8, mask_scale=0.25)
    assert torch.equal(a, ops.upsample_convex(d, 0.25 * m, 8))           # power-of-two scale: identical bits
    assert torch.allclose(a.cpu(), O.upsample_convex(g["disp"], g["mask"], 8, 0.25), rtol=1e-4, atol=1e-4)
    b = ops.upsample_convex(d, m, 8, mask_scale=0.3)
    assert torch.allclose(b.cpu(), O.upsample_convex(g["disp"], g["mask"], 8, 0.3), rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------------------------------ drop-in modules
def _load_step(g):
    import stereoformer_b200.update as U
    step = U.DisparityUpdateDLCWithMask(input_channels=24, hidden_dim=4, sample_points=20).eval()
    sd = _sd(g)
    own = step.state_dict()
    for k in own:
        if k in sd and own[k].shape == sd[k].shape:
            own[k] = sd[k]
    own["mask.2.weight"][:18] = sd["mask.2.weight"]
    own["mask.2.bias"][:18] = sd["mask.2.bias"]
    for n in ("z", "b", "g"):                                             # the Sequential aliases of update.py:29-34
        own["gru.conv_%s%s.0.weight" % (n, n)] = own["gru.conv_%s.weight" % n]
        own["gru.conv_%s%s.0.bias" % (n, n)] = own["gru.conv_%s.bias" % n]
    step.load_state_dict(own, strict=True)
    return step.cuda()


def test_update_step_golden():
    """Two chained reference DisparityUpdateDLCWithMask steps (golden, CPU reference run) through the drop-in."""
    import stereoformer_b200.modules as M
    g = load_golden("dlc_update_step")
    step = _load_step(g)
    old = M.DEFAULT_DIV_MODE
    M.DEFAULT_DIV_MODE = 0                                                # the golden is a CPU run (true division)
    try:
        with torch.no_grad():
            d1, h1, m1 = step(dev(g["cv"]), dev(g["cur"]), dev(g["left"]), dev(g["right"]), dev(g["hidden0"]), True)
            d2, h2, m2 = step(dev(g["cv"]), d1, dev(g["left"]), dev(g["right"]), h1, False)
    finally:
        M.DEFAULT_DIV_MODE = old
    assert torch.allclose(h1.cpu(), g["hidden1"], rtol=1e-4, atol=1e-5)
    assert torch.allclose(m1.cpu()[:, :18], g["mask1_head"], rtol=1e-4, atol=1e-4)
    assert (d1.cpu() - g["disp1"]).abs().max() < 1e-3                     # <= 1e-3 px (north_star)
    assert torch.allclose(h2.cpu(), g["hidden2"], rtol=1e-4, atol=2e-5)
    e2 = (d2.cpu() - g["disp2"]).abs()
    assert e2.median() < 1e-4 and (e2 < 1e-3).float().mean() > 0.99


def test_state_dict_keys_match_reference():
    """Parameter names are the reference's, so its checkpoints load with strict=True."""
    import stereoformer_b200.update as U
    g = load_golden("dlc_update_step")
    keys = set(U.DisparityUpdateDLCWithMask(24, 4, 20).state_dict().keys())
    ref = {k[3:] for k in g if k.startswith("sd_")}
    assert ref <= keys
    extra = {k for k in keys - ref if "num_batches" not in k}
    assert extra == {"gru.conv_%s%s.0.%s" % (n, n, p) for n in "zbg" for p in ("weight", "bias")}


@pytest.mark.parametrize("iters", [1, 4])
def test_refiner_graph_equals_eager(iters):
    """The graph-captured loop of baseline_dynamic_gru.py:128-139 replays to the eager loop's bits, twice, and both
    follow the oracle's step-by-step loop."""
    import stereoformer_b200.update as U
    g = load_golden("dlc_update_step")
    step = _load_step(g)
    torch.manual_seed(91)
    cv, cur = dev(g["cv"]), dev(g["cur"])
    left, right = torch.randn(2, 3, 48, 320, device="cuda"), torch.randn(2, 3, 48, 320, device="cuda")   # 8x the grid
    eager = U.DlcGruRefiner(step, use_graph=False).refine(cv, cur, left, right, iters)
    ref = U.DlcGruRefiner(step, use_graph=True)
    for rep in range(2):
        outs = ref.refine(cv, cur, left, right, iters)
        assert len(outs) == iters and outs[0].shape == (2, 1, 48, 320)
        for a, b in zip(outs, eager):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-5)
    # new inputs through the same captured graph
    cv2 = cv.flip(0).contiguous()
    outs2 = [o.clone() for o in ref.refine(cv2, cur, left, right, iters)]
    eager2 = U.DlcGruRefiner(step, use_graph=False).refine(cv2, cur, left, right, iters)
    for a, b in zip(outs2, eager2):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-5)
    # oracle loop (CPU, reciprocal division like the GPU default)
    sd = {k: v.cpu() for k, v in step.state_dict().items()}
    import torch.nn.functional as F
    lr = F.interpolate(left.cpu(), size=[6, 40], mode="bilinear", align_corners=False)
    rr = F.interpolate(right.cpu(), size=[6, 40], mode="bilinear", align_corners=False)
    d, h = g["cur"], None
    for it in range(iters):
        d, h, m = O.dlc_update_step(sd, g["cv"], d, lr, rr, h, 20, True, recip=True)
        full = O.upsample_convex(d, m, 8)
        err = (eager[it].cpu() - full).abs()
        assert err.median() < 1e-3 and (err < 2e-2).float().mean() > 0.98, (it, err.max())
