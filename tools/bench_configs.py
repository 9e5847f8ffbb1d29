"""The other BASELINE.json configs, the DCN shapes and the full model as entries of bench.py's JSON line
(`configs`: VERDICT r1 next-round #2).  Every entry is measured live: CUDA events on the launching stream around
CUDA-graph replays (or eager launches where the comparator allocates), after warm-up; `parity_max_err` compares the
kernel's output with the oracle's PyTorch-fp32 restatement evaluated on the same device tensors (checker only) or with
the unmodified reference where one exists for the entry.

entry = {"us": per call, "alg_bytes" | "flops": per call (SURVEY 8d formulas), "bound", "frac", "peak_source",
         "parity_max_err", ...}
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            return {"hbm_gbs": float(j["hbm_gbs"]), "bf16_tflops": float(j.get("bf16_tflops_sustained", j["bf16_tflops"])),
                    "source": "measured (MEASURED_PEAKS.json: hbm_gbs, bf16_tflops_sustained)"}
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1500.0, "source": "fallback (B200_PROFILING.md)"}


def graph_us(fn, inner=6, reps=None, min_ms=30.0):
    """us per call of fn(), replayed from a CUDA graph of `inner` calls until >= min_ms have been timed."""
    fn()
    torch.cuda.synchronize()
    g, s = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        s.synchronize()
        with torch.cuda.graph(g, stream=s):
            for i in range(inner):
                fn(i) if fn.__code__.co_argcount else fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    one = max(e0.elapsed_time(e1), 1e-3)
    n = reps or max(3, int(min_ms / one) + 1)
    e0.record()
    for _ in range(n):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * inner) * 1e3


def eager_us(fn, min_ms=30.0):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    n = max(3, int(min_ms / max(e0.elapsed_time(e1), 1e-3)) + 1)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def _hbm(entry, pk):
    entry["bound"] = "hbm"
    entry["achieved_gbs"] = entry["alg_bytes"] / entry["us"] / 1e3
    entry["frac"] = entry["achieved_gbs"] / pk["hbm_gbs"]
    entry["peak_source"] = pk["source"]
    return entry


def _tensor(entry, pk):
    entry["bound"] = "tensor"
    entry["achieved_tflops"] = entry["flops"] / entry["us"] / 1e6
    entry["frac"] = entry["achieved_tflops"] / pk["bf16_tflops"]
    entry["peak_source"] = pk["source"]
    return entry


def _relerr(got, want):
    """max |got - want| / max(|want|): the cost-volume tolerance of north_star (1e-4 relative to the magnitude)."""
    return float((got - want).abs().max() / want.abs().max().clamp_min(1e-30))


# ------------------------------------------------------------------------------------------------ C1
def c1_lcv_small(dev, pk):
    """BASELINE configs[0]: LocalCostVolume cost-volume + soft-argmin forward on one 1x3x256x512 pair (32x64 grid)."""
    from stereoformer_b200 import ops
    from oracle import stereo_oracle as O
    B, C, H, W, D = 1, 256, 32, 64, 24
    g = torch.Generator(device="cpu").manual_seed(7)
    sets = [(torch.randn(B, C, H, W, generator=g).to(dev), torch.randn(B, C, H, W, generator=g).to(dev))
            for _ in range(3)]
    cvb = torch.empty(B, D, H, W, device=dev)
    ob = torch.empty(B, 1, H, W, device=dev)

    def step(i=0):
        cv = ops.corr_volume(*sets[i % 3], D, out=cvb)
        return ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20, consider_valid=True,
                                div_mode=ops.SF_DIV_RECIP, out=ob)
    us = graph_us(step)
    eag = eager_us(step)
    got = step(0).clone()
    cv_ref = O.corr_volume(*sets[0], D)
    want = O.local_cost_volume(cv_ref, O.soft_argmin(cv_ref).unsqueeze(1), 2, 20, True, recip=True)
    e = {"shape": [B, C, H, W, D], "us": us, "eager_api_us": eag, "kernels": 2,
         "alg_bytes": ((2 * C + D) + (D + 2)) * H * W * 4 * B,
         "parity_max_err": float((got - want).abs().max()), "parity_epe": float((got - want).abs().mean()),
         "parity_unit": "px (1/8 scale) vs oracle chain", "note": "launch-latency bound at this size (4.6 MB)"}
    return _hbm(e, pk)


# ------------------------------------------------------------------------------------------------ C2 backward (f3)
def c2_cv_backward(dev, pk):
    """SURVEY 8f-3 at the bench shape: gradient of the correlation cost volume w.r.t. both feature maps (what autograd
    derives for build_cost_volume.py:39-47), tcgen05 band product vs the fp32 FFMA kernel; parity against autograd
    through the UNMODIFIED reference module on the same GPU (one pair) when the reference copy travelled."""
    from stereoformer_b200 import ops
    B, C, H, W, D = 8, 256, 68, 120, 24
    g = torch.Generator(device="cpu").manual_seed(9)
    sets = [tuple(torch.randn(*shp, generator=g).to(dev) for shp in ((B, C, H, W), (B, C, H, W), (B, D, H, W)))
            for _ in range(3)]
    us = graph_us(lambda i=0: ops.corr_volume_backward(*sets[i % 3], D, algo=ops.SF_ALGO_TC))
    us_ffma = graph_us(lambda i=0: ops.corr_volume_backward(*sets[i % 3], D, algo=ops.SF_ALGO_GENERIC), min_ms=15.0)
    gl, gr = ops.corr_volume_backward(*sets[0], D, algo=ops.SF_ALGO_TC)
    el, er = ops.corr_volume_backward(*sets[0], D, algo=ops.SF_ALGO_GENERIC)
    e = {"shape": [B, C, H, W, D], "us": us, "exact_ffma_us": us_ffma, "alg_bytes": (4 * C + D) * B * H * W * 4,
         "parity_max_err": max(_relerr(gl, el), _relerr(gr, er)),
         "parity_unit": "relative to max |grad| vs the fp32 FFMA kernel (all 8 pairs)"}
    try:
        from oracle import refenv
        if refenv.activate(dcn=None) is not None:
            import importlib
            cvm = importlib.import_module("models.utils.build_cost_volume").CostVolume(D, "correlation")
            l, r = sets[0][0][:1].clone().requires_grad_(True), sets[0][1][:1].clone().requires_grad_(True)
            rl, rr = torch.autograd.grad(cvm(l, r), [l, r], sets[0][2][:1])
            e["parity_vs_reference_autograd"] = max(_relerr(gl[:1], rl), _relerr(gr[:1], rr))
            e["reference_fwd_plus_bwd_us_8_pairs_eager"] = 8 * eager_us(
                lambda: torch.autograd.grad(cvm(l, r), [l, r], sets[0][2][:1]), min_ms=10.0)
    except Exception as ex:          # noqa: BLE001
        e["reference_error"] = str(ex)[:200]
    return _hbm(e, pk)


# ------------------------------------------------------------------------------------------------ C3 (RAFT-Stereo)
def c3_corr1d(dev, pk):
    """BASELINE configs[2], CorrBlock1D reading: all-pairs pyramid at (4,256,136,240) + one radius-4 lookup."""
    from stereoformer_b200 import ops
    from oracle import stereo_oracle as O
    B, C, H, W = 4, 256, 136, 240
    g = torch.Generator(device="cpu").manual_seed(8)
    sets = [[torch.randn(B, C, H, W, generator=g).to(dev) for _ in range(2)] for _ in range(2)]
    us_pyr = eager_us(lambda: ops.corr1d_pyramid(*sets[0], 4))
    us_pyr = min(us_pyr, eager_us(lambda: ops.corr1d_pyramid(*sets[1], 4)))
    f1, f2 = sets[0]
    coords = torch.cat((torch.arange(float(W), device=dev).view(1, 1, 1, W).repeat(B, 1, H, 1) -
                        torch.rand(B, 1, H, W, device=dev) * W / 4,
                        torch.arange(float(H), device=dev).view(1, 1, H, 1).repeat(B, 1, 1, W)), 1)
    pyr = ops.corr1d_pyramid(f1, f2, 4)
    us_look = graph_us(lambda: ops.corr1d_lookup(pyr, coords, 4, 4, ops.SF_DIV_RECIP))
    # parity on one pair (the oracle's einsum volume of 4 pairs is 125 MB x levels)
    want_pyr = O.corr1d_pyramid(f1[:1], f2[:1], 4)
    got_pyr = ops.corr1d_pyramid(f1[:1].contiguous(), f2[:1].contiguous(), 4)
    perr = max(_relerr(a, b) for a, b in zip(got_pyr, want_pyr))
    want_l = O.corr1d_lookup(want_pyr, coords[:1], 4, 4, recip=True)
    got_l = ops.corr1d_lookup(got_pyr, coords[:1].contiguous(), 4, 4, ops.SF_DIV_RECIP)
    by = (2 * C * H * W + 1.9375 * H * W * W) * 4 * B
    e = {"shape": [B, C, H, W], "us": us_pyr, "alg_bytes": by, "flops": 2.0 * H * W * W * C * B,
         "lookup_us": us_look, "lookup_alg_bytes": (36 + 2 + 40) * H * W * 4 * B,
         "parity_max_err": perr, "lookup_parity_max_err": _relerr(got_l, want_l),
         "parity_unit": "relative to max |volume| vs oracle (pair 0)"}
    return _hbm(e, pk)


# ------------------------------------------------------------------------------------------------ C3 (DLC-GRU)
def c3_dlc_gru(dev, pk):
    """BASELINE configs[2], DLC-GRU reading: 12 refinement iterations at 544x960 (B = 8), the loop of
    baseline_dynamic_gru.py:128-139 around DisparityUpdateDLCWithMask.  Comparator: the UNMODIFIED reference module
    (oracle/_ref) with the same weights, eager on the same GPU."""
    import torch.nn.functional as F
    from stereoformer_b200 import update as U
    B, D, H, W, ITERS = 8, 24, 68, 120, 12
    torch.manual_seed(0)
    step = U.DisparityUpdateDLCWithMask(24, 32, 20).to(dev).eval()
    cv = (torch.randn(B, D, H, W, device=dev) * 3 - 2).abs()
    disp0 = torch.rand(B, 1, H, W, device=dev) * (D - 1)
    left, right = torch.randn(B, 3, 8 * H, 8 * W, device=dev), torch.randn(B, 3, 8 * H, 8 * W, device=dev)
    graph = U.DlcGruRefiner(step, use_graph=True)
    eager = U.DlcGruRefiner(step, use_graph=False)
    e = {"shape": [B, D, H, W], "iters": ITERS}
    e["us"] = eager_us(lambda: graph.refine(cv, disp0, left, right, ITERS))
    e["eager_api_us"] = eager_us(lambda: eager.refine(cv, disp0, left, right, ITERS))
    got = graph.refine(cv, disp0, left, right, ITERS)[-1].clone()
    try:
        from tools import fullmodel
        fullmodel.activate_reference()
        import importlib
        ref_upd = importlib.import_module("models.LocalCostVolume.Attempts.update")
        ref_up8 = importlib.import_module("models.utils.upsample").upsample_convex8
        ref = ref_upd.DisparityUpdateDLCWithMask(24, 32, 20).to(dev).eval()
        ref.load_state_dict(step.state_dict(), strict=True)

        @torch.no_grad()
        def ref_loop():
            d, h, full = disp0, None, None
            for _ in range(ITERS):
                d, h, m = ref(cv, d, left, right, h, True)
                full = ref_up8(d, m)
            return full
        e["gpu_eager_reference_us"] = eager_us(ref_loop)
        want = ref_loop()
        err = (got - want).abs()
        e["parity_epe"] = float(err.mean())
        e["parity_max_err"] = float(err.max())
        e["parity_unit"] = "px (full resolution, after 12 iterations) vs the unmodified reference module on the GPU"
        e["speedup_vs_gpu_eager_reference"] = e["gpu_eager_reference_us"] / e["us"]
    except Exception as ex:          # noqa: BLE001 -- the comparator is optional (no copy of the reference on the box)
        e["gpu_eager_reference_us"] = None
        e["parity_max_err"] = None
        e["comparator_error"] = "%s: %s" % (type(ex).__name__, ex)
    e["pairs_per_s"] = B / e["us"] * 1e6
    e["bound"] = "cuDNN convolutions between the custom kernels (out of scope); custom kernels: see DESIGN 3"
    e["frac"] = None
    e["peak_source"] = pk["source"]
    return e


# ------------------------------------------------------------------------------------------------ C4
def c4_attn(dev, pk):
    """BASELINE configs[3]: epipolar band attention at 1/4 of 384x1248 (96x312), C = 128, band 48, bf16."""
    from stereoformer_b200 import ops
    from oracle import stereo_oracle as O
    B, C, H, W, band = 8, 128, 96, 312, 48
    g = torch.Generator(device="cpu").manual_seed(9)
    sets = [[torch.randn(B, C, H, W, generator=g).to(dev).bfloat16() for _ in range(3)] for _ in range(3)]
    outs = [torch.empty_like(sets[0][0]) for _ in range(3)]
    us = graph_us(lambda i=0: ops.epipolar_attention(*sets[i % 3], band, out=outs[i % 3]), inner=6)
    got = ops.epipolar_attention(*[t[:1].contiguous() for t in sets[0]], band).float()
    want = O.epipolar_attention(*[t[:1].float() for t in sets[0]], band)
    if isinstance(want, tuple):
        want = want[0]
    e = {"shape": [B, C, H, W, band], "us": us, "alg_bytes": 4 * C * H * W * 2 * B,
         "flops": 4.0 * H * W * band * C * B, "parity_max_err": _relerr(got, want.float()),
         "parity_unit": "relative to max |O| vs the SPEC-DERIVED oracle (reference file is 0 bytes: parity unpinned)"}
    e = _hbm(e, pk)
    e["band_tflops"] = e["flops"] / us / 1e6
    return e


# ------------------------------------------------------------------------------------------------ DCN
def dcn_shapes(dev, pk):
    from tools import time_dcn_ref
    res = time_dcn_ref.run(shapes=time_dcn_ref.SHAPES[:3], quiet=True)
    out = {}
    for name, r in res.items():
        e = {"shape": r["shape"], "us": r["ours_us"], "eager_api_us": r["ours_eager_us"], "exact_ffma_us": r["ours_exact_us"],
             "alg_bytes": r["alg_bytes"], "flops": r["flops"],
             "reference_extension_us": r.get("reference_us"), "torchvision_us": r.get("torchvision_us"),
             "parity_max_err": r.get("max_rel_err_vs_reference"),
             "parity_unit": "relative to max |out| vs the reference's own compiled extension (oracle/_ref/native/DCN.so)",
             "backward_us": r.get("ours_backward_us"), "reference_extension_backward_us": r.get("reference_backward_us"),
             "backward_parity_max_err": r.get("backward_max_rel_err_vs_reference")}
        out[name] = _hbm(e, pk) if name == "dcn_16" else _tensor(e, pk)
        if e["reference_extension_us"]:
            e["speedup_vs_reference_extension"] = e["reference_extension_us"] / e["us"]
    return out


# ------------------------------------------------------------------------------------------------ full model
def full_model(dev, pk, variant="lc", batch=8, height=544, width=960):
    """BASELINE configs[1] read literally: the reference's LowCNN (unmodified, oracle/_ref) on synthetic 544x960 pairs,
    stock (PyTorch eager + reference DCN extension) vs the same object with the drop-ins installed."""
    from tools import fullmodel
    from tools.synth import stereo_pair
    old = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        net = fullmodel.build_net(variant, dev)
        left, right, _ = stereo_pair(batch, height, width, device=dev)
        import stereoformer_b200 as sf
        ref = fullmodel.forward(net, variant, left, right)
        ms_ref = fullmodel.time_forward(net, variant, left, right)
        with fullmodel.dropins():
            got = fullmodel.forward(net, variant, left, right)
            ms_ours = fullmodel.time_forward(net, variant, left, right)
            with sf.precision("exact"):
                err_exact = (fullmodel.forward(net, variant, left, right) - ref).abs()
        err = (got - ref).abs()
        floor = (fullmodel.forward(net, variant, torch.nextafter(left, left + 1), right) - ref).abs()
        return {"variant": variant, "shape": [batch, 3, height, width], "us": ms_ours * 1e3,
                "parity_epe_exact_precision": float(err_exact.mean()), "parity_max_exact_precision": float(err_exact.max()),
                "stock_vs_stock_1ulp_input_epe": float(floor.mean()),
                "pairs_per_s": batch / ms_ours * 1e3, "gpu_eager_reference_us": ms_ref * 1e3,
                "gpu_eager_reference_pairs_per_s": batch / ms_ref * 1e3, "speedup_vs_gpu_eager_reference": ms_ref / ms_ours,
                "parity_epe": float(err.mean()), "parity_max_err": float(err.max()),
                "parity_unit": "px (full resolution) vs the stock reference network on the same GPU",
                "bound": "cuDNN trunk (out of scope, SURVEY 2 rows 8/11/12); the path's kernels are the other entries",
                "frac": None, "peak_source": pk["source"],
                "launch": "eager (the reference's forward, incl. its host syncs baseline.py:119)", "tf32": False}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


# ------------------------------------------------------------------------------------------------ C5 (row bands)
def c5_bands(dev, pk, world, rank, root_only=True, steps=40):
    """BASELINE configs[4]: ONE 1984x2880 pair (248x360 grid, C = 256, D = 48) split into `world` row bands; per pair:
    cost volume + fused window, guidance (1-row halo), the shipped 16->16 DCN (data-dependent halo), convex x8
    upsampling whose store also delivers the rows to rank 0 (NVLink peer stores) -- strong scaling."""
    import torch.distributed as dist
    from stereoformer_b200 import bands as sfb
    from stereoformer_b200 import dist as sfd
    from stereoformer_b200 import ops
    H, W, C, D = 248, 360, 256, 48
    g = torch.Generator(device=dev).manual_seed(1024)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)    # noqa: E731
    full = dict(left_f=r(1, C, H, W), right_f=r(1, C, H, W), left_i=r(1, 3, H, W), right_i=r(1, 3, H, W),
                x=r(1, 16, H, W), wt=r(16, 16, 3, 3) * 0.1, bias=r(16), off=r(1, 18, H, W) * 2,
                msk=torch.sigmoid(r(1, 9, H, W)), up_mask=r(1, 576, H, W))
    halo = sfb.dcn_halo_rows(full["off"], 3, 1, 1)
    band = sfb.RowBand.of_rank(H, world, rank, halo=halo)
    bands = sfd.row_bands(H, world, halo=halo)
    mine = {k: (band.interior(v) if k in ("left_f", "right_f", "left_i") else
                v if k in ("wt", "bias") else band.source(v)) for k, v in full.items()}
    b1 = sfb.RowBand(H, band.y0, band.y1, max(band.y0 - 1, 0), min(band.y1 + 1, H))
    lf1, rf1, um1 = b1.source(full["left_f"]), b1.source(full["right_f"]), b1.source(full["up_mask"])
    pg, gather_mode = None, "none (1 GPU)"
    if world > 1:
        try:
            pg = sfd.PeerGather(-(-(H * 8 * W * 8) // world), dev)
        except Exception as ex:       # noqa: BLE001
            pg, gather_mode = None, "unavailable (%s)" % type(ex).__name__
        flags = torch.tensor([1.0 if pg is not None else 0.0], device=dev)
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        if float(flags.item()) < 1.0:
            pg = None
        else:
            gather_mode = "P2P stores fused into the upsampling kernel, to %s" % ("rank 0" if root_only else "every rank")

    def band_pass():
        cv1, disp1 = sfb.lcv_band(lf1, rf1, D)
        cv, disp = b1.crop(cv1).contiguous(), b1.crop(disp1).contiguous()
        err, unc = sfb.guidance_band(cv, disp, mine["left_i"], mine["right_i"], band)
        y = sfb.dcn_band(mine["x"], mine["wt"], mine["bias"], mine["off"], mine["msk"], band, halo=halo)
        up = sfb.upsample_convex_band(disp1, um1, b1, gather=pg, root=0 if root_only else None)
        return disp, err, unc, y, up

    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(3):
            out = band_pass()
        side.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            out = band_pass()
    torch.cuda.synchronize()
    disp, err, unc, y, up = out
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    if pg is not None:
        pg.sync()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        y_full = sfd.gather_row_bands(y, bands)
        disp_full = sfd.gather_row_bands(disp, bands)
    else:
        y_full, disp_full = y, disp
    res = None
    if rank == 0:
        whole = sfb.RowBand(H, 0, H)
        cv, d_ref = sfb.lcv_band(full["left_f"], full["right_f"], D)
        y_ref = sfb.dcn_band(full["x"], full["wt"], full["bias"], full["off"], full["msk"], whole, halo=halo)
        up_ref = ops.upsample_convex(d_ref, full["up_mask"])
        up_full = up if world == 1 else pg.buffer[:H * 8 * W * 8].view(1, 1, H * 8, W * 8) if pg is not None else None
        us = float(ms.item()) * 1e3
        res = {"shape": [1, C, H, W, D], "n_bands": world, "us": us, "pairs_per_s": 1e6 / us, "scaling": "strong",
               "dcn_halo_rows": halo, "gather": gather_mode, "launch": "one CUDA graph per band pass (5 kernels)",
               "disparity_bit_exact_vs_whole_image": bool(torch.equal(disp_full, d_ref)),
               "upsampled_bit_exact_vs_whole_image": (bool(torch.equal(up_full, up_ref)) if up_full is not None else None),
               "parity_max_err": float((y_full - y_ref).abs().max()),
               "parity_unit": "max |band DCN - whole-image DCN| (the other outputs are bit-exact flags)",
               "alg_bytes": ((2 * C + D) + (D + 2) + (2 * D + 10) + (16 + 27 + 16) + (9 * 64 + 1 + 64)) * H * W * 4}
        _hbm(res, pk)
    if world > 1:
        dist.barrier()
    return res


def run_single_gpu(dev, want=None):
    """All one-GPU entries; a failing entry is reported, not fatal (the headline must still print)."""
    pk = peaks()
    table = [("c1", c1_lcv_small), ("c2_cv_backward", c2_cv_backward), ("c3_corr1d", c3_corr1d), ("c3_dlc_gru", c3_dlc_gru), ("c4_attn", c4_attn),
             ("dcn", dcn_shapes), ("full_model_lc", lambda d, p: full_model(d, p, "lc")),
             ("full_model_dlc_gru", lambda d, p: full_model(d, p, "dlc_gru", batch=2))]
    out = {}
    for name, fn in table:
        if want and name not in want:
            continue
        try:
            r = fn(dev, pk)
            if name == "dcn":
                out.update(r)
            else:
                out[name] = r
        except Exception as ex:          # noqa: BLE001
            out[name] = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300])}
        torch.cuda.synchronize()
    return out


if __name__ == "__main__":
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    res = run_single_gpu(dev, set(sys.argv[1:]) or None)
    res["c5_bands"] = c5_bands(dev, peaks(), 1, 0)
    print(json.dumps(res, indent=1))
