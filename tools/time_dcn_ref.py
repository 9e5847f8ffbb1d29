"""DCN forward (a10): ours vs the comparators on the same box (VERDICT r1 missing #3).

  reference  : the reference's own extension (oracle/_ref/native/DCN.so: modulated_deformable_im2col kernel +
               at::addmm per group + permute/contiguous, deform/src/cuda/modulated_deform_conv_cuda.cu:90-118)
  torchvision: torchvision.ops.deform_conv2d (CUDA)
  ours       : sf_dcn_forward (AUTO dispatch), plus the exact FFMA kernel

CUDA events over eager launches (the comparators allocate, so no graph), best of 3 blocks; prints one JSON line per
shape.  Also reports max |ours - reference| / max |reference|.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from stereoformer_b200 import ops  # noqa: E402

SHAPES = [("dcn_16", 8, 16, 16, 68, 120, 1, 1), ("dcn_128", 1, 128, 128, 248, 360, 1, 1),
          ("dcn_256", 1, 256, 256, 248, 360, 1, 1), ("dcn_64_128_dg2", 2, 64, 128, 128, 128, 1, 2)]


def timeit(fn, reps):
    fn()
    torch.cuda.synchronize()
    best = float("inf")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return best * 1e3


def graph_time(fn, reps=5, inner=3):
    fn()
    torch.cuda.synchronize()
    g, s = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(inner):
                fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * inner) * 1e3


def run(shapes=SHAPES, quiet=False):
    from oracle import refenv              # comparator only
    refenv.activate(dcn="reference")
    import DCN as refdcn
    has_ref = hasattr(refdcn, "modulated_deform_conv_forward") and not getattr(refdcn, "__sf_stub__", False)
    try:
        from torchvision.ops import deform_conv2d
    except Exception:
        deform_conv2d = None
    out = {}
    for name, B, Cin, Cout, H, W, G, dg in shapes:
        torch.manual_seed(0)
        x = torch.randn(B, Cin, H, W, device="cuda")
        w = torch.randn(Cout, Cin // G, 3, 3, device="cuda") * 0.05
        b = torch.randn(Cout, device="cuda")
        off = torch.randn(B, dg * 18, H, W, device="cuda") * 2
        msk = torch.sigmoid(torch.randn(B, dg * 9, H, W, device="cuda"))
        M = B * H * W
        flops = 2.0 * M * (Cin // G) * 9 * Cout
        nbytes = (Cin + 27 * dg + Cout) * M * 4
        res = {"shape": [B, Cin, Cout, H, W, G, dg], "flops": flops, "alg_bytes": nbytes}
        ours = lambda: ops.dcn_forward(x, w, b, off, msk, (1, 1), (1, 1), (1, 1), G, dg)   # noqa: E731
        res["ours_us"] = graph_time(ours)
        res["ours_eager_us"] = timeit(ours, 10)
        res["ours_exact_us"] = graph_time(lambda: ops.dcn_forward(x, w, b, off, msk, (1, 1), (1, 1), (1, 1), G, dg,
                                                                  algo=ops.SF_ALGO_GENERIC))
        got = ours()
        if has_ref:
            ref = lambda: refdcn.modulated_deform_conv_forward(x, w, b, off, msk, 3, 3, 1, 1, 1, 1, 1, 1, G, dg, 64)  # noqa: E731
            res["reference_us"] = timeit(ref, 10)
            want = ref()
            res["max_rel_err_vs_reference"] = float((got - want).abs().max() / want.abs().max())
        if deform_conv2d is not None and G == 1:
            tv = lambda: deform_conv2d(x, off, w, b, stride=1, padding=1, dilation=1, mask=msk)   # noqa: E731
            res["torchvision_us"] = timeit(tv, 10)
        # backward (all five gradients): ours vs the reference extension, eager CUDA-event timing on both sides
        go = torch.randn_like(got)
        bwd = lambda: ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), G, dg)   # noqa: E731
        res["ours_backward_us"] = timeit(bwd, 5)
        if has_ref:
            rb = lambda: refdcn.modulated_deform_conv_backward(x, w, b, off, msk, go, 3, 3, 1, 1, 1, 1, 1, 1, G, dg, 64)  # noqa: E731
            res["reference_backward_us"] = timeit(rb, 3)
            a, c = bwd(), rb()            # (grad_input, grad_offset, grad_mask, grad_weight, grad_bias) on both sides
            res["backward_max_rel_err_vs_reference"] = max(
                float((u - v).abs().max() / v.abs().max().clamp_min(1e-20)) for u, v in zip(a, c))
        res["tflops"] = flops / res["ours_us"] / 1e6
        res["gbs"] = nbytes / res["ours_us"] / 1e3
        out[name] = res
        if not quiet:
            print(json.dumps({name: res}), flush=True)
    return out


if __name__ == "__main__":
    run()
