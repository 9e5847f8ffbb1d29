"""Dev tool: CUDA-graph timing of the small hot-path kernels at the bench shape (cost volume resident in L2)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, D, H, W = 8, 24, 68, 120
torch.manual_seed(0)
cv = (torch.randn(B, D, H, W, device="cuda") * 3 - 2).abs()
disp = torch.rand(B, 1, H, W, device="cuda") * (D - 1)
img = torch.randn(B, 3, H, W, device="cuda")
mask = torch.randn(B, 576, H, W, device="cuda")


def timeit(name, fn, nbytes, nl=10, reps=20):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(nl):
                fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / (reps * nl) * 1e3
    print("%-34s %7.2f us   %6.0f GB/s (algorithmic)" % (name, us, nbytes / (us * 1e-6) / 1e9))


px = B * H * W
timeit("local_window fused soft-argmin", lambda: ops.local_window(cv, None, mode=ops.SF_WIN_FIXED, radius=2, n_points=20,
                                                                   consider_valid=True), (D + 1) * px * 4)
timeit("local_window (disp given)", lambda: ops.local_window(cv, disp, mode=ops.SF_WIN_FIXED, radius=2, n_points=20,
                                                              consider_valid=True), (D + 2) * px * 4)
timeit("local_window variance", lambda: ops.local_window(cv, disp, mode=ops.SF_WIN_VARIANCE, gamma=1.0, n_points=20,
                                                          consider_valid=False), (D + 2) * px * 4)
timeit("soft_argmin", lambda: ops.soft_argmin(cv), (D + 1) * px * 4)
timeit("disp_warp", lambda: ops.disp_warp(img, disp, check=False), (3 * 3 + 1) * px * 4)
timeit("upsample_convex8", lambda: ops.upsample_convex(disp, mask, 8), (576 + 1 + 64) * px * 4)
timeit("upsample_simple8", lambda: ops.upsample_simple(disp, 8), (1 + 64) * px * 4)

# ---- f2 / f3 kernels at the same shape (C2: B=8, 68x120, D=24, features 256 ch, hidden 64)
C = 64
zb = torch.randn(B, 2 * C, H, W, device="cuda")
gp, h0, xin = (torch.randn(B, C, H, W, device="cuda") for _ in range(3))
timeit("gru_reset", lambda: ops.gru_reset(zb[:, C:], h0, xin), 5 * C * px * 4)
timeit("gru_blend", lambda: ops.gru_blend(zb[:, :C], gp, h0), 4 * C * px * 4)
pre = torch.randn(B, 2, H, W, device="cuda") * 3
timeit("local_window head-fused (DLC)", lambda: ops.local_window(cv, disp, lo=pre, mode=ops.SF_WIN_HEAD, n_points=20,
                                                                  consider_valid=True), (D + 4) * px * 4)
left = torch.randn(B, 3, H, W, device="cuda")
timeit("guidance", lambda: ops.guidance(cv, disp, left, img), (2 * D + 10) * px * 4)
fl, fr = torch.randn(B, 256, H, W, device="cuda"), torch.randn(B, 256, H, W, device="cuda")
gcv, gd = torch.randn(B, D, H, W, device="cuda"), torch.randn(B, 1, H, W, device="cuda")
timeit("corr_volume backward", lambda: ops.corr_volume_backward(fl, fr, gcv, D), (4 * 256 + D) * px * 4, nl=3, reps=5)
timeit("soft_argmin backward", lambda: ops.soft_argmin_backward(cv, gd), (2 * D + 1) * px * 4)
timeit("local_window backward (fixed)", lambda: ops.local_window_backward(cv, disp, None, None, gd, ops.SF_WIN_FIXED, 2.0,
                                                                           20, True), (2 * D + 3) * px * 4)
timeit("local_window backward (head)", lambda: ops.local_window_backward(cv, disp, pre, None, gd, ops.SF_WIN_HEAD, 0.0, 20,
                                                                          True), (2 * D + 7) * px * 4)
timeit("disp_warp backward", lambda: ops.disp_warp_backward(img, disp, img), (3 * 3 * 2 + 2) * px * 4)
gfull = torch.randn(B, 1, 8 * H, 8 * W, device="cuda")
timeit("upsample_simple8 backward", lambda: ops.upsample_simple_backward(gfull, 8), (1 + 64) * px * 4)
timeit("upsample_convex8 backward", lambda: ops.upsample_convex_backward(disp, mask, gfull, 8), (2 * 576 + 2 + 64) * px * 4)
