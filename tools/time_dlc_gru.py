"""Dev tool: the custom-kernel part of the 12-iteration DLC-GRU refinement loop (BASELINE configs[2],
baseline_dynamic_gru.py:128-139) captured in ONE CUDA graph: per iteration guidance (soft-max -> uncertainty
volume + warp + error), window sampling with per-pixel bounds, convex x8 upsampling.  The cuDNN convolutions /
ConvGRU between them are outside the hot-path scope; their outputs (bounds, upsampling mask) are synthetic."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, D, H, W, ITERS = 8, 24, 68, 120, 12
torch.manual_seed(0)
cv = (torch.randn(B, D, H, W, device="cuda") * 3 - 2).abs()
left, right = torch.randn(B, 3, H, W, device="cuda"), torch.randn(B, 3, H, W, device="cuda")
lo = torch.relu(torch.randn(B, 1, H, W, device="cuda")) * 3
up = torch.relu(torch.randn(B, 1, H, W, device="cuda")) * 3
masks = [torch.randn(B, 576, H, W, device="cuda") for _ in range(2)]
flag = torch.zeros(1, dtype=torch.int32, device="cuda")


def loop():
    disp = ops.soft_argmin(cv).unsqueeze(1)
    outs = []
    for it in range(ITERS):
        err, unc = ops.guidance(cv, disp, left, right, flag=flag)
        disp = ops.local_window(cv, disp, lo, up, mode=ops.SF_WIN_OFFSETS, n_points=20, consider_valid=True)
        outs.append(ops.upsample_convex(disp, masks[it % 2], 8))
    return outs


loop(); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        outs = loop()
g.replay(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    g.replay()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
e0.record()
for _ in range(3):
    loop()
e1.record(); torch.cuda.synchronize()
ms_eager = e0.elapsed_time(e1) / 3
px = B * H * W
by = ITERS * ((2 * D + 10) + (D + 4) + (576 + 1 + 64)) * px * 4
print("DLC-GRU custom kernels, %d iterations, B=%d: graph %.1f us (%.1f us/iter, %.0f GB/s algorithmic), eager %.1f us" %
      (ITERS, B, ms * 1e3, ms * 1e3 / ITERS, by / ms / 1e6, ms_eager * 1e3))
