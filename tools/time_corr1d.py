"""Dev tool: CUDA-event timing of the CorrBlock1D pyramid build (BASELINE config 3 shape by default)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, C, H, W = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (4, 256, 136, 240))]
algo = int(sys.argv[5]) if len(sys.argv) > 5 else 0
sets = [[torch.randn(B, C, H, W, device="cuda") for _ in range(2)] for _ in range(2)]
for i in range(3):
    ops.corr1d_pyramid(*sets[i % 2], 4, algo=algo)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 30
e0.record()
for i in range(n):
    ops.corr1d_pyramid(*sets[i % 2], 4, algo=algo)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / n * 1e3
by = (2 * C * H * W + 1.9375 * H * W * W) * 4 * B
print("corr1d algo=%d shape=%s us=%.1f  %.0f GB/s (%.3f of HBM peak)  %.1f TFLOP/s (fp32-equivalent)" %
      (algo, (B, C, H, W), us, by / (us * 1e-6) / 1e9, by / (us * 1e-6) / 1e9 / 6549.8, 2.0 * H * W * W * C * B / (us * 1e-6) / 1e12))
