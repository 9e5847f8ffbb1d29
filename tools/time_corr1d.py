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
# lookup per GRU iteration: volume form (a9) vs the on-the-fly alternative (PytorchAlternateCorrBlock1D)
f1, f2 = sets[0]
coords = torch.cat((torch.arange(float(W), device="cuda").view(1, 1, 1, W).repeat(B, 1, H, 1) -
                    torch.rand(B, 1, H, W, device="cuda") * W / 4,
                    torch.arange(float(H), device="cuda").view(1, 1, H, 1).repeat(B, 1, 1, W)), 1)
pyr = ops.corr1d_pyramid(f1, f2, 4, algo=algo)
for name, fn in (("lookup (volume form)", lambda: ops.corr1d_lookup(pyr, coords, 4, 4)),
                 ("lookup (alternative, on the fly)", lambda: ops.corr1d_alt(f1, f2, coords, 4, 4))):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print("corr1d %-34s shape=%s us=%.1f" % (name, (B, C, H, W), e0.elapsed_time(e1) / 5 * 1e3))
