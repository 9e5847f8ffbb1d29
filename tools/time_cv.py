"""Dev tool: CUDA-event timing of the correlation cost-volume kernel (rotating inputs > L2)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, C, H, W, D = [int(x) for x in (sys.argv[1:6] if len(sys.argv) > 5 else (8, 256, 68, 120, 24))]
sets = [(torch.randn(B, C, H, W, device="cuda"), torch.randn(B, C, H, W, device="cuda")) for _ in range(3)]
for i in range(5):
    ops.corr_volume(*sets[i % 3], D)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 200
e0.record()
for i in range(n):
    ops.corr_volume(*sets[i % 3], D)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / n * 1e3
gb = (2 * C + D) * H * W * 4 * B / (us * 1e-6) / 1e9
print("variant=%s shape=%s us=%.2f GB/s=%.0f frac=%.3f" % (os.environ.get("SF_CV_VARIANT", "0"), (B, C, H, W, D), us, gb, gb / 6463))
