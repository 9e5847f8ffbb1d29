"""Dev tool: CUDA-event timing of the correlation cost-volume kernel (rotating inputs > L2).
Launches are replayed from a CUDA graph so that host launch overhead does not hide the kernel time."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, C, H, W, D = [int(x) for x in (sys.argv[1:6] if len(sys.argv) > 5 else (8, 256, 68, 120, 24))]
algo = int(os.environ.get("SF_CV_ALGO", "0"))
sets = [(torch.randn(B, C, H, W, device="cuda"), torch.randn(B, C, H, W, device="cuda")) for _ in range(3)]
for i in range(5):
    ops.corr_volume(*sets[i % 3], D, algo=algo)
torch.cuda.synchronize()
NL = 12
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    ops.corr_volume(*sets[0], D, algo=algo)
    torch.cuda.synchronize()
    with torch.cuda.graph(g, stream=s):
        outs = [ops.corr_volume(*sets[i % 3], D, algo=algo) for i in range(NL)]
g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 20
e0.record()
for _ in range(reps):
    g.replay()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / (reps * NL) * 1e3
# eager (host-launched) for comparison
e0.record()
for i in range(100):
    ops.corr_volume(*sets[i % 3], D, algo=algo)
e1.record()
torch.cuda.synchronize()
us_eager = e0.elapsed_time(e1) / 100 * 1e3
gb = (2 * C + D) * H * W * 4 * B / (us * 1e-6) / 1e9
print("variant=%s algo=%d shape=%s graph us=%.2f GB/s=%.0f frac=%.3f   (eager us=%.2f)" %
      (os.environ.get("SF_CV_VARIANT", "0"), algo, (B, C, H, W, D), us, gb, gb / 6549.8, us_eager))
