"""Dev tool: CUDA-event timing of the epipolar band attention (BASELINE config 4 shape by default)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, C, H, W, band = [int(x) for x in (sys.argv[1:6] if len(sys.argv) > 5 else (8, 128, 96, 312, 48))]
algo = int(sys.argv[6]) if len(sys.argv) > 6 else 0
sets = [[torch.randn(B, C, H, W, device="cuda").bfloat16() for _ in range(3)] for _ in range(3)]
for i in range(3):
    ops.epipolar_attention(*sets[i % 3], band, algo=algo)
torch.cuda.synchronize()
# one CUDA graph of n launches over rotating inputs: the eager ctypes path is host-bound at these kernel times
n = 30
g = torch.cuda.CUDAGraph()
st = torch.cuda.Stream()
with torch.cuda.stream(st):
    outs = [torch.empty_like(sets[0][0]) for _ in range(3)]
    with torch.cuda.graph(g, stream=st):
        for i in range(n):
            ops.epipolar_attention(*sets[i % 3], band, algo=algo, out=outs[i % 3])
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
us = e0.elapsed_time(e1) / n * 1e3
gb = 4 * C * H * W * 2 * B / (us * 1e-6) / 1e9
fl = 4.0 * H * W * band * C * B / (us * 1e-6) / 1e12
print("attn algo=%d shape=%s band=%d us=%.1f  %.0f GB/s (%.3f of HBM peak)  band-TFLOP/s=%.1f" %
      (algo, (B, C, H, W), band, us, gb, gb / 6549.8, fl))
