"""Dev tool: one DCN forward for ncu captures: tcgen05 at 128->128 @ 248x360 (default) or, with the argument `small`,
the shipped 16->16 shape at 8 x 68x120."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
small = len(sys.argv) > 1 and sys.argv[1] == "small"
B, Cin, Cout, H, W = (8, 16, 16, 68, 120) if small else (1, 128, 128, 248, 360)
torch.manual_seed(0)
x = torch.randn(B, Cin, H, W, device="cuda")
w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05
b = torch.randn(Cout, device="cuda")
off = torch.randn(B, 18, H, W, device="cuda") * 2
msk = torch.sigmoid(torch.randn(B, 9, H, W, device="cuda"))
for _ in range(3):
    ops.dcn_forward(x, w, b, off, msk, (1, 1), (1, 1), (1, 1), 1, 1, algo=ops.SF_ALGO_GENERIC if small else ops.SF_ALGO_TC)
torch.cuda.synchronize()
