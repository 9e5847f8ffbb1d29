"""Dev tool: a few DCN backward calls at 128 -> 128 @ 248x360 (for ncu launch lists / captures)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, Cin, Cout, H, W = (1, 256, 256, 248, 360) if "256" in sys.argv else (1, 128, 128, 248, 360)
torch.manual_seed(0)
x = torch.randn(B, Cin, H, W, device="cuda")
w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05
off = torch.randn(B, 18, H, W, device="cuda") * 2
msk = torch.sigmoid(torch.randn(B, 9, H, W, device="cuda"))
go = torch.randn(B, Cout, H, W, device="cuda")
for _ in range(3):
    ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), 1, 1)
torch.cuda.synchronize()
