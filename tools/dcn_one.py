import sys, os, torch
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
from stereoformer_b200 import ops
B, Cin, Cout, H, W = [int(v) for v in sys.argv[1:6]]
torch.manual_seed(0)
x = torch.randn(B, Cin, H, W, device="cuda"); w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.1; b = torch.randn(Cout, device="cuda")
off = torch.randn(B, 18, H, W, device="cuda") * 2; m = torch.sigmoid(torch.randn(B, 9, H, W, device="cuda"))
y = ops.dcn_forward(x, w, b, off, m, (1, 1), (1, 1), (1, 1), 1, 1)
torch.cuda.synchronize()
print("ok", float(y.abs().mean()))
