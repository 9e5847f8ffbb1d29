"""Dev tool: per-role clock64 timeline of CTA 0 of the tcgen05 cost-volume kernel (SF_CV_TS hook)."""
import os; os.environ.setdefault("SF_TUNING_LIB", "1")   # needs the -DSF_TUNING build (python -m stereoformer_b200.build --tuning)
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, C, H, W, D = 8, 256, 68, 120, 24
L, R = torch.randn(B, C, H, W, device="cuda"), torch.randn(B, C, H, W, device="cuda")
ops.corr_volume(L, R, D)
ts = torch.zeros(4, 128, dtype=torch.int64, device="cuda")
os.environ["SF_CV_TS"] = "%x" % ts.data_ptr()
ops.corr_volume(L, R, D)
torch.cuda.synchronize()
t = ts.cpu()
t0 = int(t[t > 0].min())
names = ["tma issue (after empty wait)", "converter w0: full | op_empty | (issue loads) | stores issued | fenced", "mma: op_full", "epilogue: acc_full | bar | ld0 | ld1 | sts done | bar2"]
for r in range(4):
    v = [int(x) - t0 for x in t[r] if x > 0]
    print(names[r], len(v))
    print("  ", v)
