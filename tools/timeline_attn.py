"""Dev tool: per-role clock64 timeline of CTA 0 of the tcgen05 band-attention kernel (SF_ATTN_TS hook)."""
import os; os.environ.setdefault("SF_TUNING_LIB", "1")   # needs the -DSF_TUNING build (python -m stereoformer_b200.build --tuning)
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
B, C, H, W, band = [int(x) for x in (sys.argv[1:6] if len(sys.argv) > 5 else (8, 128, 96, 312, 48))]
q, k, v = [torch.randn(B, C, H, W, device="cuda").bfloat16() for _ in range(3)]
ops.epipolar_attention(q, k, v, band)
ts = torch.zeros(4, 128, dtype=torch.int64, device="cuda")
os.environ["SF_ATTN_TS"] = "%x" % ts.data_ptr()
ops.epipolar_attention(q, k, v, band)
torch.cuda.synchronize()
t = ts.cpu()
t0 = int(t[t > 0].min())
names = ["tma: per tile NC x (Q/K stage free) | V buffer free", "mma: NC x QK chunk issued / PV issued (interleaved)",
         "epilogue w11: o_full | tile written", "softmax w2: s_full | max exchanged + P free | p_full arrive"]
per = [None, None, 2, 3]
for r in range(4):
    vv = [int(x) - t0 for x in t[r] if x > 0]
    print(names[r], len(vv))
    print("  ", vv[:60])
    if per[r]:
        n = per[r]
        for i in range(4 * n, min(len(vv) - n, 10 * n), n):
            print("   tile: start +%d, marks" % (vv[i] - vv[i - n]), [vv[i + k + 1] - vv[i + k] for k in range(n - 1)],
                  "to next tile start", vv[i + n] - vv[i + n - 1])
