"""Dev tool: the cost-volume backward (tcgen05 band product vs the fp32 FFMA kernel) at the BASELINE shapes.
    python tools/time_cv_bwd.py            # CUDA-graph timing, rotating inputs, JSON lines
    python tools/time_cv_bwd.py once       # a few plain launches at the bench shape (for ncu)"""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops

PEAK = 6544.7
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
SHAPES = [("c2_x8", 8, 256, 68, 120, 24), ("c5", 1, 256, 248, 360, 48), ("c3_4x136x240", 4, 256, 136, 240, 48),
          ("c2_x8_C128", 8, 128, 68, 120, 24), ("c1", 1, 256, 32, 64, 24),
          ("c5_D24", 1, 256, 248, 360, 24), ("c2_D48", 8, 256, 68, 120, 48), ("c2_W128", 8, 256, 68, 128, 24)]
if len(sys.argv) > 2 and sys.argv[1] == "tc":        # tc D1,D2,...: the bench grid with other disparity counts
    SHAPES = [("c2_D%s" % d, 8, 256, 68, 120, int(d)) for d in sys.argv[2].split(",")]
if len(sys.argv) > 1 and sys.argv[1] == "tc":        # tensor-core kernel only (quick A/B runs)
    ONLY_TC = True
else:
    ONLY_TC = False


def make(B, C, H, W, D, n):
    g = torch.Generator(device="cuda").manual_seed(3)
    return [(torch.randn(B, C, H, W, device="cuda", generator=g), torch.randn(B, C, H, W, device="cuda", generator=g),
             torch.randn(B, D, H, W, device="cuda", generator=g)) for _ in range(n)]


def graph_us(fn, inner, min_ms=40.0):
    for i in range(inner):
        fn(i)
    torch.cuda.synchronize()
    g, s = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for i in range(inner):
                fn(i)
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    n = max(3, int(min_ms / max(e0.elapsed_time(e1), 1e-3)) + 1)
    e0.record()
    for _ in range(n):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (n * inner) * 1e3


if len(sys.argv) > 1 and sys.argv[1] == "once":
    _, B, C, H, W, D = SHAPES[int(sys.argv[2]) if len(sys.argv) > 2 else 0]
    sets = make(B, C, H, W, D, 3)
    for i in range(6):
        ops.corr_volume_backward(*sets[i % 3], D, algo=ops.SF_ALGO_TC)
    torch.cuda.synchronize()
    sys.exit(0)

for name, B, C, H, W, D in SHAPES:
    sets = make(B, C, H, W, D, 3)
    by = (4 * C + D) * B * H * W * 4
    row = {"shape": name, "dims": [B, C, H, W, D], "alg_bytes": by}
    for tag, algo in (("tcgen05", ops.SF_ALGO_TC), ("ffma", ops.SF_ALGO_GENERIC))[:1 if ONLY_TC else 2]:
        us = graph_us(lambda i: ops.corr_volume_backward(*sets[i % 3], D, algo=algo), 6, 40.0 if tag == "tcgen05" else 15.0)
        row[tag + "_us"] = round(us, 2)
        row[tag + "_frac"] = round(by / us / 1e3 / PEAK, 3)
    if ONLY_TC:
        print(json.dumps(row), flush=True)
        continue
    a = ops.corr_volume_backward(*sets[0], D, algo=ops.SF_ALGO_TC)
    b = ops.corr_volume_backward(*sets[0], D, algo=ops.SF_ALGO_GENERIC)
    row["max_rel_diff"] = max(float((x - y).abs().max() / y.abs().max()) for x, y in zip(a, b))
    print(json.dumps(row), flush=True)
