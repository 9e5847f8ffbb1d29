"""Dev tool: CUDA-graph timing of the tensor-core DCN forward alone (pre-pass + weight split + main kernel).
    python tools/time_dcn2.py [shape ...] [scale=2.0]     shapes: dcn_128 dcn_256 dcn_64 dcn_128x8"""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
SH = {"dcn_128": (1, 128, 128, 248, 360), "dcn_256": (1, 256, 256, 248, 360), "dcn_64": (2, 64, 64, 136, 240),
      "dcn_128x8": (8, 128, 128, 68, 120), "dcn_128_256": (1, 128, 256, 248, 360), "dcn_16": (8, 16, 16, 68, 120),
      "dcn_32": (8, 32, 32, 68, 120), "dcn_16_c5": (1, 16, 16, 248, 360)}
ALGO = ops.SF_ALGO_TC if "algo=tc" in sys.argv else (ops.SF_ALGO_GENERIC if "algo=exact" in sys.argv else None)
names = [a for a in sys.argv[1:] if a in SH] or ["dcn_128", "dcn_256"]
scale = [float(a.split("=")[1]) for a in sys.argv[1:] if a.startswith("scale=")] or [2.0]
for name in names:
    B, Cin, Cout, H, W = SH[name]
    torch.manual_seed(0)
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05
    b = torch.randn(Cout, device="cuda")
    off = torch.randn(B, 18, H, W, device="cuda") * scale[0]
    msk = torch.sigmoid(torch.randn(B, 9, H, W, device="cuda"))
    fn = lambda: ops.dcn_forward(x, w, b, off, msk, (1, 1), (1, 1), (1, 1), 1, 1, algo=ALGO)   # noqa: E731
    ref = ops.dcn_forward(x, w, b, off, msk, (1, 1), (1, 1), (1, 1), 1, 1, algo=ops.SF_ALGO_GENERIC)
    err = float((fn() - ref).abs().max() / ref.abs().max())
    fn(); torch.cuda.synchronize()
    g, s = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(3):
                fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 30 * 1e3
    fl = 2.0 * B * H * W * Cin * 9 * Cout
    print(json.dumps({"shape": name, "offset_scale": scale[0], "us": round(us, 1), "tflops": round(fl / us / 1e6, 1), "rel_err_vs_exact": err}), flush=True)
