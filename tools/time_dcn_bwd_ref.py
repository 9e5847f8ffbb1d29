"""Dev tool: DCN backward, ours vs the reference's own extension (oracle/_ref/native/DCN.so) on the same GPU."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import refenv
refenv.activate(dcn="reference")
import DCN as refdcn
import stereoformer_b200.DCN as ours

SH = {"dcn_16": (8, 16, 16, 68, 120), "dcn_128": (1, 128, 128, 248, 360), "dcn_256": (1, 256, 256, 248, 360),
      "dcn_128x8": (8, 128, 128, 68, 120)}


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for name in [a for a in sys.argv[1:] if a in SH] or ["dcn_16", "dcn_128"]:
    B, Cin, Cout, H, W = SH[name]
    torch.manual_seed(0)
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin, 3, 3, device="cuda") * 0.05
    b = torch.randn(Cout, device="cuda")
    off = torch.randn(B, 18, H, W, device="cuda") * 2
    msk = torch.sigmoid(torch.randn(B, 9, H, W, device="cuda"))
    go = torch.randn(B, Cout, H, W, device="cuda")
    args = (3, 3, 1, 1, 1, 1, 1, 1, 1, 1, 64)
    r = {"shape": name}
    r["ours_us"] = timeit(lambda: ours.modulated_deform_conv_backward(x, w, b, off, msk, go, *args))
    r["reference_us"] = timeit(lambda: refdcn.modulated_deform_conv_backward(x, w, b, off, msk, go, *args))
    a = ours.modulated_deform_conv_backward(x, w, b, off, msk, go, *args)
    c = refdcn.modulated_deform_conv_backward(x, w, b, off, msk, go, *args)
    r["max_rel_err"] = max(float((p - q).abs().max() / q.abs().max().clamp_min(1e-20)) for p, q in zip(a, c))
    print(json.dumps(r), flush=True)
