"""Dev tool: CUDA-graph timing of the DCN forward (SURVEY 8a a10 shapes), exact FFMA kernel vs tcgen05 path."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
shapes = [(8, 16, 16, 68, 120, 1, 1), (1, 128, 128, 248, 360, 1, 1), (1, 256, 256, 248, 360, 1, 1), (2, 64, 128, 128, 128, 1, 2)]
for B, Cin, Cout, H, W, G, dg in shapes:
    torch.manual_seed(0)
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin // G, 3, 3, device="cuda") * 0.05
    b = torch.randn(Cout, device="cuda")
    off = torch.randn(B, dg * 18, H, W, device="cuda") * 2
    msk = torch.sigmoid(torch.randn(B, dg * 9, H, W, device="cuda"))
    for algo, name in ((ops.SF_ALGO_GENERIC, "ffma"), (ops.SF_ALGO_TC, "tcgen05")):
        if algo == ops.SF_ALGO_TC and Cout // G < 8:
            continue
        fn = lambda: ops.dcn_forward(x, w, b, off, msk, (1, 1), (1, 1), (1, 1), G, dg, algo=algo)
        try:
            fn()
        except ValueError as e:
            print(name, "unsupported:", e); continue
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                for _ in range(3):
                    fn()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 15 * 1e3
        M = H * W * B
        fl = 2.0 * M * (Cin // G) * 9 * Cout
        by = (Cin + 27 * dg + Cout) * M * 4
        print("dcn %-8s B=%d %d->%d %dx%d G=%d dg=%d  %8.1f us  %6.1f TFLOP/s  %6.0f GB/s (algorithmic)" %
              (name, B, Cin, Cout, H, W, G, dg, us, fl / us / 1e6, by / us / 1e3))
    # backward (f3): data (input / offset / mask) and weight gradients, event-timed eagerly (ms-scale kernels)
    go = torch.randn(B, Cout, H, W, device="cuda")
    for needs, name in (((True, True, True, False, False), "bwd data"), ((False, False, False, True, True), "bwd weight")):
        fn = lambda: ops.dcn_backward(x, w, off, msk, go, (1, 1), (1, 1), (1, 1), G, dg, needs=needs)
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            fn()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 3 * 1e3
        print("dcn %-10s B=%d %d->%d %dx%d G=%d dg=%d  %8.1f us  %6.1f TFLOP/s" %
              (name, B, Cin, Cout, H, W, G, dg, us, 2.0 * H * W * B * (Cin // G) * 9 * Cout / us / 1e6))
