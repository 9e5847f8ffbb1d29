"""Dev tool: odd shapes through the tensor-core kernels of round 2 (run under `timeout`): parity against the FFMA kernels."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereoformer_b200 import ops
torch.manual_seed(0)
bad = 0
for (B, C, H, W, D, G) in [(1, 64, 1, 4, 1, 1), (1, 32, 2, 8, 3, 1), (3, 160, 2, 132, 17, 1), (1, 448, 1, 260, 65, 1), (2, 64, 3, 1028, 9, 2),
                           (1, 96, 1, 128, 33, 3), (1, 128, 5, 124, 24, 1)]:
    L, R = torch.randn(B, C, H, W, device="cuda"), torch.randn(B, C, H, W, device="cuda")
    Gv = torch.randn(B * G * D * H * W, device="cuda")
    a = ops.corr_volume_backward(L, R, Gv, D, G, algo=ops.SF_ALGO_TC)
    b = ops.corr_volume_backward(L, R, Gv, D, G, algo=ops.SF_ALGO_GENERIC)
    e = max(float((x - y).abs().max() / y.abs().max().clamp_min(1e-20)) for x, y in zip(a, b))
    print("cv_bwd", (B, C, H, W, D, G), "%.2e" % e, flush=True); bad += e > 1e-4
for (B, Cin, Cout, H, W, G, dg, s, p, d) in [(1, 64, 8, 5, 4, 1, 1, 1, 1, 1), (1, 64, 48, 9, 12, 1, 1, 1, 1, 1), (3, 32, 40, 7, 8, 1, 2, 1, 1, 1),
                                             (1, 320, 72, 8, 20, 1, 5, 1, 1, 1), (1, 256, 512, 6, 8, 2, 4, 1, 1, 1), (2, 64, 64, 40, 4, 1, 1, 1, 0, 1),
                                             (1, 128, 24, 30, 50, 1, 1, 3, 2, 2), (1, 16, 16, 9, 8, 1, 1, 1, 1, 1), (1, 32, 16, 11, 12, 2, 1, 1, 1, 1)]:
    x = torch.randn(B, Cin, H, W, device="cuda")
    w = torch.randn(Cout, Cin // G, 3, 3, device="cuda") * 0.1
    bb = torch.randn(Cout, device="cuda")
    Ho, Wo = (H + 2 * p - (d * 2 + 1)) // s + 1, (W + 2 * p - (d * 2 + 1)) // s + 1
    off = torch.randn(B, dg * 18, Ho, Wo, device="cuda") * 2
    msk = torch.sigmoid(torch.randn(B, dg * 9, Ho, Wo, device="cuda"))
    go = torch.randn(B, Cout, Ho, Wo, device="cuda")
    f1 = ops.dcn_forward(x, w, bb, off, msk, (s, s), (p, p), (d, d), G, dg)
    f0 = ops.dcn_forward(x, w, bb, off, msk, (s, s), (p, p), (d, d), G, dg, algo=ops.SF_ALGO_GENERIC)
    e = float((f1 - f0).abs().max() / f0.abs().max())
    g1 = ops.dcn_backward(x, w, off, msk, go, (s, s), (p, p), (d, d), G, dg)
    g0 = ops.dcn_backward(x, w, off, msk, go, (s, s), (p, p), (d, d), G, dg, algo=ops.SF_ALGO_GENERIC)
    eb = max(float((u - v).abs().max() / v.abs().max().clamp_min(1e-20)) for u, v in zip(g1, g0))
    print("dcn", (B, Cin, Cout, H, W, G, dg, s, p, d), "fwd %.2e bwd %.2e" % (e, eb), flush=True); bad += (e > 1e-4) + (eb > 2e-4)
# band attention: tcgen05 kernel (P in tensor memory for C <= 128) against the FFMA kernel; many tiles per CTA, ragged last tiles,
# band 1..64, repeated calls on the same buffers (stale P / S in tensor memory must never leak into a later tile)
for (B, C, H, W, band) in [(1, 64, 1, 8, 1), (1, 128, 3, 136, 64), (2, 64, 700, 24, 5), (1, 128, 450, 392, 33), (3, 128, 7, 1000, 48),
                           (1, 64, 300, 520, 64), (1, 192, 5, 264, 17), (1, 256, 150, 312, 48), (2, 128, 96, 312, 2)]:
    q, k, v = (torch.randn(B, C, H, W, device="cuda").bfloat16() for _ in range(3))
    ref = ops.epipolar_attention(q, k, v, band, algo=ops.SF_ALGO_GENERIC).float()
    e = 0.0
    for rep in range(3):
        out = ops.epipolar_attention(q, k, v, band, algo=ops.SF_ALGO_TMA).float()
        e = max(e, float((out - ref).abs().max()))
    print("attn", (B, C, H, W, band), "%.2e" % e, flush=True); bad += e > 3e-2
torch.cuda.synchronize()
print("BAD" if bad else "all ok")
