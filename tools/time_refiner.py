"""Dev tool: the full 12-iteration DLC-GRU refinement (BASELINE configs[2]; baseline_dynamic_gru.py:128-139) at
544x960, D=24 @ 1/8: DisparityUpdateDLCWithMask(24, 32, 20) with random-init weights.  Three arms:
  torch    -- the reference's arithmetic as eager PyTorch ops (restated here, cuDNN convs)
  eager    -- stereoformer_b200 drop-ins launched eagerly
  graph    -- DlcGruRefiner: loop invariants hoisted, whole loop replayed from one CUDA graph"""
import os, sys, torch
import torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stereoformer_b200 as sf
from stereoformer_b200 import update as U

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
D, H, W, ITERS = 24, 68, 120, 12
torch.manual_seed(0)
torch.backends.cudnn.benchmark = True
step = U.DisparityUpdateDLCWithMask(24, 32, 20).cuda().eval()
cv = (torch.randn(B, D, H, W, device="cuda") * 3 - 2).abs()
disp0 = torch.rand(B, 1, H, W, device="cuda") * (D - 1)
left, right = torch.randn(B, 3, 8 * H, 8 * W, device="cuda"), torch.randn(B, 3, 8 * H, 8 * W, device="cuda")


def torch_step(cv, cur, left, right, hidden):
    """Attempts/update.py:263-317 + :71-90 + :36-49 as PyTorch ops (what the reference launches)."""
    prob = F.softmax(cv, 1)
    cand = torch.arange(D, device=cv.device).view(1, D, 1, 1).repeat(1, 1, H, W).type_as(cur)
    l = F.interpolate(left, size=[H, W], mode="bilinear", align_corners=False)
    r = F.interpolate(right, size=[H, W], mode="bilinear", align_corners=False)
    xs = torch.arange(W, device=cv.device, dtype=cur.dtype).view(1, 1, W).expand(B, H, W)
    ys = torch.arange(H, device=cv.device, dtype=cur.dtype).view(1, H, 1).expand(B, H, W)
    grid = torch.stack((2 * (xs - cur[:, 0]) / (W - 1) - 1, 2 * ys / (H - 1) - 1), -1)
    warped = F.grid_sample(r, grid, mode="bilinear", padding_mode="border", align_corners=False)
    err = warped - l
    unc = prob * cand * (cand - cur) * (cand - cur)
    feat = torch.cat([step.encoder.disparity_error_encoder(err), step.encoder.uncertain_encoder(unc)], 1)
    g = step.gru
    h0 = torch.zeros_like(feat) if hidden is None else hidden
    x1 = torch.cat((feat, h0), 1)
    z, b = g.conv_zz(x1), g.conv_bb(x1)
    gg = g.conv_gg(torch.cat((b * h0, feat), 1))
    hidden = (1 - z) * h0 + z * gg
    mask = .25 * step.mask(hidden)
    lo, up = torch.chunk(step.offset(hidden), 2, dim=1)
    lo, up = cur - lo, cur + up
    inv = torch.clamp((lo < 0).float() + (up >= D - 1).float(), max=1.0)
    c = lo + torch.arange(21, device=cv.device).type_as(cv).view(1, 21, 1, 1) * ((up - lo) * 1.0 / 20)
    c = c * (1 - inv) + inv * cur.repeat(1, 21, 1, 1)
    ce, fl = torch.clamp(torch.ceil(c), 0, D - 1), torch.clamp(torch.floor(c), 0, D - 1)
    fr = ce - c
    local = torch.gather(cv, 1, ce.long()) * (1 - fr) + torch.gather(cv, 1, fl.long()) * fr
    d = torch.sum(torch.softmax(local, 1) * c, 1).unsqueeze(1)
    m = torch.softmax(mask.view(B, 1, 9, 8, 8, H, W), 2)
    u = F.unfold(8 * d, [3, 3], padding=1).view(B, 1, 9, 1, 1, H, W)
    full = torch.sum(m * u, 2).permute(0, 1, 4, 2, 5, 3).reshape(B, 1, 8 * H, 8 * W)
    return d, hidden, full


@torch.no_grad()
def torch_loop():
    d, h, outs = disp0, None, []
    for _ in range(ITERS):
        d, h, full = torch_step(cv, d, left, right, h)
        outs.append(full)
    return outs


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


eager = U.DlcGruRefiner(step, use_graph=False)
graph = U.DlcGruRefiner(step, use_graph=True)
a = torch_loop()
b = eager.refine(cv, disp0, left, right, ITERS)
c = graph.refine(cv, disp0, left, right, ITERS)
err = [(x - y).abs() for x, y in zip(a, b)]
print("B=%d: drop-in vs torch restatement, final disparity: median |diff| %.2e px, 99%% quantile %.2e px" %
      (B, err[-1].median().item(), err[-1].flatten()[::7].quantile(0.99).item()))
print("graph vs eager max |diff| %.2e" % max((x - y).abs().max().item() for x, y in zip(b, c)))
t_t, t_e = timeit(torch_loop), timeit(lambda: eager.refine(cv, disp0, left, right, ITERS))
t_g = timeit(lambda: graph.refine(cv, disp0, left, right, ITERS))
print("12 iterations, B=%d: torch eager %.2f ms | drop-ins eager %.2f ms | graph replay %.2f ms  (%.0f pairs/s; %.1fx over torch)" %
      (B, t_t, t_e, t_g, B / t_g * 1e3, t_t / t_g))
