"""CPU oracle for the StereoFormer hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT.

Every function restates one reference function in PyTorch fp32 on the CPU and
cites the reference file:line it follows (paths relative to the reference repo
root).  The oracle is pinned: ``oracle/gen_golden.py`` ran the UNMODIFIED
reference code (imported from /root/reference in the build container) on seeded
inputs and stored its outputs under ``tests/golden/``; ``tests/test_oracle_golden.py``
checks every function here against those vectors (bit-exact for indices,
candidates and sampling coordinates).  The one exception is ``epipolar_attention``:
the reference file is empty (0 bytes), so that function is SPEC-DERIVED and its
parity is UNPINNED.

Third-party arithmetic the reference calls (F.grid_sample, F.softmax, torch.gather,
F.avg_pool2d, einsum) is PyTorch's; the oracle calls the same PyTorch ops where the
reference does and restates everything around them.

``scalar_div`` semantics: the reference divides tensors by Python scalars
(``/ N``, ``/ (w - 1)``).  PyTorch's CPU kernels do a true IEEE division, its CUDA
kernels multiply by the reciprocal of the scalar.  ``recip=False`` reproduces the
CPU behaviour (what the golden vectors were made with), ``recip=True`` the CUDA one.
"""
import math

import torch
import torch.nn.functional as F


def _sdiv(t, s, recip=False):
    """tensor / python-scalar with CPU (true division) or CUDA (x * (1/s)) rounding."""
    if recip:
        return t * (torch.tensor(1.0, dtype=t.dtype) / torch.tensor(float(s), dtype=t.dtype)).to(t.device)
    return t / float(s)


# --------------------------------------------------------------------------- a1 / a2
def corr_volume(left, right, max_disp):
    """models/utils/build_cost_volume.py:39-47 ('correlation').

    cv[b,d,y,x] = mean_c L[b,c,y,x] * R[b,c,y,x-d] for x >= d, 0 elsewhere.
    """
    b, c, h, w = left.shape
    cv = left.new_zeros(b, max_disp, h, w)
    for d in range(min(max_disp, w)):
        cv[:, d, :, d:] = (left[..., d:] * right[..., : w - d]).mean(dim=1)
    return cv


def gwc_volume(left, right, max_disp, groups):
    """models/CrossAttentionCostVolume/GWC_Cost_Volume.py:5-24 -> [B,G,D,H,W]."""
    b, c, h, w = left.shape
    assert c % groups == 0
    vol = left.new_zeros(b, groups, max_disp, h, w)
    for d in range(min(max_disp, w)):
        prod = left[..., d:] * right[..., : w - d]
        vol[:, :, d, :, d:] = prod.view(b, groups, c // groups, h, w - d).mean(dim=2)
    return vol


def diff_volume(left, right, max_disp):
    """models/utils/build_cost_volume.py:21-27 ('difference') -> [B,C,D,H,W]."""
    b, c, h, w = left.shape
    vol = left.new_zeros(b, c, max_disp, h, w)
    for d in range(min(max_disp, w)):
        vol[:, :, d, :, d:] = left[..., d:] - right[..., : w - d]
    return vol


def concat_volume(left, right, max_disp):
    """models/utils/build_cost_volume.py:29-37 ('concat') -> [B,2C,D,H,W]."""
    b, c, h, w = left.shape
    vol = left.new_zeros(b, 2 * c, max_disp, h, w)
    for d in range(min(max_disp, w)):
        vol[:, :c, d, :, d:] = left[..., d:]
        vol[:, c:, d, :, d:] = right[..., : w - d]
    return vol


# --------------------------------------------------------------------------- a3
def soft_argmin(cv, match_similarity=True, with_prob=False):
    """models/utils/estimation.py:14-33 (and :44-62 with the probability volume)."""
    assert cv.dim() == 4
    prob = F.softmax(cv if match_similarity else -cv, dim=1)
    cand = torch.arange(cv.size(1), dtype=cv.dtype, device=cv.device).view(1, -1, 1, 1)
    disp = torch.sum(prob * cand, dim=1)
    return (disp, prob) if with_prob else disp


# --------------------------------------------------------------------------- a4 / a5 / a6 sampler
def window_candidates(lower, upper, cur_disp, D, n_points, consider_valid,
                      clamp_mode="fixed", extra_invalid=None, recip=False):
    """Candidate disparities of the local window.

    models/LocalCostVolume/Attempts/fixed_local_cost_volume.py:23-45 (clamp_mode 'fixed':
    lower>=0, upper in [0,D]); variance_local_cost_volume.py:44-66 (clamp_mode 'variance':
    both in [0,D-1], plus ``extra_invalid`` = upper > x); Attempts/update.py:284-304 (as 'fixed').
    Returns (candidates [B,N+1,H,W], invalid_mask or None).
    """
    inv = None
    if consider_valid:
        inv = (lower < 0).float() + (upper >= D - 1).float()
        if extra_invalid is not None:
            inv = inv + extra_invalid
        inv = torch.clamp(inv, max=1.0)
    elif clamp_mode == "fixed":
        lower = torch.clamp(lower, min=0)
        upper = torch.clamp(upper, min=0, max=D)
    else:
        lower = torch.clamp(lower, min=0, max=D - 1)
        upper = torch.clamp(upper, min=0, max=D - 1)
    step = _sdiv((upper - lower) * 1.0, n_points, recip)
    k = torch.arange(n_points + 1, dtype=lower.dtype, device=lower.device).view(1, -1, 1, 1)
    cand = lower + k * step
    if consider_valid:
        cand = cand * (1 - inv) + inv * cur_disp.repeat(1, n_points + 1, 1, 1)
    return cand, inv


def window_indices(cand, D):
    """fixed_local_cost_volume.py:5-9,65-69: straight-through ceil/floor, clamp, int64.

    ste_ceil(x) = ceil(x) - x + x is evaluated in fp32 and is NOT always ceil(x);
    the int64 cast then truncates.  Returns (idx_floor, idx_ceil).
    """
    c = torch.ceil(cand) - cand + cand
    f = torch.floor(cand) - cand + cand
    ic = torch.clamp(c, min=0, max=D - 1).long()
    jf = torch.clamp(f, min=0, max=D - 1).long()
    return jf, ic


def sample_volume(old_volume, cand):
    """build_cost_volume_from_volume, fixed_local_cost_volume.py:61-80 (same in update.py:326-345)."""
    D = old_volume.shape[1]
    jf, ic = window_indices(cand, D)
    floor_rate = ic - cand
    ceil_rate = 1.0 - floor_rate
    return torch.gather(old_volume, 1, jf) * floor_rate + torch.gather(old_volume, 1, ic) * ceil_rate


def window_expectation(old_volume, cand):
    """fixed_local_cost_volume.py:48-58: softmax over the samples, expectation of the candidates."""
    local = sample_volume(old_volume, cand)
    score = torch.softmax(local, dim=1)
    return torch.sum(score * cand, dim=1).unsqueeze(1)


def local_cost_volume(old_volume, cur_disp, radius, n_points, consider_valid=False, recip=False):
    """LocalCostVolume.forward, fixed_local_cost_volume.py:18-58."""
    D = old_volume.shape[1]
    cand, _ = window_candidates(cur_disp - radius, cur_disp + radius, cur_disp, D, n_points,
                                consider_valid, "fixed", recip=recip)
    return window_expectation(old_volume, cand)


def disparity_sigma(prob, cur_disp):
    """GetVarince, variance_local_cost_volume.py:12-25."""
    d = torch.arange(prob.size(1), dtype=prob.dtype, device=prob.device).view(1, -1, 1, 1)
    err = (d - cur_disp) * (d - cur_disp)
    return torch.sqrt(torch.sum(prob * err, dim=1)).unsqueeze(1)


def variance_local_cost_volume(old_volume, cur_disp, gamma, n_points, consider_valid=False, recip=False):
    """Variance_Based_Local_Cost_Volume.forward, variance_local_cost_volume.py:34-78."""
    B, D, H, W = old_volume.shape
    sigma = disparity_sigma(F.softmax(old_volume, dim=1), cur_disp)
    lower = cur_disp - gamma * sigma
    upper = cur_disp + gamma * sigma
    extra = None
    if consider_valid:
        xs = torch.arange(W, dtype=cur_disp.dtype, device=cur_disp.device).view(1, 1, 1, W)
        extra = (upper > xs).float()
    cand, _ = window_candidates(lower, upper, cur_disp, D, n_points, consider_valid,
                                "variance", extra, recip=recip)
    return window_expectation(old_volume, cand)


def dlc_window(old_volume, cur_disp, lo, up, n_points, consider_valid=True, recip=False):
    """Sampler tail of DisparityUpdateDLC(WithMask).forward, Attempts/update.py:279-313:
    bounds = cur_disp -/+ ReLU-conv outputs ``lo``/``up`` (both [B,1,H,W])."""
    D = old_volume.shape[1]
    cand, _ = window_candidates(cur_disp - lo, cur_disp + up, cur_disp, D, n_points,
                                consider_valid, "fixed", recip=recip)
    return window_expectation(old_volume, cand)


def uncertainty_volume(old_volume, cur_disp):
    """BasicGuidanceNet.forward, Attempts/update.py:71-72,86: softmax(cv) * d * (d-disp) * (d-disp)."""
    prob = F.softmax(old_volume, dim=1)
    d = torch.arange(old_volume.size(1), dtype=cur_disp.dtype, device=cur_disp.device).view(1, -1, 1, 1)
    return prob * d * (d - cur_disp) * (d - cur_disp)


# --------------------------------------------------------------------------- a7
def warp_grid(disp, H, W, recip=False):
    """utils/disparity_warper.py:60-80 (meshgrid), :96-98, :48-58 (normalize_coords).

    Returns the normalised sampling grid [B,H,W,2] (x, y) exactly as the reference hands it
    to F.grid_sample: gx = 2*((x - disp)/(W-1)) - 1, gy = 2*(y/(H-1)) - 1.
    """
    dt = disp.dtype
    xs = torch.arange(W, dtype=dt, device=disp.device).view(1, 1, W).expand(1, H, W)
    ys = torch.arange(H, dtype=dt, device=disp.device).view(1, H, 1).expand(1, H, W)
    sx = xs + (-disp[:, 0])
    sy = ys + torch.zeros_like(disp[:, 0])
    gx = 2 * _sdiv(sx, W - 1, recip) - 1
    gy = 2 * _sdiv(sy, H - 1, recip) - 1
    return torch.stack((gx, gy), dim=-1)


def warp_source_coords(grid, H, W, border=True):
    """ATen grid_sampler (align_corners=False): ix = ((g + 1) * size - 1) / 2, then the
    border clip to [0, size-1].  Third-party (PyTorch) arithmetic restated so that the
    kernel's pixel-space sampling coordinates can be compared bit-for-bit."""
    ix = ((grid[..., 0] + 1) * W - 1) / 2
    iy = ((grid[..., 1] + 1) * H - 1) / 2
    if border:
        ix = torch.clamp(ix, 0, W - 1)
        iy = torch.clamp(iy, 0, H - 1)
    return ix, iy


def disp_warp(img, disp, padding_mode="border", recip=False):
    """disp_warp, utils/disparity_warper.py:83-106 -> (warped, valid_mask)."""
    assert disp.min() >= 0
    H, W = img.shape[-2:]
    grid = warp_grid(disp, H, W, recip)
    warped = F.grid_sample(img, grid, mode="bilinear", padding_mode=padding_mode, align_corners=False)
    mask = F.grid_sample(torch.ones_like(img), grid, mode="bilinear", padding_mode="zeros",
                         align_corners=False)
    mask = torch.where(mask < 0.9999, torch.zeros_like(mask), mask)
    mask = torch.where(mask > 0, torch.ones_like(mask), mask)
    return warped, mask


def guidance(old_volume, cur_disp, left_img, right_img, recip=False):
    """Loop body of BasicGuidanceNet.forward up to the conv encoders, Attempts/update.py:68-86.
    Images are expected already at the disparity resolution (the F.interpolate of :74-77 is
    loop-invariant and stays a library call).  Returns (error_map, uncertainty_volume)."""
    warped, _ = disp_warp(right_img, cur_disp, recip=recip)
    return warped - left_img, uncertainty_volume(old_volume, cur_disp)


# --------------------------------------------------------------------------- a8 / a9
def corr1d_pyramid(f1, f2, num_levels=4):
    """CorrBlock1D.__init__/corr, models/IterativeLocalCostVolume/corr/corr.py:55-76,103-111.
    Returns num_levels+1 tensors [B*H*W1,1,1,W2/2^i] (the reference stores one more level
    than it reads)."""
    B, C, H, W1 = f1.shape
    W2 = f2.shape[-1]
    corr = torch.einsum("aijk,aijh->ajkh", f1, f2).reshape(B, H, W1, 1, W2)
    corr = corr / torch.sqrt(torch.tensor(C).float()).to(corr.device)
    corr = corr.reshape(B * H * W1, 1, 1, W2)
    pyr = [corr]
    for _ in range(num_levels):
        corr = F.avg_pool2d(corr, [1, 2], stride=[1, 2])
        pyr.append(corr)
    return pyr


def corr1d_lookup_coords(coords_x, level, radius, Wl, recip=False):
    """corr.py:88-90 + utils/utils.py:62-64: x0 = dx + x/2^i, normalised 2*x0/(Wl-1) - 1.
    coords_x: [N] flattened x coordinates.  Returns (x0 [N,2r+1], xgrid [N,2r+1])."""
    dx = torch.linspace(-radius, radius, 2 * radius + 1).view(1, -1).to(coords_x.device)
    x0 = dx + coords_x.view(-1, 1) / 2 ** level
    xg = _sdiv(2 * x0, Wl - 1, recip) - 1
    return x0, xg


def corr1d_lookup(pyramid, coords, radius=4, num_levels=4, recip=False):
    """CorrBlock1D.__call__, corr.py:78-101 with bilinear_sampler, utils/utils.py:59-76."""
    B, _, H, W = coords.shape
    cx = coords[:, :1].permute(0, 2, 3, 1).reshape(-1)
    out = []
    for i in range(num_levels):
        vol = pyramid[i]
        Wl = vol.shape[-1]
        _, xg = corr1d_lookup_coords(cx, i, radius, Wl, recip)
        grid = torch.stack((xg, torch.zeros_like(xg)), dim=-1).view(-1, 2 * radius + 1, 1, 2)
        s = F.grid_sample(vol, grid, align_corners=True)
        out.append(s.view(B, H, W, -1))
    return torch.cat(out, dim=-1).permute(0, 3, 1, 2).contiguous().float()


def corr1d_alt(f1, f2, coords, radius=4, num_levels=4, recip=False):
    """PytorchAlternateCorrBlock1D.__call__ / .corr, corr/corr.py:15-51: per level, the x-pooled right map is sampled at
    (x/2^i + dx, y) with grid_sample(align_corners=True) after the reference's own 2*x/(W-1) - 1 normalisation, dotted
    with the left map over channels and divided by sqrt(C)."""
    B, C, H, W1 = f1.shape
    c = coords.permute(0, 2, 3, 1)
    dx = torch.linspace(-radius, radius, 2 * radius + 1).to(coords.device)
    out = []
    fm2 = f2
    for i in range(num_levels):
        Wl = fm2.shape[3]
        x0 = c[..., 0:1] / 2 ** i + dx.view(1, 1, 1, -1)                   # centroid + delta (:42-44)
        y0 = c[..., 1:2].expand_as(x0)
        xg = _sdiv(2 * x0, Wl - 1, recip) - 1
        yg = _sdiv(2 * y0, H - 1, recip) - 1
        lv = []
        for t in range(2 * radius + 1):
            grid = torch.stack((xg[..., t], yg[..., t]), dim=-1)           # [B,H,W1,2]
            samp = F.grid_sample(fm2, grid, align_corners=True)
            lv.append(torch.sum(samp * f1, dim=1))
        out.append(torch.stack(lv, dim=1).permute(0, 2, 3, 1) / torch.sqrt(torch.tensor(C).float()))
        fm2 = F.avg_pool2d(fm2, [1, 2], stride=[1, 2])
    return torch.cat(out, dim=-1).permute(0, 3, 1, 2).contiguous().float()


# --------------------------------------------------------------------------- a10 (python form; C form in dcn_oracle.c)
def dcn_forward(inp, weight, bias, offset, mask, stride, pad, dil, groups, dgroups):
    """deform/src/cuda/modulated_deform_conv_cuda.cu:75-118 + modulated_deform_im2col_cuda.cuh:25-54,
    140-193, restated with dense tensor ops (float64-capable).  ``mask=None`` gives DCNv1
    (deform_im2col_cuda.cuh:127-189).  All of stride/pad/dil are (h, w) pairs."""
    B, Cin, H, W = inp.shape
    Cout, Cg, kh, kw = weight.shape
    Ho = (H + 2 * pad[0] - (dil[0] * (kh - 1) + 1)) // stride[0] + 1
    Wo = (W + 2 * pad[1] - (dil[1] * (kw - 1) + 1)) // stride[1] + 1
    dt = inp.dtype
    ys = (torch.arange(Ho, dtype=dt) * stride[0] - pad[0]).view(1, 1, Ho, 1)
    xs = (torch.arange(Wo, dtype=dt) * stride[1] - pad[1]).view(1, 1, 1, Wo)
    cpg = Cin // dgroups
    cols = inp.new_zeros(B, Cin, kh * kw, Ho, Wo)
    flat = inp.reshape(B, Cin, H * W)
    for t in range(kh * kw):
        i, j = divmod(t, kw)
        for g in range(dgroups):
            oh = offset[:, g * 2 * kh * kw + 2 * t].unsqueeze(1)
            ow = offset[:, g * 2 * kh * kw + 2 * t + 1].unsqueeze(1)
            hi = ys + i * dil[0] + oh
            wi = xs + j * dil[1] + ow
            inside = (hi > -1) & (wi > -1) & (hi < H) & (wi < W)
            hl = torch.floor(hi)
            wl = torch.floor(wi)
            lh, lw = hi - hl, wi - wl
            hh, hw = 1 - lh, 1 - lw
            hl, wl = hl.long(), wl.long()
            val = 0
            src = flat[:, g * cpg:(g + 1) * cpg]
            for (yy, xx, wt) in ((hl, wl, hh * hw), (hl, wl + 1, hh * lw),
                                 (hl + 1, wl, lh * hw), (hl + 1, wl + 1, lh * lw)):
                ok = (yy >= 0) & (yy <= H - 1) & (xx >= 0) & (xx <= W - 1) & inside
                idx = (yy.clamp(0, H - 1) * W + xx.clamp(0, W - 1)).view(B, 1, Ho * Wo).expand(B, cpg, Ho * Wo)
                v = torch.gather(src, 2, idx).view(B, cpg, Ho, Wo)
                val = val + wt * v * ok.to(dt)
            if mask is not None:
                val = val * mask[:, g * kh * kw + t].unsqueeze(1)
            cols[:, g * cpg:(g + 1) * cpg, t] = val
    cols = cols.view(B, groups, (Cin // groups) * kh * kw, Ho * Wo)
    wg = weight.view(groups, Cout // groups, Cg * kh * kw)
    out = torch.einsum("gok,bgkm->bgom", wg, cols).reshape(B, Cout, Ho, Wo)
    return out + bias.view(1, -1, 1, 1)


# --------------------------------------------------------------------------- a11 (SPEC-DERIVED, UNPINNED)
def epipolar_attention(q, k, v, band):
    """models/CrossAttentionCostVolume/cross_attention_cost_volume.py is EMPTY (0 bytes) in the
    reference: there is nothing to pin against.  Definition taken from BASELINE.json north_star:
    per epipolar row, S = Q K^T / sqrt(C) restricted to 0 <= x_q - x_k < band, softmax over keys,
    O = P V.  q,k,v: [B,C,H,W] (bf16-rounded values, any float dtype); fp32 math.
    Returns (O [B,C,H,W] fp32, band scores [B,band,H,W] fp32 with -inf where x_q - d < 0)."""
    B, C, H, W = q.shape
    qf, kf, vf = q.float(), k.float(), v.float()
    s = torch.einsum("bcyq,bcyk->byqk", qf, kf) / math.sqrt(C)
    xq = torch.arange(W, device=q.device).view(W, 1)
    xk = torch.arange(W, device=q.device).view(1, W)
    valid = ((xq - xk) >= 0) & ((xq - xk) < band)
    s = s.masked_fill(~valid, float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = torch.einsum("byqk,bcyk->bcyq", p, vf)
    scores = s.new_full((B, band, H, W), float("-inf"))
    for d in range(min(band, W)):
        scores[:, d, :, d:] = torch.diagonal(s, offset=-d, dim1=2, dim2=3)
    return o, scores


# --------------------------------------------------------------------------- f1 ("next" row)
def upsample_convex(disp, mask, factor=8, mask_scale=None):
    """upsample_convex8, models/utils/upsample.py:22-38 (factor 8); RAFT upsample_flow uses 4.
    mask_scale: the `.25 * self.mask(hidden_state)` of Attempts/update.py:273 applied first."""
    N, _, H, W = disp.shape
    if mask_scale is not None:
        mask = mask_scale * mask
    m = torch.softmax(mask.view(N, 1, 9, factor, factor, H, W), dim=2)
    u = F.unfold(factor * disp, [3, 3], padding=1).view(N, 1, 9, 1, 1, H, W)
    u = torch.sum(m * u, dim=2).permute(0, 1, 4, 2, 5, 3)
    return u.reshape(N, 1, factor * H, factor * W)


def upsample_simple(disp, factor=8):
    """upsample_simple8, models/utils/upsample.py:40-43: factor * F.interpolate(bilinear, align_corners=True), with
    ATen's index / weight arithmetic written out (aten/src/ATen/native/UpSample.h: area_pixel_compute_scale,
    guard_index_and_lambda; cpu/UpSampleKernel.cpp: interpolate over the outer dimension first).  Indices and weights
    are ATen's; the accumulation order of ATen's vectorised CPU loop is not reproduced, so values agree with the
    reference to 2 ulp (pinned at 1e-4 px in tests/test_oracle_golden.py), not bit for bit."""
    N, C, H, W = disp.shape

    def axis(n_in, n_out):
        scale = (torch.tensor(float(n_in - 1)) / torch.tensor(float(n_out - 1))) if n_out > 1 else torch.tensor(0.0)
        src = scale * torch.arange(n_out, dtype=torch.float32)
        i0 = torch.clamp(src.floor().long(), max=n_in - 1)
        l1 = torch.clamp(src - i0.float(), 0.0, 1.0)
        i1 = i0 + (i0 < n_in - 1).long()
        return i0, i1, 1.0 - l1, l1

    y0, y1, ly0, ly1 = axis(H, H * factor)
    x0, x1, lx0, lx1 = axis(W, W * factor)
    top = disp[:, :, y0][..., x0] * lx0 + disp[:, :, y0][..., x1] * lx1
    bot = disp[:, :, y1][..., x0] * lx0 + disp[:, :, y1][..., x1] * lx1
    return factor * (top * ly0.view(-1, 1) + bot * ly1.view(-1, 1))


# --------------------------------------------------------------------------- f2 ("next" row): ConvGRU + DLC step
def gru_reset(b_pre, prev_h, inp):
    """ConvGRU.forward, Attempts/update.py:43-46 after conv_b: s = cat(sigmoid(b_pre) * prev_h, input)."""
    if prev_h is None:
        prev_h = torch.zeros_like(b_pre)                                  # update.py:38-39
    return torch.cat((torch.sigmoid(b_pre) * prev_h, inp), 1)


def gru_blend(z_pre, g_pre, prev_h, softsign=False):
    """ConvGRU.forward, Attempts/update.py:29-34,47-48: h = (1 - z) * prev_h + z * g."""
    if prev_h is None:
        prev_h = torch.zeros_like(g_pre)
    z = torch.sigmoid(z_pre)
    g = F.softsign(g_pre) if softsign else torch.tanh(g_pre)
    return (1 - z) * prev_h + z * g


def conv_gru(sd, prefix, inp, prev_h, softsign=False):
    """ConvGRU.forward, Attempts/update.py:36-49, from a state dict (keys prefix + conv_z.weight ...)."""
    h0 = torch.zeros(inp.size(0), sd[prefix + "conv_z.weight"].size(0), inp.size(2), inp.size(3),
                     dtype=inp.dtype, device=inp.device) if prev_h is None else prev_h
    x1 = torch.cat((inp, h0), 1)
    z_pre = F.conv2d(x1, sd[prefix + "conv_z.weight"], sd[prefix + "conv_z.bias"], 1, 1)
    b_pre = F.conv2d(x1, sd[prefix + "conv_b.weight"], sd[prefix + "conv_b.bias"], 1, 1)
    s = gru_reset(b_pre, h0, inp)
    g_pre = F.conv2d(s, sd[prefix + "conv_g.weight"], sd[prefix + "conv_g.bias"], 1, 1)
    return gru_blend(z_pre, g_pre, h0, softsign)


def head_window(old_volume, cur_disp, bounds_pre, n_points, consider_valid=True, recip=False):
    """OffsetPredictionHead's relu2 + chunk (update.py:104-105, 275-277) followed by the window tail."""
    lo, up = torch.chunk(torch.relu(bounds_pre), 2, dim=1)
    return dlc_window(old_volume, cur_disp, lo, up, n_points, consider_valid, recip)


def _conv_bn_relu(sd, prefix, x, eps=1e-5):
    y = F.conv2d(x, sd[prefix + "0.weight"], None, 1, 1)
    y = F.batch_norm(y, sd[prefix + "1.running_mean"], sd[prefix + "1.running_var"], sd[prefix + "1.weight"],
                     sd[prefix + "1.bias"], False, 0.0, eps)
    return F.relu(y)


def dlc_update_step(sd, old_volume, cur_disp, left_r, right_r, hidden, n_points, consider_valid=True, recip=False):
    """DisparityUpdateDLCWithMask.forward in eval mode, Attempts/update.py:263-317, from its state dict;
    images already at the disparity resolution.  Returns (disp, hidden_state, mask) with mask = .25 * logits."""
    err, unc = guidance(old_volume, cur_disp, left_r, right_r, recip=recip)
    feat = torch.cat([_conv_bn_relu(sd, "encoder.disparity_error_encoder.", err),
                      _conv_bn_relu(sd, "encoder.uncertain_encoder.", unc)], 1)
    hidden = conv_gru(sd, "gru.", feat, hidden)
    mask = F.conv2d(F.relu(F.conv2d(hidden, sd["mask.0.weight"], sd["mask.0.bias"], 1, 1)), sd["mask.2.weight"],
                    sd["mask.2.bias"])
    b_pre = F.conv2d(F.relu(F.conv2d(hidden, sd["offset.conv1.weight"], sd["offset.conv1.bias"], 1, 1)),
                     sd["offset.conv2.weight"], sd["offset.conv2.bias"], 1, 1)
    disp = head_window(old_volume, cur_disp, b_pre, n_points, consider_valid, recip)
    return disp, hidden, .25 * mask
