"""Host-side mirror of the reference's operator interface for the hot path.

Same class / function names, constructor arguments, forward signatures and error behaviour as the
reference's pure-PyTorch modules, so reference networks (LowCNN, RAFT_Stereo) use them as drop-ins
(see install.py).  All of them are parameter-free, so checkpoints are unaffected.  Under autograd the
correlation / group-wise volumes, soft-argmin, fixed and DLC windows, build_cost_volume_from_volume and disp_warp
run their own backward kernels; the remaining operators are forward-only and raise on backward.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.autograd.function import once_differentiable

from . import ops
from ._lib import SF_DIV_RECIP, SF_DIV_TRUE, SF_WIN_ABSOLUTE, SF_WIN_FIXED, SF_WIN_OFFSETS, SF_WIN_VARIANCE

# Scalar-division rounding used for sampling coordinates: the reference running on CUDA multiplies by the
# reciprocal (PyTorch CUDA div-by-scalar), on CPU it divides.  Default = what the reference does on a GPU.
DEFAULT_DIV_MODE = SF_DIV_RECIP


def _div(obj):
    """div_mode of a module; reference instances that carry our grafted forward (install.py) have no such attribute."""
    dm = getattr(obj, "div_mode", None)
    return DEFAULT_DIV_MODE if dm is None else dm


class _ForwardOnly(torch.autograd.Function):
    """Runs ``fn(*args)`` outside autograd; differentiating through it raises."""

    @staticmethod
    def forward(ctx, fn, name, *args):
        ctx.name = name
        with torch.no_grad():
            out = fn(*args)
        return out

    @staticmethod
    def backward(ctx, *grads):
        raise NotImplementedError("stereoformer_b200.%s: backward is not built yet (forward-only drop-in, "
                                  "SURVEY 8f-3)" % ctx.name)


class _Differentiable(torch.autograd.Function):
    """Runs ``fn(*args)`` on the CUDA kernels and ``bwd(args, grad, needs)`` (another kernel) for the backward."""

    @staticmethod
    def forward(ctx, fn, bwd, *args):
        ctx.bwd = bwd
        ctx.is_tensor = [torch.is_tensor(a) for a in args]
        ctx.consts = [None if t else a for t, a in zip(ctx.is_tensor, args)]
        ctx.save_for_backward(*[a for a in args if torch.is_tensor(a)])
        with torch.no_grad():
            out = fn(*args)
        ctx.multi = getattr(bwd, "all_outputs", False)   # bwd wants the gradients of every output
        if isinstance(out, tuple) and not ctx.multi:     # only the first output is differentiable (e.g. warp's mask)
            ctx.mark_non_differentiable(*out[1:])
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, grad, *more):
        saved = iter(ctx.saved_tensors)
        args = [next(saved) if t else c for t, c in zip(ctx.is_tensor, ctx.consts)]
        g = (grad,) + more if ctx.multi else grad.contiguous()
        grads = ctx.bwd(args, g, ctx.needs_input_grad[2:])
        return (None, None) + tuple(grads)


def _run(fn, name, *args, bwd=None):
    if torch.is_grad_enabled() and any(torch.is_tensor(a) and a.requires_grad for a in args):
        if bwd is not None:
            return _Differentiable.apply(fn, bwd, *args)
        return _ForwardOnly.apply(fn, name, *args)
    return fn(*args)


# ----------------------------------------------------------------------------- cost volumes
class CostVolume(nn.Module):
    """models/utils/build_cost_volume.py:6-54."""

    def __init__(self, max_disp, feature_similarity='correlation'):
        super().__init__()
        self.max_disp = max_disp
        self.feature_similarity = feature_similarity

    def forward(self, left_feature, right_feature):
        if self.feature_similarity == 'correlation':
            return _run(lambda l, r: ops.corr_volume(l, r, self.max_disp), "CostVolume", left_feature, right_feature,
                        bwd=lambda a, g, n: ops.corr_volume_backward(a[0], a[1], g, self.max_disp, 1, n))
        if self.feature_similarity == 'difference':
            return _run(lambda l, r: ops.shift_volume(l, r, self.max_disp, False), "CostVolume", left_feature,
                        right_feature)
        if self.feature_similarity == 'concat':
            return _run(lambda l, r: ops.shift_volume(l, r, self.max_disp, True), "CostVolume", left_feature,
                        right_feature)
        raise NotImplementedError


def build_gwc_volume(refimg_fea, targetimg_fea, maxdisp, num_groups):
    """models/CrossAttentionCostVolume/GWC_Cost_Volume.py:14-24 (dup models/utils/submodule.py:52-63)."""
    return _run(lambda l, r: ops.gwc_volume(l, r, maxdisp, num_groups), "build_gwc_volume", refimg_fea, targetimg_fea,
                bwd=lambda a, g, n: ops.corr_volume_backward(a[0], a[1], g, maxdisp, num_groups, n))


def build_corr(img_left, img_right, max_disp=40):
    """models/utils/disp_residual.py:9-19 (same correlation volume)."""
    return _run(lambda l, r: ops.corr_volume(l, r, max_disp), "build_corr", img_left, img_right,
                bwd=lambda a, g, n: ops.corr_volume_backward(a[0], a[1], g, max_disp, 1, n))


# ----------------------------------------------------------------------------- soft-argmin
class DisparityEstimation(nn.Module):
    """models/utils/estimation.py:7-33."""

    def __init__(self, max_disp, match_similarity=True):
        super().__init__()
        self.max_disp = max_disp
        self.match_similarity = match_similarity

    def forward(self, cost_volume):
        assert cost_volume.dim() == 4
        return _run(lambda cv: ops.soft_argmin(cv, self.match_similarity), "DisparityEstimation", cost_volume,
                    bwd=lambda a, g, n: (ops.soft_argmin_backward(a[0], g, self.match_similarity),))


class DisparityEstimationWithProb(DisparityEstimation):
    """models/utils/estimation.py:37-62: also returns the probability volume."""

    def forward(self, cost_volume):
        assert cost_volume.dim() == 4
        return _run(lambda cv: ops.soft_argmin(cv, self.match_similarity, with_prob=True),
                    "DisparityEstimationWithProb", cost_volume)


# ----------------------------------------------------------------------------- local windows
def build_cost_volume_from_volume(old_volume, sampling_candidates):
    """fixed_local_cost_volume.py:61-80 (identical copies in variance_/pure_learning_/update*.py)."""
    return _run(ops.sample_volume, "build_cost_volume_from_volume", old_volume, sampling_candidates,
                bwd=lambda a, g, n: ops.sample_volume_backward(a[0], a[1], g, n))


class LocalCostVolume(nn.Module):
    """models/LocalCostVolume/Attempts/fixed_local_cost_volume.py:13-58."""

    div_mode = None          # None = DEFAULT_DIV_MODE at call time

    def __init__(self, radius=4, sample_points=10) -> None:
        super().__init__()
        self.radius = radius
        self.sample_points = sample_points

    def forward(self, old_cost_volume, cur_disparity, consider_valid=False):
        dm = _div(self)
        return _run(lambda cv, d: ops.local_window(cv, d, mode=SF_WIN_FIXED, radius=self.radius,
                                                   n_points=self.sample_points, consider_valid=consider_valid,
                                                   div_mode=dm),
                    "LocalCostVolume", old_cost_volume, cur_disparity,
                    bwd=lambda a, g, n: ops.local_window_backward(
                        a[0], a[1], None, None, g, SF_WIN_FIXED, self.radius, self.sample_points, consider_valid,
                        dm, (n[0], n[1], False, False))[:2])


class Variance_Based_Local_Cost_Volume(nn.Module):
    """models/LocalCostVolume/Attempts/variance_local_cost_volume.py:28-78."""

    div_mode = None

    def __init__(self, garma=1, sample_points=10):
        super().__init__()
        self.sample_points = sample_points
        self.garma = garma

    def forward(self, old_cost_volume, cur_disparity, consider_valid=False):
        dm = _div(self)
        return _run(lambda cv, d: ops.local_window(cv, d, mode=SF_WIN_VARIANCE, gamma=self.garma,
                                                   n_points=self.sample_points, consider_valid=consider_valid,
                                                   div_mode=dm),
                    "Variance_Based_Local_Cost_Volume", old_cost_volume, cur_disparity,
                    bwd=lambda a, g, n: ops.local_window_backward(
                        a[0], a[1], None, None, g, mode=SF_WIN_VARIANCE, radius=self.garma,
                        n_points=self.sample_points, consider_valid=consider_valid, div_mode=dm,
                        needs=(n[0], n[1], False, False))[:2])


def dlc_window(old_cost_volume, cur_disp, lower_offset, upper_offset, sample_points, consider_valid=True,
               absolute=False, div_mode=None):
    """Window tail of DisparityUpdateDLC(WithMask).forward, Attempts/update.py:279-313: bounds are
    cur_disp - lower_offset / cur_disp + upper_offset (absolute=True: the tensors are the bounds themselves,
    pure_learning_method.py:112)."""
    dm = DEFAULT_DIV_MODE if div_mode is None else div_mode
    mode = SF_WIN_ABSOLUTE if absolute else SF_WIN_OFFSETS
    return _run(lambda cv, d, lo, up: ops.local_window(cv, d, lo, up, mode=mode, n_points=sample_points,
                                                       consider_valid=consider_valid, div_mode=dm),
                "dlc_window", old_cost_volume, cur_disp, lower_offset, upper_offset,
                bwd=lambda a, g, n: ops.local_window_backward(a[0], a[1], a[2], a[3], g, mode, 0.0, sample_points,
                                                              consider_valid, dm, n))


def guidance_inputs(old_cost_volume, cur_disp, cur_left_image, cur_right_image, div_mode=None, check=False):
    """Fused front half of BasicGuidanceNet.forward (update.py:71-86) -> (error_map, uncertainty_volume).  check=True
    restores the reference's ``assert disp.min() >= 0`` (disp_warp inside the prologue) at the price of a host sync."""
    dm = DEFAULT_DIV_MODE if div_mode is None else div_mode

    def bwd(a, g, n):                                    # g = (grad_error_map, grad_uncertainty_volume)
        if n[2] or n[3]:
            raise NotImplementedError("stereoformer_b200.guidance_inputs: no gradient w.r.t. the images (inputs)")
        return ops.guidance_backward(a[0], a[1], a[3], g[0], g[1], dm, (n[0], n[1])) + (None, None)
    bwd.all_outputs = True
    return _run(lambda cv, d, l, r: ops.guidance(cv, d, l, r, div_mode=dm, check=check), "guidance_inputs", old_cost_volume,
                cur_disp, cur_left_image, cur_right_image, bwd=bwd)


# ----------------------------------------------------------------------------- warp
def disp_warp(img, disp, padding_mode='border'):
    """utils/disparity_warper.py:83-106 -> (warped_img, valid_mask); asserts disp.min() >= 0."""
    return _run(lambda i, d: ops.disp_warp(i, d, padding_mode, div_mode=DEFAULT_DIV_MODE), "disp_warp", img, disp,
                bwd=lambda a, g, n: ops.disp_warp_backward(a[0], a[1], g, padding_mode, DEFAULT_DIV_MODE, n))


def LRwarp_error(imgL, disp, imgR):
    """utils/disparity_warper.py:109-115."""
    if imgL.size(-1) > disp.size(-1):
        imgL = F.interpolate(imgL, size=[disp.size(-2), disp.size(-1)], mode='bilinear', align_corners=False)
    if imgR.size(-1) > disp.size(-1):
        imgR = F.interpolate(imgR, size=[disp.size(-2), disp.size(-1)], mode='bilinear', align_corners=False)
    warped_right, _ = disp_warp(imgL, disp)
    return imgR - warped_right


# ----------------------------------------------------------------------------- RAFT-Stereo 1-D correlation
class CorrBlock1D:
    """models/IterativeLocalCostVolume/corr/corr.py:54-111."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels = num_levels
        self.radius = radius
        # forward-only: under autograd the pyramid carries a backward that raises instead of silently dropping gradients
        self.corr_pyramid = list(_run(lambda a, b: tuple(ops.corr1d_pyramid(a, b, num_levels, extra_level=True)),
                                      "CorrBlock1D", fmap1, fmap2))

    def __call__(self, coords):
        return _run(lambda c, *pyr: ops.corr1d_lookup(list(pyr), c, self.radius, self.num_levels, _div(self)),
                    "CorrBlock1D", coords, *self.corr_pyramid)

    @staticmethod
    def corr(fmap1, fmap2):
        B, D, H, W1 = fmap1.shape
        W2 = fmap2.shape[3]
        with torch.no_grad():
            lv = ops.corr1d_pyramid(fmap1, fmap2, 1, extra_level=False)
        return lv[0].view(B, H, W1, 1, W2)


class PytorchAlternateCorrBlock1D:
    """models/IterativeLocalCostVolume/corr/corr.py:8-51: the on-the-fly variant (no all-pairs volume)."""

    def __init__(self, fmap1, fmap2, num_levels=4, radius=4):
        self.num_levels = num_levels
        self.radius = radius
        self.corr_pyramid = []
        self.fmap1 = fmap1
        self.fmap2 = fmap2

    def __call__(self, coords):
        return _run(lambda c, a, b: ops.corr1d_alt(a, b, c, self.radius, self.num_levels, _div(self)),
                    "PytorchAlternateCorrBlock1D", coords, self.fmap1, self.fmap2)


# ----------------------------------------------------------------------------- attention (spec-derived)
class EpipolarCrossAttention(nn.Module):
    """Band-restricted epipolar row cross-attention.  The reference's
    models/CrossAttentionCostVolume/cross_attention_cost_volume.py is empty; defined from north_star."""

    def __init__(self, band):
        super().__init__()
        self.band = band

    def forward(self, q, k, v, return_scores=False):
        return _run(lambda a, b, c: ops.epipolar_attention(a, b, c, self.band, return_scores),
                    "EpipolarCrossAttention", q, k, v)


# ----------------------------------------------------------------------------- upsampling ("next" row)
def upsample_convex8(disp, mask):
    """models/utils/upsample.py:22-38."""
    return _run(lambda d, m: ops.upsample_convex(d, m, 8), "upsample_convex8", disp, mask,
                bwd=lambda a, g, n: ops.upsample_convex_backward(a[0], a[1], g, 8, 1.0, n))


def upsample_simple8(disp, mode='bilinear'):
    """models/utils/upsample.py:40-43: 8 * F.interpolate(disp, size=8x, mode=mode, align_corners=True)."""
    if mode != 'bilinear':
        raise NotImplementedError("upsample_simple8: only mode='bilinear' (the reference's default and only use, "
                                  "baseline.py:138) is built")
    return _run(lambda d: ops.upsample_simple(d, 8), "upsample_simple8", disp,
                bwd=lambda a, g, n: (ops.upsample_simple_backward(g, 8),))
