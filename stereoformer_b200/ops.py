"""Functional operators: thin PyTorch wrappers over the C ABI (device memory + stream plumbing only).

Every function takes CUDA tensors, launches hand-written sm_100a kernels from libsfb200.so on the
current PyTorch stream and returns freshly allocated CUDA tensors.  CPU tensors raise
``RuntimeError("Not implemented on the CPU")`` like the reference's DCN extension
(deform/src/modulated_deform_conv.h:43); nothing here computes on the host.
"""
import ctypes
import os

import torch

from . import _lib
from ._lib import (SF_ALGO_AUTO, SF_ALGO_EXACT, SF_ALGO_GENERIC, SF_ALGO_TC, SF_ALGO_TMA, SF_DIV_RECIP, SF_DIV_TRUE, SF_WIN_ABSOLUTE,
                   SF_WIN_FIXED, SF_WIN_HEAD, SF_WIN_OFFSETS, SF_WIN_VARIANCE, SF_WIN_VARIANCE_BOUNDS)

__all__ = ["corr_volume", "gwc_volume", "shift_volume", "soft_argmin", "local_window", "sample_volume", "disp_warp",
           "guidance", "corr1d_pyramid", "corr1d_lookup", "dcn_forward", "dcn_backward", "corr_volume_backward", "soft_argmin_backward",
           "local_window_backward", "sample_volume_backward", "disp_warp_backward", "gru_reset_backward",
           "gru_blend_backward", "guidance_backward", "upsample_convex_backward", "epipolar_attention", "upsample_convex",
           "gru_reset", "gru_blend"]


# ----------------------------------------------------------------------------- precision of the contraction kernels
# "fast" (default): contractions with >= 32 channels per group run on tcgen05 with a bf16 hi + mid operand split
#   (~16-bit operand mantissa, fp32 accumulation): cost / group-wise volumes and the CorrBlock1D pyramid agree with the
#   fp32 reference to ~5e-6 of the volume's magnitude (north_star allows 1e-4), DCN to ~1e-5 of max |out|.
# "exact": fp32 FFMA kernels everywhere (the reference's 1e-5 analytic DCN tests, bit-level work).
# Operators called with an explicit algo= are not affected.
_PRECISION = "fast"


def set_precision(mode):
    """'fast' or 'exact'; returns the previous mode."""
    global _PRECISION
    if mode not in ("fast", "exact"):
        raise ValueError("precision must be 'fast' or 'exact'")
    old, _PRECISION = _PRECISION, mode
    return old


def get_precision():
    return _PRECISION


class precision:
    """with ops.precision('exact'): ..."""

    def __init__(self, mode):
        self.mode = mode

    def __enter__(self):
        self.old = set_precision(self.mode)

    def __exit__(self, *exc):
        set_precision(self.old)
        return False


def _algo(algo):
    if algo is None:
        return SF_ALGO_EXACT if _PRECISION == "exact" else SF_ALGO_AUTO
    return int(algo)


# The torch C++-extension front end (csrc/torch_ext.cpp; same C ABI underneath) takes the plain calls of the operators
# that the reference's eager forward issues every step; ctypes remains the binding for everything else and the
# fallback when the extension has not been built.
try:
    from . import _sfb200_torch as _ext
except ImportError:          # pragma: no cover -- built by stereoformer_b200.build / __graft_entry__.build()
    _ext = None
if os.environ.get("SF_TUNING_LIB") == "1":
    _ext = None              # the extension is linked against the product library; tools/ want the tuning build


def _plain(*ts):
    return all(t is None or (torch.is_tensor(t) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous())
               for t in ts)


def _chk(t, name, dtype=torch.float32, dim=None):
    if not torch.is_tensor(t):
        raise TypeError("%s must be a tensor" % name)
    if not t.is_cuda:
        raise RuntimeError("Not implemented on the CPU (%s is a CPU tensor; stereoformer_b200 has no CPU path)" % name)
    if t.dtype != dtype:
        raise TypeError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if dim is not None and t.dim() != dim:
        raise ValueError("%s must be %d-D, got shape %s" % (name, dim, tuple(t.shape)))
    return t.contiguous()


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def corr_volume(left, right, max_disp, groups=1, algo=None, out=None):
    """CostVolume('correlation') / build_gwc_volume.  Returns [B,D,H,W] (groups=1) or [B,G,D,H,W]
    (written into ``out`` when given)."""
    if _ext is not None and _plain(left, right, out) and left.dim() == 4:
        return _ext.corr_volume(left, right, int(max_disp), int(groups), _algo(algo), out)
    left, right = _chk(left, "left_feature", dim=4), _chk(right, "right_feature", dim=4)
    if left.shape != right.shape:
        raise ValueError("feature shapes differ: %s vs %s" % (tuple(left.shape), tuple(right.shape)))
    B, C, H, W = left.shape
    shape = (B, max_disp, H, W) if groups == 1 else (B, groups, max_disp, H, W)
    if out is None:
        out = torch.empty(shape, device=left.device, dtype=torch.float32)
    elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous() or out.device != left.device:
        raise ValueError("out must be a contiguous fp32 CUDA tensor of shape %s" % (shape,))
    with torch.cuda.device(left.device):
        _lib.call("sf_corr_volume_f32", _ptr(left), _ptr(right), _ptr(out), B, C, H, W, int(max_disp), int(groups),
                  _algo(algo), _stream())
    return out


def gwc_volume(left, right, max_disp, num_groups):
    if left.shape[1] % num_groups != 0:
        raise AssertionError("channels not divisible by num_groups")
    return corr_volume(left, right, max_disp, groups=num_groups).view(left.shape[0], num_groups, max_disp,
                                                                      left.shape[2], left.shape[3])


def shift_volume(left, right, max_disp, concat=False):
    """CostVolume('difference' | 'concat'): [B, C or 2C, D, H, W]."""
    left, right = _chk(left, "left_feature", dim=4), _chk(right, "right_feature", dim=4)
    B, C, H, W = left.shape
    out = torch.empty((B, 2 * C if concat else C, max_disp, H, W), device=left.device, dtype=torch.float32)
    with torch.cuda.device(left.device):
        _lib.call("sf_shift_volume_f32", _ptr(left), _ptr(right), _ptr(out), B, C, H, W, int(max_disp), int(concat),
                  _stream())
    return out


def soft_argmin(cv, match_similarity=True, with_prob=False):
    if _ext is not None and not with_prob and _plain(cv) and cv.dim() == 4:
        return _ext.soft_argmin(cv, bool(match_similarity))
    cv = _chk(cv, "cost_volume", dim=4)
    B, D, H, W = cv.shape
    disp = torch.empty((B, H, W), device=cv.device, dtype=torch.float32)
    prob = torch.empty_like(cv) if with_prob else None
    with torch.cuda.device(cv.device):
        _lib.call("sf_soft_argmin_f32", _ptr(cv), _ptr(disp), _ptr(prob), B, D, H, W, int(bool(match_similarity)),
                  _stream())
    return (disp, prob) if with_prob else disp


def _check_bounds(lo, up, mode, n):
    """lo / up element counts the kernels read: HEAD takes lo as the [B,2,H,W] head output, OFFSETS / ABSOLUTE and
    VARIANCE_BOUNDS take one value per pixel each (a wrong shape would be an out-of-bounds device read)."""
    if mode == SF_WIN_HEAD:
        if lo is None or lo.numel() != 2 * n:
            raise ValueError("SF_WIN_HEAD needs lo = the offset head's [B,2,H,W] output (%d elements)" % (2 * n))
    elif mode in (SF_WIN_OFFSETS, SF_WIN_ABSOLUTE):
        if lo is None or up is None or lo.numel() != n or up.numel() != n:
            raise ValueError("lower / upper bounds must have B*H*W = %d elements each" % n)
    else:
        for t, name in ((lo, "lower"), (up, "upper")):
            if t is not None and t.numel() != n:
                raise ValueError("%s bound must have B*H*W = %d elements" % (name, n))


def local_window(cv, disp=None, lo=None, up=None, mode=SF_WIN_FIXED, radius=0.0, gamma=1.0, n_points=10,
                 consider_valid=False, div_mode=SF_DIV_TRUE, debug=False, return_disp0=False, out=None, peers=None,
                 peer_offset=0, gather=None):
    """Local disparity-window sampling + soft-max expectation -> [B,1,H,W] (written into ``out`` when given).

    debug=True additionally returns dict(cand, idx_floor, idx_ceil, local) for the bit-exact checks.
    peers: list of device pointers (ints) of every rank's gather buffer (NVLink peer mappings, see
    dist.PeerGather); the result is also stored at element ``peer_offset`` + pixel index of each of them.
    gather: a ``dist.SignalledGather``; the result goes to its destinations' slots with the completion protocol of
    sf_local_window_gather_signal_f32 (consumers call ``gather.wait()``)."""
    if (_ext is not None and not debug and not return_disp0 and not peers and gather is None and
            _plain(cv, disp, lo, up, out) and cv.dim() == 4):
        return _ext.local_window(cv, disp, lo, up, int(mode), float(radius), float(gamma), int(n_points),
                                 bool(consider_valid), int(div_mode), out)
    cv = _chk(cv, "old_cost_volume", dim=4)
    B, D, H, W = cv.shape
    if disp is not None:
        disp = _chk(disp, "cur_disparity")
        if disp.numel() != B * H * W:
            raise ValueError("cur_disparity must have B*H*W elements")
    lo = _chk(lo, "lower bound") if lo is not None else None
    up = _chk(up, "upper bound") if up is not None else None
    _check_bounds(lo, up, int(mode), B * H * W)
    if out is None:
        out = torch.empty((B, 1, H, W), device=cv.device, dtype=torch.float32)
    elif out.dtype != torch.float32 or out.numel() != B * H * W or not out.is_contiguous() or out.device != cv.device:
        raise ValueError("out must be a contiguous fp32 CUDA tensor with B*H*W elements")
    d0 = torch.empty((B, 1, H, W), device=cv.device, dtype=torch.float32) if return_disp0 else None
    dbg = {}
    if debug:
        K = n_points + 1
        dbg = dict(cand=torch.empty((B, K, H, W), device=cv.device, dtype=torch.float32),
                   idx_floor=torch.empty((B, K, H, W), device=cv.device, dtype=torch.int32),
                   idx_ceil=torch.empty((B, K, H, W), device=cv.device, dtype=torch.int32),
                   local=torch.empty((B, K, H, W), device=cv.device, dtype=torch.float32))
    if gather is not None:
        if debug or return_disp0 or peers:
            raise ValueError("gather cannot be combined with debug / return_disp0 / peers")
        if (B, H, W) != gather.shape:
            raise ValueError("gather was built for maps of shape %r, got %r" % (gather.shape, (B, H, W)))
        with torch.cuda.device(cv.device):
            _lib.call("sf_local_window_gather_signal_f32", _ptr(cv), _ptr(disp), _ptr(lo), _ptr(up), _ptr(out), B, D, H,
                      W, int(mode), float(radius), float(gamma), int(n_points), int(bool(consider_valid)),
                      int(div_mode), gather.c_dst_data, gather.c_dst_sig, gather.c_dst_ranks, len(gather.dst_ranks),
                      ctypes.c_longlong(gather.offset()), ctypes.c_longlong(gather.slot_stride),
                      ctypes.c_void_p(gather.sig_ptrs[gather.rank]), _stream())
        return out
    if peers:
        if debug or return_disp0:
            raise ValueError("peers cannot be combined with debug / return_disp0")
        arr = (ctypes.c_void_p * len(peers))(*[ctypes.c_void_p(int(q)) for q in peers])
        with torch.cuda.device(cv.device):
            _lib.call("sf_local_window_gather_f32", _ptr(cv), _ptr(disp), _ptr(lo), _ptr(up), _ptr(out), B, D, H, W,
                      int(mode), float(radius), float(gamma), int(n_points), int(bool(consider_valid)), int(div_mode),
                      arr, len(peers), ctypes.c_longlong(int(peer_offset)), _stream())
        return out
    with torch.cuda.device(cv.device):
        _lib.call("sf_local_window_f32", _ptr(cv), _ptr(disp), _ptr(lo), _ptr(up), _ptr(out), _ptr(d0),
                  _ptr(dbg.get("cand")), _ptr(dbg.get("idx_floor")), _ptr(dbg.get("idx_ceil")), _ptr(dbg.get("local")),
                  B, D, H, W, int(mode), float(radius), float(gamma), int(n_points), int(bool(consider_valid)),
                  int(div_mode), _stream())
    res = (out,)
    if return_disp0:
        res += (d0,)
    if debug:
        res += (dbg,)
    return res[0] if len(res) == 1 else res


def sample_volume(old_volume, cand, return_indices=False):
    """build_cost_volume_from_volume(old_volume, sampling_candidates) -> [B,K,H,W] (bit-exact)."""
    cv, cand = _chk(old_volume, "old_volume", dim=4), _chk(cand, "sampling_candidates", dim=4)
    B, D, H, W = cv.shape
    K = cand.shape[1]
    if cand.shape != (B, K, H, W):
        raise ValueError("sampling_candidates shape %s does not match volume %s" % (tuple(cand.shape), tuple(cv.shape)))
    out = torch.empty_like(cand)
    jf = torch.empty(cand.shape, device=cv.device, dtype=torch.int32) if return_indices else None
    ic = torch.empty(cand.shape, device=cv.device, dtype=torch.int32) if return_indices else None
    with torch.cuda.device(cv.device):
        _lib.call("sf_sample_volume_f32", _ptr(cv), _ptr(cand), _ptr(out), _ptr(jf), _ptr(ic), B, D, K, H, W, _stream())
    return (out, jf, ic) if return_indices else out


def disp_warp(img, disp, padding_mode="border", div_mode=SF_DIV_TRUE, check=True, with_mask=True, debug=False):
    """disp_warp -> (warped, valid_mask).  check=True keeps the reference's ``assert disp.min() >= 0``
    (one device->host read, like the reference); check=False is sync-free (CUDA-graph safe)."""
    img, disp = _chk(img, "img", dim=4), _chk(disp, "disp")
    B, C, H, W = img.shape
    if disp.numel() != B * H * W:
        raise ValueError("disp must be [B,1,H,W]")
    if padding_mode not in ("border", "zeros"):
        raise ValueError("padding_mode must be 'border' or 'zeros'")
    warped = torch.empty_like(img)
    mask = torch.empty_like(img) if with_mask else None
    flag = torch.zeros(1, device=img.device, dtype=torch.int32) if check else None
    grid = torch.empty((B, H, W, 2), device=img.device, dtype=torch.float32) if debug else None
    src = torch.empty((B, H, W, 2), device=img.device, dtype=torch.float32) if debug else None
    with torch.cuda.device(img.device):
        _lib.call("sf_disp_warp_f32", _ptr(img), _ptr(disp), _ptr(warped), _ptr(mask), _ptr(grid), _ptr(src), B, C, H, W,
                  int(padding_mode == "border"), int(div_mode), _ptr(flag), _stream())
    if check and int(flag.item()) != 0:
        raise AssertionError("disp.min() >= 0 violated")
    if debug:
        return warped, mask, dict(grid=grid, src=src)
    return warped, mask


def disp_warp_band(img_src, disp, full_height, y0, src_y0, padding_mode="border", div_mode=SF_DIV_TRUE, check=True,
                   with_mask=True, debug=False):
    """disp_warp on the row band [y0, y0 + rows) of an image of ``full_height`` rows (partition 2).

    ``img_src`` [B,C,src_rows,W] holds rows [src_y0, src_y0 + src_rows) and must include a 1-row halo; ``disp``
    [B,1,rows,W].  Equals the same rows of the whole-image call bit for bit."""
    img_src, disp = _chk(img_src, "img", dim=4), _chk(disp, "disp", dim=4)
    B, C, S, W = img_src.shape
    rows = disp.shape[2]
    if disp.shape != (B, 1, rows, W):
        raise ValueError("disp must be [B,1,rows,W]")
    if padding_mode not in ("border", "zeros"):
        raise ValueError("padding_mode must be 'border' or 'zeros'")
    dev = img_src.device
    warped = torch.empty((B, C, rows, W), device=dev, dtype=torch.float32)
    mask = torch.empty_like(warped) if with_mask else None
    flag = torch.zeros(1, device=dev, dtype=torch.int32) if check else None
    grid = torch.empty((B, rows, W, 2), device=dev, dtype=torch.float32) if debug else None
    src = torch.empty((B, rows, W, 2), device=dev, dtype=torch.float32) if debug else None
    with torch.cuda.device(dev):
        _lib.call("sf_disp_warp_band_f32", _ptr(img_src), _ptr(disp), _ptr(warped), _ptr(mask), _ptr(grid), _ptr(src),
                  B, C, rows, W, int(full_height), int(y0), int(src_y0), S, int(padding_mode == "border"),
                  int(div_mode), _ptr(flag), _stream())
    if check and int(flag.item()) != 0:
        raise AssertionError("disp.min() >= 0 violated")
    if debug:
        return warped, mask, dict(grid=grid, src=src)
    return warped, mask


def guidance_band(cv, disp, left, right_src, full_height, y0, src_y0, div_mode=SF_DIV_TRUE, flag=None):
    """guidance() on a row band: cv/disp/left hold the band's rows, ``right_src`` rows [src_y0, ...) with a 1-row halo."""
    cv, disp = _chk(cv, "old_cost_volume", dim=4), _chk(disp, "cur_disp")
    left, right_src = _chk(left, "left_image", dim=4), _chk(right_src, "right_image", dim=4)
    B, D, rows, W = cv.shape
    C, S = left.shape[1], right_src.shape[2]
    if left.shape != (B, C, rows, W) or right_src.shape != (B, C, S, W) or disp.numel() != B * rows * W:
        raise ValueError("band tensors must be [B,*,rows,W]; right_src [B,C,src_rows,W]")
    err = torch.empty_like(left)
    unc = torch.empty_like(cv)
    with torch.cuda.device(cv.device):
        _lib.call("sf_guidance_band_f32", _ptr(cv), _ptr(disp), _ptr(left), _ptr(right_src), _ptr(err), _ptr(unc), B,
                  D, C, rows, W, int(full_height), int(y0), int(src_y0), S, int(div_mode), _ptr(flag), _stream())
    return err, unc


def guidance(cv, disp, left, right, div_mode=SF_DIV_TRUE, flag=None, check=False):
    """Fused BasicGuidanceNet prologue -> (error_map [B,C,H,W], uncertainty_volume [B,D,H,W]).  check=True keeps the
    ``assert disp.min() >= 0`` of the reference's disp_warp (disparity_warper.py:93) that runs inside this prologue
    (one device->host read; the default is sync-free so the refinement loop can be captured in a CUDA graph)."""
    if check and flag is None:
        flag = torch.zeros(1, device=cv.device, dtype=torch.int32)
    cv, disp = _chk(cv, "old_cost_volume", dim=4), _chk(disp, "cur_disp")
    left, right = _chk(left, "left_image", dim=4), _chk(right, "right_image", dim=4)
    B, D, H, W = cv.shape
    C = left.shape[1]
    if left.shape != (B, C, H, W) or right.shape != left.shape:
        raise ValueError("images must be [B,C,H,W] at the disparity resolution")
    err = torch.empty_like(left)
    unc = torch.empty_like(cv)
    with torch.cuda.device(cv.device):
        _lib.call("sf_guidance_f32", _ptr(cv), _ptr(disp), _ptr(left), _ptr(right), _ptr(err), _ptr(unc), B, D, C, H, W,
                  int(div_mode), _ptr(flag), _stream())
    if check and int(flag.item()) != 0:
        raise AssertionError("disp.min() >= 0 violated")
    return err, unc


def corr1d_pyramid(f1, f2, num_levels=4, extra_level=True, algo=None):
    """CorrBlock1D pyramid: list of [B*H*W1,1,1,W2>>l]; extra_level keeps the reference's unused last level."""
    f1, f2 = _chk(f1, "fmap1", dim=4), _chk(f2, "fmap2", dim=4)
    B, C, H, W1 = f1.shape
    W2 = f2.shape[3]
    n_out = num_levels + (1 if extra_level else 0)
    while n_out > 1 and (W2 >> (n_out - 1)) < 1:
        n_out -= 1
    lv = [torch.empty((B * H * W1, 1, 1, W2 >> l), device=f1.device, dtype=torch.float32) for l in range(n_out)]
    arr = (ctypes.c_void_p * n_out)(*[t.data_ptr() for t in lv])
    with torch.cuda.device(f1.device):
        _lib.call("sf_corr1d_pyramid_f32", _ptr(f1), _ptr(f2), ctypes.cast(arr, ctypes.c_void_p), n_out, B, C, H, W1, W2,
                  _algo(algo), _stream())
    return lv


def corr1d_lookup(pyramid, coords, radius=4, num_levels=4, div_mode=SF_DIV_TRUE, debug=False):
    coords = _chk(coords, "coords", dim=4)
    B, _, H, W1 = coords.shape
    lv = [_chk(p, "pyramid level") for p in pyramid[:num_levels]]
    ptrs = (ctypes.c_void_p * num_levels)(*[t.data_ptr() for t in lv])
    widths = (ctypes.c_int * num_levels)(*[t.shape[-1] for t in lv])
    ch = num_levels * (2 * radius + 1)
    out = torch.empty((B, ch, H, W1), device=coords.device, dtype=torch.float32)
    x0 = torch.empty_like(out) if debug else None
    src = torch.empty_like(out) if debug else None
    with torch.cuda.device(coords.device):
        _lib.call("sf_corr1d_lookup_f32", ctypes.cast(ptrs, ctypes.c_void_p), ctypes.cast(widths, ctypes.c_void_p),
                  num_levels, int(radius), _ptr(coords), _ptr(out), _ptr(x0), _ptr(src), B, H, W1, int(div_mode),
                  _stream())
    return (out, dict(x0=x0, src=src)) if debug else out


def corr1d_alt(f1, f2, coords, radius=4, num_levels=4, div_mode=SF_DIV_TRUE, debug=False):
    """PytorchAlternateCorrBlock1D.__call__: on-the-fly correlation lookup -> [B, L*(2r+1), H, W1]."""
    f1, f2, coords = _chk(f1, "fmap1", dim=4), _chk(f2, "fmap2", dim=4), _chk(coords, "coords", dim=4)
    B, C, H, W1 = f1.shape
    W2 = f2.shape[3]
    if f2.shape[:3] != (B, C, H) or coords.shape != (B, 2, H, W1):
        raise ValueError("fmap2 must be [B,C,H,W2] and coords [B,2,H,W1]")
    taps = 2 * radius + 1
    out = torch.empty((B, num_levels * taps, H, W1), device=f1.device, dtype=torch.float32)
    src = torch.empty((B, num_levels * taps, H, W1, 2), device=f1.device, dtype=torch.float32) if debug else None
    with torch.cuda.device(f1.device):
        nws = int(_lib.lib().sf_corr1d_alt_workspace_bytes(B, C, H, W2, int(num_levels)))
        ws = torch.empty(max(nws, 1), device=f1.device, dtype=torch.uint8)
        _lib.call("sf_corr1d_alt_f32", _ptr(f1), _ptr(f2), _ptr(coords), _ptr(out), _ptr(src), _ptr(ws),
                  ctypes.c_size_t(nws), B, C, H, W1, W2, int(num_levels), int(radius), int(div_mode), _stream())
    return (out, dict(src=src)) if debug else out


def dcn_forward(inp, weight, bias, offset, mask, stride, padding, dilation, groups, deformable_groups,
                algo=None):
    """DCNv2 (mask given) / DCNv1 (mask None) forward -> [B,Cout,Ho,Wo].

    algo: SF_ALGO_AUTO (tcgen05 implicit GEMM from 32 channels per group, exact FFMA kernel below),
    SF_ALGO_GENERIC (exact FFMA kernel), SF_ALGO_TC (force the tensor-core path)."""
    dt = inp.dtype if torch.is_tensor(inp) and inp.dtype == torch.float64 else torch.float32   # fp64: verification dtype
    inp, weight = _chk(inp, "input", dt, dim=4), _chk(weight, "weight", dt, dim=4)
    bias, offset = _chk(bias, "bias", dt), _chk(offset, "offset", dt, dim=4)
    mask = _chk(mask, "mask", dt, dim=4) if mask is not None else None
    B, Cin, H, W = inp.shape
    Cout, Cg, kh, kw = weight.shape
    sh, sw = stride
    ph, pw = padding
    dh, dw = dilation
    if Cin != Cg * groups:
        raise RuntimeError("Input shape and kernel channels wont match: (%d vs %d)." % (Cin, Cg * groups))
    Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) // sh + 1
    Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) // sw + 1
    if offset.shape != (B, deformable_groups * 2 * kh * kw, Ho, Wo):
        raise RuntimeError("offset shape %s, expected %s" % (tuple(offset.shape),
                                                             (B, deformable_groups * 2 * kh * kw, Ho, Wo)))
    if mask is not None and mask.shape != (B, deformable_groups * kh * kw, Ho, Wo):
        raise RuntimeError("mask shape %s, expected %s" % (tuple(mask.shape), (B, deformable_groups * kh * kw, Ho, Wo)))
    out = torch.empty((B, Cout, Ho, Wo), device=inp.device, dtype=dt)
    if dt == torch.float64:
        with torch.cuda.device(inp.device):
            _lib.call("sf_dcn_forward_f64", _ptr(inp), _ptr(weight), _ptr(bias), _ptr(offset), _ptr(mask), _ptr(out), B,
                      Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, dh, dw, int(groups), int(deformable_groups), _stream())
        return out
    with torch.cuda.device(inp.device):
        algo = _algo(algo)
        nws = int(_lib.lib().sf_dcn_workspace_bytes_ex(B, Cin, H, W, Cout, kh, kw, int(groups), int(deformable_groups))) \
            if algo not in (SF_ALGO_GENERIC, SF_ALGO_EXACT) else 0
        ws = torch.empty(nws, device=inp.device, dtype=torch.uint8) if nws else None   # bf16 hi/mid weights (+ channels-last input copy) (scratch)
        _lib.call("sf_dcn_forward_ws_f32", _ptr(inp), _ptr(weight), _ptr(bias), _ptr(offset), _ptr(mask), _ptr(out), B,
                  Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, dh, dw, int(groups), int(deformable_groups), _ptr(ws),
                  ctypes.c_size_t(nws), algo, _stream())
    return out


# ----------------------------------------------------------------------------- backward passes
def corr_volume_backward(left, right, grad_cv, max_disp, groups=1, needs=(True, True), algo=None):
    """Gradients of corr_volume / gwc_volume w.r.t. (left, right).  algo as corr_volume: the precision switch picks the
    tcgen05 band product ("fast", from 64 channels per group) or the fp32 FFMA kernel ("exact")."""
    left, right = _chk(left, "left_feature", dim=4), _chk(right, "right_feature", dim=4)
    grad_cv = _chk(grad_cv, "grad_cost_volume")
    B, C, H, W = left.shape
    if grad_cv.numel() != B * groups * max_disp * H * W:
        raise ValueError("grad_cost_volume has %d elements, expected %d" % (grad_cv.numel(), B * groups * max_disp * H * W))
    gl = torch.empty_like(left) if needs[0] else None
    gr = torch.empty_like(right) if needs[1] else None
    with torch.cuda.device(left.device):
        _lib.call("sf_corr_volume_backward_ex_f32", _ptr(left), _ptr(right), _ptr(grad_cv), _ptr(gl), _ptr(gr), B, C, H,
                  W, int(max_disp), int(groups), _algo(algo), _stream())
    return gl, gr


def soft_argmin_backward(cv, grad_disp, match_similarity=True):
    cv, grad_disp = _chk(cv, "cost_volume", dim=4), _chk(grad_disp, "grad_disp")
    B, D, H, W = cv.shape
    if grad_disp.numel() != B * H * W:
        raise ValueError("grad_disp must have B*H*W elements")
    gcv = torch.empty_like(cv)
    with torch.cuda.device(cv.device):
        _lib.call("sf_soft_argmin_backward_f32", _ptr(cv), _ptr(grad_disp), _ptr(gcv), B, D, H, W,
                  int(bool(match_similarity)), _stream())
    return gcv


def local_window_backward(cv, disp, lo, up, grad_out, mode=SF_WIN_FIXED, radius=0.0, n_points=10, consider_valid=False,
                          div_mode=SF_DIV_TRUE, needs=(True, True, True, True)):
    """Gradients of local_window w.r.t. (cv, disp, lo, up); modes FIXED / OFFSETS / ABSOLUTE / HEAD."""
    cv, disp, grad_out = _chk(cv, "old_cost_volume", dim=4), _chk(disp, "cur_disparity"), _chk(grad_out, "grad_out")
    B, D, H, W = cv.shape
    if disp.numel() != B * H * W or grad_out.numel() != B * H * W:
        raise ValueError("cur_disparity / grad_out must have B*H*W elements")
    lo = _chk(lo, "lower bound") if lo is not None else None
    up = _chk(up, "upper bound") if up is not None else None
    _check_bounds(lo, up, int(mode), B * H * W)
    gcv = torch.empty_like(cv) if needs[0] else None
    gd = torch.empty_like(disp) if needs[1] else None
    glo = torch.empty_like(lo) if (needs[2] and lo is not None) else None
    gup = torch.empty_like(up) if (needs[3] and up is not None) else None
    with torch.cuda.device(cv.device):
        _lib.call("sf_local_window_backward_f32", _ptr(cv), _ptr(disp), _ptr(lo), _ptr(up), _ptr(grad_out), _ptr(gcv),
                  _ptr(gd), _ptr(glo), _ptr(gup), B, D, H, W, int(mode), float(radius), int(n_points),
                  int(bool(consider_valid)), int(div_mode), _stream())
    return gcv, gd, glo, gup


def sample_volume_backward(cv, cand, grad_local, needs=(True, True)):
    cv, cand, grad_local = _chk(cv, "old_volume", dim=4), _chk(cand, "sampling_candidates", dim=4), _chk(grad_local, "grad", dim=4)
    B, D, H, W = cv.shape
    K = cand.shape[1]
    gcv = torch.empty_like(cv) if needs[0] else None
    gc = torch.empty_like(cand) if needs[1] else None
    with torch.cuda.device(cv.device):
        _lib.call("sf_sample_volume_backward_f32", _ptr(cv), _ptr(cand), _ptr(grad_local), _ptr(gcv), _ptr(gc), B, D, K,
                  H, W, _stream())
    return gcv, gc


def disp_warp_backward(img, disp, grad_warped, padding_mode="border", div_mode=SF_DIV_TRUE, needs=(True, True)):
    img, disp, grad_warped = _chk(img, "img", dim=4), _chk(disp, "disp"), _chk(grad_warped, "grad_warped", dim=4)
    B, C, H, W = img.shape
    gi = torch.empty_like(img) if needs[0] else None
    gd = torch.empty_like(disp) if needs[1] else None
    with torch.cuda.device(img.device):
        _lib.call("sf_disp_warp_backward_f32", _ptr(img), _ptr(disp), _ptr(grad_warped), _ptr(gi), _ptr(gd), B, C, H, W,
                  int(padding_mode == "border"), int(div_mode), _stream())
    return gi, gd


def gru_reset_backward(b_pre, prev_h, inp_shape, grad_s, needs=(True, True, True)):
    """Gradients of gru_reset w.r.t. (b_pre, prev_h, input)."""
    grad_s = _chk(grad_s, "grad_s", dim=4)
    B, Cx, H, W = inp_shape
    C = b_pre.shape[1]
    b_pre, bs = _gate_view(b_pre, "b_pre", B, C, H, W)
    prev_h = _chk(prev_h, "prev_h", dim=4) if prev_h is not None else None
    dev = grad_s.device
    gb = torch.empty((B, C, H, W), device=dev, dtype=torch.float32) if needs[0] else None
    gh = torch.empty((B, C, H, W), device=dev, dtype=torch.float32) if (needs[1] and prev_h is not None) else None
    gx = torch.empty((B, Cx, H, W), device=dev, dtype=torch.float32) if needs[2] else None
    with torch.cuda.device(dev):
        _lib.call("sf_gru_reset_backward_f32", _ptr(b_pre), ctypes.c_longlong(bs), _ptr(prev_h), _ptr(grad_s), _ptr(gb),
                  _ptr(gh), _ptr(gx), B, C, Cx, H, W, _stream())
    return gb, gh, gx


def gru_blend_backward(z_pre, g_pre, prev_h, grad_h, softsign=False, needs=(True, True, True)):
    """Gradients of gru_blend w.r.t. (z_pre, g_pre, prev_h)."""
    g_pre, grad_h = _chk(g_pre, "g_pre", dim=4), _chk(grad_h, "grad_h", dim=4)
    B, C, H, W = g_pre.shape
    z_pre, zs = _gate_view(z_pre, "z_pre", B, C, H, W)
    prev_h = _chk(prev_h, "prev_h", dim=4) if prev_h is not None else None
    gz = torch.empty_like(g_pre) if needs[0] else None
    gg = torch.empty_like(g_pre) if needs[1] else None
    gh = torch.empty_like(g_pre) if (needs[2] and prev_h is not None) else None
    with torch.cuda.device(g_pre.device):
        _lib.call("sf_gru_blend_backward_f32", _ptr(z_pre), ctypes.c_longlong(zs), _ptr(g_pre), _ptr(prev_h), _ptr(grad_h),
                  _ptr(gz), _ptr(gg), _ptr(gh), B, C, H, W, int(bool(softsign)), _stream())
    return gz, gg, gh


def guidance_backward(cv, disp, right, grad_err, grad_unc, div_mode=SF_DIV_TRUE, needs=(True, True)):
    """Gradients of guidance w.r.t. (cv, disp); the images are inputs and get none."""
    cv, disp, right = _chk(cv, "old_cost_volume", dim=4), _chk(disp, "cur_disp"), _chk(right, "right_image", dim=4)
    grad_err = _chk(grad_err, "grad_err", dim=4) if grad_err is not None else None
    grad_unc = _chk(grad_unc, "grad_unc", dim=4) if grad_unc is not None else None
    B, D, H, W = cv.shape
    gcv = torch.empty_like(cv) if needs[0] else None
    gd = torch.empty_like(disp) if needs[1] else None
    with torch.cuda.device(cv.device):
        _lib.call("sf_guidance_backward_f32", _ptr(cv), _ptr(disp), _ptr(right), _ptr(grad_err), _ptr(grad_unc), _ptr(gcv),
                  _ptr(gd), B, D, right.shape[1], H, W, int(div_mode), _stream())
    return gcv, gd


def upsample_convex_backward(disp, mask, grad_out, factor=8, mask_scale=1.0, needs=(True, True)):
    disp, mask, grad_out = _chk(disp, "disp", dim=4), _chk(mask, "mask", dim=4), _chk(grad_out, "grad_out", dim=4)
    N, _, H, W = disp.shape
    gd = torch.empty_like(disp) if needs[0] else None
    gm = torch.empty_like(mask) if needs[1] else None
    with torch.cuda.device(disp.device):
        _lib.call("sf_upsample_convex_backward_f32", _ptr(disp), _ptr(mask), float(mask_scale), _ptr(grad_out), _ptr(gd),
                  _ptr(gm), N, H, W, int(factor), _stream())
    return gd, gm


def dcn_backward(inp, weight, offset, mask, grad_out, stride, padding, dilation, groups, deformable_groups,
                 needs=(True, True, True, True, True), algo=None):
    """DCNv2 / DCNv1 (mask None) backward -> (grad_input, grad_offset, grad_mask, grad_weight, grad_bias), the order
    of the reference's DCN.modulated_deform_conv_backward (modulated_deform_conv_cuda.cu:277-279); entries whose
    ``needs`` flag is False (or grad_mask for DCNv1) are None."""
    dt = inp.dtype if torch.is_tensor(inp) and inp.dtype == torch.float64 else torch.float32
    inp, weight = _chk(inp, "input", dt, dim=4), _chk(weight, "weight", dt, dim=4)
    offset, grad_out = _chk(offset, "offset", dt, dim=4), _chk(grad_out, "grad_output", dt, dim=4)
    mask = _chk(mask, "mask", dt, dim=4) if mask is not None else None
    B, Cin, H, W = inp.shape
    Cout, Cg, kh, kw = weight.shape
    sh, sw = stride
    ph, pw = padding
    dh, dw = dilation
    Ho = (H + 2 * ph - (dh * (kh - 1) + 1)) // sh + 1
    Wo = (W + 2 * pw - (dw * (kw - 1) + 1)) // sw + 1
    if Cin != Cg * groups:
        raise RuntimeError("Input shape and kernel channels wont match: (%d vs %d)." % (Cin, Cg * groups))
    if grad_out.shape != (B, Cout, Ho, Wo):   # modulated_deform_conv_cuda.cu:185-192
        raise RuntimeError("Input shape and grad_out shape wont match: (%s vs %s)." % ((B, Cout, Ho, Wo), tuple(grad_out.shape)))
    if offset.shape != (B, deformable_groups * 2 * kh * kw, Ho, Wo):
        raise RuntimeError("offset shape %s, expected %s" % (tuple(offset.shape),
                                                             (B, deformable_groups * 2 * kh * kw, Ho, Wo)))
    n_in, n_off, n_mask, n_w, n_b = needs
    n_mask = n_mask and mask is not None
    g_in = torch.empty_like(inp) if n_in else None
    g_off = torch.empty_like(offset) if n_off else None
    g_mask = torch.empty_like(mask) if n_mask else None
    g_w = torch.empty_like(weight) if n_w else None
    g_b = torch.empty(Cout, device=inp.device, dtype=dt) if n_b else None
    with torch.cuda.device(inp.device):
        if dt == torch.float64:
            _lib.call("sf_dcn_backward_f64", _ptr(inp), _ptr(weight), _ptr(offset), _ptr(mask), _ptr(grad_out), _ptr(g_in),
                      _ptr(g_w), _ptr(g_b), _ptr(g_off), _ptr(g_mask), B, Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, dh, dw,
                      int(groups), int(deformable_groups), _stream())
            return g_in, g_off, g_mask, g_w, g_b
        algo = _algo(algo)
        nws = int(_lib.lib().sf_dcn_backward_workspace_bytes(B, Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, dh, dw, int(groups),
                                                             int(deformable_groups))) \
            if algo not in (SF_ALGO_GENERIC, SF_ALGO_EXACT) else 0
        ws = torch.empty(nws, device=inp.device, dtype=torch.uint8) if nws else None   # channels-last copies + column gradients (scratch)
        _lib.call("sf_dcn_backward_ws_f32", _ptr(inp), _ptr(weight), _ptr(offset), _ptr(mask), _ptr(grad_out), _ptr(g_in),
                  _ptr(g_w), _ptr(g_b), _ptr(g_off), _ptr(g_mask), B, Cin, H, W, Cout, kh, kw, sh, sw, ph, pw, dh, dw,
                  int(groups), int(deformable_groups), _ptr(ws), ctypes.c_size_t(nws), algo if nws or algo != SF_ALGO_AUTO else SF_ALGO_GENERIC,
                  _stream())
    return g_in, g_off, g_mask, g_w, g_b


def epipolar_attention(q, k, v, band, return_scores=False, algo=None, out=None):
    """Band-restricted epipolar row cross-attention, bf16 [B,C,H,W] in/out (spec-derived)."""
    q, k, v = (_chk(t, n, dtype=torch.bfloat16, dim=4) for t, n in ((q, "q"), (k, "k"), (v, "v")))
    if not (q.shape == k.shape == v.shape):
        raise ValueError("q, k, v must share one shape")
    B, C, H, W = q.shape
    if out is None:
        out = torch.empty_like(q)
    elif out.shape != q.shape or out.dtype != q.dtype or not out.is_contiguous():
        raise ValueError("out must be a contiguous bf16 tensor of q's shape")
    sc = torch.empty((B, band, H, W), device=q.device, dtype=torch.float32) if return_scores else None
    with torch.cuda.device(q.device):
        _lib.call("sf_epipolar_attn_bf16", _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(sc), B, C, H, W, int(band),
                  _algo(algo), _stream())
    return (out, sc) if return_scores else out


def upsample_convex(disp, mask, factor=8, mask_scale=1.0, out=None):
    """Convex upsampling; ``mask_scale`` multiplies the mask first (the `.25 *` of Attempts/update.py:273)."""
    if _ext is not None and out is None and _plain(disp, mask) and disp.dim() == 4 and mask.dim() == 4:
        return _ext.upsample_convex(disp, mask, int(factor), float(mask_scale))
    disp, mask = _chk(disp, "disp", dim=4), _chk(mask, "mask", dim=4)
    N, _, H, W = disp.shape
    if mask.shape != (N, 9 * factor * factor, H, W):
        raise ValueError("mask must be [N,%d,H,W]" % (9 * factor * factor))
    if out is None:
        out = torch.empty((N, 1, factor * H, factor * W), device=disp.device, dtype=torch.float32)
    elif out.shape != (N, 1, factor * H, factor * W) or out.dtype != torch.float32 or not out.is_contiguous():
        raise ValueError("out must be a contiguous fp32 [N,1,F*H,F*W] tensor")
    with torch.cuda.device(disp.device):
        if mask_scale == 1.0:
            _lib.call("sf_upsample_convex_f32", _ptr(disp), _ptr(mask), _ptr(out), N, H, W, int(factor), _stream())
        else:
            _lib.call("sf_upsample_convex_scaled_f32", _ptr(disp), _ptr(mask), float(mask_scale), _ptr(out), N, H, W,
                      int(factor), _stream())
    return out


def upsample_simple(disp, factor=8):
    """upsample_simple8: factor * bilinear(align_corners=True) resize by ``factor`` -> [N,1,F*H,F*W]."""
    disp = _chk(disp, "disp", dim=4)
    N, C, H, W = disp.shape
    out = torch.empty((N, C, factor * H, factor * W), device=disp.device, dtype=torch.float32)
    with torch.cuda.device(disp.device):
        _lib.call("sf_upsample_simple_f32", _ptr(disp), _ptr(out), N * C, H, W, int(factor), _stream())
    return out


def upsample_simple_backward(grad_out, factor=8):
    grad_out = _chk(grad_out, "grad_out", dim=4)
    N, C, Ho, Wo = grad_out.shape
    gd = torch.empty((N, C, Ho // factor, Wo // factor), device=grad_out.device, dtype=torch.float32)
    with torch.cuda.device(grad_out.device):
        _lib.call("sf_upsample_simple_backward_f32", _ptr(grad_out), _ptr(gd), N * C, Ho // factor, Wo // factor,
                  int(factor), _stream())
    return gd


def upsample_convex_rows(disp_src, mask_src, y_first, rows, factor=8, mask_scale=1.0, peers=None, peer_offset=0,
                         peer_batch_stride=0):
    """Convex upsampling of the low-resolution rows [y_first, y_first + rows) of tensors that hold a band's SOURCE
    rows (band + halo) -> [N,1,rows*F,W*F].  peers: device pointers of every rank's whole-map buffer (NVLink peer
    mappings, dist.PeerGather); each result is also stored at element ``peer_offset`` + n*``peer_batch_stride`` +
    its row-major index inside the band of each of them, so the gather needs no collective."""
    disp_src, mask_src = _chk(disp_src, "disp", dim=4), _chk(mask_src, "mask", dim=4)
    N, _, H, W = disp_src.shape
    if mask_src.shape != (N, 9 * factor * factor, H, W):
        raise ValueError("mask must be [N,%d,H,W]" % (9 * factor * factor))
    out = torch.empty((N, 1, factor * rows, factor * W), device=disp_src.device, dtype=torch.float32)
    peers = list(peers or [])
    arr = (ctypes.c_void_p * max(len(peers), 1))(*[ctypes.c_void_p(int(q)) for q in peers])
    with torch.cuda.device(disp_src.device):
        _lib.call("sf_upsample_convex_gather_f32", _ptr(disp_src), _ptr(mask_src), float(mask_scale), _ptr(out), N, H,
                  W, int(factor), int(y_first), int(rows), arr if peers else None, len(peers),
                  ctypes.c_longlong(int(peer_offset)), ctypes.c_longlong(int(peer_batch_stride)), _stream())
    return out


def _gate_view(t, name, B, C, H, W):
    """A [B,C,H,W] gate pre-activation that may be a channel slice of a wider contiguous tensor -> (tensor, stride)."""
    if not torch.is_tensor(t) or not t.is_cuda:
        raise RuntimeError("Not implemented on the CPU (%s)" % name)
    if t.dtype != torch.float32 or t.shape != (B, C, H, W):
        raise ValueError("%s must be fp32 [%d,%d,%d,%d], got %s %s" % (name, B, C, H, W, t.dtype, tuple(t.shape)))
    if t.stride()[1:] != (H * W, W, 1) or (B > 1 and t.stride(0) < C * H * W):
        t = t.contiguous()
    return t, (t.stride(0) if B > 1 else C * H * W)


def gru_reset(b_pre, prev_h, inp, out=None):
    """ConvGRU reset path: cat(sigmoid(b_pre) * prev_h, input) -> [B,C+Cx,H,W]  (Attempts/update.py:43-46)."""
    inp = _chk(inp, "input", dim=4)
    B, Cx, H, W = inp.shape
    C = b_pre.shape[1]
    b_pre, bs = _gate_view(b_pre, "b_pre", B, C, H, W)
    if prev_h is not None:
        prev_h = _chk(prev_h, "prev_h", dim=4)
        if prev_h.shape != (B, C, H, W):
            raise ValueError("prev_h must be [B,C,H,W]")
    if out is None:
        out = torch.empty((B, C + Cx, H, W), device=inp.device, dtype=torch.float32)
    with torch.cuda.device(inp.device):
        _lib.call("sf_gru_reset_f32", _ptr(b_pre), ctypes.c_longlong(bs), _ptr(prev_h), _ptr(inp), _ptr(out), B, C, Cx,
                  H, W, _stream())
    return out


def gru_blend(z_pre, g_pre, prev_h, softsign=False, out=None):
    """ConvGRU output: (1 - sigmoid(z_pre)) * prev_h + sigmoid(z_pre) * tanh(g_pre)  (Attempts/update.py:47-48)."""
    g_pre = _chk(g_pre, "g_pre", dim=4)
    B, C, H, W = g_pre.shape
    z_pre, zs = _gate_view(z_pre, "z_pre", B, C, H, W)
    if prev_h is not None:
        prev_h = _chk(prev_h, "prev_h", dim=4)
        if prev_h.shape != (B, C, H, W):
            raise ValueError("prev_h must be [B,C,H,W]")
    if out is None:
        out = torch.empty_like(g_pre)
    with torch.cuda.device(g_pre.device):
        _lib.call("sf_gru_blend_f32", _ptr(z_pre), ctypes.c_longlong(zs), _ptr(g_pre), _ptr(prev_h), _ptr(out), B, C, H, W,
                  int(bool(softsign)), _stream())
    return out
