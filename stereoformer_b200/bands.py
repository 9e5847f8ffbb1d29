"""Row-band execution of the hot path for ONE large stereo pair (SURVEY 8e partition 2; BASELINE config 5:
Middlebury-shaped 1984x2880, maxdisp 384 -> 248x360 grid, D = 48, split across 2/4/8 GPUs).

What each operator needs from the neighbouring bands:

* cost volume, soft-argmin, local / variance / DLC windows, ``build_cost_volume_from_volume``, CorrBlock1D and the
  epipolar attention read and write only epipolar row ``y``: a rank runs the whole-image kernels on its rows,
  nothing else (``lcv_band``).  The results are the rows of the whole-image call bit for bit.
* ``disp_warp`` / the guidance stage sample rows ``y-1 .. y+1`` (the reference normalises with ``H-1`` but samples
  with ``align_corners=False``, utils/disparity_warper.py:96-100, so ``y_src = y*H/(H-1) - 0.5`` depends on the
  GLOBAL row index and height): 1-row halo and the band entry points ``sf_disp_warp_band_f32`` /
  ``sf_guidance_band_f32``, bit-exact.
* convex upsampling reads the 3x3 neighbourhood of the low-resolution disparity (models/utils/upsample.py:26-31):
  1-row halo, run on the haloed rows and crop.
* DCN (deform/src/cuda/modulated_deform_im2col_cuda.cuh:140-193) samples ``y - pad + i*dil + dy`` with learned,
  data-dependent ``dy``: the halo is ``pad + ceil(max |dy|) + 1`` rows (``dcn_halo_rows``), checked per call.  The
  local row index differs from the global one by an integer, so ``y + dy`` rounds differently in fp32: equal to the
  whole-image call to ~1e-5 (ulp(248) = 1.5e-5 px of sampling coordinate), not bit for bit.

The cuDNN trunk either side (its receptive-field halo) is outside the path; a rank is handed feature / image rows
``[ys0, ys1)`` that already include the halo.  The only exchange is the gather of the band's disparity rows
(``dist.gather_row_bands``).
"""
import math

import torch

from . import ops
from ._lib import SF_ALGO_AUTO, SF_DIV_RECIP, SF_WIN_FIXED


class RowBand:
    """Rows [y0, y1) of an image of ``height`` rows, held with halo as source rows [ys0, ys1)."""

    def __init__(self, height, y0, y1, ys0=None, ys1=None):
        self.height, self.y0, self.y1 = int(height), int(y0), int(y1)
        self.ys0 = self.y0 if ys0 is None else int(ys0)
        self.ys1 = self.y1 if ys1 is None else int(ys1)
        if not (0 <= self.ys0 <= self.y0 < self.y1 <= self.ys1 <= self.height):
            raise ValueError("need 0 <= ys0 <= y0 < y1 <= ys1 <= height, got %r" % ((self.ys0, self.y0, self.y1,
                                                                                      self.ys1, self.height),))

    @classmethod
    def of_rank(cls, height, world, rank, align=1, halo=0):
        from .dist import row_bands
        return cls(height, *row_bands(height, world, align=align, halo=halo)[rank])

    @property
    def rows(self):
        return self.y1 - self.y0

    def halo_top(self):
        return self.y0 - self.ys0

    def halo_bottom(self):
        return self.ys1 - self.y1

    def covers(self, halo):
        """True when the source rows hold ``halo`` rows either side of the band (or reach the image border)."""
        return ((self.halo_top() >= halo or self.ys0 == 0) and
                (self.halo_bottom() >= halo or self.ys1 == self.height))

    def source(self, full):
        """Rows [ys0, ys1) of a whole-image NCHW tensor, contiguous (what a rank holds)."""
        return full[:, :, self.ys0:self.ys1].contiguous()

    def interior(self, full):
        return full[:, :, self.y0:self.y1].contiguous()

    def crop(self, src):
        """Band rows of a tensor laid out on the source rows."""
        return src[:, :, self.halo_top():self.halo_top() + self.rows]

    def scaled(self, factor):
        return RowBand(self.height * factor, self.y0 * factor, self.y1 * factor, self.ys0 * factor, self.ys1 * factor)


def lcv_band(left_rows, right_rows, max_disp, radius=2.0, n_points=20, consider_valid=False, div_mode=SF_DIV_RECIP,
             algo=None, peers=None):
    """LocalCostVolume hot path (a1 + a3 + a4) on a band's own rows: [B,C,rows,W] features -> (cv, disp).
    Row-local: no halo, same kernels as the whole image."""
    cv = ops.corr_volume(left_rows, right_rows, max_disp, algo=algo)
    disp = ops.local_window(cv, None, mode=SF_WIN_FIXED, radius=radius, n_points=n_points,
                            consider_valid=consider_valid, div_mode=div_mode, peers=peers)
    return cv, disp


def disp_warp_band(img_src, disp_rows, band, padding_mode="border", **kw):
    """disp_warp on ``band``: ``img_src`` holds the band's source rows (>= 1-row halo), ``disp_rows`` its rows."""
    if img_src.shape[2] != band.ys1 - band.ys0 or disp_rows.shape[2] != band.rows:
        raise ValueError("img_src / disp_rows do not match the band's row counts")
    return ops.disp_warp_band(img_src, disp_rows, band.height, band.y0, band.ys0, padding_mode=padding_mode, **kw)


def guidance_band(cv_rows, disp_rows, left_rows, right_src, band, **kw):
    if right_src.shape[2] != band.ys1 - band.ys0 or cv_rows.shape[2] != band.rows:
        raise ValueError("right_src / cv_rows do not match the band's row counts")
    return ops.guidance_band(cv_rows, disp_rows, left_rows, right_src, band.height, band.y0, band.ys0, **kw)


def upsample_convex_band(disp_src, mask_src, band, factor=8, mask_scale=1.0, gather=None, root=None):
    """Convex upsampling of a band: disparity and mask on the source rows (1-row halo); returns the band's
    ``rows*factor`` output rows (the halo rows are not computed).  ``gather``: a ``dist.PeerGather`` over whole-map
    buffers [N,1,height*factor,W*factor]; the band's rows are then also stored into every rank's buffer by the same
    kernel (NVLink peer stores), which replaces the all-gather of the full-resolution map.  ``root``: store into that
    rank's buffer only (the gather-to-device-0 of the reference's ``nn.DataParallel``, trainer_DLC_GRU.py:102-106):
    1/world of the peer traffic."""
    if not band.covers(1):
        raise ValueError("convex upsampling needs a 1-row halo")
    if disp_src.shape[2] != band.ys1 - band.ys0:
        raise ValueError("disp_src does not match the band's source rows")
    N, W = disp_src.shape[0], disp_src.shape[3]
    peers, off, bstride = None, 0, band.height * factor * W * factor
    if gather is not None:
        if gather.numel_total < N * bstride:
            raise ValueError("gather buffer must hold [N,1,height*factor,W*factor] floats")
        peers = gather.peer_ptrs if root is None else [gather.peer_ptrs[int(root)]]
        off = band.y0 * factor * W * factor
    return ops.upsample_convex_rows(disp_src, mask_src, band.halo_top(), band.rows, factor=factor,
                                    mask_scale=mask_scale, peers=peers, peer_offset=off, peer_batch_stride=bstride)


def dcn_halo_rows(offset, kernel_h, padding_h, dilation_h):
    """Rows of halo a band needs for this offset field: pad + ceil(max |dy|) + 1 (one device -> host read).
    The dy channels are the even ones of each deformable group (modulated_deform_im2col_cuda.cuh:171-172)."""
    if padding_h * 2 != dilation_h * (kernel_h - 1):
        raise ValueError("band DCN supports 'same' convolutions only (2*pad == dil*(k-1))")
    dy_max = float(offset[:, 0::2].abs().amax().item()) if offset.numel() else 0.0
    if not math.isfinite(dy_max):
        raise ValueError("non-finite offsets")
    return int(padding_h) + int(math.ceil(dy_max)) + 1


def dcn_band(inp_src, weight, bias, offset_src, mask_src, band, stride=(1, 1), padding=(1, 1), dilation=(1, 1),
             groups=1, deformable_groups=1, halo=None, algo=None):
    """DCN forward on a row band.  All of input / offset / mask are given on the source rows [ys0, ys1) (the offset
    convolution of the reference's DCN pack runs on the same haloed input); the result is cropped to the band.
    Raises ``ValueError`` naming the halo the offsets need when the band's is too small."""
    if tuple(stride) != (1, 1):
        raise ValueError("band DCN supports stride 1 only")
    kh = weight.shape[2]
    off_rows = band.crop(offset_src)
    need = dcn_halo_rows(off_rows, kh, padding[0], dilation[0]) if halo is None else int(halo)
    if not band.covers(need):
        raise ValueError("DCN offsets reach %d rows beyond the band, the source rows give %d / %d"
                         % (need, band.halo_top(), band.halo_bottom()))
    out = ops.dcn_forward(inp_src, weight, bias, offset_src, mask_src, stride, padding, dilation, groups,
                          deformable_groups, algo=algo)
    return band.crop(out).contiguous()
