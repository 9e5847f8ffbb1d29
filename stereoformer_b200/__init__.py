"""stereoformer_b200 -- B200-native (sm_100a) implementation of StereoFormer's stereo-matching hot path.

The compute lives in hand-written CUDA kernels behind a C ABI (include/stereoformer_b200.h,
libsfb200.so); this package is the thin PyTorch-facing mirror of the reference's operator interface.
There is no CPU fallback: operators raise if the library is missing or a tensor is not on a GPU.
"""
from . import _lib  # noqa: F401
from . import ops  # noqa: F401
from .ops import get_precision, precision, set_precision  # noqa: F401
from .modules import (CorrBlock1D, CostVolume, DisparityEstimation, DisparityEstimationWithProb,  # noqa: F401
                      EpipolarCrossAttention, LocalCostVolume, LRwarp_error, PytorchAlternateCorrBlock1D,
                      Variance_Based_Local_Cost_Volume,
                      build_corr, build_cost_volume_from_volume, build_gwc_volume, disp_warp, dlc_window,
                      guidance_inputs, upsample_convex8, upsample_simple8)
from .dcn_modules import (DeformConv, DeformConvPack, ModulatedDeformConv, ModulatedDeformConvPack,  # noqa: F401
                          ModulatedDeformNormal)

from .update import (BasicGuidanceNet, ConvGRU, DisparityUpdateDLC, DisparityUpdateDLCWithMask,  # noqa: F401
                     DlcGruRefiner, OffsetPredictionHead)

__version__ = "0.1.0"
