"""Install the drop-ins into an importable copy of the reference (INTEGRATION.md shows the same as a patch).

    import stereoformer_b200.install as sfi
    sfi.install_dcn()                 # `import DCN` now resolves to stereoformer_b200.DCN
    sfi.install_reference_ops()       # rebinds CostVolume, DisparityEstimation, LocalCostVolume, ... in the
                                      # reference's modules (needs the reference on sys.path)

Nothing here is needed to use the operators directly; it only rebinds names in the reference's modules so
that LowCNN / RAFT_Stereo pick the B200 kernels up unchanged.
"""
import importlib
import sys

from . import DCN as _dcn
from . import modules as _m
from . import update as _u

# reference module -> {attribute: replacement}
PATCHES = {
    "models.utils.build_cost_volume": {"CostVolume": _m.CostVolume},
    "models.utils.cost": {"CostVolume": _m.CostVolume},
    "models.utils.estimation": {"DisparityEstimation": _m.DisparityEstimation,
                                "DisparityEstimationWithProb": _m.DisparityEstimationWithProb},
    "models.LocalCostVolume.Attempts.fixed_local_cost_volume": {
        "LocalCostVolume": _m.LocalCostVolume, "build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.variance_local_cost_volume": {
        "Variance_Based_Local_Cost_Volume": _m.Variance_Based_Local_Cost_Volume,
        "build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.update": {"build_cost_volume_from_volume": _m.build_cost_volume_from_volume,
                                               "disp_warp": _m.disp_warp, "ConvGRU": _u.ConvGRU,
                                               "BasicGuidanceNet": _u.BasicGuidanceNet,
                                               "OffsetPredictionHead": _u.OffsetPredictionHead,
                                               "DisparityUpdateDLC": _u.DisparityUpdateDLC,
                                               "DisparityUpdateDLCWithMask": _u.DisparityUpdateDLCWithMask},
    # LowCNN binds the class at import time (baseline_dynamic_gru.py:11)
    "models.LocalCostVolume.baseline_dynamic_gru": {"DisparityUpdateDLCWithMask": _u.DisparityUpdateDLCWithMask,
                                                    "upsample_convex8": _m.upsample_convex8},
    "models.LocalCostVolume.Attempts.updatev2": {"build_cost_volume_from_volume": _m.build_cost_volume_from_volume,
                                                 "disp_warp": _m.disp_warp},
    "models.LocalCostVolume.Attempts.pure_learning_method": {
        "build_cost_volume_from_volume": _m.build_cost_volume_from_volume, "disp_warp": _m.disp_warp},
    "models.CrossAttentionCostVolume.GWC_Cost_Volume": {"build_gwc_volume": _m.build_gwc_volume},
    "models.IterativeLocalCostVolume.corr.corr": {"CorrBlock1D": _m.CorrBlock1D,
                                                  "PytorchAlternateCorrBlock1D": _m.PytorchAlternateCorrBlock1D},
    "models.utils.upsample": {"upsample_convex8": _m.upsample_convex8, "upsample_simple8": _m.upsample_simple8},
    "utils.disparity_warper": {"disp_warp": _m.disp_warp, "LRwarp_error": _m.LRwarp_error},
}


def install_dcn():
    """Make ``import DCN`` (deform/functions/*_func.py:13) resolve to the B200 implementation."""
    sys.modules["DCN"] = _dcn
    return _dcn


def install_reference_ops(strict=False):
    """Rebind the hot-path names inside the reference's modules.  Returns the list of patched (module, attr)."""
    done = []
    for modname, attrs in PATCHES.items():
        try:
            mod = importlib.import_module(modname)
        except Exception:
            if strict:
                raise
            continue
        for name, repl in attrs.items():
            if hasattr(mod, name):
                setattr(mod, name, repl)
                done.append((modname, name))
    return done
