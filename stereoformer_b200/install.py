"""Install the drop-ins into an importable copy of the reference (INTEGRATION.md shows the same as a patch).

    import stereoformer_b200.install as sfi
    sfi.install_dcn()                 # `import DCN` now resolves to stereoformer_b200.DCN
    report = sfi.install_reference_ops()      # needs the reference on sys.path
    ...
    sfi.uninstall_reference_ops()     # back to the stock reference (used by the parity tests: same network object,
                                      # same weights, stock forward vs drop-in forward)

Order does not matter: it works before or after the reference's model modules are imported and also for networks
that are already constructed.  Two mechanisms, both reversible:

  * classes  -- the hot-path METHODS of the reference classes are replaced in place (``RefCostVolume.forward =
    stereoformer_b200.CostVolume.forward`` ...).  Class identity, constructor, parameter names and therefore state
    dicts are untouched, ``super(CostVolume, self)`` inside the reference keeps resolving, and existing instances
    (``net.low_scale_cost_volume``) pick the kernels up as well.
  * functions / the ``DCN`` extension module -- every module in ``sys.modules`` that holds the original object
    (``from utils.disparity_warper import disp_warp`` binds at import time) gets the name rebound.

Nothing here is needed to use the operators directly.
"""
import importlib
import sys
import types
import warnings

from . import DCN as _dcn
from . import modules as _m
from . import update as _u

# reference module -> {class name: (our class, methods to graft)}
CLASS_GRAFTS = {
    "models.utils.build_cost_volume": {"CostVolume": (_m.CostVolume, ("forward",))},
    "models.utils.cost": {"CostVolume": (_m.CostVolume, ("forward",))},
    "models.utils.estimation": {"DisparityEstimation": (_m.DisparityEstimation, ("forward",)),
                                "DisparityEstimationWithProb": (_m.DisparityEstimationWithProb, ("forward",))},
    "models.LocalCostVolume.Attempts.fixed_local_cost_volume": {
        "LocalCostVolume": (_m.LocalCostVolume, ("forward",))},
    "models.LocalCostVolume.Attempts.variance_local_cost_volume": {
        "Variance_Based_Local_Cost_Volume": (_m.Variance_Based_Local_Cost_Volume, ("forward",))},
    "models.LocalCostVolume.Attempts.pure_learning_method": {
        "PureLearningLocalCostVolume": (_u.PureLearningLocalCostVolume, ("forward",))},
    "models.LocalCostVolume.Attempts.pure_learning_method_w_super": {
        "PureLearningLocalCostVolume": (_u.PureLearningLocalCostVolumeWithBounds, ("forward", "with_bounds"))},
    "models.LocalCostVolume.Attempts.update": {
        "ConvGRU": (_u.ConvGRU, ("forward",)),
        "BasicGuidanceNet": (_u.BasicGuidanceNet, ("forward",)),
        "DisparityUpdateDLC": (_u.DisparityUpdateDLC, ("forward",)),
        "DisparityUpdateDLCWithMask": (_u.DisparityUpdateDLCWithMask, ("forward",))},
    "models.LocalCostVolume.Attempts.updatev2": {
        "ConvGRU": (_u.ConvGRU, ("forward",)),
        "BasicGuidanceNet": (_u.BasicGuidanceNet, ("forward",)),
        "DisparityUpdateDLC": (_u.DisparityUpdateDLC, ("forward",)),
        "DisparityUpdateDLCWithMask": (_u.DisparityUpdateDLCWithMaskV2, ("forward",))},
    "models.IterativeLocalCostVolume.corr.corr": {
        "CorrBlock1D": (_m.CorrBlock1D, ("__init__", "__call__", "corr")),
        "PytorchAlternateCorrBlock1D": (_m.PytorchAlternateCorrBlock1D, ("__init__", "__call__"))},
}

# reference module -> {function name: replacement}; rebound in every module that imported the original
FUNCTION_PATCHES = {
    "utils.disparity_warper": {"disp_warp": _m.disp_warp, "LRwarp_error": _m.LRwarp_error},
    "models.utils.disparity_warper": {"disp_warp": _m.disp_warp, "LRwarp_error": _m.LRwarp_error},
    "models.utils.upsample": {"upsample_convex8": _m.upsample_convex8, "upsample_simple8": _m.upsample_simple8},
    "models.utils.disp_residual": {"build_corr": _m.build_corr, "upsample_convex8": _m.upsample_convex8,
                                   "upsample_simple8": _m.upsample_simple8},
    "models.CrossAttentionCostVolume.GWC_Cost_Volume": {"build_gwc_volume": _m.build_gwc_volume},
    "models.utils.submodule": {"build_gwc_volume": _m.build_gwc_volume},
    "models.LocalCostVolume.Attempts.fixed_local_cost_volume": {
        "build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.variance_local_cost_volume": {
        "build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.pure_learning_method": {
        "build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.pure_learning_method_w_super": {
        "build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.update": {"build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
    "models.LocalCostVolume.Attempts.updatev2": {"build_cost_volume_from_volume": _m.build_cost_volume_from_volume},
}

# modules of the reference that bind the pybind extension by `import DCN` (deform/functions/*_func.py:13); they are
# importable as `functions.x` (deform/ on sys.path, what deform/modules/*.py:12 does) or as `deform.functions.x`
DCN_IMPORTERS = ("functions.modulated_deform_conv_func", "functions.deform_conv_func",
                 "deform.functions.modulated_deform_conv_func", "deform.functions.deform_conv_func")

_MISSING = object()
_undo = []          # (kind, holder, name, original) in application order
_reverse = {}       # id(replacement function) -> (replacement, original), for modules imported after the install
_prev_dcn = None    # what `import DCN` resolved to before install_dcn() (the reference's own extension, if built)


def _raw(cls, name):
    """The attribute as stored on the class (keeps staticmethod / classmethod wrappers)."""
    for k in cls.__mro__:
        if name in vars(k):
            return vars(k)[name]
    raise AttributeError("%s.%s" % (cls.__name__, name))


def _set(kind, holder, name, value):
    old = vars(holder).get(name, _MISSING)
    _undo.append((kind, holder, name, old))
    setattr(holder, name, value)


def install_dcn():
    """Make ``import DCN`` (deform/functions/*_func.py:13) resolve to the B200 implementation; modules that already
    imported the reference's extension get their ``DCN`` name rebound."""
    global _prev_dcn
    if sys.modules.get("DCN") is not _dcn:
        _prev_dcn = sys.modules.get("DCN")
    sys.modules["DCN"] = _dcn
    done = []
    for name in DCN_IMPORTERS:
        mod = sys.modules.get(name)
        if mod is not None and getattr(mod, "DCN", None) is not _dcn:
            _set("dcn", mod, "DCN", _dcn)
            done.append(name)
    return done


def install_reference_ops(strict=False):
    """Route the reference's hot path through the B200 kernels.  Returns a report
    ``{"grafted": [(module, class, method)], "rebound": [(holder module, name)], "skipped": [module]}``;
    ``skipped`` lists reference modules that could not be imported (strict=True raises instead)."""
    if _undo and any(k != "dcn" for k, *_ in _undo):
        uninstall_reference_ops(keep_dcn=True)
    report = {"grafted": [], "rebound": [], "skipped": []}

    def load(modname):
        try:
            return importlib.import_module(modname)
        except Exception:
            if strict:
                raise
            if modname not in report["skipped"]:
                report["skipped"].append(modname)
            return None

    for modname, classes in CLASS_GRAFTS.items():
        mod = load(modname)
        if mod is None:
            continue
        for cname, (ours, methods) in classes.items():
            ref = getattr(mod, cname, None)
            if not isinstance(ref, type) or ref is ours:
                continue
            for meth in methods:
                _set("graft", ref, meth, _raw(ours, meth))
                report["grafted"].append((modname, cname, meth))

    originals = {}                                      # id(original function) -> replacement
    for modname, funcs in FUNCTION_PATCHES.items():
        mod = load(modname)
        if mod is None:
            continue
        for fname, repl in funcs.items():
            orig = getattr(mod, fname, None)
            if isinstance(orig, types.FunctionType) and orig is not repl:
                originals[id(orig)] = (orig, repl)
    for orig, repl in originals.values():
        _reverse.setdefault(id(repl), (repl, orig))
    report["rebound"] = _rebind_everywhere(originals, record=True)
    return report


def _rebind_everywhere(table, record):
    """table: id(function) -> (function, replacement).  Rebinds every module-level name that holds `function`."""
    done = []
    if not table:
        return done
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")              # lazy-module proxies may warn when inspected
        for holder_name, holder in list(sys.modules.items()):
            if not isinstance(holder, types.ModuleType) or holder_name.startswith("stereoformer_b200"):
                continue
            try:
                items = list(vars(holder).items())
            except Exception:
                continue
            for name, val in items:
                if type(val) is not types.FunctionType:
                    continue
                hit = table.get(id(val))
                if hit is not None and hit[0] is val:
                    if record:
                        _set("rebind", holder, name, hit[1])
                    else:
                        setattr(holder, name, hit[1])
                    done.append((holder_name, name))
    return done


def uninstall_reference_ops(keep_dcn=False):
    """Undo install_reference_ops() (and install_dcn() unless keep_dcn): the reference runs stock again."""
    kept = []
    while _undo:
        kind, holder, name, old = _undo.pop()
        if keep_dcn and kind == "dcn":
            kept.append((kind, holder, name, old))
            continue
        if old is _MISSING:
            try:
                delattr(holder, name)
            except AttributeError:
                pass
        else:
            setattr(holder, name, old)
    _undo.extend(reversed(kept))
    _rebind_everywhere(dict(_reverse), record=False)     # modules that imported the replacements after the install
    _reverse.clear()
    if not keep_dcn and sys.modules.get("DCN") is _dcn:
        if _prev_dcn is not None:
            sys.modules["DCN"] = _prev_dcn
        else:
            del sys.modules["DCN"]


def is_installed():
    return any(k != "dcn" for k, *_ in _undo)
