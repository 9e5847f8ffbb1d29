"""install.py against the unmodified reference (oracle/_ref, vendored by __graft_entry__.build()): structural checks
that need no GPU.  The numerical full-model checks are in test_gpu_fullmodel.py."""
import importlib
import subprocess
import sys
import os

import pytest
import torch

from oracle import refenv

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(refenv.ref_root() is None, reason="no copy of the reference reachable")

VARIANTS = ["baseline", "baseline_var", "baseline_dynamic", "baseline_dynamic_supervised", "baseline_dynamic_gru",
            "baseline_dynamic_gru2"]

SCRIPT = r"""
import sys, warnings
warnings.filterwarnings("ignore")
sys.path.insert(0, %(root)r)
import torch
from oracle import refenv
refenv.activate()
import stereoformer_b200 as sf
import stereoformer_b200.install as sfi
order = %(order)r
mods = {}
def imp():
    import importlib
    for v in %(variants)r:
        mods[v] = importlib.import_module("models.LocalCostVolume." + v)
if order == "import_first":
    imp()
    nets = {v: m.LowCNN(cost_volume_type="correlation", upsample_type="convex", adaptive_refinement=True).eval()
            for v, m in mods.items()}
sfi.install_dcn()
rep = sfi.install_reference_ops(strict=True)
assert not rep["skipped"], rep["skipped"]
if order == "install_first":
    imp()
    nets = {v: m.LowCNN(cost_volume_type="correlation", upsample_type="convex", adaptive_refinement=True).eval()
            for v, m in mods.items()}
import functions.modulated_deform_conv_func as f
assert f.DCN is sf.DCN
for v, net in nets.items():
    assert type(net.low_scale_cost_volume).forward is sf.CostVolume.forward, v
    assert type(net.disp_estimation3).forward is sf.DisparityEstimation.forward, v
    lc = type(net.local_cost_volume)
    want = {"baseline": sf.LocalCostVolume, "baseline_var": sf.Variance_Based_Local_Cost_Volume,
            "baseline_dynamic": sf.update.PureLearningLocalCostVolume,
            "baseline_dynamic_supervised": sf.update.PureLearningLocalCostVolumeWithBounds,
            "baseline_dynamic_gru": sf.update.DisparityUpdateDLCWithMask,
            "baseline_dynamic_gru2": sf.update.DisparityUpdateDLCWithMaskV2}[v]
    assert lc.forward is want.forward, (v, lc)
    assert mods[v].upsample_convex8 is sf.upsample_convex8, v
    # the drop-ins have no CPU path: a CPU forward must fail loudly, not fall back
    x = torch.randn(1, 3, 64, 128)
    try:
        with torch.no_grad():
            net(x, x)
    except RuntimeError as e:
        assert "CPU" in str(e), e
    else:
        raise AssertionError("CPU forward did not raise with the drop-ins installed: " + v)
sfi.uninstall_reference_ops()
for v, net in nets.items():
    assert type(net.low_scale_cost_volume).forward is not sf.CostVolume.forward
    assert mods[v].upsample_convex8 is not sf.upsample_convex8
with torch.no_grad():
    out = nets["baseline"](torch.randn(1, 3, 64, 128), torch.randn(1, 3, 64, 128))
assert out[1].shape == (1, 1, 64, 128)
print("OK", len(rep["grafted"]), len(rep["rebound"]))
"""


@pytest.mark.parametrize("order", ["install_first", "import_first"])
def test_install_both_orders(order):
    """installing after `import models.LocalCostVolume.baseline` used to break LowCNN()."""
    code = SCRIPT % {"root": ROOT, "order": order, "variants": VARIANTS}
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.strip().splitlines()[-1].startswith("OK")


def test_reference_dcn_extension_is_the_reference():
    """oracle/_ref/native/DCN.so exports the six pybind functions of deform/src/vision.cpp:5-12."""
    if not refenv.available()["dcn"]:
        pytest.skip("reference DCN extension not built")
    code = ("import sys; sys.path.insert(0, %r); from oracle import refenv; refenv.activate(); import DCN; "
            "print(sorted(n for n in dir(DCN) if not n.startswith('_')))" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "modulated_deform_conv_forward" in r.stdout and "deform_psroi_pooling_backward" in r.stdout
