"""The reference's own LowCNN networks (unmodified, from oracle/_ref or /root/reference) as the CALLER of the drop-ins:
stock forward (PyTorch eager + the reference's compiled DCN extension) vs the same network object with
stereoformer_b200 installed.  Used by tests/test_gpu_fullmodel.py and by bench.py's full-model entries.
"""
import importlib
import os
import sys
import time
import warnings

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

VARIANTS = {
    # name: (module under models.LocalCostVolume, forward kwargs)
    "lc": ("baseline", {}),                                   # fixed window r=2, N=20 (BASELINE configs[1])
    "var": ("baseline_var", {}),                              # variance window
    "dlc": ("baseline_dynamic", {}),                          # SmallUNet bounds (DCN 16->16 inside)
    "dlc_super": ("baseline_dynamic_supervised", {}),
    "dlc_gru": ("baseline_dynamic_gru", {"iters": 12}),       # BASELINE configs[2]: 12 refinement iterations
    "dlc_gru2": ("baseline_dynamic_gru2", {"iters": 12}),
}


def activate_reference():
    from oracle import refenv
    root = refenv.activate(dcn="reference")
    if root is None:
        raise RuntimeError("no copy of the reference reachable (run __graft_entry__.build() in the build container)")
    return root


def build_net(variant, device, seed=1024, d8=None):
    """LowCNN of the given variant with the reference's initialisers under `seed`, eval mode.
    d8: disparity levels at 1/8 scale if not the hard-coded 192//8 (config 5: 384//8 = 48, SURVEY 7 'D is baked')."""
    activate_reference()
    modname, _ = VARIANTS[variant]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        mod = importlib.import_module("models.LocalCostVolume." + modname)
        torch.manual_seed(seed)
        if d8 is None or d8 == 24:
            net = mod.LowCNN(cost_volume_type="correlation", upsample_type="convex", adaptive_refinement=True)
        else:
            net = _build_wide(mod, d8)
    return net.to(device).eval()


def _build_wide(mod, d8):
    """The reference hard-codes 192//8 levels (baseline.py:56-66); config 5 (maxdisp 384) needs 48.  Builds the same
    network and swaps the four D-dependent members for ones constructed with the reference's own classes."""
    import torch.nn as nn
    net = mod.LowCNN(max_disp=8 * d8, cost_volume_type="correlation", upsample_type="convex",
                     adaptive_refinement=True)
    ResBlock = importlib.import_module("models.BasicBlocks.resnet").ResBlock
    net.correlation_aggreagtion = nn.Sequential(ResBlock(d8, d8, 3, 1), ResBlock(d8, d8, 3, 1), ResBlock(d8, d8, 3, 1))
    net.low_scale_cost_volume = type(net.low_scale_cost_volume)(max_disp=d8, feature_similarity="correlation")
    net.disp_estimation3 = type(net.disp_estimation3)(max_disp=d8, match_similarity=True)
    lc = net.local_cost_volume
    if hasattr(lc, "offset_prediction_network"):
        lc.offset_prediction_network = type(lc.offset_prediction_network)(input_channels=d8, hidden_layer=32)
    for m in net.correlation_aggreagtion.modules():
        if isinstance(m, nn.Conv2d):
            nn.init.kaiming_normal_(m.weight.data)
            m.bias.data.zero_()
    return net


def forward(net, variant, left, right):
    """-> the final full-resolution disparity [B,1,H,W] (last refinement for the GRU variants)."""
    _, kw = VARIANTS[variant]
    with torch.no_grad():
        if "iters" in kw:
            out = net(left, right, kw["iters"], False)
            return out[-1]
        out = net(left, right, False)
        if isinstance(out, tuple):          # dlc_super: ([pr0, pr1], [lower, upper], disp3)
            out = out[0]
        return out[-1] if isinstance(out, (list, tuple)) else out


class dropins:
    """Context manager: the reference runs on the stereoformer_b200 kernels inside, stock outside."""

    def __enter__(self):
        import stereoformer_b200.install as sfi
        sfi.install_dcn()
        self.report = sfi.install_reference_ops()
        return self.report

    def __exit__(self, *exc):
        import stereoformer_b200.install as sfi
        sfi.uninstall_reference_ops()
        return False


def _stats(err):
    err = err.abs().flatten()
    return {"epe": float(err.mean()), "max": float(err.max()),
            "p99": float(torch.quantile(err[:: max(1, err.numel() // 1000000)], 0.99)),
            "frac_over_1e-3": float((err > 1e-3).float().mean())}


def compare(variant, batch, height, width, device, seed=1024, d8=None, precisions=("fast",), sensitivity=False):
    """Stock vs drop-in forward of the same network object on the same synthetic pair.  precisions: which settings of
    stereoformer_b200.set_precision to run (the first one fills the top-level epe / max keys).  sensitivity=True also
    measures the STOCK network against itself with the left image moved by one ulp: the conditioning of the (random-
    init, untrained) network, i.e. the floor below which no implementation that is not bit-identical can stay."""
    import stereoformer_b200 as sf
    from tools.synth import stereo_pair
    net = build_net(variant, device, seed, d8)
    left, right, _ = stereo_pair(batch, height, width, maxdisp=8 * (d8 or 24), seed=seed, device=device)
    ref = forward(net, variant, left, right)
    res = {"variant": variant, "shape": [batch, height, width], "disp_mean": float(ref.mean()),
           "disp_max": float(ref.max())}
    for i, mode in enumerate(precisions):
        with sf.precision(mode), dropins():
            got = forward(net, variant, left, right)
        st = _stats(got - ref)
        res[mode] = st
        if i == 0:
            res.update(st)
            res["precision"] = mode
    if sensitivity:
        res["stock_vs_stock_1ulp_input"] = _stats(forward(net, variant, torch.nextafter(left, left + 1), right) - ref)
    return res


def time_forward(net, variant, left, right, reps=5):
    for _ in range(2):
        forward(net, variant, left, right)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        forward(net, variant, left, right)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


if __name__ == "__main__":
    import json
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda", 0)
    names = sys.argv[1:] or list(VARIANTS)
    for v in names:
        print(json.dumps(compare(v, 1, 544, 960, dev)), flush=True)
