"""Experiment: where does the full-model DLC-GRU (12 iterations, random-init weights) disparity difference come from?
stock vs drop-ins at both precisions, and the stock network's own sensitivity to a 1-ulp perturbation of its input."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import stereoformer_b200 as sf
from tools import fullmodel
from tools.synth import stereo_pair

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
dev = torch.device("cuda", 0)
for variant, B in (("dlc_gru", 2), ("lc", 2)):
    net = fullmodel.build_net(variant, dev)
    left, right, _ = stereo_pair(B, 544, 960, device=dev)
    ref = fullmodel.forward(net, variant, left, right)
    res = {"variant": variant, "batch": B}
    for mode in ("fast", "exact"):
        with sf.precision(mode), fullmodel.dropins():
            got = fullmodel.forward(net, variant, left, right)
        e = (got - ref).abs()
        res[mode] = {"epe": float(e.mean()), "max": float(e.max()), "over_1e-3": float((e > 1e-3).float().mean())}
    # the stock network against itself with the input moved by one ulp
    l2 = torch.nextafter(left, left + 1)
    e = (fullmodel.forward(net, variant, l2, right) - ref).abs()
    res["stock_vs_stock_1ulp_input"] = {"epe": float(e.mean()), "max": float(e.max()), "over_1e-3": float((e > 1e-3).float().mean())}
    # and with cuDNN free to pick other algorithms
    torch.backends.cudnn.benchmark = True
    e = (fullmodel.forward(net, variant, left, right) - ref).abs()
    torch.backends.cudnn.benchmark = False
    res["stock_vs_stock_cudnn_benchmark"] = {"epe": float(e.mean()), "max": float(e.max())}
    print(json.dumps(res), flush=True)
