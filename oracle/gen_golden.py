"""Generate tests/golden/*.npz by running the UNMODIFIED reference on seeded inputs.

TEST INFRASTRUCTURE.  Runs only in the build container (needs /root/reference, which does
not exist on the GPU box); the vectors it writes are committed.

    python oracle/gen_golden.py            # rewrites tests/golden/

Import stubs: matplotlib (models/utils/estimation.py:5) and the
``models.IterativeLocalCostVolume.utils.utils`` scipy import are satisfied by what the
container has or by empty stand-in modules; no reference arithmetic is touched.
"""
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("SF_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def import_reference():
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib.pyplot  # noqa: F401
        except Exception:
            mp = _stub("matplotlib")
            mp.pyplot = _stub("matplotlib.pyplot")
    sys.path.insert(0, REF)
    from models.utils.build_cost_volume import CostVolume
    from models.utils.estimation import DisparityEstimation, DisparityEstimationWithProb
    from models.LocalCostVolume.Attempts import fixed_local_cost_volume as flc
    from models.LocalCostVolume.Attempts import variance_local_cost_volume as vlc
    from models.LocalCostVolume.Attempts import update as upd
    from models.CrossAttentionCostVolume.GWC_Cost_Volume import build_gwc_volume
    from models.IterativeLocalCostVolume.corr.corr import CorrBlock1D
    from models.utils.upsample import upsample_convex8
    from utils.disparity_warper import disp_warp, meshgrid, normalize_coords
    return dict(CostVolume=CostVolume, DisparityEstimation=DisparityEstimation,
                DisparityEstimationWithProb=DisparityEstimationWithProb, flc=flc, vlc=vlc, upd=upd,
                build_gwc_volume=build_gwc_volume, CorrBlock1D=CorrBlock1D,
                upsample_convex8=upsample_convex8, disp_warp=disp_warp, meshgrid=meshgrid,
                normalize_coords=normalize_coords)


ONLY = set(a for a in sys.argv[1:] if not a.startswith("-"))   # `gen_golden.py NAME ...` rewrites only those files


def save(name, **arrs):
    if ONLY and name not in ONLY:
        return
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"),
                        **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v))
                           for k, v in arrs.items()})


def sampler_trace(mod, cv, cand):
    """Re-run the reference's own ste_ceil/ste_floor + clamp + long on the reference's candidates."""
    D = cv.shape[1]
    ic = torch.clamp(mod.ste_ceil(cand), min=0, max=D - 1).long()
    jf = torch.clamp(mod.ste_floor(cand), min=0, max=D - 1).long()
    return jf, ic


def main():
    R = import_reference()
    torch.manual_seed(1024)  # scripts/run.sh:31

    # ---- a1 / a1' / a2 cost volumes ------------------------------------------------
    for tag, (B, C, H, W, D) in {"small": (2, 16, 5, 40, 8), "odd": (1, 12, 3, 37, 6),
                                 "wide_d": (1, 8, 2, 12, 16)}.items():
        L, Rt = torch.randn(B, C, H, W), torch.randn(B, C, H, W)
        save("corr_volume_" + tag, left=L, right=Rt, D=D,
             corr=R["CostVolume"](D, "correlation")(L, Rt),
             diff=R["CostVolume"](D, "difference")(L, Rt),
             concat=R["CostVolume"](D, "concat")(L, Rt),
             gwc=R["build_gwc_volume"](L, Rt, D, 4))

    # ---- a3 soft-argmin ------------------------------------------------------------
    cv = torch.abs(torch.randn(2, 24, 6, 40) * 3 - 2)  # pure_learning_method.py:170 style
    d_sim = R["DisparityEstimation"](24, True)(cv)
    d_cost = R["DisparityEstimation"](24, False)(cv)
    d_p, prob = R["DisparityEstimationWithProb"](24, True)(cv)
    save("soft_argmin", cv=cv, disp_sim=d_sim, disp_cost=d_cost, prob=prob)

    # ---- a4 fixed window (radius 2, 20 points: baseline.py:38) ----------------------
    flc = R["flc"]
    cur = (torch.rand(2, 1, 6, 40) * 23)
    cur[0, 0, 0, :8] = torch.tensor([0., 1., 2., 3.5, 21., 20.999, 23., 10.])  # edges of validity
    for cvalid in (True, False):
        mod = flc.LocalCostVolume(radius=2, sample_points=20)
        out = mod(cv, cur, cvalid)
        # candidates exactly as LocalCostVolume.forward computes them (:20-45)
        lo, up = cur - 2, cur + 2
        if cvalid:
            inv = torch.clamp((lo < 0).float() + (up >= 23).float(), max=1.0)
        else:
            lo, up = torch.clamp(lo, min=0), torch.clamp(up, min=0, max=24)
        step = (up - lo) * 1.0 / 20
        cand = lo + torch.arange(21).type_as(cv).view(1, 21, 1, 1) * step
        if cvalid:
            cand = cand * (1 - inv) + inv * cur.repeat(1, 21, 1, 1)
        jf, ic = sampler_trace(flc, cv, cand)
        save("local_window_valid%d" % int(cvalid), cv=cv, cur=cur, radius=2, n_points=20,
             out=out, cand=cand, idx_floor=jf, idx_ceil=ic,
             local=flc.build_cost_volume_from_volume(cv, cand))

    # arbitrary candidates through build_cost_volume_from_volume (incl. out of range, integers)
    cand = torch.rand(2, 9, 6, 40) * 30 - 3
    cand[:, 0] = torch.round(cand[:, 0])
    jf, ic = sampler_trace(flc, cv, cand)
    save("sample_volume", cv=cv, cand=cand, idx_floor=jf, idx_ceil=ic,
         local=flc.build_cost_volume_from_volume(cv, cand))

    # ---- a5 variance window ----------------------------------------------------------
    vlc = R["vlc"]
    cvv = torch.randn(2, 24, 6, 40)
    curv = R["DisparityEstimation"](24, True)(cvv).unsqueeze(1)
    for cvalid in (True, False):
        sigma = vlc.GetVarince(torch.softmax(cvv, 1), curv)
        # bounds / candidates exactly as Variance_Based_Local_Cost_Volume.forward computes them (:41-66)
        lo0, up0 = curv - 1 * sigma, curv + 1 * sigma
        lo, up = lo0, up0
        if cvalid:
            xs = torch.arange(40).view(1, 1, 1, 40).repeat(1, 1, 6, 1).type_as(curv)
            inv = torch.clamp((lo < 0).float() + (up >= 23).float() + (up > xs).float(), max=1.0)
        else:
            lo, up = torch.clamp(lo, min=0, max=23), torch.clamp(up, min=0, max=23)
        step = (up - lo) * 1.0 / 20
        cand = lo + torch.arange(21).type_as(cvv).view(1, 21, 1, 1) * step
        if cvalid:
            cand = cand * (1 - inv) + inv * curv.repeat(1, 21, 1, 1)
        jf, ic = sampler_trace(vlc, cvv, cand)
        save("variance_window_valid%d" % int(cvalid), cv=cvv, cur=curv, gamma=1.0, n_points=20,
             out=vlc.Variance_Based_Local_Cost_Volume(1, 20)(cvv, curv, cvalid),
             sigma=sigma, lo=lo0, up=up0, cand=cand, idx_floor=jf, idx_ceil=ic,
             local=vlc.build_cost_volume_from_volume(cvv, cand))

    # ---- a6 DLC window tail + uncertainty volume (update.py:279-313, :86) -------------
    upd = R["upd"]
    lo_b, up_b = torch.relu(torch.randn(2, 1, 6, 40)) * 3, torch.relu(torch.randn(2, 1, 6, 40)) * 3
    for cvalid in (True, False):
        lo, up = cur - lo_b, cur + up_b
        if cvalid:
            inv = torch.clamp((lo < 0).float() + (up >= 23).float(), max=1.0)
        else:
            lo, up = torch.clamp(lo, min=0), torch.clamp(up, min=0, max=24)
        step = (up - lo) * 1.0 / 20
        cand = lo + torch.arange(21).type_as(cv).view(1, 21, 1, 1) * step
        if cvalid:
            cand = cand * (1 - inv) + inv * cur.repeat(1, 21, 1, 1)
        local = upd.build_cost_volume_from_volume(cv, cand)
        out = torch.sum(torch.softmax(local, 1) * cand, 1).unsqueeze(1)
        jf, ic = sampler_trace(upd, cv, cand)
        save("dlc_window_valid%d" % int(cvalid), cv=cv, cur=cur, lo=lo_b, up=up_b, n_points=20,
             out=out, cand=cand, idx_floor=jf, idx_ceil=ic)
    dcand = torch.arange(24).view(1, 24, 1, 1).repeat(1, 1, 6, 40).type_as(cur)
    unc = torch.softmax(cv, 1) * dcand * (dcand - cur) * (dcand - cur)
    save("uncertainty", cv=cv, cur=cur, unc=unc)

    # ---- a7 warp ----------------------------------------------------------------------
    img = torch.randn(2, 3, 6, 40)
    dw = torch.rand(2, 1, 6, 40) * 12
    dw[0, 0, 0, :4] = torch.tensor([0., 0.5, 39., 45.])
    grid = R["normalize_coords"](R["meshgrid"](img) + torch.cat((-dw, torch.zeros_like(dw)), 1))
    wb, mb = R["disp_warp"](img, dw)
    wz, mz = R["disp_warp"](img, dw, padding_mode="zeros")
    save("disp_warp", img=img, disp=dw, grid=grid, warped_border=wb, mask_border=mb,
         warped_zeros=wz, mask_zeros=mz)

    # ---- a8 / a9 CorrBlock1D ------------------------------------------------------------
    f1, f2 = torch.randn(2, 16, 3, 24), torch.randn(2, 16, 3, 24)
    blk = R["CorrBlock1D"](f1, f2, num_levels=3, radius=2)
    xs = torch.arange(24).float().view(1, 1, 1, 24).repeat(2, 1, 3, 1)
    coords = torch.cat((xs - torch.rand(2, 1, 3, 24) * 8, torch.zeros(2, 1, 3, 24)), 1)
    coords[0, 0, 0, :4] = torch.tensor([0., 3., 23., -2.5])
    save("corr1d", f1=f1, f2=f2, coords=coords, radius=2, num_levels=3,
         lookup=blk(coords), **{"pyr%d" % i: p for i, p in enumerate(blk.corr_pyramid)})

    # ---- f1 convex upsampling -------------------------------------------------------------
    dlow, msk = torch.rand(1, 1, 3, 5) * 10, torch.randn(1, 576, 3, 5)
    save("upsample_convex8", disp=dlow, mask=msk, out=R["upsample_convex8"](dlow, msk))

    # ---- f2 ConvGRU gates + one DLC-GRU step (update.py:16-49, 243-317) -------------------------------------
    torch.manual_seed(77)
    gru = upd.ConvGRU(nb_channel=8, softsign=False)
    with torch.no_grad():                    # the reference zero-fills the biases; make them count
        for c in (gru.conv_z, gru.conv_b, gru.conv_g):
            c.bias.normal_(0, 0.5)
    gin, gh = torch.randn(2, 8, 5, 12), torch.randn(2, 8, 5, 12)
    cap = {}
    hooks = [getattr(gru, n).register_forward_hook(lambda m, i, o, n=n: cap.update({n: o.detach().clone(),
                                                                                    n + "_in": i[0].detach().clone()}))
             for n in ("conv_z", "conv_b", "conv_g")]
    with torch.no_grad():
        gout = gru(gin, gh)
    for h in hooks:
        h.remove()
    save("conv_gru", input=gin, prev_h=gh, z_pre=cap["conv_z"], b_pre=cap["conv_b"], s=cap["conv_g_in"],
         g_pre=cap["conv_g"], h=gout, **{"sd_" + k: v for k, v in gru.state_dict().items() if "conv_" in k and
                                         k.split(".")[0] in ("conv_z", "conv_b", "conv_g")})
    gru_ss = upd.ConvGRU(nb_channel=8, softsign=True)
    gru_ss.load_state_dict(gru.state_dict())
    with torch.no_grad():
        save("conv_gru_softsign", h=gru_ss(gin, gh))

    torch.manual_seed(78)
    step = upd.DisparityUpdateDLCWithMask(input_channels=24, hidden_dim=4, sample_points=20).eval()
    with torch.no_grad():
        for m in step.modules():
            if isinstance(m, torch.nn.BatchNorm2d):
                m.running_mean.normal_(0, 0.2); m.running_var.uniform_(0.5, 1.5)
                m.weight.uniform_(0.5, 1.5); m.bias.normal_(0, 0.2)
        step.offset.conv2.bias.fill_(1.0)    # keeps a good share of the bounds positive after relu2
        lr, rr = torch.randn(2, 3, 6, 40), torch.randn(2, 3, 6, 40)
        hid0 = torch.randn(2, 8, 6, 40)
        d1, h1, m1 = step(cv, cur, lr, rr, hid0, True)
        d2, h2, m2 = step(cv, d1, lr, rr, h1, False)
    save("dlc_update_step", cv=cv, cur=cur, left=lr, right=rr, hidden0=hid0, disp1=d1, hidden1=h1,
         mask1_head=m1[:, :18], disp2=d2, hidden2=h2, mask2_head=m2[:, :18],
         **{"sd_" + k: (v[:18] if k.startswith("mask.2.") else v) for k, v in step.state_dict().items()
            if "num_batches" not in k and not k.startswith(("gru.conv_zz", "gru.conv_bb", "gru.conv_gg"))})
    # (only the first 18 of the 576 mask-head output filters are kept: the fixture checks mask[:, :18])

    # ---- f4 on-disk formats: PFM bytes written / read by the reference's utils/file_io.py ------------------
    import tempfile
    from utils import file_io as fio
    rng = np.random.RandomState(5)
    grey = (rng.rand(5, 7).astype(np.float32) * 200)
    color = rng.randn(4, 6, 3).astype(np.float32)
    blobs = {}
    with tempfile.TemporaryDirectory() as td:
        for name, pimg, sc in (("grey", grey, 1), ("color", color, 2.5), ("grey1", grey[:, :, None], 1)):
            fn = os.path.join(td, name + ".pfm")
            fio.write_pfm(fn, pimg, sc)
            blobs["pfm_" + name] = np.frombuffer(open(fn, "rb").read(), dtype=np.uint8)
            back, bsc = fio._read_pfm(fn)
            blobs["read_" + name] = np.ascontiguousarray(back)
            blobs["scale_" + name] = bsc
        big = os.path.join(td, "big.pfm")               # a big-endian file (positive scale)
        with open(big, "wb") as f:
            f.write(b"Pf\n7 5\n1.000000\n")
            f.write(np.flipud(grey).astype(">f4").tobytes())
        blobs["pfm_big"] = np.frombuffer(open(big, "rb").read(), dtype=np.uint8)
        blobs["read_big"] = np.ascontiguousarray(fio._read_pfm(big)[0]).astype(np.float32)
        from PIL import Image
        k16 = (rng.rand(4, 9) * 40000).astype(np.uint16)
        kp = os.path.join(td, "kitti.png")
        Image.fromarray(k16).save(kp)
        blobs["png_kitti"] = np.frombuffer(open(kp, "rb").read(), dtype=np.uint8)
        blobs["read_kitti"] = fio.read_disp(kp)
        blobs["read_subset"] = fio.read_disp(os.path.join(td, "grey.pfm"), subset=True)
    save("file_io", grey=grey, color=color, **blobs)

    # ---- f3 gradients: autograd through the UNMODIFIED reference modules on the inputs of the fixtures above ----
    def grads(fn, *ts):
        leaves = [t.clone().requires_grad_(True) for t in ts]
        out = fn(*leaves)
        out = out[0] if isinstance(out, tuple) else out
        gen = torch.Generator().manual_seed(123)
        go = torch.randn(out.shape, generator=gen)
        return [go] + list(torch.autograd.grad(out, leaves, go, allow_unused=True))

    gz = np.load(os.path.join(OUT, "corr_volume_small.npz"))
    Lg, Rg, Dg = torch.from_numpy(gz["left"]), torch.from_numpy(gz["right"]), int(gz["D"])
    blob = {}
    for name, res in (
            ("corr", grads(lambda l, r: R["CostVolume"](Dg, "correlation")(l, r), Lg, Rg)),
            ("gwc", grads(lambda l, r: R["build_gwc_volume"](l, r, Dg, 4), Lg, Rg)),
            ("argmin_sim", grads(lambda c: R["DisparityEstimation"](24, True)(c), cv)),
            ("argmin_cost", grads(lambda c: R["DisparityEstimation"](24, False)(c), cv)),
            ("win_valid1", grads(lambda c, d: flc.LocalCostVolume(2, 20)(c, d, True), cv, cur)),
            ("win_valid0", grads(lambda c, d: flc.LocalCostVolume(2, 20)(c, d, False), cv, cur)),
            ("sample", grads(lambda c, k: flc.build_cost_volume_from_volume(c, k), cv,
                             torch.from_numpy(np.load(os.path.join(OUT, "sample_volume.npz"))["cand"]))),
            ("warp_border", grads(lambda i, d: R["disp_warp"](i, d), img, dw)),
            ("warp_zeros", grads(lambda i, d: R["disp_warp"](i, d, padding_mode="zeros"), img, dw))):
        for k, t in enumerate(res):
            if t is not None:
                blob["%s_%d" % (name, k)] = t
    save("gradients", **blob)

    # ---- gradients of the variance window (a5) through the unmodified reference module; added after the others
    vblob = {}
    for cvalid in (0, 1):
        gv = np.load(os.path.join(OUT, "variance_window_valid%d.npz" % cvalid))
        res = grads(lambda c, d: vlc.Variance_Based_Local_Cost_Volume(float(gv["gamma"]), int(gv["n_points"]))(
            c, d, bool(cvalid)), torch.from_numpy(gv["cv"]), torch.from_numpy(gv["cur"]))
        for k, t in enumerate(res):
            vblob["var_valid%d_%d" % (cvalid, k)] = t
    save("gradients_variance", **vblob)

    # ---- a9 alternative: PytorchAlternateCorrBlock1D (corr/corr.py:8-51); own seed: added after the others
    from models.IterativeLocalCostVolume.corr.corr import PytorchAlternateCorrBlock1D
    torch.manual_seed(80)
    a1, a2 = torch.randn(2, 16, 5, 24), torch.randn(2, 16, 5, 24)
    xs = torch.arange(24).float().view(1, 1, 1, 24).repeat(2, 1, 5, 1) - torch.rand(2, 1, 5, 24) * 8
    ys = torch.arange(5).float().view(1, 1, 5, 1).repeat(2, 1, 1, 24)
    ac = torch.cat((xs, ys), 1)
    ac[0, 0, 0, :4] = torch.tensor([0., 3., 23., -2.5])
    save("corr1d_alt", f1=a1, f2=a2, coords=ac, radius=2, num_levels=3,
         out=PytorchAlternateCorrBlock1D(a1, a2, num_levels=3, radius=2)(ac))

    # ---- f1 upsample_simple8 (models/utils/upsample.py:40-43) with its gradient; own seed: added after the others
    from models.utils.upsample import upsample_simple8
    torch.manual_seed(79)
    ds = (torch.rand(2, 1, 5, 9) * 20).requires_grad_(True)
    us = upsample_simple8(ds)
    gs = torch.randn(us.shape)
    save("upsample_simple8", disp=ds, out=us, grad_out=gs, grad_disp=torch.autograd.grad(us, ds, gs)[0])
    print("golden vectors written to", os.path.normpath(OUT))


if __name__ == "__main__":
    main()
