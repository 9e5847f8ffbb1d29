"""Pins the DCN oracles (plain C and dense-PyTorch forms) with the reference's own analytic tests
(deform/test.py) and cross-checks them against torchvision.ops.deform_conv2d.  CPU only."""
import pytest
import torch
import torch.nn as nn

from oracle import dcn_c
from oracle import stereo_oracle as O

# deform/test.py:15-21
N, inC, inH, inW, outC, kH, kW = 2, 4, 4, 4, 4, 3, 3


def _conv_ref(groups=2):
    torch.manual_seed(3)
    conv = nn.Conv2d(inC, outC, (kH, kW), stride=1, padding=1, dilation=1, groups=groups, bias=True)
    return conv


@pytest.mark.parametrize("impl", ["c", "torch"])
def test_zero_offset_equals_conv2d(impl):
    """deform/test.py:69-110 (check_mdconv_zero_offset): zero offsets, mask = 1 -> nn.Conv2d(groups=2), 1e-5."""
    conv = _conv_ref(2)
    x = torch.randn(N, inC, inH, inW)
    off = torch.zeros(N, 2 * kH * kW, inH, inW)
    msk = torch.ones(N, kH * kW, inH, inW)
    ref = conv(x).detach()
    args = (x, conv.weight.detach(), conv.bias.detach(), off, msk, (1, 1), (1, 1), (1, 1), 2, 1)
    out = dcn_c.dcn_forward(*args) if impl == "c" else O.dcn_forward(*args)
    assert (out - ref).abs().max() < 1e-5
    # DCNv1 twin, deform/test.py:36-67
    args1 = (x, conv.weight.detach(), conv.bias.detach(), off, None, (1, 1), (1, 1), (1, 1), 2, 1)
    out1 = dcn_c.dcn_forward(*args1) if impl == "c" else O.dcn_forward(*args1)
    assert (out1 - ref).abs().max() < 1e-5


def test_identity_weights():
    """deform/test.py:142-181: identity kernel, mask 0.5 -> 2 * out == input."""
    x = torch.randn(N, inC, inH, inW)
    w = torch.zeros(inC, 1, kH, kW)
    w[:, 0, 1, 1] = 1.0
    off = torch.zeros(N, 2 * kH * kW, inH, inW)
    msk = torch.full((N, kH * kW, inH, inW), 0.5)
    out = dcn_c.dcn_forward(x, w, torch.zeros(inC), off, msk, (1, 1), (1, 1), (1, 1), inC, 1)
    assert (2 * out - x).abs().max() < 1e-6


@pytest.mark.parametrize("cfg", [dict(Cin=8, Cout=6, G=2, dg=2, s=1, p=1, d=1, H=7, W=9),
                                 dict(Cin=4, Cout=4, G=1, dg=1, s=2, p=2, d=2, H=9, W=8),
                                 dict(Cin=6, Cout=3, G=3, dg=3, s=1, p=0, d=1, H=6, W=6)])
def test_against_torchvision(cfg):
    tv = pytest.importorskip("torchvision.ops")
    torch.manual_seed(0)
    Cin, Cout, G, dg, s, p, d, H, W = (cfg[k] for k in ("Cin", "Cout", "G", "dg", "s", "p", "d", "H", "W"))
    x = torch.randn(2, Cin, H, W)
    w = torch.randn(Cout, Cin // G, 3, 3)
    b = torch.randn(Cout)
    Ho = (H + 2 * p - (d * 2 + 1)) // s + 1
    Wo = (W + 2 * p - (d * 2 + 1)) // s + 1
    off = torch.randn(2, dg * 18, Ho, Wo) * 2      # deform/test.py:313
    msk = torch.sigmoid(torch.randn(2, dg * 9, Ho, Wo))
    ref = tv.deform_conv2d(x, off, w, b, stride=s, padding=p, dilation=d, mask=msk)
    oc = dcn_c.dcn_forward(x, w, b, off, msk, (s, s), (p, p), (d, d), G, dg)
    ot = O.dcn_forward(x, w, b, off, msk, (s, s), (p, p), (d, d), G, dg)
    assert (oc - ref).abs().max() < 1e-4
    assert (ot - ref).abs().max() < 1e-4
    ref1 = tv.deform_conv2d(x, off, w, b, stride=s, padding=p, dilation=d)
    assert (dcn_c.dcn_forward(x, w, b, off, None, (s, s), (p, p), (d, d), G, dg) - ref1).abs().max() < 1e-4


# ------------------------------------------------------------------------------------------ backward (f3)
BWD_CFGS = [dict(Cin=8, Cout=6, G=2, dg=2, s=1, p=1, d=1, H=7, W=9, k=3, B=2),
            dict(Cin=4, Cout=4, G=1, dg=1, s=2, p=2, d=2, H=9, W=8, k=3, B=1),
            dict(Cin=6, Cout=3, G=3, dg=1, s=1, p=0, d=1, H=6, W=6, k=3, B=2),     # a deformable group spans 3 groups
            dict(Cin=4, Cout=8, G=1, dg=2, s=1, p=0, d=1, H=5, W=7, k=1, B=2)]


def _bwd_inputs(c, v2=True, seed=11):
    torch.manual_seed(seed)
    k = c["k"]
    Ho = (c["H"] + 2 * c["p"] - (c["d"] * (k - 1) + 1)) // c["s"] + 1
    Wo = (c["W"] + 2 * c["p"] - (c["d"] * (k - 1) + 1)) // c["s"] + 1
    x = torch.randn(c["B"], c["Cin"], c["H"], c["W"])
    w = torch.randn(c["Cout"], c["Cin"] // c["G"], k, k) * 0.3
    b = torch.randn(c["Cout"])
    off = torch.randn(c["B"], c["dg"] * 2 * k * k, Ho, Wo) * 2           # deform/test.py:313 scale; samples leave the image
    msk = torch.sigmoid(torch.randn(c["B"], c["dg"] * k * k, Ho, Wo)) if v2 else None
    go = torch.randn(c["B"], c["Cout"], Ho, Wo)
    return x, w, b, off, msk, go


@pytest.mark.parametrize("cfg", BWD_CFGS)
@pytest.mark.parametrize("v2", [True, False])
def test_backward_oracle_matches_autograd(cfg, v2):
    """The C restatement of col2im / col2im_coord / the weight GEMMs against autograd through the float64 dense
    forward (two independent derivations; the reference pins its kernels with gradcheck, deform/test.py:304-349)."""
    x, w, b, off, msk, go = _bwd_inputs(cfg, v2)
    sp = ((cfg["s"],) * 2, (cfg["p"],) * 2, (cfg["d"],) * 2, cfg["G"], cfg["dg"])
    leaves = [t.double().requires_grad_(True) if t is not None else None for t in (x, off, msk, w, b)]
    out = O.dcn_forward(leaves[0], leaves[3], leaves[4], leaves[1], leaves[2], *sp)
    ref = torch.autograd.grad(out, [t for t in leaves if t is not None], go.double())
    got = [g for g in dcn_c.dcn_backward(x, w, off, msk, go, *sp) if g is not None]
    assert len(got) == len(ref)
    for name, a, r in zip(("input", "offset", "mask", "weight", "bias") if v2 else ("input", "offset", "weight", "bias"),
                          got, ref):
        scale = r.abs().max().item() + 1e-6
        assert (a.double() - r).abs().max().item() < 2e-5 * max(scale, 1.0), name


def test_backward_oracle_numerical():
    """Central differences on the C forward (float32, so coarse) for a few offset / mask / input coordinates."""
    c = BWD_CFGS[0]
    x, w, b, off, msk, go = _bwd_inputs(c, True, seed=5)
    sp = ((c["s"],) * 2, (c["p"],) * 2, (c["d"],) * 2, c["G"], c["dg"])
    gin, goff, gm, gw, gb = dcn_c.dcn_backward(x, w, off, msk, go, *sp)

    def loss(x_, off_, msk_):
        return (dcn_c.dcn_forward(x_, w, b, off_, msk_, *sp).double() * go.double()).sum().item()

    eps = 1e-2
    gen = torch.Generator().manual_seed(0)
    for t, g in ((x, gin), (msk, gm)):
        for _ in range(6):
            idx = tuple(int(torch.randint(0, n, (1,), generator=gen)) for n in t.shape)
            tp, tm = t.clone(), t.clone()
            tp[idx] += eps
            tm[idx] -= eps
            num = (loss(tp if t is x else x, off, tp if t is msk else msk) -
                   loss(tm if t is x else x, off, tm if t is msk else msk)) / (2 * eps)
            assert abs(num - g[idx].item()) < 2e-2 * max(1.0, abs(num)), (idx, num, g[idx].item())
