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
