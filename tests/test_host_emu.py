"""No-GPU unit tests of the kernels' per-pixel arithmetic (stereoformer_b200/csrc/pixel_math.h compiled as
plain C++ by tests/host_emu) against the golden vectors of the unmodified reference: candidates, indices and
sampling coordinates must be bit-identical.  This is a test harness, not a CPU fallback: the package never
loads it."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import torch

from conftest import ROOT, load_golden
from oracle import stereo_oracle as O

SRC = os.path.join(ROOT, "tests", "host_emu", "emu.cpp")
SO = os.path.join(ROOT, "tests", "host_emu", "_build", "libemu.so")


@pytest.fixture(scope="module")
def emu():
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    hdr = os.path.join(ROOT, "stereoformer_b200", "csrc", "pixel_math.h")
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", SO, SRC])
    return ctypes.CDLL(SO)


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _window(emu, g, mode, valid, recip=0):
    cur = g["cur"].numpy().ravel()
    n, K = cur.size, g["n_points"] + 1
    lo = g["lo"].numpy().ravel() if "lo" in g else None
    up = g["up"].numpy().ravel() if "up" in g else None
    cand = np.empty((n, K), np.float32)
    jf, ic = np.empty((n, K), np.int32), np.empty((n, K), np.int32)
    emu.emu_window(_p(cur), _p(lo), _p(up), n, g["cv"].shape[1], g["n_points"], mode,
                   ctypes.c_float(float(g.get("radius", 0))), valid, recip, _p(cand), _p(jf), _p(ic))
    B, _, H, W = g["cur"].shape
    f = lambda a: torch.from_numpy(a).view(B, H, W, K).permute(0, 3, 1, 2)
    return f(cand), f(jf), f(ic)


@pytest.mark.parametrize("valid", [0, 1])
def test_fixed_window_bits(emu, valid):
    g = load_golden("local_window_valid%d" % valid)
    cand, jf, ic = _window(emu, g, 0, valid)
    assert torch.equal(cand, g["cand"])
    assert torch.equal(jf.long(), g["idx_floor"]) and torch.equal(ic.long(), g["idx_ceil"])
    vf, vc = torch.gather(g["cv"], 1, jf.long()), torch.gather(g["cv"], 1, ic.long())
    out = np.empty(cand.numel(), np.float32)
    emu.emu_lerp(_p(vf.contiguous().numpy()), _p(vc.contiguous().numpy()), _p(ic.contiguous().numpy()),
                 _p(cand.contiguous().numpy()), cand.numel(), _p(out))
    assert torch.equal(torch.from_numpy(out).view_as(g["local"]), g["local"])


@pytest.mark.parametrize("valid", [0, 1])
def test_dlc_window_bits(emu, valid):
    g = load_golden("dlc_window_valid%d" % valid)
    cand, jf, ic = _window(emu, g, 2, valid)
    assert torch.equal(cand, g["cand"])
    assert torch.equal(jf.long(), g["idx_floor"]) and torch.equal(ic.long(), g["idx_ceil"])


def test_reciprocal_division_mode(emu):
    g = load_golden("local_window_valid1")
    cand, _, _ = _window(emu, g, 0, 1, recip=1)
    ref, _ = O.window_candidates(g["cur"] - 2, g["cur"] + 2, g["cur"], 24, 20, True, recip=True)
    assert torch.equal(cand, ref)
    assert not torch.equal(cand, g["cand"])      # the two roundings really differ on this data


def test_arbitrary_candidate_indices(emu):
    g = load_golden("sample_volume")
    c = g["cand"].contiguous().numpy()
    jf, ic = np.empty(c.size, np.int32), np.empty(c.size, np.int32)
    emu.emu_index(_p(c), c.size, g["cv"].shape[1], _p(jf), _p(ic))
    assert torch.equal(torch.from_numpy(jf).view_as(g["idx_floor"]).long(), g["idx_floor"])
    assert torch.equal(torch.from_numpy(ic).view_as(g["idx_ceil"]).long(), g["idx_ceil"])


def test_warp_grid_bits(emu):
    g = load_golden("disp_warp")
    B, _, H, W = g["disp"].shape
    for b in range(B):
        d = g["disp"][b, 0].contiguous().numpy()
        grid, src = np.empty((H * W, 2), np.float32), np.empty((H * W, 2), np.float32)
        emu.emu_warp(_p(d), H, W, 0, _p(grid), _p(src))
        assert torch.equal(torch.from_numpy(grid).view(H, W, 2), g["grid"][b])
        ix, iy = O.warp_source_coords(g["grid"][b], H, W)
        assert torch.equal(torch.from_numpy(src).view(H, W, 2)[..., 0], ix)
        assert torch.equal(torch.from_numpy(src).view(H, W, 2)[..., 1], iy)


def test_lookup_coords_bits(emu):
    g = load_golden("corr1d")
    cx = g["coords"][:, 0].reshape(-1).contiguous().numpy()
    r = g["radius"]
    for l in range(g["num_levels"]):
        Wl = g["pyr%d" % l].shape[-1]
        x0 = np.empty((cx.size, 2 * r + 1), np.float32)
        xg, ix = np.empty_like(x0), np.empty_like(x0)
        emu.emu_lookup(_p(cx), cx.size, l, r, Wl, 0, _p(x0), _p(xg), _p(ix))
        rx0, rxg = O.corr1d_lookup_coords(torch.from_numpy(cx), l, r, Wl)
        assert torch.equal(torch.from_numpy(x0), rx0) and torch.equal(torch.from_numpy(xg), rxg)
        assert torch.equal(torch.from_numpy(ix), ((rxg + 1) / 2) * (Wl - 1))
