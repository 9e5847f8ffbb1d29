"""The C-ABI shared library loads and exports every symbol include/stereoformer_b200.h declares (no compute)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared():
    txt = open(os.path.join(ROOT, "include", "stereoformer_b200.h")).read()
    return re.findall(r"SF_API\s+[\w\s\*]+?\b(sf_\w+)\s*\(", txt)


@pytest.fixture(scope="module")
def so():
    import __graft_entry__ as ge
    ge.build()
    from stereoformer_b200 import _lib
    return _lib


def test_header_symbols_exported(so):
    names = declared()
    assert len(names) >= 14
    h = ctypes.CDLL(so.LIB_PATH)
    for n in names:
        assert hasattr(h, n), "missing export " + n


def test_python_binding_covers_header(so):
    names = set(declared()) - {"sf_last_error", "sf_abi_version", "sf_device_sm_count"}
    assert names == set(so.SIGNATURES), names ^ set(so.SIGNATURES)
    assert so.lib().sf_abi_version() == 1


def test_header_arity_matches_binding(so):
    txt = open(os.path.join(ROOT, "include", "stereoformer_b200.h")).read()
    for name, args in so.SIGNATURES.items():
        m = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, txt, re.S)
        assert m, name
        assert len([a for a in m.group(1).split(",") if a.strip() not in ("", "void")]) == len(args), name


def test_cpu_tensors_fail_loudly(so):
    import torch
    import stereoformer_b200 as sf
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        sf.CostVolume(8)(torch.randn(1, 4, 2, 8), torch.randn(1, 4, 2, 8))
    with pytest.raises(RuntimeError, match="Not implemented on the CPU"):
        sf.ModulatedDeformConvPack(4, 4, 3, 1, 1)(torch.randn(1, 4, 5, 5))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "stereoformer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "oracle/" not in src, f


def test_workspace_queries_are_host_only(so):
    """The size queries make no CUDA call, so a caller can plan its buffers without a device (sizes per DESIGN section 3)."""
    h = so.lib()
    small = int(h.sf_dcn_workspace_bytes(128, 128, 3, 3, 1))
    assert small == 2 * 128 * 1152 * 2                                   # bf16 hi + mid of [Cout][Cin*9]
    big = int(h.sf_dcn_workspace_bytes_ex(1, 128, 248, 360, 128, 3, 3, 1, 1))
    assert big == small + 248 * 360 * 128 * 4                            # + the channels-last copy of the input
    assert int(h.sf_dcn_workspace_bytes_ex(8, 16, 68, 120, 16, 3, 3, 1, 1)) == int(h.sf_dcn_workspace_bytes(16, 16, 3, 3, 1))
    assert int(h.sf_dcn_workspace_bytes_ex(1, 24, 20, 33, 24, 3, 3, 3, 3)) == int(h.sf_dcn_workspace_bytes(24, 24, 3, 3, 3))   # 8 per group
    # backward: channels-last input + grad_input copies, split weights, column gradients [pixels][Cin * 9]
    bw = int(h.sf_dcn_backward_workspace_bytes(1, 128, 248, 360, 128, 3, 3, 1, 1, 1, 1, 1, 1, 1, 1))
    tiles = 31 * 23
    assert bw == 2 * 248 * 360 * 128 * 4 + 2 * 1152 * 128 * 2 + tiles * 18 * 8192 * 4
    assert int(h.sf_dcn_backward_workspace_bytes(8, 16, 68, 120, 16, 3, 3, 1, 1, 1, 1, 1, 1, 1, 1)) > 0      # the shipped 16 -> 16 layer
    assert int(h.sf_dcn_backward_workspace_bytes(2, 24, 20, 32, 24, 3, 3, 1, 1, 1, 1, 1, 1, 1, 1)) == 0     # 24 channels: FFMA kernels
    assert int(h.sf_dcn_static_scratch_bytes()) == 4 * 1024 * 1024
    assert int(h.sf_dcn_workspace_bytes(256, 256, 3, 3, 1)) <= int(h.sf_dcn_static_scratch_bytes())


def test_product_library_has_no_tuning_hooks(so):
    """the env-driven sweeps, debug kernel variants and timeline pointers exist only in the
    -DSF_TUNING build that tools/ load; the shipped library never reads them.  Every name passed to tuning_env() in
    the sources must be absent from libsfb200.so (the hook compiles to a constant nullptr there)."""
    import glob
    names = set()
    for f in glob.glob(os.path.join(ROOT, "stereoformer_b200", "csrc", "*.cu*")):
        names |= set(re.findall(r'tuning_env\("([A-Z0-9_]+)"\)', open(f).read()))
    assert len(names) >= 10, names
    blob = open(so.LIB_PATH, "rb").read()
    leaked = sorted(n for n in names if n.encode() + b"\0" in blob)
    assert not leaked, "tuning hooks compiled into the product library: %s" % leaked
    # and nothing else in the sources reads the environment behind tuning_env's back
    for f in glob.glob(os.path.join(ROOT, "stereoformer_b200", "csrc", "*.cu*")):
        txt = open(f).read()
        if os.path.basename(f) == "common.cuh":
            continue
        assert "getenv(" not in txt, f
