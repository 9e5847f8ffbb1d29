"""The UNMODIFIED reference as a checker and a timed baseline (test infrastructure; never imported by the package).

Only tests/, __graft_entry__ and bench.py's reference / cpu_baseline legs import this module.

`/root/reference` exists only in the build container, so `vendor()` (run by `__graft_entry__.build()` there)
materialises what the checks need under the git-ignored `oracle/_ref/`, which travels to the GPU box with gpurun like
the built `.so` files:

  oracle/_ref/py/{models,utils,deform}   byte-for-byte copies of the reference's Python files (no edits)
  oracle/_ref/native/DCN.so              the reference's own pybind extension `DCN` (deform/src/vision.cpp:5-12) compiled
                                         from the sources where they lie under /root/reference/deform/src for sm_100a, with
                                         nothing but include-path shims for headers PyTorch has since removed
                                         (oracle/ref_shims/: THC/*.h -> ATen equivalents; one restored dispatch overload)

`activate()` puts them on sys.path with the import stubs SURVEY.md section 0 lists (matplotlib, timm: modules the
reference imports but never uses on this path).
"""
import glob
import os
import shutil
import subprocess
import sys
import sysconfig
import types
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_DST = os.path.join(HERE, "_ref")
PY_DST = os.path.join(REF_DST, "py")
NATIVE_DST = os.path.join(REF_DST, "native")
SHIMS = os.path.join(HERE, "ref_shims")
PY_TREES = ("models", "utils", "losses", "deform/modules", "deform/functions")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def vendor_python(force=False):
    """Copy the reference's Python files (unchanged) to oracle/_ref/py."""
    if not os.path.isdir(REF_SRC):
        return None
    n = 0
    for tree in PY_TREES:
        for src in glob.glob(os.path.join(REF_SRC, tree, "**", "*.py"), recursive=True):
            dst = os.path.join(PY_DST, os.path.relpath(src, REF_SRC))
            if force or not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                shutil.copyfile(src, dst)
                n += 1
    init = os.path.join(PY_DST, "deform", "__init__.py")
    if not os.path.exists(init):
        shutil.copyfile(os.path.join(REF_SRC, "deform", "__init__.py"), init)
    shutil.copyfile(os.path.join(REF_SRC, "deform", "test.py"), os.path.join(PY_DST, "deform", "test.py"))
    return n


def build_reference_dcn(force=False, verbose=False):
    """Compile the reference's DCN extension from /root/reference/deform/src (all of vision.cpp's six functions)."""
    out = os.path.join(NATIVE_DST, "DCN.so")
    if not os.path.isdir(REF_SRC):
        return out if os.path.exists(out) else None
    src = os.path.join(REF_SRC, "deform", "src")
    units = [os.path.join(src, p) for p in (
        "vision.cpp", "cpu/deform_cpu.cpp", "cpu/modulated_deform_cpu.cpp", "cpu/deform_psroi_pooling_cpu.cpp",
        "cuda/deform_conv_cuda.cu", "cuda/modulated_deform_conv_cuda.cu", "cuda/deform_psroi_pooling_cuda.cu")]
    if not force and os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(u) for u in units):
        return out
    import torch
    from torch.utils import cpp_extension as ce
    obj_dir = os.path.join(NATIVE_DST, "obj")
    os.makedirs(obj_dir, exist_ok=True)
    inc = ["-I" + SHIMS, "-I" + src] + ["-I" + p for p in ce.include_paths()] + \
          ["-I" + sysconfig.get_paths()["include"], "-I/usr/local/cuda/include"]
    common = ["-O2", "-std=c++17", "-DWITH_CUDA", "-DTORCH_EXTENSION_NAME=DCN", "-DTORCH_API_INCLUDE_EXTENSION_H",
              "-include", os.path.join(SHIMS, "compat.h")] + inc

    def cc(u):
        o = os.path.join(obj_dir, os.path.basename(u).rsplit(".", 1)[0] + ".o")
        if u.endswith(".cu"):
            cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-w"] + common
        else:
            cmd = ["g++", "-fPIC", "-w"] + common
        r = subprocess.run(cmd + ["-c", u, "-o", o], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("reference DCN build failed for %s:\n%s" % (u, r.stderr[-4000:]))
        return o

    with ThreadPoolExecutor(max_workers=len(units)) as ex:
        objs = list(ex.map(cc, units))
    tl = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-shared", "-o", out] + objs + ["-L" + tl, "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python",
                                                   "-lc10_cuda", "-ltorch_cuda", "-L/usr/local/cuda/lib64", "-lcudart",
                                                   "-Wl,-rpath," + tl]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("reference DCN link failed:\n%s" % r.stderr[-4000:])
    if verbose:
        print("built", out)
    return out


def vendor(force=False):
    """Everything above; a no-op (returning what is already there) when /root/reference is absent."""
    vendor_python(force)
    build_reference_dcn(force)
    return available()


def available():
    return {"python": os.path.isdir(os.path.join(PY_DST, "models")),
            "dcn": os.path.exists(os.path.join(NATIVE_DST, "DCN.so"))}


def ref_root():
    """Directory holding the reference's Python packages: the vendored copy, else /root/reference, else None."""
    if os.path.isdir(os.path.join(PY_DST, "models")):
        return PY_DST
    if os.path.isdir(os.path.join(REF_SRC, "models")):
        return REF_SRC
    return None


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__sf_stub__ = True
    sys.modules[name] = m
    return m


def activate(dcn="reference"):
    """Make `import models...` / `import utils...` / `from deform.modules...` resolve to the reference.

    dcn="reference": `import DCN` resolves to the reference's own compiled extension (oracle/_ref/native/DCN.so);
    dcn=None: leave `DCN` alone (the caller installs stereoformer_b200.DCN, or nothing needs it).
    Returns the root directory, or None when no copy of the reference is reachable."""
    root = ref_root()
    if root is None:
        return None
    import torch
    for p in (os.path.join(root, "deform"), root):
        if p not in sys.path:
            sys.path.insert(0, p)
    try:
        import matplotlib.pyplot  # noqa: F401  (models/utils/estimation.py:5 imports it and never uses it)
    except Exception:
        mp = _stub("matplotlib")
        mp.pyplot = _stub("matplotlib.pyplot")
    try:
        import timm.models.layers  # noqa: F401  (baseline*.py:16 uses trunc_normal_ only)
    except Exception:
        t = _stub("timm")
        t.models = _stub("timm.models")
        t.models.layers = _stub("timm.models.layers", trunc_normal_=torch.nn.init.trunc_normal_)
    if dcn == "reference":
        so = os.path.join(NATIVE_DST, "DCN.so")
        cur = sys.modules.get("DCN")
        if os.path.exists(so) and (cur is None or getattr(cur, "__file__", None) != so):
            import importlib.util
            spec = importlib.util.spec_from_file_location("DCN", so)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            sys.modules["DCN"] = mod
        elif not os.path.exists(so) and cur is None:
            _stub("DCN")          # import-only placeholder: networks without a DCN layer still import resnet.py
    return root


if __name__ == "__main__":
    print(vendor(force="--force" in sys.argv))
