"""In-tree build of libsfb200.so (hand-written sm_100a kernels + the C ABI).

    python -m stereoformer_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  Objects go to stereoformer_b200/csrc/_build/, the
library to stereoformer_b200/libsfb200.so (git-ignored, travels to the GPU box with gpurun).
"""
import glob
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_build")
LIB = os.path.join(HERE, "libsfb200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _deps_mtime():
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return max(os.path.getmtime(h) for h in hdrs)


def build(force=False, verbose=False, tuning=False):
    """tuning=True: the -DSF_TUNING build (env-driven sweeps, component-isolation kernels, clock64 timelines) for
    tools/, written to libsfb200_tuning.so; the product library never reads the environment."""
    flags = FLAGS + (["-DSF_TUNING"] if tuning else [])
    obj_dir = OBJ + ("_tuning" if tuning else "")
    lib = LIB.replace("libsfb200.so", "libsfb200_tuning.so") if tuning else LIB
    return _build(force, verbose, flags, obj_dir, lib)


def _build(force, verbose, FLAGS, OBJ, LIB):
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hm = _deps_mtime()
    todo, objs = [], []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(s), hm):
            todo.append((s, o))

    def cc(so):
        s, o = so
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (s, r.stdout, r.stderr))
        return s, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(todo)))) as ex:
        for s, err in ex.map(cc, todo):
            if verbose:
                print("==", os.path.basename(s))
                print(err)
    if todo or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


def build_torch_ext(force=False):
    """The torch C++-extension front end (csrc/torch_ext.cpp -> _sfb200_torch.so, in-tree): pybind11 functions that
    forward to the C ABI of libsfb200.so.  g++ only (no device code)."""
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    src = os.path.join(CSRC, "torch_ext.cpp")
    out = os.path.join(HERE, "_sfb200_torch.so")
    hdr = os.path.join(HERE, "..", "include", "stereoformer_b200.h")
    # (dynamically linked against libsfb200.so: only the front end's own source and the ABI header matter)
    if not force and os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return out
    tl = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-DTORCH_EXTENSION_NAME=_sfb200_torch",
           "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)] + \
          ["-I" + p for p in ce.include_paths()] + ["-I" + sysconfig.get_paths()["include"], "-I/usr/local/cuda/include",
                                                    src, "-o", out, "-L" + HERE, "-l:libsfb200.so", "-L" + tl, "-lc10",
                                                    "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda",
                                                    "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tl]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("torch extension build failed:\n%s" % r.stderr[-4000:])
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, tuning="--tuning" in sys.argv))
    if "--tuning" not in sys.argv:
        print(build_torch_ext(force="--force" in sys.argv))
