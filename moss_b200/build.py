"""In-tree build of the C-ABI library for sm_100a: ``python -m moss_b200.build``.

nvcc cross-compiles without a GPU; the resulting ``moss_b200/_lib/libmoss_splat.so`` is git-ignored
but travels with the repo snapshot to the GPU box.
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "_lib", "libmoss_splat.so")

NVCC_FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared", "-ldl"]       # -ldl: NVTX v3 (header-only) loads the tool's library itself


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) +
                  glob.glob(os.path.join(HERE, "..", "include", "*.h")))


def build(force=False, verbose=False):
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not force and os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(s) for s in sources()):
        return OUT
    cmd = ["nvcc", *NVCC_FLAGS, os.path.join(CSRC, "moss_splat.cu"), "-o", OUT]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
        print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return OUT


BINDING_SRC = os.path.join(CSRC, "torch_binding.cpp")
BINDING_OUT = os.path.join(HERE, "_C.so")


def build_binding(force=False, verbose=False):
    """The compiled autograd binding (csrc/torch_binding.cpp -> moss_b200/_C.so): host-only C++ over the C ABI, linked
    against libmoss_splat.so (found next to it through an $ORIGIN rpath) and the torch libraries of this interpreter."""
    build(force=False)
    deps = [BINDING_SRC, os.path.join(HERE, "..", "include", "moss_splat.h")]
    if not force and os.path.exists(BINDING_OUT) and all(os.path.getmtime(BINDING_OUT) >= os.path.getmtime(d) for d in deps):
        return BINDING_OUT
    import sysconfig
    import torch
    from torch.utils import cpp_extension as ce
    inc = ce.include_paths(device_type="cuda") if "device_type" in ce.include_paths.__code__.co_varnames else ce.include_paths(cuda=True)
    libdirs = ce.library_paths(device_type="cuda") if "device_type" in ce.library_paths.__code__.co_varnames else ce.library_paths(cuda=True)
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-Wno-deprecated-declarations",
           "-I" + sysconfig.get_paths()["include"]]
    cmd += ["-I" + p for p in inc]
    cmd += [BINDING_SRC, "-o", BINDING_OUT, "-L" + os.path.dirname(OUT), "-lmoss_splat", "-Wl,-rpath,$ORIGIN/_lib"]
    for d in libdirs:
        cmd += ["-L" + d, "-Wl,-rpath," + d]
    cmd += ["-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python"]
    if verbose:
        print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return BINDING_OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
    if "--no-binding" not in sys.argv:
        print(build_binding(force="--force" in sys.argv, verbose=True))
