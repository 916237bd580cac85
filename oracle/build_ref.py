"""Build recipe for the reference rasterizer (TEST/BENCH INFRASTRUCTURE ONLY).

Compiles the UNMODIFIED reference sources where they lie under
/root/reference/submodules/diff-gaussian-rasterization into ``oracle/_ref/``
(git-ignored, but shipped to the GPU box):

  oracle/_ref/libref_raster.so
      reference cuda_rasterizer/{forward,backward,rasterizer_impl}.cu + our own
      C-ABI shim (oracle/ref_shim.cu) that drives CudaRasterizer::Rasterizer::
      forward/backward with raw pointers and exports the internal state
      (tiles_touched, sorted key/ID lists, tile ranges ...) for bit-exact parity.
      No torch headers -> compiles in ~1 min.

  oracle/_ref/diff_gaussian_rasterization/{__init__.py,_C.so}
      the reference's own torch extension + python wrapper (an *install* of the
      reference package, like ``pip install --target``), used ONLY by
      ``bench.py --impl reference`` / the B1 timing leg.  ~6 min (torch/extension.h
      under nvcc).

Nothing under oracle/ is ever imported by the product package ``moss_b200``.
No reference source is copied into the repository history.

gcc-13 fix: rasterizer_impl.h uses fixed-width ints without <cstdint>
(DGR/cuda_rasterizer/rasterizer_impl.h:24,40-61) -> ``-include cstdint``.
"""
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/submodules/diff-gaussian-rasterization"
OUT = os.path.join(HERE, "_ref")
ARCH = ["-gencode", "arch=compute_100,code=sm_100"]


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def available():
    return os.path.isdir(REF)


def build_shim(force=False):
    """reference .cu files + oracle/ref_shim.cu -> oracle/_ref/libref_raster.so"""
    os.makedirs(OUT, exist_ok=True)
    so = os.path.join(OUT, "libref_raster.so")
    srcs = [os.path.join(REF, "cuda_rasterizer", f) for f in
            ("forward.cu", "backward.cu", "rasterizer_impl.cu")] + [os.path.join(HERE, "ref_shim.cu")]
    if not force and os.path.exists(so) and all(
            os.path.getmtime(so) >= os.path.getmtime(s) for s in srcs):
        return so
    objs = []
    for s in srcs:
        o = os.path.join(OUT, os.path.basename(s).replace(".cu", ".o"))
        _run(["nvcc", "-std=c++17", "-O3", *ARCH, "-Xcompiler", "-fPIC", "-include", "cstdint",
              "-I", os.path.join(REF, "third_party", "glm"), "-I", REF,
              "-I", os.path.join(REF, "cuda_rasterizer"), "-c", s, "-o", o])
        objs.append(o)
    _run(["nvcc", "-shared", *ARCH, "-o", so, *objs])
    return so


def build_torch_ext(force=False):
    """reference torch extension -> oracle/_ref/diff_gaussian_rasterization/_C.so"""
    import torch  # noqa: F401
    from torch.utils import cpp_extension as ce
    pkg = os.path.join(OUT, "diff_gaussian_rasterization")
    os.makedirs(pkg, exist_ok=True)
    so = os.path.join(pkg, "_C.so")
    if not force and os.path.exists(so):
        return so
    bdir = os.path.join(OUT, "build_torch")
    os.makedirs(bdir, exist_ok=True)
    inc = []
    for p in ce.include_paths("cuda"):
        inc += ["-I", p]
    inc += ["-I", sysconfig.get_paths()["include"], "-I", os.path.join(REF, "third_party", "glm")]
    defs = ["-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
            "-D__CUDA_NO_HALF_OPERATORS__", "-D__CUDA_NO_HALF_CONVERSIONS__",
            "-D__CUDA_NO_BFLOAT16_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__"]
    objs = []
    procs = []
    for s in ("cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
              "cuda_rasterizer/backward.cu", "rasterize_points.cu"):
        o = os.path.join(bdir, os.path.basename(s) + ".o")
        cmd = ["nvcc", "-std=c++17", "--expt-relaxed-constexpr", *ARCH, "-Xcompiler", "-fPIC",
               "-include", "cstdint", *defs, *inc, "-c", os.path.join(REF, s), "-o", o]
        print("+", " ".join(cmd), flush=True)
        procs.append(subprocess.Popen(cmd))
        objs.append(o)
    o = os.path.join(bdir, "ext.o")
    cmd = ["g++", "-std=c++17", "-O2", "-fPIC", *defs, *inc, "-c", os.path.join(REF, "ext.cpp"), "-o", o]
    print("+", " ".join(cmd), flush=True)
    procs.append(subprocess.Popen(cmd))
    objs.append(o)
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("reference torch extension failed to compile")
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    _run(["g++", "-shared", "-o", so, *objs, "-L", libdir, "-L", "/usr/local/cuda/lib64",
          "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch", "-ltorch_python", "-lcudart",
          "-Wl,-rpath," + libdir])
    # install step: the package's python wrapper travels with its extension (git-ignored)
    shutil.copyfile(os.path.join(REF, "diff_gaussian_rasterization", "__init__.py"),
                    os.path.join(pkg, "__init__.py"))
    return so


if __name__ == "__main__":
    if not available():
        print("reference not mounted; nothing to build")
        sys.exit(0)
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("shim", "all"):
        build_shim(force="--force" in sys.argv)
    if what in ("torch", "all"):
        build_torch_ext(force="--force" in sys.argv)
