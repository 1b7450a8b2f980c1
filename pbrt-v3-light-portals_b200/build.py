"""In-tree build of the native libraries.

  libpbrthost.so  host front end (parser, BVH build, flatten, film/image IO)   g++
  libpbrtgpu.so   sm_100a CUDA kernels + the C ABI of include/pbrtgpu.h        nvcc
  pbrt_gpu        pbrt-compatible command line (`pbrt_gpu scene.pbrt`)         g++

Arithmetic flags matter for parity: the reference is built without FMA
contraction (gcc -O3, no -march), so host code uses -ffp-contract=off and
device code -fmad=false; nothing uses fast-math.
"""
import glob
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OUT = os.path.join(PKG, "lib")

HOST_FLAGS = ["-std=c++17", "-O2", "-g", "-fPIC", "-ffp-contract=off", "-Wall", "-Wno-unused-function", "-pthread"]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "--expt-relaxed-constexpr",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout)
        raise RuntimeError("build failed: " + os.path.basename(cmd[-1]))
    return r.stdout


def _deps():
    hdrs = glob.glob(os.path.join(CSRC, "**", "*.h"), recursive=True)
    hdrs += glob.glob(os.path.join(CSRC, "**", "*.cuh"), recursive=True)
    hdrs += glob.glob(os.path.join(CSRC, "**", "*.inc"), recursive=True)
    hdrs += glob.glob(os.path.join(ROOT, "include", "*.h"))
    return hdrs


def build_host(verbose=False):
    os.makedirs(OUT, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "host", "*.cpp")))
    srcs = [s for s in srcs if os.path.basename(s) != "main.cpp"]
    target = os.path.join(OUT, "libpbrthost.so")
    if _newer(target, srcs + _deps()):
        cxx = shutil.which("g++") or "g++"
        _run([cxx] + HOST_FLAGS + ["-shared", "-I" + os.path.join(ROOT, "include")] + srcs + ["-o", target])
        if verbose:
            print("built", target)
    return target


def build_gpu(verbose=False):
    os.makedirs(OUT, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "device", "*.cu")))
    target = os.path.join(OUT, "libpbrtgpu.so")
    if srcs and _newer(target, srcs + _deps()):
        nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
        extra = os.environ.get("PBRTGPU_EXTRA_NVCC_FLAGS", "").split()  # tuning experiments, e.g. -DPG_PERSIST_MIN_BLOCKS=8
        log = _run([nvcc] + NVCC_FLAGS + extra + ["-Xptxas", "-v", "-shared", "-I" + os.path.join(ROOT, "include")] + srcs +
                   ["-o", target, "-lcudart", "-lpthread"])
        with open(os.path.join(OUT, "ptxas.log"), "w") as f:
            f.write(log)
        if verbose:
            print("built", target)
    return target


def build_cli(verbose=False):
    main = os.path.join(CSRC, "host", "main.cpp")
    if not os.path.exists(main):
        return None
    target = os.path.join(OUT, "pbrt_gpu")
    if _newer(target, [main] + _deps() + [os.path.join(OUT, "libpbrthost.so"), os.path.join(OUT, "libpbrtgpu.so")]):
        cxx = shutil.which("g++") or "g++"
        _run([cxx] + HOST_FLAGS + ["-I" + os.path.join(ROOT, "include"), main, "-L" + OUT, "-lpbrthost", "-lpbrtgpu",
                                   "-Wl,-rpath,$ORIGIN", "-o", target])
        if verbose:
            print("built", target)
    return target


def build_all(verbose=False):
    build_host(verbose)
    build_gpu(verbose)
    build_cli(verbose)


if __name__ == "__main__":
    build_all(verbose=True)
