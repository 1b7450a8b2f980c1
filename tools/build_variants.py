"""Developer aid: build libpbrtgpu.so variants with extra nvcc flags for A/B runs on the GPU box.

    python tools/build_variants.py name="-DPG_NI_BSDF=PG_HD -DPG_NI_LOBE=PG_HD" other="..."

-> pbrt-v3-light-portals_b200/lib/variants/libpbrtgpu_<name>.so (git-ignored, travels with gpurun);
select one with PBRTGPU_LIBRARY=<path> (see __init__.gpu_lib).
"""
import glob
import importlib
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b = importlib.import_module("pbrt-v3-light-portals_b200.build")


def main():
    import concurrent.futures
    out = os.path.join(b.OUT, "variants")
    os.makedirs(out, exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(b.CSRC, "device", "*.cu")))
    def one(arg):
        name, flags = arg.split("=", 1)
        target = os.path.join(out, "libpbrtgpu_%s.so" % name)
        log = b._run(["nvcc"] + b.NVCC_FLAGS + flags.split() + ["-Xptxas", "-v", "-shared", "-I" + os.path.join(ROOT, "include")] + srcs +
                     ["-o", target, "-lcudart", "-lpthread"])
        open(target + ".ptxas.log", "w").write(log)
        return name + " -> " + target
    with concurrent.futures.ThreadPoolExecutor(max_workers=4) as ex:  # one nvcc per variant, side by side
        for line in ex.map(one, sys.argv[1:]):
            print(line)


if __name__ == "__main__":
    main()
