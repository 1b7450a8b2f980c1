"""GPU-box diagnostic: which (pixel, sample number) differ between the CUDA render and the same stage functions
stepped on the CPU (tests/hostsim), and -- with --trace K -- what the first K of them do, vertex by vertex, on
both sides (needs the trace builds: tools/build_variants.py trace="-DPG_DEVICE_TRACE" for the CUDA library; the
CPU side is compiled here with -DPG_HOST_TRACE).

    python tools/gpu_diff_samples.py SCENE --res W H --spp N [--depth D] [--integrator '...'] [--trace K] [--out PREFIX]
    python tools/gpu_diff_samples.py SCENE ... --one X Y S --side gpu|cpu      (internal: one sample, traced)
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import util  # noqa: E402
from dev_compare import hostsim_lib, variant  # noqa: E402

TRACE_GPU_LIB = os.path.join(ROOT, "pbrt-v3-light-portals_b200", "lib", "variants", "libpbrtgpu_trace.so")
TRACE_CPU_LIB = os.path.join(ROOT, "tests", "_build", "libhostsim_trace.so")


def load_scene(pk, a):
    text = variant(open(a.scene).read(), a.spp, a.res, a.depth, a.integrator)
    return pk.HostScene(text=text, base_dir=os.path.dirname(os.path.abspath(a.scene)))


def one_sample(a):
    """Render exactly one (pixel, sample) with tracing; the trace goes to stdout (GPU) / stderr (CPU)."""
    x, y, s = a.one
    if a.side == "gpu":
        os.environ["PBRTGPU_LIBRARY"] = TRACE_GPU_LIB
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    hs = load_scene(pk, a)
    p = hs.params.copy()
    p.pixel_bounds[0], p.pixel_bounds[1], p.pixel_bounds[2], p.pixel_bounds[3] = x, y, x + 1, y + 1
    p.spp_begin, p.spp_end = s, s + 1
    cb = p.cropped_bounds
    if a.side == "gpu":
        ctx = pk.GpuContext(0)
        ctx.upload(hs)
        film, _ = ctx.render(p)
        ctx.close()
    else:
        if not os.path.exists(TRACE_CPU_LIB):
            os.makedirs(os.path.dirname(TRACE_CPU_LIB), exist_ok=True)
            subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", "-DPG_HOST_TRACE",
                                   os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp"), "-o", TRACE_CPU_LIB])
        lib = C.CDLL(TRACE_CPU_LIB)
        lib.hostsim_create.restype = C.c_void_p
        lib.hostsim_create.argtypes = [C.c_void_p]
        lib.hostsim_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        h = C.c_void_p(lib.hostsim_create(C.byref(hs.view)))
        film = np.zeros((cb[3] - cb[1], cb[2] - cb[0], 4), np.float32)
        st = (C.c_int64 * 4)()
        lib.hostsim_render(h, C.byref(p), film.ctypes.data, st, 1)
    sys.stdout.flush()
    print("film[%d,%d] = %r" % (y, x, [float(v) for v in film[y - cb[1], x - cb[0]]]), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene")
    ap.add_argument("--res", type=int, nargs=2)
    ap.add_argument("--spp", type=int, default=4)
    ap.add_argument("--depth", type=int)
    ap.add_argument("--integrator")
    ap.add_argument("--trace", type=int, default=0)
    ap.add_argument("--one", type=int, nargs=3)
    ap.add_argument("--side", default="gpu")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "diff"))
    a = ap.parse_args()
    if a.one:
        return one_sample(a)
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    hs = load_scene(pk, a)
    lib = hostsim_lib()
    ctx = pk.GpuContext(0)
    ctx.upload(hs)
    found, total = [], 0
    for s in range(hs.params.spp):
        p = hs.params.copy()
        p.spp_begin, p.spp_end = s, s + 1
        fg, _ = ctx.render(p)
        _, fc, _ = util.hostsim_render(lib, hs, p, threads=os.cpu_count())
        d = np.abs(fg[..., :3].astype(np.float64) - fc[..., :3]).max(axis=2)
        rel = d / np.maximum(np.abs(fc[..., :3]).max(axis=2), 1e-6)
        bad = np.argwhere(rel > 1e-3)
        total += len(bad)
        cb = hs.params.cropped_bounds
        for (y, x) in bad[:40]:
            found.append({"sample": s, "x": int(x) + cb[0], "y": int(y) + cb[1], "gpu": [float(v) for v in fg[y, x]], "cpu": [float(v) for v in fc[y, x]]})
        print("sample %d: %d pixels differ by > 1e-3" % (s, len(bad)), flush=True)
    ctx.close()
    cb = hs.params.cropped_bounds
    json.dump({"scene": a.scene, "res": a.res, "spp": a.spp, "pixels": (cb[2] - cb[0]) * (cb[3] - cb[1]), "differing": total, "examples": found},
              open(a.out + "_samples.json", "w"), indent=1)
    print("total differing (pixel, sample):", total, flush=True)
    base = [sys.executable, os.path.abspath(__file__), a.scene, "--spp", str(a.spp)]
    if a.res:
        base += ["--res", str(a.res[0]), str(a.res[1])]
    if a.depth is not None:
        base += ["--depth", str(a.depth)]
    if a.integrator:
        base += ["--integrator", a.integrator]
    for k, e in enumerate(found[:a.trace]):
        for side in ("gpu", "cpu"):
            r = subprocess.run(base + ["--one", str(e["x"]), str(e["y"]), str(e["sample"]), "--side", side], stdout=subprocess.PIPE,
                               stderr=subprocess.STDOUT, text=True)
            open("%s_trace%d_%s.log" % (a.out, k, side), "w").write("# pixel (%d, %d) sample %d\n" % (e["x"], e["y"], e["sample"]) + r.stdout)


if __name__ == "__main__":
    main()
