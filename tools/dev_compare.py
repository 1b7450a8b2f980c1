"""Developer aid (CPU only): render a scene with the reference binary
(oracle/_ref/pbrt, single thread, PFM) and with the device code stepped on the
CPU (tests/hostsim), and print image-difference statistics.
    python tools/dev_compare.py SCENE [--spp N] [--res W H] [--depth D] [--keep]
"""
import argparse
import ctypes as C
import importlib
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data[::-1].astype(np.float32)


def variant(text, spp=None, res=None, depth=None, integrator=None):
    if spp is not None:
        text = re.sub(r'"integer pixelsamples"\s*\[?\s*\d+\s*\]?', '"integer pixelsamples" [%d]' % spp, text)
    if res is not None:
        text = re.sub(r'"integer xresolution"\s*\[?\s*\d+\s*\]?', '"integer xresolution" [%d]' % res[0], text)
        text = re.sub(r'"integer yresolution"\s*\[?\s*\d+\s*\]?', '"integer yresolution" [%d]' % res[1], text)
    if integrator is not None:
        text = re.sub(r'^Integrator .*$', integrator, text, flags=re.M)
    if depth is not None:
        text = re.sub(r'^(Integrator "path").*$', r'\1 "integer maxdepth" [%d]' % depth, text, flags=re.M)
    return text


def hostsim_lib():
    out = os.path.join(ROOT, "tests", "_build", "libhostsim.so")
    src = os.path.join(ROOT, "tests", "hostsim", "hostsim.cpp")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-pthread", src, "-o", out])
    lib = C.CDLL(out)
    lib.hostsim_create.restype = C.c_void_p
    lib.hostsim_create.argtypes = [C.c_void_p]
    lib.hostsim_destroy.argtypes = [C.c_void_p]
    lib.hostsim_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    return lib


def stats(a, b):
    d = (a - b).astype(np.float64)
    rmse = np.sqrt((d ** 2).mean())
    mean = b.astype(np.float64).mean()
    rel = np.abs(d) / np.maximum(np.abs(b), 1e-3 * max(mean, 1e-6))
    return dict(rmse=rmse, rel_rmse=rmse / max(mean, 1e-12), mean_ref=mean, mean_ours=a.astype(np.float64).mean(),
                max_abs=np.abs(d).max(), frac_rel_gt_1pct=(rel.max(axis=2) > 0.01).mean(), frac_exact=(a == b).all(axis=2).mean())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene")
    ap.add_argument("--spp", type=int)
    ap.add_argument("--res", type=int, nargs=2)
    ap.add_argument("--depth", type=int)
    ap.add_argument("--integrator")
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--keep", action="store_true")
    args = ap.parse_args()
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    src = os.path.abspath(args.scene)
    text = variant(open(src).read(), args.spp, args.res, args.depth, args.integrator)
    tmp = os.path.join(os.path.dirname(src), "_dev_%d.pbrt" % os.getpid())
    open(tmp, "w").write(text)
    try:
        out = tempfile.mktemp(suffix=".pfm")
        t0 = time.time()
        subprocess.check_call([os.path.join(ROOT, "oracle", "_ref", "pbrt"), "--nthreads", "1", "--quiet", "--minloglevel", "2",
                               "--outfile", out, tmp], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        t_ref = time.time() - t0
        ref = read_pfm(out)
        os.unlink(out)
        hs = pk.HostScene(tmp)
    finally:
        if not args.keep:
            os.unlink(tmp)
    lib = hostsim_lib()
    h = C.c_void_p(lib.hostsim_create(C.byref(hs.view)))
    cb = hs.params.cropped_bounds
    W, H = cb[2] - cb[0], cb[3] - cb[1]
    film = np.zeros((H, W, 4), np.float32)
    st = (C.c_int64 * 4)()
    t0 = time.time()
    lib.hostsim_render(h, C.byref(hs.params), film.ctypes.data, st, args.threads)
    t_sim = time.time() - t0
    rgb = hs.film_to_rgb(film)
    print("ref %.1fs  hostsim %.1fs  camera %d closest %d shadow %d scrubbed %d" % (t_ref, t_sim, st[0], st[1], st[2], st[3]))
    for k, v in stats(rgb, ref).items():
        print("  %-18s %.6g" % (k, v))
    np.save("/tmp/dev_ours.npy", rgb)
    np.save("/tmp/dev_ref.npy", ref)


if __name__ == "__main__":
    main()
