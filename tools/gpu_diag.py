"""GPU-box diagnostic: per-sample comparison of the CUDA render with the CPU-stepped device code."""
import ctypes as C
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import util  # noqa: E402
from make_golden_images import CASES, scene_text  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tools"))
from dev_compare import hostsim_lib  # noqa: E402


def main():
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    lib = hostsim_lib()
    ctx = pk.GpuContext(0)
    golden = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
    cases = sys.argv[1:] or sorted(CASES)
    for case in cases:
        hs = pk.HostScene(text=scene_text(case), base_dir=os.path.join(ROOT, "tests", "scenes"))
        ctx.upload(hs)
        film, st = ctx.render(hs.params)
        rgb = hs.film_to_rgb(film)
        print(case, "gpu vs golden", util.image_stats(rgb, golden[case]), flush=True)
        np.save(os.path.join(ROOT, "gpurun_out", "gpu_%s.npy" % case), rgb)
        ndiff = 0
        for s in range(hs.params.spp):
            p = hs.params.copy()
            p.spp_begin, p.spp_end = s, s + 1
            fg, _ = ctx.render(p)
            _, fc, _ = util.hostsim_render(lib, hs, p)
            d = np.abs(fg[..., :3].astype(np.float64) - fc[..., :3]).max(axis=2)
            rel = d / np.maximum(np.abs(fc[..., :3]).max(axis=2), 1e-6)
            bad = np.argwhere(rel > 1e-3)
            ndiff += len(bad)
            for (y, x) in bad[:6]:
                print("   sample %d pixel (%d,%d): gpu %s cpu %s" % (s, x, y, fg[y, x, :3], fc[y, x, :3]))
        print(case, "samples differing by >1e-3:", ndiff, flush=True)


if __name__ == "__main__":
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    main()
