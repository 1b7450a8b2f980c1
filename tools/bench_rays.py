"""Ray-batch benchmark (config C5): Mrays/s of the closest-hit and any-hit
kernels on device-resident ray buffers, against the killeroo BVH if staged,
else tests/scenes/tri_soup.pbrt.  Prints one JSON line per ray set.
    python tools/bench_rays.py [--rays N] [--reps K]
"""
import argparse
import ctypes as C
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=20_000_000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--oracle-rays", type=int, default=2_000_000)
    args = ap.parse_args()
    import torch
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    kil = os.path.join(ROOT, "oracle", "_ref", "scenes", "killeroo-path.pbrt")
    path = kil if os.path.exists(kil) else os.path.join(ROOT, "tests", "scenes", "tri_soup.pbrt")
    hs = pk.HostScene(path)
    ctx = pk.GpuContext(0)
    ctx.upload(hs)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    n = args.rays
    box = ([-100, -100, -150], [100, 150, 50]) if path == kil else ([-10, -10, -10], [10, 10, 10])
    sets = {
        "incoherent_world": (util.rays_incoherent_closest(hs.view, n, seed=42), False),
        "toward_object": (util.rays_toward_box(n, box[0], box[1], seed=44), False),
        "shadow_world": (util.rays_shadow(hs.view, n, seed=43), True),
        "shadow_object": (util.rays_shadow(hs.view, n, seed=46, lo=np.array(box[0], float) * 1.5, hi=np.array(box[1], float) * 1.5), True),
    }
    oracle = C.CDLL(os.path.join(ROOT, "oracle", "_build", "liboracle_port.so"))
    for name, (rays, any_hit) in sets.items():
        d_rays = torch.from_numpy(rays).cuda()
        d_out = torch.empty(n * (1 if any_hit else 16), dtype=torch.uint8, device="cuda")
        ms = []
        for _ in range(args.reps + 2):
            ms.append(ctx.intersect_device(d_rays.data_ptr(), n, d_out.data_ptr(), any_hit))
        ms = ms[2:]
        m = args.oracle_rays
        if any_hit:
            _, (V, T) = util.oracle_any(oracle, hs.view, rays[:m], threads=os.cpu_count())
        else:
            _, _, _, (V, T) = util.oracle_closest(oracle, hs.view, rays[:m], threads=os.cpu_count())
        V, T = V / m, T / m
        bytes_per_ray = 32 + (1 if any_hit else 16) + 32 * V + 48 * T
        best = min(ms)
        print(json.dumps({"set": name, "scene": os.path.basename(path), "rays": n, "ms_best": best, "ms_mean": float(np.mean(ms)),
                          "mrays_per_s": n / best / 1e3, "V_per_ray": V, "T_per_ray": T,
                          "alg_bytes_per_ray": bytes_per_ray, "alg_GBps": n * bytes_per_ray / best / 1e6}), flush=True)
        del d_rays, d_out


if __name__ == "__main__":
    main()
