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
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=20_000_000)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--oracle-rays", type=int, default=2_000_000, help="rays the instrumented restatement counts V and T on")
    ap.add_argument("--check-rays", type=int, default=-1, help="rays whose results are compared with the REFERENCE's BVHAccel "
                    "(oracle/_ref/ref_harness, all host threads); -1 = the whole batch")
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
    harness = os.path.join(ROOT, "oracle", "_ref", "ref_harness")
    shm = "/dev/shm" if os.path.isdir("/dev/shm") else tempfile.gettempdir()
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
        # the timed batch through the reference's own BVHAccel::Intersect / IntersectP (bvh.cpp:662-738): every hit ID,
        # every t (bit pattern) and every occlusion flag of the launch that was timed must be equal
        k = n if args.check_rays < 0 else min(n, args.check_rays)
        verdict = {"bit_exact": None, "rays_checked": 0}
        if k > 0 and os.path.exists(harness):
            rp, op = os.path.join(shm, "_rays_%d.bin" % os.getpid()), os.path.join(shm, "_out_%d.bin" % os.getpid())
            rays[:k].tofile(rp)
            try:
                r = subprocess.run([harness, "intersect", path, rp, op, "any" if any_hit else "closest", str(os.cpu_count())],
                                   stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, check=True)
                ref_line = json.loads(r.stdout.strip().splitlines()[-1])
                got = d_out.cpu().numpy()
                if any_hit:
                    want = np.fromfile(op, np.uint8)
                    verdict = {"bit_exact": bool(np.array_equal(got[:k], want)), "rays_checked": k, "occluded": int(want.sum())}
                else:
                    want = np.fromfile(op, np.uint32).reshape(-1, 2)
                    g = got.view(np.uint32).reshape(-1, 4)[:k]
                    same_id = bool(np.array_equal(g[:, 0], want[:, 0]))
                    hit = want[:, 0] != 0xFFFFFFFF
                    same_t = bool(np.array_equal(g[hit, 1], want[hit, 1]))
                    verdict = {"bit_exact": same_id and same_t, "ids_equal": same_id, "t_bits_equal": same_t, "rays_checked": k, "hits": int(hit.sum())}
                verdict["reference_cpu_mrays_per_s"] = ref_line["mrays_per_s"]
                verdict["reference_threads"] = ref_line["threads"]
            finally:
                for f in (rp, op):
                    if os.path.exists(f):
                        os.unlink(f)
        bytes_per_ray = 32 + (1 if any_hit else 16) + 32 * V + 48 * T
        best = min(ms)
        print(json.dumps({"set": name, "scene": os.path.basename(path), "rays": n, "ms_best": best, "ms_mean": float(np.mean(ms)),
                          "mrays_per_s": n / best / 1e3, "V_per_ray": V, "T_per_ray": T,
                          "alg_bytes_per_ray": bytes_per_ray, "alg_GBps": n * bytes_per_ray / best / 1e6, **verdict}), flush=True)
        assert verdict["bit_exact"] is not False, (name, verdict)
        del d_rays, d_out


if __name__ == "__main__":
    main()
