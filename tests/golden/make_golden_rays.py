"""Generates tests/golden/tri_soup_*.npz with the REFERENCE's own answers
(oracle/_ref/ref_harness, i.e. the reference's parser, BVHAccel and shapes):
  tri_soup_scene.npz  LinearBVHNode dump + BVH-ordered primitive table
  tri_soup_rays.npz   closest-hit (origId, t) and any-hit bits for four seeded ray sets
Run in the build container (needs /root/reference to have built oracle/_ref):
    python tests/golden/make_golden_rays.py
The ray sets are regenerated from seeds by the tests; a checksum guards drift.
"""
import hashlib
import importlib
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402

HARNESS = os.path.join(ROOT, "oracle", "_ref", "ref_harness")


def ray_sets(view):
    lo, hi = util.world_bounds(view)
    return {
        "incoherent": util.rays_incoherent_closest(view, 40000, seed=42),
        "toward": util.rays_toward_box(40000, [-10, -10, -10], [10, 10, 10], seed=44),
        "edges": util.rays_edge_cases(view, 20000, seed=45),
        "shadow": util.rays_shadow(view, 40000, seed=43),
    }


def main():
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    scene = os.path.join(ROOT, "tests", "scenes", "tri_soup.pbrt")
    hs = pk.HostScene(scene)
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.check_call([HARNESS, "dump", scene, os.path.join(tmp, "d")])
        nodes = np.fromfile(os.path.join(tmp, "d.nodes.bin"), dtype=np.uint8).reshape(-1, 32)
        prims = np.fromfile(os.path.join(tmp, "d.prims.bin"), dtype=np.uint8).reshape(-1, 44)
        np.savez_compressed(os.path.join(ROOT, "tests", "golden", "tri_soup_scene.npz"), nodes=nodes, prims=prims,
                            lights=open(os.path.join(tmp, "d.lights.txt")).read())
        for name, rays in ray_sets(hs.view).items():
            rp = os.path.join(tmp, name + ".bin")
            rays.tofile(rp)
            out[name + "_sha"] = hashlib.sha256(rays.tobytes()).hexdigest()
            subprocess.check_call([HARNESS, "intersect", scene, rp, rp + ".c", "closest", "8"], stdout=subprocess.DEVNULL)
            subprocess.check_call([HARNESS, "intersect", scene, rp, rp + ".a", "any", "8"], stdout=subprocess.DEVNULL)
            ref = np.fromfile(rp + ".c", dtype=np.int32).reshape(-1, 2)
            out[name + "_id"] = ref[:, 0].copy()
            out[name + "_t"] = ref[:, 1].copy().view(np.float32)
            out[name + "_occ"] = np.fromfile(rp + ".a", dtype=np.uint8)
            print(name, len(rays), "hit frac %.3f" % (ref[:, 0] >= 0).mean(), "occ frac %.3f" % out[name + "_occ"].mean())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "tri_soup_rays.npz"), **out)


if __name__ == "__main__":
    main()
