"""SURVEY.md 8(f) rank 3 measured: Accelerator "bvh" "splitmethod" "hlbvh" built on the host (csrc/host/bvh_build.cpp, the
reference's algorithm on one core) against pbrtgpu_build_hlbvh on the GPU, and pbrtgpu_upload_scene with the layout
conversion on the host (PBRTGPU_LAYOUT=host) against the conversion kernels -- on the C4-shaped scene of tools/gen_c4.py.
Every timed result is checked: the two trees and primitive orders must be identical, and so must the device layouts.

    python tools/bench_build.py [--quads 2236] [--repeat 3]      (one JSON line per measurement)
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quads", type=int, default=2236)
    ap.add_argument("--repeat", type=int, default=3)
    args = ap.parse_args()
    pk = importlib.import_module("pbrt-v3-light-portals_b200")
    d = os.path.join(tempfile.gettempdir(), "pbrtgpu_c4_%d" % args.quads)
    if not os.path.exists(os.path.join(d, "c4.pbrt")):
        subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_c4.py"), d, "--quads", str(args.quads)], stdout=subprocess.DEVNULL)
    text = open(os.path.join(d, "c4.pbrt")).read().replace('"string splitmethod" "sah"', '"string splitmethod" "hlbvh"')
    assert '"hlbvh"' in text
    ctx = pk.GpuContext(0)

    def load(device_build):
        ctx.install_hlbvh_builder(device_build)
        try:
            t0 = time.perf_counter()
            hs = pk.HostScene(text=text, base_dir=d)
            load_ms = (time.perf_counter() - t0) * 1e3
            wall, dev = pk.last_bvh_build_ms()
        finally:
            ctx.install_hlbvh_builder(False)
        return hs, load_ms, wall, dev

    host, load_h, wall_h, _ = load(False)
    v = host.view
    tris = v.n_prims
    print(json.dumps({"what": "hlbvh build, host (one core)", "triangles": tris, "nodes": v.n_nodes, "build_ms": round(wall_h, 1), "scene_load_ms": round(load_h, 1)}), flush=True)
    best = None
    for _ in range(args.repeat):
        dev, load_d, wall_d, kern_d = load(True)
        if best is None or wall_d < best[0]:
            best = (wall_d, kern_d, load_d)
        w = dev.view
        def arr(view, f, k):
            a = np.asarray(view.array(f, view.n_nodes if f.startswith("bvh") else view.n_prims, k))
            return a.view(np.uint32) if a.dtype == np.float32 else a
        same = w.n_nodes == v.n_nodes and all(np.array_equal(arr(v, f, k), arr(w, f, k)) for f, k in
                                              (("bvh_bounds", 6), ("bvh_offset", 1), ("bvh_nprims", 1), ("bvh_axis", 1), ("prim_orig_id", 1)))
        assert same, "the device-built tree differs from the host-built one"
        if _ + 1 < args.repeat:
            dev.close()
    print(json.dumps({"what": "hlbvh build, pbrtgpu_build_hlbvh (H2D of the bounds + kernels + upper SAH on the host + D2H of the tree)", "triangles": tris,
                      "build_ms": round(best[0], 1), "kernels_ms": round(best[1], 2), "scene_load_ms": round(best[2], 1),
                      "identical_to_host_build": True, "speedup_vs_host_build": round(wall_h / best[0], 1)}), flush=True)

    def upload_ms(c, hs):
        ts = []
        for _ in range(args.repeat + 1):
            t0 = time.perf_counter()
            c.upload(hs)
            ts.append((time.perf_counter() - t0) * 1e3)
        return min(ts[1:]), ts[0]
    up_d, first_d = upload_ms(ctx, host)
    nodes_d, tris_d = ctx.read_layout(host)
    os.environ["PBRTGPU_LAYOUT"] = "host"
    other = pk.GpuContext(0)
    del os.environ["PBRTGPU_LAYOUT"]
    up_h, first_h = upload_ms(other, host)
    nodes_h, tris_h = other.read_layout(host)
    assert np.array_equal(nodes_d, nodes_h) and np.array_equal(tris_d, tris_h), "device-built layout differs from the host-built one"
    other.close()
    print(json.dumps({"what": "pbrtgpu_upload_scene", "triangles": tris, "layout_on_host_ms": round(up_h, 1), "layout_on_device_ms": round(up_d, 1),
                      "first_call_ms": {"host": round(first_h, 1), "device": round(first_d, 1)}, "identical_layouts": True,
                      "speedup": round(up_h / up_d, 2)}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
