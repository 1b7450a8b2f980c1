"""Device time of one rank's share of a bench.py step at world size N, measured on ONE GPU (shard 0 of N)."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pk = importlib.import_module("pbrt-v3-light-portals_b200")
scene_path, _ = bench.find_scene()
hs = pk.HostScene(text=bench.scene_variant(scene_path, 1024, bench.RES), base_dir=os.path.dirname(scene_path))
ctx = pk.GpuContext(0)
ctx.upload(hs)
for n in (1, 2, 4, 8):
    p = hs.params.copy()
    p.shard_index, p.shard_count = 0, n
    best = None
    for k in range(4):
        p.spp_begin, p.spp_end = 16 * n * k, 16 * n * (k + 1)
        _, st = ctx.render_device(p)
        best = st.device_ms if best is None or k == 0 else min(best, st.device_ms)
    print("world %d: shard 0 renders its step in %.2f ms, %d launches" % (n, best, st.kernel_launches), flush=True)
