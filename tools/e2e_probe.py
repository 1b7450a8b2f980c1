"""Where the end-to-end step of bench.py spends its wall time (run on the GPU box)."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pk = importlib.import_module("pbrt-v3-light-portals_b200")
scene_path, _ = bench.find_scene()
hs = pk.HostScene(text=bench.scene_variant(scene_path, 1024, bench.RES), base_dir=os.path.dirname(scene_path))
ctx = pk.GpuContext(0)
p = hs.params.copy()
for k in range(8):
    p.spp_begin, p.spp_end = 16 * k, 16 * k + 16
    t0 = time.perf_counter(); ctx.upload(hs); t1 = time.perf_counter(); film, st = ctx.render(p); t2 = time.perf_counter()
    print("iter %d upload %.1f ms render %.1f ms (device %.1f ms)" % (k, (t1 - t0) * 1e3, (t2 - t1) * 1e3, st.device_ms), flush=True)
