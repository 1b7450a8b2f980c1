"""Render config C1 (Cornell box, Integrator "path" maxdepth 5, 512x512, 64 spp) a few times; used under ncu."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pk = importlib.import_module("pbrt-v3-light-portals_b200")
ref_scenes = os.path.join(ROOT, "oracle", "_ref", "scenes")
p = os.path.join(ref_scenes, "cornell-path.pbrt")
hs = pk.HostScene(text=bench.scene_variant(p, 64, (512, 512)), base_dir=ref_scenes)
ctx = pk.GpuContext(0)
ctx.upload(hs)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    _, st = ctx.render_device(hs.params)
    print(st.device_ms)
