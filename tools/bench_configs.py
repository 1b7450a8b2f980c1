"""Device-timed throughput of the BASELINE.json configurations other than the bench.py workload
(run on the GPU box).  Prints one JSON line per configuration.

  C1  scenes/cornell_dielectric.pbrt, Integrator "path" maxdepth 5, 64 spp (512x512, gaussian filter)
  C3  tests/scenes/portal_room.pbrt at 512x512, 64 spp, strategy light | portal | projection | solidangle | solidangle_mis
  C5  100 M closest-hit and any-hit rays against the killeroo BVH (tools/bench_rays.py covers this one)
"""
import importlib, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
pk = importlib.import_module("pbrt-v3-light-portals_b200")


ONLY = [a.lower() for a in sys.argv[1:]]  # optional substrings: run only the configurations whose name contains one of them


def run(name, text, base_dir, repeats=3):
    if ONLY and not any(o in name.lower() for o in ONLY):
        return
    hs = pk.HostScene(text=text, base_dir=base_dir)
    ctx = pk.GpuContext(0)
    ctx.upload(hs)
    best = None
    for _ in range(repeats + 1):  # first pass warms up
        _, st = ctx.render_device(hs.params)
        best = st.device_ms if best is None else min(best, st.device_ms)
    ctx.set_profile(True)
    ctx.render_device(hs.params)
    prof = ctx.get_profile()
    ctx.set_profile(False)
    tot = sum(v["ms"] for v in prof.values()) or 1.0
    sb = hs.params.sample_bounds
    samples = (sb[2] - sb[0]) * (sb[3] - sb[1]) * hs.params.spp
    print(json.dumps({"config": name, "samples": samples, "device_ms": round(best, 3), "Msamples_per_s": round(samples / best / 1e3, 1),
                      "rays_per_sample": round((st.closest_rays + st.shadow_rays) / samples, 2), "launches": st.kernel_launches,
                      "kernel_ms": {k: round(v["ms"], 2) for k, v in prof.items()}}), flush=True)


if __name__ == "__main__":
    ref_scenes = os.path.join(ROOT, "oracle", "_ref", "scenes")
    p = os.path.join(ref_scenes, "cornell-path.pbrt")
    if os.path.exists(p):
        t = open(p).read()
        run("C1 cornell_dielectric path 512x512 64spp", bench.scene_variant(p, 64, (512, 512)), ref_scenes)
        run("C1 cornell_dielectric path 1024x1024 256spp", bench.scene_variant(p, 256, (1024, 1024)), ref_scenes)
        t2 = re.sub(r'PixelFilter\s+"gaussian"\s+"float xwidth"\s*\[2\]\s+"float ywidth"\s*\[2\]', 'PixelFilter "box"', bench.scene_variant(p, 256, (1024, 1024)))
        run("C1 variant: box filter, 1024x1024 256spp", t2, ref_scenes)
    room = os.path.join(ROOT, "tests", "scenes", "portal_room.pbrt")
    for s in ("light", "portal", "projection", "solidangle", "solidangle_mis"):  # the last two are not the reference's
        t = bench.scene_variant(room, 64, (512, 512)).replace("STRATEGY", s).replace('"integer maxdepth" [4]', '"integer maxdepth" [5]')
        run("C3 portal room 512x512 64spp strategy " + s, t, os.path.dirname(room))
