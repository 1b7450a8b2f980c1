"""Tier-2 parity of the CUDA path on BASELINE.json's own configurations, at sizes that exercise the batching
machinery (multi-million-slot batches, several batches per frame, the spp-chunk loop, the two-ended hit queue,
the fused traversal launch over three concatenated queues, the 16-byte film reductions, CUDA's libm).

The oracle is the reference itself, run HERE, now: oracle/_ref/pbrt (built from /root/reference by
oracle/Makefile; it travels to the GPU box, /root/reference does not), float PFM output.  It runs on all host
threads except where the reference's own data race matters (PortalArealight::selectedPortal is a
shared mutable member, so scenes whose light has more than one portal are rendered with --nthreads 1).  A
multi-threaded reference image differs from the single-threaded one only in the order tiles are merged
(~1e-7 relative).

Per-case tolerances are 4 .. 10 x the value observed on a B200 (profiles/r02_parity_stats.jsonl), not a blanket
figure.  The CUDA path differs from the reference in two things only: the order of float atomics into a pixel
(~1e-7), and the last bit of CUDA's sinf / cosf / acosf / atan2f against glibc's.  The second one moves a few
samples across DISCRETE decisions, each of which changes one pixel by 1/spp of a sample's radiance:
  * lamp / sandbox under directlighting: no such decision is reachable -> rel RMSE 2e-8 .. 2e-7;
  * glass (Cornell: FresnelSpecular picks reflection or transmission by `u < F`), microfacet lobe choice (plastic
    clutter in C3 / C4): a handful of samples -> rel RMSE 2e-5 .. 2e-4;
  * killeroo: 0.1 % of the SAMPLES, because its light is a radius-3 sphere seen from up to 900 units away and a
    shadow ray towards a sample near the sphere's limb is cut off -- or not -- by the sphere's own ill-conditioned
    quadratic (traced in tests/test_render.py::test_one_ulp_of_libm_moves_killeroo_samples, which reproduces the
    effect on the CPU with a libm nudged by one ulp).  The flips go both ways: the image mean agrees to 1e-8,
    and `mean_rel` is what the killeroo case is held to (a lobe biased by 0.05 % would move it by 1e-4).
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import util
from conftest import REF_DIR, REF_SCENES, ROOT

sys.path.insert(0, os.path.join(ROOT, "tools"))
from dev_compare import read_pfm  # noqa: E402
import gen_c3  # noqa: E402

pytestmark = pytest.mark.gpu
PBRT = os.path.join(REF_DIR, "pbrt")
CORES = os.cpu_count() or 1


def _variant(text, spp=None, res=None, integrator=None):
    if spp is not None:
        text = re.sub(r'"integer pixelsamples"\s*\[?\s*\d+\s*\]?', '"integer pixelsamples" [%d]' % spp, text)
    if res is not None:
        text = re.sub(r'"integer xresolution"\s*\[?\s*\d+\s*\]?', '"integer xresolution" [%d]' % res[0], text)
        text = re.sub(r'"integer yresolution"\s*\[?\s*\d+\s*\]?', '"integer yresolution" [%d]' % res[1], text)
    if integrator is not None:
        text = re.sub(r'^Integrator .*$', integrator, text, flags=re.M)
    return text


_REF_CACHE = {}


def reference_image(text, base_dir, threads):
    """Renders `text` with the reference binary; the scene file is written next to the files it includes.
    Images are kept for the session (the variance test reuses the parity tests' renders)."""
    key = (hash(text), base_dir, threads)
    if key not in _REF_CACHE:
        _REF_CACHE[key] = _reference_image(text, base_dir, threads)
    return _REF_CACHE[key]


def _reference_image(text, base_dir, threads):
    if not os.path.exists(PBRT):
        pytest.skip("oracle/_ref/pbrt not built (run __graft_entry__.build() where /root/reference exists)")
    tmp = os.path.join(base_dir, "_parity_%d.pbrt" % os.getpid())
    out = os.path.join(tempfile.gettempdir(), "_parity_%d.pfm" % os.getpid())
    open(tmp, "w").write(text)
    try:
        subprocess.check_call([PBRT, "--nthreads", str(threads), "--quiet", "--minloglevel", "2", "--outfile", out, tmp],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return read_pfm(out)
    finally:
        os.unlink(tmp)
        if os.path.exists(out):
            os.unlink(out)


def gpu_image(pkg, gpu, text, base_dir, batch_slots=None, monkeypatch=None):
    hs = pkg.HostScene(text=text, base_dir=base_dir)
    gpu.upload(hs)
    if batch_slots is not None:
        monkeypatch.setenv("PBRTGPU_BATCH_SLOTS", str(batch_slots))
    film, st = gpu.render(hs.params)
    if batch_slots is not None:
        monkeypatch.delenv("PBRTGPU_BATCH_SLOTS")
    sb = hs.params.sample_bounds
    assert st.camera_rays == (sb[2] - sb[0]) * (sb[3] - sb[1]) * hs.params.spp
    assert st.scrubbed == 0 and st.truncated_paths == 0 and st.nee_rays_dropped == 0
    return hs.film_to_rgb(film), st


def check(case, ours, ref, tol, extra=None):
    s = util.image_stats(ours, ref)
    rec = dict(case=case, **s, **(extra or {}))
    log_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(log_dir):
        with open(os.path.join(log_dir, "parity_stats.jsonl"), "a") as f:
            f.write(json.dumps(rec) + "\n")
    for k, v in tol.items():
        assert s[k] <= v, rec
    return s


# ---------------------------------------------------------------- C1: Cornell box, 512 x 512, 64 spp, gaussian filter
def test_c1_cornell_512_64spp(pkg, gpu, monkeypatch):
    """scenes/cornell_dielectric.pbrt:15-16 with Integrator "path" maxdepth 5 and 64 spp:
    17.0 M camera samples -- one 17 M-slot batch, and again cut into 2 M-slot batches (9 batches, 2 spp chunks)."""
    src = os.path.join(REF_SCENES, "cornell-path.pbrt")
    if not os.path.exists(src):
        pytest.skip("reference scenes not staged")
    text = _variant(open(src).read(), spp=64, res=(512, 512))
    ref = reference_image(text, REF_SCENES, CORES)
    ours, st = gpu_image(pkg, gpu, text, REF_SCENES)
    check("c1_cornell_512_64spp", ours, ref, dict(rel_rmse=3e-4, bad_frac=1e-4, mean_rel=5e-6), dict(launches=st.kernel_launches))
    many, st2 = gpu_image(pkg, gpu, text, REF_SCENES, batch_slots=2 << 20, monkeypatch=monkeypatch)
    assert st2.kernel_launches > 3 * st.kernel_launches  # really several batches
    assert (st2.camera_rays, st2.closest_rays, st2.shadow_rays) == (st.camera_rays, st.closest_rays, st.shadow_rays)
    check("c1_cornell_512_64spp_2M_batches", many, ref, dict(rel_rmse=3e-4, bad_frac=1e-4, mean_rel=5e-6), dict(launches=st2.kernel_launches))
    # the batching itself changes no path: only the order of float atomics into a pixel
    check("c1_batches_vs_one_batch", many, ours, dict(rel_rmse=4e-5))  # 25-pixel Gaussian footprint: 25 atomics per sample


# ---------------------------------------------------------------- C2: killeroo 1920 x 1080, one bench step (16 spp)
def test_c2_killeroo_1080p_bench_step(pkg, gpu, monkeypatch):
    """scenes/killeroo-simple.pbrt:13-24 with Integrator "path" at 1920 x 1080, 16 spp: exactly bench.py's step
    (one 33 M-slot batch), and the same frame in 6 M-slot batches."""
    src = os.path.join(REF_SCENES, "killeroo-path.pbrt")
    if not os.path.exists(src):
        pytest.skip("reference scenes not staged")
    text = _variant(open(src).read(), spp=16, res=(1920, 1080))
    ref = reference_image(text, REF_SCENES, CORES)
    ours, st = gpu_image(pkg, gpu, text, REF_SCENES)
    check("c2_killeroo_1080p_16spp", ours, ref, dict(rel_rmse=5e-4, bad_frac=0.03, mean_rel=1e-6), dict(launches=st.kernel_launches))
    many, st2 = gpu_image(pkg, gpu, text, REF_SCENES, batch_slots=6 << 20, monkeypatch=monkeypatch)
    assert st2.kernel_launches > 3 * st.kernel_launches
    check("c2_killeroo_1080p_16spp_6M_batches", many, ref, dict(rel_rmse=5e-4, bad_frac=0.03, mean_rel=1e-6))
    check("c2_batches_vs_one_batch", many, ours, dict(rel_rmse=4e-6, bad_frac=0.0))  # the same paths: only the order of atomics


# ---------------------------------------------------------------- the fork's shipped portal scenes, unmodified
@pytest.mark.parametrize("name,integrator", [("lamp/lamp.pbrt", None),                     # directlighting "all", maxdepth 100, 2 portals, projection
                                             ("lamp/lamp.pbrt", 'Integrator "path"'),      # the same under the path integrator, maxdepth 100
                                             ("creeper/sandbox.pbrt", None),               # directlighting maxdepth 10, point light + 3 portals
                                             ("creeper/sandbox.pbrt", 'Integrator "path"')])
def test_shipped_portal_scenes(pkg, gpu, name, integrator):
    """scenes/blender/lamp/out/lamp.pbrt:9 and scenes/blender/creeper/out/sandbox.pbrt as shipped (500 x 500 and
    700 x 700, 5 spp): several portals per light, so PortalArealight::EstimateDirect's portal selection
    (portal_arealight.cpp:39-100, Distribution1D over the in-front portals, SampleDiscrete(u1.x)) runs on the GPU
    with nPortals > 1.  Reference on one thread (its selectedPortal member is racy)."""
    src = os.path.join(REF_SCENES, name)
    if not os.path.exists(src):
        pytest.skip("reference scenes not staged")
    text = _variant(open(src).read(), integrator=integrator)
    hs = pkg.HostScene(text=text, base_dir=os.path.dirname(src))
    lpb = hs.view.array("light_portal_begin", hs.view.n_lights + 1)
    assert int(np.diff(lpb).max()) > 1, "the scene should have a light with several portals"
    del hs
    ref = reference_image(text, os.path.dirname(src), 1)
    ours, st = gpu_image(pkg, gpu, text, os.path.dirname(src))
    case = name.split("/")[1].replace(".pbrt", "") + ("_path" if integrator else "_as_shipped")
    # observed: lamp_as_shipped 1.9e-7, sandbox_as_shipped 1.6e-8, sandbox_path 7.9e-8; lamp_path 4.4e-4 with 2 of 250 000
    # pixels off (metal lampshade, maxdepth 100: two samples take another microfacet branch)
    tol = dict(rel_rmse=4e-3, bad_frac=1e-4, mean_rel=5e-6) if case == "lamp_path" else dict(rel_rmse=2e-6, bad_frac=1e-5, mean_rel=1e-7)
    check(case, ours, ref, tol, dict(launches=st.kernel_launches))


# ---------------------------------------------------------------- C3: the skylight room as SURVEY 8(d) specifies it
C3_TOL = dict(rel_rmse=6e-4, bad_frac=4e-4, mean_rel=2e-6)  # observed at 16 spp: 5e-5 .. 8e-5, 2e-5 .. 4e-5, <= 1.3e-7


@pytest.mark.parametrize("strategy", ["light", "portal", "projection"])
def test_c3_room_three_spp_levels(pkg, gpu, strategy):
    """tools/gen_c3.py: 20 000-triangle clutter (seed 7), spp in {16, 64, 256}, the three reference strategies;
    images against the reference at every level.  512 x 512 at 16 and 64 spp; the 256-spp level at 256 x 256 (the
    reference needs 140 s of 8 cores for the full frame at that level)."""
    for spp, res in ((16, 512), (64, 512), (256, 256)):
        text = gen_c3.scene_text(strategy, spp, res, "ceiling")
        ref = reference_image(text, tempfile.gettempdir(), CORES)
        ours, _ = gpu_image(pkg, gpu, text, tempfile.gettempdir())
        check("c3_%s_%dspp" % (strategy, spp), ours, ref, C3_TOL)


@pytest.mark.parametrize("window", ["wall_x", "wall_y"])
def test_c3_room_wall_window(pkg, gpu, window):
    """The wall-window variants (portal axis 0 / 1, emitter `twosided`): AAPlaneShape::Intersect's hard-coded
    dpdu / dpdv make the emitter's normal +-z whatever its axis."""
    for strategy in ("portal", "projection"):
        text = gen_c3.scene_text(strategy, 16, 512, window)
        ref = reference_image(text, tempfile.gettempdir(), CORES)
        ours, _ = gpu_image(pkg, gpu, text, tempfile.gettempdir())
        check("c3_%s_%s" % (window, strategy), ours, ref, C3_TOL)


def test_c3_portal_variance_no_worse_than_reference(pkg, gpu):
    """BASELINE.json tier 2, second half, on C3 proper: per-pixel MSE against a 16 384-spp image of strategy
    "light" (rendered here; that strategy is parity-checked above), for our 16 / 64-spp images and the reference's."""
    conv, _ = gpu_image(pkg, gpu, gen_c3.scene_text("light", 16384, 512, "ceiling"), tempfile.gettempdir())
    conv = conv.astype(np.float64)
    mse = {}
    for strategy in ("light", "portal"):
        for spp in (16, 64):
            text = gen_c3.scene_text(strategy, spp, 512, "ceiling")
            ours, _ = gpu_image(pkg, gpu, text, tempfile.gettempdir())
            ref = reference_image(text, tempfile.gettempdir(), CORES)
            a, b = float(np.mean((ours - conv) ** 2)), float(np.mean((ref - conv) ** 2))
            mse[(strategy, spp)] = (a, b)
            assert a <= b * (1 + 1e-3), (strategy, spp, a, b)
    assert mse[("portal", 16)][0] < mse[("light", 16)][0], mse  # the emitter is 16 x the hole: the portal pays
    assert mse[("portal", 64)][0] < mse[("portal", 16)][0], mse


# ---------------------------------------------------------------- C4-shaped: 500 k triangles, 4 portal skylights, >= 3 batches
def test_c4_shaped_scene_in_several_batches(pkg, gpu, monkeypatch, tmp_path):
    """tools/gen_c4.py --quads 500 (500 000 displaced-grid triangles as two PLY meshes, four portal lights, strategy
    "portal", splitmethod hlbvh as C4 allows) at 640 x 360, 16 spp, in 1 M-slot batches (4 batches)."""
    d = str(tmp_path)
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_c4.py"), d, "--quads", "500", "--res", "640", "360", "--spp", "16"],
                          stdout=subprocess.DEVNULL)
    text = open(os.path.join(d, "c4.pbrt")).read().replace('"string splitmethod" "sah"', '"string splitmethod" "hlbvh"')
    ref = reference_image(text, d, CORES)
    ours, st = gpu_image(pkg, gpu, text, d, batch_slots=1 << 20, monkeypatch=monkeypatch)
    one, st1 = gpu_image(pkg, gpu, text, d)
    assert st.kernel_launches >= 3 * st1.kernel_launches
    check("c4_500k_tris_4_batches", ours, ref, dict(rel_rmse=2e-3, bad_frac=6e-4, mean_rel=2e-6))  # observed 2.3e-4, 6e-5, 1e-7
    check("c4_500k_tris_one_batch", one, ref, dict(rel_rmse=2e-3, bad_frac=6e-4, mean_rel=2e-6))
    check("c4_batches_vs_one_batch", ours, one, dict(rel_rmse=4e-6, bad_frac=0.0))
