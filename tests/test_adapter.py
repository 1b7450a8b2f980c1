"""The drop-in adapter of INTEGRATION.md section B as compiled code: oracle/_ref/pbrt_gpu_adapter is the REFERENCE's own
front end (main/pbrt.cpp, parser, pbrt* API, shape / material / light factories, BVHAccel, Film::WriteImage) with
`Integrator "path"` / `"directlighting"` routed to libpbrtgpu.so through oracle/adapter/gpupath.cpp
(integrator.h:53-58, api.cpp:1794).  Built by `make -C oracle adapter` where /root/reference exists; it travels to the
GPU box with oracle/_ref.

CPU tests: what the adapter flattens out of the reference's own Scene objects is, array for array, what pbrthost
makes of the same scene file (so the C ABI carries everything the reference's front end knows, and pbrthost's own
parser / BVH build / flattening is pinned against the reference's live objects); without a CUDA device the adapter
fails loudly.  GPU tests: the image the reference's Film writes after a GPU render matches the reference's own image.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import util
from conftest import GOLDEN, REF_DIR, ROOT, TEST_SCENES, scene_path

sys.path.insert(0, os.path.join(ROOT, "tools"))
from dev_compare import read_pfm  # noqa: E402

ADAPTER = os.path.join(REF_DIR, "pbrt_gpu_adapter")
needs_adapter = pytest.mark.skipif(not os.path.exists(ADAPTER), reason="oracle/_ref/pbrt_gpu_adapter not built (make -C oracle adapter)")

DT = {"f4": np.float32, "i4": np.int32, "u2": np.uint16, "u1": np.uint8}


def read_dump(path):
    out = {}
    with open(path, "rb") as f:
        while True:
            line = f.readline()
            if not line:
                break
            name, dt, count = line.decode().split()
            out[name] = np.frombuffer(f.read(int(count) * np.dtype(DT[dt]).itemsize), dtype=DT[dt]).copy()
    return out


def scene_text(case):
    name, subst = case
    text = open(scene_path(name)).read()
    for k, v in subst.items():
        text = text.replace(k, v)
    return text


def run_adapter(text, base_dir, tmp_path, dump=None, outfile="out.pfm", env_extra=None):
    scene = os.path.join(base_dir, "_adapter_%d.pbrt" % os.getpid()) if base_dir != str(tmp_path) else str(tmp_path / "scene.pbrt")
    with open(scene, "w") as f:
        f.write(text)
    env = dict(os.environ)
    if dump:
        env["PBRTGPU_ADAPTER_DUMP"] = dump
    env.update(env_extra or {})
    try:
        return subprocess.run([ADAPTER, "--quiet", "--outfile", str(tmp_path / outfile), scene], capture_output=True, text=True, env=env, timeout=600)
    finally:
        if base_dir != str(tmp_path):
            os.unlink(scene)


CASES = {
    "box_glass": ("box_glass.pbrt", {}),
    "sphere_lights": ("sphere_lights.pbrt", {}),
    "tri_soup": ("tri_soup.pbrt", {}),
    "portal_room_projection": ("portal_room.pbrt", {"STRATEGY": "projection"}),
    "direct_one": ("direct_mixed.pbrt", {"DIRECT_STRATEGY": "one"}),
    "direct_glass_all": ("direct_glass.pbrt", {}),
    "null_shells": ("null_shells.pbrt", {}),
    "killeroo": ("killeroo-path.pbrt", {}),
    "cornell": ("cornell-path.pbrt", {}),
    "lamp": ("lamp/lamp.pbrt", {}),
    "sandbox": ("creeper/sandbox.pbrt", {}),
}


@needs_adapter
@pytest.mark.parametrize("case", sorted(CASES))
def test_adapter_hands_over_the_scene_pbrthost_builds(pkg, case, tmp_path):
    name, _ = CASES[case]
    path = scene_path(name)
    text = scene_text(CASES[case])
    dump = str(tmp_path / "scene.bin")
    r = run_adapter(text, os.path.dirname(path), tmp_path, dump=dump, env_extra={"CUDA_VISIBLE_DEVICES": ""})
    assert os.path.exists(dump), r.stderr
    a = read_dump(dump)
    hs = pkg.HostScene(text=text, base_dir=os.path.dirname(path))
    v = hs.view
    nn, npr, nl = v.n_nodes, v.n_prims, v.n_lights

    def same(field, count, width=1, bits=True):
        arr = v.array(field, count, width)
        theirs = a[field]
        mine = np.asarray(arr).reshape(-1) if arr is not None else np.zeros(0, theirs.dtype)
        assert mine.shape == theirs.shape, field
        if bits and mine.dtype == np.float32:
            assert np.array_equal(mine.view(np.uint32), theirs.view(np.uint32)), field
        else:
            assert np.array_equal(mine, theirs), field

    # the BVH and the primitive order are the reference's, bit for bit
    same("bvh_bounds", nn, 6), same("bvh_offset", nn), same("bvh_nprims", nn), same("bvh_axis", nn)
    same("prim_type", npr), same("prim_flags", npr), same("tri_p", npr, 9), same("prim_light", npr)
    for f, w in (("tri_n", 9), ("tri_uv", 6), ("tri_s", 9)):
        if getattr(v, f):
            same(f, npr, w)
        else:
            assert a[f].size == 0, f
    # shapes and materials are shared objects on both sides, numbered in different orders: compare what each primitive resolves to
    ptype = a["prim_type"]
    for field, stride, n_mine, kind in (("plane_params", 48, v.n_planes, 1), ("sphere_params", 48, v.n_spheres, 2)):
        mine = np.asarray(v.array(field, n_mine, stride)).reshape(-1, stride) if n_mine else np.zeros((0, stride), np.float32)
        theirs = a[field].reshape(-1, stride)
        sel = np.nonzero(ptype == kind)[0]
        ms = np.asarray(v.array("prim_shape", npr))[sel]
        assert np.array_equal(mine[ms].view(np.uint32), theirs[a["prim_shape"][sel]].view(np.uint32)), field
    mm = np.asarray(v.array("prim_material", npr))
    tm = a["prim_material"]
    assert np.array_equal(mm < 0, tm < 0)
    has = mm >= 0
    mt_mine, mt_theirs = np.asarray(v.array("mat_tag", v.n_materials)), a["mat_tag"]
    mp_mine = np.asarray(v.array("mat_params", v.n_materials, 24)).reshape(-1, 24)
    mp_theirs = a["mat_params"].reshape(-1, 24)
    assert np.array_equal(mt_mine[mm[has]], mt_theirs[tm[has]])
    assert np.array_equal(mp_mine[mm[has]].view(np.uint32), mp_theirs[tm[has]].view(np.uint32)), "material parameters"
    # lights, in Scene::lights order
    same("light_tag", nl), same("light_prim", nl), same("light_portal_begin", nl + 1), same("light_nsamples", nl)
    same("light_params", nl, 48), same("portal_params", v.n_portals, 16)
    portal = a["light_tag"] == 2
    assert np.array_equal(np.asarray(v.array("light_strategy", nl))[portal], a["light_strategy"][portal])
    assert np.array_equal(np.array([v.world_center[0], v.world_center[1], v.world_center[2], v.world_radius], np.float32).view(np.uint32),
                          a["world"].view(np.uint32))
    # camera, film, filter table, sampler and integrator parameters: the whole struct, byte for byte
    import ctypes as C
    mine = np.frombuffer(C.string_at(C.addressof(hs.params), C.sizeof(hs.params)), np.uint8)
    assert mine.size == a["render_params"].size
    diff = np.nonzero(mine != a["render_params"])[0]
    assert diff.size == 0, "pbrtgpu_render_params differs at byte offsets %s" % diff[:16]


@needs_adapter
def test_adapter_fails_loudly_without_a_device(tmp_path):
    r = run_adapter(scene_text(CASES["tri_soup"]), str(tmp_path), tmp_path, env_extra={"CUDA_VISIBLE_DEVICES": ""})
    assert r.returncode == 2
    assert "GPU path integrator" in r.stderr + r.stdout and "no CPU fallback" in r.stderr + r.stdout
    assert not os.path.exists(tmp_path / "out.pfm")


# (case, tolerance on rel RMSE): the golden images are the reference binary's own renders (tests/golden/make_golden_images.py)
GPU_CASES = [("box_glass", 2e-3), ("sphere_lights", 1e-5), ("portal_room_projection", 1e-5), ("direct_one", 1e-5), ("direct_glass_all", 1e-5)]


@needs_adapter
@pytest.mark.gpu
@pytest.mark.parametrize("case,tol", GPU_CASES)
def test_gpu_adapter_image_matches_reference(case, tol, tmp_path):
    golden = np.load(os.path.join(GOLDEN, "images.npz"))
    r = run_adapter(scene_text(CASES[case]), str(tmp_path), tmp_path)
    assert r.returncode == 0, r.stderr + r.stdout
    img = read_pfm(str(tmp_path / "out.pfm"))
    s = util.image_stats(img, golden[case])
    assert s["rel_rmse"] <= tol and s["bad_frac"] <= 0.01, s


@needs_adapter
@pytest.mark.gpu
def test_gpu_adapter_renders_the_shipped_portal_scene(tmp_path):
    """lamp.pbrt as shipped (directlighting, maxdepth 100, four portals) through the reference's own front end and the GPU,
    against the reference's own CPU render of the same file."""
    path = scene_path("lamp/lamp.pbrt")
    text = open(path).read()
    r = run_adapter(text, os.path.dirname(path), tmp_path)
    assert r.returncode == 0, r.stderr + r.stdout
    img = read_pfm(str(tmp_path / "out.pfm"))
    ref_out = str(tmp_path / "ref.pfm")
    subprocess.check_call([os.path.join(REF_DIR, "pbrt"), "--quiet", "--nthreads", "1", "--outfile", ref_out, path])
    s = util.image_stats(img, read_pfm(ref_out))
    assert s["rel_rmse"] <= 1e-5, s
