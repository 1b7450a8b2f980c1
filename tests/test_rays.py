"""Tier-1 parity: BVHAccel::Intersect / IntersectP (accelerators/bvh.cpp:662-738).

CPU (`-m "not gpu"`):
  * the host front end reproduces the reference's BVH and primitive table bit for bit
    (golden dump from oracle/_ref/ref_harness),
  * the oracle restatement reproduces the reference's hit IDs, t and occlusion bits,
  * the device traversal code, stepped on the CPU, does too.
GPU (`-m gpu`): the CUDA kernels, through the C ABI, against the golden vectors
(bit-exact IDs, t and occlusion) and against the oracle on larger batches.
"""
import hashlib
import os

import numpy as np
import pytest

import util
from conftest import GOLDEN, ROOT, scene_path

SCENE = "tri_soup.pbrt"


@pytest.fixture(scope="module")
def soup(pkg):
    return pkg.HostScene(scene_path(SCENE))


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(GOLDEN, "tri_soup_rays.npz"))


def ray_sets(view):
    return {
        "incoherent": util.rays_incoherent_closest(view, 40000, seed=42),
        "toward": util.rays_toward_box(40000, [-10, -10, -10], [10, 10, 10], seed=44),
        "edges": util.rays_edge_cases(view, 20000, seed=45),
        "shadow": util.rays_shadow(view, 40000, seed=43),
    }


def test_host_bvh_matches_reference_dump(soup):
    g = np.load(os.path.join(GOLDEN, "tri_soup_scene.npz"))
    nodes, prims = g["nodes"], g["prims"]
    v = soup.view
    assert v.n_nodes == len(nodes) and v.n_prims == len(prims)
    assert np.array_equal(util.bits(v.array("bvh_bounds", v.n_nodes, 6)), nodes[:, :24].copy().view(np.uint32))
    assert np.array_equal(v.array("bvh_offset", v.n_nodes), nodes[:, 24:28].copy().view(np.int32).ravel())
    nprims = nodes[:, 28:30].copy().view(np.uint16).ravel()
    assert np.array_equal(v.array("bvh_nprims", v.n_nodes), nprims)
    assert np.array_equal(v.array("bvh_axis", v.n_nodes)[nprims == 0], nodes[:, 30][nprims == 0])
    assert np.array_equal(v.array("prim_orig_id", v.n_prims), prims[:, :4].copy().view(np.int32).ravel())
    assert np.array_equal(v.array("prim_type", v.n_prims), prims[:, 4:8].copy().view(np.int32).ravel())
    assert np.array_equal(util.bits(v.array("tri_p", v.n_prims, 9)), prims[:, 8:].copy().view(np.uint32).reshape(-1, 9))


@pytest.mark.parametrize("method", ["hlbvh", "middle", "equal", "sah"])
def test_host_bvh_split_methods_match_live_reference(pkg, ref_harness, method, tmp_path):
    """Accelerator "bvh" "splitmethod": every method's tree, node for node and leaf slot for leaf slot, against a
    dump of the reference's BVHAccel (HLBVHBuild bvh.cpp:404-638 included: Morton codes, treelets, upper SAH)."""
    import subprocess
    if not ref_harness:
        pytest.skip("oracle/_ref not built")
    src = scene_path("tri_soup.pbrt")
    text = open(src).read().replace("WorldBegin", 'Accelerator "bvh" "string splitmethod" "%s"\nWorldBegin' % method, 1)
    tmp = os.path.join(os.path.dirname(src), "_split_%s_%d.pbrt" % (method, os.getpid()))
    open(tmp, "w").write(text)
    try:
        subprocess.check_call([ref_harness, "dump", tmp, str(tmp_path / "d")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        hs = pkg.HostScene(tmp)
    finally:
        os.unlink(tmp)
    nodes = np.fromfile(str(tmp_path / "d.nodes.bin"), dtype=np.uint8).reshape(-1, 32)
    prims = np.fromfile(str(tmp_path / "d.prims.bin"), dtype=np.uint8).reshape(-1, 44)
    v = hs.view
    assert v.n_nodes == len(nodes) and v.n_prims == len(prims)
    nprims = nodes[:, 28:30].copy().view(np.uint16).ravel()
    assert np.array_equal(util.bits(v.array("bvh_bounds", v.n_nodes, 6)), nodes[:, :24].copy().view(np.uint32))
    assert np.array_equal(v.array("bvh_offset", v.n_nodes), nodes[:, 24:28].copy().view(np.int32).ravel())
    assert np.array_equal(v.array("bvh_nprims", v.n_nodes), nprims)
    assert np.array_equal(v.array("bvh_axis", v.n_nodes)[nprims == 0], nodes[:, 30][nprims == 0])
    assert np.array_equal(v.array("prim_orig_id", v.n_prims), prims[:, :4].copy().view(np.int32).ravel())


def test_ray_generators_are_stable(soup, golden):
    for name, rays in ray_sets(soup.view).items():
        assert hashlib.sha256(rays.tobytes()).hexdigest() == str(golden[name + "_sha"]), name


@pytest.mark.parametrize("name", ["incoherent", "toward", "edges", "shadow"])
def test_oracle_matches_reference(soup, golden, oracle_lib, name):
    rays = ray_sets(soup.view)[name]
    ids, t, _, (V, T) = util.oracle_closest(oracle_lib, soup.view, rays)
    assert np.array_equal(ids, golden[name + "_id"])
    assert np.array_equal(util.bits(t), util.bits(golden[name + "_t"]))
    occ, _ = util.oracle_any(oracle_lib, soup.view, rays)
    assert np.array_equal(occ, golden[name + "_occ"])
    assert V > 0 and T > 0


@pytest.mark.parametrize("name", ["incoherent", "toward", "edges", "shadow"])
def test_device_traversal_logic_on_cpu(soup, golden, hostsim_lib, name):
    rays = ray_sets(soup.view)[name]
    ids, t, _ = util.hostsim_closest(hostsim_lib, soup.view, rays)
    assert np.array_equal(ids, golden[name + "_id"])
    assert np.array_equal(util.bits(t), util.bits(golden[name + "_t"]))
    assert np.array_equal(util.hostsim_any(hostsim_lib, soup.view, rays), golden[name + "_occ"])


def test_oracle_against_live_reference(pkg, oracle_lib, ref_harness, tmp_path):
    """Where oracle/_ref is present, also pin the oracle on the staged killeroo scene."""
    import subprocess
    if ref_harness is None:
        pytest.skip("oracle/_ref not built")
    path = scene_path("killeroo-path.pbrt")
    hs = pkg.HostScene(path)
    rays = np.concatenate([util.rays_incoherent_closest(hs.view, 100000, seed=1),
                           util.rays_toward_box(100000, [-100, -100, -150], [100, 150, 50], seed=2)])
    rp = str(tmp_path / "r.bin")
    rays.tofile(rp)
    subprocess.check_call([ref_harness, "intersect", path, rp, rp + ".c", "closest", "8"], stdout=subprocess.DEVNULL)
    subprocess.check_call([ref_harness, "intersect", path, rp, rp + ".a", "any", "8"], stdout=subprocess.DEVNULL)
    ref = np.fromfile(rp + ".c", dtype=np.int32).reshape(-1, 2)
    ids, t, _, _ = util.oracle_closest(oracle_lib, hs.view, rays)
    assert np.array_equal(ids, ref[:, 0])
    assert np.array_equal(util.bits(t), ref[:, 1].copy().view(np.uint32))
    occ, _ = util.oracle_any(oracle_lib, hs.view, rays)
    assert np.array_equal(occ, np.fromfile(rp + ".a", dtype=np.uint8))


# ---------------------------------------------------------------- GPU

@pytest.mark.gpu
@pytest.mark.parametrize("name", ["incoherent", "toward", "edges", "shadow"])
def test_gpu_matches_golden(soup, golden, gpu, name):
    gpu.upload(soup)
    rays = ray_sets(soup.view)[name]
    ids, t, _ = gpu.intersect_closest(rays)
    assert np.array_equal(ids, golden[name + "_id"])
    assert np.array_equal(util.bits(t), util.bits(golden[name + "_t"]))
    assert np.array_equal(gpu.intersect_any(rays), golden[name + "_occ"])


@pytest.mark.gpu
def test_gpu_matches_oracle_large(soup, gpu, oracle_lib):
    gpu.upload(soup)
    rays = np.concatenate([util.rays_incoherent_closest(soup.view, 1000000, seed=7),
                           util.rays_toward_box(1000000, [-10, -10, -10], [10, 10, 10], seed=8),
                           util.rays_edge_cases(soup.view, 300000, seed=9)])
    ids, t, b = gpu.intersect_closest(rays)
    oid, ot, ob, _ = util.oracle_closest(oracle_lib, soup.view, rays)
    assert np.array_equal(ids, oid)
    assert np.array_equal(util.bits(t), util.bits(ot))
    assert np.array_equal(util.bits(b), util.bits(ob))
    srays = util.rays_shadow(soup.view, 1000000, seed=10)
    occ, _ = util.oracle_any(oracle_lib, soup.view, srays)
    assert np.array_equal(gpu.intersect_any(srays), occ)


@pytest.mark.gpu
def test_gpu_killeroo_matches_oracle(pkg, gpu, oracle_lib):
    hs = pkg.HostScene(scene_path("killeroo-path.pbrt"))
    gpu.upload(hs)
    rays = np.concatenate([util.rays_incoherent_closest(hs.view, 1000000, seed=1),
                           util.rays_toward_box(2000000, [-100, -100, -150], [100, 150, 50], seed=2)])
    ids, t, _ = gpu.intersect_closest(rays)
    oid, ot, _, _ = util.oracle_closest(oracle_lib, hs.view, rays)
    assert np.array_equal(ids, oid)
    assert np.array_equal(util.bits(t), util.bits(ot))
    srays = util.rays_shadow(hs.view, 1000000, seed=3, lo=np.array([-150., -150, -150]), hi=np.array([250., 250, 100]))
    occ, _ = util.oracle_any(oracle_lib, hs.view, srays)
    assert np.array_equal(gpu.intersect_any(srays), occ)


@pytest.mark.gpu
@pytest.mark.parametrize("scene", ["soup", "killeroo"])
def test_gpu_traversal_enters_the_nodes_the_reference_enters(pkg, soup, gpu, oracle_lib, scene):
    """The visiting-order claim of bvh.h / trace_persistent.h, counted instead of argued: over a batch of closest-hit rays the
    persistent kernel tests exactly the primitives the reference's loop (bvh.cpp:662-738, restated in oracle/port) tests, and
    enters exactly the interior nodes it enters.  The reference counts one bounds test per node it visits (1 + 2 per interior
    node whose bounds it passes); the device tests both children when it enters a record, record 0 -- the synthetic parent of
    the root -- included: 2 + 2 per interior node, i.e. one more per ray.  A cull that differed, a child visited in the other
    order (a different tMax at the far child's pop) or a leaf walked twice would move one of the two tallies.  Any-hit rays stop
    at the first accepted primitive, with stacked entries left whose tMax-independent half the device has already tested, so
    only the primitive tally is compared there."""
    hs = soup if scene == "soup" else pkg.HostScene(scene_path("killeroo-path.pbrt"))
    gpu.upload(hs)
    lo, hi = ([-10, -10, -10], [10, 10, 10]) if scene == "soup" else ([-100, -100, -150], [100, 150, 50])
    rays = np.concatenate([util.rays_incoherent_closest(hs.view, 300000, seed=21), util.rays_toward_box(300000, lo, hi, seed=22),
                           util.rays_edge_cases(hs.view, 50000, seed=23)])
    gpu.set_profile(True)
    try:
        ids, t, _ = gpu.intersect_closest(rays)
        closest = gpu.get_profile()["trace_closest"]
        srays = util.rays_shadow(hs.view, 300000, seed=24)
        occ = gpu.intersect_any(srays)
        shadow = gpu.get_profile()["trace_shadow"]
    finally:
        gpu.set_profile(False)
    oid, ot, _, (nodes, prims) = util.oracle_closest(oracle_lib, hs.view, rays)
    assert np.array_equal(ids, oid) and np.array_equal(util.bits(t), util.bits(ot))
    assert closest["prim_tests"] == prims
    assert closest["node_tests"] == nodes + len(rays)
    oocc, (_, sprims) = util.oracle_any(oracle_lib, hs.view, srays)
    assert np.array_equal(occ, oocc)
    assert shadow["prim_tests"] == sprims


@pytest.mark.gpu
def test_gpu_edge_inputs(soup, gpu):
    gpu.upload(soup)
    ids, t, _ = gpu.intersect_closest(np.zeros((0, 8), np.float32))
    assert len(ids) == 0
    # zero direction, NaN origin, negative tMax: must terminate and report a miss
    bad = np.zeros((3, 8), np.float32)
    bad[0, 3] = np.inf
    bad[1, :3] = np.nan; bad[1, 3] = np.inf; bad[1, 4:7] = [0, 0, 1]
    bad[2, 3] = -1; bad[2, 4:7] = [0, 1, 0]
    ids, t, _ = gpu.intersect_closest(bad)
    assert (ids == -1).all()
    assert (gpu.intersect_any(bad) == 0).all()


def test_sphere_early_out_never_rejects_a_hit(pkg, hostsim_lib):
    """shapes.h:sphereRejects claims to be EXACT: whenever it says "rejected", the interval-arithmetic
    Sphere::Intersect restatement (sphereTest) returns false as well.  Checked on rays of every kind around the
    spheres of tests/scenes/sphere_lights.pbrt: aimed at the sphere, grazing it, starting inside it, ending
    just short of it (the shadow-ray case the early-out exists for), and missing it by far."""
    import ctypes as C
    hs = pkg.HostScene(os.path.join(os.path.dirname(__file__), "scenes", "sphere_lights.pbrt"))
    v = hs.view
    assert v.n_spheres > 0
    recs = np.ctypeslib.as_array(v.sphere_params, shape=(v.n_spheres, 48)).copy()
    rng = np.random.default_rng(3)
    total_rej = total = 0
    for rec in recs:
        radius = rec[0]
        o2w = rec[16:32].reshape(4, 4)
        centre = o2w[:3, 3].astype(np.float64)
        n = 40000
        rays = np.zeros((n, 8), np.float32)
        org = centre + rng.normal(size=(n, 3)) * radius * rng.choice([0.5, 1.5, 4.0, 50.0], size=(n, 1))
        on = rng.normal(size=(n, 3))
        on /= np.linalg.norm(on, axis=1, keepdims=True)
        target = centre + on * radius * rng.choice([0.0, 0.9, 1.0, 1.0001, 1.1, 3.0], size=(n, 1))
        rays[:, 0:3] = org
        rays[:, 4:7] = target - org
        rays[:, 3] = rng.choice([np.inf, 1.0 - 1e-4, 1.0, 0.5, 2.0], size=n)
        rej = np.zeros(n, np.uint8)
        acc = np.zeros(n, np.uint8)
        hostsim_lib.hostsim_sphere_guard(rec.ctypes.data_as(C.c_void_p), rays.ctypes.data_as(C.c_void_p), C.c_int64(n),
                                         rej.ctypes.data_as(C.c_void_p), acc.ctypes.data_as(C.c_void_p))
        assert not np.any((rej == 1) & (acc == 1)), "sphereRejects turned away a ray that Sphere::Intersect accepts"
        total_rej += int(rej.sum())
        total += int((acc == 0).sum())
    assert total_rej > 0.5 * total  # and it is worth having: it catches most of the rays the full test rejects


# ---------------------------------------------------------------- HLBVH built on the device
def _scene_arrays(hs):
    v = hs.view
    out = {"bvh_bounds": util.bits(v.array("bvh_bounds", v.n_nodes, 6)).copy()}
    for name, count in (("bvh_offset", v.n_nodes), ("bvh_nprims", v.n_nodes), ("bvh_axis", v.n_nodes), ("prim_orig_id", v.n_prims),
                        ("prim_type", v.n_prims), ("prim_material", v.n_prims), ("prim_light", v.n_prims)):
        out[name] = np.array(v.array(name, count))
    out["tri_p"] = util.bits(v.array("tri_p", v.n_prims, 9)).copy()
    return out


def _load_both_ways(pkg, gpu, text, base_dir):
    """The same scene text with the host's HLBVHBuild and with pbrtgpu_build_hlbvh installed as pbrthost's builder."""
    host = pkg.HostScene(text=text, base_dir=base_dir)
    gpu.install_hlbvh_builder(True)
    try:
        dev = pkg.HostScene(text=text, base_dir=base_dir)
        wall_ms, dev_ms = pkg.last_bvh_build_ms()
    finally:
        gpu.install_hlbvh_builder(False)
    assert dev_ms > 0, "the device builder was not called"
    return host, dev


def _random_triangles_scene(n, seed, extra="", maxnodeprims=4, clumps=0, degenerate=0):
    """n small random triangles in a 10^3 box; `clumps` > 0 puts every triangle's centroid on one of that many points, so that
    whole groups share a Morton code (emitLBVH's bitIndex == -1 leaves, bvh.cpp:481)."""
    rng = np.random.default_rng(seed)
    if clumps:
        c = rng.uniform(-5, 5, (clumps, 3))[rng.integers(0, clumps, n)]
        d = rng.normal(0, 0.05, (n, 2, 3))
        tri = np.stack([c + d[:, 0], c + d[:, 1], c - d[:, 0] - d[:, 1]], 1)  # centroid of the bounds is not c exactly, but within a cell
    else:
        c = rng.uniform(-5, 5, (n, 1, 3))
        tri = c + rng.normal(0, 0.05, (n, 3, 3))
    tri[:degenerate, 2] = tri[:degenerate, 1]  # zero-area triangles: Triangle::Intersect rejects them (triangle.cpp:309-317)
    P = " ".join("%.9g" % x for x in tri.astype(np.float32).ravel())
    idx = " ".join(str(i) for i in range(3 * n))
    return ('LookAt 0 -20 0  0 0 0  0 0 1\nCamera "perspective" "float fov" [40]\n'
            'Film "image" "integer xresolution" [16] "integer yresolution" [16] "string filename" "t.pfm"\n'
            'Sampler "halton" "integer pixelsamples" [1]\nIntegrator "path"\n'
            'Accelerator "bvh" "string splitmethod" "hlbvh" "integer maxnodeprims" [%d]\nWorldBegin\n%s'
            'LightSource "point" "point from" [0 -15 5]\nMaterial "matte"\n'
            'Shape "trianglemesh" "integer indices" [%s] "point P" [%s]\nWorldEnd\n' % (maxnodeprims, extra, idx, P))


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["one", "three", "soup", "random_60k", "random_60k_leaf1", "random_60k_leaf255", "clumps", "killeroo"])
def test_gpu_hlbvh_build_matches_host_build(pkg, gpu, case):
    """pbrtgpu_build_hlbvh (device/bvh_build.cu) against the host's HLBVHBuild, which is pinned node for node to the
    reference's (test_host_bvh_split_methods_match_live_reference): the LinearBVHNode array bit for bit, the primitive order,
    and everything flattened in that order.  From one primitive to 133 000, maxnodeprims 1 / 4 / 255, shared Morton codes."""
    base = os.path.join(ROOT, "tests", "scenes")
    if case == "soup":
        text = open(scene_path("tri_soup.pbrt")).read().replace("WorldBegin", 'Accelerator "bvh" "string splitmethod" "hlbvh"\nWorldBegin', 1)
    elif case == "killeroo":
        p = scene_path("killeroo-path.pbrt")
        base = os.path.dirname(p)
        text = open(p).read().replace("WorldBegin", 'Accelerator "bvh" "string splitmethod" "hlbvh"\nWorldBegin', 1)
    elif case == "one":
        text = _random_triangles_scene(1, 1)
    elif case == "three":
        text = _random_triangles_scene(3, 2)
    elif case == "clumps":
        text = _random_triangles_scene(20000, 3, clumps=37)
    else:
        text = _random_triangles_scene(60000, 4, maxnodeprims={"random_60k": 4, "random_60k_leaf1": 1, "random_60k_leaf255": 255}[case])
    host, dev = _load_both_ways(pkg, gpu, text, base)
    a, b = _scene_arrays(host), _scene_arrays(dev)
    assert host.view.n_nodes == dev.view.n_nodes and host.view.n_prims == dev.view.n_prims
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    if case == "clumps":
        assert a["bvh_nprims"].max() > 4  # leaves made because the Morton bits ran out, not because the range was small


@pytest.mark.gpu
def test_gpu_hlbvh_tree_traces_like_the_reference(pkg, gpu, ref_harness, tmp_path):
    """End to end: a scene whose tree was built on the GPU, traced on the GPU, against the reference's own BVHAccel
    (splitmethod hlbvh) on the same rays: hit IDs and t bit for bit."""
    import subprocess
    if not ref_harness:
        pytest.skip("oracle/_ref not built")
    text = _random_triangles_scene(60000, 9)
    path = str(tmp_path / "s.pbrt")
    open(path, "w").write(text)
    _, dev = _load_both_ways(pkg, gpu, text, str(tmp_path))
    rays = np.concatenate([util.rays_incoherent_closest(dev.view, 200000, seed=3), util.rays_toward_box(100000, [-6, -6, -6], [6, 6, 6], seed=4)])
    rays.tofile(str(tmp_path / "r.bin"))
    subprocess.check_call([ref_harness, "intersect", path, str(tmp_path / "r.bin"), str(tmp_path / "o.bin"), "closest", str(os.cpu_count() or 1)],
                          stdout=subprocess.DEVNULL)
    ref = np.fromfile(str(tmp_path / "o.bin"), dtype=[("id", "<i4"), ("t", "<f4")])
    gpu.upload(dev)
    ids, t, _ = gpu.intersect_closest(rays)
    assert np.array_equal(ids, ref["id"]) and np.array_equal(util.bits(t), util.bits(ref["t"]))
    assert (ids >= 0).sum() > 10000


@pytest.mark.gpu
def test_gpu_hlbvh_build_rejects_bad_arguments(pkg, gpu):
    import ctypes as C
    lib = pkg.gpu_lib()
    pb = np.zeros((4, 6), np.float32)
    out = np.zeros(64, np.float32)
    m = C.c_int32(0)
    rc = lib.pbrtgpu_build_hlbvh(gpu._h, pb.ctypes.data, 4, 4, 7, out.ctypes.data, out.ctypes.data, out.ctypes.data, out.ctypes.data, out.ctypes.data,
                                 C.byref(m), None)
    assert rc == -3 and b"node_capacity" in lib.pbrtgpu_last_error(gpu._h)
    assert lib.pbrtgpu_build_hlbvh(gpu._h, None, 4, 4, 8, out.ctypes.data, out.ctypes.data, out.ctypes.data, out.ctypes.data, out.ctypes.data,
                                   C.byref(m), None) == -3


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["soup", "killeroo", "box_glass", "sphere_lights", "portal_room", "random_60k"])
def test_gpu_layout_conversion_matches_host_restatement(pkg, gpu, case, monkeypatch):
    """pbrtgpu_upload_scene builds the device layout (64-byte two-child records, 48-byte primitive slots with their tags:
    type, degenerate, last-in-leaf, shape index) with kernels (device/layout_device.cu); PBRTGPU_LAYOUT=host builds it with the
    CPU restatement the hostsim tests run on (device/layout_build.h).  Same bytes."""
    base = os.path.join(ROOT, "tests", "scenes")
    if case == "random_60k":
        hs = pkg.HostScene(text=_random_triangles_scene(60000, 4, degenerate=70), base_dir=base)
    elif case == "portal_room":
        hs = pkg.HostScene(text=open(scene_path("portal_room.pbrt")).read().replace("STRATEGY", "portal"), base_dir=base)
    else:
        hs = pkg.HostScene(scene_path({"soup": "tri_soup.pbrt", "killeroo": "killeroo-path.pbrt", "box_glass": "box_glass.pbrt",
                                       "sphere_lights": "sphere_lights.pbrt"}[case]))
    gpu.upload(hs)
    nodes_d, tris_d = gpu.read_layout(hs)
    monkeypatch.setenv("PBRTGPU_LAYOUT", "host")
    other = pkg.GpuContext(0)
    try:
        other.upload(hs)
        nodes_h, tris_h = other.read_layout(hs)
    finally:
        other.close()
    assert np.array_equal(nodes_d, nodes_h)
    assert np.array_equal(tris_d, tris_h)
    assert ((tris_d[:, 3] & 8) != 0).sum() == (np.asarray(hs.view.array("bvh_nprims", hs.view.n_nodes)) != 0).sum()  # one kTagLast per leaf
    if case == "random_60k":
        assert ((tris_d[:, 3] & 4) != 0).sum() == 70  # kTagDegenerate
