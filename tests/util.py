"""Helpers shared by the tests, the golden-vector generator and bench.py:
seeded ray batches and thin callers for the
oracle / hostsim libraries."""
import ctypes as C

import numpy as np

SHADOW_TMAX = np.float32(1) - np.float32(0.0001)  # 1 - ShadowEpsilon, interaction.h:70-74


def world_bounds(view):
    b = view.array("bvh_bounds", 1, 6)[0]
    return b[:3].astype(np.float64), b[3:].astype(np.float64)


def rays_incoherent_closest(view, n, seed=42, lo=None, hi=None):
    """Origins uniform in the world bound x1.1, directions uniform on the sphere, tMax = inf."""
    rng = np.random.default_rng(seed)
    if lo is None:
        lo, hi = world_bounds(view)
    c, ext = (lo + hi) / 2, (hi - lo) * 1.1 / 2
    o = c + (rng.random((n, 3)) * 2 - 1) * ext
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((n, 8), np.float32)
    r[:, :3] = o
    r[:, 3] = np.inf
    r[:, 4:7] = d
    return r


def rays_shadow(view, n, seed=43, lo=None, hi=None):
    """o = p0, d = p1 - p0 (un-normalised), tMax = 1 - ShadowEpsilon: Interaction::SpawnRayTo."""
    rng = np.random.default_rng(seed)
    if lo is None:
        lo, hi = world_bounds(view)
    c, ext = (lo + hi) / 2, (hi - lo) * 1.1 / 2
    p0 = c + (rng.random((n, 3)) * 2 - 1) * ext
    p1 = c + (rng.random((n, 3)) * 2 - 1) * ext
    r = np.zeros((n, 8), np.float32)
    r[:, :3] = p0
    r[:, 3] = SHADOW_TMAX
    r[:, 4:7] = (p1 - p0).astype(np.float32)
    return r


def rays_toward_box(n, lo, hi, seed=44, radius_scale=2.0):
    """Origins on a sphere around [lo,hi], aimed at uniform points inside it (most rays hit)."""
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    c = (lo + hi) / 2
    rad = np.linalg.norm(hi - lo) / 2 * radius_scale
    o = rng.normal(size=(n, 3))
    o = c + rad * o / np.linalg.norm(o, axis=1, keepdims=True)
    tgt = lo + rng.random((n, 3)) * (hi - lo)
    d = tgt - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((n, 8), np.float32)
    r[:, :3] = o
    r[:, 3] = np.inf
    r[:, 4:7] = d
    return r


def rays_edge_cases(view, n, seed=45):
    """Rays aimed exactly at triangle vertices and edge midpoints, axis-parallel
    directions (zero components -> infinite invDir), and zero-length offsets."""
    rng = np.random.default_rng(seed)
    tri = view.array("tri_p", view.n_prims, 9)
    types = view.array("prim_type", view.n_prims)
    tri = tri[types == 0].reshape(-1, 3, 3).astype(np.float64)
    pick = rng.integers(0, len(tri), n)
    kind = rng.integers(0, 4, n)
    tgt = np.where((kind == 0)[:, None], tri[pick, rng.integers(0, 3, n)],
                   (tri[pick, 0] + tri[pick, 1]) / 2)
    lo, hi = world_bounds(view)
    o = lo + rng.random((n, 3)) * (hi - lo)
    d = tgt - o
    axis = rng.integers(0, 3, n)
    par = kind == 3
    d[par] = 0
    d[par, axis[par]] = rng.choice([-1.0, 1.0], par.sum())
    o[par] = tgt[par] - d[par] * 5.0
    nz = np.linalg.norm(d, axis=1) > 0
    d[nz] /= np.linalg.norm(d[nz], axis=1, keepdims=True)
    r = np.zeros((n, 8), np.float32)
    r[:, :3] = o
    r[:, 3] = np.inf
    r[:, 4:7] = d
    return r[nz]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle_closest(lib, view, rays, threads=8):
    n = len(rays)
    ids, t, b = np.empty(n, np.int32), np.empty(n, np.float32), np.empty((n, 2), np.float32)
    cnt = (C.c_int64 * 2)()
    lib.oracle_intersect_closest(C.byref(view), _ptr(rays), C.c_int64(n), _ptr(ids), _ptr(t), _ptr(b), threads, cnt)
    return ids, t, b, (cnt[0], cnt[1])


def oracle_any(lib, view, rays, threads=8):
    n = len(rays)
    occ = np.empty(n, np.uint8)
    cnt = (C.c_int64 * 2)()
    lib.oracle_intersect_any(C.byref(view), _ptr(rays), C.c_int64(n), _ptr(occ), threads, cnt)
    return occ, (cnt[0], cnt[1])


def hostsim_closest(lib, view, rays, threads=8):
    h = C.c_void_p(lib.hostsim_create(C.byref(view)))
    n = len(rays)
    ids, t, b = np.empty(n, np.int32), np.empty(n, np.float32), np.empty((n, 2), np.float32)
    lib.hostsim_intersect_closest(h, _ptr(rays), C.c_int64(n), _ptr(ids), _ptr(t), _ptr(b), threads)
    lib.hostsim_destroy(h)
    return ids, t, b


def hostsim_any(lib, view, rays, threads=8):
    h = C.c_void_p(lib.hostsim_create(C.byref(view)))
    n = len(rays)
    occ = np.empty(n, np.uint8)
    lib.hostsim_intersect_any(h, _ptr(rays), C.c_int64(n), _ptr(occ), threads)
    lib.hostsim_destroy(h)
    return occ


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


# ---------------------------------------------------------------- images (tier 2)

def image_stats(ours, ref):
    """rel_rmse = RMSE / mean(ref); bad_frac = share of pixels whose worst channel is off by
    more than 1 % (relative to max(|ref|, 0.1 % of the image mean))."""
    d = ours.astype(np.float64) - ref.astype(np.float64)
    mean = max(float(ref.astype(np.float64).mean()), 1e-12)
    rel = np.abs(d) / np.maximum(np.abs(ref), 1e-3 * mean)
    return {"rel_rmse": float(np.sqrt((d ** 2).mean()) / mean), "bad_frac": float((rel.max(axis=2) > 0.01).mean()),
            "mean_rel": float(abs(ours.astype(np.float64).mean() - mean) / mean)}


def hostsim_render(lib, scene, params=None, threads=8):
    """The device stage functions stepped on the CPU (tests/hostsim): returns (rgb, stats4)."""
    lib.hostsim_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    params = params if params is not None else scene.params
    h = C.c_void_p(lib.hostsim_create(C.byref(scene.view)))
    cb = params.cropped_bounds
    film = np.zeros((cb[3] - cb[1], cb[2] - cb[0], 4), np.float32)
    st = (C.c_int64 * 4)()
    lib.hostsim_render(h, C.byref(params), film.ctypes.data, st, threads)
    lib.hostsim_destroy(h)
    return scene.film_to_rgb(film), film, list(st)
