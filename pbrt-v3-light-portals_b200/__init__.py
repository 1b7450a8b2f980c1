"""pbrt-v3-light-portals on B200: Python bindings over the two C ABIs.

  include/pbrthost.h  -> lib/libpbrthost.so  scene files, BVH build, flattening, film output (CPU)
  include/pbrtgpu.h   -> lib/libpbrtgpu.so   sm_100a kernels: ray batches and the path integrator

Nothing here computes: the classes are thin ctypes mirrors used by the tests,
bench.py and the `pbrt_gpu` command line.  There is no CPU fallback -- GpuContext
raises if the CUDA library is missing or no device is usable.

The package directory name contains dashes; import it with
    importlib.import_module("pbrt-v3-light-portals_b200")
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_PKG, "lib")

MAT_STRIDE, LIGHT_STRIDE, PLANE_STRIDE, SPHERE_STRIDE, PORTAL_STRIDE = 24, 48, 48, 48, 16
N_HALTON_PERMS = 3682913
NOALLOC_MAX_RAYS = 32768  # PBRTGPU_NOALLOC_MAX_RAYS


class PbrtGpuError(RuntimeError):
    pass


def _p(t):
    return C.POINTER(t)


class SceneView(C.Structure):  # pbrtgpu_scene
    _fields_ = [
        ("n_nodes", C.c_int32), ("bvh_bounds", _p(C.c_float)), ("bvh_offset", _p(C.c_int32)),
        ("bvh_nprims", _p(C.c_uint16)), ("bvh_axis", _p(C.c_uint8)),
        ("n_prims", C.c_int32), ("prim_type", _p(C.c_uint8)), ("prim_flags", _p(C.c_uint8)),
        ("prim_shape", _p(C.c_int32)), ("prim_material", _p(C.c_int32)), ("prim_light", _p(C.c_int32)),
        ("prim_orig_id", _p(C.c_int32)), ("tri_p", _p(C.c_float)), ("tri_n", _p(C.c_float)),
        ("tri_uv", _p(C.c_float)), ("tri_s", _p(C.c_float)),
        ("n_planes", C.c_int32), ("plane_params", _p(C.c_float)),
        ("n_spheres", C.c_int32), ("sphere_params", _p(C.c_float)),
        ("n_materials", C.c_int32), ("mat_tag", _p(C.c_int32)), ("mat_params", _p(C.c_float)),
        ("n_lights", C.c_int32), ("light_tag", _p(C.c_int32)), ("light_prim", _p(C.c_int32)),
        ("light_strategy", _p(C.c_int32)), ("light_portal_begin", _p(C.c_int32)), ("light_params", _p(C.c_float)),
        ("n_portals", C.c_int32), ("portal_params", _p(C.c_float)),
        ("world_center", C.c_float * 3), ("world_radius", C.c_float), ("light_nsamples", _p(C.c_int32)),
    ]

    def array(self, name, count, width=1):
        ptr = getattr(self, name)
        if not ptr or count == 0:
            return None
        a = np.ctypeslib.as_array(ptr, (count * width,))
        return a.reshape(count, width) if width > 1 else a


class RenderParams(C.Structure):  # pbrtgpu_render_params
    _fields_ = [
        ("camera_to_world", C.c_float * 16), ("raster_to_camera", C.c_float * 16),
        ("lens_radius", C.c_float), ("focal_distance", C.c_float),
        ("shutter_open", C.c_float), ("shutter_close", C.c_float),
        ("full_resolution", C.c_int32 * 2), ("cropped_bounds", C.c_int32 * 4),
        ("sample_bounds", C.c_int32 * 4), ("pixel_bounds", C.c_int32 * 4),
        ("filter_radius", C.c_float * 2), ("filter_table", C.c_float * 256),
        ("max_sample_luminance", C.c_float),
        ("spp", C.c_int32), ("sample_at_pixel_center", C.c_int32),
        ("max_depth", C.c_int32), ("rr_threshold", C.c_float), ("light_strategy", C.c_int32),
        ("shard_index", C.c_int32), ("shard_count", C.c_int32),
        ("spp_begin", C.c_int32), ("spp_end", C.c_int32),
        ("integrator", C.c_int32), ("direct_strategy", C.c_int32),
    ]

    def copy(self):
        out = RenderParams()
        C.memmove(C.byref(out), C.byref(self), C.sizeof(RenderParams))
        return out


class Stats(C.Structure):  # pbrtgpu_stats
    _fields_ = [("camera_rays", C.c_int64), ("closest_rays", C.c_int64), ("shadow_rays", C.c_int64),
                ("scrubbed", C.c_int64), ("kernel_launches", C.c_int64), ("nee_closest_rays", C.c_int64), ("device_ms", C.c_float), ("nee_rays_culled", C.c_int64),
                ("truncated_paths", C.c_int64), ("nee_rays_dropped", C.c_int64)]


K_NAMES = ["raygen", "trace_closest", "shade", "trace_shadow", "trace_nee", "film"]


class Profile(C.Structure):  # pbrtgpu_profile
    _fields_ = [("ms", C.c_double * 6), ("launches", C.c_int64 * 6), ("node_tests", C.c_int64 * 6), ("prim_tests", C.c_int64 * 6)]


_host = None
_gpu = None


def build(verbose=False):
    """Compile the native libraries in-tree (see build.py)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("_pbrt_b200_build", os.path.join(_PKG, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build_all(verbose=verbose)


def host_lib():
    global _host
    if _host is None:
        path = os.path.join(_LIB, "libpbrthost.so")
        if not os.path.exists(path):
            raise PbrtGpuError("libpbrthost.so is not built; run __graft_entry__.build()")
        lib = C.CDLL(path)
        lib.pbrthost_load_file.argtypes = [C.c_char_p, C.c_int, _p(C.c_void_p)]
        lib.pbrthost_load_string.argtypes = [C.c_char_p, C.c_char_p, C.c_int, _p(C.c_void_p)]
        lib.pbrthost_free.argtypes = [C.c_void_p]
        lib.pbrthost_scene_view.argtypes = [C.c_void_p]
        lib.pbrthost_scene_view.restype = _p(SceneView)
        lib.pbrthost_render_params.argtypes = [C.c_void_p]
        lib.pbrthost_render_params.restype = _p(RenderParams)
        lib.pbrthost_output_filename.argtypes = [C.c_void_p]
        lib.pbrthost_output_filename.restype = C.c_char_p
        lib.pbrthost_integrator_name.argtypes = [C.c_void_p]
        lib.pbrthost_integrator_name.restype = C.c_char_p
        lib.pbrthost_set_spp.argtypes = [C.c_void_p, C.c_int]
        lib.pbrthost_film_to_rgb.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.pbrthost_write_image.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
        lib.pbrthost_halton_permutations.argtypes = [C.c_void_p, C.c_int64]
        lib.pbrthost_set_hlbvh_builder.argtypes = [C.c_void_p, C.c_void_p]
        lib.pbrthost_set_hlbvh_builder.restype = None
        lib.pbrthost_last_bvh_build_ms.argtypes = [_p(C.c_double), _p(C.c_float)]
        lib.pbrthost_last_bvh_build_ms.restype = None
        _host = lib
    return _host


def gpu_lib():
    """Load libpbrtgpu.so.  Loading needs libcudart but no device."""
    global _gpu
    if _gpu is None:
        path = os.environ.get("PBRTGPU_LIBRARY") or os.path.join(_LIB, "libpbrtgpu.so")  # the override is for A/B builds (tools/build_variants.py)
        if not os.path.exists(path):
            raise PbrtGpuError("libpbrtgpu.so is not built; run __graft_entry__.build() (there is no CPU fallback)")
        lib = C.CDLL(path)
        lib.pbrtgpu_create.argtypes = [C.c_int, _p(C.c_void_p)]
        lib.pbrtgpu_destroy.argtypes = [C.c_void_p]
        lib.pbrtgpu_last_error.argtypes = [C.c_void_p]
        lib.pbrtgpu_last_error.restype = C.c_char_p
        lib.pbrtgpu_upload_scene.argtypes = [C.c_void_p, _p(SceneView)]
        lib.pbrtgpu_intersect_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.pbrtgpu_intersect_any.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]
        lib.pbrtgpu_intersect_closest_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, _p(C.c_float)]
        lib.pbrtgpu_intersect_any_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, _p(C.c_float)]
        lib.pbrtgpu_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        lib.pbrtgpu_debug_read_layout.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        lib.pbrtgpu_build_hlbvh.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_void_p, _p(C.c_int32), _p(C.c_float)]
        if hasattr(lib, "pbrtgpu_intersect_closest_noalloc"):
            lib.pbrtgpu_intersect_closest_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        if hasattr(lib, "pbrtgpu_render"):
            lib.pbrtgpu_render.argtypes = [C.c_void_p, _p(RenderParams), C.c_void_p, _p(Stats)]
            lib.pbrtgpu_render_device.argtypes = [C.c_void_p, _p(RenderParams), _p(C.c_void_p), _p(Stats)]
        if hasattr(lib, "pbrtgpu_render_device_into"):
            lib.pbrtgpu_render_device_into.argtypes = [C.c_void_p, _p(RenderParams), C.c_void_p, _p(Stats)]
            lib.pbrtgpu_kernel_launches.argtypes = [C.c_void_p]
            lib.pbrtgpu_kernel_launches.restype = C.c_int64
        if hasattr(lib, "pbrtgpu_set_profile"):
            lib.pbrtgpu_set_profile.argtypes = [C.c_void_p, C.c_int]
            lib.pbrtgpu_get_profile.argtypes = [C.c_void_p, _p(Profile)]
            lib.pbrtgpu_copy_film_device.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        if hasattr(lib, "pbrtgpu_halton_samples"):
            lib.pbrtgpu_halton_samples.argtypes = [C.c_void_p, _p(RenderParams), C.c_void_p, C.c_void_p, C.c_void_p,
                                                   C.c_int64, C.c_void_p]
        _gpu = lib
    return _gpu


class HostScene:
    """A parsed, BVH-built, flattened scene (pbrthost_scene)."""

    def __init__(self, path=None, text=None, base_dir=None, quiet=True):
        lib = host_lib()
        h = C.c_void_p()
        if path is not None:
            rc = lib.pbrthost_load_file(os.fsencode(path), int(quiet), C.byref(h))
        else:
            rc = lib.pbrthost_load_string(text.encode(), os.fsencode(base_dir) if base_dir else None, int(quiet), C.byref(h))
        if rc != 0:
            raise PbrtGpuError("scene load failed (status %d); see stderr" % rc)
        self._h = h
        self.view = lib.pbrthost_scene_view(h).contents
        self.params = lib.pbrthost_render_params(h).contents

    def close(self):
        if getattr(self, "_h", None):
            host_lib().pbrthost_free(self._h)
            self._h = None

    __del__ = close

    @property
    def output_filename(self):
        return host_lib().pbrthost_output_filename(self._h).decode()

    def set_spp(self, spp):
        host_lib().pbrthost_set_spp(self._h, int(spp))

    def film_to_rgb(self, film_xyzw):
        cb = self.params.cropped_bounds
        w, h = cb[2] - cb[0], cb[3] - cb[1]
        film = np.ascontiguousarray(film_xyzw, dtype=np.float32).reshape(h, w, 4)
        rgb = np.empty((h, w, 3), np.float32)
        rc = host_lib().pbrthost_film_to_rgb(self._h, film.ctypes.data, rgb.ctypes.data)
        if rc != 0:
            raise PbrtGpuError("film_to_rgb failed")
        return rgb

    def write_image(self, filename, rgb):
        rgb = np.ascontiguousarray(rgb, dtype=np.float32)
        if host_lib().pbrthost_write_image(self._h, os.fsencode(filename), rgb.ctypes.data) != 0:
            raise PbrtGpuError("write_image(%s) failed" % filename)


def last_bvh_build_ms():
    """(wall ms, builder's device ms) of the BVH build of the scene loaded last."""
    w, b = C.c_double(0), C.c_float(0)
    host_lib().pbrthost_last_bvh_build_ms(C.byref(w), C.byref(b))
    return w.value, b.value


def halton_permutations():
    out = np.empty(N_HALTON_PERMS, np.uint16)
    if host_lib().pbrthost_halton_permutations(out.ctypes.data, N_HALTON_PERMS) != 0:
        raise PbrtGpuError("halton_permutations failed")
    return out


class GpuContext:
    """pbrtgpu_ctx: one per GPU."""

    def __init__(self, device=0):
        self._lib = gpu_lib()
        h = C.c_void_p()
        rc = self._lib.pbrtgpu_create(int(device), C.byref(h))
        if rc != 0:
            raise PbrtGpuError("pbrtgpu_create failed (%d): %s" % (rc, self._lib.pbrtgpu_last_error(None).decode()))
        self._h = h
        self.device = device

    def close(self):
        if getattr(self, "_h", None):
            self._lib.pbrtgpu_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            raise PbrtGpuError("pbrtgpu call failed (%d): %s" % (rc, self._lib.pbrtgpu_last_error(self._h).decode()))

    def upload(self, scene):
        view = scene.view if isinstance(scene, HostScene) else scene
        self._check(self._lib.pbrtgpu_upload_scene(self._h, C.byref(view)))

    def build_hlbvh(self, prim_bounds, max_prims_in_node=4):
        """BVHAccel::HLBVHBuild on the device.  prim_bounds: (n, 6) world bounds in creation order.  Returns (bounds (m, 6),
        offset, nprims, axis, ordered_prims, device_ms): the LinearBVHNode arrays and the primitive order."""
        pb = np.ascontiguousarray(prim_bounds, dtype=np.float32).reshape(-1, 6)
        n = len(pb)
        cap = 2 * n
        bounds, offset = np.zeros((cap, 6), np.float32), np.zeros(cap, np.int32)
        nprims, axis, order = np.zeros(cap, np.uint16), np.zeros(cap, np.uint8), np.zeros(n, np.int32)
        m, ms = C.c_int32(0), C.c_float(0)
        self._check(gpu_lib().pbrtgpu_build_hlbvh(self._h, pb.ctypes.data, n, max_prims_in_node, cap, bounds.ctypes.data, offset.ctypes.data,
                                                  nprims.ctypes.data, axis.ctypes.data, order.ctypes.data, C.byref(m), C.byref(ms)))
        return bounds[:m.value], offset[:m.value], nprims[:m.value], axis[:m.value], order, ms.value

    def read_layout(self, scene):
        """The device layout of the uploaded scene as (records (m, 16) uint32, slots (n, 12) uint32)."""
        v = scene.view
        m = 1 + int((np.asarray(v.array("bvh_nprims", v.n_nodes)) == 0).sum())
        nodes, tris = np.zeros((m, 16), np.float32), np.zeros((v.n_prims, 12), np.float32)
        self._check(gpu_lib().pbrtgpu_debug_read_layout(self._h, nodes.ctypes.data, m, tris.ctypes.data, v.n_prims))
        return nodes.view(np.uint32), tris.view(np.uint32)

    def install_hlbvh_builder(self, on=True):
        """Scenes loaded from now on with Accelerator "bvh" "string splitmethod" "hlbvh" get their tree from this GPU
        (pbrthost_set_hlbvh_builder); on=False restores the host build."""
        fn = C.cast(gpu_lib().pbrtgpu_build_hlbvh, C.c_void_p) if on else None
        host_lib().pbrthost_set_hlbvh_builder(fn, self._h if on else None)

    def set_stream(self, cuda_stream_ptr):
        self._check(self._lib.pbrtgpu_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def intersect_closest(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        b = np.empty((n, 2), np.float32)
        self._check(self._lib.pbrtgpu_intersect_closest(self._h, rays.ctypes.data, n, ids.ctypes.data, t.ctypes.data, b.ctypes.data))
        return ids, t, b

    def intersect_closest_noalloc(self, rays):
        """pbrtgpu_intersect_closest_noalloc: at most NOALLOC_MAX_RAYS rays, scratch in a fixed device array."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        b = np.empty((n, 2), np.float32)
        self._check(self._lib.pbrtgpu_intersect_closest_noalloc(self._h, rays.ctypes.data, n, ids.ctypes.data, t.ctypes.data, b.ctypes.data))
        return ids, t, b

    def intersect_any(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        occ = np.empty(n, np.uint8)
        self._check(self._lib.pbrtgpu_intersect_any(self._h, rays.ctypes.data, n, occ.ctypes.data))
        return occ

    def intersect_device(self, d_rays_ptr, n, d_out_ptr, any_hit=False):
        ms = C.c_float()
        fn = self._lib.pbrtgpu_intersect_any_device if any_hit else self._lib.pbrtgpu_intersect_closest_device
        self._check(fn(self._h, C.c_void_p(d_rays_ptr), n, C.c_void_p(d_out_ptr), C.byref(ms)))
        return ms.value

    def render(self, params, want_stats=True, out=None):
        """Integrator::Render: returns (film[h, w, 4] = sum w*X, sum w*Y, sum w*Z, sum w; stats).
        `out` may be a caller-owned float32 array of that shape (e.g. a view of pinned memory)."""
        cb = params.cropped_bounds
        w, h = cb[2] - cb[0], cb[3] - cb[1]
        if out is not None:
            if out.dtype != np.float32 or out.shape != (h, w, 4) or not out.flags["C_CONTIGUOUS"]:
                raise ValueError("out must be a C-contiguous float32 array of shape (%d, %d, 4)" % (h, w))
            film = out
        else:
            film = np.empty((h, w, 4), np.float32)
        stats = Stats()
        self._check(self._lib.pbrtgpu_render(self._h, C.byref(params), film.ctypes.data, C.byref(stats)))
        return film, stats

    def set_profile(self, on):
        self._check(self._lib.pbrtgpu_set_profile(self._h, int(bool(on))))

    def get_profile(self):
        """{kernel: {ms, launches, node_tests, prim_tests}} summed since set_profile(True)."""
        p = Profile()
        self._check(self._lib.pbrtgpu_get_profile(self._h, C.byref(p)))
        return {K_NAMES[k]: {"ms": p.ms[k], "launches": p.launches[k], "node_tests": p.node_tests[k], "prim_tests": p.prim_tests[k]}
                for k in range(6)}

    def copy_film_device(self, dst_ptr, nbytes):
        self._check(self._lib.pbrtgpu_copy_film_device(self._h, C.c_void_p(dst_ptr), int(nbytes)))

    def render_device_into(self, params, d_film_ptr, want_stats=False):
        """Render into a caller-owned device film.  Without stats the call only enqueues (no host synchronisation)."""
        stats = Stats() if want_stats else None
        self._check(self._lib.pbrtgpu_render_device_into(self._h, C.byref(params), C.c_void_p(d_film_ptr), C.byref(stats) if want_stats else None))
        return stats

    def kernel_launches(self):
        return int(self._lib.pbrtgpu_kernel_launches(self._h))

    def render_device(self, params):
        d_film = C.c_void_p()
        stats = Stats()
        self._check(self._lib.pbrtgpu_render_device(self._h, C.byref(params), C.byref(d_film), C.byref(stats)))
        return d_film.value, stats
