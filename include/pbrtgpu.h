/* pbrtgpu.h -- C ABI of libpbrtgpu.so, the B200 (sm_100a) device side of the
 * pbrt-v3-light-portals path integrator.
 *
 * This is the drop-in boundary.  In the reference the
 * seam is `Integrator::Render(const Scene &)` (src/core/integrator.h:53-58),
 * selected by the name "path" in RenderOptions::MakeIntegrator
 * (src/core/api.cpp:1794) and called once from pbrtWorldEnd (api.cpp:1731).
 * A reference-side `GpuPathIntegrator : Integrator` flattens
 * the scene it is handed and calls exactly the entry points below.  All
 * pointers are plain host memory owned by the caller and copied during the
 * call unless a name says `_device`.  No C++/torch types cross this ABI.
 *
 * Every function returns 0 on success and a negative pbrtgpu_status on
 * failure; pbrtgpu_last_error() then describes it.  There is no CPU fallback:
 * without a usable CUDA device pbrtgpu_create fails with PBRTGPU_ERR_NO_DEVICE.
 */
#ifndef PBRTGPU_H
#define PBRTGPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PBRTGPU_ABI_VERSION 3

typedef enum {
    PBRTGPU_OK = 0,
    PBRTGPU_ERR_NO_DEVICE = -1, /* no CUDA device / driver */
    PBRTGPU_ERR_CUDA = -2,      /* a CUDA call failed; see last_error */
    PBRTGPU_ERR_INVALID = -3,   /* bad argument or scene description */
    PBRTGPU_ERR_STATE = -4      /* call order (e.g. render before upload) */
} pbrtgpu_status;

typedef struct pbrtgpu_ctx pbrtgpu_ctx; /* opaque; one per GPU, not thread-safe */

/* primitive kinds met in a BVH leaf (GeometricPrimitive::shape, primitive.cpp:116) */
enum { PBRTGPU_PRIM_TRIANGLE = 0, PBRTGPU_PRIM_AAPLANE = 1, PBRTGPU_PRIM_SPHERE = 2 };

/* prim_flags bits */
enum {
    PBRTGPU_PF_REVERSE_ORIENTATION = 1, /* Shape::reverseOrientation            shape.h */
    PBRTGPU_PF_SWAPS_HANDEDNESS = 2,    /* Shape::transformSwapsHandedness      shape.cpp:49 */
    PBRTGPU_PF_HAS_NORMALS = 4,         /* TriangleMesh::n != nullptr           triangle.h:64 */
    PBRTGPU_PF_HAS_UV = 8,              /* TriangleMesh::uv != nullptr          triangle.h:66 */
    PBRTGPU_PF_HAS_TANGENTS = 16        /* TriangleMesh::s != nullptr           triangle.h:65 */
};

/* material tags: Material::ComputeScatteringFunctions implementations */
enum {
    PBRTGPU_MAT_NONE = 0,
    PBRTGPU_MAT_MATTE = 1,            /* materials/matte.cpp:45-62   params Kd[3] sigma */
    PBRTGPU_MAT_PLASTIC = 2,          /* materials/plastic.cpp:45-70 params Kd[3] Ks[3] alpha */
    PBRTGPU_MAT_GLASS = 3,            /* materials/glass.cpp:45-92   params Kr[3] Kt[3] ualpha valpha eta */
    PBRTGPU_MAT_DISPERSIVE_GLASS = 4, /* materials/dispersive_glass.cpp:48-123 params Kr[3] Kt[3] ualpha valpha etaMin etaMax */
    PBRTGPU_MAT_MIRROR = 5,           /* materials/mirror.cpp        params Kr[3] */
    PBRTGPU_MAT_METAL = 6,            /* materials/metal.cpp         params eta[3] k[3] ualpha valpha */
    PBRTGPU_MAT_UBER = 7              /* materials/uber.cpp:45-105   params Kd[3] Ks[3] Kr[3] Kt[3] opacity[3] ualpha valpha eta (all clamped) */
};
#define PBRTGPU_MAT_STRIDE 24 /* floats per material in mat_params */

/* light tags */
enum {
    PBRTGPU_LIGHT_DIFFUSE_AREA = 1, /* lights/diffuse.cpp          params L[3] twoSided area; shape = BVH-order prim */
    PBRTGPU_LIGHT_PORTAL_AREA = 2,  /* lights/portal_arealight.cpp same + strategy, portal range */
    PBRTGPU_LIGHT_INFINITE = 3,     /* lights/infinite.cpp (constant L, no map) params L[3] */
    PBRTGPU_LIGHT_POINT = 4,        /* lights/point.cpp            params I[3] pos[3] */
    PBRTGPU_LIGHT_DISTANT = 5,      /* lights/distant.cpp          params L[3] wLight[3] */
    PBRTGPU_LIGHT_SPOT = 6          /* lights/spot.cpp             params I[3] pos[3] cosTotal cosFalloff + W2L */
};
#define PBRTGPU_LIGHT_STRIDE 48 /* floats per light in light_params: [0..7] scalars, [8..23] LightToWorld, [24..39] WorldToLight */

/* PortalStrategy, lights/portal_arealight.h:12 (same numbering) */
enum { PBRTGPU_PORTAL_UNIFORM_PORTAL = 0, PBRTGPU_PORTAL_UNIFORM_LIGHT = 1, PBRTGPU_PORTAL_PROJECTION = 2,
       /* not in the reference: solid-angle sampling of the portal rectangle, alone / one-sample MIS with the BSDF (device/portal.h) */
       PBRTGPU_PORTAL_SOLIDANGLE = 3, PBRTGPU_PORTAL_SOLIDANGLE_MIS = 4,
       /* portal / BSDF / light-area sampling combined with the fork's PowerHeuristic3 (sampling.h:177-180), AAPortal::InFrustum
          (aaportal.cpp:101-112, disabled in the reference) re-enabled as a cull of the portal technique */
       PBRTGPU_PORTAL_SOLIDANGLE_MIS3 = 5 };

#define PBRTGPU_PLANE_STRIDE 48  /* floats per AAPlaneShape: lo[3] hi[3] ax facingFw area pad[7] O2W[16] W2O[16]   shapes/plane.h:12-40 */
#define PBRTGPU_SPHERE_STRIDE 48 /* floats per Sphere: radius zMin zMax thetaMin thetaMax phiMax pad[10] O2W[16] W2O[16]  shapes/sphere.h */
#define PBRTGPU_PORTAL_STRIDE 16 /* floats per AAPortal: lo[3] hi[3] ax facingFw area pad[7]; transforms are its light's (aaportal.cpp:12) */

typedef struct {
    /* BVH in the reference's own flattened form: LinearBVHNode[], DFS order,
     * first child at i+1 (accelerators/bvh.cpp:95-104, 640-658). */
    int32_t n_nodes;
    const float *bvh_bounds;    /* n_nodes*6: pMin.xyz pMax.xyz */
    const int32_t *bvh_offset;  /* n_nodes: primitivesOffset (leaf) | secondChildOffset (interior) */
    const uint16_t *bvh_nprims; /* n_nodes: 0 = interior */
    const uint8_t *bvh_axis;    /* n_nodes */

    /* primitives in BVH leaf order (BVHAccel::primitives after the build) */
    int32_t n_prims;
    const uint8_t *prim_type;     /* PBRTGPU_PRIM_* */
    const uint8_t *prim_flags;    /* PBRTGPU_PF_* */
    const int32_t *prim_shape;    /* index into plane_params / sphere_params; unused for triangles */
    const int32_t *prim_material; /* index into materials, -1 = none (GeometricPrimitive::material == nullptr) */
    const int32_t *prim_light;    /* index into lights, -1 = not emissive (primitive.cpp:132) */
    const int32_t *prim_orig_id;  /* creation-order primitive number, the tier-1 ID */
    const float *tri_p;           /* n_prims*9 world-space p0 p1 p2 (triangle.cpp:73-75); zeros for non-triangles */
    const float *tri_n;           /* n_prims*9 or NULL */
    const float *tri_uv;          /* n_prims*6 or NULL */
    const float *tri_s;           /* n_prims*9 or NULL */

    int32_t n_planes;
    const float *plane_params;
    int32_t n_spheres;
    const float *sphere_params;

    int32_t n_materials;
    const int32_t *mat_tag;
    const float *mat_params; /* n_materials * PBRTGPU_MAT_STRIDE */

    /* lights in Scene::lights order (api.cpp:1434-1436 appends area lights as shapes are made) */
    int32_t n_lights;
    const int32_t *light_tag;
    const int32_t *light_prim;     /* BVH-order primitive that carries an area light, else -1 */
    const int32_t *light_strategy; /* PBRTGPU_PORTAL_* for portal lights */
    const int32_t *light_portal_begin; /* n_lights+1 offsets into portal_params */
    const float *light_params;     /* n_lights * PBRTGPU_LIGHT_STRIDE */
    int32_t n_portals;
    const float *portal_params;

    float world_center[3]; /* Scene::WorldBound().BoundingSphere, for infinite/distant lights (infinite.cpp:68-70) */
    float world_radius;
    const int32_t *light_nsamples; /* Light::nSamples ("samples" / "nsamples", diffuse.cpp:110-111, infinite.cpp:158; 1 for delta lights):
                                      read by Integrator "directlighting" strategy "all" only; NULL = 1 each */
} pbrtgpu_scene;

typedef struct {
    /* camera: cameras/perspective.cpp:45-98, core/camera.h:99-119 */
    float camera_to_world[16];
    float raster_to_camera[16];
    float lens_radius, focal_distance;
    float shutter_open, shutter_close;
    /* film: core/film.cpp:45-86, core/film.h:121-161 */
    int32_t full_resolution[2];
    int32_t cropped_bounds[4]; /* x0 y0 x1 y1 */
    int32_t sample_bounds[4];  /* Film::GetSampleBounds */
    int32_t pixel_bounds[4];   /* Integrator "pixelbounds" ∩ sample bounds (path.cpp:197-208) */
    float filter_radius[2];
    float filter_table[256];   /* filterTableWidth = 16 (film.h:83) */
    float max_sample_luminance;
    /* sampler: samplers/halton.cpp:65-138 */
    int32_t spp;
    int32_t sample_at_pixel_center;
    /* integrator: integrators/path.cpp:191-214 */
    int32_t max_depth;
    float rr_threshold;
    int32_t light_strategy; /* 0 uniform, 1 power (lightdistrib.cpp:48-66) */
    /* multi-GPU: this context renders the 16x16 sample tiles t with
     * t % shard_count == shard_index (tiles as in integrator.cpp:533-553) */
    int32_t shard_index, shard_count;
    /* sample range [spp_begin, spp_end) of every pixel; 0,0 = all spp */
    int32_t spp_begin, spp_end;
    /* which SamplerIntegrator::Li: PBRTGPU_INTEGRATOR_PATH (path.cpp:64-189), or
     * PBRTGPU_INTEGRATOR_DIRECT (directlighting.cpp:63-98: direct light at each vertex + specular
     * reflection; max_depth is its "maxdepth", rr_threshold / light_strategy are unused) */
    int32_t integrator;
    int32_t direct_strategy; /* PBRTGPU_DIRECT_ALL (UniformSampleAllLights, integrator.cpp:69-98) or PBRTGPU_DIRECT_ONE */
} pbrtgpu_render_params;
enum { PBRTGPU_INTEGRATOR_PATH = 0, PBRTGPU_INTEGRATOR_DIRECT = 1 };
enum { PBRTGPU_DIRECT_ALL = 0, PBRTGPU_DIRECT_ONE = 1 };

typedef struct {
    int64_t camera_rays;    /* "Integrator/Camera rays traced" (integrator.cpp:52) */
    int64_t closest_rays;   /* Scene::Intersect calls  (scene.cpp:45) */
    int64_t shadow_rays;    /* Scene::IntersectP calls (scene.cpp:51) */
    int64_t scrubbed;       /* NaN / negative / infinite samples zeroed (integrator.cpp:592-613) */
    int64_t kernel_launches;
    int64_t nee_closest_rays; /* the part of closest_rays traced for direct lighting (BSDF-sampled MIS rays, portal rays) */
    float device_ms;        /* CUDA-event time of the render, device only */
    int64_t nee_rays_culled; /* BSDF-sampled MIS rays the reference would trace but which miss the sampled light's own shape:
                                their closest hit cannot be that light, they contribute nothing and are not traced (not in closest_rays) */
    int64_t truncated_paths; /* paths still alive when an open-ended batch reached its iteration cap (a chain of more than ~4000
                                material-less boundaries or pending branches); their radiance so far is splatted.  0 in every parity test */
    int64_t nee_rays_dropped; /* light-sample rays that found their queue full (cannot happen while the queue is sized
                                 nSlots x light samples per vertex; counted so that it can never be a silent bias).  Must be 0 */
} pbrtgpu_stats;

/* Kernels of the wavefront, for per-kernel timing (pbrtgpu_profile). */
enum {
    PBRTGPU_K_RAYGEN = 0,        /* Sampler::GetCameraSample + PerspectiveCamera::GenerateRay */
    PBRTGPU_K_TRACE_CLOSEST = 1, /* BVHAccel::Intersect for path rays */
    PBRTGPU_K_SHADE = 2,         /* the body of PathIntegrator::Li for one vertex */
    PBRTGPU_K_TRACE_SHADOW = 3,  /* BVHAccel::IntersectP for light samples + resolve */
    PBRTGPU_K_TRACE_NEE = 4,     /* BVHAccel::Intersect for BSDF-sampled / portal rays + resolve */
    PBRTGPU_K_FILM = 5,          /* FilmTile::AddSample */
    PBRTGPU_K_COUNT = 6
};
typedef struct {
    double ms[PBRTGPU_K_COUNT];          /* CUDA-event time summed over launches */
    int64_t launches[PBRTGPU_K_COUNT];
    int64_t node_tests[PBRTGPU_K_COUNT]; /* Bounds3::IntersectP calls (trace kernels only) */
    int64_t prim_tests[PBRTGPU_K_COUNT]; /* Shape::Intersect(P) calls (trace kernels only) */
} pbrtgpu_profile;

int pbrtgpu_abi_version(void);
int pbrtgpu_create(int cuda_device, pbrtgpu_ctx **out);
void pbrtgpu_destroy(pbrtgpu_ctx *ctx);
const char *pbrtgpu_last_error(const pbrtgpu_ctx *ctx); /* ctx may be NULL: last create error */

int pbrtgpu_upload_scene(pbrtgpu_ctx *ctx, const pbrtgpu_scene *scene);

/* BVHAccel::HLBVHBuild on the device (accelerators/bvh.cpp:404-638: Morton codes :107-137, 411-428; the stable LSD radix
 * sort :139-180; one emitLBVH treelet per 12-bit Morton prefix :433-455, 474-536; the SAH tree over the treelet roots
 * :538-638, which stays on the host; flattenBVHTree :640-658).  prim_bounds: n*6 world bounds {pMin.xyz pMax.xyz} in creation
 * order (Primitive::WorldBound).  Results, in the reference's flattened form: the LinearBVHNode arrays (node_capacity >= 2 n
 * entries each, *n_nodes_out used) and ordered_prims[slot] = creation-order number of the primitive in BVH slot `slot`.
 * The tree equals the reference's (and pbrthost's host build with "splitmethod" "hlbvh") node for node; leaf slots are numbered
 * as the reference numbers them with --nthreads 1.  elapsed_ms (may be NULL): CUDA-event time of the build kernels.
 * The signature is pbrthost.h's pbrthost_hlbvh_builder with `user` = ctx. */
int pbrtgpu_build_hlbvh(pbrtgpu_ctx *ctx, const float *prim_bounds, int32_t n, int32_t max_prims_in_node, int32_t node_capacity,
                        float *bvh_bounds, int32_t *bvh_offset, uint16_t *bvh_nprims, uint8_t *bvh_axis, int32_t *ordered_prims,
                        int32_t *n_nodes_out, float *elapsed_ms);

/* BVHAccel::Intersect / IntersectP on a ray batch (accelerators/bvh.cpp:662-738).
 * rays: n*8 floats {ox,oy,oz,tMax,dx,dy,dz,pad}.  Host-buffer forms copy in
 * and out; the _device forms work on device pointers and report the kernel's
 * CUDA-event time.  prim_orig_id: -1 = miss.  t: ray.tMax after the call.
 * b0b1: triangle barycentrics, may be NULL. */
int pbrtgpu_intersect_closest(pbrtgpu_ctx *ctx, const float *rays, int64_t n,
                              int32_t *prim_orig_id, float *t, float *b0b1);
int pbrtgpu_intersect_any(pbrtgpu_ctx *ctx, const float *rays, int64_t n, uint8_t *occluded);
/* pbrtgpu_intersect_closest for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same arguments and results as pbrtgpu_intersect_closest, but the device-side staging of rays and hit
 * records and the kernel's queue cursor live in one file-scope device array (csrc/device/noalloc.cu) that every GPU
 * loading the library holds until exit: 2 MiB + 256 B.  No cudaMalloc / cudaFree, no host allocation on the success
 * path.  n > PBRTGPU_NOALLOC_MAX_RAYS is refused with PBRTGPU_ERR_INVALID.  All calls on one GPU share the array (it
 * keeps the last call's contents; the cursor is cleared by every call): calls that overlap on a GPU race -- the
 * caller serialises them. */
#define PBRTGPU_NOALLOC_MAX_RAYS 32768
int pbrtgpu_intersect_closest_noalloc(pbrtgpu_ctx *ctx, const float *rays, int64_t n,
                                      int32_t *prim_orig_id, float *t, float *b0b1);
int pbrtgpu_intersect_closest_device(pbrtgpu_ctx *ctx, const void *d_rays, int64_t n,
                                     void *d_hits /* n*16 B {origId,t,b0,b1} */, float *elapsed_ms);
int pbrtgpu_intersect_any_device(pbrtgpu_ctx *ctx, const void *d_rays, int64_t n,
                                 void *d_occluded /* n bytes */, float *elapsed_ms);

/* SamplerIntegrator::Render + PathIntegrator::Li (integrator.cpp:526-637,
 * path.cpp:64-189).  film_xyzw: cropped W*H*4 floats {sum w*X, sum w*Y,
 * sum w*Z, sum w} exactly as Film::Pixel accumulates them (film.h:94-101);
 * normalisation and XYZ->RGB stay with the caller's Film::WriteImage. */
int pbrtgpu_render(pbrtgpu_ctx *ctx, const pbrtgpu_render_params *params,
                   float *film_xyzw, pbrtgpu_stats *stats_or_null);
/* Same, film left on the device (for the NCCL reduce); *d_film is owned by ctx. */
int pbrtgpu_render_device(pbrtgpu_ctx *ctx, const pbrtgpu_render_params *params,
                          void **d_film, pbrtgpu_stats *stats_or_null);
/* Same, accumulated straight in a caller-owned device buffer (cropped W*H*4 floats, 16-byte aligned; e.g. the tensor
 * handed to NCCL -- Film::MergeFilmTile film.cpp:117-130 across GPUs): no staging copy.  With stats_or_null == NULL the
 * call does not wait for the GPU: it returns once the render is enqueued on the ctx's stream, so the caller can order an
 * ncclReduce of this film behind it and start the next render while the reduce runs (renders whose iteration count is
 * not known in advance -- material-less primitives, directlighting over specular transmission -- still synchronise
 * between rounds).  The buffer must stay valid until that work has run. */
int pbrtgpu_render_device_into(pbrtgpu_ctx *ctx, const pbrtgpu_render_params *params, void *d_film,
                               pbrtgpu_stats *stats_or_null);
/* Kernels launched by this context so far (host-side count, no device read-back). */
int64_t pbrtgpu_kernel_launches(const pbrtgpu_ctx *ctx);
/* Copy the film left by pbrtgpu_render_device into a caller-owned device buffer (e.g. a tensor handed to NCCL). */
int pbrtgpu_copy_film_device(pbrtgpu_ctx *ctx, void *d_dst, int64_t bytes);
/* Profile mode: renders record a CUDA-event pair around every kernel launch and count
 * traversal steps; pbrtgpu_get_profile returns the sums since profile mode was last set. */
int pbrtgpu_set_profile(pbrtgpu_ctx *ctx, int on);
int pbrtgpu_get_profile(pbrtgpu_ctx *ctx, pbrtgpu_profile *out);
/* Run the render on this CUDA stream (cudaStream_t) instead of the ctx's own. */
int pbrtgpu_set_stream(pbrtgpu_ctx *ctx, void *cuda_stream);

/* Unit-parity hooks: evaluate device functions on small batches so tests can
 * compare them with the oracle one function at a time. */
int pbrtgpu_halton_samples(pbrtgpu_ctx *ctx, const pbrtgpu_render_params *params,
                           const int32_t *pixel_xy, const int64_t *sample_num, const int32_t *dim,
                           int64_t n, float *out); /* HaltonSampler::SampleDimension */

/* The device layout of the uploaded scene -- 64-byte two-child records (n_records = interior nodes + 1) and 48-byte
 * primitive slots, csrc/device/bvh.h -- copied back, so that the conversion kernels run by pbrtgpu_upload_scene can be
 * compared byte for byte with the host restatement (PBRTGPU_LAYOUT=host). */
int pbrtgpu_debug_read_layout(pbrtgpu_ctx *ctx, float *nodes /* n_records*16 */, int64_t n_records, float *tris /* n_prims*12 */,
                              int64_t n_prims);

#ifdef __cplusplus
}
#endif
#endif /* PBRTGPU_H */
