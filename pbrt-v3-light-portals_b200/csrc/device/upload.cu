// Context lifetime and scene upload: reference-layout arrays (pbrtgpu_scene)
// -> device layout (bvh.h).  Host code only; no kernels here.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../common/halton_tables.h"
#include "ctx.h"
#include "layout_build.h"

using namespace pg;

static std::string gCreateError;

namespace pg {
int buildLayoutOnDevice(pbrtgpu_ctx *ctx, const pbrtgpu_scene *s, int nInterior, const float *d_triUV, const unsigned char *d_primFlags,
                        F4 *d_nodes, F4 *d_tris, int *firstBadHost);
}

namespace pg {
int ensureBuffer(pbrtgpu_ctx *ctx, DeviceBuffer &b, size_t bytes) {
    if (b.bytes >= bytes) return PBRTGPU_OK;
    if (b.p) PG_CUDA(ctx, cudaFree(b.p));
    b.p = nullptr;
    b.bytes = 0;
    PG_CUDA(ctx, cudaMalloc(&b.p, bytes));
    b.bytes = bytes;
    return PBRTGPU_OK;
}
}  // namespace pg

static void drainScenePool(pbrtgpu_ctx *ctx) {
    for (auto &b : ctx->scenePool) cudaFree(b.p);
    ctx->scenePool.clear();
}
// keepForReuse: a re-upload of a scene of the same shape then needs no cudaMalloc / cudaFree
static void freeScene(pbrtgpu_ctx *ctx, bool keepForReuse = false) {
    drainScenePool(ctx);
    if (keepForReuse) ctx->scenePool.swap(ctx->sceneAllocs);
    else
        for (auto &b : ctx->sceneAllocs) cudaFree(b.p);
    ctx->sceneAllocs.clear();
    ctx->haveScene = false;
}

extern "C" {

int pbrtgpu_abi_version(void) { return PBRTGPU_ABI_VERSION; }

const char *pbrtgpu_last_error(const pbrtgpu_ctx *ctx) { return ctx ? ctx->lastError.c_str() : gCreateError.c_str(); }

int pbrtgpu_create(int cuda_device, pbrtgpu_ctx **out) {
    if (!out) return PBRTGPU_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        gCreateError = std::string("no CUDA device available: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                       " (libpbrtgpu has no CPU fallback)";
        return PBRTGPU_ERR_NO_DEVICE;
    }
    if (cuda_device < 0 || cuda_device >= count) {
        gCreateError = "cuda_device out of range";
        return PBRTGPU_ERR_INVALID;
    }
    pbrtgpu_ctx *ctx = new pbrtgpu_ctx;
    ctx->device = cuda_device;
    auto bail = [&](cudaError_t err, const char *what) {
        gCreateError = std::string(what) + ": " + cudaGetErrorString(err);
        delete ctx;
        return PBRTGPU_ERR_CUDA;
    };
    if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    ctx->ownStream = true;
    cudaDeviceGetAttribute(&ctx->smCount, cudaDevAttrMultiProcessorCount, cuda_device);
    if (const char *v = getenv("PBRTGPU_TRACE_PERSISTENT")) ctx->tracePersistent = atoi(v) != 0;
    if (const char *v = getenv("PBRTGPU_TRACE_SHARED_SCENE")) ctx->traceSharedScene = atoi(v) != 0;
    if (const char *v = getenv("PBRTGPU_TRACE_FUSED")) ctx->traceFused = atoi(v) != 0;
    if (const char *v = getenv("PBRTGPU_SHADE_SYNC")) ctx->shadeSync = atoi(v) != 0;
    if (const char *v = getenv("PBRTGPU_LIGHT_STAGE_BLOCKS")) ctx->lightStageBlocksPerSM = std::min(32, std::max(1, atoi(v)));
    if (const char *v = getenv("PBRTGPU_LAYOUT")) ctx->layoutOnDevice = strcmp(v, "host") != 0;
    if (const char *v = getenv("PBRTGPU_SORT_HITS")) ctx->sortHits = atoi(v) < 0 ? 0 : (atoi(v) > 2 ? 2 : atoi(v));
    if (const char *v = getenv("PBRTGPU_SHADE_LAMBERT")) ctx->shadeLambert = atoi(v) != 0;
    if (const char *v = getenv("PBRTGPU_LAMBERT_BLOCK")) ctx->lambertBlock = atoi(v);
    if (const char *v = getenv("PBRTGPU_FETCH_THRESHOLD")) ctx->fetchThreshold = std::min(32, std::max(1, atoi(v)));
    if (const char *v = getenv("PBRTGPU_SHADE_BLOCK")) ctx->shadeBlock = atoi(v);
    if (const char *v = getenv("PBRTGPU_PHASE_BIAS")) ctx->phaseBias = std::min(1024, std::max(0, atoi(v)));
    if (const char *v = getenv("PBRTGPU_PERSIST_BLOCKS")) ctx->persistBlocksPerSM = std::min(16, std::max(1, atoi(v)));
    if ((e = cudaMalloc((void **)&ctx->dFetch, sizeof(int))) != cudaSuccess) return bail(e, "cudaMalloc(fetch)");
    // Halton digit permutations (7.4 MB of uint16), generated, not stored
    std::vector<uint16_t> perms = radicalInversePermutations(firstPrimes());
    if ((e = cudaMalloc((void **)&ctx->dPerms, perms.size() * sizeof(uint16_t))) != cudaSuccess) return bail(e, "cudaMalloc(perms)");
    if ((e = cudaMemcpy(ctx->dPerms, perms.data(), perms.size() * sizeof(uint16_t), cudaMemcpyHostToDevice)) != cudaSuccess)
        return bail(e, "cudaMemcpy(perms)");
    {
        std::vector<int> primes = firstPrimes(), sums(primes.size());
        int acc = 0;
        for (size_t i = 0; i < primes.size(); ++i) sums[i] = acc, acc += primes[i];
        if ((e = cudaMalloc((void **)&ctx->dPrimes, primes.size() * sizeof(int))) != cudaSuccess) return bail(e, "cudaMalloc(primes)");
        if ((e = cudaMalloc((void **)&ctx->dPrimeSums, primes.size() * sizeof(int))) != cudaSuccess) return bail(e, "cudaMalloc(primeSums)");
        if ((e = cudaMalloc((void **)&ctx->dTotals, 8 * sizeof(unsigned long long))) != cudaSuccess) return bail(e, "cudaMalloc(totals)");
        cudaMemcpy(ctx->dPrimes, primes.data(), primes.size() * sizeof(int), cudaMemcpyHostToDevice);
        cudaMemcpy(ctx->dPrimeSums, sums.data(), primes.size() * sizeof(int), cudaMemcpyHostToDevice);
        std::vector<unsigned long long> magic = primeMagics(primes);
        if ((e = cudaMalloc((void **)&ctx->dPrimeMagic, magic.size() * 8)) != cudaSuccess) return bail(e, "cudaMalloc(primeMagic)");
        cudaMemcpy(ctx->dPrimeMagic, magic.data(), magic.size() * 8, cudaMemcpyHostToDevice);
        std::vector<HaltonDim> dims(primes.size());
        fillHaltonDims(primes.data(), sums.data(), perms.data(), (int)primes.size(), dims.data());
        if ((e = cudaMalloc((void **)&ctx->dHaltonDims, dims.size() * sizeof(HaltonDim))) != cudaSuccess) return bail(e, "cudaMalloc(haltonDims)");
        cudaMemcpy(ctx->dHaltonDims, dims.data(), dims.size() * sizeof(HaltonDim), cudaMemcpyHostToDevice);
    }
    *out = ctx;
    return PBRTGPU_OK;
}

void pbrtgpu_destroy(pbrtgpu_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    freeScene(ctx);
    for (auto &b : ctx->scratch)
        if (b.p) cudaFree(b.p);
    if (ctx->film.p) cudaFree(ctx->film.p);
    if (ctx->layoutScratch.p) cudaFree(ctx->layoutScratch.p);
    if (ctx->dPerms) cudaFree(ctx->dPerms);
    if (ctx->dPrimes) cudaFree(ctx->dPrimes);
    if (ctx->dPrimeSums) cudaFree(ctx->dPrimeSums);
    if (ctx->dPrimeMagic) cudaFree(ctx->dPrimeMagic);
    if (ctx->dHaltonDims) cudaFree(ctx->dHaltonDims);
    if (ctx->dTotals) cudaFree(ctx->dTotals);
    if (ctx->dFetch) cudaFree(ctx->dFetch);
    if (ctx->dTraversalCounts) cudaFree(ctx->dTraversalCounts);
    for (auto &b : ctx->wave)
        if (b.p) cudaFree(b.p);
    if (ctx->waveParams.p) cudaFree(ctx->waveParams.p);
    if (ctx->lightDist.p) cudaFree(ctx->lightDist.p);
    if (ctx->filterTable.p) cudaFree(ctx->filterTable.p);
    if (ctx->ownStream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int pbrtgpu_set_stream(pbrtgpu_ctx *ctx, void *cuda_stream) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (ctx->ownStream && ctx->stream) cudaStreamDestroy(ctx->stream);
    ctx->stream = (cudaStream_t)cuda_stream;
    ctx->ownStream = false;
    return PBRTGPU_OK;
}

int pbrtgpu_upload_scene(pbrtgpu_ctx *ctx, const pbrtgpu_scene *s) {
    if (!ctx || !s) return PBRTGPU_ERR_INVALID;
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    if (s->n_nodes <= 0 || s->n_prims <= 0) return ctx->fail(PBRTGPU_ERR_INVALID, "scene has no primitives");
    if (!s->bvh_bounds || !s->bvh_offset || !s->bvh_nprims || !s->bvh_axis || !s->prim_type || !s->tri_p || !s->prim_orig_id)
        return ctx->fail(PBRTGPU_ERR_INVALID, "scene arrays missing");
    // validate the tree shape, so that traversal cannot run away on bad input ...
    int nInterior = 0;
    for (int i = 0; i < s->n_nodes; ++i) {
        if (s->bvh_nprims[i] == 0) {
            ++nInterior;
            if (i + 1 >= s->n_nodes || s->bvh_offset[i] <= i || s->bvh_offset[i] >= s->n_nodes || s->bvh_axis[i] > 2)
                return ctx->fail(PBRTGPU_ERR_INVALID, "BVH node %d: bad child offset / axis", i);
        } else if (s->bvh_offset[i] < 0 || s->bvh_offset[i] + (int)s->bvh_nprims[i] > s->n_prims)
            return ctx->fail(PBRTGPU_ERR_INVALID, "BVH leaf %d: primitive range out of bounds", i);
    }
    // ... and every index the kernels use to address a table, so that a bad caller gets PBRTGPU_ERR_INVALID, not a device fault
    if ((s->n_materials > 0 && (!s->mat_tag || !s->mat_params)) || (s->n_lights > 0 && (!s->light_tag || !s->light_params || !s->light_prim)))
        return ctx->fail(PBRTGPU_ERR_INVALID, "material / light tables missing");
    for (int i = 0; i < s->n_materials; ++i)
        if (s->mat_tag[i] < PBRTGPU_MAT_NONE || s->mat_tag[i] > PBRTGPU_MAT_UBER) return ctx->fail(PBRTGPU_ERR_INVALID, "material %d: unknown tag %d", i, s->mat_tag[i]);
    for (int i = 0; i < s->n_prims; ++i) {
        if (s->prim_type[i] > PBRTGPU_PRIM_SPHERE) return ctx->fail(PBRTGPU_ERR_INVALID, "primitive %d: unknown type %d", i, (int)s->prim_type[i]);
        if (s->prim_material && (s->prim_material[i] < -1 || s->prim_material[i] >= s->n_materials))
            return ctx->fail(PBRTGPU_ERR_INVALID, "primitive %d: material index %d out of range", i, s->prim_material[i]);
        if (s->prim_light && (s->prim_light[i] < -1 || s->prim_light[i] >= s->n_lights))
            return ctx->fail(PBRTGPU_ERR_INVALID, "primitive %d: light index %d out of range", i, s->prim_light[i]);
    }
    for (int i = 0; i < s->n_lights; ++i) {
        const int tag = s->light_tag[i];
        if (tag < PBRTGPU_LIGHT_DIFFUSE_AREA || tag > PBRTGPU_LIGHT_SPOT) return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: unknown tag %d", i, tag);
        const bool area = tag == PBRTGPU_LIGHT_DIFFUSE_AREA || tag == PBRTGPU_LIGHT_PORTAL_AREA;
        if (area && (s->light_prim[i] < 0 || s->light_prim[i] >= s->n_prims))
            return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: area light without a valid primitive (%d)", i, s->light_prim[i]);
        if (tag == PBRTGPU_LIGHT_PORTAL_AREA) {
            if (s->prim_type[s->light_prim[i]] != PBRTGPU_PRIM_AAPLANE) return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: a portal light's shape must be an aaplane", i);
            if (!s->light_strategy || s->light_strategy[i] < PBRTGPU_PORTAL_UNIFORM_PORTAL || s->light_strategy[i] > PBRTGPU_PORTAL_SOLIDANGLE_MIS3)
                return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: unknown portal strategy", i);
            if (!s->light_portal_begin) return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: light_portal_begin missing", i);
        }
    }
    if (s->light_portal_begin) {
        if (s->n_portals > 0 && !s->portal_params) return ctx->fail(PBRTGPU_ERR_INVALID, "portal_params missing");
        if (s->n_lights > 0 && s->light_portal_begin[0] < 0) return ctx->fail(PBRTGPU_ERR_INVALID, "light_portal_begin must start at 0 or above");
        for (int i = 0; i < s->n_lights; ++i)
            if (s->light_portal_begin[i + 1] < s->light_portal_begin[i] || s->light_portal_begin[i + 1] > s->n_portals)
                return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: portal range [%d, %d) not inside [0, %d)", i, s->light_portal_begin[i], s->light_portal_begin[i + 1], s->n_portals);
    }
    freeScene(ctx, /*keepForReuse=*/true);
    DevScene &d = ctx->scene;
    memset(&d, 0, sizeof(d));

    int rc;
#define UP(host, count, dev) \
    if ((rc = uploadArray(ctx, host, (size_t)(count), dev)) != PBRTGPU_OK) return rc;
    UP(s->prim_flags, s->prim_flags ? s->n_prims : 0, &d.primFlags);
    UP(s->tri_uv, s->tri_uv ? 6 * (size_t)s->n_prims : 0, &d.triUV);
    int firstBadPrim = 0x7fffffff;
    ctx->nWideRecords = 1 + nInterior;
    if (ctx->layoutOnDevice) {
        // the raw arrays go up as they are and kernels write the records (layout_device.cu)
        void *nodes = nullptr, *tris = nullptr;
        if ((rc = allocSceneArray(ctx, 64 * (size_t)(1 + nInterior), &nodes)) != PBRTGPU_OK) return rc;
        if ((rc = allocSceneArray(ctx, 48 * (size_t)s->n_prims, &tris)) != PBRTGPU_OK) return rc;
        if ((rc = buildLayoutOnDevice(ctx, s, nInterior, d.triUV, d.primFlags, (F4 *)nodes, (F4 *)tris, &firstBadPrim)) != PBRTGPU_OK) return rc;
        d.bvh.nodes = (const F4 *)nodes;
        d.bvh.tris = (const F4 *)tris;
    } else {
        // host conversion (layout_build.h): primitives on a second thread while this one does the nodes
        std::vector<F4> wide, tris;
        std::string err;
        bool trisOk = true;
        std::thread trisThread([&] { trisOk = buildTris(s, &tris, &err); });
        buildWideNodes(s, &wide);
        trisThread.join();
        if (!trisOk) return ctx->fail(PBRTGPU_ERR_INVALID, "%s", err.c_str());
        UP(wide.data(), wide.size(), &d.bvh.nodes);
        UP(tris.data(), tris.size(), &d.bvh.tris);
        PG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // the staging vectors go out of scope
    }
    UP(s->plane_params, (size_t)s->n_planes * PBRTGPU_PLANE_STRIDE, &d.bvh.planeParams);
    UP(s->sphere_params, (size_t)s->n_spheres * PBRTGPU_SPHERE_STRIDE, &d.bvh.sphereParams);
    d.nPrims = s->n_prims;
    d.nLights = s->n_lights;
    d.nMaterials = s->n_materials;
    UP(s->prim_orig_id, s->n_prims, &d.primOrigId);
    UP(s->prim_material, s->prim_material ? s->n_prims : 0, &d.primMaterial);
    UP(s->prim_light, s->prim_light ? s->n_prims : 0, &d.primLight);
    UP(s->tri_n, s->tri_n ? 9 * (size_t)s->n_prims : 0, &d.triN);
    UP(s->tri_s, s->tri_s ? 9 * (size_t)s->n_prims : 0, &d.triS);
    UP(s->mat_tag, s->n_materials, &d.matTag);
    UP(s->mat_params, (size_t)s->n_materials * PBRTGPU_MAT_STRIDE, &d.matParams);
    UP(s->light_tag, s->n_lights, &d.lightTag);
    UP(s->light_prim, s->n_lights, &d.lightPrim);
    UP(s->light_strategy, s->n_lights, &d.lightStrategy);
    ctx->extPortalStrategy = false;
    ctx->hasPortalLight = false;
    for (int i = 0; i < s->n_lights; ++i)
        if (s->light_tag[i] == PBRTGPU_LIGHT_PORTAL_AREA) ctx->hasPortalLight = true;
    for (int i = 0; i < s->n_lights; ++i)
        if (s->light_tag[i] == PBRTGPU_LIGHT_PORTAL_AREA && s->light_strategy && s->light_strategy[i] >= PBRTGPU_PORTAL_SOLIDANGLE) ctx->extPortalStrategy = true;
    {
        std::vector<MatLobes> lobes;
        buildMaterialLobes(s, &lobes);
        UP(lobes.data(), lobes.size(), &d.matLobes);
        ctx->hasDispersiveMaterial = false;
        for (const MatLobes &m : lobes)
            if (m.dispersive) ctx->hasDispersiveMaterial = true;
        ctx->hasSpecularOnlyMaterial = false;
        for (int i = 0; i < s->n_materials; ++i)  // the path integrator's half of the table
            if (lobes[i].shadeClass == 2) ctx->hasSpecularOnlyMaterial = true;
        ctx->maxSpecTransLobes = maxSpecularTransmissionLobes(lobes, s->n_materials, &ctx->specTransDispersive);
    }
    UP(s->light_portal_begin, s->light_portal_begin ? s->n_lights + 1 : 0, &d.lightPortalBegin);
    UP(s->light_nsamples, s->light_nsamples ? s->n_lights : 0, &d.lightNSamples);
    ctx->sumLightSamples = 0;
    for (int i = 0; i < s->n_lights; ++i) {
        int n = s->light_nsamples ? s->light_nsamples[i] : 1;
        if (n < 1) return ctx->fail(PBRTGPU_ERR_INVALID, "light %d: nsamples must be positive", i);
        ctx->sumLightSamples += n;
    }
    std::vector<float> lightParams(s->light_params, s->light_params + (size_t)s->n_lights * PBRTGPU_LIGHT_STRIDE);
    for (int i = 0; i < s->n_lights; ++i)
        if (s->light_tag[i] == PBRTGPU_LIGHT_INFINITE) finishInfiniteLight(lightParams.data() + (size_t)i * PBRTGPU_LIGHT_STRIDE);
    UP(lightParams.data(), lightParams.size(), &d.lightParams);
    ctx->hasNullMaterial = false;
    ctx->hasInfiniteLight = false;
    for (int i = 0; i < s->n_lights; ++i)
        if (s->light_tag[i] == PBRTGPU_LIGHT_INFINITE) ctx->hasInfiniteLight = true;
    if (s->prim_material)
        for (int i = 0; i < s->n_prims; ++i)
            if (s->prim_material[i] < 0) ctx->hasNullMaterial = true;
    UP(s->portal_params, (size_t)s->n_portals * PBRTGPU_PORTAL_STRIDE, &d.portalParams);
#undef UP
    {   // light-pick distributions for both "lightsamplestrategy" values (uniform, power)
        std::vector<float> all;
        for (int strategy = 0; strategy < 2; ++strategy) {
            std::vector<float> func, cdf;
            pbrtgpu_scene tmp = *s;
            tmp.light_params = lightParams.data();
            buildLightDistribution(&tmp, strategy, &func, &cdf, &ctx->lightFuncInt[strategy]);
            all.insert(all.end(), func.begin(), func.end());
            all.insert(all.end(), cdf.begin(), cdf.end());
        }
        if ((rc = ensureBuffer(ctx, ctx->lightDist, all.size() * sizeof(float))) != PBRTGPU_OK) return rc;
        PG_CUDA(ctx, cudaMemcpy(ctx->lightDist.p, all.data(), all.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    memcpy(d.worldCenter, s->world_center, sizeof(d.worldCenter));
    d.worldRadius = s->world_radius;
    PG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    drainScenePool(ctx);
    if (firstBadPrim != 0x7fffffff) return ctx->fail(PBRTGPU_ERR_INVALID, "primitive %d: shape index out of range", firstBadPrim);
    ctx->haveScene = true;
    return PBRTGPU_OK;
}

// Unit-parity hook: the device layout of the uploaded scene (bvh.h), for the byte-for-byte test of the device-side
// conversion against layout_build.h.  nodes: 16 floats per record, tris: 12 floats per primitive slot.
int pbrtgpu_debug_read_layout(pbrtgpu_ctx *ctx, float *nodes, int64_t n_records, float *tris, int64_t n_prims) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!ctx->haveScene) return ctx->fail(PBRTGPU_ERR_STATE, "no scene uploaded");
    if (n_records != ctx->nWideRecords || n_prims != ctx->scene.nPrims)
        return ctx->fail(PBRTGPU_ERR_INVALID, "the uploaded scene has %lld records and %d primitives", ctx->nWideRecords, ctx->scene.nPrims);
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    PG_CUDA(ctx, cudaMemcpy(nodes, ctx->scene.bvh.nodes, 64 * (size_t)n_records, cudaMemcpyDeviceToHost));
    PG_CUDA(ctx, cudaMemcpy(tris, ctx->scene.bvh.tris, 48 * (size_t)n_prims, cudaMemcpyDeviceToHost));
    return PBRTGPU_OK;
}

}  // extern "C"
