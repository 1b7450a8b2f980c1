// Device context behind the opaque pbrtgpu_ctx handle.
#ifndef PG_DEVICE_CTX_H
#define PG_DEVICE_CTX_H

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include "../../../include/pbrtgpu.h"
#include "ctx_scene.h"
#include "kernels.h"

namespace pg {

struct DeviceBuffer {
    void *p = nullptr;
    size_t bytes = 0;
};

}  // namespace pg

struct pbrtgpu_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool ownStream = false;
    std::string lastError;
    bool haveScene = false;
    pg::DevScene scene;
    struct Block {
        void *p;
        size_t bytes;
    };
    std::vector<Block> sceneAllocs;  // the scene's device arrays
    std::vector<Block> scenePool;    // the previous scene's arrays during a re-upload: reused when a size matches
    // Halton tables (uploaded at create)
    unsigned short *dPerms = nullptr;
    unsigned long long *dPrimeMagic = nullptr;
    pg::HaltonDim *dHaltonDims = nullptr;  // per-dimension records of the Halton sampler (sampler.h)
    int *dPrimes = nullptr, *dPrimeSums = nullptr;  // first 1000 primes and their running sums
    // wavefront state of the path integrator (render.cu), grown on demand
    pg::DeviceBuffer wave[24];
    pg::DeviceBuffer waveParams;  // the current batch's WaveParams, read by the shade / raygen / resolve kernels
    pg::DeviceBuffer lightDist;                     // {func[n], cdf[n+1]} x {uniform, power}: the light-pick distributions
    float lightFuncInt[2] = {0, 0};
    pg::DeviceBuffer filterTable;
    unsigned long long *dTotals = nullptr;          // cameraRays, closestRays, shadowRays, scrubbed
    bool profile = false;                           // pbrtgpu_set_profile: event pairs per launch + traversal counters
    double profMs[PBRTGPU_K_COUNT] = {0};
    long long profLaunches[PBRTGPU_K_COUNT] = {0};
    unsigned long long *dTraversalCounts = nullptr;
    int maxSpecTransLobes = 0;                      // per material, allowMultipleLobes = false: > 0 makes directlighting a tree
    bool specTransDispersive = false;
    long long sumLightSamples = 0;                  // sum of Light::nSamples (directlighting "all")
    bool hasInfiniteLight = false;
    bool hasDispersiveMaterial = false;             // some material's index of refraction follows the path's wavelength
    bool hasNullMaterial = false;                   // some primitive has no material (path.cpp:108-113)
    // render scratch, grown on demand
    pg::DeviceBuffer scratch[16];
    pg::DeviceBuffer film;
    pg::DeviceBuffer layoutScratch;  // raw node / primitive arrays of the upload in progress (layout_device.cu), kept for the next one
    long long nWideRecords = 0;      // interior records of the uploaded scene, the synthetic root parent included
    bool layoutOnDevice = true;      // PBRTGPU_LAYOUT=host: build the device layout with layout_build.h on the CPU instead
    void *filmTarget = nullptr;  // pbrtgpu_render_device_into: the caller's device film for the render in progress (else `film`)
    bool asyncRender = false;    // ... and, without a stats request, no host synchronisation at the end of the render
    int smCount = 148;
    // trace kernel variant (tuning knobs, env PBRTGPU_TRACE_PERSISTENT / _FETCH_THRESHOLD / _PERSIST_BLOCKS)
    bool tracePersistent = true;
    bool traceFused = true;   // one traversal launch per depth for path + next-event + shadow rays
    bool traceSharedScene = true;  // ... which stages the whole scene in shared memory when it fits (trace_persistent.h: kSharedSceneBytes)
    bool shadeSync = true;
    int lightStageBlocksPerSM = 16;  // grid of k_split_hits / k_post_trace, in 128-thread blocks per SM
    bool extPortalStrategy = false;  // a portal light of the uploaded scene uses a strategy the reference does not have
    int sortHits = 2;           // hit queue split by material class (render.cu:splitHits): 0 = off, 1 = Lambert | rest, 2 = Lambert | rest | all-specular
    bool hasSpecularOnlyMaterial = false;  // some material of the uploaded scene is MatLobes::shadeClass 2
    bool shadeLambert = true;   // the front run of the split queue is shaded by the Lambert instantiation of the shade kernel
    int lambertBlock = 5122;    // its block shape: 5122 = 512 threads x 2 per SM, 5121, 3842, 2563, 2562 (A/B knob)
    bool hasPortalLight = false;
    bool shadeSmemSet = false;  // k_shade_sync's shared-memory opt-in done on this device
    int shadeBlock = 5122;      // 512-thread blocks, two per SM (A/B in profiles/r01g_ab_shade.log)
    int fetchThreshold = 12;
    int phaseBias = 16;
    int persistBlocksPerSM = pg::kPersistMinBlocks;
    int *dFetch = nullptr;  // queue cursor of the ray-batch entry points
    long long launches = 0;

    int fail(int code, const char *fmt, ...) {
        char buf[1024];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        lastError = buf;
        return code;
    }
};

#define PG_CUDA(ctx, call)                                                                              \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return (ctx)->fail(PBRTGPU_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), \
                               __FILE__, __LINE__);                                                     \
    } while (0)

namespace pg {
int ensureBuffer(pbrtgpu_ctx *ctx, DeviceBuffer &b, size_t bytes);
// a device array that belongs to the uploaded scene: taken from the previous scene's arrays when a size matches
inline int allocSceneArray(pbrtgpu_ctx *ctx, size_t bytes, void **out) {
    void *d = nullptr;
    for (size_t i = 0; i < ctx->scenePool.size(); ++i)
        if (ctx->scenePool[i].bytes == bytes) {
            d = ctx->scenePool[i].p;
            ctx->scenePool.erase(ctx->scenePool.begin() + i);
            break;
        }
    if (!d) PG_CUDA(ctx, cudaMalloc(&d, bytes));
    ctx->sceneAllocs.push_back({d, bytes});
    *out = d;
    return PBRTGPU_OK;
}
template <typename T>
int uploadArray(pbrtgpu_ctx *ctx, const T *host, size_t count, const T **dev) {
    *dev = nullptr;
    if (!host || count == 0) return PBRTGPU_OK;
    void *d = nullptr;
    const size_t bytes = count * sizeof(T);
    for (size_t i = 0; i < ctx->scenePool.size(); ++i)
        if (ctx->scenePool[i].bytes == bytes) {
            d = ctx->scenePool[i].p;
            ctx->scenePool.erase(ctx->scenePool.begin() + i);
            break;
        }
    if (!d) PG_CUDA(ctx, cudaMalloc(&d, bytes));
    ctx->sceneAllocs.push_back({d, bytes});
    PG_CUDA(ctx, cudaMemcpyAsync(d, host, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    *dev = (const T *)d;
    return PBRTGPU_OK;
}
}  // namespace pg

#endif
