// pbrtgpu_intersect_closest_noalloc: BVHAccel::Intersect on a small ray batch for callers that must not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.  Nothing else in the library calls into this file.
//
// pbrtgpu_intersect_closest stages the rays and the hit records in cudaMalloc'ed scratch (ctx->scratch) and hands
// the persistent kernel a queue cursor from the heap.  Here all of that lives in ONE file-scope device array of a
// fixed size, which every GPU that loads the library holds until the process exits:
//
//     [ cursor (256 B) | rays: kMax x 32 B | hits: kMax x 16 B | ids: kMax x 4 B | t: kMax x 4 B | b0b1: kMax x 8 B ]
//
// kMax = PBRTGPU_NOALLOC_MAX_RAYS = 32 768 rays -> 2 MiB + 256 B per GPU.  The call makes no cudaMalloc / cudaFree and
// no host allocation on its success path: H2D of the caller's rays into the array, memset of the cursor, the same
// persistent traversal (trace_persistent.h) as k_pintersect_closest, a kernel that splits the hit records into the
// three output columns, three D2H copies straight into the caller's arrays.
//
// Every call on a device shares the array and finds the previous call's contents in it: the cursor -- the only part
// that must start at zero -- is cleared by every call, everything else is overwritten before it is read.  Calls that
// overlap on one GPU (two contexts, two host threads) therefore RACE; the caller serialises them.
#include <algorithm>
#include <mutex>

#include "ctx.h"
#include "kernels.h"
#include "trace_persistent.h"

using namespace pg;

namespace {

constexpr long long kMax = PBRTGPU_NOALLOC_MAX_RAYS;
constexpr size_t kCursorBytes = 256;
constexpr size_t kRaysOff = kCursorBytes;
constexpr size_t kHitsOff = kRaysOff + (size_t)kMax * 32;
constexpr size_t kIdsOff = kHitsOff + (size_t)kMax * 16;
constexpr size_t kTOff = kIdsOff + (size_t)kMax * 4;
constexpr size_t kBaryOff = kTOff + (size_t)kMax * 4;
constexpr size_t kScratchBytes = kBaryOff + (size_t)kMax * 8;

// In HOST code this name is a host-side stand-in, not a device address: a kernel handed `gNoAllocScratch` faults (or,
// where the GPU can reach host memory, silently uses the wrong memory).  Host code gets the device address from
// cudaGetSymbolAddress, once per device, with that device current (scratchOf below).
static __device__ __align__(256) unsigned char gNoAllocScratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex gScratchMutex;
unsigned char *gScratchPtr[kMaxDevices] = {nullptr};

// the device address of this device's copy of the array; ctx->device is current
int scratchOf(pbrtgpu_ctx *ctx, unsigned char **out) {
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return ctx->fail(PBRTGPU_ERR_INVALID, "device index %d not supported by the no-alloc entry point", ctx->device);
    std::lock_guard<std::mutex> lock(gScratchMutex);
    if (!gScratchPtr[ctx->device]) {
        void *p = nullptr;
        PG_CUDA(ctx, cudaGetSymbolAddress(&p, gNoAllocScratch));
        gScratchPtr[ctx->device] = (unsigned char *)p;
    }
    *out = gScratchPtr[ctx->device];
    return PBRTGPU_OK;
}

__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_pintersect_closest_static(DevScene sc, const float4 *__restrict__ rays, int n,
                                                                                             float4 *__restrict__ hits, int *next,
                                                                                             int fetchThreshold, int phaseBias) {
    tracePersistent<kTraceClosest, false>(sc.bvh, SingleQueue{rays, 0}, n, next, fetchThreshold, phaseBias, nullptr,
                                          [&](int i, int, const Ray &ray, const Hit &h, bool found) {
                                              float4 out;
                                              out.x = __int_as_float(found ? __ldg(sc.primOrigId + h.prim) : -1);
                                              out.y = ray.tMax;
                                              out.z = found ? h.b0 : 0.f;
                                              out.w = found ? h.b1 : 0.f;
                                              hits[i] = out;
                                          });
}

// {origId, t, b0, b1} records -> the three columns the C ABI returns
__global__ void k_split_hit_columns(const float4 *__restrict__ hits, int n, int *__restrict__ ids, float *__restrict__ t, float2 *__restrict__ b0b1) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 h = hits[i];
        ids[i] = __float_as_int(h.x);
        t[i] = h.y;
        b0b1[i] = make_float2(h.z, h.w);
    }
}

}  // namespace

extern "C" int pbrtgpu_intersect_closest_noalloc(pbrtgpu_ctx *ctx, const float *rays, int64_t n, int32_t *prim_orig_id, float *t, float *b0b1) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!ctx->haveScene) return ctx->fail(PBRTGPU_ERR_STATE, "intersect before upload_scene");
    if (n < 0 || (n > 0 && (!rays || !prim_orig_id))) return ctx->fail(PBRTGPU_ERR_INVALID, "null ray/output buffer");
    if (n > kMax)
        return ctx->fail(PBRTGPU_ERR_INVALID, "pbrtgpu_intersect_closest_noalloc: %lld rays exceed PBRTGPU_NOALLOC_MAX_RAYS = %lld (its scratch is a "
                         "fixed device array; use pbrtgpu_intersect_closest, or split the batch)", (long long)n, kMax);
    if (n == 0) return PBRTGPU_OK;
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    if (int rc = scratchOf(ctx, &base)) return rc;
    int *cursor = (int *)base;
    float4 *dRays = (float4 *)(base + kRaysOff), *dHits = (float4 *)(base + kHitsOff);
    int *dIds = (int *)(base + kIdsOff);
    float *dT = (float *)(base + kTOff);
    float2 *dBary = (float2 *)(base + kBaryOff);
    const int m = (int)n;
    PG_CUDA(ctx, cudaMemsetAsync(cursor, 0, kCursorBytes, ctx->stream));  // the one part that must start at zero
    PG_CUDA(ctx, cudaMemcpyAsync(dRays, rays, (size_t)m * 32, cudaMemcpyHostToDevice, ctx->stream));
    const int blocks = (int)std::min<long long>(((long long)m + kTraceBlock - 1) / kTraceBlock, (long long)ctx->smCount * ctx->persistBlocksPerSM);
    k_pintersect_closest_static<<<blocks, kTraceBlock, 0, ctx->stream>>>(ctx->scene, dRays, m, dHits, cursor, ctx->fetchThreshold, ctx->phaseBias);
    k_split_hit_columns<<<std::min(ctx->smCount * 4, (m + 255) / 256), 256, 0, ctx->stream>>>(dHits, m, dIds, dT, dBary);
    ctx->launches += 2;
    PG_CUDA(ctx, cudaGetLastError());
    PG_CUDA(ctx, cudaMemcpyAsync(prim_orig_id, dIds, (size_t)m * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (t) PG_CUDA(ctx, cudaMemcpyAsync(t, dT, (size_t)m * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (b0b1) PG_CUDA(ctx, cudaMemcpyAsync(b0b1, dBary, (size_t)m * 8, cudaMemcpyDeviceToHost, ctx->stream));
    PG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PBRTGPU_OK;
}
