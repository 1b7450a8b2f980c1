// Ray-batch entry points: BVHAccel::Intersect / IntersectP over device ray
// buffers (config C5 and the tier-1 parity tests), plus the trace kernels the
// wavefront renderer reuses.
#include <algorithm>

#include "ctx.h"
#include "kernels.h"
#include "trace_persistent.h"

using namespace pg;

namespace pg {

// rays: 2 x float4 per ray {ox,oy,oz,tMax} {dx,dy,dz,-}; hits: float4 {origId(int), t, b0, b1}
__global__ void __launch_bounds__(kTraceBlock) k_intersect_closest(DevScene sc, const float4 *__restrict__ rays,
                                                                   long long n, float4 *__restrict__ hits) {
    long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
        Ray ray;
        ray.o = v3(a.x, a.y, a.z);
        ray.tMax = a.w;
        ray.d = v3(b.x, b.y, b.z);
        Hit h;
        traverse<false>(sc.bvh, ray, &h);
        float4 out;
        out.x = __int_as_float(h.prim >= 0 ? __ldg(sc.primOrigId + h.prim) : -1);
        out.y = ray.tMax;
        out.z = h.prim >= 0 ? h.b0 : 0.f;
        out.w = h.prim >= 0 ? h.b1 : 0.f;
        hits[i] = out;
    }
}

__global__ void __launch_bounds__(kTraceBlock) k_intersect_any(DevScene sc, const float4 *__restrict__ rays, long long n,
                                                               unsigned char *__restrict__ occluded) {
    long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float4 a = __ldg(rays + 2 * i), b = __ldg(rays + 2 * i + 1);
        Ray ray;
        ray.o = v3(a.x, a.y, a.z);
        ray.tMax = a.w;
        ray.d = v3(b.x, b.y, b.z);
        occluded[i] = traverse<true>(sc.bvh, ray, nullptr) ? 1 : 0;
    }
}

// Persistent variants (trace_persistent.h); `rays` / outputs already offset to the chunk, n < 2^31.
// COUNT (profile mode, pbrtgpu_set_profile): box tests and primitive tests of the batch are added to counts[0], counts[1] --
// what tests/test_rays.py compares with the tallies of the reference's own loop (the same nodes entered, the same primitives
// tested: "same visiting order" checked, not argued).
template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_pintersect_closest(DevScene sc, const float4 *__restrict__ rays, int n,
                                                                    float4 *__restrict__ hits, int *next, int fetchThreshold, int phaseBias,
                                                                    unsigned long long *counts) {
    TraverseCounters c = {0, 0};
    tracePersistent<kTraceClosest, COUNT>(sc.bvh, SingleQueue{rays, 0}, n, next, fetchThreshold, phaseBias, &c, [&](int i, int, const Ray &ray, const Hit &h, bool found) {
        float4 out;
        out.x = __int_as_float(found ? __ldg(sc.primOrigId + h.prim) : -1);
        out.y = ray.tMax;
        out.z = found ? h.b0 : 0.f;
        out.w = found ? h.b1 : 0.f;
        hits[i] = out;
    });
    if (COUNT) {
        atomicAdd(counts, (unsigned long long)c.nodes);
        atomicAdd(counts + 1, (unsigned long long)c.prims);
    }
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_pintersect_any(DevScene sc, const float4 *__restrict__ rays, int n,
                                                                unsigned char *__restrict__ occluded, int *next, int fetchThreshold, int phaseBias,
                                                                unsigned long long *counts) {
    TraverseCounters c = {0, 0};
    tracePersistent<kTraceAny, COUNT>(sc.bvh, SingleQueue{rays, 1}, n, next, fetchThreshold, phaseBias, &c,
                                      [&](int i, int, const Ray &, const Hit &, bool occ) { occluded[i] = occ ? 1 : 0; });
    if (COUNT) {
        atomicAdd(counts, (unsigned long long)c.nodes);
        atomicAdd(counts + 1, (unsigned long long)c.prims);
    }
}

int traceGrid(const pbrtgpu_ctx *ctx, long long n) {
    long long blocks = (n + kTraceBlock - 1) / kTraceBlock;
    long long cap = (long long)ctx->smCount * kTraceBlocksPerSM;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace pg

static int timedLaunch(pbrtgpu_ctx *ctx, bool any, const void *dRays, long long n, void *dOut, float *ms) {
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms) {
        PG_CUDA(ctx, cudaEventCreate(&e0));
        PG_CUDA(ctx, cudaEventCreate(&e1));
        PG_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    }
    if (ctx->tracePersistent) {
        const long long kChunk = 1ll << 30;
        for (long long off = 0; off < n; off += kChunk) {
            int m = (int)std::min(kChunk, n - off);
            long long blocks = std::min<long long>(((long long)m + kTraceBlock - 1) / kTraceBlock, (long long)ctx->smCount * ctx->persistBlocksPerSM);
            PG_CUDA(ctx, cudaMemsetAsync(ctx->dFetch, 0, sizeof(int), ctx->stream));
            const bool count = ctx->profile && ctx->dTraversalCounts;  // {closest, shadow, nee} x {node tests, prim tests}
            if (any && count)
                k_pintersect_any<true><<<(int)blocks, kTraceBlock, 0, ctx->stream>>>(ctx->scene, (const float4 *)dRays + 2 * off, m, (unsigned char *)dOut + off,
                                                                                     ctx->dFetch, ctx->fetchThreshold, ctx->phaseBias, ctx->dTraversalCounts + 2);
            else if (any)
                k_pintersect_any<false><<<(int)blocks, kTraceBlock, 0, ctx->stream>>>(ctx->scene, (const float4 *)dRays + 2 * off, m, (unsigned char *)dOut + off,
                                                                                      ctx->dFetch, ctx->fetchThreshold, ctx->phaseBias, nullptr);
            else if (count)
                k_pintersect_closest<true><<<(int)blocks, kTraceBlock, 0, ctx->stream>>>(ctx->scene, (const float4 *)dRays + 2 * off, m, (float4 *)dOut + off,
                                                                                         ctx->dFetch, ctx->fetchThreshold, ctx->phaseBias, ctx->dTraversalCounts);
            else
                k_pintersect_closest<false><<<(int)blocks, kTraceBlock, 0, ctx->stream>>>(ctx->scene, (const float4 *)dRays + 2 * off, m, (float4 *)dOut + off,
                                                                                          ctx->dFetch, ctx->fetchThreshold, ctx->phaseBias, nullptr);
        }
    } else {
        int grid = traceGrid(ctx, n);
        if (any) k_intersect_any<<<grid, kTraceBlock, 0, ctx->stream>>>(ctx->scene, (const float4 *)dRays, n, (unsigned char *)dOut);
        else k_intersect_closest<<<grid, kTraceBlock, 0, ctx->stream>>>(ctx->scene, (const float4 *)dRays, n, (float4 *)dOut);
    }
    ctx->launches++;
    PG_CUDA(ctx, cudaGetLastError());
    if (ms) {
        PG_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
        PG_CUDA(ctx, cudaEventSynchronize(e1));
        PG_CUDA(ctx, cudaEventElapsedTime(ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
    }
    return PBRTGPU_OK;
}

extern "C" {

int pbrtgpu_intersect_closest_device(pbrtgpu_ctx *ctx, const void *d_rays, int64_t n, void *d_hits, float *elapsed_ms) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!ctx->haveScene) return ctx->fail(PBRTGPU_ERR_STATE, "intersect before upload_scene");
    if (n < 0 || (n > 0 && (!d_rays || !d_hits))) return ctx->fail(PBRTGPU_ERR_INVALID, "null ray/hit buffer");
    if (n == 0) { if (elapsed_ms) *elapsed_ms = 0; return PBRTGPU_OK; }
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    return timedLaunch(ctx, false, d_rays, n, d_hits, elapsed_ms);
}

int pbrtgpu_intersect_any_device(pbrtgpu_ctx *ctx, const void *d_rays, int64_t n, void *d_occluded, float *elapsed_ms) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!ctx->haveScene) return ctx->fail(PBRTGPU_ERR_STATE, "intersect before upload_scene");
    if (n < 0 || (n > 0 && (!d_rays || !d_occluded))) return ctx->fail(PBRTGPU_ERR_INVALID, "null ray/output buffer");
    if (n == 0) { if (elapsed_ms) *elapsed_ms = 0; return PBRTGPU_OK; }
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    return timedLaunch(ctx, true, d_rays, n, d_occluded, elapsed_ms);
}

int pbrtgpu_intersect_closest(pbrtgpu_ctx *ctx, const float *rays, int64_t n, int32_t *prim_orig_id, float *t, float *b0b1) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!ctx->haveScene) return ctx->fail(PBRTGPU_ERR_STATE, "intersect before upload_scene");
    if (n < 0 || (n > 0 && (!rays || !prim_orig_id))) return ctx->fail(PBRTGPU_ERR_INVALID, "null ray/output buffer");
    if (n == 0) return PBRTGPU_OK;
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc;
    if ((rc = ensureBuffer(ctx, ctx->scratch[0], (size_t)n * 32)) != PBRTGPU_OK) return rc;
    if ((rc = ensureBuffer(ctx, ctx->scratch[1], (size_t)n * 16)) != PBRTGPU_OK) return rc;
    PG_CUDA(ctx, cudaMemcpyAsync(ctx->scratch[0].p, rays, (size_t)n * 32, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = timedLaunch(ctx, false, ctx->scratch[0].p, n, ctx->scratch[1].p, nullptr)) != PBRTGPU_OK) return rc;
    std::vector<float> hits((size_t)n * 4);
    PG_CUDA(ctx, cudaMemcpyAsync(hits.data(), ctx->scratch[1].p, (size_t)n * 16, cudaMemcpyDeviceToHost, ctx->stream));
    PG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int64_t i = 0; i < n; ++i) {
        memcpy(&prim_orig_id[i], &hits[4 * i], 4);
        if (t) t[i] = hits[4 * i + 1];
        if (b0b1) b0b1[2 * i] = hits[4 * i + 2], b0b1[2 * i + 1] = hits[4 * i + 3];
    }
    return PBRTGPU_OK;
}

int pbrtgpu_intersect_any(pbrtgpu_ctx *ctx, const float *rays, int64_t n, uint8_t *occluded) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!ctx->haveScene) return ctx->fail(PBRTGPU_ERR_STATE, "intersect before upload_scene");
    if (n < 0 || (n > 0 && (!rays || !occluded))) return ctx->fail(PBRTGPU_ERR_INVALID, "null ray/output buffer");
    if (n == 0) return PBRTGPU_OK;
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc;
    if ((rc = ensureBuffer(ctx, ctx->scratch[0], (size_t)n * 32)) != PBRTGPU_OK) return rc;
    if ((rc = ensureBuffer(ctx, ctx->scratch[1], (size_t)n)) != PBRTGPU_OK) return rc;
    PG_CUDA(ctx, cudaMemcpyAsync(ctx->scratch[0].p, rays, (size_t)n * 32, cudaMemcpyHostToDevice, ctx->stream));
    if ((rc = timedLaunch(ctx, true, ctx->scratch[0].p, n, ctx->scratch[1].p, nullptr)) != PBRTGPU_OK) return rc;
    PG_CUDA(ctx, cudaMemcpyAsync(occluded, ctx->scratch[1].p, (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    PG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PBRTGPU_OK;
}

}  // extern "C"
