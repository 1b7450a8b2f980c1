// The wavefront path integrator: SamplerIntegrator::Render + PathIntegrator::Li
// (core/integrator.cpp:526-637, integrators/path.cpp:64-189) as a sequence of
// kernels over queues of path vertices.  Per batch of paths:
//
//   k_raygen                       camera samples -> closest-hit queue 0
//   for depth = 0 .. maxDepth:
//     k_trace_closest              BVHAccel::Intersect for the queue
//     k_shade                      emission, BSDF, next-event ray generation, BSDF sampling,
//                                  Russian roulette; survivors are compacted into the other queue
//     k_trace_shadow               BVHAccel::IntersectP for the light samples, resolved in place
//     k_trace_nee                  BVHAccel::Intersect for BSDF-sampled / portal rays, resolved in place
//   k_film                         scrub + FilmTile::AddSample (float atomics)
//
// Queue lengths live in device memory (WaveCounters): no kernel's launch
// geometry depends on them, so a batch runs without a single host sync.
// The stage bodies are in integrator.h; this file only maps threads to entries.
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "../common/halton_tables.h"
#include "ctx.h"
#include "integrator.h"
#include "kernels.h"
#include "trace_persistent.h"

using namespace pg;

namespace pg {

constexpr int kStageBlock = 128;

// The kernels that call out-of-line device functions (Halton digits, hit records, BSDFs, lights) hand them
// w.sc / w.hal BY REFERENCE.  A by-value __global__ parameter whose address is taken is copied by every
// thread to its own stack first (ptxas: ~1.4 KB of "stack frame" = sizeof(WaveParams) + the real locals), and
// the callees then read scene pointers through local memory: 1024 threads x 0.8 KB per SM, far more than
// the L1 holds.  Those kernels therefore read ONE copy of the parameters from global memory (WaveParams *,
// written per batch by k_set_wave, stream-ordered); every lane loads the same address, which is one sector.
__global__ void k_set_wave(WaveParams w, WaveParams *dst) {
    const int *src = reinterpret_cast<const int *>(&w);
    int *d = reinterpret_cast<int *>(dst);
    for (int i = threadIdx.x; i < (int)(sizeof(WaveParams) / sizeof(int)); i += blockDim.x) d[i] = src[i];
}
static_assert(kHaltonPrimes == kPrimeTableSize, "sampler.h and halton_tables.h disagree on the prime table size");
static_assert(sizeof(WaveParams) % sizeof(int) == 0, "WaveParams is copied word by word");

__global__ void __launch_bounds__(kStageBlock) k_raygen(const WaveParams *__restrict__ wp) {
    const WaveParams &w = *wp;
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w.nSlots; i += stride) stageRaygen(w, i);
}

// COUNT = profile mode: also accumulate node / primitive tests into w.countTraversal[2*which..]
template <bool COUNT>
__device__ __forceinline__ void flushCounts(const WaveParams &w, int which, const TraverseCounters &c) {
    if (COUNT) {
        atomicAdd(w.countTraversal + 2 * which, (unsigned long long)c.nodes);
        atomicAdd(w.countTraversal + 2 * which + 1, (unsigned long long)c.prims);
    }
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock) k_trace_closest(WaveParams w, int depth, int cur) {
    int n = w.cnt->nActive[depth];
    int stride = gridDim.x * blockDim.x;
    const F4 *rays = w.rays[cur];
    TraverseCounters c = {0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Ray ray = loadRay(rays, i);
        Hit h;
        traverse<false, COUNT>(w.sc.bvh, ray, &h, &c);
        if (h.prim >= 0) {
            w.hits[i] = f4(__int_as_float(h.prim), h.b0, h.b1, h.b2);
            w.hitIdx[atomicIncInt(&w.cnt->nHit[depth])] = i;
        } else
            w.missIdx[atomicIncInt(&w.cnt->nMiss[depth])] = i;
    }
    flushCounts<COUNT>(w, 0, c);
}

// Persistent variants (trace_persistent.h): same results, rays handed to lanes dynamically.
// They only record what was hit; queue compaction and the resolve arithmetic run as
// separate dense kernels (few lanes finish a ray in the same round).
// SMEM: the scene's records and primitive slots are staged into shared memory by every block (trace_persistent.h:
// stageSceneShared; scenes of at most kSharedSceneBytes), nRecords = interior records of the device layout.
template <bool COUNT, bool SMEM = false>
__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_ptrace_closest(WaveParams w, int depth, int cur, int nRecords) {
    TraverseCounters c = {0, 0};
    auto sink = [&](int i, int, const Ray &, const Hit &h, bool) { w.hits[i] = f4(__int_as_float(h.prim), h.b0, h.b1, h.b2); };
    if (SMEM)
        tracePersistentFrom<kTraceClosest, COUNT>(w.sc.bvh, stageSceneShared(w.sc.bvh, nRecords, w.sc.nPrims), SingleQueue{w.rays[cur], 0},
                                                  w.cnt->nActive[depth], &w.cnt->fetch[0][depth], w.fetchThreshold, w.phaseBias, &c, sink);
    else
        tracePersistent<kTraceClosest, COUNT>(w.sc.bvh, SingleQueue{w.rays[cur], 0}, w.cnt->nActive[depth], &w.cnt->fetch[0][depth],
                                              w.fetchThreshold, w.phaseBias, &c, sink);
    flushCounts<COUNT>(w, 0, c);
}

// Everything shade(depth) asked for, in ONE launch: the path rays of depth + 1 (closest hit), the
// BSDF-sampled / portal next-event rays of this depth (closest hit) and its light-sample rays (any
// hit) form one queue, heaviest kind first.  One launch has one tail (the last rays in flight keep
// few lanes busy) instead of three, and the short any-hit rays fill the lanes the long ones vacate.
struct FusedQueue {
    const F4 *path, *nee, *shadow;
    int nPath, nNee;
    __device__ __forceinline__ int operator()(int idx, Ray *ray) const {
        // one load site: the sub-queue only selects the address (three load sites made the compiler keep the ray in local memory)
        const F4 *q = path;
        int kind = 0;
        if (idx >= nPath) {
            idx -= nPath;
            q = nee, kind = 2;
            if (idx >= nNee) idx -= nNee, q = shadow, kind = 1;
        }
        *ray = loadRayF4(q, idx);
        return kind;
    }
    __device__ __forceinline__ void prefetch(int idx) const {
        const F4 *q = path;
        if (idx >= nPath) {
            idx -= nPath;
            q = nee;
            if (idx >= nNee) idx -= nNee, q = shadow;
        }
        asm volatile("prefetch.global.L1 [%0];" ::"l"(q + 2 * (size_t)idx));
    }
};

// Lengths of the light-sample queues as their consumers see them: a push past the capacity is counted
// (WaveCounters::dropped) and not stored, so the counter can exceed what the queue holds.
__device__ __forceinline__ int neeQueueLen(const WaveParams &w, int depth) { return min(w.cnt->nNee[depth], w.neeCapacity); }
__device__ __forceinline__ int shadowQueueLen(const WaveParams &w, int depth) { return min(w.cnt->nShadow[depth], w.neeCapacity); }

template <bool COUNT, bool SMEM = false>
__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_ptrace_fused(WaveParams w, int depth, int cur, int nRecords) {
    TraverseCounters c[3] = {{0, 0}, {0, 0}, {0, 0}};
    const int nPath = w.cnt->nActive[depth + 1], nNee = neeQueueLen(w, depth), nShadow = shadowQueueLen(w, depth);
    FusedQueue q = {w.rays[cur ^ 1], w.neeRays, w.shadowRays, nPath, nNee};
    auto sink = [&](int i, int kind, const Ray &, const Hit &h, bool found) {
        if (kind == 0) w.hits[i] = f4(__int_as_float(h.prim), h.b0, h.b1, h.b2);
        else if (kind == 2) w.neeHits[i - nPath] = f4(__int_as_float(h.prim), h.b0, h.b1, h.b2);
        else w.shadowOcc[i - nPath - nNee] = found ? 1 : 0;
    };
    if (SMEM)
        tracePersistentFrom<kTraceMixed, COUNT>(w.sc.bvh, stageSceneShared(w.sc.bvh, nRecords, w.sc.nPrims), q, nPath + nNee + nShadow,
                                                &w.cnt->fetch[1][depth], w.fetchThreshold, w.phaseBias, c, sink);
    else
        tracePersistent<kTraceMixed, COUNT>(w.sc.bvh, q, nPath + nNee + nShadow, &w.cnt->fetch[1][depth], w.fetchThreshold, w.phaseBias, c, sink);
    flushCounts<COUNT>(w, 0, c[0]);
    flushCounts<COUNT>(w, 1, c[1]);
    flushCounts<COUNT>(w, 2, c[2]);
}

// hit / miss positions of the depth-d queue, in queue order (so that k_shade reads its inputs nearly in order)
// Hits on plain Lambertian materials fill hitIdx from the front, hits on everything else (plastic, uber, glass,
// metal: microfacet or specular lobes, several lobes of which one is picked per sample) from the back: the
// shade kernel walks the two runs one after the other, so all but one of its warps shade a single class
// instead of every warp paying for the longest branch of a mixed one.  A third run holds the hits on materials all of whose
// lobes are specular (smooth glass, mirror): they never sample a light, so their instantiation of the shade kernel has four
// phases instead of nine, and they no longer halve the lanes of the warps that shade rough glass or plastic.  That run fills
// missIdx from its end downwards (hits + misses = queue length <= nSlots, so the two ends never meet).
__device__ __forceinline__ void splitHits(const WaveParams &w, int depth, bool wantMiss, int i) {
    int prim = __float_as_int(w.hits[i].x);
    if (prim >= 0) {
        int cls = 0;
        if (w.sortHits && w.sc.primMaterial) {
            int m = __ldg(w.sc.primMaterial + prim);
            if (m >= 0) cls = __ldg(&w.sc.matLobes[m].shadeClass);
            if (cls == 2 && w.sortHits < 2) cls = 1;  // sortHits 1: two runs only (A/B)
        }
        if (cls == 1) w.hitIdx[w.nSlots - 1 - atomicIncInt(&w.cnt->nHitB[depth])] = i;
        else if (cls == 2) w.missIdx[w.nSlots - 1 - atomicIncInt(&w.cnt->nHitC[depth])] = i;
        else w.hitIdx[atomicIncInt(&w.cnt->nHit[depth])] = i;
    } else if (wantMiss)
        w.missIdx[atomicIncInt(&w.cnt->nMiss[depth])] = i;
}
__device__ __forceinline__ int hitQueueLen(const WaveParams &w, int depth) { return w.cnt->nHit[depth] + w.cnt->nHitB[depth] + w.cnt->nHitC[depth]; }
__device__ __forceinline__ int hitQueueAt(const WaveParams &w, int depth, int i) {
    int nA = w.cnt->nHit[depth], nB = w.cnt->nHitB[depth];
    if (i < nA) return w.hitIdx[i];
    if (i < nA + nB) return w.hitIdx[w.nSlots - 1 - (i - nA)];
    return w.missIdx[w.nSlots - 1 - (i - nA - nB)];
}
__global__ void __launch_bounds__(kStageBlock) k_split_hits(WaveParams w, int depth, bool wantMiss) {
    int n = w.cnt->nActive[depth];
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) splitHits(w, depth, wantMiss, i);
}

// after k_ptrace_fused(depth): the light samples of this depth count where unoccluded, and the
// next depth's queue is split into hits and misses
__global__ void __launch_bounds__(kStageBlock) k_post_trace(WaveParams w, int depth, bool wantMiss) {
    int nS = shadowQueueLen(w, depth), nA = w.cnt->nActive[depth + 1];
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nS; i += stride) stageShadowResolve(w, i, w.shadowOcc[i] != 0);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nA; i += stride) splitHits(w, depth + 1, wantMiss, i);
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_ptrace_shadow(WaveParams w, int depth) {
    TraverseCounters c = {0, 0};
    tracePersistent<kTraceAny, COUNT>(w.sc.bvh, SingleQueue{w.shadowRays, 1}, shadowQueueLen(w, depth), &w.cnt->fetch[1][depth], w.fetchThreshold,
                                      w.phaseBias, &c, [&](int i, int, const Ray &, const Hit &, bool occluded) { w.shadowOcc[i] = occluded ? 1 : 0; });
    flushCounts<COUNT>(w, 1, c);
}

__global__ void __launch_bounds__(kStageBlock) k_shadow_resolve(WaveParams w, int depth) {
    int n = shadowQueueLen(w, depth);
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) stageShadowResolve(w, i, w.shadowOcc[i] != 0);
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock, kPersistMinBlocks) k_ptrace_nee(WaveParams w, int depth) {
    TraverseCounters c = {0, 0};
    tracePersistent<kTraceClosest, COUNT>(w.sc.bvh, SingleQueue{w.neeRays, 0}, neeQueueLen(w, depth), &w.cnt->fetch[2][depth], w.fetchThreshold,
                                          w.phaseBias, &c, [&](int i, int, const Ray &, const Hit &h, bool) {
                                              w.neeHits[i] = f4(__int_as_float(h.prim), h.b0, h.b1, h.b2);
                                          });
    flushCounts<COUNT>(w, 2, c);
}

// the emitter lookup / MIS weight of the rays k_ptrace_nee traced
__global__ void __launch_bounds__(kStageBlock) k_nee_resolve(const WaveParams *__restrict__ wp, int depth) {
    const WaveParams &w = *wp;
    int n = neeQueueLen(w, depth);
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) stageNeeResolve(w, i);
}

__global__ void __launch_bounds__(kStageBlock) k_shade(const WaveParams *__restrict__ wp, int depth, int cur) {
    const WaveParams &w = *wp;
    int n = hitQueueLen(w, depth);
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) stageShade(w, depth, cur, hitQueueAt(w, depth, i));
}

// One 384-thread block per SM, all of whose warps walk the phases of stageShadePhased together
// (__syncthreads between phases): the shading code is far larger than the instruction cache.
#ifndef PG_SYNC_MASK
#define PG_SYNC_MASK 0xff
#endif
struct BlockSync {
    // i = the phase that just ended; PG_SYNC_MASK (A/B builds) drops individual barriers
    __device__ __forceinline__ void operator()(int i) const {
        if ((PG_SYNC_MASK >> i) & 1) __syncthreads();
    }
};
// EXT: with the non-reference portal strategies (direct.h).  LAMBERT / PORTAL: see stageShadePhased.
// run: which part of the hit queue -- 0 = all of it, 1 = the front run (hits on single-Lambert-lobe materials and on
// material-less primitives, render.cu:splitHits), 2 = the back run (every other material but the all-specular ones),
// 3 = the all-specular run (SPEC).
template <int BLOCK, int PER_SM, bool EXT = false, bool LAMBERT = false, bool PORTAL = true, bool SPEC = false>
__global__ void __launch_bounds__(BLOCK, PER_SM) k_shade_sync(const WaveParams *__restrict__ wp, int depth, int cur, int run) {
    const WaveParams &w = *wp;
    const int nA = w.cnt->nHit[depth], nB = w.cnt->nHitB[depth], nC = w.cnt->nHitC[depth];
    const int begin = run == 2 ? nA : (run == 3 ? nA + nB : 0), end = run == 1 ? nA : (run == 2 ? nA + nB : nA + nB + nC);
    int stride = gridDim.x * blockDim.x;
#ifdef PG_SHADE_LOCAL_SCRATCH
    ShadeScratch scr;
#else
    extern __shared__ __align__(16) unsigned char pgShadeSmem[];  // BLOCK * sizeof(ShadeScratch), see shadeSmemBytes
    ShadeScratch &scr = reinterpret_cast<ShadeScratch *>(pgShadeSmem)[threadIdx.x];
#endif
    for (int base = begin + blockIdx.x * blockDim.x; base < end; base += stride) {
        int i = base + threadIdx.x;
        bool active = i < end;
        stageShadePhased<BlockSync, EXT, LAMBERT, PORTAL, SPEC>(w, depth, cur, active ? hitQueueAt(w, depth, i) : 0, active, BlockSync(), scr);
    }
}

// Integrator "directlighting": one vertex per thread, the block re-synchronised at every vertex
__global__ void __launch_bounds__(512, 1) k_shade_direct(const WaveParams *__restrict__ wp, int depth, int cur) {
    const WaveParams &w = *wp;
    int n = hitQueueLen(w, depth);
    int stride = gridDim.x * blockDim.x;
    for (int base = blockIdx.x * blockDim.x; base < n; base += stride) {
        int i = base + threadIdx.x;
        __syncthreads();
        if (i < n) stageShadeDirect(w, depth, cur, hitQueueAt(w, depth, i));
    }
}

// escaped rays: only launched when the scene has an infinite light or paths keep pending branches (directlighting tree mode)
__global__ void __launch_bounds__(kStageBlock) k_miss(WaveParams w, int depth, int cur) {
    int n = w.cnt->nMiss[depth];
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) stageMiss(w, depth, cur, w.missIdx[i]);
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock) k_trace_shadow(WaveParams w, int depth) {
    int n = shadowQueueLen(w, depth);
    int stride = gridDim.x * blockDim.x;
    TraverseCounters c = {0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Ray ray = loadRay(w.shadowRays, i);
        bool occ = traverse<true, COUNT>(w.sc.bvh, ray, nullptr, &c);
        stageShadowResolve(w, i, occ);
    }
    flushCounts<COUNT>(w, 1, c);
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceBlock) k_trace_nee(WaveParams w, int depth) {
    int n = neeQueueLen(w, depth);
    int stride = gridDim.x * blockDim.x;
    TraverseCounters c = {0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        Ray ray = loadRay(w.neeRays, i);
        Hit h;
        traverse<false, COUNT>(w.sc.bvh, ray, &h, &c);
        w.neeHits[i] = f4(__int_as_float(h.prim), h.b0, h.b1, h.b2);
        stageNeeResolve(w, i);
    }
    flushCounts<COUNT>(w, 2, c);
}

__global__ void __launch_bounds__(kStageBlock) k_film(WaveParams w) {
    int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < w.nSlots; i += stride) stageFilm(w, i);
}

// fold one batch's counters into the running totals, then clear them for the next batch.
// carry != 0 (directlighting tree mode, paths still hold pending branches after `depths` iterations): the
// queue of depth `depths` becomes the queue of depth 0 of another round of iterations over the same batch.
__global__ void k_batch_stats(WaveCounters *cnt, unsigned long long *totals, int depths, int carry, int firstRound) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long closest = 0, shadow = 0, nee = 0, culled = 0;
    for (int d = 0; d < depths + (carry ? 0 : 1); ++d) closest += (unsigned long long)cnt->nActive[d];
    for (int d = 0; d < depths; ++d)
        nee += (unsigned long long)cnt->nNee[d], shadow += (unsigned long long)cnt->nShadow[d], culled += (unsigned long long)cnt->nNeeCulled[d];
    totals[5] += culled;
    if (firstRound) totals[0] += (unsigned long long)cnt->nActive[0];
    totals[1] += closest + nee;
    totals[4] += nee;
    totals[2] += shadow;
    totals[3] += cnt->scrubbed;
    totals[6] += cnt->dropped;
    const int a = cnt->nActive[depths], h = depths < kMaxDepthSlots ? cnt->nHit[depths] : 0, m = depths < kMaxDepthSlots ? cnt->nMiss[depths] : 0;
    const int hb = depths < kMaxDepthSlots ? cnt->nHitB[depths] : 0, hc = depths < kMaxDepthSlots ? cnt->nHitC[depths] : 0;
    for (int d = 0; d <= kMaxDepthSlots; ++d) cnt->nActive[d] = 0;
    for (int d = 0; d < kMaxDepthSlots; ++d)
        cnt->nShadow[d] = cnt->nNee[d] = cnt->nHit[d] = cnt->nHitB[d] = cnt->nHitC[d] = cnt->nMiss[d] = cnt->fetch[0][d] = cnt->fetch[1][d] = cnt->fetch[2][d] = cnt->nNeeCulled[d] = 0;
    cnt->scrubbed = 0;
    cnt->dropped = 0;
    if (carry) cnt->nActive[0] = a, cnt->nHit[0] = h, cnt->nHitB[0] = hb, cnt->nHitC[0] = hc, cnt->nMiss[0] = m;
}

// Film::MergeFilmTile's RGB -> XYZ (film.cpp:125-127), once per pixel at the end
__global__ void k_film_to_xyz(float *rgbw, long long nPixels) {
    long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nPixels; i += stride) {
        float4 v = reinterpret_cast<float4 *>(rgbw)[i];
        float rgb[3] = {v.x, v.y, v.z}, xyz[3];
        rgbToXYZ(rgb, xyz);
        reinterpret_cast<float4 *>(rgbw)[i] = make_float4(xyz[0], xyz[1], xyz[2], v.w);
    }
}

// HaltonSampler::SampleDimension on a batch, for unit parity against the oracle
__global__ void k_halton(HaltonParams h, const int *pixelXY, const long long *sampleNum, const int *dim, long long n, float *out) {
    long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        long long index = haltonPixelOffset(h, pixelXY[2 * i], pixelXY[2 * i + 1]) + sampleNum[i] * h.sampleStride;
        out[i] = haltonSample(h, index, dim[i]);
    }
}

}  // namespace pg

#include "render_host.inc"

extern "C" int pbrtgpu_halton_samples(pbrtgpu_ctx *ctx, const pbrtgpu_render_params *params, const int32_t *pixel_xy,
                                      const int64_t *sample_num, const int32_t *dim, int64_t n, float *out) {
    if (!ctx || !params || (n > 0 && (!pixel_xy || !sample_num || !dim || !out))) return PBRTGPU_ERR_INVALID;
    if (n <= 0) return PBRTGPU_OK;
    for (int64_t i = 0; i < n; ++i)
        if (dim[i] < 0 || dim[i] >= kPrimeTableSize) return ctx->fail(PBRTGPU_ERR_INVALID, "HaltonSampler can only sample %d dimensions.", kPrimeTableSize);
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    HaltonParams h = makeHaltonParams(params->sample_bounds, params->sample_at_pixel_center);
    h.primes = ctx->dPrimes, h.primeSums = ctx->dPrimeSums, h.perms = ctx->dPerms, h.primeMagic = ctx->dPrimeMagic, h.dims = ctx->dHaltonDims;
    int rc;
    size_t bytes = (size_t)n * (8 + 8 + 4 + 4);
    if ((rc = ensureBuffer(ctx, ctx->scratch[2], bytes)) != PBRTGPU_OK) return rc;
    char *base = (char *)ctx->scratch[2].p;
    long long *dNum = (long long *)base;
    int *dXY = (int *)(base + 8 * (size_t)n);
    int *dDim = (int *)(base + 16 * (size_t)n);
    float *dOut = (float *)(base + 20 * (size_t)n);
    PG_CUDA(ctx, cudaMemcpyAsync(dNum, sample_num, 8 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    PG_CUDA(ctx, cudaMemcpyAsync(dXY, pixel_xy, 8 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    PG_CUDA(ctx, cudaMemcpyAsync(dDim, dim, 4 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    k_halton<<<stageGrid(ctx, n, 128, 8), 128, 0, ctx->stream>>>(h, dXY, dNum, dDim, n, dOut);
    ctx->launches++;
    PG_CUDA(ctx, cudaGetLastError());
    PG_CUDA(ctx, cudaMemcpyAsync(out, dOut, 4 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    PG_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return PBRTGPU_OK;
}
