// Persistent-warp BVH traversal with ray replacement (CUDA only).
//
// Same visiting order and arithmetic as traverse<> in bvh.h (and therefore as
// BVHAccel::Intersect / IntersectP, accelerators/bvh.cpp:662-738); only the
// mapping of rays to lanes differs:
//   * a warp keeps running until the ray queue is drained; a lane whose ray
//     has finished picks up the next queued ray (queue positions are claimed 32
//     at a time, one chunk ahead, their rays prefetched) instead of idling until
//     the slowest ray of its warp is done;
//   * every round of the loop the warp does ONE kind of work, chosen by a vote:
//     either each lane that is at an interior record tests its two children,
//     or each lane that is at a primitive tests it; box tests and triangle
//     tests never serialise against each other inside a warp, and a lane waits
//     only until its kind of work has the majority;
//   * the triangle shear (triangle.cpp:205-222) is computed once per ray.
// The result of a ray does not depend on which lane traced it or when, so the
// output is bit-identical to the one-thread-per-ray loop.
#ifndef PG_DEVICE_TRACE_PERSISTENT_H
#define PG_DEVICE_TRACE_PERSISTENT_H

#include "bvh.h"

#if defined(__CUDACC__)
namespace pg {

__device__ __forceinline__ Ray loadRayF4(const F4 *__restrict__ rays, int i) {
    F4 a = __ldg(rays + 2 * (size_t)i), b = __ldg(rays + 2 * (size_t)i + 1);
    Ray r;
    r.o = v3(a.x, a.y, a.z);
    r.tMax = a.w;
    r.d = v3(b.x, b.y, b.z);
    return r;
}

// slabPartial (bvh.h) without the early returns: the same comparisons and the same
// conditional updates, combined at the end (values computed past a failed test are unused).
// The three dirIsNeg flags come as values (the traversal derives them from invDir's signs each round: three compares
// instead of three bytes of per-ray state that the 64-register budget keeps in local memory).
__device__ __forceinline__ bool slabPartialFlat(V3 bmin, V3 bmax, V3 o, V3 invDir, bool negX, bool negY, bool negZ, float *tMinOut) {
    const float k = 1 + 2 * gammaf(3);
    float tMin = ((negX ? bmax.x : bmin.x) - o.x) * invDir.x;
    float tMax = ((negX ? bmin.x : bmax.x) - o.x) * invDir.x;
    float tyMin = ((negY ? bmax.y : bmin.y) - o.y) * invDir.y;
    float tyMax = ((negY ? bmin.y : bmax.y) - o.y) * invDir.y;
    float tzMin = ((negZ ? bmax.z : bmin.z) - o.z) * invDir.z;
    float tzMax = ((negZ ? bmin.z : bmax.z) - o.z) * invDir.z;
    tMax *= k;
    tyMax *= k;
    tzMax *= k;
    bool missXY = tMin > tyMax || tyMin > tMax;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    bool missZ = tMin > tzMax || tzMin > tMax;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    *tMinOut = tMin;
    return !missXY && !missZ && tMax > 0;
}

// sink(idx, kind, ray, hit, found): called once per ray by the lane that finished it.
// `next` is a device counter that is 0 at launch.  fetchThreshold: refill when
// at least this many lanes of the warp are idle (1..32).  phaseBias (in 1/16):
// a round runs box tests when 16 * (#lanes at an interior record) >=
// phaseBias * (#lanes at a primitive), else one primitive test per such lane.
//
// MODE: kTraceClosest / kTraceAny for a queue of one kind, kTraceMixed when the queue holds both kinds.
// source(idx, &ray) loads ray `idx` and returns its kind (0 = closest hit, 1 = any hit; in mixed mode
// any other small integer >= 2 is a closest-hit ray of another sub-queue); the kind indexes cnt[] and
// is handed back to sink(idx, kind, ray, hit, found).
enum { kTraceClosest = 0, kTraceAny = 1, kTraceMixed = 2 };

struct SingleQueue {  // one ray array, one kind
    const F4 *rays;
    int kind;
    __device__ __forceinline__ int operator()(int idx, Ray *ray) const {
        *ray = loadRayF4(rays, idx);
        return kind;
    }
    __device__ __forceinline__ void prefetch(int idx) const { asm volatile("prefetch.global.L1 [%0];" ::"l"(rays + 2 * (size_t)idx)); }
};

// Where a round's interior record / primitive slot comes from.  GlobalFetch: the scene arrays through the read-only path
// (L1 / L2).  SharedFetch: a copy of BOTH arrays that the block staged into shared memory at kernel start (stageSceneShared);
// only scenes whose records and slots fit kSharedSceneBytes qualify (the Cornell box: 20 records + 32 slots = 2.8 KB) --
// every fetch of a round is then one LDS.128 and the L1 is left to the ray queues.
constexpr int kSharedSceneBytes = 16 * 1024;  // per block; 8 resident blocks per SM
struct GlobalFetch {
    const F4 *nodes, *tris;
    __device__ __forceinline__ F4 node(size_t i) const { return __ldg(nodes + i); }
    __device__ __forceinline__ F4 tri(size_t i) const { return __ldg(tris + i); }
};
struct SharedFetch {
    int triBase;  // in F4: the primitive slots follow the records
    __device__ __forceinline__ F4 node(size_t i) const {
        extern __shared__ __align__(16) F4 pgSceneSmem[];
        return pgSceneSmem[i];
    }
    __device__ __forceinline__ F4 tri(size_t i) const {
        extern __shared__ __align__(16) F4 pgSceneSmem[];
        return pgSceneSmem[triBase + i];
    }
};
// every thread of the block calls this before tracePersistent
__device__ __forceinline__ SharedFetch stageSceneShared(const DevBVH &bvh, int nRecords, int nSlots) {
    extern __shared__ __align__(16) F4 pgSceneSmem[];
    const int nN = 4 * nRecords, nT = 3 * nSlots;
    for (int i = threadIdx.x; i < nN; i += blockDim.x) pgSceneSmem[i] = __ldg(bvh.nodes + i);
    for (int i = threadIdx.x; i < nT; i += blockDim.x) pgSceneSmem[nN + i] = __ldg(bvh.tris + i);
    __syncthreads();
    return SharedFetch{nN};
}

template <int MODE, bool COUNT, class Fetch, class Source, class Sink>
__device__ __forceinline__ void tracePersistentFrom(const DevBVH &bvh, Fetch fetch, Source source, int n, int *next, int fetchThreshold,
                                                    int phaseBias, TraverseCounters *cnt, Sink sink) {
    const unsigned kFull = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int2 stack[64];  // {ref, bits of the entry distance}: one 8-byte local access per push / pop
    // ref >= 0: interior record to visit; ref < 0: ~ref is the next primitive slot to test; kEmptyRef: no ray
    int sp = 0, ref = kEmptyRef, idx = -1, kind = MODE == kTraceAny ? 1 : 0;
    Ray ray;
    ray.o = ray.d = v3(0.f);
    ray.tMax = 0;
    V3 invDir = v3(0.f);
    int topRef = 0;    // the top stack entry lives in registers: a pop does not wait for local memory
    float topT = 0;
    RayShear shear = {0, 1, 2, 0.f, 0.f, 0.f};
    Hit hit;
    hit.prim = -1;
    hit.t = hit.b0 = hit.b1 = hit.b2 = 0;
    bool exhausted = false;  // warp-uniform: the queue has been handed out completely
    // The first 32 rays of every warp are assigned statically (queue positions [32 * warp, 32 * warp + 32));
    // the shared cursor `next` hands out what lies beyond all warps' first chunks, 32 positions per atomicAdd.
    // A launch over a short queue therefore costs no atomics at all.
    const int warpsInGrid = (gridDim.x * blockDim.x) >> 5;
    const int warpId = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int dynamicBegin = warpsInGrid * 32;
    // Queue positions are claimed one chunk AHEAD of their use: [aheadBase, aheadBase + aheadLeft) belongs to this warp and
    // its rays are on their way into L1 (one prefetch per lane at claim time), so the ray loads of a refill -- 5 % of the
    // kernel's stall samples in the r02l source profile, with every lane of the warp waiting -- find them there, and most
    // refills need no atomic.  +1 % on killeroo, Cornell and the portal room (profiles/r02m_ab_trace_variants.log).
    int aheadBase = warpId * 32, aheadLeft = 32;
    for (;;) {
        unsigned idle = __ballot_sync(kFull, ref == kEmptyRef);
        if (idle) {
            int nIdle = __popc(idle);
            if (!exhausted && nIdle >= fetchThreshold) {
                const int rank = __popc(idle & ((1u << lane) - 1));
                const int k = min(nIdle, aheadLeft);
                int mine = aheadBase + rank;  // lanes of rank < k take the rest of the claimed chunk
                if (nIdle >= aheadLeft) {     // ... which is used up: claim the next one, lanes of rank >= k start it
                    int nb = n;
                    if (dynamicBegin < n) {
                        if (lane == 0) nb = dynamicBegin + atomicAdd(next, 32);
                        nb = __shfl_sync(kFull, nb, 0);
                    }
                    if (rank >= k) mine = nb + (rank - k);
                    aheadBase = nb + (nIdle - k);
                    aheadLeft = 32 - (nIdle - k);
                    const int p = aheadBase + lane;
                    if (lane < aheadLeft && p < n) source.prefetch(p);
                } else {
                    aheadBase += k;
                    aheadLeft -= k;
                }
                if (ref == kEmptyRef && mine < n) {
                    idx = mine;
                    kind = source(idx, &ray);
                    invDir = v3(1 / ray.d.x, 1 / ray.d.y, 1 / ray.d.z);
                    shear = rayShear(ray.d);
                    hit.prim = -1;
                    sp = 0;
                    ref = 0;
                }
                exhausted = aheadBase >= n;  // nothing claimed is left to start, and nothing is left to claim
            }
            if (exhausted && __all_sync(kFull, ref == kEmptyRef)) return;
        }
        const bool any = MODE == kTraceMixed ? kind == 1 : MODE == kTraceAny;
        const bool atInterior = ref >= 0, atPrim = ref < 0 && ref != kEmptyRef;
        int nI = __popc(__ballot_sync(kFull, atInterior)), nP = __popc(__ballot_sync(kFull, atPrim));
        bool popNext = false, occluded = false;
        if (nI > 0 && 16 * nI >= phaseBias * nP) {
            // ---- one interior record: both children's bounds
            if (atInterior) {
                const size_t nd = 4 * (size_t)ref;
                F4 q0 = fetch.node(nd), q1 = fetch.node(nd + 1), q2 = fetch.node(nd + 2), q3 = fetch.node(nd + 3);
                if (COUNT) cnt[MODE == kTraceMixed ? kind : 0].nodes += 2;
                int ref0 = f2i(q3.x), ref1 = f2i(q3.y), axis = f2i(q3.z);
                // dirIsNeg (bvh.cpp:667-668) from the signs of invDir; both slab tests run unconditionally (an empty child --
                // only the second child of record 0 -- is masked afterwards), so t0 / t1 stay in registers
                const bool negX = invDir.x < 0, negY = invDir.y < 0, negZ = invDir.z < 0;
                float t0 = 0, t1 = 0;
                bool ok0 = slabPartialFlat(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), ray.o, invDir, negX, negY, negZ, &t0);
                bool ok1 = slabPartialFlat(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), ray.o, invDir, negX, negY, negZ, &t1);
                ok0 = ok0 & (ref0 != kEmptyRef);
                ok1 = ok1 & (ref1 != kEmptyRef);
                bool nearIs1 = axis == 0 ? negX : (axis == 1 ? negY : negZ);
                int refNear = nearIs1 ? ref1 : ref0, refFar = nearIs1 ? ref0 : ref1;
                bool okNear = nearIs1 ? ok1 : ok0, okFar = nearIs1 ? ok0 : ok1;
                float tNear = nearIs1 ? t1 : t0, tFar = nearIs1 ? t0 : t1;
                if (okFar) {
                    if (sp > 0) stack[sp - 1] = make_int2(topRef, __float_as_int(topT));
                    topRef = refFar;
                    topT = tFar;
                    ++sp;
                }
                if (okNear && tNear < ray.tMax) ref = refNear;
                else popNext = true;
            }
        } else if (atPrim) {
            // ---- one primitive of the leaf
            int slot = ~ref;
            const size_t tp = 3 * (size_t)slot;
            F4 a = fetch.tri(tp);
            int tag = f2i(a.w);
            int type = tag & kTagTypeMask;
            if (COUNT) cnt[MODE == kTraceMixed ? kind : 0].prims += 1;
            float t, b0 = 0, b1 = 0, b2 = 0;
            bool h;
            if (type == 0) {
                F4 b = fetch.tri(tp + 1), c = fetch.tri(tp + 2);
                h = triangleTestSheared(ray, shear, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &b0, &b1, &b2);
                if (!any && (tag & kTagDegenerate)) h = false;
            } else if (type == 1) {
                h = aaplaneTest(ray, bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift), &t, nullptr);
            } else {
                h = sphereTestGuarded(ray, bvh.sphereParams + 48 * (size_t)(tag >> kTagShapeShift), &t);
            }
            if (h) {
                if (any) occluded = true;
                else {
                    ray.tMax = t;
                    hit.prim = slot;
                    hit.t = t;
                    hit.b0 = b0;
                    hit.b1 = b1;
                    hit.b2 = b2;
                }
            }
            if (occluded) ref = kEmptyRef;
            else if (tag & kTagLast) popNext = true;
            else ref = ~(slot + 1);
        }
        if (popNext) {  // the next stacked node whose entry distance is still inside the ray
            ref = kEmptyRef;
            while (sp > 0) {
                --sp;
                int r = topRef;
                float t = topT;
                if (sp > 0) {
                    int2 e = stack[sp - 1];
                    topRef = e.x;
                    topT = __int_as_float(e.y);
                }
                if (t < ray.tMax) {
                    ref = r;
                    break;
                }
            }
        }
        if ((atInterior || atPrim) && ref == kEmptyRef) sink(idx, kind, ray, hit, any ? occluded : hit.prim >= 0);
    }
}

template <int MODE, bool COUNT, class Source, class Sink>
__device__ __forceinline__ void tracePersistent(const DevBVH &bvh, Source source, int n, int *next, int fetchThreshold,
                                                int phaseBias, TraverseCounters *cnt, Sink sink) {
    tracePersistentFrom<MODE, COUNT>(bvh, GlobalFetch{bvh.nodes, bvh.tris}, source, n, next, fetchThreshold, phaseBias, cnt, sink);
}

}  // namespace pg
#endif  // __CUDACC__
#endif
