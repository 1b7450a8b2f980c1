// PathIntegrator::Li as a wavefront state machine: one function per stage,
// each called for one queue entry by a thin CUDA kernel (render.cu) -- or, in
// tests/hostsim, by a plain CPU loop for debugging.
//   SamplerIntegrator::Render (per-sample part)   core/integrator.cpp:561-623
//   PathIntegrator::Li                            integrators/path.cpp:64-189
//   SurfaceInteraction::Le                        core/interaction.cpp:148-151
//   FilmTile::AddSample / Film::MergeFilmTile     core/film.h:121-161, film.cpp:117-130
#ifndef PG_DEVICE_INTEGRATOR_H
#define PG_DEVICE_INTEGRATOR_H

#include "direct.h"

namespace pg {

// Device-side counters of one batch.  One slot per depth so that nothing has
// to be reset (or read back by the host) between the stages of a batch.
constexpr int kMaxDepthSlots = 64;
struct WaveCounters {
    int nActive[kMaxDepthSlots + 1];  // closest-hit queue length entering depth d
    int nShadow[kMaxDepthSlots];      // any-hit NEE rays generated at depth d
    int nNee[kMaxDepthSlots];         // closest-hit NEE rays generated at depth d
    int nHit[kMaxDepthSlots];         // of the nActive[d] rays: how many hit something (queue hitIdx) ...
    int nMiss[kMaxDepthSlots];        // ... and how many escaped (queue missIdx)
    int nHitB[kMaxDepthSlots];        // hits on multi-lobe materials, kept apart: they fill hitIdx from its END downwards (render.cu:splitHits)
    int nHitC[kMaxDepthSlots];        // hits on all-specular materials (MatLobes::shadeClass 2): they fill missIdx from its END downwards
    int fetch[3][kMaxDepthSlots];     // persistent trace kernels: next unclaimed queue position {closest, shadow, nee}
    int nNeeCulled[kMaxDepthSlots];   // MIS rays not traced because they miss the sampled light's shape (bvh.h:primitiveAccepts)
    unsigned long long cameraRays, closestRays, shadowRays, scrubbed;
    unsigned long long dropped;       // light-sample rays that found their queue full (pushDirectRays); must stay 0
};

struct FilmParams {
    int cropped[4], sampleBounds[4], pixelBounds[4];
    float radius[2], invRadius[2];
    const float *filterTable;  // 16 x 16
    float maxSampleLuminance;
    float *rgbw;  // cropped W*H*4: sum of L*w*filter (RGB, as FilmTilePixel::contribSum), sum of filter weights
};

struct WaveParams {
    DevScene sc;
    HaltonParams hal;
    CameraParams cam;
    FilmParams film;
    int maxDepth;
    float rrThreshold;
    int integrator, directStrategy;  // PBRTGPU_INTEGRATOR_*, PBRTGPU_DIRECT_*
    int nSampleArrays;               // 2D sample arrays the integrator requested (DirectLightingIntegrator::Preprocess)
    int neeCapacity;                 // entries of the shadow / next-event queues (nSlots x light samples per vertex)
    int directTree;                  // directlighting over materials with specular transmission: Li's recursion is a tree
    int branchCapacity;              // pending branches a path can hold (= maxdepth)
    F4 *branch;                      // 3 x float4 per entry, branchCapacity entries per slot: {o, valid} {d, depth} {beta, -}
    // batch geometry: slot -> (pixel, sample number)
    int tilesX;            // sample-bounds tiles per row (16 x 16 tiles as integrator.cpp:533-553)
    int shardIndex, shardCount;
    int tileBegin;         // first shard-local tile of this batch
    int sppBegin, sppCount;  // sample numbers [sppBegin, sppBegin + sppCount) of every pixel in the batch
    int nSlots;
    // per-slot path state
    long long *hIndex;     // Halton sample index
    int *dim;              // next sampler dimension
    F4 *L;                 // rgb = radiance so far, w = hero wavelength (nm)
    F4 *beta;              // rgb = throughput, w = etaScale
    int *flags;            // bits 0-7 bounces, bit 8 specularBounce, bit 9 alive (has a film sample),
                           // bits 10-25 vertices shaded so far (directlighting: which sample arrays are next),
                           // bits 26-30 pending branches (directlighting over specular transmission)
    float *pFilm;          // 2 per slot
    // queues (position-indexed)
    F4 *rays[2];           // closest-hit rays of the current / next depth, 2 x float4 each
    int *qSlot[2];
    F4 *hits;              // {prim slot (int), b0, b1, b2}
    int *hitIdx, *missIdx; // positions of the current queue that hit / escaped, compacted by the trace kernel
    F4 *shadowRays;        // 2 x float4 each
    F4 *shadowContrib;     // rgb, slot (int)
    F4 *neeRays;
    F4 *neeContrib;        // rgb, slot (int)
    int *neeInfo;          // 2 per entry: light, mode
    F4 *neeHits;
    unsigned char *shadowOcc;
    WaveCounters *cnt;
    int skipWavelength;    // no dispersive material: the camera sample's wavelength dimension is stepped over, not computed
    int sortHits;          // split the hit queue by material class so that a warp shades one class: 1 = Lambert | rest, 2 = ... | all-specular
    int phaseBias;         // persistent trace kernels: box-test rounds vs primitive-test rounds, see trace_persistent.h
    int fetchThreshold;    // persistent trace kernels refill a warp when this many lanes are idle
    unsigned long long *countTraversal;  // profile mode: {closest, shadow, nee} x {node tests, prim tests}; else null
};

#if defined(__CUDA_ARCH__)
PG_HD int atomicIncInt(int *p) {  // warp-aggregated: one atomic per warp
    unsigned mask = __activemask();
    int leader = __ffs(mask) - 1;
    int lane = threadIdx.x & 31;
    int base = 0;
    if (lane == leader) base = atomicAdd(p, __popc(mask));
    base = __shfl_sync(mask, base, leader);
    return base + __popc(mask & ((1u << lane) - 1));
}
PG_HD void atomicAddFloat(float *p, float v) { atomicAdd(p, v); }
// one 16-byte reduction (red.global.add.v4.f32, sm_90+) instead of four scalar ones; p is 16-byte aligned
PG_HD void atomicAddFloat4(float *p, float x, float y, float z, float w) { atomicAdd(reinterpret_cast<float4 *>(p), make_float4(x, y, z, w)); }
PG_HD void atomicAddU64(unsigned long long *p, unsigned long long v) { atomicAdd(p, v); }
#else
PG_HD int atomicIncInt(int *p) { return __atomic_fetch_add(p, 1, __ATOMIC_RELAXED); }
PG_HD void atomicAddFloat(float *p, float v) {
    unsigned int *ip = (unsigned int *)p;
    unsigned int old = __atomic_load_n(ip, __ATOMIC_RELAXED), want;
    do {
        want = floatToBits(bitsToFloat(old) + v);
    } while (!__atomic_compare_exchange_n(ip, &old, want, false, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}
PG_HD void atomicAddFloat4(float *p, float x, float y, float z, float w) {
    atomicAddFloat(p, x), atomicAddFloat(p + 1, y), atomicAddFloat(p + 2, z), atomicAddFloat(p + 3, w);
}
PG_HD void atomicAddU64(unsigned long long *p, unsigned long long v) { __atomic_fetch_add(p, v, __ATOMIC_RELAXED); }
#endif

PG_HD F4 f4(float x, float y, float z, float w) {
    F4 r;
    r.x = x, r.y = y, r.z = z, r.w = w;
    return r;
}
PG_HD void storeRay(F4 *q, int pos, const Ray &r) {
    q[2 * (size_t)pos] = f4(r.o.x, r.o.y, r.o.z, r.tMax);
    q[2 * (size_t)pos + 1] = f4(r.d.x, r.d.y, r.d.z, 0.f);
}
PG_HD Ray loadRay(const F4 *q, int pos) {
    F4 a = q[2 * (size_t)pos], b = q[2 * (size_t)pos + 1];
    Ray r;
    r.o = v3(a.x, a.y, a.z);
    r.tMax = a.w;
    r.d = v3(b.x, b.y, b.z);
    return r;
}

// slot -> pixel and sample number.  Slots run tile by tile (16 x 16 pixels,
// this shard's tiles only), pixel by pixel inside a tile, sample by sample
// inside a pixel: a warp works on neighbouring samples of one or two pixels.
PG_HD bool slotToSample(const WaveParams &w, int slot, int *px, int *py, int *sampleNum) {
    int s = slot % w.sppCount;
    int pix = slot / w.sppCount;
    int tileLocal = w.tileBegin + pix / 256, inTile = pix % 256;
    int tile = tileLocal * w.shardCount + w.shardIndex;
    int tx = tile % w.tilesX, ty = tile / w.tilesX;
    *px = w.film.sampleBounds[0] + tx * 16 + (inTile & 15);
    *py = w.film.sampleBounds[1] + ty * 16 + (inTile >> 4);
    *sampleNum = w.sppBegin + s;
    if (*px >= w.film.sampleBounds[2] || *py >= w.film.sampleBounds[3]) return false;
    // integrator.cpp:571: pixels outside pixelBounds are skipped
    return *px >= w.film.pixelBounds[0] && *px < w.film.pixelBounds[2] && *py >= w.film.pixelBounds[1] && *py < w.film.pixelBounds[3];
}

// ---- stage 1: camera sample + camera ray
PG_HD void stageRaygen(const WaveParams &w, int slot) {
    int px, py, sampleNum;
    if (!slotToSample(w, slot, &px, &py, &sampleNum)) {
        w.flags[slot] = 0;
        return;
    }
    PathSampler smp;
    smp.index = haltonPixelOffset(w.hal, px, py) + (long long)sampleNum * w.hal.sampleStride;
    smp.dim = 0;
    CameraSample cs = getCameraSample(w.hal, smp, px, py, w.cam.lensRadius > 0, !w.skipWavelength);
    float wvl0;
    Ray ray = generateCameraRay(w.cam, cs, &wvl0);
    w.hIndex[slot] = smp.index;
    w.dim[slot] = smp.dim;
    w.L[slot] = f4(0, 0, 0, wvl0);
    w.beta[slot] = f4(1, 1, 1, 1);
    w.flags[slot] = 1 << 9;
    w.pFilm[2 * (size_t)slot] = cs.pFilmX;
    w.pFilm[2 * (size_t)slot + 1] = cs.pFilmY;
    int pos = atomicIncInt(&w.cnt->nActive[0]);
    storeRay(w.rays[0], pos, ray);
    w.qSlot[0][pos] = slot;
}

// Interaction::n and emission of a hit, for SurfaceInteraction::Le(w)
PG_NI_INTERACTION SurfInt hitInteraction(const DevScene &sc, int prim, float b0, float b1, float b2, const Ray &ray) {
    const F4 *tp = sc.bvh.tris + 3 * (size_t)prim;
    int tag = f2i(PG_LDG(tp).w);
    int type = tag & kTagTypeMask;
    int flags = sc.primFlags ? PG_LD(sc.primFlags + prim) : 0;
    Ray r = ray;
    r.tMax = kInfinity;
    if (type == PBRTGPU_PRIM_TRIANGLE)
        return triangleInteraction(tp, flags, sc.triN ? sc.triN + 9 * (size_t)prim : nullptr,
                                   sc.triUV ? sc.triUV + 6 * (size_t)prim : nullptr,
                                   sc.triS ? sc.triS + 9 * (size_t)prim : nullptr, b0, b1, b2, ray.d);
    if (type == PBRTGPU_PRIM_AAPLANE) return aaplaneInteraction(r, sc.bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift), flags);
    return sphereInteraction(r, sc.bvh.sphereParams + 48 * (size_t)(tag >> kTagShapeShift), flags);
}
PG_HD Spec hitLe(const DevScene &sc, int prim, V3 n, V3 w) {
    int li = sc.primLight ? PG_LD(sc.primLight + prim) : -1;
    return li >= 0 ? areaLightL(sc, li, n, w) : spec(0.f);
}

}  // namespace pg
#include "integrator_stages.h"
#endif
