// Host-side setup shared by the CUDA driver (render.cu) and tests/hostsim:
// pbrtgpu_render_params -> the value part of WaveParams, the batch plan, and
// the per-slot buffer table.  Plain C++.
#ifndef PG_DEVICE_WAVE_SETUP_H
#define PG_DEVICE_WAVE_SETUP_H

#include <cstring>

#include "../../../include/pbrtgpu.h"
#include "integrator.h"

namespace pg {

inline void fillWaveParams(const pbrtgpu_render_params *rp, WaveParams *w) {
    memcpy(w->cam.cameraToWorld.m, rp->camera_to_world, 64);
    memcpy(w->cam.rasterToCamera.m, rp->raster_to_camera, 64);
    w->cam.lensRadius = rp->lens_radius;
    w->cam.focalDistance = rp->focal_distance;
    for (int i = 0; i < 4; ++i) {
        w->film.cropped[i] = rp->cropped_bounds[i];
        w->film.sampleBounds[i] = rp->sample_bounds[i];
        w->film.pixelBounds[i] = rp->pixel_bounds[i];
    }
    for (int i = 0; i < 2; ++i) {
        w->film.radius[i] = rp->filter_radius[i];
        w->film.invRadius[i] = 1 / rp->filter_radius[i];  // FilmTile: invFilterRadius(1 / filterRadius.x, ...)
    }
    w->film.maxSampleLuminance = rp->max_sample_luminance;
    w->hal = makeHaltonParams(rp->sample_bounds, rp->sample_at_pixel_center);
    w->maxDepth = rp->max_depth;
    w->rrThreshold = rp->rr_threshold;
    w->integrator = rp->integrator;
    w->directStrategy = rp->direct_strategy;
    w->directTree = 0;
    w->branchCapacity = 0;
    int sw = rp->sample_bounds[2] - rp->sample_bounds[0];
    w->tilesX = (sw + 15) / 16;
    w->shardIndex = rp->shard_index;
    w->shardCount = rp->shard_count > 0 ? rp->shard_count : 1;
}

// What both drivers of the stage functions (render_host.inc on the GPU, tests/hostsim on the CPU) refuse.
// The bounce count of a path is an 8-bit field of its flags word; the reference has no limit, its shipped
// scenes go up to maxdepth 100 (scenes/blender/lamp/out/lamp.pbrt:9).
constexpr int kMaxMaxDepth = 255;
constexpr int kMaxIterationsPerBatch = 4096;  // open-ended batches stop here; what is left is counted in pbrtgpu_stats.truncated_paths
inline const char *validateRenderParams(const pbrtgpu_render_params *rp) {
    if (rp->spp <= 0) return "spp must be positive";
    if (rp->max_depth < 0 || rp->max_depth > kMaxMaxDepth) return "maxdepth out of range (0 .. 255)";
    if (rp->shard_count > 1 && (rp->shard_index < 0 || rp->shard_index >= rp->shard_count)) return "shard_index out of range";
    const int *cb = rp->cropped_bounds;
    if (cb[2] <= cb[0] || cb[3] <= cb[1]) return "empty film";
    if (rp->integrator != PBRTGPU_INTEGRATOR_PATH && rp->integrator != PBRTGPU_INTEGRATOR_DIRECT) return "unknown integrator";
    return nullptr;
}

struct BatchPlan {
    int tilesX, tilesY, shardTiles;  // shardTiles = tiles t with t % shardCount == shardIndex
    int sppBegin, sppEnd;
    int sppChunk;        // samples of a pixel per batch
    int tilesPerBatch;
    int maxSlots;
    int depthIters;      // wavefront iterations a batch is known to need (a path makes at most maxDepth + 1 scene intersections)
    int roundIters;      // iterations per round: WaveCounters has one slot per iteration of a round; between rounds the counters are
                         // folded and the live queue becomes iteration 0 of the next round (k_batch_stats, carry)
    int openEnded;       // the iteration count is not known in advance (directlighting tree walk; vertices on material-less
                         // primitives, which do not count as bounces): rounds continue until the path queue is empty
    int branchPerSlot;   // directlighting tree mode: pending-branch entries per path slot (0 otherwise)
    int neeMult;         // light samples per vertex: the shadow / next-event queues hold maxSlots * neeMult entries
};

// Light samples a vertex can ask for: 1 for the path integrator and directlighting "one", the sum of the
// lights' nSamples for directlighting "all".  Sizes the shadow / next-event queues.
inline int lightSamplesPerVertex(const pbrtgpu_render_params *rp, const pbrtgpu_scene *s) {
    if (rp->integrator != PBRTGPU_INTEGRATOR_DIRECT || rp->direct_strategy != PBRTGPU_DIRECT_ALL) return 1;
    long long sum = 0;
    for (int i = 0; i < s->n_lights; ++i) sum += s->light_nsamples ? s->light_nsamples[i] : 1;
    return (int)(sum < 1 ? 1 : sum);
}
// DirectLightingIntegrator::Preprocess (directlighting.cpp:45-61): two 2D arrays per light per depth
inline int sampleArrayCount(const pbrtgpu_render_params *rp, int nLights) {
    if (rp->integrator != PBRTGPU_INTEGRATOR_DIRECT || rp->direct_strategy != PBRTGPU_DIRECT_ALL) return 0;
    return 2 * nLights * (rp->max_depth > 0 ? rp->max_depth : 0);
}

inline BatchPlan planBatches(const pbrtgpu_render_params *rp, long long targetSlots, bool hasNullMaterial, int neeMult = 1,
                             int branchPerSlot = 0) {
    BatchPlan p;
    targetSlots /= neeMult;
    targetSlots = targetSlots * 290 / (290 + 48LL * branchPerSlot);  // keep the batch's memory roughly constant
    p.branchPerSlot = branchPerSlot;
    int sw = rp->sample_bounds[2] - rp->sample_bounds[0], sh = rp->sample_bounds[3] - rp->sample_bounds[1];
    p.tilesX = (sw + 15) / 16;
    p.tilesY = (sh + 15) / 16;
    int total = p.tilesX * p.tilesY;
    int sc = rp->shard_count > 0 ? rp->shard_count : 1;
    p.shardTiles = total > rp->shard_index ? (total - rp->shard_index + sc - 1) / sc : 0;
    p.sppBegin = rp->spp_begin;
    p.sppEnd = rp->spp_end > rp->spp_begin ? rp->spp_end : rp->spp;
    int spp = p.sppEnd - p.sppBegin;
    p.sppChunk = spp < 32 ? (spp > 0 ? spp : 1) : 32;
    bool wholeShard = false;  // every batch takes all of the shard's tiles
    // Few tiles (one shard of many, a small image): take more sample numbers of every pixel into a batch, so that it
    // still fills targetSlots -- half-size batches pay the tail of every traversal launch twice (8 GPUs: 4 batches
    // of 8 M slots per step instead of 2 of 16 M cost 2 ms of 48).  The sample numbers are split evenly.
    if (p.shardTiles > 0 && spp > 32) {
        long long want = targetSlots / (256LL * p.shardTiles);
        if (want > 32) {
            long long most = want + want / 4;  // a batch may exceed the target by a quarter rather than leave a small remainder
            int nChunks = (int)((spp + most - 1) / most);
            p.sppChunk = (spp + nChunks - 1) / nChunks;
            wholeShard = true;
        }
    }
    long long t = targetSlots / (256LL * p.sppChunk);
    if (t < 1) t = 1;
    if (t > p.shardTiles || wholeShard) t = p.shardTiles > 0 ? p.shardTiles : 1;
    p.tilesPerBatch = (int)t;
    p.maxSlots = p.tilesPerBatch * 256 * p.sppChunk;
    // a path makes at most maxDepth + 1 scene intersections; vertices on
    // material-less primitives do not count as bounces (path.cpp:108-113), so
    // leave room for a few of those when the scene has any
    p.neeMult = neeMult;
    p.depthIters = rp->max_depth + 1 + (hasNullMaterial ? 8 : 0);
    if (rp->integrator == PBRTGPU_INTEGRATOR_DIRECT) p.depthIters = (rp->max_depth > 1 ? rp->max_depth : 1) + (hasNullMaterial ? 8 : 0);
    // tree mode: a path walks up to 2^maxdepth - 1 vertices, one per wavefront iteration.  The iterations run in
    // rounds of this many; after each round the host looks whether any path is left (render_host.inc, anotherRound)
    if (branchPerSlot > 0) p.depthIters = 2 * rp->max_depth + 2 + (hasNullMaterial ? 8 : 0);
    p.roundIters = p.depthIters < kMaxDepthSlots - 1 ? p.depthIters : kMaxDepthSlots - 1;  // stage d touches the counters of depth d + 1
    p.openEnded = (branchPerSlot > 0 || hasNullMaterial) ? 1 : 0;
    return p;
}

// Bytes per slot of each wave buffer, in the order render.cu / hostsim bind them.
enum WaveBuf {
    WB_HINDEX, WB_DIM, WB_L, WB_BETA, WB_FLAGS, WB_PFILM, WB_RAYS0, WB_RAYS1, WB_Q0, WB_Q1, WB_HITS, WB_SHADOW_RAYS,
    WB_SHADOW_CONTRIB, WB_NEE_RAYS, WB_NEE_CONTRIB, WB_NEE_INFO, WB_NEE_HITS, WB_SHADOW_OCC, WB_HIT_IDX, WB_MISS_IDX, WB_BRANCH, WB_COUNT
};
inline size_t waveBufBytesPerSlot(int b) {
    static const size_t sz[WB_COUNT] = {8, 4, 16, 16, 4, 8, 32, 32, 4, 4, 16, 32, 16, 32, 16, 8, 16, 1, 4, 4, 48};
    return sz[b];
}
inline bool waveBufIsLightSampleQueue(int b) {
    return b == WB_SHADOW_RAYS || b == WB_SHADOW_CONTRIB || b == WB_NEE_RAYS || b == WB_NEE_CONTRIB || b == WB_NEE_INFO || b == WB_NEE_HITS ||
           b == WB_SHADOW_OCC;
}
inline size_t waveBufBytes(int b, const BatchPlan &p) {
    if (b == WB_BRANCH) return waveBufBytesPerSlot(b) * (size_t)p.maxSlots * (size_t)p.branchPerSlot + 48;
    return waveBufBytesPerSlot(b) * (size_t)p.maxSlots * (waveBufIsLightSampleQueue(b) ? (size_t)p.neeMult : 1);
}
inline void bindWaveBuffers(WaveParams *w, void *const bufs[WB_COUNT]) {
    w->hIndex = (long long *)bufs[WB_HINDEX];
    w->dim = (int *)bufs[WB_DIM];
    w->L = (F4 *)bufs[WB_L];
    w->beta = (F4 *)bufs[WB_BETA];
    w->flags = (int *)bufs[WB_FLAGS];
    w->pFilm = (float *)bufs[WB_PFILM];
    w->rays[0] = (F4 *)bufs[WB_RAYS0];
    w->rays[1] = (F4 *)bufs[WB_RAYS1];
    w->qSlot[0] = (int *)bufs[WB_Q0];
    w->qSlot[1] = (int *)bufs[WB_Q1];
    w->hits = (F4 *)bufs[WB_HITS];
    w->shadowRays = (F4 *)bufs[WB_SHADOW_RAYS];
    w->shadowContrib = (F4 *)bufs[WB_SHADOW_CONTRIB];
    w->neeRays = (F4 *)bufs[WB_NEE_RAYS];
    w->neeContrib = (F4 *)bufs[WB_NEE_CONTRIB];
    w->neeInfo = (int *)bufs[WB_NEE_INFO];
    w->neeHits = (F4 *)bufs[WB_NEE_HITS];
    w->shadowOcc = (unsigned char *)bufs[WB_SHADOW_OCC];
    w->hitIdx = (int *)bufs[WB_HIT_IDX];
    w->missIdx = (int *)bufs[WB_MISS_IDX];
    w->branch = (F4 *)bufs[WB_BRANCH];
}

}  // namespace pg
#endif
