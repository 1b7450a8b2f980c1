// TEST INFRASTRUCTURE ONLY.
//
// There is no GPU in the development container, so the `-m "not gpu"` suite
// compiles the device headers (csrc/device/*.h, all PG_HD = __host__ __device__
// functions) with g++ and steps them on the CPU, one "thread" at a time, to
// catch logic errors before GPU time is spent.  This is a debugging aid for
// the kernels' device functions; it is built only into tests/_build/ and is
// never loaded by the product library, bench.py or the CLI (those fail loudly
// without a CUDA device).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

// -DPG_PERTURB_LIBM: the same code with a libm whose sinf / cosf / acosf / atan2f are off by one ulp in half of the
// calls -- the difference CUDA's libm has against glibc's.  tests/test_render.py uses this build to measure how
// much of an image depends on the last bit of those functions (a 1-ulp change of a sampled direction can flip a
// shadow ray that grazes the limb of a small, distant sphere light), i.e. what tolerance a GPU image can be held to.
#ifdef PG_PERTURB_LIBM
static inline float pgNudge(float v) {
    uint32_t u;
    memcpy(&u, &v, 4);
    if (!std::isfinite(v) || v == 0.f || !(u & 2u)) return v;
    return std::nextafterf(v, (u & 4u) ? INFINITY : -INFINITY);
}
#define sinf(x) pgNudge(sinf(x))
#define cosf(x) pgNudge(cosf(x))
#define acosf(x) pgNudge(acosf(x))
#define atan2f(y, x) pgNudge(atan2f(y, x))
#endif

#include "../../pbrt-v3-light-portals_b200/csrc/device/layout_build.h"

using namespace pg;

namespace {
struct SimScene {
    std::vector<F4> wide, tris;
    DevBVH bvh;
    const pbrtgpu_scene *src;
};

template <typename F>
void parallelFor(int64_t n, int threads, F f) {
    if (threads <= 1) { f(0, n); return; }
    std::vector<std::thread> pool;
    int64_t chunk = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        int64_t a = t * chunk, b = std::min<int64_t>(n, a + chunk);
        if (a >= b) break;
        pool.emplace_back([=] { f(a, b); });
    }
    for (auto &th : pool) th.join();
}
}  // namespace

extern "C" {

void *hostsim_create(const pbrtgpu_scene *s) {
    auto *sim = new SimScene;
    sim->src = s;
    buildWideNodes(s, &sim->wide);
    std::string err;
    if (!buildTris(s, &sim->tris, &err)) { delete sim; return nullptr; }
    sim->bvh.nodes = sim->wide.data();
    sim->bvh.tris = sim->tris.data();
    sim->bvh.planeParams = s->plane_params;
    sim->bvh.sphereParams = s->sphere_params;
    return sim;
}
void hostsim_destroy(void *p) { delete (SimScene *)p; }

int hostsim_intersect_closest(void *p, const float *rays, int64_t n, int32_t *ids, float *t, float *b0b1, int threads) {
    SimScene *sim = (SimScene *)p;
    parallelFor(n, threads, [&](int64_t a, int64_t b) {
        for (int64_t i = a; i < b; ++i) {
            const float *r = rays + 8 * i;
            Ray ray;
            ray.o = v3(r[0], r[1], r[2]); ray.tMax = r[3]; ray.d = v3(r[4], r[5], r[6]);
            Hit h;
            traverse<false>(sim->bvh, ray, &h);
            ids[i] = h.prim >= 0 ? sim->src->prim_orig_id[h.prim] : -1;
            if (t) t[i] = ray.tMax;
            if (b0b1) { b0b1[2 * i] = h.prim >= 0 ? h.b0 : 0; b0b1[2 * i + 1] = h.prim >= 0 ? h.b1 : 0; }
        }
    });
    return 0;
}

int hostsim_intersect_any(void *p, const float *rays, int64_t n, uint8_t *occ, int threads) {
    SimScene *sim = (SimScene *)p;
    parallelFor(n, threads, [&](int64_t a, int64_t b) {
        for (int64_t i = a; i < b; ++i) {
            const float *r = rays + 8 * i;
            Ray ray;
            ray.o = v3(r[0], r[1], r[2]); ray.tMax = r[3]; ray.d = v3(r[4], r[5], r[6]);
            occ[i] = traverse<true>(sim->bvh, ray, nullptr) ? 1 : 0;
        }
    });
    return 0;
}

}  // extern "C"

// ---- the wavefront path integrator, stepped on the CPU (same stage functions
// as render.cu's kernels, same batch plan, same buffers)
#include "../../pbrt-v3-light-portals_b200/csrc/common/halton_tables.h"
#include "../../pbrt-v3-light-portals_b200/csrc/device/wave_setup.h"

extern "C" int hostsim_render(void *p, const pbrtgpu_render_params *rp, float *film_xyzw, int64_t *stats4, int threads) {
    SimScene *sim = (SimScene *)p;
    const pbrtgpu_scene *s = sim->src;
    DevScene sc;
    memset(&sc, 0, sizeof(sc));
    sc.bvh = sim->bvh;
    sc.nPrims = s->n_prims, sc.nLights = s->n_lights, sc.nMaterials = s->n_materials;
    sc.primOrigId = s->prim_orig_id, sc.primMaterial = s->prim_material, sc.primLight = s->prim_light;
    sc.primFlags = s->prim_flags, sc.triN = s->tri_n, sc.triUV = s->tri_uv, sc.triS = s->tri_s;
    sc.matTag = s->mat_tag, sc.matParams = s->mat_params;
    sc.lightTag = s->light_tag, sc.lightPrim = s->light_prim, sc.lightStrategy = s->light_strategy;
    sc.lightPortalBegin = s->light_portal_begin, sc.portalParams = s->portal_params;
    sc.lightNSamples = s->light_nsamples;
    std::vector<MatLobes> matLobes;
    buildMaterialLobes(s, &matLobes);
    sc.matLobes = matLobes.data();
    std::vector<float> lightParams(s->light_params, s->light_params + (size_t)s->n_lights * PBRTGPU_LIGHT_STRIDE);
    for (int i = 0; i < s->n_lights; ++i)
        if (s->light_tag[i] == PBRTGPU_LIGHT_INFINITE) finishInfiniteLight(lightParams.data() + (size_t)i * PBRTGPU_LIGHT_STRIDE);
    sc.lightParams = lightParams.data();
    std::vector<float> lfunc, lcdf;
    pbrtgpu_scene tmp = *s;
    tmp.light_params = lightParams.data();
    buildLightDistribution(&tmp, (rp->light_strategy == 1 && s->n_lights > 1) ? 1 : 0, &lfunc, &lcdf, &sc.lightFuncInt);
    sc.lightFunc = lfunc.data(), sc.lightCdf = lcdf.data();
    memcpy(sc.worldCenter, s->world_center, 12);
    sc.worldRadius = s->world_radius;
    bool nullMat = false;
    if (s->prim_material)
        for (int i = 0; i < s->n_prims; ++i) nullMat |= s->prim_material[i] < 0;

    std::vector<int> primes = firstPrimes(), sums(primes.size());
    int acc = 0;
    for (size_t i = 0; i < primes.size(); ++i) sums[i] = acc, acc += primes[i];
    std::vector<uint16_t> perms = radicalInversePermutations(primes);

    if (validateRenderParams(rp)) return 1;  // the same refusals as pbrtgpu_render (wave_setup.h)
    WaveParams w;
    memset(&w, 0, sizeof(w));
    fillWaveParams(rp, &w);
    w.sc = sc;
    std::vector<unsigned long long> magic = primeMagics(primes);
    w.hal.primes = primes.data(), w.hal.primeSums = sums.data(), w.hal.perms = perms.data(), w.hal.primeMagic = magic.data();
    std::vector<HaltonDim> haltonDims(primes.size());
    fillHaltonDims(primes.data(), sums.data(), perms.data(), (int)primes.size(), haltonDims.data());
    w.hal.dims = haltonDims.data();
    bool dispersiveTrans = false;
    const bool tree = rp->integrator == PBRTGPU_INTEGRATOR_DIRECT && rp->max_depth > 1 &&
                      maxSpecularTransmissionLobes(matLobes, s->n_materials, &dispersiveTrans) > 0;
    BatchPlan pl = planBatches(rp, 1 << 18, nullMat, lightSamplesPerVertex(rp, s), tree ? rp->max_depth : 0);
    w.directTree = tree ? 1 : 0;
    w.branchCapacity = tree ? rp->max_depth : 0;
    if (const char *e = getenv("PBRTGPU_TREE_ROUND_ITERS"))  // tests: short rounds, so that the carry between rounds is exercised
        if (tree && atoi(e) > 0 && atoi(e) < pl.roundIters) pl.roundIters = atoi(e);
    w.nSampleArrays = sampleArrayCount(rp, s->n_lights);
    w.hal.arrayEndDim = kArrayStartDim + 2 * w.nSampleArrays;
    const int *cb = rp->cropped_bounds;
    size_t nPix = (size_t)(cb[2] - cb[0]) * (cb[3] - cb[1]);
    std::vector<float> film(nPix * 4, 0.f);
    w.film.rgbw = film.data();
    w.film.filterTable = rp->filter_table;
    std::vector<std::vector<char>> store(WB_COUNT);
    void *bufs[WB_COUNT];
    for (int b = 0; b < WB_COUNT; ++b) {
        store[b].resize(waveBufBytes(b, pl) + 16);
        bufs[b] = store[b].data();
    }
    bindWaveBuffers(&w, bufs);
    WaveCounters cnt;
    w.cnt = &cnt;
    const bool lambertPath = getenv("PBRTGPU_HOSTSIM_LAMBERT") && atoi(getenv("PBRTGPU_HOSTSIM_LAMBERT")) != 0;
    int64_t totals[4] = {0, 0, 0, 0};
    for (int s0 = pl.sppBegin; s0 < pl.sppEnd; s0 += pl.sppChunk) {
        w.sppBegin = s0;
        w.sppCount = std::min(pl.sppChunk, pl.sppEnd - s0);
        for (int t0 = 0; t0 < pl.shardTiles; t0 += pl.tilesPerBatch) {
            int nt = std::min(pl.tilesPerBatch, pl.shardTiles - t0);
            w.tileBegin = t0;
            w.nSlots = nt * 256 * w.sppCount;
            w.neeCapacity = w.nSlots * pl.neeMult;
            memset(&cnt, 0, sizeof(cnt));
            parallelFor(w.nSlots, threads, [&](int64_t a, int64_t b) { for (int64_t i = a; i < b; ++i) stageRaygen(w, (int)i); });
            int cur = 0;
            bool firstRound = true;
            int itersDone = 0;
            for (bool again = true; again;) {  // rounds, as render_host.inc runs them
            const int roundLen = pl.openEnded ? pl.roundIters : std::min(pl.roundIters, pl.depthIters - itersDone);
            for (int d = 0; d < roundLen; ++d) {
                int n = cnt.nActive[d];
                parallelFor(n, threads, [&](int64_t a, int64_t b) {
                    for (int64_t i = a; i < b; ++i) {
                        Ray ray = loadRay(w.rays[cur], (int)i);
                        Hit h;
                        traverse<false>(w.sc.bvh, ray, &h);
                        if (h.prim >= 0) {
                            w.hits[i] = f4(bitsToFloat((unsigned)h.prim), h.b0, h.b1, h.b2);
                            w.hitIdx[atomicIncInt(&cnt.nHit[d])] = (int)i;
                        } else
                            w.missIdx[atomicIncInt(&cnt.nMiss[d])] = (int)i;
                    }
                });
                parallelFor(cnt.nMiss[d], threads, [&](int64_t a, int64_t b) { for (int64_t i = a; i < b; ++i) stageMiss(w, d, cur, w.missIdx[i]); });
                parallelFor(cnt.nHit[d], threads, [&](int64_t a, int64_t b) { 
                    for (int64_t i = a; i < b; ++i) {
                        if (w.integrator == PBRTGPU_INTEGRATOR_DIRECT) stageShadeDirect(w, d, cur, w.hitIdx[i]);
                        else if (lambertPath) {  // as the CUDA driver does: single-Lambert-lobe materials through the Lambert instantiation
                            int prim = f2i(w.hits[w.hitIdx[i]].x);
                            int m = w.sc.primMaterial ? w.sc.primMaterial[prim] : -1;
                            if (m < 0 || w.sc.matLobes[m].shadeClass == 0) stageShadeLambert(w, d, cur, w.hitIdx[i]);
                            else if (w.sc.matLobes[m].shadeClass == 2) stageShadeSpecular(w, d, cur, w.hitIdx[i]);  // all-specular materials
                            else stageShade(w, d, cur, w.hitIdx[i]);
                        } else stageShade(w, d, cur, w.hitIdx[i]);
                    }
                });
                parallelFor(cnt.nShadow[d], threads, [&](int64_t a, int64_t b) {
                    for (int64_t i = a; i < b; ++i) {
                        Ray ray = loadRay(w.shadowRays, (int)i);
                        stageShadowResolve(w, (int)i, traverse<true>(w.sc.bvh, ray, nullptr));
                    }
                });
                parallelFor(cnt.nNee[d], threads, [&](int64_t a, int64_t b) {
                    for (int64_t i = a; i < b; ++i) {
                        Ray ray = loadRay(w.neeRays, (int)i);
                        Hit h;
                        traverse<false>(w.sc.bvh, ray, &h);
                        w.neeHits[i] = f4(bitsToFloat((unsigned)h.prim), h.b0, h.b1, h.b2);
                        stageNeeResolve(w, (int)i);
                    }
                });
                totals[1] += n + cnt.nNee[d];
                totals[2] += cnt.nShadow[d];
                cur ^= 1;
            }
            // tree mode: paths that still hold pending branches go through another round (render_host.inc, anotherRound)
            if (firstRound) totals[0] += cnt.nActive[0];
            firstRound = false;
            itersDone += roundLen;
            int left = cnt.nActive[roundLen];
            again = left > 0 && (pl.openEnded ? itersDone < kMaxIterationsPerBatch : itersDone < pl.depthIters);
            memset(&cnt, 0, sizeof(cnt));
            cnt.nActive[0] = again ? left : 0;
            }
            parallelFor(w.nSlots, threads, [&](int64_t a, int64_t b) { for (int64_t i = a; i < b; ++i) stageFilm(w, (int)i); });
            totals[3] += (int64_t)cnt.scrubbed;
        }
    }
    for (size_t i = 0; i < nPix; ++i) {
        float xyz[3];
        rgbToXYZ(&film[4 * i], xyz);
        film_xyzw[4 * i] = xyz[0], film_xyzw[4 * i + 1] = xyz[1], film_xyzw[4 * i + 2] = xyz[2], film_xyzw[4 * i + 3] = film[4 * i + 3];
    }
    if (stats4) memcpy(stats4, totals, sizeof(totals));
    return 0;
}

// the digit permutation of one dimension (for tests that restate ScrambledRadicalInverse independently)
extern "C" int hostsim_halton_perm(int dim, int32_t *base, uint16_t *out) {
    static std::vector<int> primes = firstPrimes();
    static std::vector<uint16_t> perms = radicalInversePermutations(primes);
    if (dim < 0 || dim >= (int)primes.size()) return 1;
    size_t off = 0;
    for (int i = 0; i < dim; ++i) off += primes[i];
    *base = primes[dim];
    for (int i = 0; i < primes[dim]; ++i) out[i] = perms[off + i];
    return 0;
}

// sphereRejects (the cheap exact early-out) next to sphereTest (the interval version) on the same rays:
// rays = 8 floats each (o.xyz, tMax, d.xyz, -), rec = one PBRTGPU_SPHERE_STRIDE record
extern "C" int hostsim_sphere_guard(const float *rec, const float *rays, int64_t n, uint8_t *rejects, uint8_t *accepts) {
    for (int64_t i = 0; i < n; ++i) {
        Ray r;
        r.o = v3(rays[8 * i], rays[8 * i + 1], rays[8 * i + 2]);
        r.tMax = rays[8 * i + 3];
        r.d = v3(rays[8 * i + 4], rays[8 * i + 5], rays[8 * i + 6]);
        float t;
        rejects[i] = sphereRejects(r, rec) ? 1 : 0;
        accepts[i] = sphereTest(r, rec, &t, nullptr, nullptr, nullptr) ? 1 : 0;
    }
    return 0;
}

extern "C" int hostsim_halton(const pbrtgpu_render_params *rp, const int32_t *xy, const int64_t *num, const int32_t *dim, int64_t n, float *out) {
    static std::vector<int> primes = firstPrimes(), sums;
    static std::vector<uint16_t> perms;
    static std::vector<unsigned long long> magic = primeMagics(primes);
    if (sums.empty()) {
        sums.resize(primes.size());
        int acc = 0;
        for (size_t i = 0; i < primes.size(); ++i) sums[i] = acc, acc += primes[i];
        perms = radicalInversePermutations(primes);
    }
    HaltonParams h = makeHaltonParams(rp->sample_bounds, rp->sample_at_pixel_center);
    h.primes = primes.data(), h.primeSums = sums.data(), h.perms = perms.data(), h.primeMagic = magic.data();
    static std::vector<HaltonDim> haltonDims;
    if (haltonDims.empty()) {
        haltonDims.resize(primes.size());
        fillHaltonDims(primes.data(), sums.data(), perms.data(), (int)primes.size(), haltonDims.data());
    }
    h.dims = haltonDims.data();
    for (int64_t i = 0; i < n; ++i)
        out[i] = haltonSample(h, haltonPixelOffset(h, xy[2 * i], xy[2 * i + 1]) + num[i] * h.sampleStride, dim[i]);
    return 0;
}

// AAPortal::InFrustum as portal.h restates it (strategy "solidangle_mis3"), for the geometric test against brute-force
// visibility: out[i] = 1 if points[3i..] lies in the viewing frustum of portal `portal` of light `light`.
extern "C" int hostsim_portal_in_frustum(const pbrtgpu_scene *s, int light, int portal, const float *points, int64_t n, uint8_t *out) {
    if (light < 0 || light >= s->n_lights || s->light_tag[light] != PBRTGPU_LIGHT_PORTAL_AREA) return -1;
    int slot = s->light_prim[light];
    const float *lightRec = s->plane_params + PBRTGPU_PLANE_STRIDE * (size_t)s->prim_shape[slot];
    int pb = s->light_portal_begin[light];
    if (portal < 0 || pb + portal >= s->light_portal_begin[light + 1]) return -1;
    PortalRec po = loadPortal(s->portal_params + PBRTGPU_PORTAL_STRIDE * (size_t)(pb + portal));
    PlaneParams lp = loadPlane(lightRec);
    M44 w2o = loadM44(lightRec + 32);
    for (int64_t i = 0; i < n; ++i) {
        bool front = portalInFront(po, w2o, v3(points[3 * i], points[3 * i + 1], points[3 * i + 2]));
        out[i] = (front ? 2 : 0) | (portalInFrustum(po, lp, w2o, v3(points[3 * i], points[3 * i + 1], points[3 * i + 2])) ? 1 : 0);
    }
    return 0;
}
