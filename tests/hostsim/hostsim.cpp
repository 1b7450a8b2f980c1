// TEST INFRASTRUCTURE ONLY.
//
// There is no GPU in the development container, so the `-m "not gpu"` suite
// compiles the device headers (csrc/device/*.h, all PG_HD = __host__ __device__
// functions) with g++ and steps them on the CPU, one "thread" at a time, to
// catch logic errors before GPU time is spent.  This is a debugging aid for
// the kernels' device functions; it is built only into tests/_build/ and is
// never loaded by the product library, bench.py or the CLI (those fail loudly
// without a CUDA device).
#include <cstdint>
#include <string>
#include <thread>
#include <vector>

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
