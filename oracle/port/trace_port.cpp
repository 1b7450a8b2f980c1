// TEST INFRASTRUCTURE ONLY -- the oracle.  Never imported, linked or called by
// the product (pbrt-v3-light-portals_b200/); only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline leg may use it, and only as the checker.
//
// CPU restatement of the reference's ray/scene intersection over the flat
// arrays that cross the C ABI (pbrtgpu_scene, reference node layout):
//   BVHAccel::Intersect / IntersectP      src/accelerators/bvh.cpp:662-738
//   Bounds3::IntersectP                   src/core/geometry.h:1580-1606
//   Triangle::Intersect / IntersectP      src/shapes/triangle.cpp:189-292, 427-520
//   AAPlaneShape::Intersect               src/shapes/plane.cpp:15-55
//   Sphere::Intersect / IntersectP        src/shapes/sphere.cpp:49-104, 158-215
//   EFloat, Quadratic                     src/core/efloat.h
// It walks the reference's own LinearBVHNode array (not the device layout),
// and counts node visits V and primitive tests T per batch -- the figures
// SURVEY.md 8(d) bases the algorithmic bytes on.
// Parity pin: tests/test_oracle_pins.py checks this file's output against
// hit IDs and t values dumped from the reference itself (oracle/_ref/ref_harness)
// and committed under tests/golden/.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "../../include/pbrtgpu.h"

namespace {

const float kEps = 5.9604644775390625e-08f;
inline float gam(int n) { return (n * kEps) / (1 - n * kEps); }
inline float fmin2(float a, float b) { return b < a ? b : a; }
inline float fmax2(float a, float b) { return a < b ? b : a; }
inline uint32_t bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
inline float fromBits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
inline float up(float v) {
    if (std::isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t u = bits(v);
    if (v >= 0) ++u; else --u;
    return fromBits(u);
}
inline float down(float v) {
    if (std::isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t u = bits(v);
    if (v > 0) --u; else ++u;
    return fromBits(u);
}

struct RayP {
    float o[3], d[3];
    float tMax;
};

// geometry.h:1580-1606
bool boxHit(const float *b, const RayP &r, const float invDir[3], const int neg[3]) {
    const float *lo = b, *hi = b + 3;
    float tMin = ((neg[0] ? hi : lo)[0] - r.o[0]) * invDir[0];
    float tMax = ((neg[0] ? lo : hi)[0] - r.o[0]) * invDir[0];
    float tyMin = ((neg[1] ? hi : lo)[1] - r.o[1]) * invDir[1];
    float tyMax = ((neg[1] ? lo : hi)[1] - r.o[1]) * invDir[1];
    tMax *= 1 + 2 * gam(3);
    tyMax *= 1 + 2 * gam(3);
    if (tMin > tyMax || tyMin > tMax) return false;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = ((neg[2] ? hi : lo)[2] - r.o[2]) * invDir[2];
    float tzMax = ((neg[2] ? lo : hi)[2] - r.o[2]) * invDir[2];
    tzMax *= 1 + 2 * gam(3);
    if (tMin > tzMax || tzMin > tMax) return false;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    return (tMin < r.tMax) && (tMax > 0);
}

inline void cross3(const float a[3], const float b[3], float out[3]) {  // geometry.h:1110-1116
    double ax = a[0], ay = a[1], az = a[2], bx = b[0], by = b[1], bz = b[2];
    out[0] = (float)((ay * bz) - (az * by));
    out[1] = (float)((az * bx) - (ax * bz));
    out[2] = (float)((ax * by) - (ay * bx));
}

// triangle.cpp:199-292; `closest` adds the degenerate-triangle rejection of :309-317
bool triHit(const RayP &r, const float *P, const float *uvOrNull, bool closest, float *tOut, float *b0Out, float *b1Out) {
    float pt[3][3];
    for (int k = 0; k < 3; ++k)
        for (int a = 0; a < 3; ++a) pt[k][a] = P[3 * k + a] - r.o[a];
    float ad[3] = {std::fabs(r.d[0]), std::fabs(r.d[1]), std::fabs(r.d[2])};
    int kz = (ad[0] > ad[1]) ? ((ad[0] > ad[2]) ? 0 : 2) : ((ad[1] > ad[2]) ? 1 : 2);
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    float d[3] = {r.d[kx], r.d[ky], r.d[kz]};
    float q[3][3];
    for (int k = 0; k < 3; ++k) { q[k][0] = pt[k][kx]; q[k][1] = pt[k][ky]; q[k][2] = pt[k][kz]; }
    float Sx = -d[0] / d[2], Sy = -d[1] / d[2], Sz = 1.f / d[2];
    for (int k = 0; k < 3; ++k) { q[k][0] += Sx * q[k][2]; q[k][1] += Sy * q[k][2]; }
    float e0 = q[1][0] * q[2][1] - q[1][1] * q[2][0];
    float e1 = q[2][0] * q[0][1] - q[2][1] * q[0][0];
    float e2 = q[0][0] * q[1][1] - q[0][1] * q[1][0];
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {
        double a = (double)q[2][0] * (double)q[1][1], b = (double)q[2][1] * (double)q[1][0];
        e0 = (float)(b - a);
        a = (double)q[0][0] * (double)q[2][1]; b = (double)q[0][1] * (double)q[2][0];
        e1 = (float)(b - a);
        a = (double)q[1][0] * (double)q[0][1]; b = (double)q[1][1] * (double)q[0][0];
        e2 = (float)(b - a);
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    q[0][2] *= Sz; q[1][2] *= Sz; q[2][2] *= Sz;
    float tScaled = e0 * q[0][2] + e1 * q[1][2] + e2 * q[2][2];
    if (det < 0 && (tScaled >= 0 || tScaled < r.tMax * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > r.tMax * det)) return false;
    float invDet = 1 / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, t = tScaled * invDet;
    float maxZt = fmax2(std::fabs(q[0][2]), fmax2(std::fabs(q[1][2]), std::fabs(q[2][2])));
    float deltaZ = gam(3) * maxZt;
    float maxXt = fmax2(std::fabs(q[0][0]), fmax2(std::fabs(q[1][0]), std::fabs(q[2][0])));
    float maxYt = fmax2(std::fabs(q[0][1]), fmax2(std::fabs(q[1][1]), std::fabs(q[2][1])));
    float deltaX = gam(5) * (maxXt + maxZt), deltaY = gam(5) * (maxYt + maxZt);
    float deltaE = 2 * (gam(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = fmax2(std::fabs(e0), fmax2(std::fabs(e1), std::fabs(e2)));
    float deltaT = 3 * (gam(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * std::fabs(invDet);
    if (t <= deltaT) return false;
    if (closest) {
        float uv[3][2] = {{0, 0}, {1, 0}, {1, 1}};
        if (uvOrNull) for (int i = 0; i < 3; ++i) { uv[i][0] = uvOrNull[2 * i]; uv[i][1] = uvOrNull[2 * i + 1]; }
        float duv02[2] = {uv[0][0] - uv[2][0], uv[0][1] - uv[2][1]}, duv12[2] = {uv[1][0] - uv[2][0], uv[1][1] - uv[2][1]};
        float dp02[3], dp12[3];
        for (int a = 0; a < 3; ++a) { dp02[a] = P[a] - P[6 + a]; dp12[a] = P[3 + a] - P[6 + a]; }
        float determinant = duv02[0] * duv12[1] - duv02[1] * duv12[0];
        bool degenerateUV = std::fabs(determinant) < 1e-8;
        float dpdu[3] = {0, 0, 0}, dpdv[3] = {0, 0, 0};
        if (!degenerateUV) {
            float invdet = 1 / determinant;
            for (int a = 0; a < 3; ++a) {
                dpdu[a] = (duv12[1] * dp02[a] - duv02[1] * dp12[a]) * invdet;
                dpdv[a] = (-duv12[0] * dp02[a] + duv02[0] * dp12[a]) * invdet;
            }
        }
        float c[3];
        cross3(dpdu, dpdv, c);
        if (degenerateUV || (c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) == 0) {
            float e20[3], e10[3], ng[3];
            for (int a = 0; a < 3; ++a) { e20[a] = P[6 + a] - P[a]; e10[a] = P[3 + a] - P[a]; }
            cross3(e20, e10, ng);
            if ((ng[0] * ng[0] + ng[1] * ng[1] + ng[2] * ng[2]) == 0) return false;
        }
    }
    *tOut = t; *b0Out = b0; *b1Out = b1;
    return true;
}

// Transform::operator()(const Ray&, Vector3f*, Vector3f*), transform.h:402-414
void rayToObject(const float *m /*row-major 4x4*/, const RayP &r, RayP *out, float oErr[3], float dErr[3]) {
    float x = r.o[0], y = r.o[1], z = r.o[2];
    float p[4];
    for (int i = 0; i < 4; ++i) p[i] = (m[4 * i] * x + m[4 * i + 1] * y) + (m[4 * i + 2] * z + m[4 * i + 3]);
    for (int i = 0; i < 3; ++i)
        oErr[i] = gam(3) * (std::fabs(m[4 * i] * x) + std::fabs(m[4 * i + 1] * y) + std::fabs(m[4 * i + 2] * z) + std::fabs(m[4 * i + 3]));
    if (p[3] == 1) { out->o[0] = p[0]; out->o[1] = p[1]; out->o[2] = p[2]; }
    else { float inv = 1.f / p[3]; out->o[0] = p[0] * inv; out->o[1] = p[1] * inv; out->o[2] = p[2] * inv; }
    float vx = r.d[0], vy = r.d[1], vz = r.d[2];
    for (int i = 0; i < 3; ++i) {
        dErr[i] = gam(3) * (std::fabs(m[4 * i] * vx) + std::fabs(m[4 * i + 1] * vy) + std::fabs(m[4 * i + 2] * vz));
        out->d[i] = m[4 * i] * vx + m[4 * i + 1] * vy + m[4 * i + 2] * vz;
    }
    out->tMax = r.tMax;
    float l2 = out->d[0] * out->d[0] + out->d[1] * out->d[1] + out->d[2] * out->d[2];
    if (l2 > 0) {
        float dt = (std::fabs(out->d[0]) * oErr[0] + std::fabs(out->d[1]) * oErr[1] + std::fabs(out->d[2]) * oErr[2]) / l2;
        for (int i = 0; i < 3; ++i) out->o[i] += out->d[i] * dt;
    }
}

// plane.cpp:15-31
bool planeHit(const RayP &r, const float *rec, float *tOut) {
    RayP ro; float oe[3], de[3];
    rayToObject(rec + 32, r, &ro, oe, de);
    int ax = (int)rec[6];
    int ax0 = ax == 2 ? 0 : (ax == 0 ? 1 : 2), ax1 = ax == 2 ? 1 : (ax == 0 ? 2 : 0);
    const float *lo = rec, *hi = rec + 3;
    float t = (lo[ax] - ro.o[ax]) / ro.d[ax];
    float ph[3];
    for (int i = 0; i < 3; ++i) ph[i] = ro.o[i] + t * ro.d[i];
    if (ph[ax0] > lo[ax0] && ph[ax0] < hi[ax0] && ph[ax1] > lo[ax1] && ph[ax1] < hi[ax1] && t < ro.tMax) {
        *tOut = t;
        return true;
    }
    return false;
}

struct EF { float v, lo, hi; };
EF mk(float v, float err = 0.f) { EF r; r.v = v; if (err == 0.f) r.lo = r.hi = v; else { r.lo = down(v - err); r.hi = up(v + err); } return r; }
EF add(EF a, EF b) { EF r; r.v = a.v + b.v; r.lo = down(a.lo + b.lo); r.hi = up(a.hi + b.hi); return r; }
EF sub(EF a, EF b) { EF r; r.v = a.v - b.v; r.lo = down(a.lo - b.hi); r.hi = up(a.hi - b.lo); return r; }
EF mul(EF a, EF b) {
    EF r; r.v = a.v * b.v;
    float p[4] = {a.lo * b.lo, a.hi * b.lo, a.lo * b.hi, a.hi * b.hi};
    r.lo = down(fmin2(fmin2(p[0], p[1]), fmin2(p[2], p[3])));
    r.hi = up(fmax2(fmax2(p[0], p[1]), fmax2(p[2], p[3])));
    return r;
}
EF dvd(EF a, EF b) {
    EF r; r.v = a.v / b.v;
    if (b.lo < 0 && b.hi > 0) { r.lo = -INFINITY; r.hi = INFINITY; }
    else {
        float p[4] = {a.lo / b.lo, a.hi / b.lo, a.lo / b.hi, a.hi / b.hi};
        r.lo = down(fmin2(fmin2(p[0], p[1]), fmin2(p[2], p[3])));
        r.hi = up(fmax2(fmax2(p[0], p[1]), fmax2(p[2], p[3])));
    }
    return r;
}

// sphere.cpp:49-104 / 158-215
bool sphereHit(const RayP &r, const float *rec, float *tOut) {
    float radius = rec[0], zMin = rec[1], zMax = rec[2], phiMax = rec[5];
    RayP ray; float oe[3], de[3];
    rayToObject(rec + 32, r, &ray, oe, de);
    EF ox = mk(ray.o[0], oe[0]), oy = mk(ray.o[1], oe[1]), oz = mk(ray.o[2], oe[2]);
    EF dx = mk(ray.d[0], de[0]), dy = mk(ray.d[1], de[1]), dz = mk(ray.d[2], de[2]);
    EF a = add(add(mul(dx, dx), mul(dy, dy)), mul(dz, dz));
    EF b = mul(mk(2.f), add(add(mul(dx, ox), mul(dy, oy)), mul(dz, oz)));
    EF c = sub(add(add(mul(ox, ox), mul(oy, oy)), mul(oz, oz)), mul(mk(radius), mk(radius)));
    double discrim = (double)b.v * (double)b.v - 4. * (double)a.v * (double)c.v;
    if (discrim < 0.) return false;
    double rootDiscrim = std::sqrt(discrim);
    EF frd = mk((float)rootDiscrim, (float)(kEps * rootDiscrim));
    EF q = b.v < 0 ? mul(mk(-.5f), sub(b, frd)) : mul(mk(-.5f), add(b, frd));
    EF t0 = dvd(q, a), t1 = dvd(c, q);
    if (t0.v > t1.v) std::swap(t0, t1);
    if (t0.hi > ray.tMax || t1.lo <= 0) return false;
    EF th = t0;
    if (th.lo <= 0) { th = t1; if (th.hi > ray.tMax) return false; }
    auto refine = [&](float t, float ph[3], float *phi) {
        for (int i = 0; i < 3; ++i) ph[i] = ray.o[i] + ray.d[i] * t;
        float dist = std::sqrt(ph[0] * ph[0] + ph[1] * ph[1] + ph[2] * ph[2]);
        float s = radius / dist;
        for (int i = 0; i < 3; ++i) ph[i] *= s;
        if (ph[0] == 0 && ph[1] == 0) ph[0] = 1e-5f * radius;
        *phi = std::atan2(ph[1], ph[0]);
        if (*phi < 0) *phi += 2 * 3.14159265358979323846f;
    };
    float ph[3], phi;
    refine(th.v, ph, &phi);
    if ((zMin > -radius && ph[2] < zMin) || (zMax < radius && ph[2] > zMax) || phi > phiMax) {
        if (th.v == t1.v) return false;
        if (t1.hi > ray.tMax) return false;
        th = t1;
        refine(th.v, ph, &phi);
        if ((zMin > -radius && ph[2] < zMin) || (zMax < radius && ph[2] > zMax) || phi > phiMax) return false;
    }
    *tOut = th.v;
    return true;
}

struct Counters { long long nodes = 0, prims = 0; };

// bvh.cpp:662-738 over the reference node layout
template <bool ANY>
bool walk(const pbrtgpu_scene *s, RayP &ray, int *primSlot, float *b0, float *b1, Counters *cnt) {
    bool hit = false;
    float invDir[3] = {1 / ray.d[0], 1 / ray.d[1], 1 / ray.d[2]};
    int neg[3] = {invDir[0] < 0, invDir[1] < 0, invDir[2] < 0};
    int toVisit = 0, cur = 0, stack[64];
    while (true) {
        cnt->nodes++;
        if (boxHit(s->bvh_bounds + 6 * (size_t)cur, ray, invDir, neg)) {
            int np = s->bvh_nprims[cur];
            if (np > 0) {
                for (int i = 0; i < np; ++i) {
                    int slot = s->bvh_offset[cur] + i;
                    cnt->prims++;
                    float t, bb0 = 0, bb1 = 0;
                    bool h;
                    int type = s->prim_type[slot];
                    if (type == PBRTGPU_PRIM_TRIANGLE) {
                        const float *uv = (s->tri_uv && s->prim_flags && (s->prim_flags[slot] & PBRTGPU_PF_HAS_UV)) ? s->tri_uv + 6 * (size_t)slot : nullptr;
                        h = triHit(ray, s->tri_p + 9 * (size_t)slot, uv, !ANY, &t, &bb0, &bb1);
                    } else if (type == PBRTGPU_PRIM_AAPLANE)
                        h = planeHit(ray, s->plane_params + PBRTGPU_PLANE_STRIDE * (size_t)s->prim_shape[slot], &t);
                    else
                        h = sphereHit(ray, s->sphere_params + PBRTGPU_SPHERE_STRIDE * (size_t)s->prim_shape[slot], &t);
                    if (h) {
                        if (ANY) return true;
                        hit = true;
                        ray.tMax = t;  // GeometricPrimitive::Intersect, primitive.cpp:121
                        *primSlot = slot; *b0 = bb0; *b1 = bb1;
                    }
                }
                if (toVisit == 0) break;
                cur = stack[--toVisit];
            } else {
                if (neg[s->bvh_axis[cur]]) { stack[toVisit++] = cur + 1; cur = s->bvh_offset[cur]; }
                else { stack[toVisit++] = s->bvh_offset[cur]; cur = cur + 1; }
            }
        } else {
            if (toVisit == 0) break;
            cur = stack[--toVisit];
        }
    }
    return hit;
}

template <typename F>
void parallelFor(int64_t n, int threads, F f) {
    if (threads <= 1) { f(0, n, 0); return; }
    std::vector<std::thread> pool;
    int64_t chunk = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
        int64_t a = t * chunk, b = std::min<int64_t>(n, a + chunk);
        if (a >= b) break;
        pool.emplace_back([=] { f(a, b, t); });
    }
    for (auto &th : pool) th.join();
}

}  // namespace

extern "C" {

// counters_out[0] = node visits, [1] = primitive tests (summed over the batch)
int oracle_intersect_closest(const pbrtgpu_scene *s, const float *rays, int64_t n, int32_t *ids, float *t, float *b0b1,
                             int threads, int64_t *counters_out) {
    std::vector<Counters> cs(threads > 0 ? threads : 1);
    parallelFor(n, threads, [&](int64_t a, int64_t b, int tid) {
        for (int64_t i = a; i < b; ++i) {
            const float *r = rays + 8 * i;
            RayP ray = {{r[0], r[1], r[2]}, {r[4], r[5], r[6]}, r[3]};
            int slot = -1; float b0 = 0, b1 = 0;
            bool h = walk<false>(s, ray, &slot, &b0, &b1, &cs[tid]);
            ids[i] = h ? s->prim_orig_id[slot] : -1;
            if (t) t[i] = ray.tMax;
            if (b0b1) { b0b1[2 * i] = h ? b0 : 0; b0b1[2 * i + 1] = h ? b1 : 0; }
        }
    });
    if (counters_out) { counters_out[0] = counters_out[1] = 0; for (auto &c : cs) { counters_out[0] += c.nodes; counters_out[1] += c.prims; } }
    return 0;
}

int oracle_intersect_any(const pbrtgpu_scene *s, const float *rays, int64_t n, uint8_t *occ, int threads, int64_t *counters_out) {
    std::vector<Counters> cs(threads > 0 ? threads : 1);
    parallelFor(n, threads, [&](int64_t a, int64_t b, int tid) {
        for (int64_t i = a; i < b; ++i) {
            const float *r = rays + 8 * i;
            RayP ray = {{r[0], r[1], r[2]}, {r[4], r[5], r[6]}, r[3]};
            int slot; float b0, b1;
            occ[i] = walk<true>(s, ray, &slot, &b0, &b1, &cs[tid]) ? 1 : 0;
        }
    });
    if (counters_out) { counters_out[0] = counters_out[1] = 0; for (auto &c : cs) { counters_out[0] += c.nodes; counters_out[1] += c.prims; } }
    return 0;
}

}  // extern "C"
