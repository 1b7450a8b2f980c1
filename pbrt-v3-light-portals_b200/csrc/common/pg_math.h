// Shared vector / matrix arithmetic for the host flattener and the device
// kernels.  Every function states the reference routine whose rounding
// behaviour it reproduces, because tier-1 parity (bit-exact hit IDs) depends
// on the order of operations, not just on the formulas:
//   * Vector/Point/Normal division multiplies by a rounded reciprocal
//     (src/core/geometry.h:244-248),
//   * Cross() is evaluated in double and rounded once (geometry.h:1110-1116),
//   * Dot/LengthSquared accumulate left to right in float.
// Compile device code with -fmad=false and host code with -ffp-contract=off so
// that no multiply-add is fused behind our back.
#ifndef PG_MATH_H
#define PG_MATH_H

#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define PG_HD __host__ __device__ __forceinline__
#define PG_HD_NOINLINE __host__ __device__ __noinline__ inline
#else
#define PG_HD inline
#define PG_HD_NOINLINE inline
#endif

// Call boundaries of the shade stage, by group (tuning: -DPG_NI_BSDF=PG_HD inlines that group)
#ifndef PG_NI_INTERACTION
#define PG_NI_INTERACTION PG_HD_NOINLINE
#endif
#ifndef PG_NI_SAMPLER
#define PG_NI_SAMPLER PG_HD_NOINLINE
#endif
#ifndef PG_NI_LOBE
#define PG_NI_LOBE PG_HD_NOINLINE
#endif
#ifndef PG_NI_BSDF
#define PG_NI_BSDF PG_HD_NOINLINE
#endif
#ifndef PG_NI_DIRECT
#define PG_NI_DIRECT PG_HD_NOINLINE
#endif

// read-only global loads (scalar); on the host a plain dereference
#if defined(__CUDA_ARCH__)
#define PG_LD(p) __ldg(p)
#else
#define PG_LD(p) (*(p))
#endif

namespace pg {

constexpr float kInfinity = INFINITY;
constexpr float kMachineEpsilon = 5.9604644775390625e-08f;  // 2^-24, pbrt.h:204-207
constexpr float kShadowEpsilon = 0.0001f;                    // pbrt.h:209
constexpr float kPi = 3.14159265358979323846f;
constexpr float kInvPi = 0.31830988618379067154f;
constexpr float kInv2Pi = 0.15915494309189533577f;
constexpr float kInv4Pi = 0.07957747154594766788f;
constexpr float kPiOver2 = 1.57079632679489661923f;
constexpr float kPiOver4 = 0.78539816339744830961f;
constexpr float kOneMinusEpsilon = 0x1.fffffep-1f;  // rng.h:50

// pbrt.h:294-296
PG_HD constexpr float gammaf(int n) {
    return (n * kMachineEpsilon) / (1 - n * kMachineEpsilon);
}

PG_HD uint32_t floatToBits(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}
PG_HD float bitsToFloat(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

// pbrt.h:246-270
PG_HD float nextFloatUp(float v) {
    if (std::isinf(v) && v > 0.f) return v;
    if (v == -0.f) v = 0.f;
    uint32_t ui = floatToBits(v);
    if (v >= 0) ++ui; else --ui;
    return bitsToFloat(ui);
}
PG_HD float nextFloatDown(float v) {
    if (std::isinf(v) && v < 0.f) return v;
    if (v == 0.f) v = -0.f;
    uint32_t ui = floatToBits(v);
    if (v > 0) --ui; else ++ui;
    return bitsToFloat(ui);
}

PG_HD bool isInfF(float v) { return v == kInfinity || v == -kInfinity; }
PG_HD float clampf(float v, float lo, float hi) {
    return v < lo ? lo : (v > hi ? hi : v);
}
PG_HD float minf(float a, float b) { return b < a ? b : a; }  // std::min
PG_HD float maxf(float a, float b) { return a < b ? b : a; }  // std::max
PG_HD float radians(float deg) { return (kPi / 180) * deg; }  // pbrt.h

struct V3 {
    float x, y, z;
    PG_HD float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    PG_HD void set(int i, float v) { if (i == 0) x = v; else if (i == 1) y = v; else z = v; }
};

PG_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
PG_HD V3 v3(float s) { return v3(s, s, s); }
PG_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
PG_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
PG_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
PG_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
PG_HD V3 operator*(float s, V3 a) { return v3(a.x * s, a.y * s, a.z * s); }
PG_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
// geometry.h:244-248: multiply by the rounded reciprocal.
PG_HD V3 operator/(V3 a, float f) { float inv = 1.f / f; return v3(a.x * inv, a.y * inv, a.z * inv); }
PG_HD V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
PG_HD V3 &operator*=(V3 &a, float s) { a = a * s; return a; }
PG_HD V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
PG_HD bool operator==(V3 a, V3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
PG_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
PG_HD float absDot(V3 a, V3 b) { return fabsf(dot(a, b)); }
PG_HD float lengthSquared(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
PG_HD float length(V3 a) { return sqrtf(lengthSquared(a)); }
PG_HD V3 normalize(V3 a) { return a / length(a); }
PG_HD V3 vabs(V3 a) { return v3(fabsf(a.x), fabsf(a.y), fabsf(a.z)); }
PG_HD V3 vmin(V3 a, V3 b) { return v3(minf(a.x, b.x), minf(a.y, b.y), minf(a.z, b.z)); }
PG_HD V3 vmax(V3 a, V3 b) { return v3(maxf(a.x, b.x), maxf(a.y, b.y), maxf(a.z, b.z)); }
PG_HD float maxComponent(V3 a) { return maxf(a.x, maxf(a.y, a.z)); }
PG_HD float distanceSquared(V3 a, V3 b) { return lengthSquared(a - b); }
PG_HD float distance(V3 a, V3 b) { return length(a - b); }
// geometry.h:1110-1116: double products, one rounding per component.
PG_HD V3 cross(V3 a, V3 b) {
    double ax = a.x, ay = a.y, az = a.z, bx = b.x, by = b.y, bz = b.z;
    return v3((float)((ay * bz) - (az * by)), (float)((az * bx) - (ax * bz)),
              (float)((ax * by) - (ay * bx)));
}
// geometry.h:1151-1153
PG_HD int maxDimension(V3 v) {
    return (v.x > v.y) ? ((v.x > v.z) ? 0 : 2) : ((v.y > v.z) ? 1 : 2);
}
PG_HD V3 permute(V3 v, int x, int y, int z) { return v3(v[x], v[y], v[z]); }
// permute(v, kx, ky, kz) for the triples Triangle::Intersect uses (kz = the ray's largest direction component, kx = kz + 1,
// ky = kx + 1, both modulo 3; triangle.cpp:205-210): a rotation of the components, so two predicates select it
// (six conditional moves) where the indexed form compiles to a chain of branches per component.
PG_HD V3 rotateToZ(V3 v, int kz) {
    const bool z0 = kz == 0, z1 = kz == 1;
    return v3(z0 ? v.y : (z1 ? v.z : v.x), z0 ? v.z : (z1 ? v.x : v.y), z0 ? v.x : (z1 ? v.y : v.z));
}
PG_HD V3 faceforward(V3 n, V3 v) { return (dot(n, v) < 0.f) ? -n : n; }
// geometry.h:1173-1180
PG_HD void coordinateSystem(V3 v1, V3 *v2, V3 *v3o) {
    if (fabsf(v1.x) > fabsf(v1.y))
        *v2 = v3(-v1.z, 0, v1.x) / sqrtf(v1.x * v1.x + v1.z * v1.z);
    else
        *v2 = v3(0, v1.z, -v1.y) / sqrtf(v1.y * v1.y + v1.z * v1.z);
    *v3o = cross(v1, *v2);
}

// geometry.h:1608-1622
PG_HD V3 offsetRayOrigin(V3 p, V3 pError, V3 n, V3 w) {
    float d = dot(vabs(n), pError);
    V3 offset = d * n;
    if (dot(w, n) < 0) offset = -offset;
    V3 po = p + offset;
    if (offset.x > 0) po.x = nextFloatUp(po.x); else if (offset.x < 0) po.x = nextFloatDown(po.x);
    if (offset.y > 0) po.y = nextFloatUp(po.y); else if (offset.y < 0) po.y = nextFloatDown(po.y);
    if (offset.z > 0) po.z = nextFloatUp(po.z); else if (offset.z < 0) po.z = nextFloatDown(po.z);
    return po;
}

// Row-major 4x4, as Matrix4x4 (transform.h).  A Transform is {m, mInv}.
struct M44 {
    float m[4][4];
};

struct Xform {
    M44 m, mInv;
};

// transform.h:218-233
PG_HD V3 xformPoint(const M44 &t, V3 p) {
    float x = p.x, y = p.y, z = p.z;
    float xp = t.m[0][0] * x + t.m[0][1] * y + t.m[0][2] * z + t.m[0][3];
    float yp = t.m[1][0] * x + t.m[1][1] * y + t.m[1][2] * z + t.m[1][3];
    float zp = t.m[2][0] * x + t.m[2][1] * y + t.m[2][2] * z + t.m[2][3];
    float wp = t.m[3][0] * x + t.m[3][1] * y + t.m[3][2] * z + t.m[3][3];
    if (wp == 1) return v3(xp, yp, zp);
    return v3(xp, yp, zp) / wp;
}
// transform.h:235-241
PG_HD V3 xformVector(const M44 &t, V3 v) {
    float x = v.x, y = v.y, z = v.z;
    return v3(t.m[0][0] * x + t.m[0][1] * y + t.m[0][2] * z,
              t.m[1][0] * x + t.m[1][1] * y + t.m[1][2] * z,
              t.m[2][0] * x + t.m[2][1] * y + t.m[2][2] * z);
}
// transform.h:243-249: normals use the transpose of the inverse.
PG_HD V3 xformNormal(const M44 &tInv, V3 n) {
    float x = n.x, y = n.y, z = n.z;
    return v3(tInv.m[0][0] * x + tInv.m[1][0] * y + tInv.m[2][0] * z,
              tInv.m[0][1] * x + tInv.m[1][1] * y + tInv.m[2][1] * z,
              tInv.m[0][2] * x + tInv.m[1][2] * y + tInv.m[2][2] * z);
}
// transform.h:276-297: point transform with the paired-sum rounding order and
// its absolute error bound.
PG_HD V3 xformPointErr(const M44 &t, V3 p, V3 *pError) {
    float x = p.x, y = p.y, z = p.z;
    float xp = (t.m[0][0] * x + t.m[0][1] * y) + (t.m[0][2] * z + t.m[0][3]);
    float yp = (t.m[1][0] * x + t.m[1][1] * y) + (t.m[1][2] * z + t.m[1][3]);
    float zp = (t.m[2][0] * x + t.m[2][1] * y) + (t.m[2][2] * z + t.m[2][3]);
    float wp = (t.m[3][0] * x + t.m[3][1] * y) + (t.m[3][2] * z + t.m[3][3]);
    float xs = (fabsf(t.m[0][0] * x) + fabsf(t.m[0][1] * y) + fabsf(t.m[0][2] * z) + fabsf(t.m[0][3]));
    float ys = (fabsf(t.m[1][0] * x) + fabsf(t.m[1][1] * y) + fabsf(t.m[1][2] * z) + fabsf(t.m[1][3]));
    float zs = (fabsf(t.m[2][0] * x) + fabsf(t.m[2][1] * y) + fabsf(t.m[2][2] * z) + fabsf(t.m[2][3]));
    *pError = gammaf(3) * v3(xs, ys, zs);
    if (wp == 1) return v3(xp, yp, zp);
    return v3(xp, yp, zp) / wp;
}
// transform.h:299-331
PG_HD V3 xformPointErr2(const M44 &t, V3 p, V3 pe, V3 *absError) {
    float x = p.x, y = p.y, z = p.z;
    float xp = (t.m[0][0] * x + t.m[0][1] * y) + (t.m[0][2] * z + t.m[0][3]);
    float yp = (t.m[1][0] * x + t.m[1][1] * y) + (t.m[1][2] * z + t.m[1][3]);
    float zp = (t.m[2][0] * x + t.m[2][1] * y) + (t.m[2][2] * z + t.m[2][3]);
    float wp = (t.m[3][0] * x + t.m[3][1] * y) + (t.m[3][2] * z + t.m[3][3]);
    absError->x = (gammaf(3) + 1.f) * (fabsf(t.m[0][0]) * pe.x + fabsf(t.m[0][1]) * pe.y + fabsf(t.m[0][2]) * pe.z) +
                  gammaf(3) * (fabsf(t.m[0][0] * x) + fabsf(t.m[0][1] * y) + fabsf(t.m[0][2] * z) + fabsf(t.m[0][3]));
    absError->y = (gammaf(3) + 1.f) * (fabsf(t.m[1][0]) * pe.x + fabsf(t.m[1][1]) * pe.y + fabsf(t.m[1][2]) * pe.z) +
                  gammaf(3) * (fabsf(t.m[1][0] * x) + fabsf(t.m[1][1] * y) + fabsf(t.m[1][2] * z) + fabsf(t.m[1][3]));
    absError->z = (gammaf(3) + 1.f) * (fabsf(t.m[2][0]) * pe.x + fabsf(t.m[2][1]) * pe.y + fabsf(t.m[2][2]) * pe.z) +
                  gammaf(3) * (fabsf(t.m[2][0] * x) + fabsf(t.m[2][1] * y) + fabsf(t.m[2][2] * z) + fabsf(t.m[2][3]));
    if (wp == 1.f) return v3(xp, yp, zp);
    return v3(xp, yp, zp) / wp;
}
// transform.h:333-349
PG_HD V3 xformVectorErr(const M44 &t, V3 v, V3 *absError) {
    float x = v.x, y = v.y, z = v.z;
    absError->x = gammaf(3) * (fabsf(t.m[0][0] * x) + fabsf(t.m[0][1] * y) + fabsf(t.m[0][2] * z));
    absError->y = gammaf(3) * (fabsf(t.m[1][0] * x) + fabsf(t.m[1][1] * y) + fabsf(t.m[1][2] * z));
    absError->z = gammaf(3) * (fabsf(t.m[2][0] * x) + fabsf(t.m[2][1] * y) + fabsf(t.m[2][2] * z));
    return xformVector(t, v);
}

}  // namespace pg

#endif  // PG_MATH_H
