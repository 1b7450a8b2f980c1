// Ray / primitive tests of the BVH leaf loop.  Each is the arithmetic of the
// reference routine it cites, operation for operation (no FMA contraction:
// this file is compiled with -fmad=false), recording only (t, b0, b1); the
// hit record is rebuilt once per path vertex in shading (see interaction.h).
#ifndef PG_DEVICE_SHAPES_H
#define PG_DEVICE_SHAPES_H

#include "../common/pg_math.h"

namespace pg {

struct Ray {
    V3 o, d;
    float tMax;
};

// The shear that takes the ray to +z depends on the ray alone; the traversal
// computes it once per ray instead of once per triangle (same values).
struct RayShear {
    int kx, ky, kz;
    float Sx, Sy, Sz;
};
PG_HD RayShear rayShear(V3 rd) {
    RayShear s;
    s.kz = maxDimension(vabs(rd));
    s.kx = s.kz + 1;
    if (s.kx == 3) s.kx = 0;
    s.ky = s.kx + 1;
    if (s.ky == 3) s.ky = 0;
    V3 d = rotateToZ(rd, s.kz);
    s.Sx = -d.x / d.z, s.Sy = -d.y / d.z, s.Sz = 1.f / d.z;
    return s;
}

// ---- Triangle::Intersect / IntersectP test part, shapes/triangle.cpp:199-292
// (identical in 427-520).  Returns true when the hit is accepted for
// ray.tMax; does not update tMax.
PG_HD bool triangleTestSheared(const Ray &ray, const RayShear &sh, V3 p0, V3 p1, V3 p2, float *tOut, float *b0Out, float *b1Out,
                               float *b2Out = nullptr) {
    V3 p0t = p0 - ray.o, p1t = p1 - ray.o, p2t = p2 - ray.o;
    p0t = rotateToZ(p0t, sh.kz);  // = permute(p, kx, ky, kz)
    p1t = rotateToZ(p1t, sh.kz);
    p2t = rotateToZ(p2t, sh.kz);
    float Sx = sh.Sx, Sy = sh.Sy, Sz = sh.Sz;
    p0t.x += Sx * p0t.z;
    p0t.y += Sy * p0t.z;
    p1t.x += Sx * p1t.z;
    p1t.y += Sy * p1t.z;
    p2t.x += Sx * p2t.z;
    p2t.y += Sy * p2t.z;
    float e0 = p1t.x * p2t.y - p1t.y * p2t.x;
    float e1 = p2t.x * p0t.y - p2t.y * p0t.x;
    float e2 = p0t.x * p1t.y - p0t.y * p1t.x;
    if (e0 == 0.0f || e1 == 0.0f || e2 == 0.0f) {  // double fallback on edges, triangle.cpp:234-246
        double p2txp1ty = (double)p2t.x * (double)p1t.y;
        double p2typ1tx = (double)p2t.y * (double)p1t.x;
        e0 = (float)(p2typ1tx - p2txp1ty);
        double p0txp2ty = (double)p0t.x * (double)p2t.y;
        double p0typ2tx = (double)p0t.y * (double)p2t.x;
        e1 = (float)(p0typ2tx - p0txp2ty);
        double p1txp0ty = (double)p1t.x * (double)p0t.y;
        double p1typ0tx = (double)p1t.y * (double)p0t.x;
        e2 = (float)(p1typ0tx - p1txp0ty);
    }
    if ((e0 < 0 || e1 < 0 || e2 < 0) && (e0 > 0 || e1 > 0 || e2 > 0)) return false;
    float det = e0 + e1 + e2;
    if (det == 0) return false;
    p0t.z *= Sz;
    p1t.z *= Sz;
    p2t.z *= Sz;
    float tScaled = e0 * p0t.z + e1 * p1t.z + e2 * p2t.z;
    if (det < 0 && (tScaled >= 0 || tScaled < ray.tMax * det)) return false;
    else if (det > 0 && (tScaled <= 0 || tScaled > ray.tMax * det)) return false;
    float invDet = 1 / det;
    float b0 = e0 * invDet, b1 = e1 * invDet, b2 = e2 * invDet;
    float t = tScaled * invDet;
    // conservative t > 0 check, triangle.cpp:271-292
    float maxZt = maxComponent(vabs(v3(p0t.z, p1t.z, p2t.z)));
    float deltaZ = gammaf(3) * maxZt;
    float maxXt = maxComponent(vabs(v3(p0t.x, p1t.x, p2t.x)));
    float maxYt = maxComponent(vabs(v3(p0t.y, p1t.y, p2t.y)));
    float deltaX = gammaf(5) * (maxXt + maxZt);
    float deltaY = gammaf(5) * (maxYt + maxZt);
    float deltaE = 2 * (gammaf(2) * maxXt * maxYt + deltaY * maxXt + deltaX * maxYt);
    float maxE = maxComponent(vabs(v3(e0, e1, e2)));
    float deltaT = 3 * (gammaf(3) * maxE * maxZt + deltaE * maxZt + deltaZ * maxE) * fabsf(invDet);
    if (t <= deltaT) return false;
    *tOut = t;
    *b0Out = b0;
    *b1Out = b1;
    if (b2Out) *b2Out = b2;
    return true;
}

PG_HD bool triangleTest(const Ray &ray, V3 p0, V3 p1, V3 p2, float *tOut, float *b0Out, float *b1Out, float *b2Out = nullptr) {
    return triangleTestSheared(ray, rayShear(ray.d), p0, p1, p2, tOut, b0Out, b1Out, b2Out);
}

// ---- Ray to object space with error bounds, transform.h:402-414
PG_HD Ray xformRayErr(const M44 &w2o, const Ray &r, V3 *oErr, V3 *dErr) {
    Ray out;
    out.o = xformPointErr(w2o, r.o, oErr);
    out.d = xformVectorErr(w2o, r.d, dErr);
    out.tMax = r.tMax;
    float lengthSquared_ = lengthSquared(out.d);
    if (lengthSquared_ > 0) {
        float dt = dot(vabs(out.d), *oErr) / lengthSquared_;
        out.o += out.d * dt;
    }
    return out;
}

struct PlaneParams {  // AAPlaneShape, shapes/plane.h:12-40
    V3 lo, hi;
    int ax, ax0, ax1;
    bool facingFw;
    float area;
};
PG_HD PlaneParams loadPlane(const float *rec) {
    PlaneParams p;
    p.lo = v3(rec[0], rec[1], rec[2]);
    p.hi = v3(rec[3], rec[4], rec[5]);
    p.ax = (int)rec[6];
    p.ax0 = p.ax == 2 ? 0 : (p.ax == 0 ? 1 : 2);
    p.ax1 = p.ax == 2 ? 1 : (p.ax == 0 ? 2 : 0);
    p.facingFw = rec[7] != 0.f;
    p.area = rec[8];
    return p;
}
PG_HD M44 loadM44(const float *m) {
    M44 r;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) r.m[i][j] = m[4 * i + j];
    return r;
}

// ---- AAPlaneShape::Intersect test part, shapes/plane.cpp:15-31.  Note: no
// t > 0 test (hits behind the origin are accepted) and strict bounds.
PG_HD_NOINLINE bool aaplaneTest(const Ray &ray, const float *rec, float *tOut, V3 *pHitObj) {
    PlaneParams pl = loadPlane(rec);
    M44 w2o = loadM44(rec + 32);
    V3 oErr, dErr;
    Ray rayT = xformRayErr(w2o, ray, &oErr, &dErr);
    float t = (pl.lo[pl.ax] - rayT.o[pl.ax]) / rayT.d[pl.ax];
    V3 pHit = rayT.o + t * rayT.d;
    if (pHit[pl.ax0] > pl.lo[pl.ax0] && pHit[pl.ax0] < pl.hi[pl.ax0] && pHit[pl.ax1] > pl.lo[pl.ax1] &&
        pHit[pl.ax1] < pl.hi[pl.ax1] && t < rayT.tMax) {
        *tOut = t;
        if (pHitObj) *pHitObj = pHit;
        return true;
    }
    return false;
}

// ---- EFloat, core/efloat.h (release build: value + interval, no long double)
struct EFloat {
    float v, low, high;
};
PG_HD EFloat ef(float v, float err = 0.f) {
    EFloat r;
    r.v = v;
    if (err == 0.f) r.low = r.high = v;
    else {
        r.low = nextFloatDown(v - err);
        r.high = nextFloatUp(v + err);
    }
    return r;
}
PG_HD EFloat operator+(EFloat a, EFloat b) {
    EFloat r;
    r.v = a.v + b.v;
    r.low = nextFloatDown(a.low + b.low);
    r.high = nextFloatUp(a.high + b.high);
    return r;
}
PG_HD EFloat operator-(EFloat a, EFloat b) {
    EFloat r;
    r.v = a.v - b.v;
    r.low = nextFloatDown(a.low - b.high);
    r.high = nextFloatUp(a.high - b.low);
    return r;
}
PG_HD EFloat operator*(EFloat a, EFloat b) {
    EFloat r;
    r.v = a.v * b.v;
    float p0 = a.low * b.low, p1 = a.high * b.low, p2 = a.low * b.high, p3 = a.high * b.high;
    r.low = nextFloatDown(minf(minf(p0, p1), minf(p2, p3)));
    r.high = nextFloatUp(maxf(maxf(p0, p1), maxf(p2, p3)));
    return r;
}
PG_HD EFloat operator/(EFloat a, EFloat b) {
    EFloat r;
    r.v = a.v / b.v;
    if (b.low < 0 && b.high > 0) {
        r.low = -kInfinity;
        r.high = kInfinity;
    } else {
        float d0 = a.low / b.low, d1 = a.high / b.low, d2 = a.low / b.high, d3 = a.high / b.high;
        r.low = nextFloatDown(minf(minf(d0, d1), minf(d2, d3)));
        r.high = nextFloatUp(maxf(maxf(d0, d1), maxf(d2, d3)));
    }
    return r;
}
// efloat.h:257-272
PG_HD bool efQuadratic(EFloat A, EFloat B, EFloat C, EFloat *t0, EFloat *t1) {
    double discrim = (double)B.v * (double)B.v - 4. * (double)A.v * (double)C.v;
    if (discrim < 0.) return false;
    double rootDiscrim = sqrt(discrim);
    EFloat floatRootDiscrim = ef((float)rootDiscrim, (float)(kMachineEpsilon * rootDiscrim));
    EFloat q;
    if (B.v < 0) q = ef(-.5f) * (B - floatRootDiscrim);
    else q = ef(-.5f) * (B + floatRootDiscrim);
    *t0 = q / A;
    *t1 = C / q;
    if (t0->v > t1->v) {
        EFloat tmp = *t0;
        *t0 = *t1;
        *t1 = tmp;
    }
    return true;
}

struct SphereParams {
    float radius, zMin, zMax, thetaMin, thetaMax, phiMax;
};
PG_HD SphereParams loadSphere(const float *rec) {
    SphereParams s;
    s.radius = rec[0], s.zMin = rec[1], s.zMax = rec[2], s.thetaMin = rec[3], s.thetaMax = rec[4], s.phiMax = rec[5];
    return s;
}

// ---- Sphere::Intersect / IntersectP test part, shapes/sphere.cpp:49-104 and
// 158-215.  Outputs t, the refined object-space hit point and phi.
PG_HD_NOINLINE bool sphereTest(const Ray &r, const float *rec, float *tOut, V3 *pHitObj, float *phiOut, Ray *rayObj) {
    SphereParams sp = loadSphere(rec);
    M44 w2o = loadM44(rec + 32);
    V3 oErr, dErr;
    Ray ray = xformRayErr(w2o, r, &oErr, &dErr);
    EFloat ox = ef(ray.o.x, oErr.x), oy = ef(ray.o.y, oErr.y), oz = ef(ray.o.z, oErr.z);
    EFloat dx = ef(ray.d.x, dErr.x), dy = ef(ray.d.y, dErr.y), dz = ef(ray.d.z, dErr.z);
    EFloat a = dx * dx + dy * dy + dz * dz;
    EFloat b = ef(2.f) * (dx * ox + dy * oy + dz * oz);
    EFloat c = ox * ox + oy * oy + oz * oz - ef(sp.radius) * ef(sp.radius);
    EFloat t0, t1;
    if (!efQuadratic(a, b, c, &t0, &t1)) return false;
    if (t0.high > ray.tMax || t1.low <= 0) return false;
    EFloat tShapeHit = t0;
    if (tShapeHit.low <= 0) {
        tShapeHit = t1;
        if (tShapeHit.high > ray.tMax) return false;
    }
    V3 pHit = ray.o + ray.d * tShapeHit.v;
    pHit *= sp.radius / distance(pHit, v3(0, 0, 0));
    if (pHit.x == 0 && pHit.y == 0) pHit.x = 1e-5f * sp.radius;
    float phi = atan2f(pHit.y, pHit.x);
    if (phi < 0) phi += 2 * kPi;
    if ((sp.zMin > -sp.radius && pHit.z < sp.zMin) || (sp.zMax < sp.radius && pHit.z > sp.zMax) || phi > sp.phiMax) {
        if (tShapeHit.v == t1.v) return false;
        if (t1.high > ray.tMax) return false;
        tShapeHit = t1;
        pHit = ray.o + ray.d * tShapeHit.v;
        pHit *= sp.radius / distance(pHit, v3(0, 0, 0));
        if (pHit.x == 0 && pHit.y == 0) pHit.x = 1e-5f * sp.radius;
        phi = atan2f(pHit.y, pHit.x);
        if (phi < 0) phi += 2 * kPi;
        if ((sp.zMin > -sp.radius && pHit.z < sp.zMin) || (sp.zMax < sp.radius && pHit.z > sp.zMax) || phi > sp.phiMax)
            return false;
    }
    *tOut = tShapeHit.v;
    if (pHitObj) *pHitObj = pHit;
    if (phiOut) *phiOut = phi;
    if (rayObj) *rayObj = ray;
    return true;
}

// Cheap, EXACT rejections in front of sphereTest.  EFloat carries a value v and an interval [low, high] with
// low <= v <= high (every efloat.h operation rounds the interval outwards and computes v from the operands'
// v alone), and sphereTest's early exits are
//     efQuadratic: discrim < 0                (double, from A.v, B.v, C.v only; efloat.h:257-259)
//     t0.high > ray.tMax  ||  t1.low <= 0     (sphere.cpp:70)
// The value parts are plain float expressions of the transformed ray, evaluated here in the same order, so
// "discrim < 0", "t0.v > tMax" (=> t0.high > tMax) and "t1.v <= 0" (=> t1.low <= 0) each imply that
// sphereTest returns false.  A NaN on either side makes a comparison false and sends the ray to the full test.
// Why it matters: a shadow ray aimed at a sampled point of a sphere light ends just short of it
// (tMax = 1 - ShadowEpsilon, t0 ~ 1), so EVERY such ray used to run the ~1100-instruction interval version
// only to be turned away by `t0.high > tMax` (ncu r01g: a quarter of k_ptrace_fused's instructions).
PG_HD bool sphereRejects(const Ray &r, const float *rec) {
    float radius = rec[0];
    M44 w2o = loadM44(rec + 32);
    V3 oErr, dErr;
    Ray ray = xformRayErr(w2o, r, &oErr, &dErr);
    float a = ray.d.x * ray.d.x + ray.d.y * ray.d.y + ray.d.z * ray.d.z;
    float b = 2.f * (ray.d.x * ray.o.x + ray.d.y * ray.o.y + ray.d.z * ray.o.z);
    float c = ray.o.x * ray.o.x + ray.o.y * ray.o.y + ray.o.z * ray.o.z - radius * radius;
    double discrim = (double)b * (double)b - 4. * (double)a * (double)c;
    if (discrim < 0.) return true;
    float root = (float)sqrt(discrim);
    float q = (b < 0) ? -.5f * (b - root) : -.5f * (b + root);
    float t0 = q / a, t1 = c / q;
    if (t0 > t1) {
        float tmp = t0;
        t0 = t1;
        t1 = tmp;
    }
    return t0 > ray.tMax || t1 <= 0;
}
PG_HD bool sphereTestGuarded(const Ray &r, const float *rec, float *tOut) {
    if (sphereRejects(r, rec)) return false;
    return sphereTest(r, rec, tOut, nullptr, nullptr, nullptr);
}

}  // namespace pg
#endif
