// Light sampling for EstimateDirect.
//   Shape::Sample(ref,u) / Shape::Pdf(ref,wi)      core/shape.cpp:56-91
//   Triangle::Sample, UniformSampleTriangle        shapes/triangle.cpp:584-609, core/sampling.cpp:106-111
//   AAPlaneShape::Sample                           shapes/plane.cpp:57-72
//   Sphere::Sample(u), Sample(ref,u), Pdf(ref,wi)  shapes/sphere.cpp:219-315
//   DiffuseAreaLight::L / Sample_Li / Pdf_Li       lights/diffuse.h:58-60, diffuse.cpp:69-90
//   InfiniteAreaLight (constant, no map)           lights/infinite.cpp:43-132
//   PointLight / DistantLight Sample_Li            lights/point.cpp:45-55, distant.cpp:52-63
//   Distribution1D / Distribution2D                core/sampling.h:55-148
#ifndef PG_DEVICE_LIGHTS_H
#define PG_DEVICE_LIGHTS_H

#include "bsdf.h"

namespace pg {

struct DevScene;

struct ShapeSample {  // the Interaction a shape's Sample() returns
    V3 p, pError, n;
};

// ---- area sampling: Shape::Sample(u, &pdf)
PG_HD ShapeSample triangleSampleArea(const F4 *tri, int flags, const float *nrm, float u0, float u1, float *pdf) {
    F4 a = PG_LDG(tri), bq = PG_LDG(tri + 1), c = PG_LDG(tri + 2);
    V3 p0 = v3(a.x, a.y, a.z), p1 = v3(bq.x, bq.y, bq.z), p2 = v3(c.x, c.y, c.z);
    float su0 = sqrtf(u0);  // UniformSampleTriangle
    float b0 = 1 - su0, b1 = u1 * su0;
    ShapeSample it;
    it.p = b0 * p0 + b1 * p1 + (1 - b0 - b1) * p2;
    it.n = normalize(cross(p1 - p0, p2 - p0));
    if (flags & PBRTGPU_PF_HAS_NORMALS) {
        V3 ns = b0 * loadV3(nrm) + b1 * loadV3(nrm + 3) + (1 - b0 - b1) * loadV3(nrm + 6);
        it.n = faceforward(it.n, ns);
    } else if (((flags & PBRTGPU_PF_REVERSE_ORIENTATION) != 0) != ((flags & PBRTGPU_PF_SWAPS_HANDEDNESS) != 0))
        it.n *= -1.f;
    V3 pAbsSum = vabs(b0 * p0) + vabs(b1 * p1) + vabs((1 - b0 - b1) * p2);
    it.pError = gammaf(6) * pAbsSum;
    float area = (float)(0.5 * (double)length(cross(p1 - p0, p2 - p0)));  // Triangle::Area: 0.5 is a double literal
    *pdf = 1 / area;
    return it;
}

PG_HD V3 aaplaneNormal(const PlaneParams &pl) {  // AAPlaneShape::Normal, plane.cpp:74-83
    V3 n = v3(0, 0, 0);
    n.set(pl.ax, 1.f);
    if (!pl.facingFw) n *= -1.f;
    return n;
}

PG_HD ShapeSample aaplaneSampleArea(const float *rec, float u0, float u1, float *pdf) {
    PlaneParams pl = loadPlane(rec);
    M44 o2w = loadM44(rec + 16);
    V3 loW = xformPoint(o2w, pl.lo), hiW = xformPoint(o2w, pl.hi);
    ShapeSample it;
    it.p = v3(0, 0, 0);
    it.p.set(pl.ax, loW[pl.ax]);
    it.p.set(pl.ax0, loW[pl.ax0] + (hiW[pl.ax0] - loW[pl.ax0]) * u0);
    it.p.set(pl.ax1, loW[pl.ax1] + (hiW[pl.ax1] - loW[pl.ax1]) * u1);
    it.n = aaplaneNormal(pl);
    it.pError = v3(0.1f, 0.1f, 0.1f);
    *pdf = 1 / pl.area;
    return it;
}

PG_HD V3 uniformSampleSphere(float u0, float u1) {
    float z = 1 - 2 * u0;
    float r = sqrtf(maxf(0.f, 1.f - z * z));
    float phi = 2 * kPi * u1;
    return v3(r * cosf(phi), r * sinf(phi), z);
}

PG_HD ShapeSample sphereSampleArea(const float *rec, int flags, float u0, float u1, float *pdf) {
    SphereParams sp = loadSphere(rec);
    M44 o2w = loadM44(rec + 16), w2o = loadM44(rec + 32);
    V3 pObj = v3(0, 0, 0) + sp.radius * uniformSampleSphere(u0, u1);
    ShapeSample it;
    it.n = normalize(xformNormal(w2o, pObj));
    if (flags & PBRTGPU_PF_REVERSE_ORIENTATION) it.n *= -1.f;
    pObj *= sp.radius / distance(pObj, v3(0, 0, 0));
    V3 pObjError = gammaf(5) * vabs(pObj);
    it.p = xformPointErr2(o2w, pObj, pObjError, &it.pError);
    *pdf = 1 / (sp.phiMax * sp.radius * (sp.zMax - sp.zMin));
    return it;
}

// SphericalDirection(sinTheta, cosTheta, phi, x, y, z), geometry.h:1629-1634
PG_HD V3 sphericalDirection(float sinT, float cosT, float phi, V3 x, V3 y, V3 z) {
    return sinT * cosf(phi) * x + sinT * sinf(phi) * y + cosT * z;
}

// Shape::Sample(ref, u, &pdf): area sample converted to solid angle; Sphere overrides it.
PG_NI_DIRECT ShapeSample shapeSampleRef(const DevBVH &bvh, int slot, int flags, const float *triN, V3 refP, V3 refPError, V3 refN,
                                 float u0, float u1, float *pdf) {
    const F4 *tp = bvh.tris + 3 * (size_t)slot;
    int tag = f2i(PG_LDG(tp).w);
    int type = tag & kTagTypeMask;
    ShapeSample intr;
    if (type == PBRTGPU_PRIM_SPHERE) {
        const float *rec = bvh.sphereParams + 48 * (size_t)(tag >> kTagShapeShift);
        SphereParams sp = loadSphere(rec);
        V3 pCenter = xformPoint(loadM44(rec + 16), v3(0, 0, 0));
        V3 pOrigin = offsetRayOrigin(refP, refPError, refN, pCenter - refP);
        if (distanceSquared(pOrigin, pCenter) <= sp.radius * sp.radius) {
            intr = sphereSampleArea(rec, flags, u0, u1, pdf);
            V3 wi = intr.p - refP;
            if (lengthSquared(wi) == 0) *pdf = 0;
            else {
                wi = normalize(wi);
                *pdf *= distanceSquared(refP, intr.p) / absDot(intr.n, -wi);
            }
            if (isInfF(*pdf)) *pdf = 0.f;
            return intr;
        }
        float dc = distance(refP, pCenter);
        float invDc = 1 / dc;
        V3 wc = (pCenter - refP) * invDc;
        V3 wcX, wcY;
        coordinateSystem(wc, &wcX, &wcY);
        float sinThetaMax = sp.radius * invDc;
        float sinThetaMax2 = sinThetaMax * sinThetaMax;
        float invSinThetaMax = 1 / sinThetaMax;
        float cosThetaMax = sqrtf(maxf(0.f, 1 - sinThetaMax2));
        float cosT = (cosThetaMax - 1) * u0 + 1;
        float sinTheta2 = 1 - cosT * cosT;
        if (sinThetaMax2 < 0.00068523f) {
            sinTheta2 = sinThetaMax2 * u0;
            cosT = sqrtf(1 - sinTheta2);
        }
        float cosAlpha = sinTheta2 * invSinThetaMax +
                         cosT * sqrtf(maxf(0.f, 1.f - sinTheta2 * invSinThetaMax * invSinThetaMax));
        float sinAlpha = sqrtf(maxf(0.f, 1.f - cosAlpha * cosAlpha));
        float phi = u1 * 2 * kPi;
        V3 nWorld = sphericalDirection(sinAlpha, cosAlpha, phi, -wcX, -wcY, -wc);
        V3 pWorld = pCenter + sp.radius * nWorld;
        intr.p = pWorld;
        intr.pError = gammaf(5) * vabs(pWorld);
        intr.n = nWorld;
        if (flags & PBRTGPU_PF_REVERSE_ORIENTATION) intr.n *= -1.f;
        *pdf = 1 / (2 * kPi * (1 - cosThetaMax));
        return intr;
    }
    if (type == PBRTGPU_PRIM_TRIANGLE) intr = triangleSampleArea(tp, flags, triN ? triN + 9 * (size_t)slot : nullptr, u0, u1, pdf);
    else intr = aaplaneSampleArea(bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift), u0, u1, pdf);
    V3 wi = intr.p - refP;
    if (lengthSquared(wi) == 0) *pdf = 0;
    else {
        wi = normalize(wi);
        *pdf *= distanceSquared(refP, intr.p) / absDot(intr.n, -wi);
        if (isInfF(*pdf)) *pdf = 0.f;
    }
    return intr;
}

}  // namespace pg
#include "lights_eval.h"
#endif
