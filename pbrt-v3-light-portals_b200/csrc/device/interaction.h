// The hit record, rebuilt ONCE per path vertex from (primitive slot, b0, b1)
// instead of once per candidate hit as the reference's leaf loop does.
//   Triangle::Intersect, record part          shapes/triangle.cpp:294-420
//   AAPlaneShape::Intersect, record part      shapes/plane.cpp:32-50
//   Sphere::Intersect, record part            shapes/sphere.cpp:107-152
//   SurfaceInteraction ctor / SetShadingGeometry   core/interaction.cpp:44-88
//   Transform::operator()(SurfaceInteraction) core/transform.cpp:262-297
//   Interaction::SpawnRay / SpawnRayTo        core/interaction.h:66-80
// Only what the BSDF frame, ray spawning and emission need is kept: p, pError,
// n, shading.n, shading.dpdu, wo.  (dndu/dndv and uv feed bump mapping and
// textures, which this path does not have.)
#ifndef PG_DEVICE_INTERACTION_H
#define PG_DEVICE_INTERACTION_H

#include "bvh.h"

namespace pg {

struct SurfInt {
    V3 p, pError;
    V3 n;      // geometric normal (Interaction::n)
    V3 ns;     // shading.n
    V3 dpdus;  // shading.dpdu
    V3 wo;
    int wvlsLost;   // the record went through Transform::operator()(SurfaceInteraction), which drops wvls
};

PG_HD V3 loadV3(const float *p) { return v3(PG_LD(p), PG_LD(p + 1), PG_LD(p + 2)); }

// Triangle hit record.  tri = the slot's three float4 (p0,tag) (p1,-) (p2,-).
PG_NI_INTERACTION SurfInt triangleInteraction(const F4 *tri, int flags, const float *nrm, const float *uvs, const float *tan, float b0,
                                  float b1, float b2, V3 rayD) {
    F4 a = PG_LDG(tri), b = PG_LDG(tri + 1), c = PG_LDG(tri + 2);
    V3 p0 = v3(a.x, a.y, a.z), p1 = v3(b.x, b.y, b.z), p2 = v3(c.x, c.y, c.z);
    SurfInt it;
    it.wvlsLost = false;
    float uv[3][2] = {{0, 0}, {1, 0}, {1, 1}};
    if (flags & PBRTGPU_PF_HAS_UV)
        for (int i = 0; i < 3; ++i) uv[i][0] = PG_LD(uvs + 2 * i), uv[i][1] = PG_LD(uvs + 2 * i + 1);
    float duv02[2] = {uv[0][0] - uv[2][0], uv[0][1] - uv[2][1]}, duv12[2] = {uv[1][0] - uv[2][0], uv[1][1] - uv[2][1]};
    V3 dp02 = p0 - p2, dp12 = p1 - p2;
    float determinant = duv02[0] * duv12[1] - duv02[1] * duv12[0];
    bool degenerateUV = fabsf(determinant) < 1e-8;
    V3 dpdu = v3(0, 0, 0), dpdv = v3(0, 0, 0);
    if (!degenerateUV) {
        float invdet = 1 / determinant;
        dpdu = (duv12[1] * dp02 - duv02[1] * dp12) * invdet;
        dpdv = (-duv12[0] * dp02 + duv02[0] * dp12) * invdet;
    }
    if (degenerateUV || lengthSquared(cross(dpdu, dpdv)) == 0) {
        V3 ng = cross(p2 - p0, p1 - p0);
        coordinateSystem(normalize(ng), &dpdu, &dpdv);  // zero-area triangles never get here (kTagDegenerate)
    }
    float xAbsSum = (fabsf(b0 * p0.x) + fabsf(b1 * p1.x) + fabsf(b2 * p2.x));
    float yAbsSum = (fabsf(b0 * p0.y) + fabsf(b1 * p1.y) + fabsf(b2 * p2.y));
    float zAbsSum = (fabsf(b0 * p0.z) + fabsf(b1 * p1.z) + fabsf(b2 * p2.z));
    it.pError = gammaf(7) * v3(xAbsSum, yAbsSum, zAbsSum);
    it.p = b0 * p0 + b1 * p1 + b2 * p2;
    it.wo = normalize(-rayD);
    it.n = it.ns = normalize(cross(dp02, dp12));
    bool flip = ((flags & PBRTGPU_PF_REVERSE_ORIENTATION) != 0) != ((flags & PBRTGPU_PF_SWAPS_HANDEDNESS) != 0);
    if (flip) it.n = it.ns = -it.n;
    it.dpdus = dpdu;
    if (flags & (PBRTGPU_PF_HAS_NORMALS | PBRTGPU_PF_HAS_TANGENTS)) {
        V3 ns;
        if (flags & PBRTGPU_PF_HAS_NORMALS) {
            ns = b0 * loadV3(nrm) + b1 * loadV3(nrm + 3) + b2 * loadV3(nrm + 6);
            if (lengthSquared(ns) > 0) ns = normalize(ns);
            else ns = it.n;
        } else
            ns = it.n;
        V3 ss;
        if (flags & PBRTGPU_PF_HAS_TANGENTS) {
            ss = b0 * loadV3(tan) + b1 * loadV3(tan + 3) + b2 * loadV3(tan + 6);
            if (lengthSquared(ss) > 0) ss = normalize(ss);
            else ss = normalize(dpdu);
        } else
            ss = normalize(dpdu);
        V3 ts = cross(ss, ns);
        if (lengthSquared(ts) > 0.f) {
            ts = normalize(ts);
            ss = cross(ts, ns);
        } else
            coordinateSystem(ns, &ss, &ts);
        if (flags & PBRTGPU_PF_REVERSE_ORIENTATION) ts = -ts;
        // SetShadingGeometry(ss, ts, ..., orientationIsAuthoritative = true)
        it.ns = normalize(cross(ss, ts));
        it.n = faceforward(it.n, it.ns);
        it.dpdus = ss;
    }
    return it;
}

// Transform::operator()(SurfaceInteraction) for the members kept here.
PG_HD SurfInt transformInteraction(const M44 &o2w, const M44 &w2o, V3 pObj, V3 pErrObj, V3 nObj, V3 nsObj, V3 dpduObj,
                                   V3 woObj) {
    SurfInt it;
    it.p = xformPointErr2(o2w, pObj, pErrObj, &it.pError);
    it.n = normalize(xformNormal(w2o, nObj));
    it.wo = normalize(xformVector(o2w, woObj));
    it.dpdus = xformVector(o2w, dpduObj);
    it.ns = normalize(xformNormal(w2o, nsObj));
    it.ns = faceforward(it.ns, it.n);
    it.wvlsLost = true;
    return it;
}

// AAPlaneShape hit record: hard-coded dpdu/dpdv/pError (plane.cpp:32-50).
PG_NI_INTERACTION SurfInt aaplaneInteraction(const Ray &ray, const float *rec, int flags) {
    float t;
    V3 pHit;
    aaplaneTest(ray, rec, &t, &pHit);  // recompute the object-space hit point; ray.tMax must admit the recorded hit
    V3 dpdu = v3(-1, 0, 0), dpdv = v3(0, 1, 0);
    V3 n = normalize(cross(dpdu, dpdv));
    bool flip = ((flags & PBRTGPU_PF_REVERSE_ORIENTATION) != 0) != ((flags & PBRTGPU_PF_SWAPS_HANDEDNESS) != 0);
    if (flip) n = -n;
    V3 wo = normalize(-ray.d);  // Interaction ctor normalises the WORLD-space -ray.d, then it is transformed
    return transformInteraction(loadM44(rec + 16), loadM44(rec + 32), pHit, v3(0.01f, 0.01f, 0.01f), n, n, dpdu, wo);
}

// Sphere hit record (sphere.cpp:107-152).
PG_NI_INTERACTION SurfInt sphereInteraction(const Ray &ray, const float *rec, int flags) {
    SphereParams sp = loadSphere(rec);
    float t, phi;
    V3 pHit;
    sphereTest(ray, rec, &t, &pHit, &phi, nullptr);
    V3 dpdu = v3(-sp.phiMax * pHit.y, sp.phiMax * pHit.x, 0);
    float theta = acosf(clampf(pHit.z / sp.radius, -1, 1));
    float zRadius = sqrtf(pHit.x * pHit.x + pHit.y * pHit.y);
    float invZRadius = 1 / zRadius;
    float cosPhi = pHit.x * invZRadius, sinPhi = pHit.y * invZRadius;
    V3 dpdv = (sp.thetaMax - sp.thetaMin) * v3(pHit.z * cosPhi, pHit.z * sinPhi, -sp.radius * sinf(theta));
    V3 pError = gammaf(5) * vabs(pHit);
    V3 n = normalize(cross(dpdu, dpdv));
    bool flip = ((flags & PBRTGPU_PF_REVERSE_ORIENTATION) != 0) != ((flags & PBRTGPU_PF_SWAPS_HANDEDNESS) != 0);
    if (flip) n = -n;
    V3 wo = normalize(-ray.d);
    return transformInteraction(loadM44(rec + 16), loadM44(rec + 32), pHit, pError, n, n, dpdu, wo);
}

// Interaction::SpawnRay, interaction.h:66-69
PG_HD Ray spawnRay(V3 p, V3 pError, V3 n, V3 d) {
    Ray r;
    r.o = offsetRayOrigin(p, pError, n, d);
    r.d = d;
    r.tMax = kInfinity;
    return r;
}
// Interaction::SpawnRayTo(const Interaction &), interaction.h:75-80
PG_HD Ray spawnRayTo(V3 p, V3 pError, V3 n, V3 p2, V3 p2Error, V3 n2) {
    Ray r;
    V3 origin = offsetRayOrigin(p, pError, n, p2 - p);
    V3 target = offsetRayOrigin(p2, p2Error, n2, origin - p2);
    r.o = origin;
    r.d = target - origin;
    r.tMax = 1 - kShadowEpsilon;
    return r;
}

}  // namespace pg
#endif
