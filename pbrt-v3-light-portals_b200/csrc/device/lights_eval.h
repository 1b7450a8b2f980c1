// Second half of lights.h: Shape::Pdf, the lights' Sample_Li / Pdf_Li / Le.
#ifndef PG_DEVICE_LIGHTS_EVAL_H
#define PG_DEVICE_LIGHTS_EVAL_H

namespace pg {

// light_params layout (pbrtgpu.h): [0..2] L or I, [3] twoSided | pos.x ..., [4] area
// infinite: [8..23] LightToWorld, [24..39] WorldToLight, [40..47] the 2x2
// sampling distribution filled in at upload (upload.cu:finishInfiniteLight).
// row v of the 2x2 image at [40 + 4v]: func[0], func[1], cdf[1], funcInt; marginal cdf[1] at [4], funcInt at [5]
enum { kInfRow = 40, kInfMargCdf = 4, kInfMargInt = 5 };

// Shape::Pdf(ref, wi) for one primitive slot (Sphere::Pdf overrides it outside the sphere).
PG_NI_DIRECT float shapePdfRef(const DevBVH &bvh, int slot, V3 refP, V3 refPError, V3 refN, V3 wi) {
    const F4 *tp = bvh.tris + 3 * (size_t)slot;
    F4 a = PG_LDG(tp);
    int tag = f2i(a.w);
    int type = tag & kTagTypeMask;
    Ray ray = spawnRay(refP, refPError, refN, wi);
    V3 pL, nL;
    float area;
    if (type == PBRTGPU_PRIM_TRIANGLE) {
        if (tag & kTagDegenerate) return 0;
        F4 b = PG_LDG(tp + 1), c = PG_LDG(tp + 2);
        V3 p0 = v3(a.x, a.y, a.z), p1 = v3(b.x, b.y, b.z), p2 = v3(c.x, c.y, c.z);
        float t, b0, b1, b2;
        if (!triangleTest(ray, p0, p1, p2, &t, &b0, &b1, &b2)) return 0;
        pL = b0 * p0 + b1 * p1 + b2 * p2;
        nL = normalize(cross(p0 - p2, p1 - p2));  // sign is irrelevant under AbsDot
        area = (float)(0.5 * (double)length(cross(p1 - p0, p2 - p0)));
    } else if (type == PBRTGPU_PRIM_AAPLANE) {
        const float *rec = bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift);
        float t;
        if (!aaplaneTest(ray, rec, &t, nullptr)) return 0;
        SurfInt li = aaplaneInteraction(ray, rec, 0);
        pL = li.p, nL = li.n;
        area = PG_LD(rec + 8);
    } else {
        const float *rec = bvh.sphereParams + 48 * (size_t)(tag >> kTagShapeShift);
        SphereParams sp = loadSphere(rec);
        V3 pCenter = xformPoint(loadM44(rec + 16), v3(0, 0, 0));
        V3 pOrigin = offsetRayOrigin(refP, refPError, refN, pCenter - refP);
        if (!(distanceSquared(pOrigin, pCenter) <= sp.radius * sp.radius)) {
            float sinThetaMax2 = sp.radius * sp.radius / distanceSquared(refP, pCenter);
            float cosThetaMax = sqrtf(maxf(0.f, 1 - sinThetaMax2));
            return 1 / (2 * kPi * (1 - cosThetaMax));  // UniformConePdf
        }
        float t;
        if (!sphereTest(ray, rec, &t, nullptr, nullptr, nullptr)) return 0;
        SurfInt li = sphereInteraction(ray, rec, 0);
        pL = li.p, nL = li.n;
        area = sp.phiMax * sp.radius * (sp.zMax - sp.zMin);
    }
    float pdf = distanceSquared(refP, pL) / (absDot(nL, -wi) * area);
    if (isInfF(pdf)) pdf = 0.f;
    return pdf;
}

// ---- constant InfiniteAreaLight: Lmap is one texel, looked up through MIPMap::triangle
PG_HD Spec infiniteLookup(Spec T, float s0, float s1) {  // mipmap.h:263-274 on a 1x1 level
    float s = s0 * 1 - 0.5f, t = s1 * 1 - 0.5f;
    float fs0 = floorf(s), ft0 = floorf(t);
    float ds = s - fs0, dt = t - ft0;
    return (1 - ds) * (1 - dt) * T + (1 - ds) * dt * T + ds * (1 - dt) * T + ds * dt * T;
}
PG_HD float sphericalTheta(V3 v) { return acosf(clampf(v.z, -1, 1)); }
PG_HD float sphericalPhi(V3 v) {
    float p = atan2f(v.y, v.x);
    return (p < 0) ? (p + 2 * kPi) : p;
}
PG_HD Spec infiniteLe(const float *lp, V3 rayD) {  // InfiniteAreaLight::Le, infinite.cpp:92-96
    V3 w = normalize(xformVector(loadM44(lp + 24), rayD));
    return infiniteLookup(loadSpec(lp), sphericalPhi(w) * kInv2Pi, sphericalTheta(w) * kInvPi);
}
// Distribution1D::SampleContinuous over two cells with cdf = {0, c1, 1}
PG_HD float sampleContinuous2(float f0, float f1, float c1, float funcInt, float u, float *pdf, int *off) {
    // FindInterval(3, cdf[i] <= u) clamped to [0,1]
    int offset = (c1 <= u) ? 1 : 0;
    if (1.f <= u) offset = 1;
    float c0 = offset ? c1 : 0.f, cN = offset ? 1.f : c1;
    *off = offset;
    float du = u - c0;
    if ((cN - c0) > 0) du /= (cN - c0);
    *pdf = (funcInt > 0) ? (offset ? f1 : f0) / funcInt : 0;
    return (offset + du) / 2;
}
PG_HD_NOINLINE Spec infiniteSampleLi(const float *lp, float u0, float u1, V3 *wi, float *pdf, float *uvOut) {
    float pdfs[2];
    int v, dummy;
    // Distribution2D::SampleContinuous: marginal over rows (its func = the rows' funcInt), then the row
    float d1 = sampleContinuous2(PG_LD(lp + kInfRow + 3), PG_LD(lp + kInfRow + 7), PG_LD(lp + kInfMargCdf), PG_LD(lp + kInfMargInt), u1,
                                 &pdfs[1], &v);
    const float *row = lp + kInfRow + 4 * v;
    float d0 = sampleContinuous2(PG_LD(row), PG_LD(row + 1), PG_LD(row + 2), PG_LD(row + 3), u0, &pdfs[0], &dummy);
    float mapPdf = pdfs[0] * pdfs[1];
    *pdf = 0;
    if (mapPdf == 0) return spec(0.f);
    float theta = d1 * kPi, phi = d0 * 2 * kPi;
    float cosT = cosf(theta), sinT = sinf(theta);
    float sinP = sinf(phi), cosP = cosf(phi);
    *wi = xformVector(loadM44(lp + 8), v3(sinT * cosP, sinT * sinP, cosT));
    *pdf = mapPdf / (2 * kPi * kPi * sinT);
    if (sinT == 0) *pdf = 0;
    uvOut[0] = d0, uvOut[1] = d1;
    return infiniteLookup(loadSpec(lp), d0, d1);
}
PG_HD float infinitePdfLi(const float *lp, V3 w) {
    V3 wi = xformVector(loadM44(lp + 24), w);
    float theta = sphericalTheta(wi), phi = sphericalPhi(wi);
    float sinT = sinf(theta);
    if (sinT == 0) return 0;
    // Distribution2D::Pdf
    int iu = (int)(phi * kInv2Pi * 2), iv = (int)(theta * kInvPi * 2);
    iu = iu < 0 ? 0 : (iu > 1 ? 1 : iu);
    iv = iv < 0 ? 0 : (iv > 1 ? 1 : iv);
    return (PG_LD(lp + kInfRow + 4 * iv + iu) / PG_LD(lp + kInfMargInt)) / (2 * kPi * kPi * sinT);
}

}  // namespace pg
#endif
