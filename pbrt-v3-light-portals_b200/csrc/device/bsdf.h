// BSDF evaluation and sampling as a flat tagged record (no virtual calls, no
// arena): at most two lobes per material on this path.
//   BSDF::f / Sample_f / Pdf / NumComponents     core/reflection.cpp:713-829
//   BxDF::Sample_f / Pdf (cosine)                core/reflection.cpp:416-427
//   LambertianReflection, OrenNayar              core/reflection.cpp:211-252, reflection.h:472-478
//   MicrofacetReflection / MicrofacetTransmission   core/reflection.cpp:259-304, 443-491
//   FresnelSpecular, SpecularReflection          core/reflection.cpp:136-143, 520-554
//   FrDielectric, Reflect, Refract               core/reflection.cpp:47-68, reflection.h:93-109
//   TrowbridgeReitzDistribution                  core/microfacet.cpp:155-163, 176-184, 238-344
//   materials: matte.cpp:45-62, plastic.cpp:45-70, glass.cpp:45-92,
//              dispersive_glass.cpp:48-123, mirror.cpp:45-55
#ifndef PG_DEVICE_BSDF_H
#define PG_DEVICE_BSDF_H

#include "camera.h"
#include "interaction.h"

namespace pg {

enum {  // BxDFType, reflection.h:119-127
    BSDF_REFLECTION = 1 << 0,
    BSDF_TRANSMISSION = 1 << 1,
    BSDF_DIFFUSE = 1 << 2,
    BSDF_GLOSSY = 1 << 3,
    BSDF_SPECULAR = 1 << 4,
    BSDF_ALL = 31
};

enum LobeKind { LOBE_LAMBERT, LOBE_OREN_NAYAR, LOBE_MICRO_REFL, LOBE_MICRO_TRANS, LOBE_FRESNEL_SPECULAR, LOBE_SPEC_REFL_NOOP,
                LOBE_MICRO_REFL_COND /* MicrofacetReflection with FresnelConductor(1, T, K): metal */,
                LOBE_SPEC_REFL_DIEL /* SpecularReflection with FresnelDielectric(etaA, etaB) */,
                LOBE_SPEC_TRANS /* SpecularTransmission(R, etaA, etaB, Radiance) */ };

struct Lobe {
    int kind;
    int type;      // BxDFType bits
    Spec R;        // R (reflection lobes) or T (MICRO_TRANS); FRESNEL_SPECULAR keeps T in T
    Spec T;
    Spec K;        // conductor absorption (metal)
    float a, b;    // OrenNayar A,B | alphax, alphay
    float etaA, etaB;  // dielectric Fresnel (etaI, etaT) / transmission indices
};

constexpr int kMaxLobes = 5;  // UberMaterial adds up to five BxDFs (uber.cpp:45-105)

struct MatLobes {  // what ComputeScatteringFunctions adds for one material (bsdf_eval.h:makeMaterialLobes)
    int nLobes;
    float eta;       // BSDF::eta
    int dispersive;  // eta (and every lobe's etaT) follows the path's wavelength
    int shadeClass;  // 0: one Lambertian lobe; 2: every lobe is specular (no light sampling: integrator.cpp:109-111 finds no
                     // non-specular component); 1: anything else -- the hit queue is split by it (render.cu:splitHits)
    Lobe lobes[kMaxLobes];
};

struct BSDF {
    V3 ns, ng, ss, ts;
    float eta;
    float etaOverride;  // 0, or etaT of all lobes (dispersive_glass)
    int nLobes;
    // pointer into the material table, kept as two words so that the record is 4-byte aligned and an odd
    // number of words long with its neighbours (k_shade_sync keeps one per thread in shared memory)
    unsigned lobesLo, lobesHi;
    PG_HD const Lobe *lobePtr() const { return (const Lobe *)(((unsigned long long)lobesHi << 32) | lobesLo); }
    PG_HD void setLobes(const Lobe *p) {
        unsigned long long v = (unsigned long long)p;
        lobesLo = (unsigned)v, lobesHi = (unsigned)(v >> 32);
    }
};

// ---- local-frame trigonometry, reflection.h:52-92
PG_HD float cosTheta(V3 w) { return w.z; }
PG_HD float cos2Theta(V3 w) { return w.z * w.z; }
PG_HD float absCosTheta(V3 w) { return fabsf(w.z); }
PG_HD float sin2Theta(V3 w) { return maxf(0.f, 1.f - cos2Theta(w)); }
PG_HD float sinTheta(V3 w) { return sqrtf(sin2Theta(w)); }
PG_HD float tanTheta(V3 w) { return sinTheta(w) / cosTheta(w); }
PG_HD float tan2Theta(V3 w) { return sin2Theta(w) / cos2Theta(w); }
PG_HD float cosPhi(V3 w) {
    float s = sinTheta(w);
    return (s == 0) ? 1 : clampf(w.x / s, -1, 1);
}
PG_HD float sinPhi(V3 w) {
    float s = sinTheta(w);
    return (s == 0) ? 0 : clampf(w.y / s, -1, 1);
}
PG_HD float cos2Phi(V3 w) { return cosPhi(w) * cosPhi(w); }
PG_HD float sin2Phi(V3 w) { return sinPhi(w) * sinPhi(w); }
PG_HD bool sameHemisphere(V3 w, V3 wp) { return w.z * wp.z > 0; }
PG_HD V3 reflectAbout(V3 wo, V3 n) { return -wo + 2 * dot(wo, n) * n; }
PG_HD bool refract(V3 wi, V3 n, float eta, V3 *wt) {
    float cosThetaI = dot(n, wi);
    float sin2ThetaI = maxf(0.f, 1 - cosThetaI * cosThetaI);
    float sin2ThetaT = eta * eta * sin2ThetaI;
    if (sin2ThetaT >= 1) return false;
    float cosThetaT = sqrtf(1 - sin2ThetaT);
    *wt = eta * -wi + (eta * cosThetaI - cosThetaT) * n;
    return true;
}

PG_HD float frDielectric(float cosThetaI, float etaI, float etaT) {
    cosThetaI = clampf(cosThetaI, -1, 1);
    bool entering = cosThetaI > 0.f;
    if (!entering) {
        float tmp = etaI;
        etaI = etaT;
        etaT = tmp;
        cosThetaI = fabsf(cosThetaI);
    }
    float sinThetaI = sqrtf(maxf(0.f, 1 - cosThetaI * cosThetaI));
    float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1) return 1;
    float cosThetaT = sqrtf(maxf(0.f, 1 - sinThetaT * sinThetaT));
    float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (Rparl * Rparl + Rperp * Rperp) / 2;
}

PG_HD Spec specSqrt(Spec s) { return spec(sqrtf(s.r), sqrtf(s.g), sqrtf(s.b)); }
PG_HD Spec operator+(Spec a, float b) { return spec(a.r + b, a.g + b, a.b + b); }
PG_HD Spec operator-(Spec a, float b) { return spec(a.r - b, a.g - b, a.b - b); }
// FrConductor, reflection.cpp:71-94
PG_HD Spec frConductor(float cosThetaI, Spec etai, Spec etat, Spec k) {
    cosThetaI = clampf(cosThetaI, -1, 1);
    Spec eta = etat / etai;
    Spec etak = k / etai;
    float cosThetaI2 = cosThetaI * cosThetaI;
    float sinThetaI2 = (float)(1. - (double)cosThetaI2);
    Spec eta2 = eta * eta;
    Spec etak2 = etak * etak;
    Spec t0 = eta2 - etak2 - sinThetaI2;
    Spec a2plusb2 = specSqrt(t0 * t0 + 4 * eta2 * etak2);
    Spec t1 = a2plusb2 + cosThetaI2;
    Spec a = specSqrt(0.5f * (a2plusb2 + t0));
    Spec t2 = (float)2 * cosThetaI * a;
    Spec Rs = (t1 - t2) / (t1 + t2);
    Spec t3 = cosThetaI2 * a2plusb2 + sinThetaI2 * sinThetaI2;
    Spec t4 = t2 * sinThetaI2;
    Spec Rp = Rs * (t3 - t4) / (t3 + t4);
    return 0.5f * (Rp + Rs);
}

// ---- TrowbridgeReitzDistribution (sampleVisibleArea = true, the default)
PG_HD float trD(float ax, float ay, V3 wh) {
    float t2 = tan2Theta(wh);
    if (isInfF(t2)) return 0.f;
    const float cos4Theta = cos2Theta(wh) * cos2Theta(wh);
    float e = (cos2Phi(wh) / (ax * ax) + sin2Phi(wh) / (ay * ay)) * t2;
    return 1 / (kPi * ax * ay * cos4Theta * (1 + e) * (1 + e));
}
PG_HD float trLambda(float ax, float ay, V3 w) {
    float absTanTheta = fabsf(tanTheta(w));
    if (isInfF(absTanTheta)) return 0.f;
    float alpha = sqrtf(cos2Phi(w) * ax * ax + sin2Phi(w) * ay * ay);
    float alpha2Tan2Theta = (alpha * absTanTheta) * (alpha * absTanTheta);
    return (-1 + sqrtf(1.f + alpha2Tan2Theta)) / 2;
}
PG_HD float trG1(float ax, float ay, V3 w) { return 1 / (1 + trLambda(ax, ay, w)); }
PG_HD float trG(float ax, float ay, V3 wo, V3 wi) { return 1 / (1 + trLambda(ax, ay, wo) + trLambda(ax, ay, wi)); }
PG_HD float trPdf(float ax, float ay, V3 wo, V3 wh) {  // MicrofacetDistribution::Pdf
    return trD(ax, ay, wh) * trG1(ax, ay, wo) * absDot(wo, wh) / absCosTheta(wo);
}
PG_HD void trSample11(float cosThetaV, float U1, float U2, float *slope_x, float *slope_y) {
    if (cosThetaV > .9999) {
        float r = sqrtf(U1 / (1 - U1));
        float phi = (float)(6.28318530718 * (double)U2);
        *slope_x = r * cosf(phi);
        *slope_y = r * sinf(phi);
        return;
    }
    float sinThetaV = sqrtf(maxf(0.f, 1.f - cosThetaV * cosThetaV));
    float tanThetaV = sinThetaV / cosThetaV;
    float a = 1 / tanThetaV;
    float G1 = 2 / (1 + sqrtf(1.f + 1.f / (a * a)));
    float A = 2 * U1 / G1 - 1;
    float tmp = 1.f / (A * A - 1.f);
    if (tmp > 1e10) tmp = 1e10;
    float B = tanThetaV;
    float D = sqrtf(maxf(B * B * tmp * tmp - (A * A - B * B) * tmp, 0.f));
    float slope_x_1 = B * tmp - D;
    float slope_x_2 = B * tmp + D;
    *slope_x = (A < 0 || slope_x_2 > 1.f / tanThetaV) ? slope_x_1 : slope_x_2;
    float S;
    if (U2 > 0.5f) {
        S = 1.f;
        U2 = 2.f * (U2 - .5f);
    } else {
        S = -1.f;
        U2 = 2.f * (.5f - U2);
    }
    float z = (U2 * (U2 * (U2 * 0.27385f - 0.73369f) + 0.46341f)) /
              (U2 * (U2 * (U2 * 0.093073f + 0.309420f) - 1.000000f) + 0.597999f);
    *slope_y = S * z * sqrtf(1.f + *slope_x * *slope_x);
}
PG_HD V3 trSampleWh(float ax, float ay, V3 wo, float u0, float u1) {
    bool flip = wo.z < 0;
    V3 wi = flip ? -wo : wo;
    V3 wiStretched = normalize(v3(ax * wi.x, ay * wi.y, wi.z));
    float slope_x, slope_y;
    trSample11(cosTheta(wiStretched), u0, u1, &slope_x, &slope_y);
    float tmp = cosPhi(wiStretched) * slope_x - sinPhi(wiStretched) * slope_y;
    slope_y = sinPhi(wiStretched) * slope_x + cosPhi(wiStretched) * slope_y;
    slope_x = tmp;
    slope_x = ax * slope_x;
    slope_y = ay * slope_y;
    V3 wh = normalize(v3(-slope_x, -slope_y, 1.f));
    if (flip) wh = -wh;
    return wh;
}

PG_HD V3 cosineSampleHemisphere(float u0, float u1) {
    float dx, dy;
    concentricSampleDisk(u0, u1, &dx, &dy);
    float z = sqrtf(maxf(0.f, 1 - dx * dx - dy * dy));
    return v3(dx, dy, z);
}

}  // namespace pg
#include "bsdf_eval.h"
#endif
