// Direct lighting at a path vertex, split in two: this file GENERATES the
// next-event rays (what the reference does before it calls Scene::Intersect /
// IntersectP) and states how each is RESOLVED once traced.
//   UniformSampleOneLight                    core/integrator.cpp:100-122
//   EstimateDirect (stock MIS)               core/integrator.cpp:124-258
//   PortalArealight::EstimateDirect{,Light,Portal,Proj}   lights/portal_arealight.cpp:29-239
//   Distribution1D::SampleDiscrete, FindInterval          core/sampling.h:90-101, core/pbrt.h:408-420
#ifndef PG_DEVICE_DIRECT_H
#define PG_DEVICE_DIRECT_H

#include "ctx_scene.h"
#include "portal.h"

namespace pg {

// How a traced next-event ray turns into radiance.
enum NeeMode {
    NEE_MIS_LIGHT = 0,     // closest hit; counts only if the hit primitive carries `light`; a miss counts light->Le(ray) (infinite lights)
    NEE_ANY_EMITTER = 1,   // closest hit; whatever emitter is hit counts (portal estimators, Q6); a miss counts nothing
    NEE_ANY_OR_LIGHT = 2   // as 1, but a miss keeps the sampled light's own L (EstimateDirectLight, portal_arealight.cpp:139-145)
};

struct NeeRequest {
    bool hasShadow;  // any-hit ray: add shadowContrib if unoccluded
    Ray shadowRay;
    Spec shadowContrib;
    bool hasClosest;  // closest-hit ray: add closestContrib * Li(resolved by mode)
    Ray closestRay;
    Spec closestContrib;
    int light, mode;
};

PG_HD int findIntervalCdf(const float *cdf, int size, float u) {  // FindInterval(size, cdf[i] <= u)
    int first = 0, len = size;
    while (len > 0) {
        int half = len >> 1, middle = first + half;
        if (PG_LD(cdf + middle) <= u) {
            first = middle + 1;
            len -= half + 1;
        } else
            len = half;
    }
    int r = first - 1;
    return r < 0 ? 0 : (r > size - 2 ? size - 2 : r);
}

// DiffuseAreaLight::L for light `li` seen from direction w at a point with normal n
PG_HD Spec areaLightL(const DevScene &sc, int li, V3 n, V3 w) {
    const float *lp = sc.lightParams + PBRTGPU_LIGHT_STRIDE * (size_t)li;
    bool twoSided = PG_LD(lp + 3) != 0.f;
    return (twoSided || dot(n, w) > 0) ? loadSpec(lp) : spec(0.f);
}

// Light::Sample_Li.  Returns Li; fills wi, pdf and the point the visibility ray aims at.
PG_NI_DIRECT Spec lightSampleLi(const DevScene &sc, int li, const SurfInt &it, float u0, float u1, V3 *wi, float *pdf, ShapeSample *pL) {
    int tag = PG_LD(sc.lightTag + li);
    const float *lp = sc.lightParams + PBRTGPU_LIGHT_STRIDE * (size_t)li;
    pL->pError = v3(0, 0, 0);
    pL->n = v3(0, 0, 0);
    if (tag == PBRTGPU_LIGHT_DIFFUSE_AREA || tag == PBRTGPU_LIGHT_PORTAL_AREA) {
        int slot = PG_LD(sc.lightPrim + li);
        int flags = sc.primFlags ? PG_LD(sc.primFlags + slot) : 0;
        *pL = shapeSampleRef(sc.bvh, slot, flags, sc.triN, it.p, it.pError, it.n, u0, u1, pdf);
        if (*pdf == 0 || lengthSquared(pL->p - it.p) == 0) {
            *pdf = 0;
            return spec(0.f);
        }
        *wi = normalize(pL->p - it.p);
        return areaLightL(sc, li, pL->n, -*wi);
    }
    if (tag == PBRTGPU_LIGHT_INFINITE) {
        float uv[2];
        Spec L = infiniteSampleLi(lp, u0, u1, wi, pdf, uv);
        pL->p = it.p + *wi * (2 * sc.worldRadius);
        return L;
    }
    if (tag == PBRTGPU_LIGHT_SPOT) {  // SpotLight::Sample_Li + Falloff, spot.cpp:51-73
        V3 pLight = loadV3(lp + 3);
        *wi = normalize(pLight - it.p);
        *pdf = 1.f;
        pL->p = pLight;
        V3 w = -*wi;
        const float *m = lp + 24;  // WorldToLight applied to a vector
        V3 wl = normalize(v3(m[0] * w.x + m[1] * w.y + m[2] * w.z, m[4] * w.x + m[5] * w.y + m[6] * w.z, m[8] * w.x + m[9] * w.y + m[10] * w.z));
        float cosTheta = wl.z, cosTotalWidth = PG_LD(lp + 6), cosFalloffStart = PG_LD(lp + 7);
        float falloff;
        if (cosTheta < cosTotalWidth) falloff = 0;
        else if (cosTheta >= cosFalloffStart) falloff = 1;
        else {
            float delta = (cosTheta - cosTotalWidth) / (cosFalloffStart - cosTotalWidth);
            falloff = (delta * delta) * (delta * delta);
        }
        return loadSpec(lp) * falloff / distanceSquared(pLight, it.p);
    }
    if (tag == PBRTGPU_LIGHT_POINT) {
        V3 pLight = loadV3(lp + 3);
        *wi = normalize(pLight - it.p);
        *pdf = 1.f;
        pL->p = pLight;
        return loadSpec(lp) / distanceSquared(pLight, it.p);
    }
    // PBRTGPU_LIGHT_DISTANT
    V3 wLight = loadV3(lp + 3);
    *wi = wLight;
    *pdf = 1;
    pL->p = it.p + wLight * (2 * sc.worldRadius);
    return loadSpec(lp);
}

PG_NI_DIRECT float lightPdfLi(const DevScene &sc, int li, const SurfInt &it, V3 wi) {
    int tag = PG_LD(sc.lightTag + li);
    if (tag == PBRTGPU_LIGHT_DIFFUSE_AREA || tag == PBRTGPU_LIGHT_PORTAL_AREA)
        return shapePdfRef(sc.bvh, PG_LD(sc.lightPrim + li), it.p, it.pError, it.n, wi);
    if (tag == PBRTGPU_LIGHT_INFINITE) return infinitePdfLi(sc.lightParams + PBRTGPU_LIGHT_STRIDE * (size_t)li, wi);
    return 0;
}

PG_HD float powerHeuristic(float fPdf, float gPdf) {  // PowerHeuristic(1, f, 1, g)
    float f = 1 * fPdf, g = 1 * gPdf;
    return (f * f) / (f * f + g * g);
}

// The portal branch of EstimateDirect.  u1 is what the reference passes as u1:
// the SCATTERING sample (Q5); u2 is unused there as well.
// The two strategies that are NOT the reference's ("solidangle", "solidangle_mis"; portal.h).  One next-event
// ray per vertex, counted only if its closest hit is this light (NEE_MIS_LIGHT).  p_A = 1 / (sum of the solid
// angles of the portals in front) for a direction through one of them, else 0.
//   solidangle      the direction is drawn through a portal:  f |cos| Le / p_A
//   solidangle_mis  one-sample MIS (Veach 9.2.4) of that technique and BSDF sampling, each taken with
//                   probability 1/2, balance heuristic:       f |cos| Le / (p_A / 2 + p_B / 2)
//   solidangle_mis3 portal, BSDF and light-area sampling under PowerHeuristic3, AAPortal::InFrustum re-enabled (below)
// u1 = the scattering sample, u2 = the light sample of UniformSampleOneLight / AllLights.
PG_HD_NOINLINE bool estimateDirectPortalSolidAngle(const DevScene &sc, int li, int strat, const float *lightRec, const float *portals, int nPortals,
                                                   const SurfInt &it, const BSDF &bsdf, float u1x, float u1y, float u2x, float u2y, Spec scale,
                                                   NeeRequest *out) {
    const int flagsNS = BSDF_ALL & ~BSDF_SPECULAR;
    M44 o2w = loadM44(lightRec + 16), w2o = loadM44(lightRec + 32);
    int sel;
    SphRect rect;
    if (strat == PBRTGPU_PORTAL_SOLIDANGLE_MIS3) {
        // Three techniques -- A: a direction through a portal, uniform in solid angle (portals that fail AAPortal's frustum
        // test left out); B: BSDF sampling; L: a point of the light, uniform in area (Shape::Sample(ref, u), shape.cpp:56-77)
        // -- one of them drawn per vertex with equal probability c (one-sample model, Veach 9.2.4) and weighted with the
        // fork's PowerHeuristic3 (sampling.h:177-180; the equal c cancel): w_i = p_i^2 / (p_A^2 + p_B^2 + p_L^2), estimate
        // w_i f |cos| Le / (c p_i).  Where no portal is usable (behind all of them, outside every frustum) only B and L remain.
        PlaneParams lightPl = loadPlane(lightRec);
        float sumOmega = portalSolidAngles(portals, nPortals, o2w, w2o, it.p, u2x, &sel, &rect, &lightPl);
        const bool haveA = sumOmega > 0 && sel >= 0;
        const float pAin = haveA ? 1 / sumOmega : 0.f;
        const int nTech = haveA ? 3 : 2;
        int tech = (int)(u2y * nTech);
        if (tech > nTech - 1) tech = nTech - 1;
        tech += haveA ? 0 : 1;  // 0 = A, 1 = B, 2 = L
        V3 wi = v3(0, 0, 0);
        Spec f;
        float pA = 0, pB = 0, pL = 0;
        if (tech == 0) {
            wi = normalize(sphRectSample(rect, u1x, u1y) - it.p);
            pA = pAin;
        } else if (tech == 1) {
            int sampledType;
            f = bsdfSampleF(bsdf, it.wo, &wi, u1x, u1y, &pB, flagsNS, &sampledType);
            if (isBlack(f) || !(pB > 0)) return true;
        } else {
            ShapeSample pLs;
            Spec Li = lightSampleLi(sc, li, it, u1x, u1y, &wi, &pL, &pLs);
            if (isBlack(Li) || !(pL > 0)) return true;
        }
        if (tech != 1) {
            f = bsdfF(bsdf, it.wo, wi, flagsNS);
            pB = bsdfPdf(bsdf, it.wo, wi, flagsNS);
        }
        f *= absDot(wi, it.ns);
        if (isBlack(f)) return true;
        if (tech != 0) pA = (haveA && portalsContainDirection(portals, nPortals, o2w, w2o, it.p, wi, &lightPl)) ? pAin : 0.f;
        if (tech != 2) pL = lightPdfLi(sc, li, it, wi);
        float pSel = tech == 0 ? pA : (tech == 1 ? pB : pL);
        float denom = pA * pA + pB * pB + pL * pL;
        if (!(pSel > 0) || !(denom > 0)) return true;
        float weight = (pSel * pSel) / denom;  // PowerHeuristic3(1, pSel, 1, ., 1, .)
        out->hasClosest = true;
        out->closestRay = spawnRay(it.p, it.pError, it.n, wi);
        out->closestContrib = scale * (f * (weight * (float)nTech / pSel));
        out->light = li;
        out->mode = NEE_MIS_LIGHT;
        return true;
    }
    float sumOmega = portalSolidAngles(portals, nPortals, o2w, w2o, it.p, u2x, &sel, &rect);
    if (!(sumOmega > 0) || sel < 0) return false;  // behind every portal: the caller falls back to sampling the light
    const float pA = 1 / sumOmega;
    const bool mis = strat == PBRTGPU_PORTAL_SOLIDANGLE_MIS;
    V3 wi;
    Spec f;
    float pdf;
    if (!mis || u2y < 0.5f) {
        wi = normalize(sphRectSample(rect, u1x, u1y) - it.p);
        f = bsdfF(bsdf, it.wo, wi, flagsNS) * absDot(wi, it.ns);
        pdf = mis ? 0.5f * pA + 0.5f * bsdfPdf(bsdf, it.wo, wi, flagsNS) : pA;
    } else {
        int sampledType;
        float pB = 0;
        wi = v3(0, 0, 0);
        f = bsdfSampleF(bsdf, it.wo, &wi, u1x, u1y, &pB, flagsNS, &sampledType);
        if (isBlack(f) || !(pB > 0)) return true;
        f *= absDot(wi, it.ns);
        pdf = 0.5f * pB + (portalsContainDirection(portals, nPortals, o2w, w2o, it.p, wi) ? 0.5f * pA : 0.f);
    }
    if (isBlack(f) || !(pdf > 0)) return true;
    out->hasClosest = true;
    out->closestRay = spawnRay(it.p, it.pError, it.n, wi);
    out->closestContrib = scale * (f / pdf);
    out->light = li;
    out->mode = NEE_MIS_LIGHT;
    return true;
}

// EXT = false compiles the two non-reference strategies out: the path integrator's shade kernel is bound by
// instruction fetch, and carrying their code costs scenes that do not use them 2 % (r01h A/B).  The host
// launches the EXT kernel only when a light of the scene asks for one of them.
template <bool EXT>
PG_HD_NOINLINE void estimateDirectPortal(const DevScene &sc, int li, const SurfInt &it, const BSDF &bsdf, float u1x, float u1y,
                                float u2x, float u2y, Spec scale, NeeRequest *out) {
    const int flagsNS = BSDF_ALL & ~BSDF_SPECULAR;
    int strat = PG_LD(sc.lightStrategy + li);
    int slot = PG_LD(sc.lightPrim + li);
    int tag = f2i(PG_LDG(sc.bvh.tris + 3 * (size_t)slot).w);
    const float *lightRec = sc.bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift);
    int sel = -2;
    float portalPdf = 1;
    int pBegin = PG_LD(sc.lightPortalBegin + li), nPortals = PG_LD(sc.lightPortalBegin + li + 1) - pBegin;
    const float *portals = sc.portalParams + PBRTGPU_PORTAL_STRIDE * (size_t)pBegin;
    M44 w2o = loadM44(lightRec + 32);
    if (strat >= PBRTGPU_PORTAL_SOLIDANGLE) {
        if (EXT && nPortals > 0 && estimateDirectPortalSolidAngle(sc, li, strat, lightRec, portals, nPortals, it, bsdf, u1x, u1y, u2x, u2y, scale, out)) return;
        strat = PBRTGPU_PORTAL_UNIFORM_LIGHT;  // behind every portal
    }
    if (strat != PBRTGPU_PORTAL_UNIFORM_LIGHT) {
        // with no portal at all the loop leaves behindAll = true -> light sampling
        sel = nPortals > 0 ? portalSelect(portals, nPortals, w2o, it.p, u1x, &portalPdf) : -2;
        if (sel == -1) return;
        if (sel >= 0 && !portalInFront(loadPortal(portals + PBRTGPU_PORTAL_STRIDE * sel), w2o, it.p)) sel = -2;
    }
    V3 wi = v3(0, 0, 0);
    float pdf = 0;
    if (sel == -2) {  // EstimateDirectLight
        ShapeSample pL;
        Spec Li = lightSampleLi(sc, li, it, u1x, u1y, &wi, &pdf, &pL);
        if (isBlack(Li) || !(pdf > 0)) return;
        out->mode = NEE_ANY_OR_LIGHT;
    } else if (strat == PBRTGPU_PORTAL_UNIFORM_PORTAL) {
        portalSampleUniform(loadPortal(portals + PBRTGPU_PORTAL_STRIDE * sel), loadM44(lightRec + 16), it.p, u1x, u1y, &wi, &pdf);
        if (!(pdf > 0)) return;
        out->mode = NEE_ANY_EMITTER;
    } else {  // projection
        portalSampleProj(loadPortal(portals + PBRTGPU_PORTAL_STRIDE * sel), loadPlane(lightRec), w2o, it.p, u1x, &wi, &pdf);
        if (!(pdf > 0)) return;
        out->mode = NEE_ANY_EMITTER;
    }
    Spec f = bsdfF(bsdf, it.wo, wi, flagsNS) * absDot(wi, it.ns);
    if (isBlack(f)) return;
    Spec Ld = f / pdf;  // times Li when the ray comes back
    if (sel >= 0 && strat == PBRTGPU_PORTAL_PROJECTION) Ld = Ld / portalPdf;  // only this strategy divides (Q8)
    out->hasClosest = true;
    out->closestRay = spawnRay(it.p, it.pError, it.n, wi);
    out->closestContrib = scale * Ld;
    out->light = li;
}

// EstimateDirect for light `li` up to its rays (integrator.cpp:124-258; the portal hook :130-133),
// in one piece.  The path integrator's shade stage has the same steps spread over its phases
// (integrator_stages.h); this form serves Integrator "directlighting", which calls it once per
// light sample.  `scale` multiplies both contributions (throughput / pick pdf or / nSamples).
PG_NI_DIRECT void estimateDirectRays(const DevScene &sc, int li, const SurfInt &it, const BSDF &bsdf, float uL0, float uL1, float uS0,
                                       float uS1, Spec scale, NeeRequest *out, bool *culled) {
    out->hasShadow = out->hasClosest = false;
    out->light = -1;
    out->mode = NEE_MIS_LIGHT;
    *culled = false;
    int tag = PG_LD(sc.lightTag + li);
    if (tag == PBRTGPU_LIGHT_PORTAL_AREA) {
        estimateDirectPortal<true>(sc, li, it, bsdf, uS0, uS1, uL0, uL1, spec(1.f), out);
        if (out->hasClosest) out->closestContrib = scale * out->closestContrib;
        return;
    }
    const int flagsNS = BSDF_ALL & ~BSDF_SPECULAR;
    bool isDelta = tag == PBRTGPU_LIGHT_POINT || tag == PBRTGPU_LIGHT_DISTANT || tag == PBRTGPU_LIGHT_SPOT;
    V3 wi = v3(0, 0, 0);  // Vector3f() is zero-initialised
    float lightPdf = 0, scatteringPdf = 0;
    ShapeSample pL;
    Spec Li = lightSampleLi(sc, li, it, uL0, uL1, &wi, &lightPdf, &pL);
    if (lightPdf > 0 && !isBlack(Li)) {
        Spec f = bsdfF(bsdf, it.wo, wi, flagsNS) * absDot(wi, it.ns);
        scatteringPdf = bsdfPdf(bsdf, it.wo, wi, flagsNS);
        if (!isBlack(f)) {
            Spec Ld;
            if (isDelta) Ld = f * Li / lightPdf;
            else Ld = f * Li * powerHeuristic(lightPdf, scatteringPdf) / lightPdf;
            out->hasShadow = true;
            out->shadowRay = spawnRayTo(it.p, it.pError, it.n, pL.p, pL.pError, pL.n);
            out->shadowContrib = scale * Ld;
        }
    }
    if (!isDelta) {
        int sampledType;
        Spec f = bsdfSampleF(bsdf, it.wo, &wi, uS0, uS1, &scatteringPdf, flagsNS, &sampledType);
        f *= absDot(wi, it.ns);
        bool sampledSpecular = (sampledType & BSDF_SPECULAR) != 0;
        if (!isBlack(f) && scatteringPdf > 0) {
            float weight = 1;
            if (!sampledSpecular) {
                lightPdf = lightPdfLi(sc, li, it, wi);
                if (lightPdf == 0) return;
                weight = powerHeuristic(scatteringPdf, lightPdf);
            }
            Ray nr = spawnRay(it.p, it.pError, it.n, wi);
            // the ray can only count if its closest hit is the light's own shape (bvh.h:primitiveAccepts)
            if (tag == PBRTGPU_LIGHT_DIFFUSE_AREA && !primitiveAccepts(sc.bvh, PG_LD(sc.lightPrim + li), nr)) {
                *culled = true;
                return;
            }
            out->hasClosest = true;
            out->closestRay = nr;
            out->closestContrib = scale * (f * weight / scatteringPdf);
            out->light = li;
            out->mode = NEE_MIS_LIGHT;
        }
    }
}

}  // namespace pg
#endif
