// Second half of bsdf.h: per-lobe f / Pdf / Sample_f and the BSDF container.
#ifndef PG_DEVICE_BSDF_EVAL_H
#define PG_DEVICE_BSDF_EVAL_H

namespace pg {

PG_HD bool lobeMatches(const Lobe &l, int flags) { return (l.type & flags) == l.type; }  // BxDF::MatchesFlags

PG_HD Spec microReflF(const Lobe &l, float etaB, V3 wo, V3 wi) {  // MicrofacetReflection::f
    float cosThetaO = absCosTheta(wo), cosThetaI = absCosTheta(wi);
    V3 wh = wi + wo;
    if (cosThetaI == 0 || cosThetaO == 0) return spec(0.f);
    if (wh.x == 0 && wh.y == 0 && wh.z == 0) return spec(0.f);
    wh = normalize(wh);
    float cosI = dot(wi, faceforward(wh, v3(0, 0, 1)));
    Spec F = l.kind == LOBE_MICRO_REFL_COND ? frConductor(fabsf(cosI), spec(1.f), l.T, l.K) : spec(frDielectric(cosI, l.etaA, etaB));
    return l.R * trD(l.a, l.b, wh) * trG(l.a, l.b, wo, wi) * F / (4 * cosThetaI * cosThetaO);
}

PG_HD Spec microTransF(const Lobe &l, float etaB, V3 wo, V3 wi) {  // MicrofacetTransmission::f, TransportMode::Radiance
    if (sameHemisphere(wo, wi)) return spec(0.f);
    float cosThetaO = cosTheta(wo), cosThetaI = cosTheta(wi);
    if (cosThetaI == 0 || cosThetaO == 0) return spec(0.f);
    float eta = cosTheta(wo) > 0 ? (etaB / l.etaA) : (l.etaA / etaB);
    V3 wh = normalize(wo + wi * eta);
    if (wh.z < 0) wh = -wh;
    if (dot(wo, wh) * dot(wi, wh) > 0) return spec(0.f);
    float F = frDielectric(dot(wo, wh), l.etaA, etaB);
    float sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
    float factor = 1 / eta;
    return (spec(1.f) - spec(F)) * l.R *
           fabsf(trD(l.a, l.b, wh) * trG(l.a, l.b, wo, wi) * eta * eta * absDot(wi, wh) * absDot(wo, wh) * factor * factor /
                 (cosThetaI * cosThetaO * sqrtDenom * sqrtDenom));
}

PG_HD float microTransPdf(const Lobe &l, float etaB, V3 wo, V3 wi) {
    if (sameHemisphere(wo, wi)) return 0;
    float eta = cosTheta(wo) > 0 ? (etaB / l.etaA) : (l.etaA / etaB);
    V3 wh = normalize(wo + wi * eta);
    if (dot(wo, wh) * dot(wi, wh) > 0) return 0;
    float sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
    float dwh_dwi = fabsf((eta * eta * dot(wi, wh)) / (sqrtDenom * sqrtDenom));
    return trPdf(l.a, l.b, wo, wh) * dwh_dwi;
}

PG_HD Spec orenNayarF(const Lobe &l, V3 wo, V3 wi) {
    float sinThetaI = sinTheta(wi), sinThetaO = sinTheta(wo);
    float maxCos = 0;
    if (sinThetaI > 1e-4 && sinThetaO > 1e-4) {
        float sinPhiI = sinPhi(wi), cosPhiI = cosPhi(wi);
        float sinPhiO = sinPhi(wo), cosPhiO = cosPhi(wo);
        float dCos = cosPhiI * cosPhiO + sinPhiI * sinPhiO;
        maxCos = maxf(0.f, dCos);
    }
    float sinAlpha, tanBeta;
    if (absCosTheta(wi) > absCosTheta(wo)) {
        sinAlpha = sinThetaO;
        tanBeta = sinThetaI / absCosTheta(wi);
    } else {
        sinAlpha = sinThetaI;
        tanBeta = sinThetaO / absCosTheta(wo);
    }
    return l.R * kInvPi * (l.a + l.b * maxCos * sinAlpha * tanBeta);
}

PG_NI_LOBE Spec lobeF(const Lobe &l, float etaB, V3 wo, V3 wi) {
    switch (l.kind) {
    case LOBE_LAMBERT: return l.R * kInvPi;
    case LOBE_OREN_NAYAR: return orenNayarF(l, wo, wi);
    case LOBE_MICRO_REFL:
    case LOBE_MICRO_REFL_COND: return microReflF(l, etaB, wo, wi);
    case LOBE_MICRO_TRANS: return microTransF(l, etaB, wo, wi);
    default: return spec(0.f);  // specular lobes
    }
}

PG_NI_LOBE float lobePdf(const Lobe &l, float etaB, V3 wo, V3 wi) {
    switch (l.kind) {
    case LOBE_LAMBERT:
    case LOBE_OREN_NAYAR: return sameHemisphere(wo, wi) ? absCosTheta(wi) * kInvPi : 0;
    case LOBE_MICRO_REFL:
    case LOBE_MICRO_REFL_COND: {
        if (!sameHemisphere(wo, wi)) return 0;
        V3 wh = normalize(wo + wi);
        return trPdf(l.a, l.b, wo, wh) / (4 * dot(wo, wh));
    }
    case LOBE_MICRO_TRANS: return microTransPdf(l, etaB, wo, wi);
    default: return 0;
    }
}

// BxDF::Sample_f of each lobe; *pdf must be 0 on entry (BSDF::Sample_f sets it).
PG_NI_LOBE Spec lobeSample(const Lobe &l, float etaB, V3 wo, float u0, float u1, V3 *wi, float *pdf, int *sampledType) {
    switch (l.kind) {
    case LOBE_LAMBERT:
    case LOBE_OREN_NAYAR:
        *wi = cosineSampleHemisphere(u0, u1);
        if (wo.z < 0) wi->z *= -1;
        *pdf = lobePdf(l, etaB, wo, *wi);
        return lobeF(l, etaB, wo, *wi);
    case LOBE_MICRO_REFL:
    case LOBE_MICRO_REFL_COND: {
        if (wo.z == 0) return spec(0.f);
        V3 wh = trSampleWh(l.a, l.b, wo, u0, u1);
        if (dot(wo, wh) < 0) return spec(0.f);
        *wi = reflectAbout(wo, wh);
        if (!sameHemisphere(wo, *wi)) return spec(0.f);
        *pdf = trPdf(l.a, l.b, wo, wh) / (4 * dot(wo, wh));
        return microReflF(l, etaB, wo, *wi);
    }
    case LOBE_MICRO_TRANS: {
        if (wo.z == 0) return spec(0.f);
        V3 wh = trSampleWh(l.a, l.b, wo, u0, u1);
        if (dot(wo, wh) < 0) return spec(0.f);
        float eta = cosTheta(wo) > 0 ? (l.etaA / etaB) : (etaB / l.etaA);
        if (!refract(wo, wh, eta, wi)) return spec(0.f);
        *pdf = microTransPdf(l, etaB, wo, *wi);
        return microTransF(l, etaB, wo, *wi);
    }
    case LOBE_FRESNEL_SPECULAR: {
        float F = frDielectric(cosTheta(wo), l.etaA, etaB);
        if (u0 < F) {
            *wi = v3(-wo.x, -wo.y, wo.z);
            *sampledType = BSDF_SPECULAR | BSDF_REFLECTION;
            *pdf = F;
            return F * l.R / absCosTheta(*wi);
        }
        bool entering = cosTheta(wo) > 0;
        float etaI = entering ? l.etaA : etaB;
        float etaT = entering ? etaB : l.etaA;
        if (!refract(wo, faceforward(v3(0, 0, 1), wo), etaI / etaT, wi)) return spec(0.f);
        Spec ft = l.T * (1 - F);
        ft *= (etaI * etaI) / (etaT * etaT);
        *sampledType = BSDF_SPECULAR | BSDF_TRANSMISSION;
        *pdf = 1 - F;
        return ft / absCosTheta(*wi);
    }
    case LOBE_SPEC_REFL_DIEL: {  // SpecularReflection::Sample_f with FresnelDielectric, reflection.cpp:131-141
        *wi = v3(-wo.x, -wo.y, wo.z);
        *pdf = 1;
        return frDielectric(cosTheta(*wi), l.etaA, etaB) * l.R / absCosTheta(*wi);
    }
    case LOBE_SPEC_TRANS: {  // SpecularTransmission::Sample_f, reflection.cpp:148-163 (TransportMode::Radiance)
        bool entering = cosTheta(wo) > 0;
        float etaI = entering ? l.etaA : etaB;
        float etaT = entering ? etaB : l.etaA;
        if (!refract(wo, faceforward(v3(0, 0, 1), wo), etaI / etaT, wi)) return spec(0.f);
        *pdf = 1;
        Spec ft = l.R * (spec(1.f) - spec(frDielectric(cosTheta(*wi), l.etaA, etaB)));
        ft *= (etaI * etaI) / (etaT * etaT);
        return ft / absCosTheta(*wi);
    }
    default: {  // LOBE_SPEC_REFL_NOOP: SpecularReflection with FresnelNoOp
        *wi = v3(-wo.x, -wo.y, wo.z);
        *pdf = 1;
        return spec(1.f) * l.R / absCosTheta(*wi);
    }
    }
}

PG_HD V3 bsdfWorldToLocal(const BSDF &b, V3 v) { return v3(dot(v, b.ss), dot(v, b.ts), dot(v, b.ns)); }
PG_HD V3 bsdfLocalToWorld(const BSDF &b, V3 v) {
    return v3(b.ss.x * v.x + b.ts.x * v.y + b.ns.x * v.z, b.ss.y * v.x + b.ts.y * v.y + b.ns.y * v.z,
              b.ss.z * v.x + b.ts.z * v.y + b.ns.z * v.z);
}
// etaT of lobe i: the material table's, or the per-path value of a dispersive material
PG_HD float lobeEtaB(const BSDF &b, int i) { return b.etaOverride != 0.f ? b.etaOverride : b.lobePtr()[i].etaB; }
PG_HD int bsdfNumComponents(const BSDF &b, int flags) {
    int n = 0;
    for (int i = 0; i < b.nLobes; ++i)
        if (lobeMatches(b.lobePtr()[i], flags)) ++n;
    return n;
}

PG_NI_BSDF Spec bsdfF(const BSDF &b, V3 woW, V3 wiW, int flags) {
    V3 wi = bsdfWorldToLocal(b, wiW), wo = bsdfWorldToLocal(b, woW);
    if (wo.z == 0) return spec(0.f);
    bool reflect = dot(wiW, b.ng) * dot(woW, b.ng) > 0;
    Spec f = spec(0.f);
    for (int i = 0; i < b.nLobes; ++i)
        if (lobeMatches(b.lobePtr()[i], flags) && ((reflect && (b.lobePtr()[i].type & BSDF_REFLECTION)) ||
                                               (!reflect && (b.lobePtr()[i].type & BSDF_TRANSMISSION))))
            f += lobeF(b.lobePtr()[i], lobeEtaB(b, i), wo, wi);
    return f;
}

PG_NI_BSDF float bsdfPdf(const BSDF &b, V3 woW, V3 wiW, int flags) {
    if (b.nLobes == 0) return 0.f;
    V3 wo = bsdfWorldToLocal(b, woW), wi = bsdfWorldToLocal(b, wiW);
    if (wo.z == 0) return 0.f;
    float pdf = 0.f;
    int matchingComps = 0;
    for (int i = 0; i < b.nLobes; ++i)
        if (lobeMatches(b.lobePtr()[i], flags)) {
            ++matchingComps;
            pdf += lobePdf(b.lobePtr()[i], lobeEtaB(b, i), wo, wi);
        }
    return matchingComps > 0 ? pdf / matchingComps : 0.f;
}

PG_NI_BSDF Spec bsdfSampleF(const BSDF &b, V3 woW, V3 *wiW, float u0, float u1, float *pdf, int type, int *sampledType) {
    int matchingComps = bsdfNumComponents(b, type);
    if (matchingComps == 0) {
        *pdf = 0;
        *sampledType = 0;
        return spec(0.f);
    }
    int comp = (int)floorf(u0 * matchingComps);
    if (comp > matchingComps - 1) comp = matchingComps - 1;
    int which = -1, count = comp;
    for (int i = 0; i < b.nLobes; ++i)
        if (lobeMatches(b.lobePtr()[i], type) && count-- == 0) {
            which = i;
            break;
        }
    const Lobe &bx = b.lobePtr()[which];
    float uR = minf(u0 * matchingComps - comp, kOneMinusEpsilon);
    V3 wi, wo = bsdfWorldToLocal(b, woW);
    // the reference returns here without touching *pdf / *sampledType (reflection.cpp:779);
    // the caller's values are then whatever they held -- zero both so the path ends the same way
    if (wo.z == 0) {
        *pdf = 0;
        *sampledType = 0;
        return spec(0.f);
    }
    *pdf = 0;
    *sampledType = bx.type;
    Spec f = lobeSample(bx, lobeEtaB(b, which), wo, uR, u1, &wi, pdf, sampledType);
    if (*pdf == 0) {
        *sampledType = 0;
        return spec(0.f);
    }
    *wiW = bsdfLocalToWorld(b, wi);
    if (!(bx.type & BSDF_SPECULAR) && matchingComps > 1)
        for (int i = 0; i < b.nLobes; ++i)
            if (i != which && lobeMatches(b.lobePtr()[i], type)) *pdf += lobePdf(b.lobePtr()[i], lobeEtaB(b, i), wo, wi);
    if (matchingComps > 1) *pdf /= matchingComps;
    if (!(bx.type & BSDF_SPECULAR)) {
        bool reflect = dot(*wiW, b.ng) * dot(woW, b.ng) > 0;
        f = spec(0.f);
        for (int i = 0; i < b.nLobes; ++i)
            if (lobeMatches(b.lobePtr()[i], type) && ((reflect && (b.lobePtr()[i].type & BSDF_REFLECTION)) ||
                                                  (!reflect && (b.lobePtr()[i].type & BSDF_TRANSMISSION))))
                f += lobeF(b.lobePtr()[i], lobeEtaB(b, i), wo, wi);
    }
    return f;
}

// ---- a BSDF that is ONE LambertianReflection (MatLobes::shadeClass 0): bsdfF / bsdfPdf / bsdfSampleF above with the lobe
// loops, the lobe switch and the calls folded away.  Every arithmetic operation the general routines perform on such
// a BSDF is performed here, in the same order, so the results are bit-identical (the shade kernel's Lambert
// instantiation and its general one give the same image; tests/test_render.py knob PBRTGPU_SHADE_LAMBERT).
//   BSDF::f        reflection.cpp:713-726   one matching lobe: f = reflect ? R / pi : 0
//   BSDF::Pdf      reflection.cpp:814-829   pdf = SameHemisphere ? |cos| / pi : 0, divided by one matching component
//   BSDF::Sample_f reflection.cpp:747-812   comp = 0, uRemapped = min(u0, 1 - eps), cosine-sampled wi, f re-evaluated
PG_HD Spec lambertF(const BSDF &b, Spec R, V3 woW, V3 wiW) {
    V3 wo = bsdfWorldToLocal(b, woW);
    if (wo.z == 0) return spec(0.f);
    bool reflect = dot(wiW, b.ng) * dot(woW, b.ng) > 0;
    Spec f = spec(0.f);
    if (reflect) f += R * kInvPi;
    return f;
}
PG_HD float lambertPdf(const BSDF &b, V3 woW, V3 wiW) {
    V3 wo = bsdfWorldToLocal(b, woW), wi = bsdfWorldToLocal(b, wiW);
    if (wo.z == 0) return 0.f;
    float pdf = 0.f;
    pdf += sameHemisphere(wo, wi) ? absCosTheta(wi) * kInvPi : 0;
    return pdf / 1;
}
PG_HD Spec lambertSampleF(const BSDF &b, Spec R, V3 woW, V3 *wiW, float u0, float u1, float *pdf, int *sampledType) {
    int comp = (int)floorf(u0 * 1);
    if (comp > 0) comp = 0;
    float uR = minf(u0 * 1 - comp, kOneMinusEpsilon);
    V3 wo = bsdfWorldToLocal(b, woW);
    if (wo.z == 0) {
        *pdf = 0;
        *sampledType = 0;
        return spec(0.f);
    }
    V3 wi = cosineSampleHemisphere(uR, u1);
    if (wo.z < 0) wi.z *= -1;
    *pdf = sameHemisphere(wo, wi) ? absCosTheta(wi) * kInvPi : 0;
    *sampledType = BSDF_REFLECTION | BSDF_DIFFUSE;
    if (*pdf == 0) {
        *sampledType = 0;
        return spec(0.f);
    }
    *wiW = bsdfLocalToWorld(b, wi);
    bool reflect = dot(*wiW, b.ng) * dot(woW, b.ng) > 0;
    Spec f = spec(0.f);
    if (reflect) f += R * kInvPi;
    return f;
}

// Material::ComputeScatteringFunctions for the material table of pbrtgpu.h
// (allowMultipleLobes = true, TransportMode::Radiance, as path.cpp:107 calls it).
//
// allowMultipleLobes: true for the path integrator (path.cpp:107), false for SamplerIntegrator-derived
// Li's such as directlighting (integrator.cpp:615 default) -- it only matters for specular glass.
// Textures are constants, so which BxDFs a material adds and with what parameters is the same at
// every vertex: makeMaterialLobes runs ONCE per material (at upload, on the host) and the kernels
// read the result from a small table.  Per vertex only the shading frame is built -- and, for
// dispersive_glass, the index of refraction of the path's hero wavelength.
PG_HD void makeMaterialLobes(int tag, const float *mp, bool allowMultipleLobes, MatLobes *m) {
    m->nLobes = 0;
    m->eta = 1;
    m->dispersive = 0;
    if (tag == PBRTGPU_MAT_MATTE) {
        Spec r = loadSpec(mp);
        float sig = mp[3];
        if (!isBlack(r)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.type = BSDF_REFLECTION | BSDF_DIFFUSE;
            l.R = r;
            if (sig == 0) l.kind = LOBE_LAMBERT;
            else {
                l.kind = LOBE_OREN_NAYAR;
                float sigma = radians(sig);
                float sigma2 = sigma * sigma;
                l.a = 1.f - (sigma2 / (2.f * (sigma2 + 0.33f)));
                l.b = 0.45f * sigma2 / (sigma2 + 0.09f);
            }
        }
    } else if (tag == PBRTGPU_MAT_PLASTIC) {
        Spec kd = loadSpec(mp), ks = loadSpec(mp + 3);
        if (!isBlack(kd)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_LAMBERT;
            l.type = BSDF_REFLECTION | BSDF_DIFFUSE;
            l.R = kd;
        }
        if (!isBlack(ks)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_MICRO_REFL;
            l.type = BSDF_REFLECTION | BSDF_GLOSSY;
            l.R = ks;
            l.a = l.b = mp[6];
            l.etaA = 1.5f, l.etaB = 1.f;
        }
    } else if (tag == PBRTGPU_MAT_GLASS || tag == PBRTGPU_MAT_DISPERSIVE_GLASS) {
        Spec R = loadSpec(mp), T = loadSpec(mp + 3);
        float eta;
        if (tag == PBRTGPU_MAT_GLASS) eta = mp[8];
        else eta = 1.f, m->dispersive = 1;  // eta = cauchyB + cauchyC / lambda^2 per path: BSDF::etaOverride (makeBSDF)
        m->eta = eta;
        if (isBlack(R) && isBlack(T)) return;
        bool isSpecular = mp[9] != 0.f;
        if (isSpecular && !allowMultipleLobes) {  // SamplerIntegrator::Li's ComputeScatteringFunctions (glass.cpp:74-87)
            if (!isBlack(R)) {
                Lobe &l = m->lobes[m->nLobes++];
                l.kind = LOBE_SPEC_REFL_DIEL;
                l.type = BSDF_REFLECTION | BSDF_SPECULAR;
                l.R = R;
                l.etaA = 1.f, l.etaB = eta;
            }
            if (!isBlack(T)) {
                Lobe &l = m->lobes[m->nLobes++];
                l.kind = LOBE_SPEC_TRANS;
                l.type = BSDF_TRANSMISSION | BSDF_SPECULAR;
                l.R = T;
                l.etaA = 1.f, l.etaB = eta;
            }
        } else if (isSpecular) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_FRESNEL_SPECULAR;
            l.type = BSDF_REFLECTION | BSDF_TRANSMISSION | BSDF_SPECULAR;
            l.R = R, l.T = T;
            l.etaA = 1.f, l.etaB = eta;
        } else {
            float ax = mp[6], ay = mp[7];
            if (!isBlack(R)) {
                Lobe &l = m->lobes[m->nLobes++];
                l.kind = LOBE_MICRO_REFL;
                l.type = BSDF_REFLECTION | BSDF_GLOSSY;
                l.R = R;
                l.a = ax, l.b = ay;
                l.etaA = 1.f, l.etaB = eta;
            }
            if (!isBlack(T)) {
                Lobe &l = m->lobes[m->nLobes++];
                l.kind = LOBE_MICRO_TRANS;
                l.type = BSDF_TRANSMISSION | BSDF_GLOSSY;
                l.R = T;
                l.a = ax, l.b = ay;
                l.etaA = 1.f, l.etaB = eta;
            }
        }
    } else if (tag == PBRTGPU_MAT_METAL) {  // materials/metal.cpp:60-82
        Lobe &l = m->lobes[m->nLobes++];
        l.kind = LOBE_MICRO_REFL_COND;
        l.type = BSDF_REFLECTION | BSDF_GLOSSY;
        l.R = spec(1.f);
        l.T = loadSpec(mp), l.K = loadSpec(mp + 3);
        l.a = mp[6], l.b = mp[7];
    } else if (tag == PBRTGPU_MAT_UBER) {  // materials/uber.cpp:45-105
        Spec Kd = loadSpec(mp), Ks = loadSpec(mp + 3), Kr = loadSpec(mp + 6), Kt = loadSpec(mp + 9), op = loadSpec(mp + 12);
        float e = mp[17];
        Spec t = clampSpec(-op + spec(1.f));
        if (!isBlack(t)) {
            m->eta = 1.f;
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_SPEC_TRANS;
            l.type = BSDF_TRANSMISSION | BSDF_SPECULAR;
            l.R = t;
            l.etaA = 1.f, l.etaB = 1.f;
        } else
            m->eta = e;
        Spec kd = op * Kd;
        if (!isBlack(kd)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_LAMBERT;
            l.type = BSDF_REFLECTION | BSDF_DIFFUSE;
            l.R = kd;
        }
        Spec ks = op * Ks;
        if (!isBlack(ks)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_MICRO_REFL;
            l.type = BSDF_REFLECTION | BSDF_GLOSSY;
            l.R = ks;
            l.a = mp[15], l.b = mp[16];
            l.etaA = 1.f, l.etaB = e;
        }
        Spec kr = op * Kr;
        if (!isBlack(kr)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_SPEC_REFL_DIEL;
            l.type = BSDF_REFLECTION | BSDF_SPECULAR;
            l.R = kr;
            l.etaA = 1.f, l.etaB = e;
        }
        Spec kt = op * Kt;
        if (!isBlack(kt)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_SPEC_TRANS;
            l.type = BSDF_TRANSMISSION | BSDF_SPECULAR;
            l.R = kt;
            l.etaA = 1.f, l.etaB = e;
        }
    } else if (tag == PBRTGPU_MAT_MIRROR) {
        Spec R = loadSpec(mp);
        if (!isBlack(R)) {
            Lobe &l = m->lobes[m->nLobes++];
            l.kind = LOBE_SPEC_REFL_NOOP;
            l.type = BSDF_REFLECTION | BSDF_SPECULAR;
            l.R = R;
        }
    }
}

PG_NI_BSDF void makeBSDF(const SurfInt &it, const MatLobes *table, int mat, const float *mp, float wvl0, BSDF *b) {
    b->ns = it.ns;
    b->ng = it.n;
    b->ss = normalize(it.dpdus);
    b->ts = cross(b->ns, b->ss);
    const MatLobes *m = table + mat;
    b->setLobes(m->lobes);
    b->nLobes = PG_LD(&m->nLobes);
    b->eta = PG_LD(&m->eta);
    b->etaOverride = 0.f;
    if (PG_LD(&m->dispersive)) {
        float eta = PG_LD(mp + 8) + PG_LD(mp + 10) / (wvl0 * wvl0);  // cauchyB + cauchyC / lambda^2 (dispersive_glass.cpp:63-65)
        b->eta = eta;
        b->etaOverride = eta;
    }
}

}  // namespace pg
#endif
