// SceneApi, world-block part: materials, lights, shapes, WorldEnd.
// References: MakeMaterial api.cpp:548-620, TextureParams (core/paramset.cpp:
// 690-800), MakeLight/MakeAreaLight api.cpp:742-785, pbrtShape api.cpp:1348-1436,
// CreateAAPortal lights/portal_arealight.cpp:245-300.
#include <cmath>
#include <cstring>

#include "api.h"
#include "xform.h"

namespace pg {

namespace {

// TextureParams lookups restricted to constant textures: a "texture" parameter
// naming a constant texture, else the material/shape value, shape overriding
// material (paramset.cpp: GetSpectrumTexture / GetFloatTexture / FindFloat).
struct TexParams {
    const ParamSet &geom, &mat;
    const std::map<std::string, float> &ftex;
    const std::map<std::string, RGB> &stex;
    RGB spectrum(const std::string &n, RGB def) const {
        std::string name = geom.findTexture(n);
        if (name.empty()) name = mat.findTexture(n);
        if (!name.empty()) {
            auto it = stex.find(name);
            if (it != stex.end()) return it->second;
            hostError("Couldn't find spectrum texture named \"%s\" for parameter \"%s\"", name.c_str(), n.c_str());
        }
        RGB val = mat.findOneSpectrum(n, def);
        return geom.findOneSpectrum(n, val);
    }
    float scalar(const std::string &n, float def) const {
        std::string name = geom.findTexture(n);
        if (name.empty()) name = mat.findTexture(n);
        if (!name.empty()) {
            auto it = ftex.find(name);
            if (it != ftex.end()) return it->second;
            hostError("Couldn't find float texture named \"%s\" for parameter \"%s\"", name.c_str(), n.c_str());
        }
        float val = mat.findOneFloat(n, def);
        return geom.findOneFloat(n, val);
    }
    bool has(const std::string &n) const {  // GetFloatTextureOrNull != nullptr
        if (!geom.findTexture(n).empty() || !mat.findTexture(n).empty()) return true;
        return geom.find(n, ParamType::Float) || mat.find(n, ParamType::Float);
    }
    bool boolean(const std::string &n, bool def) const { return geom.findOneBool(n, mat.findOneBool(n, def)); }
};

// TrowbridgeReitzDistribution::RoughnessToAlpha, core/microfacet.h:126-132.
// The reference evaluates this per hit; with constant textures the value is
// the same every time, so it is folded here with the same libm call.
float roughnessToAlpha(float roughness) {
    roughness = std::max(roughness, (float)1e-3);
    float x = std::log(roughness);
    return 1.62142f + 0.819955f * x + 0.1734f * x * x + 0.0171201f * x * x * x + 0.000640711f * x * x * x * x;
}

// measured copper, the default of Material "metal" (materials/metal.cpp:84-118): wavelength (nm), n, k
const int kCopperSamples = 56;
const float kCopperWavelengths[kCopperSamples] = {
    298.7570554, 302.4004341, 306.1337728, 309.960445, 313.8839949, 317.9081487, 322.036826, 326.2741526, 330.6244747, 335.092373,
    339.6826795, 344.4004944, 349.2512056, 354.2405086, 359.374429, 364.6593471, 370.1020239, 375.7096303, 381.4897785, 387.4505563,
    393.6005651, 399.9489613, 406.5055016, 413.2805933, 420.2853492, 427.5316483, 435.0322035, 442.8006357, 450.8515564, 459.2006593,
    467.8648226, 476.8622231, 486.2124627, 495.936712, 506.0578694, 516.6007417, 527.5922468, 539.0616435, 551.0407911, 563.5644455,
    576.6705953, 590.4008476, 604.8008683, 619.92089, 635.8162974, 652.5483053, 670.1847459, 688.8009889, 708.4810171, 729.3186941,
    751.4192606, 774.9011125, 799.8979226, 826.5611867, 855.0632966, 885.6012714};
const float kCopperN[kCopperSamples] = {
    1.400313, 1.38, 1.358438, 1.34, 1.329063, 1.325, 1.3325, 1.34, 1.334375, 1.325, 1.317812, 1.31, 1.300313, 1.29, 1.281563, 1.27,
    1.249062, 1.225, 1.2, 1.18, 1.174375, 1.175, 1.1775, 1.18, 1.178125, 1.175, 1.172812, 1.17, 1.165312, 1.16, 1.155312, 1.15,
    1.142812, 1.135, 1.131562, 1.12, 1.092437, 1.04, 0.950375, 0.826, 0.645875, 0.468, 0.35125, 0.272, 0.230813, 0.214, 0.20925, 0.213,
    0.21625, 0.223, 0.2365, 0.25, 0.254188, 0.26, 0.28, 0.3};
const float kCopperK[kCopperSamples] = {
    1.662125, 1.687, 1.703313, 1.72, 1.744563, 1.77, 1.791625, 1.81, 1.822125, 1.834, 1.85175, 1.872, 1.89425, 1.916, 1.931688, 1.95,
    1.972438, 2.015, 2.121562, 2.21, 2.177188, 2.13, 2.160063, 2.21, 2.249938, 2.289, 2.326, 2.362, 2.397625, 2.433, 2.469187, 2.504,
    2.535875, 2.564, 2.589625, 2.605, 2.595562, 2.583, 2.5765, 2.599, 2.678062, 2.809, 3.01075, 3.24, 3.458187, 3.67, 3.863125, 4.05,
    4.239563, 4.43, 4.619563, 4.817, 5.034125, 5.26, 5.485625, 5.717};

void put3(float *dst, const RGB &c) { dst[0] = c[0], dst[1] = c[1], dst[2] = c[2]; }

}  // namespace

int SceneApi::makeMaterial(const std::string &name, const ParamSet &shapeParams, const ParamSet &matParams) {
    if (name.empty() || name == "none") return -1;
    TexParams mp{shapeParams, matParams, gs.floatTextures, gs.spectrumTextures};
    MaterialData m;
    if (mp.has("bumpmap")) hostWarning("bump mapping is not supported by the GPU path integrator; ignored");
    if (name == "plastic") {  // materials/plastic.cpp:45-85
        m.tag = PBRTGPU_MAT_PLASTIC;
        put3(m.params + 0, mp.spectrum("Kd", RGB(0.25f)).clamp());
        put3(m.params + 3, mp.spectrum("Ks", RGB(0.25f)).clamp());
        float rough = mp.scalar("roughness", .1f);
        if (mp.boolean("remaproughness", true)) rough = roughnessToAlpha(rough);
        m.params[6] = rough;
    } else if (name == "glass") {  // materials/glass.cpp:45-113
        m.tag = PBRTGPU_MAT_GLASS;
        put3(m.params + 0, mp.spectrum("Kr", RGB(1.f)).clamp());
        put3(m.params + 3, mp.spectrum("Kt", RGB(1.f)).clamp());
        float eta = mp.has("eta") ? mp.scalar("eta", 1.5f) : mp.scalar("index", 1.5f);
        float ur = mp.scalar("uroughness", 0.f), vr = mp.scalar("vroughness", 0.f);
        bool remap = mp.boolean("remaproughness", true);
        bool isSpecular = ur == 0 && vr == 0;
        if (!isSpecular && remap) ur = roughnessToAlpha(ur), vr = roughnessToAlpha(vr);
        m.params[6] = ur, m.params[7] = vr, m.params[8] = eta, m.params[9] = isSpecular ? 1.f : 0.f;
    } else if (name == "dispersive_glass") {  // materials/dispersive_glass.cpp:48-145
        m.tag = PBRTGPU_MAT_DISPERSIVE_GLASS;
        put3(m.params + 0, mp.spectrum("Kr", RGB(1.f)).clamp());
        put3(m.params + 3, mp.spectrum("Kt", RGB(1.f)).clamp());
        float etaMin = mp.has("etaMin") ? mp.scalar("etaMin", 1.5f) : mp.scalar("indexMin", 1.5f);
        float etaMax = mp.has("etaMax") ? mp.scalar("etaMax", 1.5f) : mp.scalar("indexMax", 1.5f);
        float ur = mp.scalar("uroughness", 0.f), vr = mp.scalar("vroughness", 0.f);
        bool remap = mp.boolean("remaproughness", false);
        bool isSpecular = ur == 0 && vr == 0;
        // path uses bsdf[0] only, for which the remap has been applied once (dispersive_glass.cpp:98-101)
        if (!isSpecular && remap) ur = roughnessToAlpha(ur), vr = roughnessToAlpha(vr);
        // Cauchy coefficients (dispersive_glass.cpp:44-45, 63-65); lambda in nm
        const float lambdaMinSq = 400 * 400, lambdaMaxSq = 700 * 700;
        const float cauchyB = (lambdaMinSq * etaMax - lambdaMaxSq * etaMin) / (lambdaMinSq - lambdaMaxSq);
        const float cauchyC = lambdaMinSq * (etaMax - cauchyB);
        m.params[6] = ur, m.params[7] = vr, m.params[8] = cauchyB, m.params[9] = isSpecular ? 1.f : 0.f;
        m.params[10] = cauchyC, m.params[11] = etaMin, m.params[12] = etaMax;
    } else if (name == "metal") {  // materials/metal.cpp:60-135
        m.tag = PBRTGPU_MAT_METAL;
        static const RGB copperN = rgbFromSampled(kCopperWavelengths, kCopperN, kCopperSamples);
        static const RGB copperK = rgbFromSampled(kCopperWavelengths, kCopperK, kCopperSamples);
        put3(m.params + 0, mp.spectrum("eta", copperN));
        put3(m.params + 3, mp.spectrum("k", copperK));
        float rough = mp.scalar("roughness", .01f);
        float ur = mp.has("uroughness") ? mp.scalar("uroughness", 0.f) : rough;
        float vr = mp.has("vroughness") ? mp.scalar("vroughness", 0.f) : rough;
        if (mp.boolean("remaproughness", true)) ur = roughnessToAlpha(ur), vr = roughnessToAlpha(vr);
        m.params[6] = ur, m.params[7] = vr;
    } else if (name == "uber") {  // materials/uber.cpp:107-135
        m.tag = PBRTGPU_MAT_UBER;
        put3(m.params + 0, mp.spectrum("Kd", RGB(0.25f)).clamp());
        put3(m.params + 3, mp.spectrum("Ks", RGB(0.25f)).clamp());
        put3(m.params + 6, mp.spectrum("Kr", RGB(0.f)).clamp());
        put3(m.params + 9, mp.spectrum("Kt", RGB(0.f)).clamp());
        put3(m.params + 12, mp.spectrum("opacity", RGB(1.f)).clamp());
        float rough = mp.scalar("roughness", .1f);
        float ru = mp.has("uroughness") ? mp.scalar("uroughness", 0.f) : rough;
        float rv = mp.has("vroughness") ? mp.scalar("vroughness", 0.f) : ru;
        if (mp.boolean("remaproughness", true)) ru = roughnessToAlpha(ru), rv = roughnessToAlpha(rv);
        m.params[15] = ru, m.params[16] = rv;
        m.params[17] = mp.has("eta") ? mp.scalar("eta", 1.5f) : mp.scalar("index", 1.5f);
    } else if (name == "mirror") {  // materials/mirror.cpp:45-64
        m.tag = PBRTGPU_MAT_MIRROR;
        put3(m.params + 0, mp.spectrum("Kr", RGB(0.9f)).clamp());
    } else {
        if (name != "matte") {
            // MakeMaterial (api.cpp:548-617): a name the reference does not know becomes matte with a warning, there and here.
            // A material the reference DOES implement but this library does not is refused: rendering it as matte would be a
            // silently different image.
            for (const char *known : {"translucent", "hair", "disney", "mix", "substrate", "subsurface", "kdsubsurface", "fourier"})
                if (name == known) {
                    hostError("Material \"%s\" is not supported by the GPU path integrator", name.c_str());
                    failed = true;
                    return -1;
                }
            hostWarning("Material \"%s\" unknown. Using \"matte\".", name.c_str());
        }
        m.tag = PBRTGPU_MAT_MATTE;  // materials/matte.cpp:45-73
        put3(m.params + 0, mp.spectrum("Kd", RGB(0.5f)).clamp());
        m.params[3] = clampf(mp.scalar("sigma", 0.f), 0, 90);
    }
    for (size_t i = 0; i < sc->materials.size(); ++i)
        if (sc->materials[i].tag == m.tag && !memcmp(sc->materials[i].params, m.params, sizeof(m.params))) return (int)i;
    sc->materials.push_back(m);
    return (int)sc->materials.size() - 1;
}

int SceneApi::materialForShape(const ParamSet &shapeParams) {  // GraphicsState::GetMaterialForShape
    return makeMaterial(gs.material.name, shapeParams, gs.material.params);
}

void SceneApi::LightSource(const std::string &name, const ParamSet &params) {
    if (state == State::Options) {
        hostError("Scene description must be inside world block; \"LightSource\" not allowed. Ignoring.");
        return;
    }
    LightData l;
    if (name == "infinite" || name == "exinfinite") {  // lights/infinite.cpp:176-186
        RGB L = params.findOneSpectrum("L", RGB(1.f)), s = params.findOneSpectrum("scale", RGB(1.f));
        std::string texmap = params.findOneString("mapname", "");
        l.nSamples = std::max(1, params.findOneInt("samples", params.findOneInt("nsamples", 1)));
        if (!texmap.empty()) {
            hostError("LightSource \"infinite\": environment maps (\"mapname\") are not supported by the GPU path integrator");
            failed = true;
            return;
        }
        l.tag = PBRTGPU_LIGHT_INFINITE;
        put3(l.params, L * s);
        memcpy(l.params + 8, &ctm.m.m[0][0], 16 * sizeof(float));      // LightToWorld
        memcpy(l.params + 24, &ctm.mInv.m[0][0], 16 * sizeof(float));  // WorldToLight
    } else if (name == "point") {  // lights/point.cpp:80-88
        RGB I = params.findOneSpectrum("I", RGB(1.f)), s = params.findOneSpectrum("scale", RGB(1.f));
        V3 P = params.findOnePoint3("from", v3(0, 0, 0));
        Xform l2w = xfMul(xfTranslate(P), ctm);
        V3 pLight = xformPoint(l2w.m, v3(0, 0, 0));
        l.tag = PBRTGPU_LIGHT_POINT;
        put3(l.params, I * s);
        l.params[3] = pLight.x, l.params[4] = pLight.y, l.params[5] = pLight.z;
    } else if (name == "spot") {  // lights/spot.cpp:105-125
        RGB I = params.findOneSpectrum("I", RGB(1.f)), s = params.findOneSpectrum("scale", RGB(1.f));
        float coneangle = params.findOneFloat("coneangle", 30.f), conedelta = params.findOneFloat("conedeltaangle", 5.f);
        V3 from = params.findOnePoint3("from", v3(0, 0, 0)), to = params.findOnePoint3("to", v3(0, 0, 1));
        V3 dir = normalize(to - from), du, dv;
        coordinateSystem(dir, &du, &dv);
        M44 dz = m44Identity();
        dz.m[0][0] = du.x, dz.m[0][1] = du.y, dz.m[0][2] = du.z;
        dz.m[1][0] = dv.x, dz.m[1][1] = dv.y, dz.m[1][2] = dv.z;
        dz.m[2][0] = dir.x, dz.m[2][1] = dir.y, dz.m[2][2] = dir.z;
        Xform l2w = xfMul(xfMul(ctm, xfTranslate(from)), xfInverse(xfFromMatrix(dz)));
        V3 pLight = xformPoint(l2w.m, v3(0, 0, 0));
        l.tag = PBRTGPU_LIGHT_SPOT;
        put3(l.params, I * s);
        l.params[3] = pLight.x, l.params[4] = pLight.y, l.params[5] = pLight.z;
        l.params[6] = std::cos(coneangle * (kPi / 180));               // cosTotalWidth = cos(Radians(totalWidth))
        l.params[7] = std::cos((coneangle - conedelta) * (kPi / 180));  // cosFalloffStart
        memcpy(l.params + 8, &l2w.m.m[0][0], 16 * sizeof(float));
        memcpy(l.params + 24, &l2w.mInv.m[0][0], 16 * sizeof(float));
    } else if (name == "distant") {  // lights/distant.cpp
        RGB L = params.findOneSpectrum("L", RGB(1.f)), s = params.findOneSpectrum("scale", RGB(1.f));
        V3 from = params.findOnePoint3("from", v3(0, 0, 0)), to = params.findOnePoint3("to", v3(0, 0, 1));
        V3 dir = from - to;
        V3 w = normalize(xformVector(ctm.m, dir));
        l.tag = PBRTGPU_LIGHT_DISTANT;
        put3(l.params, L * s);
        l.params[3] = w.x, l.params[4] = w.y, l.params[5] = w.z;
    } else {
        hostError("LightSource: light type \"%s\" is not supported by the GPU path integrator.", name.c_str());
        failed = true;
        return;
    }
    params.reportUnused();
    sc->lights.push_back(l);
}

// "((AA lox loy loz hix hiy hiz axis +|-) ...)" -> portals.  The reference
// hands the string to sexpresso and reads child(0)'s children
// (portal_arealight.cpp:256-279); tokens are whitespace/parenthesis separated.
static std::vector<PortalData> parsePortalData(const std::string &s, const PlaneData &lightShape) {
    std::vector<PortalData> out;
    std::vector<std::string> tok;
    std::string cur;
    for (char c : s) {
        if (c == '(' || c == ')' || c == ' ' || c == '\t' || c == '\n') {
            if (!cur.empty()) tok.push_back(cur), cur.clear();
            if (c == '(' || c == ')') tok.push_back(std::string(1, c));
        } else
            cur.push_back(c);
    }
    if (!cur.empty()) tok.push_back(cur);
    int depth = 0;
    for (size_t i = 0; i < tok.size(); ++i) {
        if (tok[i] == "(") {
            ++depth;
            if (depth == 2 && i + 9 < tok.size() && tok[i + 1] == "AA") {
                PortalData p;
                p.lo = v3(std::stof(tok[i + 2]), std::stof(tok[i + 3]), std::stof(tok[i + 4]));
                p.hi = v3(std::stof(tok[i + 5]), std::stof(tok[i + 6]), std::stof(tok[i + 7]));
                p.ax = std::stoi(tok[i + 8]);
                p.facingFw = tok[i + 9] == "+";
                // AAPortal builds its plane with the LIGHT's transforms (aaportal.cpp:12)
                int ax0 = p.ax == 2 ? 0 : (p.ax == 0 ? 1 : 2), ax1 = p.ax == 2 ? 1 : (p.ax == 0 ? 2 : 0);
                V3 loW = xformPoint(lightShape.o2w.m, p.lo), hiW = xformPoint(lightShape.o2w.m, p.hi);
                p.area = (hiW[ax0] - loW[ax0]) * (hiW[ax1] - loW[ax1]);
                out.push_back(p);
            }
        } else if (tok[i] == ")")
            --depth;
    }
    return out;
}

static float shapeArea(const SceneDesc &sc, const PrimDesc &p) {
    if (p.type == PBRTGPU_PRIM_TRIANGLE) {  // Triangle::Area, triangle.cpp:576-582
        const MeshData &m = sc.meshes[p.shape];
        V3 p0 = m.p[m.indices[3 * p.tri]], p1 = m.p[m.indices[3 * p.tri + 1]], p2 = m.p[m.indices[3 * p.tri + 2]];
        return 0.5 * length(cross(p1 - p0, p2 - p0));
    }
    if (p.type == PBRTGPU_PRIM_SPHERE) {  // Sphere::Area, sphere.cpp:217
        const SphereData &s = sc.spheres[p.shape];
        return s.phiMax * s.radius * (s.zMax - s.zMin);
    }
    return sc.planes[p.shape].area;
}

void SceneApi::Shape(const std::string &name, const ParamSet &params) {
    if (state == State::Options) {
        hostError("Scene description must be inside world block; \"Shape\" not allowed. Ignoring.");
        return;
    }
    if (inObject) return;
    ShapeResult sr;
    if (!makeShape(*sc, name, ctm, gs.reverseOrientation, params, &sr)) {
        failed = true;
        return;
    }
    int mtl = materialForShape(params);
    params.reportUnused();
    for (int i = 0; i < sr.count; ++i) {
        PrimDesc p;
        p.type = (uint8_t)sr.type;
        p.shape = sr.mesh;
        p.tri = i;
        p.material = mtl;
        if (!gs.areaLight.empty()) {
            const ParamSet &lp = gs.areaLightParams;
            LightData l;
            RGB L = lp.findOneSpectrum("L", RGB(1.f)), s = lp.findOneSpectrum("scale", RGB(1.f));
            l.nSamples = std::max(1, lp.findOneInt("samples", lp.findOneInt("nsamples", 1)));
            bool twoSided = lp.findOneBool("twosided", false);
            bool made = true;
            if (gs.areaLight == "area" || gs.areaLight == "diffuse") {
                l.tag = PBRTGPU_LIGHT_DIFFUSE_AREA;
            } else if (gs.areaLight == "portal") {
                if (sr.type != PBRTGPU_PRIM_AAPLANE) {
                    // the reference dereferences a null AAPlaneShape here (api.cpp:778-779)
                    hostError("AreaLightSource \"portal\" requires Shape \"aaplane\"");
                    failed = true;
                    return;
                }
                l.tag = PBRTGPU_LIGHT_PORTAL_AREA;
                l.portals = parsePortalData(lp.findOneString("portalData", ""), sc->planes[sr.mesh]);
                std::string st = lp.findOneString("strategy", "light");
                if (st == "light") l.strategy = PBRTGPU_PORTAL_UNIFORM_LIGHT;
                else if (st == "portal") l.strategy = PBRTGPU_PORTAL_UNIFORM_PORTAL;
                else if (st == "projection") l.strategy = PBRTGPU_PORTAL_PROJECTION;
                else if (st == "solidangle") l.strategy = PBRTGPU_PORTAL_SOLIDANGLE;  // the two strategies the reference does not have
                else if (st == "solidangle_mis") l.strategy = PBRTGPU_PORTAL_SOLIDANGLE_MIS;
                else if (st == "solidangle_mis3") l.strategy = PBRTGPU_PORTAL_SOLIDANGLE_MIS3;
                else {
                    hostWarning("AAPortal strategy unknown");
                    l.strategy = PBRTGPU_PORTAL_UNIFORM_LIGHT;  // the reference leaves it uninitialised
                }
            } else {
                hostWarning("Area light \"%s\" unknown.", gs.areaLight.c_str());
                made = false;
            }
            if (made) {
                put3(l.params, L * s);
                l.params[3] = twoSided ? 1.f : 0.f;
                l.params[4] = shapeArea(*sc, p);
                l.prim = (int)sc->prims.size();
                p.light = (int)sc->lights.size();
                sc->lights.push_back(l);
            }
            if (i == sr.count - 1) lp.reportUnused();
        }
        sc->prims.push_back(p);
    }
}

void SceneApi::WorldEnd() {
    if (state == State::Options) {
        hostError("Scene description must be inside world block; \"WorldEnd\" not allowed. Ignoring.");
        return;
    }
    while (!pushedStates.empty()) {
        hostWarning("Missing end to pbrtAttributeBegin()");
        pushedStates.pop_back();
        pushedTransforms.pop_back();
    }
    while (!pushedTransforms.empty()) {
        hostWarning("Missing end to pbrtTransformBegin()");
        pushedTransforms.pop_back();
    }
    sc->worldEnded = true;
    state = State::Options;
}

}  // namespace pg
