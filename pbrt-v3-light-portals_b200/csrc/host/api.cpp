// SceneApi: the graphics-state machine behind the .pbrt directives.  Follows
// src/core/api.cpp of the reference directive by directive (cited inline);
// instead of building Shape/Material/Light objects it fills a SceneDesc.
#include "api.h"

#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "xform.h"

namespace pg {

static int gWarnings = 0, gErrors = 0;
static bool gQuiet = false;
void setHostQuiet(bool q) { gQuiet = q; }
int hostErrorCount() { return gErrors; }

void hostWarning(const char *fmt, ...) {
    ++gWarnings;
    if (gQuiet) return;
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "Warning: ");
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
}
void hostError(const char *fmt, ...) {
    ++gErrors;
    va_list ap;
    va_start(ap, fmt);
    fprintf(stderr, "Error: ");
    vfprintf(stderr, fmt, ap);
    fprintf(stderr, "\n");
    va_end(ap);
}

std::string resolveFilename(const SceneDesc &sc, const std::string &name) {
    // ResolveFilename/AbsolutePath, src/core/fileutil.cpp: relative names are
    // taken relative to the directory of the top-level scene file.
    if (name.empty() || name[0] == '/' || sc.sceneDir.empty()) return name;
    return sc.sceneDir + "/" + name;
}

SceneApi::SceneApi(SceneDesc *out) : sc(out), ctm(xfIdentity()) { sc->cameraToWorld = xfIdentity(); }

void SceneApi::fail(const char *what) {
    hostError("%s", what);
    failed = true;
}

#define VERIFY_OPTIONS(func)                                                                          \
    if (state == State::World) {                                                                      \
        hostError("Options cannot be set inside world block; \"%s\" not allowed.  Ignoring.", func); \
        return;                                                                                       \
    }
#define VERIFY_WORLD(func)                                                                                 \
    if (state == State::Options) {                                                                         \
        hostError("Scene description must be inside world block; \"%s\" not allowed. Ignoring.", func);   \
        return;                                                                                            \
    }

// ---- transforms: api.cpp:906-1012.  Only the start transform exists here
// (no motion blur on the GPU path); ActiveTransform/TransformTimes are
// accepted and ignored with a warning.
void SceneApi::Identity() { ctm = xfIdentity(); }
void SceneApi::Translate(float dx, float dy, float dz) { ctm = xfMul(ctm, xfTranslate(v3(dx, dy, dz))); }
void SceneApi::Rotate(float angle, float ax, float ay, float az) { ctm = xfMul(ctm, xfRotate(angle, v3(ax, ay, az))); }
void SceneApi::Scale(float sx, float sy, float sz) { ctm = xfMul(ctm, xfScale(sx, sy, sz)); }
void SceneApi::LookAt(float ex, float ey, float ez, float lx, float ly, float lz, float ux, float uy, float uz) {
    ctm = xfMul(ctm, xfLookAt(v3(ex, ey, ez), v3(lx, ly, lz), v3(ux, uy, uz)));
}
static M44 fromColumnMajor(const float tr[16]) {  // api.cpp:936-940
    return m44(tr[0], tr[4], tr[8], tr[12], tr[1], tr[5], tr[9], tr[13], tr[2], tr[6], tr[10], tr[14], tr[3], tr[7],
               tr[11], tr[15]);
}
void SceneApi::ConcatTransform(const float tr[16]) { ctm = xfMul(ctm, xfFromMatrix(fromColumnMajor(tr))); }
void SceneApi::Transform(const float tr[16]) { ctm = xfFromMatrix(fromColumnMajor(tr)); }
void SceneApi::CoordinateSystem(const std::string &name) { namedCoordinateSystems[name] = ctm; }
void SceneApi::CoordSysTransform(const std::string &name) {
    auto it = namedCoordinateSystems.find(name);
    if (it != namedCoordinateSystems.end()) ctm = it->second;
    else hostWarning("Couldn't find named coordinate system \"%s\"", name.c_str());
}
void SceneApi::ActiveTransform(const std::string &which) {
    if (which != "All") hostWarning("ActiveTransform %s: animated transforms are not supported on the GPU path", which.c_str());
}
void SceneApi::TransformTimes(float, float) {}

// ---- options block: api.cpp:1040-1106
void SceneApi::PixelFilter(const std::string &name, const ParamSet &params) {
    VERIFY_OPTIONS("PixelFilter");
    sc->filterName = name;
    sc->filterParams = params;
}
void SceneApi::Film(const std::string &type, const ParamSet &params) {
    VERIFY_OPTIONS("Film");
    sc->filmName = type;
    sc->filmParams = params;
}
void SceneApi::Sampler(const std::string &name, const ParamSet &params) {
    VERIFY_OPTIONS("Sampler");
    sc->samplerName = name;
    sc->samplerParams = params;
}
void SceneApi::Accelerator(const std::string &name, const ParamSet &params) {
    VERIFY_OPTIONS("Accelerator");
    sc->acceleratorName = name;
    sc->acceleratorParams = params;
}
void SceneApi::Integrator(const std::string &name, const ParamSet &params) {
    VERIFY_OPTIONS("Integrator");
    sc->integratorName = name;
    sc->integratorParams = params;
}
void SceneApi::Camera(const std::string &name, const ParamSet &params) {
    VERIFY_OPTIONS("Camera");
    sc->cameraName = name;
    sc->cameraParams = params;
    sc->cameraToWorld = xfInverse(ctm);
    namedCoordinateSystems["camera"] = sc->cameraToWorld;
}
void SceneApi::MakeNamedMedium(const std::string &name, const ParamSet &) {
    hostWarning("MakeNamedMedium \"%s\": participating media are outside the GPU path integrator; ignored", name.c_str());
}
void SceneApi::MediumInterface(const std::string &, const std::string &) {
    hostWarning("MediumInterface: participating media are outside the GPU path integrator; ignored");
}

// ---- world block
void SceneApi::WorldBegin() {
    VERIFY_OPTIONS("WorldBegin");
    state = State::World;
    ctm = xfIdentity();
    namedCoordinateSystems["world"] = ctm;
}
void SceneApi::AttributeBegin() {
    VERIFY_WORLD("AttributeBegin");
    pushedStates.push_back(gs);
    pushedTransforms.push_back(ctm);
}
void SceneApi::AttributeEnd() {
    VERIFY_WORLD("AttributeEnd");
    if (pushedStates.empty()) {
        hostError("Unmatched pbrtAttributeEnd() encountered. Ignoring it.");
        return;
    }
    gs = pushedStates.back();
    pushedStates.pop_back();
    ctm = pushedTransforms.back();
    pushedTransforms.pop_back();
}
void SceneApi::TransformBegin() {
    VERIFY_WORLD("TransformBegin");
    pushedTransforms.push_back(ctm);
}
void SceneApi::TransformEnd() {
    VERIFY_WORLD("TransformEnd");
    if (pushedTransforms.empty()) {
        hostError("Unmatched pbrtTransformEnd() encountered. Ignoring it.");
        return;
    }
    ctm = pushedTransforms.back();
    pushedTransforms.pop_back();
}
void SceneApi::ReverseOrientation() {
    VERIFY_WORLD("ReverseOrientation");
    gs.reverseOrientation = !gs.reverseOrientation;
}
void SceneApi::ObjectBegin(const std::string &name) {
    VERIFY_WORLD("ObjectBegin");
    AttributeBegin();
    inObject = true;
    hostError("ObjectBegin \"%s\": object instancing is not supported by the GPU path integrator; the instance is dropped",
              name.c_str());
}
void SceneApi::ObjectEnd() {
    VERIFY_WORLD("ObjectEnd");
    inObject = false;
    AttributeEnd();
}
void SceneApi::ObjectInstance(const std::string &name) {
    VERIFY_WORLD("ObjectInstance");
    hostError("ObjectInstance \"%s\": object instancing is not supported by the GPU path integrator", name.c_str());
}

// Texture directive, api.cpp:1187-1250: only constant textures can be folded
// into material parameters on this path.
void SceneApi::Texture(const std::string &name, const std::string &type, const std::string &texname,
                       const ParamSet &params) {
    VERIFY_WORLD("Texture");
    if (texname != "constant") {
        hostError("Texture \"%s\": texture class \"%s\" is not supported by the GPU path integrator (constant only)",
                  name.c_str(), texname.c_str());
        return;
    }
    if (type == "float") gs.floatTextures[name] = params.findOneFloat("value", 1.f);
    else if (type == "color" || type == "spectrum") gs.spectrumTextures[name] = params.findOneSpectrum("value", RGB(1.f));
    else hostError("Texture type \"%s\" unknown.", type.c_str());
    params.reportUnused();
}

void SceneApi::Material(const std::string &name, const ParamSet &params) {
    VERIFY_WORLD("Material");
    gs.material.name = name;
    gs.material.params = params;
}
void SceneApi::MakeNamedMaterial(const std::string &name, const ParamSet &params) {
    VERIFY_WORLD("MakeNamedMaterial");
    MaterialInstance mi;
    mi.name = params.findOneString("type", "");
    mi.params = params;
    if (mi.name.empty()) {
        hostError("No parameter string \"type\" found in MakeNamedMaterial");
        return;
    }
    if (gs.namedMaterials.count(name)) hostWarning("Named material \"%s\" redefined.", name.c_str());
    gs.namedMaterials[name] = mi;
}
void SceneApi::NamedMaterial(const std::string &name) {
    VERIFY_WORLD("NamedMaterial");
    auto it = gs.namedMaterials.find(name);
    if (it == gs.namedMaterials.end()) {
        hostError("NamedMaterial \"%s\" unknown.", name.c_str());
        return;
    }
    gs.material = it->second;
}
void SceneApi::AreaLightSource(const std::string &name, const ParamSet &params) {
    VERIFY_WORLD("AreaLightSource");
    gs.areaLight = name;
    gs.areaLightParams = params;
}

}  // namespace pg
