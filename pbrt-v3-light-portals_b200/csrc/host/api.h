// Scene-description API of the host front end: one method per .pbrt directive,
// same names (minus the pbrt prefix), argument meaning and state machine as the
// reference's src/core/api.h / api.cpp.  The parser calls these; at WorldEnd
// the accumulated SceneDesc is complete.
#ifndef PG_HOST_API_H
#define PG_HOST_API_H

#include <map>
#include <string>
#include <vector>

#include "params.h"
#include "scene.h"

namespace pg {

struct ShapeResult {
    int type = PBRTGPU_PRIM_TRIANGLE;
    int mesh = -1;  // mesh / plane / sphere index
    int count = 0;  // primitives created (triangles in the mesh, else 1)
};

bool makeShape(SceneDesc &sc, const std::string &name, const Xform &o2w, bool reverseOrientation,
               const ParamSet &params, ShapeResult *out);
std::string resolveFilename(const SceneDesc &sc, const std::string &name);

class SceneApi {
  public:
    explicit SceneApi(SceneDesc *out);
    // transforms (api.cpp:906-1012)
    void Identity();
    void Translate(float dx, float dy, float dz);
    void Rotate(float angle, float ax, float ay, float az);
    void Scale(float sx, float sy, float sz);
    void LookAt(float ex, float ey, float ez, float lx, float ly, float lz, float ux, float uy, float uz);
    void ConcatTransform(const float tr[16]);
    void Transform(const float tr[16]);
    void CoordinateSystem(const std::string &name);
    void CoordSysTransform(const std::string &name);
    void ActiveTransform(const std::string &which);
    void TransformTimes(float start, float end);
    // options block (api.cpp:1040-1106)
    void PixelFilter(const std::string &name, const ParamSet &params);
    void Film(const std::string &type, const ParamSet &params);
    void Sampler(const std::string &name, const ParamSet &params);
    void Accelerator(const std::string &name, const ParamSet &params);
    void Integrator(const std::string &name, const ParamSet &params);
    void Camera(const std::string &name, const ParamSet &params);
    void MakeNamedMedium(const std::string &name, const ParamSet &params);
    void MediumInterface(const std::string &inside, const std::string &outside);
    // world block (api.cpp:1137-1700)
    void WorldBegin();
    void AttributeBegin();
    void AttributeEnd();
    void TransformBegin();
    void TransformEnd();
    void Texture(const std::string &name, const std::string &type, const std::string &texname, const ParamSet &params);
    void Material(const std::string &name, const ParamSet &params);
    void MakeNamedMaterial(const std::string &name, const ParamSet &params);
    void NamedMaterial(const std::string &name);
    void LightSource(const std::string &name, const ParamSet &params);
    void AreaLightSource(const std::string &name, const ParamSet &params);
    void Shape(const std::string &name, const ParamSet &params);
    void ReverseOrientation();
    void ObjectBegin(const std::string &name);
    void ObjectEnd();
    void ObjectInstance(const std::string &name);
    void WorldEnd();

    bool ok() const { return !failed; }

  private:
    struct MaterialInstance {
        std::string name = "matte";
        ParamSet params;
    };
    struct GraphicsState {
        MaterialInstance material;
        std::map<std::string, MaterialInstance> namedMaterials;
        std::map<std::string, float> floatTextures;  // constant textures only
        std::map<std::string, RGB> spectrumTextures;
        ParamSet areaLightParams;
        std::string areaLight;
        bool reverseOrientation = false;
    };
    enum class State { Options, World };
    int materialForShape(const ParamSet &shapeParams);
    int makeMaterial(const std::string &name, const ParamSet &shapeParams, const ParamSet &matParams);
    void fail(const char *what);

    SceneDesc *sc;
    State state = State::Options;
    Xform ctm;
    std::map<std::string, Xform> namedCoordinateSystems;
    GraphicsState gs;
    std::vector<GraphicsState> pushedStates;
    std::vector<Xform> pushedTransforms;
    bool inObject = false;
    bool failed = false;
};

// parser entry points (src/core/parser.h: pbrtParseFile / pbrtParseString)
bool parseFile(const std::string &filename, SceneApi &api, SceneDesc &sc);
bool parseString(const std::string &text, SceneApi &api, SceneDesc &sc);

}  // namespace pg
#endif
