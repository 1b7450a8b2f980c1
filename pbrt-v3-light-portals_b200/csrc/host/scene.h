// Host-side scene description: what the .pbrt front end produces (SceneDesc,
// creation order, one entry per reference object) and the flattened form that
// crosses the C ABI (FlatScene, BVH leaf order, plain arrays).
#ifndef PG_HOST_SCENE_H
#define PG_HOST_SCENE_H

#include <map>
#include <memory>
#include <string>
#include <vector>

#include "../../../include/pbrtgpu.h"
#include "../common/pg_math.h"
#include "params.h"

namespace pg {

struct Bounds3 {
    V3 pMin = v3(3.402823466e+38f), pMax = v3(-3.402823466e+38f);  // Bounds3 default ctor, geometry.h:870-875
    V3 diagonal() const { return pMax - pMin; }
    int maximumExtent() const {  // geometry.h:938-946
        V3 d = diagonal();
        if (d.x > d.y && d.x > d.z) return 0;
        if (d.y > d.z) return 1;
        return 2;
    }
    float surfaceArea() const {  // geometry.h:930-933
        V3 d = diagonal();
        return 2 * (d.x * d.y + d.x * d.z + d.y * d.z);
    }
    V3 offset(V3 p) const {  // geometry.h:953-959
        V3 o = p - pMin;
        if (pMax.x > pMin.x) o.x /= pMax.x - pMin.x;
        if (pMax.y > pMin.y) o.y /= pMax.y - pMin.y;
        if (pMax.z > pMin.z) o.z /= pMax.z - pMin.z;
        return o;
    }
};
inline Bounds3 unionB(const Bounds3 &b, V3 p) {
    Bounds3 r;
    r.pMin = vmin(b.pMin, p);
    r.pMax = vmax(b.pMax, p);
    return r;
}
inline Bounds3 unionB(const Bounds3 &a, const Bounds3 &b) {
    Bounds3 r;
    r.pMin = vmin(a.pMin, b.pMin);
    r.pMax = vmax(a.pMax, b.pMax);
    return r;
}
inline Bounds3 boundsOf(V3 a, V3 b) {  // Bounds3(p1, p2), geometry.h:877-881
    Bounds3 r;
    r.pMin = vmin(a, b);
    r.pMax = vmax(a, b);
    return r;
}

// One TriangleMesh (shapes/triangle.h:52-70): vertices already in world space.
struct MeshData {
    std::vector<V3> p, n, s;
    std::vector<float> uv;  // 2 per vertex, empty if none
    std::vector<int> indices;
    bool reverseOrientation = false, swapsHandedness = false;
};

struct PlaneData {  // AAPlaneShape, shapes/plane.h:12-40
    V3 lo, hi;
    int ax = 2;
    bool facingFw = true;  // = !reverseOrientation (plane.h:24)
    float area = 0;
    bool reverseOrientation = false, swapsHandedness = false;
    Xform o2w;  // m = ObjectToWorld, mInv = WorldToObject
};

struct SphereData {  // Sphere, shapes/sphere.h
    float radius, zMin, zMax, thetaMin, thetaMax, phiMax;
    bool reverseOrientation = false, swapsHandedness = false;
    Xform o2w;
};

struct PrimDesc {  // one GeometricPrimitive, in creation order
    uint8_t type = PBRTGPU_PRIM_TRIANGLE;
    int shape = 0;  // mesh index | plane index | sphere index
    int tri = 0;    // triangle number inside the mesh
    int material = -1;
    int light = -1;
};

struct MaterialData {
    int tag = PBRTGPU_MAT_NONE;
    float params[PBRTGPU_MAT_STRIDE] = {0};
};

struct PortalData {  // AAPortal's own AAPlaneShape (portals/aaportal.cpp:8-13)
    V3 lo, hi;
    int ax;
    bool facingFw;
    float area;
};

struct LightData {
    int tag = 0;
    int prim = -1;  // creation-order primitive carrying this area light
    int strategy = PBRTGPU_PORTAL_UNIFORM_LIGHT;
    int nSamples = 1;  // Light::nSamples (light.cpp:52: max(1, nSamples))
    float params[PBRTGPU_LIGHT_STRIDE] = {0};
    std::vector<PortalData> portals;
};

struct SceneDesc {
    // render options (RenderOptions, api.cpp:164-191)
    std::string filterName = "box", filmName = "image", samplerName = "halton",
                acceleratorName = "bvh", integratorName = "path", cameraName = "perspective";
    ParamSet filterParams, filmParams, samplerParams, acceleratorParams, integratorParams, cameraParams;
    Xform cameraToWorld;
    // world
    std::vector<MeshData> meshes;
    std::vector<PlaneData> planes;
    std::vector<SphereData> spheres;
    std::vector<PrimDesc> prims;
    std::vector<MaterialData> materials;
    std::vector<LightData> lights;
    bool worldEnded = false;
    std::string sceneDir;  // for relative Include / plymesh paths
};

// Flattened arrays in the layout of pbrtgpu_scene.
struct FlatScene {
    std::vector<float> bvhBounds;
    std::vector<int32_t> bvhOffset;
    std::vector<uint16_t> bvhNPrims;
    std::vector<uint8_t> bvhAxis;
    std::vector<uint8_t> primType, primFlags;
    std::vector<int32_t> primShape, primMaterial, primLight, primOrigId;
    std::vector<float> triP, triN, triUV, triS;
    std::vector<float> planeParams, sphereParams;
    std::vector<int32_t> matTag;
    std::vector<float> matParams;
    std::vector<int32_t> lightTag, lightPrim, lightStrategy, lightPortalBegin, lightNSamples;
    std::vector<float> lightParams, portalParams;
    float worldCenter[3] = {0, 0, 0}, worldRadius = 0;
    Bounds3 worldBound;
    pbrtgpu_scene view() const;
};

// BVH build (accelerators/bvh.cpp:183-225, 236-402, 640-658)
enum class SplitMethod { SAH, HLBVH, Middle, EqualCounts };
struct BuiltBVH {
    std::vector<float> bounds;  // 6 per node
    std::vector<int32_t> offset;
    std::vector<uint16_t> nPrims;
    std::vector<uint8_t> axis;
    std::vector<int> orderedPrims;  // creation-order prim number per BVH slot
};
BuiltBVH buildBVH(const std::vector<Bounds3> &primBounds, int maxPrimsInNode, SplitMethod method);

Bounds3 primWorldBound(const SceneDesc &s, const PrimDesc &p);
bool flattenScene(const SceneDesc &desc, FlatScene *out, std::string *err);

}  // namespace pg
#endif
