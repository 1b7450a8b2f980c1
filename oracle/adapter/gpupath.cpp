// TEST INFRASTRUCTURE / INTEGRATION PROOF -- compiled only where /root/reference is present
// (oracle/Makefile, target `adapter`), never linked into the product libraries.
//
// GpuPathIntegrator: the reference-side binding of INTEGRATION.md section B, as real code.  It is
// an `Integrator` (core/integrator.h:53-58) that lives INSIDE the reference's own process: the
// reference's parser, API state machine, shape / material / light factories and BVHAccel build the
// Scene exactly as they always do; this class flattens that Scene into the plain arrays of
// include/pbrtgpu.h, renders through the C ABI of libpbrtgpu.so, merges the result into the
// reference's Film (film.cpp:117-130) and lets Film::WriteImage (film.cpp:169-211) finish.
//
// The "one changed line" of api.cpp:1794 (and :1797 for "directlighting") is made by the
// preprocessor: core/api.cpp is compiled from where it lies with CreatePathIntegrator /
// CreateDirectLightingIntegrator renamed to the two factories below, so `Integrator "path"` in a
// scene file reaches this class and PathIntegrator (integrators/path.cpp) is not even linked.
// LinearBVHNode and BVHAccel's arrays are private to accelerators/bvh.cpp, so that file is
// included textually too (option 2: the adapter is a friend of the accelerator);
// the file is compiled with -fno-access-control, which is what "friend" amounts to here.
// Nothing of the reference is copied into this repository.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "pbrtgpu.h"

#include "core/integrator.h"
#include "core/camera.h"
#include "core/paramset.h"
#include "core/sampler.h"
#include "integrators/directlighting.h"
#include "integrators/path.h"
namespace pbrt {
Integrator *CreateGpuPathIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler,
                                    std::shared_ptr<const Camera> camera);
Integrator *CreateGpuDirectLightingIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler,
                                              std::shared_ptr<const Camera> camera);
}  // namespace pbrt
#define CreatePathIntegrator CreateGpuPathIntegrator                      // api.cpp:1794
#define CreateDirectLightingIntegrator CreateGpuDirectLightingIntegrator  // api.cpp:1797
#include "core/api.cpp"
#undef CreatePathIntegrator
#undef CreateDirectLightingIntegrator
#include "accelerators/bvh.cpp"

namespace pbrt {
namespace {

struct Flat {  // owns the arrays a pbrtgpu_scene points into
    std::vector<float> bvhBounds, triP, triN, triUV, triS, planeParams, sphereParams, matParams, lightParams, portalParams;
    std::vector<int32_t> bvhOffset, primShape, primMaterial, primLight, primOrigId, matTag, lightTag, lightPrim, lightStrategy,
        lightPortalBegin, lightNSamples;
    std::vector<uint16_t> bvhNPrims;
    std::vector<uint8_t> bvhAxis, primType, primFlags;
    float worldCenter[3], worldRadius;
};

// Leaving from inside Render(): the reference's worker threads are parked in ParallelFor's pool (parallel.cpp) and
// static destructors would wait for them; flush and leave without running them.
[[noreturn]] void leave(int code) {
    fflush(stdout);
    fflush(stderr);
    _exit(code);
}

bool gFailed = false;
void fail(const char *what) {
    Error("GPU path integrator: %s", what);
    gFailed = true;
}

void put16(float *dst, const Matrix4x4 &m) { memcpy(dst, &m.m[0][0], 16 * sizeof(float)); }

// Textures on this path are constants (textures/constant.h); anything else is refused, as pbrthost refuses it.
template <typename T>
T constantOf(const std::shared_ptr<Texture<T>> &tex, const char *what) {
    auto c = dynamic_cast<const ConstantTexture<T> *>(tex.get());
    if (!c) {
        fail((std::string("texture for \"") + what + "\" is not a constant").c_str());
        return T(0.f);
    }
    SurfaceInteraction si;
    return c->Evaluate(si);
}
void putRGB(float *dst, const Spectrum &s) {
    Float rgb[3];
    s.ToRGB(rgb);
    dst[0] = rgb[0], dst[1] = rgb[1], dst[2] = rgb[2];
}
void putClamped(float *dst, const std::shared_ptr<Texture<Spectrum>> &tex, const char *what) { putRGB(dst, constantOf(tex, what).Clamp()); }

// One Material -> (tag, parameter block) of pbrtgpu.h.  The per-hit arithmetic of the reference's
// ComputeScatteringFunctions that depends only on the (constant) textures is done here once:
// Clamp(), RoughnessToAlpha, the Cauchy coefficients (materials/*.cpp).
int flattenMaterial(const Material *m, Flat &f, std::map<const Material *, int> &seen) {
    if (!m) return -1;
    auto it = seen.find(m);
    if (it != seen.end()) return it->second;
    float p[PBRTGPU_MAT_STRIDE] = {0};
    int tag = PBRTGPU_MAT_NONE;
    auto alpha = [](Float r) { return TrowbridgeReitzDistribution::RoughnessToAlpha(r); };
    if (auto mm = dynamic_cast<const MatteMaterial *>(m)) {  // matte.cpp:45-62
        tag = PBRTGPU_MAT_MATTE;
        putClamped(p, mm->Kd, "Kd");
        p[3] = Clamp(constantOf(mm->sigma, "sigma"), 0, 90);
        if (mm->bumpMap) fail("bump mapping is not supported");
    } else if (auto pm = dynamic_cast<const PlasticMaterial *>(m)) {  // plastic.cpp:45-70
        tag = PBRTGPU_MAT_PLASTIC;
        putClamped(p, pm->Kd, "Kd");
        putClamped(p + 3, pm->Ks, "Ks");
        Float rough = constantOf(pm->roughness, "roughness");
        p[6] = pm->remapRoughness ? alpha(rough) : rough;
    } else if (auto dg = dynamic_cast<const DispersiveGlassMaterial *>(m)) {  // dispersive_glass.cpp:48-123
        tag = PBRTGPU_MAT_DISPERSIVE_GLASS;
        putClamped(p, dg->Kr, "Kr");
        putClamped(p + 3, dg->Kt, "Kt");
        Float ur = constantOf(dg->uRoughness, "uroughness"), vr = constantOf(dg->vRoughness, "vroughness");
        Float etaMin = constantOf(dg->indexMin, "etaMin"), etaMax = constantOf(dg->indexMax, "etaMax");
        bool isSpecular = ur == 0 && vr == 0;
        if (!isSpecular && dg->remapRoughness) ur = alpha(ur), vr = alpha(vr);
        const Float lambdaMinSq = sampledLambdaStart * sampledLambdaStart, lambdaMaxSq = sampledLambdaEnd * sampledLambdaEnd;
        const Float cauchyB = (lambdaMinSq * etaMax - lambdaMaxSq * etaMin) / (lambdaMinSq - lambdaMaxSq);
        const Float cauchyC = lambdaMinSq * (etaMax - cauchyB);
        p[6] = ur, p[7] = vr, p[8] = cauchyB, p[9] = isSpecular ? 1.f : 0.f, p[10] = cauchyC, p[11] = etaMin, p[12] = etaMax;
    } else if (auto gm = dynamic_cast<const GlassMaterial *>(m)) {  // glass.cpp:45-92
        tag = PBRTGPU_MAT_GLASS;
        putClamped(p, gm->Kr, "Kr");
        putClamped(p + 3, gm->Kt, "Kt");
        Float ur = constantOf(gm->uRoughness, "uroughness"), vr = constantOf(gm->vRoughness, "vroughness");
        bool isSpecular = ur == 0 && vr == 0;
        if (!isSpecular && gm->remapRoughness) ur = alpha(ur), vr = alpha(vr);
        p[6] = ur, p[7] = vr, p[8] = constantOf(gm->index, "eta"), p[9] = isSpecular ? 1.f : 0.f;
    } else if (auto mr = dynamic_cast<const MirrorMaterial *>(m)) {  // mirror.cpp:45-56
        tag = PBRTGPU_MAT_MIRROR;
        putClamped(p, mr->Kr, "Kr");
    } else if (auto mt = dynamic_cast<const MetalMaterial *>(m)) {  // metal.cpp:60-82
        tag = PBRTGPU_MAT_METAL;
        putRGB(p, constantOf(mt->eta, "eta"));
        putRGB(p + 3, constantOf(mt->k, "k"));
        Float ur = mt->uRoughness ? constantOf(mt->uRoughness, "uroughness") : constantOf(mt->roughness, "roughness");
        Float vr = mt->vRoughness ? constantOf(mt->vRoughness, "vroughness") : constantOf(mt->roughness, "roughness");
        if (mt->remapRoughness) ur = alpha(ur), vr = alpha(vr);
        p[6] = ur, p[7] = vr;
    } else if (auto ub = dynamic_cast<const UberMaterial *>(m)) {  // uber.cpp:45-105
        tag = PBRTGPU_MAT_UBER;
        putClamped(p, ub->Kd, "Kd");
        putClamped(p + 3, ub->Ks, "Ks");
        putClamped(p + 6, ub->Kr, "Kr");
        putClamped(p + 9, ub->Kt, "Kt");
        putClamped(p + 12, ub->opacity, "opacity");
        Float ru = ub->roughnessu ? constantOf(ub->roughnessu, "uroughness") : constantOf(ub->roughness, "roughness");
        Float rv = ub->roughnessv ? constantOf(ub->roughnessv, "vroughness") : ru;
        if (ub->remapRoughness) ru = alpha(ru), rv = alpha(rv);
        p[15] = ru, p[16] = rv, p[17] = constantOf(ub->eta, "eta");
    } else {
        fail("a material of this scene is not one the GPU path implements (matte, plastic, glass, dispersive_glass, mirror, metal, uber)");
        return -1;
    }
    int idx = (int)f.matTag.size();
    f.matTag.push_back(tag);
    f.matParams.insert(f.matParams.end(), p, p + PBRTGPU_MAT_STRIDE);
    seen[m] = idx;
    return idx;
}

void putPlaneRecord(std::vector<float> &dst, const AAPlaneShape &pl, int stride, bool withTransforms) {
    std::vector<float> rec(stride, 0.f);
    rec[0] = pl.lo.x, rec[1] = pl.lo.y, rec[2] = pl.lo.z, rec[3] = pl.hi.x, rec[4] = pl.hi.y, rec[5] = pl.hi.z;
    rec[6] = (float)pl.ax, rec[7] = pl.facingFw ? 1.f : 0.f, rec[8] = pl.area;
    if (withTransforms) {
        put16(&rec[16], pl.ObjectToWorld->m);
        put16(&rec[32], pl.WorldToObject->m);
    }
    dst.insert(dst.end(), rec.begin(), rec.end());
}

int countNodes(const LinearBVHNode *nodes) {  // DFS order, first child at i+1 (bvh.cpp:640-658)
    int maxIdx = 0;
    std::vector<int> stack{0};
    while (!stack.empty()) {
        int i = stack.back();
        stack.pop_back();
        maxIdx = std::max(maxIdx, i);
        if (nodes[i].nPrimitives == 0) {
            stack.push_back(i + 1);
            stack.push_back(nodes[i].secondChildOffset);
        }
    }
    return maxIdx + 1;
}

// Scene -> pbrtgpu_scene: LinearBVHNode[] as BVHAccel left it, primitives in BVHAccel::primitives order.
bool flatten(const Scene &scene, Flat &f, pbrtgpu_scene *out) {
    auto bvh = dynamic_cast<const BVHAccel *>(scene.aggregate.get());
    if (!bvh) {
        fail("the scene's aggregate is not Accelerator \"bvh\"");
        return false;
    }
    if (!bvh->nodes || bvh->primitives.empty()) {
        fail("empty scene");
        return false;
    }
    int nNodes = countNodes(bvh->nodes);
    for (int i = 0; i < nNodes; ++i) {
        const LinearBVHNode &n = bvh->nodes[i];
        const float b[6] = {n.bounds.pMin.x, n.bounds.pMin.y, n.bounds.pMin.z, n.bounds.pMax.x, n.bounds.pMax.y, n.bounds.pMax.z};
        f.bvhBounds.insert(f.bvhBounds.end(), b, b + 6);
        f.bvhOffset.push_back(n.primitivesOffset);
        f.bvhNPrims.push_back(n.nPrimitives);
        f.bvhAxis.push_back(n.axis);
    }

    std::map<const Light *, int> lightIndex;
    for (size_t i = 0; i < scene.lights.size(); ++i) lightIndex[scene.lights[i].get()] = (int)i;
    std::map<const Light *, int> slotOfLight;
    std::map<const Material *, int> matIndex;
    std::map<const Shape *, int> planeIndex, sphereIndex;

    size_t n = bvh->primitives.size();
    bool anyN = false, anyUV = false, anyS = false;
    for (auto &p : bvh->primitives)
        if (auto gp = dynamic_cast<const GeometricPrimitive *>(p.get()))
            if (auto tri = dynamic_cast<const Triangle *>(gp->shape.get()))
                anyN |= tri->mesh->n != nullptr, anyUV |= tri->mesh->uv != nullptr, anyS |= tri->mesh->s != nullptr;
    f.triP.assign(9 * n, 0.f);
    if (anyN) f.triN.assign(9 * n, 0.f);
    if (anyUV) f.triUV.assign(6 * n, 0.f);
    if (anyS) f.triS.assign(9 * n, 0.f);
    for (size_t slot = 0; slot < n; ++slot) {
        auto gp = dynamic_cast<const GeometricPrimitive *>(bvh->primitives[slot].get());
        if (!gp) {
            fail("object instances / animated primitives are not supported");
            return false;
        }
        if (gp->mediumInterface.inside || gp->mediumInterface.outside) {
            fail("participating media are not supported");
            return false;
        }
        const Shape *sh = gp->shape.get();
        uint8_t flags = (sh->reverseOrientation ? PBRTGPU_PF_REVERSE_ORIENTATION : 0) |
                        (sh->transformSwapsHandedness ? PBRTGPU_PF_SWAPS_HANDEDNESS : 0);
        int type, shapeIdx = 0;
        if (auto tri = dynamic_cast<const Triangle *>(sh)) {
            type = PBRTGPU_PRIM_TRIANGLE;
            const TriangleMesh &m = *tri->mesh;
            if (m.alphaMask || m.shadowAlphaMask) {
                fail("alpha-masked triangle meshes are not supported");
                return false;
            }
            for (int k = 0; k < 3; ++k) {
                int vi = tri->v[k];
                f.triP[9 * slot + 3 * k] = m.p[vi].x, f.triP[9 * slot + 3 * k + 1] = m.p[vi].y, f.triP[9 * slot + 3 * k + 2] = m.p[vi].z;
                if (m.n) f.triN[9 * slot + 3 * k] = m.n[vi].x, f.triN[9 * slot + 3 * k + 1] = m.n[vi].y, f.triN[9 * slot + 3 * k + 2] = m.n[vi].z;
                if (m.s) f.triS[9 * slot + 3 * k] = m.s[vi].x, f.triS[9 * slot + 3 * k + 1] = m.s[vi].y, f.triS[9 * slot + 3 * k + 2] = m.s[vi].z;
                if (m.uv) f.triUV[6 * slot + 2 * k] = m.uv[vi].x, f.triUV[6 * slot + 2 * k + 1] = m.uv[vi].y;
            }
            if (m.n) flags |= PBRTGPU_PF_HAS_NORMALS;
            if (m.uv) flags |= PBRTGPU_PF_HAS_UV;
            if (m.s) flags |= PBRTGPU_PF_HAS_TANGENTS;
        } else if (auto pl = dynamic_cast<const AAPlaneShape *>(sh)) {
            type = PBRTGPU_PRIM_AAPLANE;
            auto it = planeIndex.find(sh);
            if (it == planeIndex.end()) {
                it = planeIndex.emplace(sh, (int)planeIndex.size()).first;
                putPlaneRecord(f.planeParams, *pl, PBRTGPU_PLANE_STRIDE, true);
            }
            shapeIdx = it->second;
        } else if (auto sp = dynamic_cast<const Sphere *>(sh)) {
            type = PBRTGPU_PRIM_SPHERE;
            auto it = sphereIndex.find(sh);
            if (it == sphereIndex.end()) {
                it = sphereIndex.emplace(sh, (int)sphereIndex.size()).first;
                float rec[PBRTGPU_SPHERE_STRIDE] = {0};
                rec[0] = sp->radius, rec[1] = sp->zMin, rec[2] = sp->zMax, rec[3] = sp->thetaMin, rec[4] = sp->thetaMax, rec[5] = sp->phiMax;
                put16(rec + 16, sp->ObjectToWorld->m);
                put16(rec + 32, sp->WorldToObject->m);
                f.sphereParams.insert(f.sphereParams.end(), rec, rec + PBRTGPU_SPHERE_STRIDE);
            }
            shapeIdx = it->second;
        } else {
            fail("a shape of this scene is not a triangle, aaplane or sphere");
            return false;
        }
        f.primType.push_back((uint8_t)type);
        f.primFlags.push_back(flags);
        f.primShape.push_back(shapeIdx);
        f.primMaterial.push_back(flattenMaterial(gp->material.get(), f, matIndex));
        int li = -1;
        if (gp->areaLight) {
            auto it = lightIndex.find(gp->areaLight.get());
            if (it != lightIndex.end()) li = it->second, slotOfLight[gp->areaLight.get()] = (int)slot;
        }
        f.primLight.push_back(li);
        f.primOrigId.push_back((int32_t)slot);  // creation order is gone by now (api.cpp:1763 moves the list); the render does not read it
    }

    f.lightPortalBegin.push_back(0);
    for (auto &lp : scene.lights) {
        const Light *l = lp.get();
        float p[PBRTGPU_LIGHT_STRIDE] = {0};
        int tag = 0, strategy = PBRTGPU_PORTAL_UNIFORM_LIGHT, prim = -1;
        if (auto dl = dynamic_cast<const DiffuseAreaLight *>(l)) {  // diffuse.cpp:45-60; PortalArealight derives from it
            tag = PBRTGPU_LIGHT_DIFFUSE_AREA;
            putRGB(p, dl->Lemit);
            p[3] = dl->twoSided ? 1.f : 0.f;
            p[4] = dl->area;
            auto it = slotOfLight.find(l);
            if (it == slotOfLight.end()) {
                fail("an area light without a primitive in the accelerator");
                return false;
            }
            prim = it->second;
            if (auto pa = dynamic_cast<const PortalArealight *>(l)) {  // portal_arealight.cpp:12-27
                tag = PBRTGPU_LIGHT_PORTAL_AREA;
                strategy = (int)pa->strat;  // same numbering as PBRTGPU_PORTAL_*
                for (const AAPortal &po : pa->portals) putPlaneRecord(f.portalParams, po.portal, PBRTGPU_PORTAL_STRIDE, false);
            }
        } else if (auto il = dynamic_cast<const InfiniteAreaLight *>(l)) {  // infinite.cpp:45-107 without a map: one texel
            if (il->Lmap->Width() != 1 || il->Lmap->Height() != 1) {
                fail("environment maps (\"mapname\") are not supported");
                return false;
            }
            tag = PBRTGPU_LIGHT_INFINITE;
            putRGB(p, Spectrum(il->Lmap->Texel(0, 0, 0), SpectrumType::Illuminant));
            put16(p + 8, il->LightToWorld.m);
            put16(p + 24, il->WorldToLight.m);
        } else if (auto sl = dynamic_cast<const SpotLight *>(l)) {  // spot.cpp:45-60
            tag = PBRTGPU_LIGHT_SPOT;
            putRGB(p, sl->I);
            p[3] = sl->pLight.x, p[4] = sl->pLight.y, p[5] = sl->pLight.z, p[6] = sl->cosTotalWidth, p[7] = sl->cosFalloffStart;
            put16(p + 8, sl->LightToWorld.m);
            put16(p + 24, sl->WorldToLight.m);
        } else if (auto pt = dynamic_cast<const PointLight *>(l)) {  // point.h:50-57
            tag = PBRTGPU_LIGHT_POINT;
            putRGB(p, pt->I);
            p[3] = pt->pLight.x, p[4] = pt->pLight.y, p[5] = pt->pLight.z;
        } else if (auto di = dynamic_cast<const DistantLight *>(l)) {  // distant.cpp:45-50
            tag = PBRTGPU_LIGHT_DISTANT;
            putRGB(p, di->L);
            p[3] = di->wLight.x, p[4] = di->wLight.y, p[5] = di->wLight.z;
        } else {
            fail("a light of this scene is not one the GPU path implements (area, portal, infinite, point, spot, distant)");
            return false;
        }
        f.lightTag.push_back(tag);
        f.lightPrim.push_back(prim);
        f.lightStrategy.push_back(strategy);
        f.lightNSamples.push_back(l->nSamples);
        f.lightParams.insert(f.lightParams.end(), p, p + PBRTGPU_LIGHT_STRIDE);
        f.lightPortalBegin.push_back((int32_t)(f.portalParams.size() / PBRTGPU_PORTAL_STRIDE));
    }
    if (gFailed) return false;

    Point3f c;
    Float r;
    scene.WorldBound().BoundingSphere(&c, &r);  // infinite.cpp:68-70, distant.cpp Preprocess
    f.worldCenter[0] = c.x, f.worldCenter[1] = c.y, f.worldCenter[2] = c.z, f.worldRadius = r;

    pbrtgpu_scene &s = *out;
    memset(&s, 0, sizeof(s));
    s.n_nodes = nNodes;
    s.bvh_bounds = f.bvhBounds.data(), s.bvh_offset = f.bvhOffset.data(), s.bvh_nprims = f.bvhNPrims.data(), s.bvh_axis = f.bvhAxis.data();
    s.n_prims = (int32_t)n;
    s.prim_type = f.primType.data(), s.prim_flags = f.primFlags.data(), s.prim_shape = f.primShape.data();
    s.prim_material = f.primMaterial.data(), s.prim_light = f.primLight.data(), s.prim_orig_id = f.primOrigId.data();
    s.tri_p = f.triP.data();
    s.tri_n = anyN ? f.triN.data() : nullptr, s.tri_uv = anyUV ? f.triUV.data() : nullptr, s.tri_s = anyS ? f.triS.data() : nullptr;
    s.n_planes = (int32_t)planeIndex.size(), s.plane_params = f.planeParams.data();
    s.n_spheres = (int32_t)sphereIndex.size(), s.sphere_params = f.sphereParams.data();
    s.n_materials = (int32_t)f.matTag.size(), s.mat_tag = f.matTag.data(), s.mat_params = f.matParams.data();
    s.n_lights = (int32_t)f.lightTag.size();
    s.light_tag = f.lightTag.data(), s.light_prim = f.lightPrim.data(), s.light_strategy = f.lightStrategy.data();
    s.light_portal_begin = f.lightPortalBegin.data(), s.light_params = f.lightParams.data(), s.light_nsamples = f.lightNSamples.data();
    s.n_portals = (int32_t)(f.portalParams.size() / PBRTGPU_PORTAL_STRIDE), s.portal_params = f.portalParams.data();
    memcpy(s.world_center, f.worldCenter, sizeof(f.worldCenter));
    s.world_radius = f.worldRadius;
    return true;
}

// PBRTGPU_ADAPTER_DUMP=<file>: write what would cross the C ABI as "name dtype count\n" + raw bytes records, so that a test
// without a GPU can compare it with what pbrthost makes of the same scene file (tests/test_adapter.py).
void dumpArray(FILE *fp, const char *name, const char *dtype, const void *data, size_t count, size_t elemSize) {
    fprintf(fp, "%s %s %zu\n", name, dtype, data ? count : (size_t)0);
    if (data && count) fwrite(data, elemSize, count, fp);
}
void dumpScene(const char *path, const pbrtgpu_scene &s, const pbrtgpu_render_params &rp) {
    FILE *fp = fopen(path, "wb");
    if (!fp) return;
    size_t nn = s.n_nodes, np = s.n_prims, nl = s.n_lights;
#define DUMP(field, dtype, count) dumpArray(fp, #field, dtype, s.field, count, sizeof(*s.field))
    DUMP(bvh_bounds, "f4", 6 * nn); DUMP(bvh_offset, "i4", nn); DUMP(bvh_nprims, "u2", nn); DUMP(bvh_axis, "u1", nn);
    DUMP(prim_type, "u1", np); DUMP(prim_flags, "u1", np); DUMP(prim_shape, "i4", np); DUMP(prim_material, "i4", np);
    DUMP(prim_light, "i4", np); DUMP(tri_p, "f4", 9 * np); DUMP(tri_n, "f4", 9 * np); DUMP(tri_uv, "f4", 6 * np); DUMP(tri_s, "f4", 9 * np);
    DUMP(plane_params, "f4", (size_t)s.n_planes * PBRTGPU_PLANE_STRIDE); DUMP(sphere_params, "f4", (size_t)s.n_spheres * PBRTGPU_SPHERE_STRIDE);
    DUMP(mat_tag, "i4", (size_t)s.n_materials); DUMP(mat_params, "f4", (size_t)s.n_materials * PBRTGPU_MAT_STRIDE);
    DUMP(light_tag, "i4", nl); DUMP(light_prim, "i4", nl); DUMP(light_strategy, "i4", nl); DUMP(light_portal_begin, "i4", nl + 1);
    DUMP(light_params, "f4", nl * PBRTGPU_LIGHT_STRIDE); DUMP(light_nsamples, "i4", nl);
    DUMP(portal_params, "f4", (size_t)s.n_portals * PBRTGPU_PORTAL_STRIDE);
#undef DUMP
    float world[4] = {s.world_center[0], s.world_center[1], s.world_center[2], s.world_radius};
    dumpArray(fp, "world", "f4", world, 4, 4);
    dumpArray(fp, "render_params", "u1", &rp, sizeof(rp), 1);
    fclose(fp);
}

}  // namespace

class GpuPathIntegrator : public Integrator {
  public:
    GpuPathIntegrator(int integrator, int directStrategy, int maxDepth, std::shared_ptr<const Camera> camera,
                      std::shared_ptr<Sampler> sampler, const Bounds2i &pixelBounds, Float rrThreshold, int lightStrategy)
        : integrator(integrator), directStrategy(directStrategy), maxDepth(maxDepth), camera(camera), sampler(sampler),
          pixelBounds(pixelBounds), rrThreshold(rrThreshold), lightStrategy(lightStrategy) {}

    // Camera / Film / Sampler / integrator parameters -> pbrtgpu_render_params.
    bool params(pbrtgpu_render_params *out) const {
        pbrtgpu_render_params &rp = *out;
        memset(&rp, 0, sizeof(rp));
        auto pc = dynamic_cast<const PerspectiveCamera *>(camera.get());
        if (!pc) return fail("Camera is not \"perspective\""), false;
        if (pc->CameraToWorld.actuallyAnimated) return fail("animated cameras are not supported"), false;
        if (pc->medium) return fail("participating media are not supported"), false;
        auto hs = dynamic_cast<const HaltonSampler *>(sampler.get());
        if (!hs) return fail("Sampler is not \"halton\""), false;
        put16(rp.camera_to_world, pc->CameraToWorld.startTransform->m);
        put16(rp.raster_to_camera, pc->RasterToCamera.m);
        rp.lens_radius = pc->lensRadius, rp.focal_distance = pc->focalDistance;
        rp.shutter_open = pc->shutterOpen, rp.shutter_close = pc->shutterClose;
        const Film &film = *camera->film;
        rp.full_resolution[0] = film.fullResolution.x, rp.full_resolution[1] = film.fullResolution.y;
        auto putB = [](int32_t *dst, const Bounds2i &b) { dst[0] = b.pMin.x, dst[1] = b.pMin.y, dst[2] = b.pMax.x, dst[3] = b.pMax.y; };
        putB(rp.cropped_bounds, film.croppedPixelBounds);
        putB(rp.sample_bounds, film.GetSampleBounds());
        putB(rp.pixel_bounds, pixelBounds);
        rp.filter_radius[0] = film.filter->radius.x, rp.filter_radius[1] = film.filter->radius.y;
        static_assert(Film::filterTableWidth == 16, "pbrtgpu_render_params::filter_table is 16 x 16");
        memcpy(rp.filter_table, film.filterTable, sizeof(rp.filter_table));
        rp.max_sample_luminance = film.maxSampleLuminance;
        rp.spp = (int32_t)hs->samplesPerPixel;
        rp.sample_at_pixel_center = hs->sampleAtPixelCenter ? 1 : 0;
        rp.max_depth = maxDepth, rp.rr_threshold = rrThreshold, rp.light_strategy = lightStrategy;
        rp.shard_index = 0, rp.shard_count = 1;
        rp.integrator = integrator, rp.direct_strategy = directStrategy;
        return true;
    }

    void Render(const Scene &scene) override {
        Flat flat;
        pbrtgpu_scene s;
        pbrtgpu_render_params rp;
        if (!flatten(scene, flat, &s) || !params(&rp)) leave(3);
        if (integrator == PBRTGPU_INTEGRATOR_DIRECT)  // what the device's tree walk cannot reproduce, as pbrthost refuses it
            for (size_t i = 0; i < flat.matTag.size(); ++i) {
                const float *mp = &flat.matParams[i * PBRTGPU_MAT_STRIDE];
                bool twoTransLobes = flat.matTag[i] == PBRTGPU_MAT_UBER && (mp[9] != 0 || mp[10] != 0 || mp[11] != 0) &&
                                     (mp[12] < 1 || mp[13] < 1 || mp[14] < 1);
                if (flat.matTag[i] == PBRTGPU_MAT_DISPERSIVE_GLASS || twoTransLobes) {
                    fail("Integrator \"directlighting\" with dispersive_glass, or uber with both Kt and opacity < 1, is not supported");
                    leave(3);
                }
            }
        if (const char *dump = getenv("PBRTGPU_ADAPTER_DUMP")) dumpScene(dump, s, rp);
        pbrtgpu_ctx *ctx = nullptr;
        const char *dev = getenv("PBRTGPU_DEVICE");
        if (pbrtgpu_create(dev ? atoi(dev) : 0, &ctx) != PBRTGPU_OK) {  // no CPU fallback: fail loudly
            Error("GPU path integrator: %s", pbrtgpu_last_error(nullptr));
            leave(2);
        }
        Film *film = camera->film;
        Bounds2i cb = film->croppedPixelBounds;
        size_t nPix = (size_t)cb.Area();
        std::vector<float> xyzw(4 * nPix);
        pbrtgpu_stats st;
        if (pbrtgpu_upload_scene(ctx, &s) != PBRTGPU_OK || pbrtgpu_render(ctx, &rp, xyzw.data(), &st) != PBRTGPU_OK) {
            Error("GPU path integrator: %s", pbrtgpu_last_error(ctx));
            leave(3);
        }
        // Film::MergeFilmTile (film.cpp:117-130) for the whole frame: the device returns what Film::Pixel accumulates
        for (size_t i = 0; i < nPix; ++i) {
            Film::Pixel &px = film->pixels[i];
            for (int c = 0; c < 3; ++c) px.xyz[c] += xyzw[4 * i + c];
            px.filterWeightSum += xyzw[4 * i + 3];
        }
        if (!PbrtOptions.quiet)
            printf("GPU path integrator: %lld camera rays, %lld closest-hit rays, %lld shadow rays, %.1f ms on the device\n",
                   (long long)st.camera_rays, (long long)st.closest_rays, (long long)st.shadow_rays, st.device_ms);
        pbrtgpu_destroy(ctx);
        film->WriteImage();  // film.cpp:169-211, unchanged
    }

  private:
    const int integrator, directStrategy, maxDepth;
    std::shared_ptr<const Camera> camera;
    std::shared_ptr<Sampler> sampler;
    const Bounds2i pixelBounds;
    const Float rrThreshold;
    const int lightStrategy;
};

static Bounds2i pixelBoundsParam(const ParamSet &params, const Camera &camera) {  // path.cpp:195-208, directlighting.cpp:116-130
    int np;
    const int *pb = params.FindInt("pixelbounds", &np);
    Bounds2i pixelBounds = camera.film->GetSampleBounds();
    if (pb) {
        if (np != 4)
            Error("Expected four values for \"pixelbounds\" parameter. Got %d.", np);
        else {
            pixelBounds = Intersect(pixelBounds, Bounds2i{{pb[0], pb[2]}, {pb[1], pb[3]}});
            if (pixelBounds.Area() == 0) Error("Degenerate \"pixelbounds\" specified.");
        }
    }
    return pixelBounds;
}

// Same ParamSet, same defaults, same messages as CreatePathIntegrator (path.cpp:191-214).
Integrator *CreateGpuPathIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler, std::shared_ptr<const Camera> camera) {
    int maxDepth = params.FindOneInt("maxdepth", 5);
    Bounds2i pixelBounds = pixelBoundsParam(params, *camera);
    Float rrThreshold = params.FindOneFloat("rrthreshold", 1.);
    std::string ls = params.FindOneString("lightsamplestrategy", "uniform");
    int lightStrategy = 0;
    if (ls == "power") lightStrategy = 1;
    else if (ls != "uniform")
        Error("Light sample distribution type \"%s\" is not supported by the GPU path integrator. Using \"uniform\".", ls.c_str());
    return new GpuPathIntegrator(PBRTGPU_INTEGRATOR_PATH, 0, maxDepth, camera, sampler, pixelBounds, rrThreshold, lightStrategy);
}

// CreateDirectLightingIntegrator (directlighting.cpp:100-137).
Integrator *CreateGpuDirectLightingIntegrator(const ParamSet &params, std::shared_ptr<Sampler> sampler,
                                              std::shared_ptr<const Camera> camera) {
    int maxDepth = params.FindOneInt("maxdepth", 5);
    int strategy = PBRTGPU_DIRECT_ALL;
    std::string st = params.FindOneString("strategy", "all");
    if (st == "one") strategy = PBRTGPU_DIRECT_ONE;
    else if (st != "all")
        Warning("Strategy \"%s\" for direct lighting unknown. Using \"all\".", st.c_str());
    Bounds2i pixelBounds = pixelBoundsParam(params, *camera);
    return new GpuPathIntegrator(PBRTGPU_INTEGRATOR_DIRECT, strategy, maxDepth, camera, sampler, pixelBounds, 1.f, 0);
}

}  // namespace pbrt
