// Host-side construction of the device BVH / primitive layout described in
// bvh.h from the reference-layout arrays of pbrtgpu_scene.  Plain C++ (no CUDA
// calls) so that the upload path and the CPU test harness build the very same
// bytes.
#ifndef PG_DEVICE_LAYOUT_BUILD_H
#define PG_DEVICE_LAYOUT_BUILD_H

#include <cstring>
#include <string>
#include <vector>

#include "../../../include/pbrtgpu.h"
#include "bvh.h"
#include "lights.h"

namespace pg {

// Is this triangle one that Triangle::Intersect rejects after the hit test
// (degenerate partial derivatives AND zero-area geometry, triangle.cpp:294-317)?
// Depends only on the vertices and uvs, so it is decided once here.
PG_HD bool triangleIsDegenerate(const float *p, const float *uvOrNull) {
    V3 p0 = v3(p[0], p[1], p[2]), p1 = v3(p[3], p[4], p[5]), p2 = v3(p[6], p[7], p[8]);
    float uv[3][2] = {{0, 0}, {1, 0}, {1, 1}};  // Triangle::GetUVs default, triangle.h:114-123
    if (uvOrNull)
        for (int i = 0; i < 3; ++i) uv[i][0] = uvOrNull[2 * i], uv[i][1] = uvOrNull[2 * i + 1];
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
        if (lengthSquared(ng) == 0) return true;
    }
    return false;
}

inline float i2f(int i) {
    float f;
    memcpy(&f, &i, 4);
    return f;
}

// LinearBVHNode[] (DFS, first child at i+1) -> interior records of bvh.h.
inline void buildWideNodes(const pbrtgpu_scene *s, std::vector<F4> *out) {
    std::vector<int> wideIndex(s->n_nodes, -1);
    int nInterior = 0;
    for (int i = 0; i < s->n_nodes; ++i)
        if (s->bvh_nprims[i] == 0) wideIndex[i] = 1 + nInterior++;
    out->assign(4 * (size_t)(1 + nInterior), F4{0, 0, 0, 0});
    auto refOf = [&](int node) { return s->bvh_nprims[node] == 0 ? wideIndex[node] : ~s->bvh_offset[node]; };
    auto setChild = [&](int wide, int which, int node) {
        const float *b = s->bvh_bounds + 6 * (size_t)node;
        F4 *q = out->data() + 4 * (size_t)wide;
        if (which == 0) {
            q[0].x = b[0], q[0].y = b[1], q[0].z = b[2], q[0].w = b[3];
            q[1].x = b[4], q[1].y = b[5];
            q[3].x = i2f(refOf(node));
        } else {
            q[1].z = b[0], q[1].w = b[1];
            q[2].x = b[2], q[2].y = b[3], q[2].z = b[4], q[2].w = b[5];
            q[3].y = i2f(refOf(node));
        }
    };
    // record 0: synthetic parent of the root; second child empty
    setChild(0, 0, 0);
    (*out)[3].y = i2f(kEmptyRef);
    (*out)[3].z = i2f(0);
    for (int i = 0; i < s->n_nodes; ++i) {
        if (s->bvh_nprims[i] != 0) continue;
        int w = wideIndex[i];
        setChild(w, 0, i + 1);
        setChild(w, 1, s->bvh_offset[i]);
        (*out)[4 * (size_t)w + 3].z = i2f((int)s->bvh_axis[i]);
    }
}


// (p0,tag) (p1,-) (p2,-) per BVH-order slot; tag = type | degenerate | last-in-leaf | shape index
inline bool buildTris(const pbrtgpu_scene *s, std::vector<F4> *out, std::string *err) {
    std::vector<F4> &tris = *out;
    tris.assign(3 * (size_t)s->n_prims, F4{0, 0, 0, 0});
    for (int i = 0; i < s->n_prims; ++i) {
        const float *p = s->tri_p + 9 * (size_t)i;
        int tag = s->prim_type[i] & kTagTypeMask;
        if (s->prim_type[i] == PBRTGPU_PRIM_TRIANGLE) {
            if (triangleIsDegenerate(p, s->tri_uv && s->prim_flags && (s->prim_flags[i] & PBRTGPU_PF_HAS_UV) ? s->tri_uv + 6 * (size_t)i : nullptr))
                tag |= kTagDegenerate;
        } else {
            int idx = s->prim_shape ? s->prim_shape[i] : 0;
            int limit = s->prim_type[i] == PBRTGPU_PRIM_AAPLANE ? s->n_planes : s->n_spheres;
            if (idx < 0 || idx >= limit) {
                *err = "primitive " + std::to_string(i) + ": shape index out of range";
                return false;
            }
            tag |= idx << kTagShapeShift;
        }
        tris[3 * (size_t)i + 0] = F4{p[0], p[1], p[2], i2f(tag)};
        tris[3 * (size_t)i + 1] = F4{p[3], p[4], p[5], 0.f};
        tris[3 * (size_t)i + 2] = F4{p[6], p[7], p[8], 0.f};
    }
    for (int i = 0; i < s->n_nodes; ++i)
        if (s->bvh_nprims[i] != 0) {
            size_t last = (size_t)s->bvh_offset[i] + s->bvh_nprims[i] - 1;
            int tag;
            memcpy(&tag, &tris[3 * last].w, 4);
            tris[3 * last].w = i2f(tag | kTagLast);
        }
    return true;
}

// InfiniteAreaLight's constructor for a constant light (lights/infinite.cpp:43-82):
// a 1x1 Lmap, a 2x2 scalar image of Lmap->Lookup(st, fwidth).y() * sinTheta and its
// Distribution2D (core/sampling.h:55-70, sampling.cpp Distribution2D ctor).
// Stored into the spare floats of the light record (layout: lights_eval.h).
inline void finishInfiniteLight(float *lp) {
    const int width = 2, height = 2;
    float img[4];
    for (int v = 0; v < height; ++v) {
        float vp = (v + .5f) / (float)height;
        float sinTheta = std::sin(kPi * (v + .5f) / height);
        for (int u = 0; u < width; ++u) {
            float up = (u + .5f) / (float)width;
            Spec t = infiniteLookup(spec(lp[0], lp[1], lp[2]), up, vp);  // MIPMap::Lookup(st, 0.25) -> level < 0 -> triangle(0, st)
            img[u + v * width] = luminance(t);
            img[u + v * width] *= sinTheta;
        }
    }
    auto dist1D = [](const float *func, int n, float *cdf, float *funcInt) {
        cdf[0] = 0;
        for (int i = 1; i < n + 1; ++i) cdf[i] = cdf[i - 1] + func[i - 1] / n;
        *funcInt = cdf[n];
        if (*funcInt == 0)
            for (int i = 1; i < n + 1; ++i) cdf[i] = float(i) / float(n);
        else
            for (int i = 1; i < n + 1; ++i) cdf[i] /= *funcInt;
    };
    float marginalFunc[2];
    for (int v = 0; v < height; ++v) {
        float cdf[3], funcInt;
        dist1D(&img[v * width], width, cdf, &funcInt);
        float *row = lp + kInfRow + 4 * v;
        row[0] = img[v * width], row[1] = img[v * width + 1], row[2] = cdf[1], row[3] = funcInt;
        marginalFunc[v] = funcInt;
    }
    float cdf[3], funcInt;
    dist1D(marginalFunc, height, cdf, &funcInt);
    lp[kInfMargCdf] = cdf[1];
    lp[kInfMargInt] = funcInt;
}


// The per-material BxDF table (bsdf_eval.h:makeMaterialLobes): entries [0, n) with allowMultipleLobes = true
// (path integrator), entries [n, 2n) with false (directlighting).
inline void buildMaterialLobes(const pbrtgpu_scene *s, std::vector<MatLobes> *out) {
    const int n = s->n_materials;
    out->resize(n > 0 ? 2 * n : 2);
    memset(out->data(), 0, out->size() * sizeof(MatLobes));
    for (int i = 0; i < n; ++i) {
        const float *mp = s->mat_params + (size_t)i * PBRTGPU_MAT_STRIDE;
        makeMaterialLobes(s->mat_tag[i], mp, true, &(*out)[i]);
        makeMaterialLobes(s->mat_tag[i], mp, false, &(*out)[n + i]);
        for (MatLobes *m : {&(*out)[i], &(*out)[n + i]}) {
            bool allSpecular = m->nLobes > 0;
            for (int k = 0; k < m->nLobes; ++k) allSpecular = allSpecular && (m->lobes[k].type & BSDF_SPECULAR) != 0;
            m->shadeClass = (m->nLobes == 1 && m->lobes[0].kind == LOBE_LAMBERT) ? 0 : (allSpecular ? 2 : 1);
        }
    }
}
// Specular-transmission lobes per material under allowMultipleLobes = false: 0 everywhere = directlighting's
// recursion is a path; otherwise it is a tree (SpecularReflect + SpecularTransmit, directlighting.cpp:91-95).
inline int maxSpecularTransmissionLobes(const std::vector<MatLobes> &table, int nMaterials, bool *dispersive) {
    int worst = 0;
    *dispersive = false;
    for (int i = 0; i < nMaterials; ++i) {
        const MatLobes &m = table[nMaterials + i];
        int c = 0;
        for (int k = 0; k < m.nLobes; ++k)
            if ((m.lobes[k].type & (BSDF_TRANSMISSION | BSDF_SPECULAR)) == (BSDF_TRANSMISSION | BSDF_SPECULAR)) ++c;
        if (c > 0 && m.dispersive) *dispersive = true;
        worst = c > worst ? c : worst;
    }
    return worst;
}

// PathIntegrator::Preprocess -> CreateLightSampleDistribution (core/lightdistrib.cpp:48-66,
// 68-88; ComputeLightPowerDistribution core/integrator.cpp:44-50): the
// Distribution1D the light is picked from.  strategy 0 = uniform, 1 = power.
inline void buildLightDistribution(const pbrtgpu_scene *s, int strategy, std::vector<float> *func, std::vector<float> *cdf,
                                   float *funcInt) {
    int n = s->n_lights;
    func->assign(n > 0 ? n : 1, 1.f);
    if (strategy == 1 && n > 1) {
        for (int i = 0; i < n; ++i) {
            const float *lp = s->light_params + (size_t)i * PBRTGPU_LIGHT_STRIDE;
            Spec L = spec(lp[0], lp[1], lp[2]), P;
            float R = s->world_radius;
            switch (s->light_tag[i]) {
            case PBRTGPU_LIGHT_DIFFUSE_AREA:
            case PBRTGPU_LIGHT_PORTAL_AREA: P = (lp[3] != 0.f ? 2 : 1) * L * lp[4] * kPi; break;  // diffuse.cpp:65-67
            case PBRTGPU_LIGHT_INFINITE: P = kPi * R * R * infiniteLookup(L, .5f, .5f); break;    // infinite.cpp:86-90
            case PBRTGPU_LIGHT_POINT: P = 4 * kPi * L; break;                                     // point.cpp:57
            case PBRTGPU_LIGHT_SPOT: P = L * 2 * kPi * (1 - .5f * (lp[7] + lp[6])); break;        // spot.cpp:75-77
            default: P = L * kPi * R * R; break;                                                  // distant.cpp:65-67
            }
            (*func)[i] = luminance(P);
        }
    }
    cdf->assign(n + 1 > 1 ? n + 1 : 2, 0.f);
    if (n == 0) {
        *funcInt = 0;
        return;
    }
    (*cdf)[0] = 0;
    for (int i = 1; i < n + 1; ++i) (*cdf)[i] = (*cdf)[i - 1] + (*func)[i - 1] / n;
    *funcInt = (*cdf)[n];
    if (*funcInt == 0)
        for (int i = 1; i < n + 1; ++i) (*cdf)[i] = float(i) / float(n);
    else
        for (int i = 1; i < n + 1; ++i) (*cdf)[i] /= *funcInt;
}

}  // namespace pg
#endif
