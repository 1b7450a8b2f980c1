// BVH traversal over the device layout.
//
// Layout (built at upload from the reference's LinearBVHNode array, see
// kernels.cu:buildWideNodes): one 64-byte record per INTERIOR node holding the
// bounds of both children, so that one 4x128-bit fetch decides both:
//     q0 = c0.min.xyz, c0.max.x      q1 = c0.max.yz, c1.min.xy
//     q2 = c1.min.z,  c1.max.xyz     q3 = ref0, ref1, axis, -        (ints)
// ref >= 0: index of an interior record; ref < 0: ~ref is the first slot of a
// leaf's primitives, which run until the slot whose tag carries kTagLast.
// Record 0 is a synthetic parent of the reference's root so that the root's
// own bounds are tested like any other node's.
//
// Equivalence with BVHAccel::Intersect/IntersectP (accelerators/bvh.cpp:
// 662-738).  The reference tests a node's bounds when it VISITS the node,
// against the ray's current tMax, and visits the near child (by
// dirIsNeg[axis]) first.  Bounds3::IntersectP (geometry.h:1580-1606) is
//     slabs(tMin, tMax') pass  &&  tMax' > 0  &&  tMin < ray.tMax
// and only the last term depends on ray.tMax.  Here both children are tested
// when the parent is visited: the near child at once (ray.tMax has not
// changed since), the far child's tMax-independent part now and its
// `tMin < ray.tMax` when it is popped, with the then-current tMax.  The set
// and order of visited leaves, and therefore every accepted hit including
// t == tMax ties, are the reference's.
#ifndef PG_DEVICE_BVH_H
#define PG_DEVICE_BVH_H

#include "shapes.h"

namespace pg {

#if defined(__CUDACC__)
typedef float4 F4;
#else
struct alignas(16) F4 {
    float x, y, z, w;
};
#endif

#if defined(__CUDA_ARCH__)
#define PG_LDG(p) __ldg(p)
#else
#define PG_LDG(p) (*(p))
#endif

constexpr int kEmptyRef = (int)0x80000000;
// primitive tag (int bits of tris[3*slot].w)
constexpr int kTagTypeMask = 3;   // PBRTGPU_PRIM_*
constexpr int kTagDegenerate = 4; // Triangle::Intersect rejects it (triangle.cpp:309-317); IntersectP does not
constexpr int kTagLast = 8;       // last primitive of its leaf
constexpr int kTagShapeShift = 8; // index into plane/sphere params

struct DevBVH {
    const F4 *nodes;  // 4 per interior record
    const F4 *tris;   // 3 per primitive slot: (p0, tag) (p1, -) (p2, -)
    const float *planeParams;
    const float *sphereParams;
};

struct Hit {
    int prim;  // BVH-order slot, -1 = miss
    float t, b0, b1, b2;
};

struct TraverseCounters {  // filled only by the instrumented oracle build
    long long nodes, prims;
};

PG_HD int f2i(float f) { return (int)floatToBits(f); }

// tMax-independent part of Bounds3::IntersectP; *tMinOut is the entry distance.
PG_HD bool slabPartial(V3 bmin, V3 bmax, V3 o, V3 invDir, const int dirIsNeg[3], float *tMinOut) {
    const float k = 1 + 2 * gammaf(3);
    float tMin = ((dirIsNeg[0] ? bmax.x : bmin.x) - o.x) * invDir.x;
    float tMax = ((dirIsNeg[0] ? bmin.x : bmax.x) - o.x) * invDir.x;
    float tyMin = ((dirIsNeg[1] ? bmax.y : bmin.y) - o.y) * invDir.y;
    float tyMax = ((dirIsNeg[1] ? bmin.y : bmax.y) - o.y) * invDir.y;
    tMax *= k;
    tyMax *= k;
    if (tMin > tyMax || tyMin > tMax) return false;
    if (tyMin > tMin) tMin = tyMin;
    if (tyMax < tMax) tMax = tyMax;
    float tzMin = ((dirIsNeg[2] ? bmax.z : bmin.z) - o.z) * invDir.z;
    float tzMax = ((dirIsNeg[2] ? bmin.z : bmax.z) - o.z) * invDir.z;
    tzMax *= k;
    if (tMin > tzMax || tzMin > tMax) return false;
    if (tzMin > tMin) tMin = tzMin;
    if (tzMax < tMax) tMax = tzMax;
    *tMinOut = tMin;
    return tMax > 0;
}

// One primitive against one ray, with the leaf loop's closest-hit acceptance.  Used to discard
// next-event rays that cannot contribute: a BSDF-sampled MIS ray counts only if its CLOSEST hit is
// the sampled light's own shape (integrator.cpp:226-233), and whatever the traversal has found
// before it reaches that shape only shrinks ray.tMax, which never turns a rejection into an
// acceptance (triangle.cpp:262-266, sphere.cpp:83-89, plane.cpp:27-29).  So "this test fails
// with the ray's initial tMax" implies "the closest hit is not this shape".
PG_HD bool primitiveAccepts(const DevBVH &bvh, int slot, const Ray &ray) {
    const F4 *tp = bvh.tris + 3 * (size_t)slot;
    F4 a = PG_LDG(tp);
    int tag = f2i(a.w);
    int type = tag & kTagTypeMask;
    float t, b0, b1;
    if (type == 0) {
        if (tag & kTagDegenerate) return false;
        F4 b = PG_LDG(tp + 1), c = PG_LDG(tp + 2);
        return triangleTest(ray, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &b0, &b1);
    }
    if (type == 1) return aaplaneTest(ray, bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift), &t, nullptr);
    return sphereTestGuarded(ray, bvh.sphereParams + 48 * (size_t)(tag >> kTagShapeShift), &t);
}

// Closest hit (ANY = false): ray.tMax shrinks as in GeometricPrimitive::
// Intersect (primitive.cpp:116-130), the last accepted primitive wins.
// Any hit (ANY = true): returns at the first accepted primitive.
template <bool ANY, bool COUNT = false>
PG_HD bool traverse(const DevBVH &bvh, Ray &ray, Hit *hit, TraverseCounters *cnt = nullptr) {
    V3 invDir = v3(1 / ray.d.x, 1 / ray.d.y, 1 / ray.d.z);
    int dirIsNeg[3] = {invDir.x < 0, invDir.y < 0, invDir.z < 0};
    RayShear shear = rayShear(ray.d);
    int stackRef[64];
    float stackT[64];
    int sp = 0;
    bool found = false;
    if (!ANY) hit->prim = -1;
    int ref = 0;
    for (;;) {
        if (ref >= 0) {
            const F4 *n = bvh.nodes + 4 * (size_t)ref;
            F4 q0 = PG_LDG(n), q1 = PG_LDG(n + 1), q2 = PG_LDG(n + 2), q3 = PG_LDG(n + 3);
            if (COUNT) cnt->nodes += 2;
            int ref0 = f2i(q3.x), ref1 = f2i(q3.y), axis = f2i(q3.z);
            float t0, t1;
            bool ok0 = ref0 != kEmptyRef && slabPartial(v3(q0.x, q0.y, q0.z), v3(q0.w, q1.x, q1.y), ray.o, invDir, dirIsNeg, &t0);
            bool ok1 = ref1 != kEmptyRef && slabPartial(v3(q1.z, q1.w, q2.x), v3(q2.y, q2.z, q2.w), ray.o, invDir, dirIsNeg, &t1);
            bool nearIs1 = dirIsNeg[axis] != 0;
            int refNear = nearIs1 ? ref1 : ref0, refFar = nearIs1 ? ref0 : ref1;
            bool okNear = nearIs1 ? ok1 : ok0, okFar = nearIs1 ? ok0 : ok1;
            float tNear = nearIs1 ? t1 : t0, tFar = nearIs1 ? t0 : t1;
            if (okFar) {
                stackRef[sp] = refFar;
                stackT[sp] = tFar;
                ++sp;
            }
            if (okNear && tNear < ray.tMax) {
                ref = refNear;
                continue;
            }
        } else {
            int slot = ~ref;
            for (;; ++slot) {
                const F4 *tp = bvh.tris + 3 * (size_t)slot;
                F4 a = PG_LDG(tp);
                int tag = f2i(a.w);
                int type = tag & kTagTypeMask;
                if (COUNT) cnt->prims += 1;
                float t, b0 = 0, b1 = 0, b2 = 0;
                bool h;
                if (type == 0) {
                    F4 b = PG_LDG(tp + 1), c = PG_LDG(tp + 2);
                    h = triangleTestSheared(ray, shear, v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), &t, &b0, &b1, &b2);
                    if (!ANY && (tag & kTagDegenerate)) h = false;
                } else if (type == 1) {
                    h = aaplaneTest(ray, bvh.planeParams + 48 * (size_t)(tag >> kTagShapeShift), &t, nullptr);
                } else {
                    h = sphereTestGuarded(ray, bvh.sphereParams + 48 * (size_t)(tag >> kTagShapeShift), &t);
                }
                if (h) {
                    if (ANY) return true;
                    found = true;
                    ray.tMax = t;
                    hit->prim = slot;
                    hit->t = t;
                    hit->b0 = b0;
                    hit->b1 = b1;
                    hit->b2 = b2;
                }
                if (tag & kTagLast) break;
            }
        }
        // pop the next node whose entry distance is still inside the ray
        for (;;) {
            if (sp == 0) return found;
            --sp;
            if (stackT[sp] < ray.tMax) {
                ref = stackRef[sp];
                break;
            }
        }
    }
}

}  // namespace pg
#endif
