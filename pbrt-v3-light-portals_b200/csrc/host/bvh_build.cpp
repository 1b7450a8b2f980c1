// Host BVH construction.  The tree must be the reference's tree, node for
// node: closest-hit ties (t == tMax is accepted, shapes/triangle.cpp:259-262)
// are resolved by visiting order, so tier-1 parity needs the same topology,
// the same split axes and the same primitive order inside every leaf.  This
// file therefore makes the same decisions as BVHAccel::recursiveBuild
// (accelerators/bvh.cpp:236-402) with the same float arithmetic and the same
// libstdc++ algorithms (std::partition / std::nth_element), and emits nodes in
// the DFS order of BVHAccel::flattenBVHTree (bvh.cpp:640-658).
#include <algorithm>
#include <cstdio>

#include "../common/upper_sah.h"
#include "scene.h"

namespace pg {

namespace {

struct PrimInfo {  // BVHPrimitiveInfo, bvh.cpp:46-58
    size_t primitiveNumber;
    Bounds3 bounds;
    V3 centroid;
};

struct BuildNode {
    Bounds3 bounds;
    int child[2] = {-1, -1};
    int splitAxis = 0, firstPrimOffset = 0, nPrimitives = 0;
};

struct Builder {
    std::vector<PrimInfo> info;
    std::vector<BuildNode> nodes;
    std::vector<int> ordered;
    int maxPrimsInNode;
    SplitMethod method;

    int makeLeaf(int nodeIdx, int start, int end, const Bounds3 &bounds) {
        nodes[nodeIdx].firstPrimOffset = (int)ordered.size();
        for (int i = start; i < end; ++i) ordered.push_back((int)info[i].primitiveNumber);
        nodes[nodeIdx].nPrimitives = end - start;
        nodes[nodeIdx].bounds = bounds;
        return nodeIdx;
    }

    int build(int start, int end) {
        int nodeIdx = (int)nodes.size();
        nodes.emplace_back();
        Bounds3 bounds;
        for (int i = start; i < end; ++i) bounds = unionB(bounds, info[i].bounds);
        int nPrimitives = end - start;
        if (nPrimitives == 1) return makeLeaf(nodeIdx, start, end, bounds);

        Bounds3 centroidBounds;
        for (int i = start; i < end; ++i) centroidBounds = unionB(centroidBounds, info[i].centroid);
        int dim = centroidBounds.maximumExtent();
        int mid = (start + end) / 2;
        if (centroidBounds.pMax[dim] == centroidBounds.pMin[dim]) return makeLeaf(nodeIdx, start, end, bounds);

        auto equalCounts = [&]() {
            mid = (start + end) / 2;
            std::nth_element(&info[start], &info[mid], &info[end - 1] + 1,
                             [dim](const PrimInfo &a, const PrimInfo &b) { return a.centroid[dim] < b.centroid[dim]; });
        };
        bool done = false;
        if (method == SplitMethod::Middle) {
            float pmid = (centroidBounds.pMin[dim] + centroidBounds.pMax[dim]) / 2;
            PrimInfo *midPtr = std::partition(&info[start], &info[end - 1] + 1,
                                              [dim, pmid](const PrimInfo &pi) { return pi.centroid[dim] < pmid; });
            mid = int(midPtr - &info[0]);
            if (mid != start && mid != end) done = true;
        }
        if (!done && (method == SplitMethod::Middle || method == SplitMethod::EqualCounts)) {
            equalCounts();
            done = true;
        }
        if (!done) {  // SAH (also the default branch of the reference's switch)
            if (nPrimitives <= 2) {
                equalCounts();
            } else {
                constexpr int nBuckets = 12;
                struct Bucket { int count = 0; Bounds3 bounds; } buckets[nBuckets];
                for (int i = start; i < end; ++i) {
                    int b = nBuckets * centroidBounds.offset(info[i].centroid)[dim];
                    if (b == nBuckets) b = nBuckets - 1;
                    buckets[b].count++;
                    buckets[b].bounds = unionB(buckets[b].bounds, info[i].bounds);
                }
                float cost[nBuckets - 1];
                for (int i = 0; i < nBuckets - 1; ++i) {
                    Bounds3 b0, b1;
                    int count0 = 0, count1 = 0;
                    for (int j = 0; j <= i; ++j) {
                        b0 = unionB(b0, buckets[j].bounds);
                        count0 += buckets[j].count;
                    }
                    for (int j = i + 1; j < nBuckets; ++j) {
                        b1 = unionB(b1, buckets[j].bounds);
                        count1 += buckets[j].count;
                    }
                    cost[i] = 1 + (count0 * b0.surfaceArea() + count1 * b1.surfaceArea()) / bounds.surfaceArea();
                }
                float minCost = cost[0];
                int minCostSplitBucket = 0;
                for (int i = 1; i < nBuckets - 1; ++i)
                    if (cost[i] < minCost) {
                        minCost = cost[i];
                        minCostSplitBucket = i;
                    }
                float leafCost = nPrimitives;
                if (nPrimitives > maxPrimsInNode || minCost < leafCost) {
                    PrimInfo *pmid = std::partition(&info[start], &info[end - 1] + 1, [=](const PrimInfo &pi) {
                        int b = nBuckets * centroidBounds.offset(pi.centroid)[dim];
                        if (b == nBuckets) b = nBuckets - 1;
                        return b <= minCostSplitBucket;
                    });
                    mid = int(pmid - &info[0]);
                } else {
                    return makeLeaf(nodeIdx, start, end, bounds);
                }
            }
        }
        // The reference writes both recursive calls as arguments of one call
        // (bvh.cpp:392-396); with gcc the second argument is evaluated first,
        // so the right subtree's leaves receive the lower primitive offsets.
        // Only the numbering of leaf slots depends on this, never a hit result,
        // but we keep gcc's order so that node dumps compare equal.
        int c1 = build(mid, end);
        int c0 = build(start, mid);
        BuildNode &n = nodes[nodeIdx];
        n.child[0] = c0;
        n.child[1] = c1;
        n.bounds = unionB(nodes[c0].bounds, nodes[c1].bounds);
        n.splitAxis = dim;
        n.nPrimitives = 0;
        return nodeIdx;
    }
};

// ---- HLBVH (BVHAccel::HLBVHBuild, bvh.cpp:404-464; emitLBVH :466-537; buildUpperSAH :539-638).
// Morton codes of the centroids (10 bits per axis), a stable sort on them (the reference's LSD radix
// sort is stable), one LBVH treelet per distinct value of the top 12 bits, a SAH tree over the treelet
// roots.  The reference builds treelets in parallel and hands out leaf slots with an atomic counter,
// so the NUMBERING of leaf slots depends on its thread schedule; the tree itself does not.  Treelets
// are built here in order, which is the reference's numbering with --nthreads 1.
inline uint32_t leftShift3(uint32_t x) {  // bvh.cpp:107-130
    if (x == (1 << 10)) --x;
    x = (x | (x << 16)) & 0x30000ff;
    x = (x | (x << 8)) & 0x300f00f;
    x = (x | (x << 4)) & 0x30c30c3;
    x = (x | (x << 2)) & 0x9249249;
    return x;
}
inline uint32_t encodeMorton3(V3 v) { return (leftShift3((uint32_t)v.z) << 2) | (leftShift3((uint32_t)v.y) << 1) | leftShift3((uint32_t)v.x); }

struct MortonPrim {
    int primitiveIndex;
    uint32_t mortonCode;
};

struct HLBuilder {
    const std::vector<PrimInfo> *info;
    std::vector<BuildNode> nodes;
    std::vector<int> ordered;
    int maxPrimsInNode;

    int emitLBVH(const MortonPrim *mp, int nPrimitives, int bitIndex) {
        if (bitIndex == -1 || nPrimitives < maxPrimsInNode) {
            int nodeIdx = (int)nodes.size();
            nodes.emplace_back();
            Bounds3 bounds;
            int first = (int)ordered.size();
            for (int i = 0; i < nPrimitives; ++i) {
                ordered.push_back(mp[i].primitiveIndex);
                bounds = unionB(bounds, (*info)[mp[i].primitiveIndex].bounds);
            }
            nodes[nodeIdx].firstPrimOffset = first;
            nodes[nodeIdx].nPrimitives = nPrimitives;
            nodes[nodeIdx].bounds = bounds;
            return nodeIdx;
        }
        uint32_t mask = 1u << bitIndex;
        if ((mp[0].mortonCode & mask) == (mp[nPrimitives - 1].mortonCode & mask)) return emitLBVH(mp, nPrimitives, bitIndex - 1);
        int searchStart = 0, searchEnd = nPrimitives - 1;
        while (searchStart + 1 != searchEnd) {
            int mid = (searchStart + searchEnd) / 2;
            if ((mp[searchStart].mortonCode & mask) == (mp[mid].mortonCode & mask)) searchStart = mid;
            else searchEnd = mid;
        }
        int splitOffset = searchEnd;
        int nodeIdx = (int)nodes.size();
        nodes.emplace_back();
        int c0 = emitLBVH(mp, splitOffset, bitIndex - 1);
        int c1 = emitLBVH(mp + splitOffset, nPrimitives - splitOffset, bitIndex - 1);
        BuildNode &n = nodes[nodeIdx];
        n.child[0] = c0, n.child[1] = c1;
        n.bounds = unionB(nodes[c0].bounds, nodes[c1].bounds);
        n.splitAxis = bitIndex % 3;
        n.nPrimitives = 0;
        return nodeIdx;
    }

    // buildUpperSAH (bvh.cpp:539-638) lives in common/upper_sah.h, shared with the device builder
    int buildUpper(const std::vector<int> &treeletRoots) {
        std::vector<UBox> rb(treeletRoots.size());
        for (size_t t = 0; t < treeletRoots.size(); ++t) {
            const Bounds3 &b = nodes[treeletRoots[t]].bounds;
            rb[t].mn[0] = b.pMin.x, rb[t].mn[1] = b.pMin.y, rb[t].mn[2] = b.pMin.z;
            rb[t].mx[0] = b.pMax.x, rb[t].mx[1] = b.pMax.y, rb[t].mx[2] = b.pMax.z;
        }
        UpperSAH up;
        up.rootBounds = &rb;
        std::vector<int> order(treeletRoots.size());
        for (size_t t = 0; t < order.size(); ++t) order[t] = (int)t;
        int rootRef = up.build(order, 0, (int)order.size());
        int base = (int)nodes.size();
        auto nodeOf = [&](int ref) { return ref >= 0 ? base + ref : treeletRoots[~ref]; };
        for (const UpperNode &u : up.nodes) {
            BuildNode n;
            n.bounds.pMin = v3(u.bounds.mn[0], u.bounds.mn[1], u.bounds.mn[2]);
            n.bounds.pMax = v3(u.bounds.mx[0], u.bounds.mx[1], u.bounds.mx[2]);
            n.child[0] = nodeOf(u.child[0]), n.child[1] = nodeOf(u.child[1]);
            n.splitAxis = u.axis;
            n.nPrimitives = 0;
            nodes.push_back(n);
        }
        return nodeOf(rootRef);
    }

    int build() {
        const std::vector<PrimInfo> &pi = *info;
        Bounds3 bounds;
        for (const PrimInfo &p : pi) bounds = unionB(bounds, p.centroid);
        std::vector<MortonPrim> mp(pi.size());
        const int mortonScale = 1 << 10;
        for (size_t i = 0; i < pi.size(); ++i) {
            mp[i].primitiveIndex = (int)pi[i].primitiveNumber;
            V3 o = bounds.offset(pi[i].centroid);
            mp[i].mortonCode = encodeMorton3(o * (float)mortonScale);
        }
        std::stable_sort(mp.begin(), mp.end(), [](const MortonPrim &a, const MortonPrim &b) { return a.mortonCode < b.mortonCode; });
        std::vector<int> roots;
        const uint32_t mask = 0x3ffc0000;
        for (int start = 0, end = 1; end <= (int)mp.size(); ++end)
            if (end == (int)mp.size() || ((mp[start].mortonCode & mask) != (mp[end].mortonCode & mask))) {
                roots.push_back(emitLBVH(&mp[start], end - start, 29 - 12));
                start = end;
            }
        return buildUpper(roots);
    }
};

template <class B>
void flatten(const B &b, int nodeIdx, BuiltBVH *out) {
    const BuildNode &n = b.nodes[nodeIdx];
    int my = (int)out->offset.size();
    const float bb[6] = {n.bounds.pMin.x, n.bounds.pMin.y, n.bounds.pMin.z, n.bounds.pMax.x, n.bounds.pMax.y, n.bounds.pMax.z};
    out->bounds.insert(out->bounds.end(), bb, bb + 6);
    out->offset.push_back(0);
    out->nPrims.push_back(0);
    out->axis.push_back(0);
    if (n.nPrimitives > 0) {
        out->offset[my] = n.firstPrimOffset;
        out->nPrims[my] = (uint16_t)n.nPrimitives;
    } else {
        out->axis[my] = (uint8_t)n.splitAxis;
        flatten(b, n.child[0], out);
        out->offset[my] = (int)out->offset.size();
        flatten(b, n.child[1], out);
    }
}

}  // namespace

BuiltBVH buildBVH(const std::vector<Bounds3> &primBounds, int maxPrimsInNode, SplitMethod method) {
    BuiltBVH out;
    if (primBounds.empty()) return out;
    Builder b;
    b.maxPrimsInNode = std::min(255, maxPrimsInNode);
    b.method = method;
    b.info.resize(primBounds.size());
    for (size_t i = 0; i < primBounds.size(); ++i) {
        b.info[i].primitiveNumber = i;
        b.info[i].bounds = primBounds[i];
        b.info[i].centroid = .5f * primBounds[i].pMin + .5f * primBounds[i].pMax;
    }
    if (method == SplitMethod::HLBVH) {
        HLBuilder h;
        h.info = &b.info;
        h.maxPrimsInNode = b.maxPrimsInNode;
        h.nodes.reserve(2 * primBounds.size());
        h.ordered.reserve(primBounds.size());
        int root = h.build();
        out.bounds.reserve(6 * h.nodes.size());
        flatten(h, root, &out);
        out.orderedPrims = std::move(h.ordered);
        return out;
    }
    b.nodes.reserve(2 * primBounds.size());
    b.ordered.reserve(primBounds.size());
    b.build(0, (int)primBounds.size());
    out.bounds.reserve(6 * b.nodes.size());
    flatten(b, 0, &out);
    out.orderedPrims = std::move(b.ordered);
    return out;
}

}  // namespace pg
