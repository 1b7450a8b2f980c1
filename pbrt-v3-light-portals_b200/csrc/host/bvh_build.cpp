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

void flatten(const Builder &b, int nodeIdx, BuiltBVH *out) {
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
    b.method = method == SplitMethod::HLBVH ? SplitMethod::SAH : method;
    b.info.resize(primBounds.size());
    for (size_t i = 0; i < primBounds.size(); ++i) {
        b.info[i].primitiveNumber = i;
        b.info[i].bounds = primBounds[i];
        b.info[i].centroid = .5f * primBounds[i].pMin + .5f * primBounds[i].pMax;
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
