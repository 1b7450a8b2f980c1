// BVHAccel::buildUpperSAH (accelerators/bvh.cpp:539-638): the SAH tree that HLBVHBuild puts over the roots of its
// Morton treelets.  A few thousand roots at most, recursive, std::partition -- it stays on the host for both builders
// (host/bvh_build.cpp builds the treelets on the CPU, device/bvh_build.cu on the GPU) and is shared so that the two
// cannot drift apart.  Same float arithmetic and the same libstdc++ partition as the reference: the tree is the
// reference's, node for node.
#ifndef PG_COMMON_UPPER_SAH_H
#define PG_COMMON_UPPER_SAH_H

#include <algorithm>
#include <vector>

namespace pg {

struct UBox {  // Bounds3f; the default is the reference's empty box (geometry.h:870-875)
    float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f}, mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
};
inline UBox uUnion(const UBox &a, const UBox &b) {
    UBox r;
    for (int k = 0; k < 3; ++k) r.mn[k] = std::min(a.mn[k], b.mn[k]), r.mx[k] = std::max(a.mx[k], b.mx[k]);
    return r;
}
inline UBox uUnionPoint(const UBox &a, const float p[3]) {
    UBox r;
    for (int k = 0; k < 3; ++k) r.mn[k] = std::min(a.mn[k], p[k]), r.mx[k] = std::max(a.mx[k], p[k]);
    return r;
}
inline float uSurfaceArea(const UBox &b) {  // geometry.h:930-933
    float dx = b.mx[0] - b.mn[0], dy = b.mx[1] - b.mn[1], dz = b.mx[2] - b.mn[2];
    return 2 * (dx * dy + dx * dz + dy * dz);
}
inline int uMaximumExtent(const UBox &b) {  // geometry.h:938-946
    float dx = b.mx[0] - b.mn[0], dy = b.mx[1] - b.mn[1], dz = b.mx[2] - b.mn[2];
    if (dx > dy && dx > dz) return 0;
    if (dy > dz) return 1;
    return 2;
}

// One node of the upper tree.  child[k] >= 0: another upper node; child[k] < 0: treelet ~child[k].
struct UpperNode {
    UBox bounds;
    int child[2];
    int axis;
};

struct UpperSAH {
    const std::vector<UBox> *rootBounds;  // bounds of treelet t's root
    std::vector<UpperNode> nodes;         // in creation order: a node comes before its subtrees (pre-order)

    const UBox &boundsOf(int ref) const { return ref >= 0 ? nodes[ref].bounds : (*rootBounds)[~ref]; }

    // roots[start, end) are treelet numbers; returns a reference in the encoding of UpperNode::child
    int build(std::vector<int> &roots, int start, int end) {
        int nNodes = end - start;
        if (nNodes == 1) return ~roots[start];
        const std::vector<UBox> &rb = *rootBounds;
        UBox bounds;
        for (int i = start; i < end; ++i) bounds = uUnion(bounds, rb[roots[i]]);
        UBox centroidBounds;
        for (int i = start; i < end; ++i) {
            const UBox &b = rb[roots[i]];
            float c[3] = {(b.mn[0] + b.mx[0]) * 0.5f, (b.mn[1] + b.mx[1]) * 0.5f, (b.mn[2] + b.mx[2]) * 0.5f};
            centroidBounds = uUnionPoint(centroidBounds, c);
        }
        int dim = uMaximumExtent(centroidBounds);
        // the reference CHECKs centroidBounds.pMax[dim] != pMin[dim] here and aborts otherwise (bvh.cpp:562)
        constexpr int nBuckets = 12;
        struct Bucket {
            int count = 0;
            UBox bounds;
        } buckets[nBuckets];
        const float cMin = centroidBounds.mn[dim], cMax = centroidBounds.mx[dim];
        auto bucketOf = [&](int root) {
            const UBox &b = rb[root];
            float centroid = (b.mn[dim] + b.mx[dim]) * 0.5f;
            int k = (int)(nBuckets * ((centroid - cMin) / (cMax - cMin)));
            if (k == nBuckets) k = nBuckets - 1;
            return k;
        };
        for (int i = start; i < end; ++i) {
            int k = bucketOf(roots[i]);
            buckets[k].count++;
            buckets[k].bounds = uUnion(buckets[k].bounds, rb[roots[i]]);
        }
        float cost[nBuckets - 1];
        for (int i = 0; i < nBuckets - 1; ++i) {
            UBox b0, b1;
            int count0 = 0, count1 = 0;
            for (int j = 0; j <= i; ++j) b0 = uUnion(b0, buckets[j].bounds), count0 += buckets[j].count;
            for (int j = i + 1; j < nBuckets; ++j) b1 = uUnion(b1, buckets[j].bounds), count1 += buckets[j].count;
            cost[i] = .125f + (count0 * uSurfaceArea(b0) + count1 * uSurfaceArea(b1)) / uSurfaceArea(bounds);
        }
        float minCost = cost[0];
        int minCostSplitBucket = 0;
        for (int i = 1; i < nBuckets - 1; ++i)
            if (cost[i] < minCost) minCost = cost[i], minCostSplitBucket = i;
        int *pmid = std::partition(&roots[start], &roots[end - 1] + 1, [&](int root) { return bucketOf(root) <= minCostSplitBucket; });
        int mid = (int)(pmid - &roots[0]);
        int nodeIdx = (int)nodes.size();
        nodes.emplace_back();
        int c0 = build(roots, start, mid);
        int c1 = build(roots, mid, end);
        UpperNode &n = nodes[nodeIdx];
        n.child[0] = c0, n.child[1] = c1;
        n.bounds = uUnion(boundsOf(c0), boundsOf(c1));
        n.axis = dim;
        return nodeIdx;
    }
};

}  // namespace pg
#endif
