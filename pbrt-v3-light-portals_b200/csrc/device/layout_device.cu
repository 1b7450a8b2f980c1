// The device BVH / primitive layout of bvh.h, built ON the device from the reference-layout arrays of pbrtgpu_scene
// (BVHAccel::nodes as flattenBVHTree leaves them, bvh.cpp:640-658, and the primitives in BVHAccel::primitives order).
// layout_build.h does the same on the host (kept: the CPU test harness uses it, and PBRTGPU_LAYOUT=host selects it for
// A/B and for the byte-for-byte test); this is what pbrtgpu_upload_scene runs.  The raw arrays are copied as they are
// (24 B per node + 7 B of node fields, 36 B + tag inputs per primitive) and three kernels write the records:
//   k_wide_nodes   one 64-byte two-child record per interior node; record index = 1 + number of interior nodes before it
//                  (exclusive scan of the interior flags)
//   k_prim_records (p0, tag) (p1, -) (p2, -) per BVH slot; tag = type | degenerate | shape index
//   k_tag_last     the last slot of every leaf gets kTagLast
// Same arithmetic as the host version (triangleIsDegenerate is one function compiled for both), same bytes.
#include "ctx.h"
#include "layout_build.h"

namespace pg {
int exclusiveScanInt(pbrtgpu_ctx *ctx, const int *in, int *out, long long n, int *sums, int *totalOut);
long long scanScratchInts(long long n);
}  // namespace pg

using namespace pg;

namespace {

__global__ void k_interior_flags(const unsigned short *nPrims, int n, int *flag) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) flag[i] = nPrims[i] == 0 ? 1 : 0;
}

__device__ __forceinline__ void setChild(F4 *q, int which, const float *b, int ref) {
    if (which == 0) {
        q[0] = make_float4(b[0], b[1], b[2], b[3]);
        q[1].x = b[4], q[1].y = b[5];
        q[3].x = __int_as_float(ref);
    } else {
        q[1].z = b[0], q[1].w = b[1];
        q[2] = make_float4(b[2], b[3], b[4], b[5]);
        q[3].y = __int_as_float(ref);
    }
}
// nodes[] is zeroed beforehand; interiorBefore = exclusive scan of the interior flags
__global__ void k_wide_nodes(const float *bounds, const int *offset, const unsigned short *nPrims, const unsigned char *axis, const int *interiorBefore,
                             int n, F4 *nodes) {
    auto refOf = [&](int node) { return nPrims[node] == 0 ? 1 + interiorBefore[node] : ~offset[node]; };
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (i == 0) {  // record 0: synthetic parent of the root; second child empty
            setChild(nodes, 0, bounds, refOf(0));
            nodes[3].y = __int_as_float(kEmptyRef);
            nodes[3].z = __int_as_float(0);
        }
        if (nPrims[i] != 0) continue;
        F4 *q = nodes + 4 * (size_t)(1 + interiorBefore[i]);
        const int c0 = i + 1, c1 = offset[i];
        setChild(q, 0, bounds + 6 * (size_t)c0, refOf(c0));
        setChild(q, 1, bounds + 6 * (size_t)c1, refOf(c1));
        q[3].z = __int_as_float((int)axis[i]);
    }
}

__global__ void k_prim_records(const float *triP, const unsigned char *primType, const unsigned char *primFlags, const int *primShape,
                               const float *triUV, int n, int nPlanes, int nSpheres, F4 *tris, int *firstBad) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float p[9];
        for (int k = 0; k < 9; ++k) p[k] = triP[9 * (size_t)i + k];
        const int type = primType[i];
        int tag = type & kTagTypeMask;
        if (type == PBRTGPU_PRIM_TRIANGLE) {
            float uv[6];
            const bool hasUV = triUV && primFlags && (primFlags[i] & PBRTGPU_PF_HAS_UV);
            if (hasUV)
                for (int k = 0; k < 6; ++k) uv[k] = triUV[6 * (size_t)i + k];
            if (triangleIsDegenerate(p, hasUV ? uv : nullptr)) tag |= kTagDegenerate;
        } else {
            const int idx = primShape ? primShape[i] : 0;
            const int limit = type == PBRTGPU_PRIM_AAPLANE ? nPlanes : nSpheres;
            if (idx < 0 || idx >= limit) atomicMin(firstBad, i);
            tag |= idx << kTagShapeShift;
        }
        tris[3 * (size_t)i + 0] = make_float4(p[0], p[1], p[2], __int_as_float(tag));
        tris[3 * (size_t)i + 1] = make_float4(p[3], p[4], p[5], 0.f);
        tris[3 * (size_t)i + 2] = make_float4(p[6], p[7], p[8], 0.f);
    }
}
__global__ void k_tag_last(const int *offset, const unsigned short *nPrims, int n, F4 *tris) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        if (nPrims[i] != 0) {
            size_t last = (size_t)offset[i] + nPrims[i] - 1;
            tris[3 * last].w = __int_as_float(__float_as_int(tris[3 * last].w) | kTagLast);
        }
}

}  // namespace

namespace pg {

// d_nodes: 4 * (1 + nInterior) F4, d_tris: 3 * n_prims F4 (both allocated by the caller); d_triUV / d_primFlags: the scene's own
// device copies (or null).  Everything is enqueued on ctx->stream; *firstBadHost is valid after the stream has been synchronised.
int buildLayoutOnDevice(pbrtgpu_ctx *ctx, const pbrtgpu_scene *s, int nInterior, const float *d_triUV, const unsigned char *d_primFlags,
                        F4 *d_nodes, F4 *d_tris, int *firstBadHost) {
    cudaStream_t st = ctx->stream;
    const size_t nn = (size_t)s->n_nodes, np = (size_t)s->n_prims;
    // raw inputs and scan scratch live in ctx->layoutScratch (grown on demand, kept for the next upload)
    size_t off[10], total = 0;
    auto take = [&](int k, size_t bytes) { off[k] = total, total += (bytes + 255) & ~(size_t)255; };
    take(0, 24 * nn);                                        // bvh_bounds
    take(1, 4 * nn);                                         // bvh_offset
    take(2, 2 * nn);                                         // bvh_nprims
    take(3, nn);                                             // bvh_axis
    take(4, 4 * nn);                                         // interior flags -> their exclusive scan
    take(5, 4 * (size_t)scanScratchInts((long long)nn));    // scan scratch
    take(6, 36 * np);                                        // tri_p
    take(7, np);                                             // prim_type
    take(8, s->prim_shape ? 4 * np : 0);                     // prim_shape
    take(9, 256);                                            // first bad primitive
    int rc = ensureBuffer(ctx, ctx->layoutScratch, total);
    if (rc != PBRTGPU_OK) return rc;
    char *base = (char *)ctx->layoutScratch.p;
    float *dBounds = (float *)(base + off[0]);
    int *dOffset = (int *)(base + off[1]);
    unsigned short *dNPrims = (unsigned short *)(base + off[2]);
    unsigned char *dAxis = (unsigned char *)(base + off[3]);
    int *dFlag = (int *)(base + off[4]), *dSums = (int *)(base + off[5]);
    float *dTriP = (float *)(base + off[6]);
    unsigned char *dType = (unsigned char *)(base + off[7]);
    int *dShape = s->prim_shape ? (int *)(base + off[8]) : nullptr;
    int *dBad = (int *)(base + off[9]);
    PG_CUDA(ctx, cudaMemcpyAsync(dBounds, s->bvh_bounds, 24 * nn, cudaMemcpyHostToDevice, st));
    PG_CUDA(ctx, cudaMemcpyAsync(dOffset, s->bvh_offset, 4 * nn, cudaMemcpyHostToDevice, st));
    PG_CUDA(ctx, cudaMemcpyAsync(dNPrims, s->bvh_nprims, 2 * nn, cudaMemcpyHostToDevice, st));
    PG_CUDA(ctx, cudaMemcpyAsync(dAxis, s->bvh_axis, nn, cudaMemcpyHostToDevice, st));
    PG_CUDA(ctx, cudaMemcpyAsync(dTriP, s->tri_p, 36 * np, cudaMemcpyHostToDevice, st));
    PG_CUDA(ctx, cudaMemcpyAsync(dType, s->prim_type, np, cudaMemcpyHostToDevice, st));
    if (dShape) PG_CUDA(ctx, cudaMemcpyAsync(dShape, s->prim_shape, 4 * np, cudaMemcpyHostToDevice, st));
    const int big = 0x7fffffff;
    PG_CUDA(ctx, cudaMemcpyAsync(dBad, &big, sizeof(int), cudaMemcpyHostToDevice, st));
    PG_CUDA(ctx, cudaMemsetAsync(d_nodes, 0, 64 * (size_t)(1 + nInterior), st));

    const int grid = ctx->smCount * 8;
    k_interior_flags<<<grid, 256, 0, st>>>(dNPrims, s->n_nodes, dFlag);
    if ((rc = exclusiveScanInt(ctx, dFlag, dFlag, (long long)nn, dSums, nullptr)) != PBRTGPU_OK) return rc;
    k_wide_nodes<<<grid, 256, 0, st>>>(dBounds, dOffset, dNPrims, dAxis, dFlag, s->n_nodes, d_nodes);
    k_prim_records<<<grid, 256, 0, st>>>(dTriP, dType, d_primFlags, dShape, d_triUV, s->n_prims, s->n_planes, s->n_spheres, d_tris, dBad);
    k_tag_last<<<grid, 256, 0, st>>>(dOffset, dNPrims, s->n_nodes, d_tris);
    ctx->launches += 4;
    PG_CUDA(ctx, cudaGetLastError());
    PG_CUDA(ctx, cudaMemcpyAsync(firstBadHost, dBad, sizeof(int), cudaMemcpyDeviceToHost, st));
    return PBRTGPU_OK;
}

}  // namespace pg
