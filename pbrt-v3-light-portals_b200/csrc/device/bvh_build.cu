// BVHAccel::HLBVHBuild on the device (accelerators/bvh.cpp:404-472, 107-180, 474-536).
//
//   reference (CPU)                                   here
//   centroid bounds, Morton codes   bvh.cpp:404-428   k_centroid_bounds, k_morton        one thread per primitive
//   RadixSort (LSD, 6 bits x 5)     bvh.cpp:139-180   k_radix_hist / scan / k_radix_scatter, LSD, 8 bits x 4, stable
//   treelet ranges (top 12 bits)    bvh.cpp:433-455   k_treelet_starts                   one flag per sorted position
//   emitLBVH, recursive per treelet bvh.cpp:474-536   k_emit_level                       one thread per open range, level by level
//   buildUpperSAH over the roots    bvh.cpp:538-638   host, common/upper_sah.h           (a few thousand roots)
//   flattenBVHTree                  bvh.cpp:640-658   k_up_level (subtree sizes, bounds), k_down_level (DFS numbers)
//
// The result is the reference's tree node for node (tests/test_rays.py::test_gpu_hlbvh_*): min / max are exact (minf / maxf of
// pg_math.h, the reference's std::min / std::max with its argument order, so even the sign of a zero agrees), the
// Morton arithmetic is the same float sequence (no FMA, IEEE division), any stable sort of the codes is THE stable
// sort, and emitLBVH's decisions depend only on the sorted codes.  Leaf slots are numbered in Morton order, which is
// the reference's numbering with --nthreads 1 (it hands out leaf slots with an atomic counter while it builds
// treelets in parallel, so its numbering -- not its tree -- depends on the thread schedule).
//
// Level by level instead of recursion: a range that is to become a node is a record in a temporary node array; the
// kernel for level t turns every record of that level into a leaf or an interior node and appends its two children as
// the records of level t + 1 (one atomicAdd per split).  A node at depth d has consumed at least d of the 18 Morton bits
// below the treelet prefix, so 19 levels always suffice.  Two more sweeps give the reference's DFS layout: bottom-up
// (subtree sizes, interior bounds), top-down (final index = parent's + 1 for the first child, + 1 + size(first
// subtree) for the second; flattenBVHTree's order).
#include <algorithm>
#include <vector>

#include "../common/pg_math.h"
#include "../common/upper_sah.h"
#include "ctx.h"

using namespace pg;

namespace {

constexpr int kLevels = 19;  // bits 17 .. 0 below the 12-bit treelet prefix, plus the leaf level

// ---------------------------------------------------------------- exclusive scan of int32 (n up to 2^31)
constexpr int kScanThreads = 256, kScanItems = 8, kScanChunk = kScanThreads * kScanItems;

__device__ __forceinline__ int blockExclusiveScan(int v, int *total, int *sh /* 33 ints */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
    for (int d = 1; d < 32; d <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) sh[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < (blockDim.x >> 5) ? sh[lane] : 0, winc = w;
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += t;
        }
        sh[lane] = winc - w;
        if (lane == 31) sh[32] = winc;
    }
    __syncthreads();
    int r = sh[warp] + inc - v;
    *total = sh[32];
    __syncthreads();
    return r;
}

__global__ void k_scan_reduce(const int *in, long long n, int *blockSums) {
    __shared__ int sh[33];
    long long base = (long long)blockIdx.x * kScanChunk + (long long)threadIdx.x * kScanItems;
    int s = 0;
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) s += in[base + k];
    int total;
    blockExclusiveScan(s, &total, sh);
    if (threadIdx.x == 0) blockSums[blockIdx.x] = total;
}
// one block: exclusive scan of the block sums in place; the grand total goes to *totalOut when it is not null
__global__ void k_scan_sums(int *blockSums, int nBlocks, int *totalOut) {
    __shared__ int sh[33];
    int carry = 0;
    for (int base = 0; base < nBlocks; base += blockDim.x) {
        int i = base + threadIdx.x;
        int v = i < nBlocks ? blockSums[i] : 0, total;
        int ex = blockExclusiveScan(v, &total, sh);
        if (i < nBlocks) blockSums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0 && totalOut) *totalOut = carry;
}
__global__ void k_scan_apply(const int *in, long long n, const int *blockSums, int *out) {
    __shared__ int sh[33];
    long long base = (long long)blockIdx.x * kScanChunk + (long long)threadIdx.x * kScanItems;
    int v[kScanItems], s = 0;
    for (int k = 0; k < kScanItems; ++k) {
        v[k] = base + k < n ? in[base + k] : 0;
        s += v[k];
    }
    int total;
    int ex = blockExclusiveScan(s, &total, sh) + blockSums[blockIdx.x];
    for (int k = 0; k < kScanItems; ++k) {
        if (base + k < n) out[base + k] = ex;
        ex += v[k];
    }
}

}  // namespace

namespace pg {
// out[i] = in[0] + ... + in[i-1]; `sums` holds ceil(n / kScanChunk) ints of scratch; in == out is allowed
int exclusiveScanInt(pbrtgpu_ctx *ctx, const int *in, int *out, long long n, int *sums, int *totalOut) {
    if (n <= 0) return PBRTGPU_OK;
    int nBlocks = (int)((n + kScanChunk - 1) / kScanChunk);
    k_scan_reduce<<<nBlocks, kScanThreads, 0, ctx->stream>>>(in, n, sums);
    k_scan_sums<<<1, 1024, 0, ctx->stream>>>(sums, nBlocks, totalOut);
    k_scan_apply<<<nBlocks, kScanThreads, 0, ctx->stream>>>(in, n, sums, out);
    ctx->launches += 3;
    PG_CUDA(ctx, cudaGetLastError());
    return PBRTGPU_OK;
}
long long scanScratchInts(long long n) { return (n + kScanChunk - 1) / kScanChunk + 1; }
}  // namespace pg

namespace {

// ---------------------------------------------------------------- Morton codes (bvh.cpp:107-137, 411-428)
__device__ __forceinline__ unsigned orderedBits(float f) {  // monotone float -> uint, for atomicMin / atomicMax
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float fromOrderedBits(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

// centroid = .5f * pMin + .5f * pMax (bvh.cpp:51-52); bounds of all centroids -> cb[0..2] min, cb[3..5] max (ordered bits)
__global__ void k_centroid_bounds(const float *primBounds, int n, unsigned *cb) {
    float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f}, mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float *b = primBounds + 6 * (size_t)i;
        for (int k = 0; k < 3; ++k) {
            float c = .5f * b[k] + .5f * b[3 + k];
            mn[k] = minf(mn[k], c);
            mx[k] = maxf(mx[k], c);
        }
    }
    for (int k = 0; k < 3; ++k) {
        for (int d = 16; d; d >>= 1) {
            mn[k] = minf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], d));
            mx[k] = maxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], d));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&cb[k], orderedBits(mn[k]));
            atomicMax(&cb[3 + k], orderedBits(mx[k]));
        }
    }
}
__device__ __forceinline__ unsigned leftShift3(unsigned x) {
    if (x == (1u << 10)) --x;
    x = (x | (x << 16)) & 0x30000ffu;
    x = (x | (x << 8)) & 0x300f00fu;
    x = (x | (x << 4)) & 0x30c30c3u;
    x = (x | (x << 2)) & 0x9249249u;
    return x;
}
__global__ void k_morton(const float *primBounds, int n, const unsigned *cb, unsigned *keys, int *vals) {
    float mn[3], mx[3];
    for (int k = 0; k < 3; ++k) mn[k] = fromOrderedBits(cb[k]), mx[k] = fromOrderedBits(cb[3 + k]);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float *b = primBounds + 6 * (size_t)i;
        unsigned q[3];
        for (int k = 0; k < 3; ++k) {
            float c = .5f * b[k] + .5f * b[3 + k];
            float o = c - mn[k];  // Bounds3::Offset, geometry.h:953-959
            if (mx[k] > mn[k]) o /= mx[k] - mn[k];
            q[k] = (unsigned)(o * 1024.f);  // mortonScale = 1 << 10
        }
        keys[i] = (leftShift3(q[2]) << 2) | (leftShift3(q[1]) << 1) | leftShift3(q[0]);
        vals[i] = i;
    }
}

// ---------------------------------------------------------------- stable LSD radix sort, 8 bits per pass
constexpr int kSortThreads = 256, kSortWarps = kSortThreads / 32, kSortPerWarp = 256, kSortTile = kSortWarps * kSortPerWarp;

__global__ void k_radix_hist(const unsigned *keys, int n, int shift, int nTiles, int *hist /* [256][nTiles] */) {
    __shared__ int h[256];
    for (int tile = blockIdx.x; tile < nTiles; tile += gridDim.x) {
        h[threadIdx.x] = 0;
        __syncthreads();
        int base = tile * kSortTile;
        for (int k = threadIdx.x; k < kSortTile; k += kSortThreads)
            if (base + k < n) atomicAdd(&h[(keys[base + k] >> shift) & 255], 1);
        __syncthreads();
        hist[(size_t)threadIdx.x * nTiles + tile] = h[threadIdx.x];
        __syncthreads();
    }
}
// `offsets` = exclusive scan of hist in (digit, tile) order: where tile `tile`'s keys with digit d start.
// Inside a tile the keys keep their order: warp w owns keys [256 w, 256 w + 256), walks them 32 at a time and numbers
// the keys of one digit with __match_any_sync (lane order = key order); warps are then offset by a per-digit prefix.
__global__ void k_radix_scatter(const unsigned *keys, const int *vals, int n, int shift, int nTiles, const int *offsets, unsigned *keysOut,
                                int *valsOut) {
    __shared__ int warpCount[kSortWarps][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int tile = blockIdx.x; tile < nTiles; tile += gridDim.x) {
        for (int k = threadIdx.x; k < kSortWarps * 256; k += kSortThreads) (&warpCount[0][0])[k] = 0;
        __syncthreads();
        int base = tile * kSortTile + warp * kSortPerWarp;
        unsigned key[kSortPerWarp / 32];
        int rank[kSortPerWarp / 32];
        for (int s = 0; s < kSortPerWarp / 32; ++s) {
            int i = base + 32 * s + lane;
            bool valid = i < n;
            key[s] = valid ? keys[i] : 0u;
            int d = valid ? (int)((key[s] >> shift) & 255u) : 256 + lane;  // lanes past the end match nobody
            unsigned peers = __match_any_sync(0xffffffffu, d);
            int before = __popc(peers & ((1u << lane) - 1u));
            int leader = __ffs(peers) - 1;
            int start = 0;
            if (valid && lane == leader) {
                start = warpCount[warp][d];
                warpCount[warp][d] = start + __popc(peers);
            }
            start = __shfl_sync(0xffffffffu, start, leader);
            rank[s] = start + before;
            __syncwarp();
        }
        __syncthreads();
        {   // per digit: exclusive prefix over the warps, plus the tile's global start
            int d = threadIdx.x, acc = offsets[(size_t)d * nTiles + tile];
            for (int w = 0; w < kSortWarps; ++w) {
                int c = warpCount[w][d];
                warpCount[w][d] = acc;
                acc += c;
            }
        }
        __syncthreads();
        for (int s = 0; s < kSortPerWarp / 32; ++s) {
            int i = base + 32 * s + lane;
            if (i < n) {
                int dst = warpCount[warp][(key[s] >> shift) & 255u] + rank[s];
                keysOut[dst] = key[s];
                valsOut[dst] = vals[i];
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------- treelets and emitLBVH
// a treelet starts where the top 12 of the 30 Morton bits change (mask 0x3ffc0000, bvh.cpp:436-441); sorted codes, so each
// prefix has one run and start[prefix] names it
__global__ void k_treelet_starts(const unsigned *keys, int n, int *start /* [4096], preset to -1 */) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned p = keys[i] >> 18;
        if (i == 0 || (keys[i - 1] >> 18) != p) start[p] = i;
    }
}

struct TempNodes {  // one record per range that becomes a node; children of a split are adjacent (left, left + 1)
    int *first, *count, *bit;  // the range in the sorted order and the Morton bit to look at next (in); bit becomes the split axis (out)
    int *left;                 // -1: leaf
    int *size;                 // nodes in the subtree
    int *finalIdx;             // index in the flattened array
    float *bounds;             // 6 per record
};

__global__ void k_emit_level(TempNodes t, const int *levelStart, int level, const unsigned *keys, const int *vals, const float *primBounds,
                             int maxPrimsInNode, int *nTemp) {
    const int begin = levelStart[level], end = levelStart[level + 1];
    for (int slot = begin + blockIdx.x * blockDim.x + threadIdx.x; slot < end; slot += gridDim.x * blockDim.x) {
        const int first = t.first[slot], count = t.count[slot];
        int bit = t.bit[slot];
        for (;;) {
            if (bit == -1 || count < maxPrimsInNode) {  // bvh.cpp:481-498
                float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f}, mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
                for (int i = 0; i < count; ++i) {
                    const float *b = primBounds + 6 * (size_t)vals[first + i];
                    for (int k = 0; k < 3; ++k) mn[k] = minf(mn[k], b[k]), mx[k] = maxf(mx[k], b[3 + k]);
                }
                float *o = t.bounds + 6 * (size_t)slot;
                o[0] = mn[0], o[1] = mn[1], o[2] = mn[2], o[3] = mx[0], o[4] = mx[1], o[5] = mx[2];
                t.left[slot] = -1;
                t.size[slot] = 1;
                break;
            }
            const unsigned mask = 1u << bit;
            if ((keys[first] & mask) == (keys[first + count - 1] & mask)) {  // bvh.cpp:501-505: this bit does not split the range
                --bit;
                continue;
            }
            int searchStart = 0, searchEnd = count - 1;  // bvh.cpp:508-518
            while (searchStart + 1 != searchEnd) {
                int mid = (searchStart + searchEnd) / 2;
                if ((keys[first + searchStart] & mask) == (keys[first + mid] & mask)) searchStart = mid;
                else searchEnd = mid;
            }
            const int split = searchEnd;
            const int c = atomicAdd(nTemp, 2);
            t.left[slot] = c;
            t.bit[slot] = bit % 3;  // the split axis (bvh.cpp:533)
            t.first[c] = first, t.count[c] = split, t.bit[c] = bit - 1;
            t.first[c + 1] = first + split, t.count[c + 1] = count - split, t.bit[c + 1] = bit - 1;
            break;
        }
    }
}
__global__ void k_level_end(int *levelStart, int level, const int *nTemp) { levelStart[level + 2] = *nTemp; }

__global__ void k_up_level(TempNodes t, const int *levelStart, int level) {
    const int begin = levelStart[level], end = levelStart[level + 1];
    for (int slot = begin + blockIdx.x * blockDim.x + threadIdx.x; slot < end; slot += gridDim.x * blockDim.x) {
        int l = t.left[slot];
        if (l < 0) continue;
        t.size[slot] = 1 + t.size[l] + t.size[l + 1];
        const float *a = t.bounds + 6 * (size_t)l, *b = a + 6;
        float *o = t.bounds + 6 * (size_t)slot;
        for (int k = 0; k < 3; ++k) o[k] = minf(a[k], b[k]), o[3 + k] = maxf(a[3 + k], b[3 + k]);
    }
}
struct FlatNodes {
    float *bounds;
    int *offset;
    unsigned short *nPrims;
    unsigned char *axis;
};
__global__ void k_down_level(TempNodes t, const int *levelStart, int level, FlatNodes out) {
    const int begin = levelStart[level], end = levelStart[level + 1];
    for (int slot = begin + blockIdx.x * blockDim.x + threadIdx.x; slot < end; slot += gridDim.x * blockDim.x) {
        const int f = t.finalIdx[slot], l = t.left[slot];
        const float *b = t.bounds + 6 * (size_t)slot;
        float *o = out.bounds + 6 * (size_t)f;
        for (int k = 0; k < 6; ++k) o[k] = b[k];
        if (l < 0) {  // leaf: slots are handed out in Morton order, so the first slot is the range's first position
            out.offset[f] = t.first[slot];
            out.nPrims[f] = (unsigned short)t.count[slot];
            out.axis[f] = 0;
        } else {
            const int second = f + 1 + t.size[l];
            t.finalIdx[l] = f + 1;
            t.finalIdx[l + 1] = second;
            out.offset[f] = second;
            out.nPrims[f] = 0;
            out.axis[f] = (unsigned char)t.bit[slot];
        }
    }
}

struct Scratch {  // freed on every way out
    std::vector<void *> ptrs;
    ~Scratch() {
        for (void *p : ptrs) cudaFree(p);
    }
    template <typename T>
    cudaError_t alloc(T **p, size_t count) {
        cudaError_t e = cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
};

}  // namespace

extern "C" int pbrtgpu_build_hlbvh(pbrtgpu_ctx *ctx, const float *prim_bounds, int32_t n, int32_t max_prims_in_node, int32_t node_capacity,
                                   float *bvh_bounds, int32_t *bvh_offset, uint16_t *bvh_nprims, uint8_t *bvh_axis, int32_t *ordered_prims,
                                   int32_t *n_nodes_out, float *elapsed_ms) {
    if (!ctx) return PBRTGPU_ERR_INVALID;
    if (!prim_bounds || n <= 0 || !bvh_bounds || !bvh_offset || !bvh_nprims || !bvh_axis || !ordered_prims || !n_nodes_out)
        return ctx->fail(PBRTGPU_ERR_INVALID, "pbrtgpu_build_hlbvh: null argument or no primitives");
    if ((long long)node_capacity < 2ll * n) return ctx->fail(PBRTGPU_ERR_INVALID, "pbrtgpu_build_hlbvh: node_capacity must be at least 2 n");
    PG_CUDA(ctx, cudaSetDevice(ctx->device));
    const int maxPrims = std::min(255, max_prims_in_node);  // BVHAccel's constructor, bvh.cpp:184
    cudaStream_t st = ctx->stream;
    const int grid = ctx->smCount * 8;

    Scratch sc;
    float *dBounds;
    unsigned *dKeys[2], *dCB;
    int *dVals[2], *dHist, *dSums, *dStart, *dLevelStart, *dNTemp;
    const int nTiles = (n + kSortTile - 1) / kSortTile;
    const size_t cap = 2 * (size_t)n;
    TempNodes t;
    FlatNodes flat;
    PG_CUDA(ctx, sc.alloc(&dBounds, 6 * (size_t)n));
    for (int k = 0; k < 2; ++k) {
        PG_CUDA(ctx, sc.alloc(&dKeys[k], n));
        PG_CUDA(ctx, sc.alloc(&dVals[k], n));
    }
    PG_CUDA(ctx, sc.alloc(&dCB, 6));
    PG_CUDA(ctx, sc.alloc(&dHist, 256 * (size_t)nTiles));
    PG_CUDA(ctx, sc.alloc(&dSums, (size_t)scanScratchInts(256ll * nTiles)));
    PG_CUDA(ctx, sc.alloc(&dStart, 4096));
    PG_CUDA(ctx, sc.alloc(&dLevelStart, kLevels + 2));
    PG_CUDA(ctx, sc.alloc(&dNTemp, 1));
    PG_CUDA(ctx, sc.alloc(&t.first, cap));
    PG_CUDA(ctx, sc.alloc(&t.count, cap));
    PG_CUDA(ctx, sc.alloc(&t.bit, cap));
    PG_CUDA(ctx, sc.alloc(&t.left, cap));
    PG_CUDA(ctx, sc.alloc(&t.size, cap));
    PG_CUDA(ctx, sc.alloc(&t.finalIdx, cap));
    PG_CUDA(ctx, sc.alloc(&t.bounds, 6 * cap));
    PG_CUDA(ctx, sc.alloc(&flat.bounds, 6 * cap));
    PG_CUDA(ctx, sc.alloc(&flat.offset, cap));
    PG_CUDA(ctx, sc.alloc(&flat.nPrims, cap));
    PG_CUDA(ctx, sc.alloc(&flat.axis, cap));

    cudaEvent_t ev[2] = {nullptr, nullptr};
    struct EvGuard {
        cudaEvent_t *e;
        ~EvGuard() {
            for (int k = 0; k < 2; ++k)
                if (e[k]) cudaEventDestroy(e[k]);
        }
    } evGuard{ev};
    PG_CUDA(ctx, cudaEventCreate(&ev[0]));
    PG_CUDA(ctx, cudaEventCreate(&ev[1]));
    float msDevice = 0;

    PG_CUDA(ctx, cudaMemcpyAsync(dBounds, prim_bounds, 24 * (size_t)n, cudaMemcpyHostToDevice, st));
    // ---- Morton codes, sort, treelet starts
    PG_CUDA(ctx, cudaEventRecord(ev[0], st));
    {
        const unsigned init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
        PG_CUDA(ctx, cudaMemcpyAsync(dCB, init, sizeof(init), cudaMemcpyHostToDevice, st));
    }
    k_centroid_bounds<<<grid, 256, 0, st>>>(dBounds, n, dCB);
    k_morton<<<grid, 256, 0, st>>>(dBounds, n, dCB, dKeys[0], dVals[0]);
    int cur = 0;
    for (int shift = 0; shift < 32; shift += 8) {  // the codes have 30 bits
        k_radix_hist<<<std::min(nTiles, grid), kSortThreads, 0, st>>>(dKeys[cur], n, shift, nTiles, dHist);
        ctx->launches += 1;
        int rc = exclusiveScanInt(ctx, dHist, dHist, 256ll * nTiles, dSums, nullptr);
        if (rc != PBRTGPU_OK) return rc;
        k_radix_scatter<<<std::min(nTiles, grid), kSortThreads, 0, st>>>(dKeys[cur], dVals[cur], n, shift, nTiles, dHist, dKeys[cur ^ 1], dVals[cur ^ 1]);
        ctx->launches += 1;
        cur ^= 1;
    }
    const unsigned *keys = dKeys[cur];
    const int *vals = dVals[cur];
    PG_CUDA(ctx, cudaMemsetAsync(dStart, 0xff, 4096 * sizeof(int), st));
    k_treelet_starts<<<grid, 256, 0, st>>>(keys, n, dStart);
    ctx->launches += 3;
    PG_CUDA(ctx, cudaEventRecord(ev[1], st));
    std::vector<int> startByPrefix(4096);
    PG_CUDA(ctx, cudaMemcpyAsync(startByPrefix.data(), dStart, 4096 * sizeof(int), cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaStreamSynchronize(st));
    {
        float ms = 0;
        PG_CUDA(ctx, cudaEventElapsedTime(&ms, ev[0], ev[1]));
        msDevice += ms;
    }
    std::vector<int> tFirst, tCount;
    for (int p = 0; p < 4096; ++p)
        if (startByPrefix[p] >= 0) tFirst.push_back(startByPrefix[p]);
    const int nT = (int)tFirst.size();
    if (nT == 0) return ctx->fail(PBRTGPU_ERR_CUDA, "pbrtgpu_build_hlbvh: no treelet found (sort failed?)");
    for (int k = 0; k < nT; ++k) tCount.push_back((k + 1 < nT ? tFirst[k + 1] : n) - tFirst[k]);
    for (int k = 0; k < nT; ++k)
        if (tCount[k] <= 0) return ctx->fail(PBRTGPU_ERR_CUDA, "pbrtgpu_build_hlbvh: treelet ranges are not increasing (sort failed?)");

    // ---- emitLBVH for every treelet, level by level; then sizes and bounds bottom-up
    PG_CUDA(ctx, cudaEventRecord(ev[0], st));
    {
        std::vector<int> bit(nT, 29 - 12), levelStart(kLevels + 2, 0);
        levelStart[1] = nT;
        PG_CUDA(ctx, cudaMemcpyAsync(t.first, tFirst.data(), nT * sizeof(int), cudaMemcpyHostToDevice, st));
        PG_CUDA(ctx, cudaMemcpyAsync(t.count, tCount.data(), nT * sizeof(int), cudaMemcpyHostToDevice, st));
        PG_CUDA(ctx, cudaMemcpyAsync(t.bit, bit.data(), nT * sizeof(int), cudaMemcpyHostToDevice, st));
        PG_CUDA(ctx, cudaMemcpyAsync(dLevelStart, levelStart.data(), levelStart.size() * sizeof(int), cudaMemcpyHostToDevice, st));
        PG_CUDA(ctx, cudaMemcpyAsync(dNTemp, &nT, sizeof(int), cudaMemcpyHostToDevice, st));
        PG_CUDA(ctx, cudaStreamSynchronize(st));  // the staging vectors go out of scope
    }
    for (int level = 0; level < kLevels; ++level) {
        k_emit_level<<<grid, 128, 0, st>>>(t, dLevelStart, level, keys, vals, dBounds, maxPrims, dNTemp);
        k_level_end<<<1, 1, 0, st>>>(dLevelStart, level, dNTemp);
    }
    for (int level = kLevels - 1; level >= 0; --level) k_up_level<<<grid, 128, 0, st>>>(t, dLevelStart, level);
    ctx->launches += 3 * kLevels;
    PG_CUDA(ctx, cudaEventRecord(ev[1], st));
    // ---- the upper tree (host): roots' bounds and subtree sizes come back, DFS numbers of the roots go out
    std::vector<float> rootB(6 * (size_t)nT);
    std::vector<int> rootSize(nT), levelStartHost(kLevels + 2);
    PG_CUDA(ctx, cudaMemcpyAsync(rootB.data(), t.bounds, rootB.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaMemcpyAsync(rootSize.data(), t.size, nT * sizeof(int), cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaMemcpyAsync(levelStartHost.data(), dLevelStart, levelStartHost.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaStreamSynchronize(st));
    {
        float ms = 0;
        PG_CUDA(ctx, cudaEventElapsedTime(&ms, ev[0], ev[1]));
        msDevice += ms;
    }
    if (levelStartHost[kLevels + 1] != levelStartHost[kLevels])
        return ctx->fail(PBRTGPU_ERR_CUDA, "pbrtgpu_build_hlbvh: ranges left open after %d levels", kLevels);
    std::vector<UBox> rb(nT);
    for (int k = 0; k < nT; ++k)
        for (int a = 0; a < 3; ++a) rb[k].mn[a] = rootB[6 * (size_t)k + a], rb[k].mx[a] = rootB[6 * (size_t)k + 3 + a];
    UpperSAH up;
    up.rootBounds = &rb;
    std::vector<int> order(nT);
    for (int k = 0; k < nT; ++k) order[k] = k;
    const int rootRef = up.build(order, 0, nT);
    // flattenBVHTree over the upper nodes; a treelet is a block of rootSize[k] consecutive indices
    std::vector<int> rootFinal(nT, -1);
    std::vector<float> upBounds;
    std::vector<int> upIndex, upOffset;
    std::vector<unsigned char> upAxis;
    int total = 0;
    {
        struct Frame {
            int ref, parentEntry;  // parentEntry: which upper entry waits for this subtree's index as its second child
        };
        std::vector<Frame> stack{{rootRef, -1}};
        while (!stack.empty()) {
            Frame f = stack.back();
            stack.pop_back();
            if (f.parentEntry >= 0) upOffset[f.parentEntry] = total;
            if (f.ref < 0) {
                rootFinal[~f.ref] = total;
                total += rootSize[~f.ref];
                continue;
            }
            const UpperNode &u = up.nodes[f.ref];
            int entry = (int)upIndex.size();
            upIndex.push_back(total++);
            upOffset.push_back(-1);
            upAxis.push_back((unsigned char)u.axis);
            for (int a = 0; a < 3; ++a) upBounds.push_back(u.bounds.mn[a]);
            for (int a = 0; a < 3; ++a) upBounds.push_back(u.bounds.mx[a]);
            stack.push_back({u.child[1], entry});  // second child: after the whole first subtree
            stack.push_back({u.child[0], -1});
        }
    }
    if (total > node_capacity) return ctx->fail(PBRTGPU_ERR_INVALID, "pbrtgpu_build_hlbvh: %d nodes do not fit node_capacity", total);
    // ---- DFS numbers top-down, final arrays
    PG_CUDA(ctx, cudaEventRecord(ev[0], st));
    PG_CUDA(ctx, cudaMemcpyAsync(t.finalIdx, rootFinal.data(), nT * sizeof(int), cudaMemcpyHostToDevice, st));
    for (int level = 0; level < kLevels; ++level) k_down_level<<<grid, 128, 0, st>>>(t, dLevelStart, level, flat);
    ctx->launches += kLevels;
    PG_CUDA(ctx, cudaEventRecord(ev[1], st));
    PG_CUDA(ctx, cudaMemcpyAsync(bvh_bounds, flat.bounds, 24 * (size_t)total, cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaMemcpyAsync(bvh_offset, flat.offset, 4 * (size_t)total, cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaMemcpyAsync(bvh_nprims, flat.nPrims, 2 * (size_t)total, cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaMemcpyAsync(bvh_axis, flat.axis, (size_t)total, cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaMemcpyAsync(ordered_prims, vals, 4 * (size_t)n, cudaMemcpyDeviceToHost, st));
    PG_CUDA(ctx, cudaStreamSynchronize(st));
    PG_CUDA(ctx, cudaGetLastError());
    {
        float ms = 0;
        PG_CUDA(ctx, cudaEventElapsedTime(&ms, ev[0], ev[1]));
        msDevice += ms;
    }
    for (size_t e = 0; e < upIndex.size(); ++e) {  // the upper nodes were made on the host: written straight into the result
        int f = upIndex[e];
        for (int a = 0; a < 6; ++a) bvh_bounds[6 * (size_t)f + a] = upBounds[6 * e + a];
        bvh_offset[f] = upOffset[e];
        bvh_nprims[f] = 0;
        bvh_axis[f] = upAxis[e];
    }
    *n_nodes_out = total;
    if (elapsed_ms) *elapsed_ms = msDevice;
    return PBRTGPU_OK;
}
