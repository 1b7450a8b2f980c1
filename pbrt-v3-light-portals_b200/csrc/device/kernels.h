// Launch geometry shared by the kernels.
#ifndef PG_DEVICE_KERNELS_H
#define PG_DEVICE_KERNELS_H

struct pbrtgpu_ctx;

namespace pg {

// 148 SMs on a B200; grids are sized in whole multiples of the SM count and
// loop over the work (grid-stride), so no tail wave is left partly filled.
constexpr int kTraceBlock = 128;
constexpr int kTraceBlocksPerSM = 8;
#ifndef PG_PERSIST_MIN_BLOCKS
#define PG_PERSIST_MIN_BLOCKS 8
#endif
constexpr int kPersistMinBlocks = PG_PERSIST_MIN_BLOCKS;  // persistent trace kernels: resident blocks per SM the register budget allows

int traceGrid(const pbrtgpu_ctx *ctx, long long n);

}  // namespace pg
#endif
