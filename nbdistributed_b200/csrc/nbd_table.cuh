// nbd_table.cuh — many small broadcasts in ONE launch with ONE flag barrier.
//
// The reference's own broadcast idiom is a per-parameter loop (ref:README.md:122-123:
// `for param in model.parameters(): dist.broadcast(param.data, src=0)`); BASELINE config 3 is 149
// such calls, 98 of them <= 12 KB, each paying a launch and a flag round trip.  Issued inside a
// coalescing block (c10d startCoalescing/endCoalescing, $TORCH/.../Backend.hpp:109-123) the small
// ones are collected into a table and run here: every rank packs the items it is root of into
// its peers' staging at the item's offset in a common packed layout (16-byte units), one flag
// barrier, every rank unpacks the items it is not root of.  Items may have different roots.
#pragma once

#include "nbd_kernels.cuh"

namespace nbd {

constexpr int kTableMax = 128;  // items per launch

struct TableDesc {
  int nitems;
  int pad;
  u32 first[kTableMax + 1];  // packed offset of item k in units; first[nitems] = total units
  u64 bytes[kTableMax];
  unsigned char root[kTableMax];
};

template <int NCTX>
struct PackTable {
  RankCtx ctx[NCTX];
  TableDesc d;
  char* buf[NCTX][kTableMax];
};

// index of the item that holds packed unit u
__device__ __forceinline__ int table_find(const TableDesc& d, u32 u) {
  int lo = 0, hi = d.nitems - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (d.first[mid] <= u) lo = mid;
    else hi = mid - 1;
  }
  return lo;
}

template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_bcast_table(const __grid_constant__ PackTable<NCTX> pk) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const TableDesc& d = pk.d;
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  const u64 half = (clk.seq & 1) * c.half_bytes;
  const u64 total = d.first[d.nitems];
  u64 lo, hi;
  block_span(total, total, lo, hi);
  for (u64 u = lo + threadIdx.x; u < hi; u += kThreads) {
    const int k = table_find(d, (u32)u);
    if (d.root[k] != r) continue;
    const Region src = make_region(pk.buf[ci][k], d.bytes[k]);
    const uint4 v = region_ld(src, region_aligned(src), u - d.first[k]);
#pragma unroll
    for (int q = 1; q < kMaxRanks; ++q) {
      if (q < N) {
        int p = r + q;
        if (p >= N) p -= N;
        st16(c.stage[p] + half + u * kUnit, v);
      }
    }
  }
  flag_barrier(c, ++clk.epoch);
  for (u64 u = lo + threadIdx.x; u < hi; u += kThreads) {
    const int k = table_find(d, (u32)u);
    if (d.root[k] == r) continue;
    const Region dst = make_region(pk.buf[ci][k], d.bytes[k]);
    region_st(dst, region_aligned(dst), u - d.first[k], ld16(c.stage[r] + half + u * kUnit));
  }
  kernel_end(c, 1, 1);
}

}  // namespace nbd
