// nbd_table.cuh — many small broadcasts in ONE launch with ONE flag barrier.
//
// The reference's own broadcast idiom is a per-parameter loop (ref:README.md:122-123:
// `for param in model.parameters(): dist.broadcast(param.data, src=0)`); BASELINE config 3 is 149
// such calls, 98 of them <= 12 KB, each paying a launch and a flag round trip.  Issued inside a
// coalescing block (c10d startCoalescing/endCoalescing, $TORCH/.../Backend.hpp:109-123) the small
// ones are collected into a table and run here: every rank packs the items it is root of into
// its peers' staging at the item's offset in a common packed layout (16-byte units), one flag
// barrier, every rank unpacks the items it is not root of.  Items may have different roots.
//
// k_bcast_flat is the same idea for tables too large for the direct push (root egress (N-1)*S):
// the packed layout of the items is treated as ONE message and moved the way k_broadcast's
// "scatter" algorithm moves a single tensor (chunked, each chunk cut into N-1 shards, root writes
// shard j into its owner's staging, every non-root rank pulls the other shards), with every user
// access translated from packed unit to (item, offset).  One root per launch.  Config 3's 51
// tensors of 2-154 MB are then one pipeline instead of 51 launches with their own barriers.
#pragma once

#include "nbd_kernels.cuh"

namespace nbd {

constexpr int kTableMax = 256;  // items per launch (BASELINE config 3, GPT-2 small, has 149)

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

// Packed unit -> the item that holds it, remembered between look-ups: consecutive accesses of a
// thread mostly stay inside one item, so the binary search runs only when they leave it.
struct FlatCursor {
  int k;
  u32 lo, hi;
};

template <int NCTX>
__device__ __forceinline__ Region flat_seek(const PackTable<NCTX>& pk, int ci, FlatCursor& cur, u64 u, u64& local) {
  if (u < cur.lo || u >= cur.hi) {
    cur.k = table_find(pk.d, (u32)u);
    cur.lo = pk.d.first[cur.k];
    cur.hi = pk.d.first[cur.k + 1];
  }
  local = u - cur.lo;
  return make_region(pk.buf[ci][cur.k], pk.d.bytes[cur.k]);
}

// op.bytes = packed size of the table (a multiple of kUnit), op.root = the root of every item,
// op.chunk_bytes a multiple of kUnit*(N-1) that fits (N-1) staging halves (plan_flat).
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_bcast_flat(const __grid_constant__ PackTable<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const int N = c.world, r = c.rank, root = op.root;
  KernelClock clk = kernel_begin(c);
  const u64 nsh = (u64)(N - 1);
  const u64 chunk_units = op.chunk_bytes / kUnit, total_units = op.bytes / kUnit;
  const u64 full_su = (chunk_units + nsh - 1) / nsh;
  FlatCursor cur = {0, 0, 0};
  auto shard_of = [&](int p) { int j = p - root - 1; if (j < 0) j += N; return (u64)j; };

  // units [lo, hi) of shard j of chunk cidx, as this CTA sees them; v0 = packed unit of shard unit 0
  auto shard_span = [&](u32 cidx, u64 j, u64 lo, u64 hi, u64& slo, u64& shi, u64& v0) {
    const u64 off = (u64)cidx * chunk_units;
    const u64 cu = (total_units - off < chunk_units) ? total_units - off : chunk_units;
    const u64 su = (cu + nsh - 1) / nsh;
    const u64 ru = (j * su >= cu) ? 0 : ((cu - j * su < su) ? cu - j * su : su);
    slo = lo < ru ? lo : ru;
    shi = hi < ru ? hi : ru;
    v0 = off + j * su;
  };

  auto root_push = [&](u32 cidx) {
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(full_su, full_su, lo, hi);
#pragma unroll 1
    for (u64 base = lo; base < hi; base += 4ull * kThreads) {
      const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
      for (int k = 1; k < N; ++k) {
        int p = r + k;
        if (p >= N) p -= N;
        u64 slo, shi, v0;
        shard_span(cidx, shard_of(p), base, bhi, slo, shi, v0);
        char* dst = c.stage[p] + half;
        copy_units<4>(slo, shi,
                      [&](u64 i) {
                        u64 li;
                        const Region src = flat_seek(pk, ci, cur, v0 + i, li);
                        return region_ld(src, region_aligned(src), li);
                      },
                      [&](u64 i, const uint4& v) { st16(dst + i * kUnit, v); });
      }
    }
  };

  if (r == root && op.nchunks) root_push(0);
  for (u32 cidx = 0; cidx < op.nchunks; ++cidx) {
    flag_barrier(c, ++clk.epoch);
    if (r == root) {
      if (cidx + 1 < op.nchunks) root_push(cidx + 1);
      continue;
    }
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(full_su, full_su, lo, hi);
#pragma unroll 1
    for (u64 base = lo; base < hi; base += 4ull * kThreads) {
      const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
      for (int k = 0; k < N; ++k) {  // k == 0: own shard from own staging
        int p = r + k;
        if (p >= N) p -= N;
        if (p == root) continue;
        u64 slo, shi, v0;
        shard_span(cidx, shard_of(p), base, bhi, slo, shi, v0);
        const char* src = c.stage[p] + half;
        copy_units<4>(slo, shi, [&](u64 i) { return ld16(src + i * kUnit); },
                      [&](u64 i, const uint4& v) {
                        u64 li;
                        const Region dst = flat_seek(pk, ci, cur, v0 + i, li);
                        region_st(dst, region_aligned(dst), li, v);
                      });
      }
    }
  }
  kernel_end(c, op.nchunks, op.nchunks);
}

}  // namespace nbd
