// nbd_kernels.cuh — the collective kernels.  Each is ONE launch per rank that
// moves data between the caller's (unregistered) buffers and the IPC-mapped
// staging arenas of every rank with 128-bit accesses, synchronising CTA b of
// every rank through release/acquire flags.  Large messages are software-
// pipelined through the two staging halves in chunks:
//
//   all_reduce one-shot : push(0); for c: barrier; push(c+1); reduce(c)
//   all_reduce two-shot : push(0); for c: barrierA; push(c+1); reduce(c);
//                                          barrierB; pull(c)
//   reduce_scatter      : push(0); for c: barrier; push(c+1); reduce(c)
//   all_gather          : copyin(0); for c: barrier; copyin(c+1); pull(c)
//   broadcast (root)    : push(0); for c: barrier; push(c+1)
//   broadcast (others)  :          for c: barrier; pull/copyout(c)
//
// Staging-half reuse is safe because a rank signals barrier(c) only after it
// finished every access of step c-1, and nobody touches half (c+1)%2 for step
// c+1 before it has waited on barrier(c).  Barriers are per CTA index, so CTA b
// must own the SAME staging offsets in every chunk of a launch: spans are cut
// from the full chunk size and merely clipped for a ragged last chunk.
// See DESIGN.md "Flag protocol".
#pragma once

#include "nbd_device.cuh"

namespace nbd {

template <int NCTX>
__device__ __forceinline__ int ctx_index() {
  return NCTX == 1 ? 0 : (int)blockIdx.y;
}

__device__ __forceinline__ u64 clip_bytes(u64 total, u64 begin, u64 len) {
  if (begin >= total) return 0;
  const u64 rest = total - begin;
  return rest < len ? rest : len;
}

// N-way reduce of units [lo,hi): source k is this rank's staging slot k, except
// k == rank which is read straight from the user's send buffer.  Ascending k.
template <typename T, bool SUM, bool TO_STAGE>
__device__ __forceinline__ void reduce_units(const RankCtx& c, const Region& own_src, const char* slots,
                                             u64 slot_stride, const Region& dst, char* stage_out, u64 lo,
                                             u64 hi, int op, int avg) {
  const int N = c.world, r = c.rank;
  const bool als = region_aligned(own_src), ald = region_aligned(dst);
  for (u64 i = lo + threadIdx.x; i < hi; i += kThreads) {
    uint4 v[kMaxRanks];
#pragma unroll
    for (int k = 0; k < kMaxRanks; ++k) {
      if (k < N) v[k] = (k == r) ? region_ld(own_src, als, i) : ld16(slots + (u64)k * slot_stride + i * kUnit);
    }
    Reducer<T> R;
    R.init(v[0]);
#pragma unroll
    for (int k = 1; k < kMaxRanks; ++k) {
      if (k < N) R.template add<SUM>(v[k], op);
    }
    const uint4 out = R.result(avg, N);
    region_st(dst, ald, i, out);
    if (TO_STAGE) st16(stage_out + i * kUnit, out);
  }
}

// ---------------------------------------------------------------------------
// all_reduce, one-shot: every rank receives every peer's whole chunk.
// Wire bytes per GPU: (N-1)*S each way; one barrier per chunk.  Latency path
// and the N == 2 path (where it moves no more than two-shot).
// ---------------------------------------------------------------------------
template <typename T, bool SUM, int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_allreduce_oneshot(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  const u64 slot_stride = units_of(op.chunk_bytes) * kUnit;

  auto push = [&](u32 cidx) {
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const Region src = make_region(ub.send + off, cb);
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    const bool al = region_aligned(src);
    // one read of the user chunk, N-1 remote writes, peers interleaved
    for (u64 base = lo + threadIdx.x; base < hi; base += 2ull * kThreads) {
      uint4 v[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const u64 i = base + (u64)j * kThreads;
        if (i < hi) v[j] = region_ld(src, al, i);
      }
#pragma unroll
      for (int k = 1; k < kMaxRanks; ++k) {
        if (k < N) {
          int p = r + k;
          if (p >= N) p -= N;
          char* dst = c.stage[p] + half + (u64)r * slot_stride;
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const u64 i = base + (u64)j * kThreads;
            if (i < hi) st16(dst + i * kUnit, v[j]);
          }
        }
      }
    }
  };

  if (op.nchunks) push(0);
  for (u32 cidx = 0; cidx < op.nchunks; ++cidx) {
    flag_barrier(c, ++clk.epoch);
    if (cidx + 1 < op.nchunks) push(cidx + 1);
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    reduce_units<T, SUM, false>(c, make_region(ub.send + off, cb), c.stage[r] + half, slot_stride,
                                make_region(ub.recv + off, cb), nullptr, lo, hi, op.op, op.avg);
  }
  kernel_end(c, op.nchunks, op.nchunks);
}

// ---------------------------------------------------------------------------
// all_reduce, two-shot: reduce-scatter into the shard owner, then every rank
// pulls the reduced shards.  Wire bytes per GPU: (N-1)/N*S each way per phase.
// Half layout: IN slot k (k = source rank) at k*slot_stride, OUT (this rank's
// reduced shard) at N*slot_stride.
// ---------------------------------------------------------------------------
template <typename T, bool SUM, int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_allreduce_twoshot(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  const u64 slot_stride = (units_of(op.chunk_bytes) + N - 1) / N * kUnit;  // max shard bytes

  auto shard_units = [&](u64 cb) { return (units_of(cb) + N - 1) / N; };

  auto push = [&](u32 cidx) {
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 su = shard_units(cb);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(shard_units(op.chunk_bytes), su, lo, hi);
#pragma unroll 1
    for (u64 base = lo; base < hi; base += 4ull * kThreads) {
      const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
      for (int k = 1; k < N; ++k) {
        int p = r + k;
        if (p >= N) p -= N;
        const u64 sb = clip_bytes(cb, (u64)p * su * kUnit, su * kUnit);
        const Region src = make_region(ub.send + off + (u64)p * su * kUnit, sb);
        const u64 ru = units_of(sb);
        copy_region_to_stage<4>(c.stage[p] + half + (u64)r * slot_stride, src, base < ru ? base : ru,
                                bhi < ru ? bhi : ru);
      }
    }
  };

  if (op.nchunks) push(0);
  for (u32 cidx = 0; cidx < op.nchunks; ++cidx) {
    flag_barrier(c, ++clk.epoch);
    if (cidx + 1 < op.nchunks) push(cidx + 1);

    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 su = shard_units(cb);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(shard_units(op.chunk_bytes), su, lo, hi);
    {
      const u64 sb = clip_bytes(cb, (u64)r * su * kUnit, su * kUnit);
      const u64 ru = units_of(sb);
      reduce_units<T, SUM, true>(c, make_region(ub.send + off + (u64)r * su * kUnit, sb), c.stage[r] + half,
                                 slot_stride, make_region(ub.recv + off + (u64)r * su * kUnit, sb),
                                 c.stage[r] + half + (u64)N * slot_stride, lo < ru ? lo : ru, hi < ru ? hi : ru,
                                 op.op, op.avg);
    }
    flag_barrier(c, ++clk.epoch);
#pragma unroll 1
    for (u64 base = lo; base < hi; base += 4ull * kThreads) {
      const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
      for (int k = 1; k < N; ++k) {
        int p = r + k;
        if (p >= N) p -= N;
        const u64 sb = clip_bytes(cb, (u64)p * su * kUnit, su * kUnit);
        const u64 ru = units_of(sb);
        copy_stage_to_region<4>(make_region(ub.recv + off + (u64)p * su * kUnit, sb),
                                c.stage[p] + half + (u64)N * slot_stride, base < ru ? base : ru,
                                bhi < ru ? bhi : ru);
      }
    }
  }
  kernel_end(c, 2ull * op.nchunks, op.nchunks);
}

// ---------------------------------------------------------------------------
// reduce_scatter: rank r pushes its copy of shard p to rank p, rank p reduces.
// op.bytes = bytes of one shard (= recv bytes).  Half layout: IN slot k.
// ---------------------------------------------------------------------------
template <typename T, bool SUM, int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_reduce_scatter(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  const u64 slot_stride = units_of(op.chunk_bytes) * kUnit;

  auto push = [&](u32 cidx) {
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
#pragma unroll 1
    for (u64 base = lo; base < hi; base += 4ull * kThreads) {
      const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
      for (int k = 1; k < N; ++k) {
        int p = r + k;
        if (p >= N) p -= N;
        copy_region_to_stage<4>(c.stage[p] + half + (u64)r * slot_stride,
                                make_region(ub.send + (u64)p * op.bytes + off, cb), base, bhi);
      }
    }
  };

  if (op.nchunks) push(0);
  for (u32 cidx = 0; cidx < op.nchunks; ++cidx) {
    flag_barrier(c, ++clk.epoch);
    if (cidx + 1 < op.nchunks) push(cidx + 1);
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
    reduce_units<T, SUM, false>(c, make_region(ub.send + (u64)r * op.bytes + off, cb), c.stage[r] + half,
                                slot_stride, make_region(ub.recv + off, cb), nullptr, lo, hi, op.op, op.avg);
  }
  kernel_end(c, op.nchunks, op.nchunks);
}

// ---------------------------------------------------------------------------
// all_gather: stage own shard chunk locally, peers pull it.  op.bytes = bytes
// of one rank's shard (= send bytes).
// ---------------------------------------------------------------------------
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_all_gather(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);

  auto copyin = [&](u32 cidx) {
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
    copy_region_to_stage<4>(c.stage[r] + half, make_region(ub.send + off, cb), lo, hi);
  };

  if (op.nchunks) copyin(0);
  for (u32 cidx = 0; cidx < op.nchunks; ++cidx) {
    flag_barrier(c, ++clk.epoch);
    if (cidx + 1 < op.nchunks) copyin(cidx + 1);
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
#pragma unroll 1
    for (u64 base = lo; base < hi; base += 4ull * kThreads) {
      const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
      for (int k = 1; k < N; ++k) {
        int p = r + k;
        if (p >= N) p -= N;
        copy_stage_to_region<4>(make_region(ub.recv + (u64)p * op.bytes + off, cb), c.stage[p] + half, base, bhi);
      }
    }
    // own shard: straight from send to its slot of recv (no-op when aliased)
    if (ub.send + off != ub.recv + (u64)r * op.bytes + off) {
      copy_region_to_region<4>(make_region(ub.recv + (u64)r * op.bytes + off, cb), make_region(ub.send + off, cb),
                               lo, hi);
    }
  }
  kernel_end(c, op.nchunks, op.nchunks);
}

// ---------------------------------------------------------------------------
// broadcast.  algo 0 ("direct"): root writes the chunk into every peer's
// staging half; peers copy it out.  Root egress (N-1)*S: latency path.
// algo 1 ("scatter"): the chunk is cut into N-1 shards, one per non-root rank;
// root writes shard j into its owner's staging, then every non-root rank pulls
// the other N-2 shards from their owners.  Every link carries <= S.
// ---------------------------------------------------------------------------
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_broadcast(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank, root = op.root;
  KernelClock clk = kernel_begin(c);
  const int nsh = (op.algo == 1) ? N - 1 : 1;  // shards per chunk
  // shard index owned by rank p (p != root) in scatter mode
  auto shard_of = [&](int p) { int j = p - root - 1; if (j < 0) j += N; return j; };
  auto shard_units = [&](u64 cb) { return (units_of(cb) + nsh - 1) / nsh; };

  auto root_push = [&](u32 cidx) {
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    if (op.algo == 0) {
      const Region src = make_region(ub.recv + off, cb);
      const bool al = region_aligned(src);
      u64 lo, hi;
      block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
      for (u64 base = lo + threadIdx.x; base < hi; base += 2ull * kThreads) {
        uint4 v[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const u64 i = base + (u64)j * kThreads;
          if (i < hi) v[j] = region_ld(src, al, i);
        }
#pragma unroll
        for (int k = 1; k < kMaxRanks; ++k) {
          if (k < N) {
            int p = r + k;
            if (p >= N) p -= N;
            char* dst = c.stage[p] + half;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const u64 i = base + (u64)j * kThreads;
              if (i < hi) st16(dst + i * kUnit, v[j]);
            }
          }
        }
      }
    } else {
      const u64 su = shard_units(cb);
      u64 lo, hi;
      block_span(shard_units(op.chunk_bytes), su, lo, hi);
#pragma unroll 1
      for (u64 base = lo; base < hi; base += 4ull * kThreads) {
        const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
        for (int k = 1; k < N; ++k) {
          int p = r + k;
          if (p >= N) p -= N;
          const u64 j = (u64)shard_of(p);
          const u64 sb = clip_bytes(cb, j * su * kUnit, su * kUnit);
          const u64 ru = units_of(sb);
          copy_region_to_stage<4>(c.stage[p] + half, make_region(ub.recv + off + j * su * kUnit, sb),
                                  base < ru ? base : ru, bhi < ru ? bhi : ru);
        }
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
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    if (op.algo == 0) {
      u64 lo, hi;
      block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
      copy_stage_to_region<4>(make_region(ub.recv + off, cb), c.stage[r] + half, lo, hi);
    } else {
      const u64 su = shard_units(cb);
      u64 lo, hi;
      block_span(shard_units(op.chunk_bytes), su, lo, hi);
#pragma unroll 1
      for (u64 base = lo; base < hi; base += 4ull * kThreads) {
        const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
        for (int k = 0; k < N; ++k) {  // k == 0: own shard from own staging
          int p = r + k;
          if (p >= N) p -= N;
          if (p == root) continue;
          const u64 j = (u64)shard_of(p);
          const u64 sb = clip_bytes(cb, j * su * kUnit, su * kUnit);
          const u64 ru = units_of(sb);
          copy_stage_to_region<4>(make_region(ub.recv + off + j * su * kUnit, sb), c.stage[p] + half,
                                  base < ru ? base : ru, bhi < ru ? bhi : ru);
        }
      }
    }
  }
  kernel_end(c, op.nchunks, op.nchunks);
}


// ---------------------------------------------------------------------------
// Zero-copy kernels: user buffers of every rank are peer-addressable, so there is no
// staging and no chunk pipeline — barrier, one streaming pass, barrier.  Bytes moved are the
// algorithmic minimum; both NVLink directions are busy for the whole pass.
// ---------------------------------------------------------------------------

// Ranks must pass the same registered region/offset/size: CTA 0 publishes this rank's
// descriptor to every peer before the opening barrier and compares afterwards.
template <int NCTX>
__device__ __forceinline__ void zc_publish_desc(const RankCtx& c, const ZcPtrs& z, u64 tag) {
  if (NCTX != 1 || blockIdx.x != 0) return;
  const int t = threadIdx.x;
  if (t < c.world && t != c.rank) {
    volatile u64* d = c.flags[t] + (u64)(kDescRow + c.rank) * kMaxRanks;
#pragma unroll
    for (int j = 0; j < 6; ++j) d[j] = z.desc[j];
    d[6] = tag;
  }
}

template <int NCTX>
__device__ __forceinline__ void zc_check_desc(const RankCtx& c, const ZcPtrs& z, u64 tag) {
  if (NCTX != 1 || blockIdx.x != 0) return;
  const int t = threadIdx.x;
  if (t < c.world && t != c.rank) {
    const volatile u64* d = c.flags[c.rank] + (u64)(kDescRow + t) * kMaxRanks;
    bool same = (d[6] == tag);
#pragma unroll
    for (int j = 0; j < 6; ++j) same = same && (d[j] == z.desc[j]);
    if (!same) {
      volatile HostMailbox* mb = c.mbox;
      mb->error = 8;  // -NBD_ERR_MISMATCH
    }
  }
}

// all_reduce: rank r reduces shard r straight out of every rank's send buffer and writes the
// result straight into every rank's recv buffer.  Wire per GPU: (N-1)/N*S in, (N-1)/N*S out,
// concurrently.  In place is safe: only rank r reads or writes shard r of any buffer.
template <typename T, bool SUM, int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_allreduce_zc(const __grid_constant__ PackZc<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const ZcPtrs& z = pk.z;
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  zc_publish_desc<NCTX>(c, z, clk.epoch + 1);
  flag_barrier(c, ++clk.epoch);
  zc_check_desc<NCTX>(c, z, clk.epoch);

  const u64 units = units_of(op.bytes);
  const u64 su = (units + N - 1) / N;
  const u64 first = (u64)r * su;
  const u64 mine = first >= units ? 0 : (units - first < su ? units - first : su);
  u64 lo, hi;
  block_span(su, mine, lo, hi);
  bool al = true;
#pragma unroll
  for (int k = 0; k < kMaxRanks; ++k)
    if (k < N) al = al && ((reinterpret_cast<uintptr_t>(z.send[k]) | reinterpret_cast<uintptr_t>(z.recv[k])) & (kUnit - 1)) == 0;
  for (u64 i = first + lo + threadIdx.x; i < first + hi; i += kThreads) {
    uint4 v[kMaxRanks];
#pragma unroll
    for (int k = 0; k < kMaxRanks; ++k)
      if (k < N) v[k] = region_ld(make_region(z.send[k], op.bytes), al, i);
    Reducer<T> R;
    R.init(v[0]);
#pragma unroll
    for (int k = 1; k < kMaxRanks; ++k)
      if (k < N) R.template add<SUM>(v[k], op.op);
    const uint4 out = R.result(op.avg, N);
#pragma unroll
    for (int k = 0; k < kMaxRanks; ++k) {
      if (k < N) {
        int p = r + k;  // own buffer first, then peers in rotated order
        if (p >= N) p -= N;
        region_st(make_region(z.recv[p], op.bytes), al, i, out);
      }
    }
  }
  flag_barrier(c, ++clk.epoch);
  kernel_end(c, 2, 0);
}

// reduce_scatter: rank r reduces its shard out of every rank's send buffer into its own
// (local, possibly unregistered) recv.  op.bytes = shard bytes.
template <typename T, bool SUM, int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_reduce_scatter_zc(const __grid_constant__ PackZc<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const ZcPtrs& z = pk.z;
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  zc_publish_desc<NCTX>(c, z, clk.epoch + 1);
  flag_barrier(c, ++clk.epoch);
  zc_check_desc<NCTX>(c, z, clk.epoch);

  const u64 units = units_of(op.bytes);
  u64 lo, hi;
  block_span(units, units, lo, hi);
  const Region dst = make_region(z.recv[r], op.bytes);
  const bool ald = region_aligned(dst);
  for (u64 i = lo + threadIdx.x; i < hi; i += kThreads) {
    uint4 v[kMaxRanks];
#pragma unroll
    for (int k = 0; k < kMaxRanks; ++k) {
      if (k < N) {
        const Region src = make_region(z.send[k] + (u64)r * op.bytes, op.bytes);
        v[k] = region_ld(src, region_aligned(src), i);
      }
    }
    Reducer<T> R;
    R.init(v[0]);
#pragma unroll
    for (int k = 1; k < kMaxRanks; ++k)
      if (k < N) R.template add<SUM>(v[k], op.op);
    region_st(dst, ald, i, R.result(op.avg, N));
  }
  flag_barrier(c, ++clk.epoch);  // nobody may overwrite its send buffer before all have read it
  kernel_end(c, 2, 0);
}

// all_gather.  algo 0 ("pull", send buffers registered): read every peer's shard straight
// from its send buffer.  algo 1 ("push", recv buffers registered): write own shard straight
// into every peer's recv buffer.  op.bytes = shard bytes.
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_all_gather_zc(const __grid_constant__ PackZc<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const ZcPtrs& z = pk.z;
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  zc_publish_desc<NCTX>(c, z, clk.epoch + 1);
  flag_barrier(c, ++clk.epoch);
  zc_check_desc<NCTX>(c, z, clk.epoch);

  const u64 units = units_of(op.bytes);
  u64 lo, hi;
  block_span(units, units, lo, hi);
  const Region own_src = make_region(z.send[r], op.bytes);
#pragma unroll 1
  for (u64 base = lo; base < hi; base += 4ull * kThreads) {
    const u64 bhi = (base + 4ull * kThreads < hi) ? base + 4ull * kThreads : hi;
#pragma unroll 1
    for (int k = 0; k < N; ++k) {
      int p = r + k;
      if (p >= N) p -= N;
      if (op.algo == 0) {
        if (p == r && z.send[r] == z.recv[r] + (u64)r * op.bytes) continue;  // in place
        copy_region_to_region<4>(make_region(z.recv[r] + (u64)p * op.bytes, op.bytes),
                                 make_region(z.send[p], op.bytes), base, bhi);
      } else {
        if (p == r && z.send[r] == z.recv[r] + (u64)r * op.bytes) continue;
        copy_region_to_region<4>(make_region(z.recv[p] + (u64)r * op.bytes, op.bytes), own_src, base, bhi);
      }
    }
  }
  flag_barrier(c, ++clk.epoch);
  kernel_end(c, 2, 0);
}

// barrier: one CTA per rank, one flag round trip.
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_barrier(const __grid_constant__ Pack<NCTX> pk) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  KernelClock clk = kernel_begin(c);
  flag_barrier(c, ++clk.epoch);
  kernel_end(c, 1, 0);
}

}  // namespace nbd
