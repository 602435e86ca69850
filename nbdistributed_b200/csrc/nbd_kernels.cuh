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

// Auto-registration gate (see nbd_device.cuh): true if this launch is not the one the agreed
// decision selected.  Every CTA of every rank evaluates the same decision, so either the whole
// collective runs in this kernel or none of it does.
template <int NCTX>
__device__ __forceinline__ bool gated_out(const RankCtx& c, const OpParams& op) {
  if (NCTX != 1 || op.zc_mode == 0) return false;
  const volatile ZcDecision* d = &reinterpret_cast<LocalBlock*>(c.state)->dec;
  const bool ok = d->ok != 0;
  return op.zc_mode == 1 ? !ok : ok;
}

template <int NCTX>
__device__ __forceinline__ const ZcPtrs& zc_ptrs(const RankCtx& c, const PackZc<NCTX>& pk, const OpParams& op) {
  if (NCTX == 1 && op.zc_mode == 1) return reinterpret_cast<const LocalBlock*>(c.state)->dec.z;
  return pk.z;
}

// Ranks must run the same KIND of kernel for a collective.  A rank whose buffers are registered
// (symmetric memory) runs the zero-copy kernel and publishes a descriptor tagged with its epoch
// before the opening barrier (zc_publish_desc); a rank that passed ordinary buffers runs the
// staged kernel.  The zero-copy side notices (the staged peer published nothing: stale tag);
// this is the staged side's half: after its first barrier, a descriptor carrying OUR epoch's tag
// means that peer is in a zero-copy kernel -> NBD_ERR_MISMATCH instead of silent garbage.
template <int NCTX>
__device__ __forceinline__ void staged_check_peers(const RankCtx& c, u64 tag) {
  if (NCTX != 1 || blockIdx.x != 0) return;
  const int t = threadIdx.x;
  if (t < c.world && t != c.rank) {
    const volatile u64* d = c.flags[c.rank] + (u64)(kDescRow + t) * kMaxRanks;
    if (d[6] == tag) {
      volatile HostMailbox* mb = c.mbox;
      mb->error = 8;  // -NBD_ERR_MISMATCH
      reinterpret_cast<volatile LocalState*>(c.state)->failed = 1;
    }
  }
}

// N-way reduce of units [lo,hi) with U units per thread in flight: src(k, i) loads unit i of
// rank k's contribution, dst(i, v) stores the result.  Ascending k; U*N <= 8 loads outstanding
// per thread, so few ranks get a deeper unroll (remote loads need the bytes in flight).
template <typename T, bool SUM, int U, typename Src, typename Dst>
__device__ __forceinline__ void reduce_loop_u(int N, u64 lo, u64 hi, int op, int avg, Src src, Dst dst) {
  constexpr int KMAX = kMaxRanks / U;
  for (u64 base = lo + threadIdx.x; base < hi; base += (u64)U * kThreads) {
    uint4 v[U][KMAX];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) {
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < N) v[j][k] = src(k, i);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) {
        Reducer<T> R;
        R.init(v[j][0]);
#pragma unroll
        for (int k = 1; k < KMAX; ++k)
          if (k < N) R.template add<SUM>(v[j][k], op);
        dst(i, R.result(avg, N));
      }
    }
  }
}

// DEEP: unroll further when there are few ranks.  Pays off when the loop only reads remote
// memory (reduce_scatter: 489 -> 633 GB/s at N = 2) or only local memory; the zero-copy
// all_reduce, which reads AND writes remote memory in the same loop, measured 7-10 % slower
// with the deeper unroll (bursts of reads then bursts of writes), so it keeps U = 1.
template <typename T, bool SUM, bool DEEP, typename Src, typename Dst>
__device__ __forceinline__ void reduce_loop(int N, u64 lo, u64 hi, int op, int avg, Src src, Dst dst) {
  if (DEEP && N <= 2) reduce_loop_u<T, SUM, 4>(N, lo, hi, op, avg, src, dst);
  else if (DEEP && N <= 4) reduce_loop_u<T, SUM, 2>(N, lo, hi, op, avg, src, dst);
  else reduce_loop_u<T, SUM, 1>(N, lo, hi, op, avg, src, dst);
}

// N-way reduce of units [lo,hi): source k is this rank's staging slot k, except
// k == rank which is read straight from the user's send buffer.  Ascending k.
template <typename T, bool SUM, bool TO_STAGE>
__device__ __forceinline__ void reduce_units(const RankCtx& c, const Region& own_src, const char* slots,
                                             u64 slot_stride, const Region& dst, char* stage_out, u64 lo,
                                             u64 hi, int op, int avg) {
  const int N = c.world, r = c.rank;
  const bool als = region_aligned(own_src), ald = region_aligned(dst);
  reduce_loop<T, SUM, true>(
      N, lo, hi, op, avg,
      [&](int k, u64 i) { return (k == r) ? region_ld(own_src, als, i) : ld16(slots + (u64)k * slot_stride + i * kUnit); },
      [&](u64 i, const uint4& out) {
        region_st(dst, ald, i, out);
        if (TO_STAGE) st16(stage_out + i * kUnit, out);
      });
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
  if (gated_out<NCTX>(c, op)) return;
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
    if (cidx == 0) staged_check_peers<NCTX>(c, clk.epoch);
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
// all_reduce, LL one-shot (latency path, up to the communicator's LL limit): every 4-byte word travels with its
// own 4-byte flag in one 8-byte packet, so the receiver needs no separate flag, no fence and
// no barrier: it spins on the packets themselves.  A unit of 16 payload bytes is two 16-byte
// stores {w0,f,w1,f} {w2,f,w3,f}; each 8-byte half is written atomically.  flag = operation
// sequence number (never 0), buffers alternate by its parity: a rank can run at most one LL
// operation ahead of a peer (it needs the peer's packets of operation s+1, which the peer sends
// only after finishing s), so parity s is free again when s+2 writes it.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void st_ll(char* p, u32 a, u32 b, u32 flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(flag), "r"(b), "r"(flag));
}
__device__ __forceinline__ uint4 ld_ll(const char* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// Before writing packets of LL operation `seq`: every receiver has finished operation seq - 2.
// Only needed when operation seq - 1 was a broadcast (see kLLAckRow): if seq - 2 was one too its
// ack says so directly, otherwise the ack of seq - 1 implies it.
__device__ __forceinline__ void ll_wait_acks(const RankCtx& c, u64 seq, u64 hist) {
  if (!(hist & 1)) return;  // uniform across the CTA
  const int t = threadIdx.x;
  if (t < c.world && t != c.rank) {
    // no acquire fence: nothing is READ on the strength of the ack, we only write afterwards,
    // and a store is never performed ahead of the load it depends on
    const u64 need = (hist & 2) ? seq - 1 : seq;
    const u64* f = c.flags[c.rank] + (u64)kLLAckRow * kMaxRanks + t;
    if (ld_relaxed_sys(f) < need) wait_flag_slow(c.state, c.mbox, c.timeout_ns, f, need);
  }
  __syncthreads();
}

// End of an LL operation: the last CTA of this rank advances the LL clock; after a broadcast it
// also tells every peer that this rank has finished reading it.
__device__ __forceinline__ void ll_finish(const RankCtx& c, u64 seq, u64 hist, bool bcast) {
  __syncthreads();
  if (threadIdx.x == 0) {
    LocalState* s = c.state;
    const u32 prev = atomicAdd(&s->done, 1u);
    if (prev == gridDim.x - 1) {
      s->done = 0;
      s->ll_seq = seq + 1;
      s->ll_hist = ((hist << 1) & 2) | (bcast ? 1 : 0);
      __threadfence();
      if (bcast)
        for (int p = 0; p < c.world; ++p)
          if (p != c.rank) *reinterpret_cast<volatile u64*>(c.flags[p] + (u64)kLLAckRow * kMaxRanks + c.rank) = seq + 1;
    }
  }
}

template <typename T, bool SUM, int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_allreduce_ll(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank;
  const volatile LocalState* st = c.state;
  const u64 seq = st->ll_seq, hist = st->ll_hist;
  const u32 flag = (u32)(seq % 0xfffffffeull) + 1u;
  const u64 par = (seq & 1) * (u64)kMaxRanks * c.ll_slot;
  const Region src = make_region(ub.send, op.bytes), dst = make_region(ub.recv, op.bytes);
  const bool als = region_aligned(src), ald = region_aligned(dst);
  const u64 units = units_of(op.bytes);
  u64 lo, hi;
  block_span(units, units, lo, hi);
  char* my_ll = c.stage[r] + c.ll_off + par;
  ll_wait_acks(c, seq, hist);
  // Phase 1: every packet of this CTA's span goes out before anything is awaited, so a span of
  // several units per thread costs one flight time, not one per unit.
  const bool single = hi - lo <= (u64)kThreads;  // one unit per thread: keep it in registers
  uint4 first = make_uint4(0, 0, 0, 0);
  for (u64 i = lo + threadIdx.x; i < hi; i += kThreads) {
    const uint4 mine = region_ld(src, als, i);
    if (single) first = mine;
#pragma unroll
    for (int k = 1; k < kMaxRanks; ++k) {
      if (k < N) {
        int p = r + k;
        if (p >= N) p -= N;
        char* d = c.stage[p] + c.ll_off + par + (u64)r * c.ll_slot + i * 32;
        st_ll(d, mine.x, mine.y, flag);
        st_ll(d + 16, mine.z, mine.w, flag);
      }
    }
  }
  // Phase 2: spin on the peers' packets, reduce in ascending rank order.  (In place: unit i of the
  // send buffer is read again here before unit i of the result is written, by the same thread.)
  for (u64 i = lo + threadIdx.x; i < hi; i += kThreads) {
    uint4 v[kMaxRanks];
    const uint4 mine = single ? first : region_ld(src, als, i);
#pragma unroll
    for (int k = 0; k < kMaxRanks; ++k) {
      if (k < N) {
        if (k == r) {
          v[k] = mine;
        } else {
          const char* s = my_ll + (u64)k * c.ll_slot + i * 32;
          uint4 a = ld_ll(s), b = ld_ll(s + 16);
          u32 spins = 0;
          u64 t0 = 0;
          while (a.y != flag || a.w != flag || b.y != flag || b.w != flag) {
            a = ld_ll(s);
            b = ld_ll(s + 16);
            if ((++spins & 0xFFFFu) == 0) {  // watchdog, as in wait_flag_slow
              const u64 now = globaltimer_ns();
              if (t0 == 0) t0 = now;
              volatile HostMailbox* mb = c.mbox;
              volatile LocalState* ls = c.state;
              if (mb->abort) { mb->error = 5; ls->failed = 1; break; }
              if (c.timeout_ns != 0 && now - t0 > c.timeout_ns) { mb->error = 4; ls->failed = 1; break; }
              if (ls->failed != 0) break;
            }
          }
          v[k] = make_uint4(a.x, a.z, b.x, b.z);
        }
      }
    }
    Reducer<T> R;
    R.init(v[0]);
#pragma unroll
    for (int k = 1; k < kMaxRanks; ++k)
      if (k < N) R.template add<SUM>(v[k], op.op);
    region_st(dst, ald, i, R.result(op.avg, N));
  }
  ll_finish(c, seq, hist, false);
}

// broadcast, LL (up to the communicator's LL limit): the root writes {data, flag} packets into its slot of every
// peer's LL area, the peers spin on the packets and copy them out.  No barrier, no fence.
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_broadcast_ll(const __grid_constant__ Pack<NCTX> pk, const OpParams op) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const int N = c.world, r = c.rank, root = op.root;
  const volatile LocalState* st = c.state;
  const u64 seq = st->ll_seq, hist = st->ll_hist;
  const u32 flag = (u32)(seq % 0xfffffffeull) + 1u;
  const u64 par = (seq & 1) * (u64)kMaxRanks * c.ll_slot;
  const Region buf = make_region(ub.recv, op.bytes);
  const bool al = region_aligned(buf);
  const u64 units = units_of(op.bytes);
  u64 lo, hi;
  block_span(units, units, lo, hi);
  if (r == root) {
    ll_wait_acks(c, seq, hist);
    for (u64 i = lo + threadIdx.x; i < hi; i += kThreads) {
      const uint4 mine = region_ld(buf, al, i);
#pragma unroll
      for (int k = 1; k < kMaxRanks; ++k) {
        if (k < N) {
          int p = r + k;
          if (p >= N) p -= N;
          char* d = c.stage[p] + c.ll_off + par + (u64)r * c.ll_slot + i * 32;
          st_ll(d, mine.x, mine.y, flag);
          st_ll(d + 16, mine.z, mine.w, flag);
        }
      }
    }
  } else {
    const char* slot = c.stage[r] + c.ll_off + par + (u64)root * c.ll_slot;
    for (u64 i = lo + threadIdx.x; i < hi; i += kThreads) {
      const char* s = slot + i * 32;
      uint4 a = ld_ll(s), b = ld_ll(s + 16);
      u32 spins = 0;
      u64 t0 = 0;
      while (a.y != flag || a.w != flag || b.y != flag || b.w != flag) {
        a = ld_ll(s);
        b = ld_ll(s + 16);
        if ((++spins & 0xFFFFu) == 0) {  // watchdog, as in wait_flag_slow
          const u64 now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          volatile HostMailbox* mb = c.mbox;
          volatile LocalState* ls = c.state;
          if (mb->abort) { mb->error = 5; ls->failed = 1; break; }
          if (c.timeout_ns != 0 && now - t0 > c.timeout_ns) { mb->error = 4; ls->failed = 1; break; }
          if (ls->failed != 0) break;
        }
      }
      region_st(buf, al, i, make_uint4(a.x, a.z, b.x, b.z));
    }
  }
  ll_finish(c, seq, hist, true);
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
  if (gated_out<NCTX>(c, op)) return;
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
    if (cidx == 0) staged_check_peers<NCTX>(c, clk.epoch);
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
  if (gated_out<NCTX>(c, op)) return;
  KernelClock clk = kernel_begin(c);
  const u64 slot_stride = units_of(op.chunk_bytes) * kUnit;
  // this rank's contribution to rank p's result: a slice of the send buffer, or its own buffer
  const bool listed = pk.shard[ci][0] != nullptr;
  auto shard = [&](int p) { return listed ? pk.shard[ci][p] : ub.send + (u64)p * op.bytes; };

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
        copy_region_to_stage<4>(c.stage[p] + half + (u64)r * slot_stride, make_region(shard(p) + off, cb), base, bhi);
      }
    }
  };

  if (op.nchunks) push(0);
  for (u32 cidx = 0; cidx < op.nchunks; ++cidx) {
    flag_barrier(c, ++clk.epoch);
    if (cidx == 0) staged_check_peers<NCTX>(c, clk.epoch);
    if (cidx + 1 < op.nchunks) push(cidx + 1);
    const u64 off = (u64)cidx * op.chunk_bytes;
    const u64 cb = clip_bytes(op.bytes, off, op.chunk_bytes);
    const u64 half = ((clk.seq + cidx) & 1) * c.half_bytes;
    u64 lo, hi;
    block_span(units_of(op.chunk_bytes), units_of(cb), lo, hi);
    reduce_units<T, SUM, false>(c, make_region(shard(r) + off, cb), c.stage[r] + half,
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
  if (gated_out<NCTX>(c, op)) return;
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
    if (cidx == 0) staged_check_peers<NCTX>(c, clk.epoch);
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
      reinterpret_cast<volatile LocalState*>(c.state)->failed = 1;  // later flag waits of this launch drain
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
  if (gated_out<NCTX>(c, op)) return;
  const ZcPtrs& z = zc_ptrs<NCTX>(c, pk, op);
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  if (op.zc_mode == 0) zc_publish_desc<NCTX>(c, z, clk.epoch + 1);
  flag_barrier(c, ++clk.epoch);
  if (op.zc_mode == 0) zc_check_desc<NCTX>(c, z, clk.epoch);

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
  // [first+lo, fh): whole units of aligned buffers -> plain 128-bit accesses, no per-access
  // checks in the hot loop; [fh, first+hi): ragged last unit or misaligned buffers (checked)
  const u64 full = al ? op.bytes / kUnit : 0;
  const u64 a0 = first + lo, b0 = first + hi;
  const u64 fh = b0 < full ? b0 : (a0 > full ? a0 : full);
  auto push_all = [&](u64 i, const uint4& out) {
#pragma unroll
    for (int k = 0; k < kMaxRanks; ++k) {
      if (k < N) {
        int p = r + k;  // own buffer first, then peers in rotated order
        if (p >= N) p -= N;
        st16(z.recv[p] + i * kUnit, out);
      }
    }
  };
  if (op.algo == 0) {
    // fused: every unit is reduced and immediately written to all N buffers
    reduce_loop<T, SUM, false>(N, a0, fh, op.op, op.avg, [&](int k, u64 i) { return ld16(z.send[k] + i * kUnit); }, push_all);
  } else {
    // tiled: reduce a tile of the span into our own recv (remote reads only), then push that
    // tile to the peers (remote writes only).  Loops that only read or only write remote
    // memory sustain more NVLink bandwidth than one loop doing both (632 / 667 vs 557 GB/s
    // at N = 8); the tile is re-read from L2.  chunk_bytes = tile bytes (0: the whole span).
    const u64 tile = op.chunk_bytes ? (op.chunk_bytes / kUnit + kThreads - 1) / kThreads * kThreads : (fh - a0);
    char* mine_out = z.recv[r];
#pragma unroll 1
    for (u64 t0 = a0; t0 < fh; t0 += (tile ? tile : 1)) {
      const u64 t1 = (t0 + tile < fh) ? t0 + tile : fh;
      reduce_loop<T, SUM, true>(N, t0, t1, op.op, op.avg, [&](int k, u64 i) { return ld16(z.send[k] + i * kUnit); },
                                [&](u64 i, const uint4& out) { st16(mine_out + i * kUnit, out); });
      __syncthreads();  // the tile is complete in our recv before any thread forwards it
#pragma unroll 1
      for (u64 base = t0; base < t1; base += 4ull * kThreads) {
        const u64 bhi = (base + 4ull * kThreads < t1) ? base + 4ull * kThreads : t1;
#pragma unroll 1
        for (int k = 1; k < N; ++k) {
          int p = r + k;
          if (p >= N) p -= N;
          char* dp = z.recv[p];
          copy_units<4>(base, bhi, [&](u64 i) { return ld16(mine_out + i * kUnit); },
                        [&](u64 i, const uint4& v) { st16(dp + i * kUnit, v); });
        }
      }
    }
  }
  if (fh < b0) {
    reduce_loop_u<T, SUM, 1>(
        N, fh, b0, op.op, op.avg, [&](int k, u64 i) { return region_ld(make_region(z.send[k], op.bytes), al, i); },
        [&](u64 i, const uint4& out) {
          for (int k = 0; k < N; ++k) region_st(make_region(z.recv[k], op.bytes), al, i, out);
        });
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
  if (gated_out<NCTX>(c, op)) return;
  const ZcPtrs& z = zc_ptrs<NCTX>(c, pk, op);
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  if (op.zc_mode == 0) zc_publish_desc<NCTX>(c, z, clk.epoch + 1);
  flag_barrier(c, ++clk.epoch);
  if (op.zc_mode == 0) zc_check_desc<NCTX>(c, z, clk.epoch);

  const u64 units = units_of(op.bytes);
  u64 lo, hi;
  block_span(units, units, lo, hi);
  const Region dst = make_region(z.recv[r], op.bytes);
  const bool ald = region_aligned(dst);
  bool als = true;
#pragma unroll
  for (int k = 0; k < kMaxRanks; ++k)
    if (k < N) als = als && region_aligned(make_region(z.send[k] + (u64)r * op.bytes, op.bytes));
  const u64 full = (als && ald) ? op.bytes / kUnit : 0;
  const u64 fh = hi < full ? hi : (lo > full ? lo : full);
  const u64 soff = (u64)r * op.bytes;
  char* dp = dst.p;
  reduce_loop<T, SUM, true>(N, lo, fh, op.op, op.avg, [&](int k, u64 i) { return ld16(z.send[k] + soff + i * kUnit); },
                            [&](u64 i, const uint4& out) { st16(dp + i * kUnit, out); });
  if (fh < hi)
    reduce_loop_u<T, SUM, 1>(
        N, fh, hi, op.op, op.avg,
        [&](int k, u64 i) { return region_ld(make_region(z.send[k] + soff, op.bytes), als, i); },
        [&](u64 i, const uint4& out) { region_st(dst, ald, i, out); });
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
  if (gated_out<NCTX>(c, op)) return;
  const ZcPtrs& z = zc_ptrs<NCTX>(c, pk, op);
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  if (op.zc_mode == 0) zc_publish_desc<NCTX>(c, z, clk.epoch + 1);
  flag_barrier(c, ++clk.epoch);
  if (op.zc_mode == 0) zc_check_desc<NCTX>(c, z, clk.epoch);

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

// ---------------------------------------------------------------------------
// k_zc_agree: one CTA per rank.  Round 1 exchanges {export slot, offset} descriptors of this
// operation's buffers; each rank resolves its peers' slots through its import table.  Round 2
// exchanges "I can address everybody" bits; the AND is the decision, identical on all ranks.
// Slots a rank could not resolve are reported to its host, which imports them (CUDA IPC) before
// a later operation.  desc: {send slot, send offset, recv slot, recv offset, bytes, op tag};
// need: bit 0 = peers' send buffers are needed, bit 1 = peers' recv buffers are needed.
// ---------------------------------------------------------------------------
struct AgreeArgs {
  u64 desc[6];
  const char* local_send;
  char* local_recv;
  int need;
};

static __global__ void __launch_bounds__(64, 1) k_zc_agree(const __grid_constant__ Pack<1> pk, const AgreeArgs a) {
  const RankCtx& c = pk.ctx[0];
  const int N = c.world, r = c.rank, t = threadIdx.x;
  LocalBlock* blk = reinterpret_cast<LocalBlock*>(c.state);
  KernelClock clk = kernel_begin(c);
  const u64 tag1 = clk.epoch + 1, tag2 = clk.epoch + 2;

  if (t < N && t != r) {
    volatile u64* d = c.flags[t] + (u64)(kDescRow + r) * kMaxRanks;
#pragma unroll
    for (int j = 0; j < 6; ++j) d[j] = a.desc[j];
    d[6] = tag1;
  }
  flag_barrier(c, ++clk.epoch);

  __shared__ u64 sm_miss[2 * kMaxRanks];
  __shared__ int sm_nmiss;
  if (t == 0) sm_nmiss = 0;
  __syncthreads();
  int ok = 1, same = 1;
  if (t < N) {
    // steady state (a training loop re-reducing the same buffers): every rank sees the same N
    // descriptors as last time and last time's answer was yes -> the answer is still yes, and
    // every rank knows it without a second exchange
    const volatile u64* dsc = (t == r) ? nullptr : c.flags[r] + (u64)(kDescRow + t) * kMaxRanks;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const u64 v = (t == r) ? a.desc[j] : dsc[j];
      same = same && (blk->prev_desc[t][j] == v);
      blk->prev_desc[t][j] = v;
    }
    const char* ps = nullptr;
    char* pr = nullptr;
    if (t == r) {
      ps = a.local_send;
      pr = a.local_recv;
      // our own buffers must be exported for the peers to reach them
      if ((a.need & 1) && a.desc[0] == kNoSlot) ok = 0;
      if ((a.need & 2) && a.desc[2] == kNoSlot) ok = 0;
    } else {
      const volatile u64* d = c.flags[r] + (u64)(kDescRow + t) * kMaxRanks;
      const u64 es = d[0], os = d[1], er = d[2], orr = d[3];
      if (d[6] != tag1 || d[4] != a.desc[4] || d[5] != a.desc[5]) ok = 0;
      if (ok && (a.need & 1)) {
        const u64 m = es < (u64)kMaxExports ? blk->imp[t][es] : 0;
        if (m) ps = reinterpret_cast<const char*>(m) + os;
        else {
          ok = 0;
          if (es < (u64)kMaxExports) sm_miss[atomicAdd(&sm_nmiss, 1)] = ((u64)t << 32) | es;
        }
      }
      if ((a.need & 2) && (ok || !(a.need & 1))) {
        const u64 m = er < (u64)kMaxExports ? blk->imp[t][er] : 0;
        if (m) pr = reinterpret_cast<char*>(m) + orr;
        else {
          ok = 0;
          if (er < (u64)kMaxExports && er != es) sm_miss[atomicAdd(&sm_nmiss, 1)] = ((u64)t << 32) | er;
        }
      }
    }
    blk->dec.z.send[t] = ps;
    blk->dec.z.recv[t] = pr;
  }
  const int ok_local = __syncthreads_and(ok);
  const int unchanged = __syncthreads_and(same) && ok_local && (blk->dec.ok == 1u);
  if (unchanged) {  // uniform across ranks: all hold the same descriptors and the same history
    kernel_end(c, 1, 0);
    return;
  }
  if (t == 0 && sm_nmiss > 0) {  // report unresolved slots to the host (plain stores: pinned host memory)
    volatile HostMailbox* mb = c.mbox;
    u32 n = mb->miss_count;
    for (int j = 0; j < sm_nmiss; ++j) mb->miss[(n + j) % kMissRing] = sm_miss[j];
    __threadfence_system();
    mb->miss_count = n + (u32)sm_nmiss;
  }

  if (t < N && t != r) {
    volatile u64* d = c.flags[t] + (u64)(kDescRow + r) * kMaxRanks;
    d[7] = (tag2 << 1) | (u64)(ok_local ? 1 : 0);
  }
  flag_barrier(c, ++clk.epoch);
  int okp = 1;
  if (t < N && t != r) {
    const volatile u64* d = c.flags[r] + (u64)(kDescRow + t) * kMaxRanks;
    okp = (d[7] == ((tag2 << 1) | 1ull));
  }
  const int ok_all = __syncthreads_and(okp && ok_local);
  if (t == 0) {
    blk->dec.ok = ok_all ? 1u : 0u;
    __threadfence();
  }
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
