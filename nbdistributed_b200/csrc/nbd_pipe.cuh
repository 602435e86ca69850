// nbd_pipe.cuh — tile-pipeline collectives: warp-specialised CTAs that stream a message
// through a staging buffer in small tiles, with per-tile flags instead of per-chunk barriers.
//
// Why: the staged kernels in nbd_kernels.cuh run whole 64 MiB chunks between flag barriers that
// every warp of the CTA takes part in, so copy-in, reduction and copy-out of neighbouring
// chunks cannot overlap inside a CTA and every barrier drains the CTA's in-flight NVLink
// writes (~25 us per step at 8 ranks, profiles/r01_tune_bcast_n8.jsonl).  Here a CTA is three
// independent actors that meet only through flags:
//
//   warp 0       IN   user send -> own staging slot, as a TMA ring (cp.async.bulk global ->
//                     shared -> global driven by one lane), then signals "A(i)" to every rank
//   warps 2..15  RED  waits A(i) of every rank, reduces ITS sub-shard of tile i across ranks
//                     and writes the result into every rank's slot, signals "B(i)"
//   warp 1       OUT  waits B(i) of every rank, staging slot -> user recv (TMA ring)
//
// Tile i of CTA b covers message units [(i*G + b)*Tu, +Tu) (G = grid size, Tu = tile units) and
// lives in slot i % depth of CTA b's part of the staging.  The transport of RED is a
// template switch:
//   MC = true   NVSwitch multicast: multimem.ld_reduce (the switch fetches and sums the N
//               copies) + multimem.st (the switch writes all N copies).  Wire bytes per GPU
//               ~ S each way instead of 2(N-1)/N*S.  The switch's summation order is not the
//               ascending-rank order of the other kernels (float results agree within the
//               tolerance stated in tests/helpers.py; all ranks still receive identical bits).
//   MC = false  peer-to-peer loads/stores on the IPC-mapped arenas, ascending rank order,
//               bit-identical with the other kernels; also what emulated ranks run.
//
// Flags are the arena's monotonic u64 clocks: A(i) and B(i) both carry the
// value epoch+1+i, A in flag row b, B in flag row kPipeRowB + b, word = source rank.  A slot is
// reused for tile i+depth only after OUT has finished tile i, which (through B(i)) implies
// that every rank has finished reading and writing that slot on every rank.
//
// No counterpart in the reference; semantics are those of dist.all_reduce / dist.broadcast as
// called from ref:README.md:109,122-123 (see include/nbdist_b200.h).
#pragma once

#include "nbd_kernels.cuh"

namespace nbd {

constexpr int kPipeDepth = 4;                       // default staging slots per CTA (OpParams::depth)
constexpr int kPipeRowB = 256;                      // B flags live in rows [256, 256 + grid)
constexpr int kPipeMaxBlocks = 256;                 // rows 512.. hold export slots
constexpr int kPipeRedThreads = kThreads - 64;      // warps 2..15
constexpr u32 kTmaPiece = 8192;                     // bytes per bulk copy
constexpr int kTmaStages = 12;                      // bulk copies in flight per ring (96 KiB)
constexpr size_t kPipeSmemBytes = 2 * (size_t)kTmaStages * kTmaPiece;  // IN ring + OUT ring

struct PipeBufs {
  char* local;  // this rank's staging (MC: unicast mapping of the multicast-bound buffer; else arena staging)
  char* mc;     // multicast address of the same staging (MC only)
};

template <int NCTX>
struct PackPipe {
  RankCtx ctx[NCTX];
  RankBufs buf[NCTX];
  PipeBufs pb[NCTX];
};

// ---- bulk async copies (TMA), driven by one lane ---------------------------------------------
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, u32 parity) {
  u32 ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, u32 bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, u32 bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// orders this thread's earlier generic-proxy accesses (the acquire of a flag, plain stores)
// with later async-proxy accesses (bulk copies) and vice versa
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

struct TmaRing {
  unsigned char* buf;  // kTmaStages * kTmaPiece bytes of shared memory, 128-byte aligned
  uint64_t* bar;       // kTmaStages mbarriers
  u32 phase;           // bit s = parity the next wait on stage s expects
};

// Copies nbytes (a multiple of 16; both addresses 16-byte aligned) global -> shared -> global.
// One thread calls it.  On return every store has been issued and has finished READING shared
// memory (the ring is reusable); completion of the global writes needs bulk_wait_all0().
__device__ __forceinline__ void tma_copy(TmaRing& R, char* dst, const char* src, u64 nbytes) {
  const u64 np = (nbytes + kTmaPiece - 1) / kTmaPiece;
  auto piece = [&](u64 j) { return (u32)((j + 1) * kTmaPiece <= nbytes ? kTmaPiece : nbytes - j * kTmaPiece); };
  for (u64 j = 0; j < np && j < (u64)kTmaStages; ++j) {
    mbar_expect_tx(&R.bar[j], piece(j));
    bulk_load(R.buf + j * kTmaPiece, src + j * kTmaPiece, piece(j), &R.bar[j]);
  }
  for (u64 j = 0; j < np; ++j) {
    const int s = (int)(j % kTmaStages);
    while (!mbar_try_wait(&R.bar[s], (R.phase >> s) & 1u)) {
    }
    R.phase ^= 1u << s;
    bulk_store(dst + j * kTmaPiece, R.buf + (size_t)s * kTmaPiece, piece(j));
    bulk_commit();
    const u64 nxt = j + kTmaStages;
    if (nxt < np) {
      bulk_wait_read0();  // the store has drained stage s
      mbar_expect_tx(&R.bar[s], piece(nxt));
      bulk_load(R.buf + (size_t)s * kTmaPiece, src + nxt * kTmaPiece, piece(nxt), &R.bar[s]);
    }
  }
  bulk_wait_read0();
}

// ---- multimem (NVSwitch multicast) accessors for 16-byte units --------------------------------
template <typename T>
struct McSum;  // ld: sum of the N copies of one unit; only instantiated where the switch can add T

template <>
struct McSum<float> {
  static __device__ __forceinline__ uint4 ld(const char* p) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
  }
};
template <>
struct McSum<__nv_bfloat16> {
  static __device__ __forceinline__ uint4 ld(const char* p) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
  }
};
template <>
struct McSum<__half> {
  static __device__ __forceinline__ uint4 ld(const char* p) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
  }
};
// types the switch is not asked to add (the host never launches MC = true for them)
template <typename T>
struct McSum {
  static __device__ __forceinline__ uint4 ld(const char*) { return make_uint4(0, 0, 0, 0); }
};

__device__ __forceinline__ void mc_st(char* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// divide the lanes of one unit by `world` (AVG); one extra rounding for bf16/f16 on this path
template <typename T>
__device__ __forceinline__ uint4 unit_div(const uint4& v, int world) {
  Reducer<T> R;
  R.init(v);
  return R.result(1, world);
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---- geometry ---------------------------------------------------------------------------------
struct PipeGeom {
  u64 units;   // 16-byte units of the message
  u64 tu;      // units per tile
  u32 rounds;  // tiles per CTA
  u32 depth;   // staging slots per CTA
};

__device__ __forceinline__ u64 tile_first(const PipeGeom& g, u32 i) { return ((u64)i * gridDim.x + blockIdx.x) * g.tu; }
__device__ __forceinline__ u64 tile_units(const PipeGeom& g, u32 i) {
  const u64 f = tile_first(g, i);
  return f >= g.units ? 0 : (g.units - f < g.tu ? g.units - f : g.tu);
}
__device__ __forceinline__ u64 slot_offset(const PipeGeom& g, u32 i) {
  return ((u64)blockIdx.x * g.depth + (i % g.depth)) * g.tu * kUnit;
}

// Staging hand-over with the neighbouring kernels of the stream.  The chunked kernels of
// nbd_kernels.cuh alternate between two staging halves and never end with a barrier: a peer
// may still be pulling from our arena when our next kernel starts.  A pipeline kernel whose
// slots live in the arena (op.algo bit 0) therefore opens with a flag barrier; every pipeline
// kernel closes with one, so that whatever runs next may write any rank's staging at once.
// Returns the base epoch: A(i) and B(i) carry base + 1 + i.
__device__ __forceinline__ u64 pipe_open(const RankCtx& c, KernelClock& clk, const OpParams& op) {
  if (op.algo & 1) flag_barrier(c, ++clk.epoch);
  else __syncthreads();
  return clk.epoch;
}

__device__ __forceinline__ void pipe_close(const RankCtx& c, u64 base, u32 rounds, const OpParams& op) {
  flag_barrier(c, base + rounds + 1);
  kernel_end(c, (u64)rounds + 1 + (op.algo & 1), 0);
}

// Tile of a user region <-> staging, by ONE WARP.  Whole units of a 16-byte aligned region go
// through the TMA ring (lane 0); a misaligned region or the ragged last unit go through the
// byte-wise accessors, lanes striding (slow, correct, rare).  `to_stage`: direction.
__device__ __forceinline__ void warp_tile_copy(TmaRing& R, const Region& user, u64 first, u64 tu, char* slot, bool to_stage,
                                               int lane) {
  const bool al = region_aligned(user);
  const u64 whole = user.n / kUnit;  // units of the region that are complete
  u64 fast = 0;
  if (al && whole > first) fast = whole - first < tu ? whole - first : tu;
  if (fast && lane == 0) {
    if (to_stage) tma_copy(R, slot, user.p + first * kUnit, fast * kUnit);
    else tma_copy(R, user.p + first * kUnit, slot, fast * kUnit);
  }
  for (u64 u = fast + lane; u < tu; u += 32) {
    if (to_stage) st16(slot + u * kUnit, region_ld_slow(user.p, user.n, first + u));
    else region_st_slow(user.p, user.n, first + u, ld16(slot + u * kUnit));
  }
}

// reduce_loop_u of nbd_kernels.cuh for the reducing warps only (thread rt of kPipeRedThreads)
template <typename T, bool SUM, int U, typename Src, typename Dst>
__device__ __forceinline__ void pipe_reduce_loop(int N, u64 lo, u64 hi, int rt, int nthr, int op, int avg, Src src, Dst dst) {
  constexpr int KMAX = kMaxRanks / U;
  for (u64 base = lo + rt; base < hi; base += (u64)U * nthr) {
    uint4 v[U][KMAX];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * nthr;
      if (i < hi) {
#pragma unroll
        for (int k = 0; k < KMAX; ++k)
          if (k < N) v[j][k] = src(k, i);
      }
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * nthr;
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

// In-switch reduction of units [lo, hi) of a slot: U loads in flight per thread, then U stores.
template <typename T, int U>
__device__ __forceinline__ void mc_reduce_loop(char* base, u64 lo, u64 hi, int gt, int nthr, int avg, int world) {
  for (u64 b0 = lo + gt; b0 < hi; b0 += (u64)U * nthr) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 u = b0 + (u64)j * nthr;
      if (u < hi) v[j] = McSum<T>::ld(base + u * kUnit);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 u = b0 + (u64)j * nthr;
      if (u < hi) mc_st(base + u * kUnit, avg ? unit_div<T>(v[j], world) : v[j]);
    }
  }
}

// ---------------------------------------------------------------------------
// all_reduce through the tile pipeline.  op.chunk_bytes = tile bytes (multiple of 16),
// op.nchunks = rounds.  Grid <= kPipeMaxBlocks; every CTA of every rank must be co-resident.
// ---------------------------------------------------------------------------
template <typename T, bool SUM, int NCTX, bool MC>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_allreduce_pipe(const __grid_constant__ PackPipe<NCTX> pk, const OpParams op) {
  extern __shared__ __align__(128) unsigned char pipe_smem[];
  __shared__ uint64_t bars[2 * kTmaStages];
  __shared__ volatile u32 out_done;  // tiles this CTA's OUT warp has finished
  __shared__ volatile u32 b_next;    // next tile whose B flag may be written (keeps the two RED groups in order)

  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const PipeBufs& pb = pk.pb[ci];
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  PipeGeom g;
  g.units = units_of(op.bytes);
  g.tu = op.chunk_bytes / kUnit;
  g.rounds = op.nchunks;
  g.depth = (u32)op.depth;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const u64 rowA = (u64)blockIdx.x * kMaxRanks, rowB = (u64)(kPipeRowB + blockIdx.x) * kMaxRanks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2 * kTmaStages; ++s) mbar_init(&bars[s], 1);
    out_done = 0;
    b_next = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const u64 E = pipe_open(c, clk, op);

  if (warp == 0) {
    // ---------------- IN: user send -> staging slot, signal A(i) ----------------
    TmaRing R = {pipe_smem, bars, 0};
    const Region src = make_region(ub.send, op.bytes);
    for (u32 i = 0; i < g.rounds; ++i) {
      if (i >= g.depth) {
        if (lane == 0) {
          while (out_done < i - g.depth + 1) {
          }
        }
        __syncwarp();
      }
      const u64 tu = tile_units(g, i);
      if (tu && !(op.algo & 8)) warp_tile_copy(R, src, tile_first(g, i), tu, pb.local + slot_offset(g, i), true, lane);
      if (lane == 0) {
        bulk_wait_all0();     // the bulk stores into the slot are complete ...
        fence_proxy_async();  // ... and ordered before the generic-proxy release below
      }
      __syncwarp();
      if (lane < N) st_release_sys(c.flags[lane] + rowA + r, E + 1 + i);
    }
  } else if (warp == 1) {
    // ---------------- OUT: wait B(i), staging slot -> user recv ----------------
    TmaRing R = {pipe_smem + (size_t)kTmaStages * kTmaPiece, bars + kTmaStages, 0};
    const Region dst = make_region(ub.recv, op.bytes);
    for (u32 i = 0; i < g.rounds; ++i) {
      if (lane < N) wait_flag(c, c.flags[r] + rowB + lane, E + 1 + i);
      __syncwarp();
      if (lane == 0) fence_proxy_async();  // acquired data is read by the async proxy next
      __syncwarp();
      const u64 tu = tile_units(g, i);
      if (tu && !(op.algo & 16)) warp_tile_copy(R, dst, tile_first(g, i), tu, pb.local + slot_offset(g, i), false, lane);
      if (lane == 0) bulk_wait_all0();  // slot fully read, user memory written
      __syncwarp();
      if (lane == 0) {
        __threadfence_block();
        out_done = i + 1;
      }
    }
  } else {
    // ---------------- RED: wait A(i), reduce own sub-shard, write it everywhere, signal B(i) ----
    // op.algo bit 1: the 14 warps work as TWO groups on alternate tiles, so that one group's
    // drain (the release before B(i) waits for its multicast stores) overlaps the other's loads.
    const int ngroups = (op.algo & 2) ? 2 : 1;
    const int nthr = kPipeRedThreads / ngroups;
    const int rt = threadIdx.x - 64;
    const int grp = rt / nthr, gt = rt - grp * nthr;
    for (u32 i = (u32)grp; i < g.rounds; i += (u32)ngroups) {
      if (gt < N) wait_flag(c, c.flags[r] + rowA + gt, E + 1 + i);
      named_bar_sync(1 + grp, nthr);
      const u64 tu = tile_units(g, i);
      const u64 su = (tu + N - 1) / N;
      u64 lo = (u64)r * su;
      if (lo > tu) lo = tu;
      u64 hi = lo + su;
      if (hi > tu) hi = tu;
      const u64 soff = slot_offset(g, i);
      if (op.algo & 32) {
        // timing diagnostics only (pipe_mode bits 3-5 skip copy-in / copy-out / reduction)
      } else if (MC) {
        if (op.algo & 4) mc_reduce_loop<T, 8>(pb.mc + soff, lo, hi, gt, nthr, op.avg, N);
        else mc_reduce_loop<T, 4>(pb.mc + soff, lo, hi, gt, nthr, op.avg, N);
      } else {
        // peer-to-peer transport: ascending-rank reduction of the N staged copies, result
        // written into every rank's slot (own first, then rotated); few ranks -> deeper unroll
        // (remote loads need the bytes in flight)
        auto src = [&](int k, u64 u) { return ld16((k == r ? pb.local : c.stage[k]) + soff + u * kUnit); };
        auto dst = [&](u64 u, const uint4& out) {
#pragma unroll
          for (int k = 0; k < kMaxRanks; ++k) {
            if (k < N) {
              int p = r + k;
              if (p >= N) p -= N;
              st16((p == r ? pb.local : c.stage[p]) + soff + u * kUnit, out);
            }
          }
        };
        if (N <= 2) pipe_reduce_loop<T, SUM, 4>(N, lo, hi, gt, nthr, op.op, op.avg, src, dst);
        else if (N <= 4) pipe_reduce_loop<T, SUM, 2>(N, lo, hi, gt, nthr, op.op, op.avg, src, dst);
        else pipe_reduce_loop<T, SUM, 1>(N, lo, hi, gt, nthr, op.op, op.avg, src, dst);
      }
      named_bar_sync(1 + grp, nthr);
      if (gt < N) {
        // flags only grow and B(i) must imply B(i-1): the two groups signal in tile order
        while (b_next != i) {
        }
        st_release_sys(c.flags[gt] + rowB + r, E + 1 + i);
        __syncwarp((1u << N) - 1u);
        if (gt == 0) {
          __threadfence_block();
          b_next = i + 1;
        }
      }
    }
  }
  pipe_close(c, E, g.rounds, op);
}

// ---------------------------------------------------------------------------
// broadcast through the tile pipeline.  The root's RED warps read its user buffer and write
// each tile into EVERY rank's slot (multimem.st: once over the wire, the switch replicates;
// without multicast: one store per peer), then signal A(i); the other ranks' OUT warp copies the
// slot out and signals B(i) back to the root, which gates the reuse of a slot.
// ---------------------------------------------------------------------------
template <int NCTX, bool MC>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_broadcast_pipe(const __grid_constant__ PackPipe<NCTX> pk, const OpParams op) {
  extern __shared__ __align__(128) unsigned char pipe_smem[];
  __shared__ uint64_t bars[kTmaStages];

  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const PipeBufs& pb = pk.pb[ci];
  const int N = c.world, r = c.rank, root = op.root;
  KernelClock clk = kernel_begin(c);
  PipeGeom g;
  g.units = units_of(op.bytes);
  g.tu = op.chunk_bytes / kUnit;
  g.rounds = op.nchunks;
  g.depth = (u32)op.depth;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const u64 rowA = (u64)blockIdx.x * kMaxRanks, rowB = (u64)(kPipeRowB + blockIdx.x) * kMaxRanks;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTmaStages; ++s) mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const u64 E = pipe_open(c, clk, op);

  if (r == root) {
    if (warp >= 2) {
      const int rt = threadIdx.x - 64;
      const Region src = make_region(ub.recv, op.bytes);
      const bool al = region_aligned(src);
      for (u32 i = 0; i < g.rounds; ++i) {
        if (i >= g.depth) {  // every reader has copied tile i - depth out of this slot
          if (rt < N && rt != r) wait_flag(c, c.flags[r] + rowB + rt, E + 1 + (i - g.depth));
          named_bar_sync(1, kPipeRedThreads);
        }
        const u64 tu = tile_units(g, i), first = tile_first(g, i), soff = slot_offset(g, i);
        for (u64 b0 = rt; b0 < tu; b0 += 4ull * kPipeRedThreads) {
          uint4 v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const u64 u = b0 + (u64)j * kPipeRedThreads;
            if (u < tu) v[j] = region_ld(src, al, first + u);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const u64 u = b0 + (u64)j * kPipeRedThreads;
            if (u < tu) {
              if (MC) {
                mc_st(pb.mc + soff + u * kUnit, v[j]);
              } else {
#pragma unroll
                for (int k = 1; k < kMaxRanks; ++k) {
                  if (k < N) {
                    int p = r + k;
                    if (p >= N) p -= N;
                    st16(c.stage[p] + soff + u * kUnit, v[j]);
                  }
                }
              }
            }
          }
        }
        named_bar_sync(1, kPipeRedThreads);
        if (rt < N && rt != r) st_release_sys(c.flags[rt] + rowA + r, E + 1 + i);
      }
    }
  } else if (warp == 1) {
    TmaRing R = {pipe_smem, bars, 0};
    const Region dst = make_region(ub.recv, op.bytes);
    for (u32 i = 0; i < g.rounds; ++i) {
      if (lane == 0) {
        wait_flag(c, c.flags[r] + rowA + root, E + 1 + i);
        fence_proxy_async();
      }
      __syncwarp();
      const u64 tu = tile_units(g, i);
      if (tu) warp_tile_copy(R, dst, tile_first(g, i), tu, pb.local + slot_offset(g, i), false, lane);
      if (lane == 0) bulk_wait_all0();
      __syncwarp();
      if (lane == 0) st_release_sys(c.flags[root] + rowB + r, E + 1 + i);
    }
  }
  pipe_close(c, E, g.rounds, op);
}

// ---------------------------------------------------------------------------
// broadcast as a relay through the tile pipeline (peer-to-peer transport, any world size).
// k_broadcast's "scatter" algorithm with per-tile flags instead of whole-grid barriers: tile i of
// CTA b is cut into N-1 sub-shards, one per non-root rank.
//   root      PUSH warps (2..15) read the tile from its user buffer and write sub-shard j into
//             ITS OWNER's slot only (root egress = S), then signal A(i) to every other rank
//   non-root  PUSH warps wait A(i), read the own sub-shard from the local slot and write it into
//             the slots of the other N-2 non-root ranks, signal B(i) to them
//             OUT warp (1) waits A(i) and the N-2 B(i), copies the whole tile slot -> user (TMA
//             ring), then signals D(i) = "my slot of tile i is free" to every other rank
// A writer (root or relaying peer) reuses a slot for tile i+depth after D(i) of its owner.
// Every transfer is a store and every wait polls local memory.  Flag words: A and D share flag
// row b (word = source rank: the root's word carries A, a non-root's word D), B lives in row
// kPipeRowB + b; all carry E+1+i, below the closing barrier's E+rounds+1.
// op.algo bit 1: the 14 PUSH warps work as two groups on alternate tiles (the release before a
// signal drains the group's remote stores; the other group keeps the links busy meanwhile).
// ---------------------------------------------------------------------------
template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_broadcast_relay(const __grid_constant__ PackPipe<NCTX> pk, const OpParams op) {
  extern __shared__ __align__(128) unsigned char pipe_smem[];
  __shared__ uint64_t bars[kTmaStages];
  __shared__ volatile u32 sig_next;  // tiles [0, sig_next) have been pushed and signalled, in order

  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const RankBufs& ub = pk.buf[ci];
  const PipeBufs& pb = pk.pb[ci];
  const int N = c.world, r = c.rank, root = op.root;
  KernelClock clk = kernel_begin(c);
  PipeGeom g;
  g.units = units_of(op.bytes);
  g.tu = op.chunk_bytes / kUnit;
  g.rounds = op.nchunks;
  g.depth = (u32)op.depth;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const u64 rowA = (u64)blockIdx.x * kMaxRanks, rowB = (u64)(kPipeRowB + blockIdx.x) * kMaxRanks;
  const int nsh = N - 1;
  auto owner_of = [&](int j) { int p = root + 1 + j; if (p >= N) p -= N; return p; };

  if (threadIdx.x == 0) {
    for (int s = 0; s < kTmaStages; ++s) mbar_init(&bars[s], 1);
    sig_next = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const u64 E = pipe_open(c, clk, op);

  if (warp >= 2) {
    const int ngroups = (op.algo & 2) ? 2 : 1;
    const int nthr = kPipeRedThreads / ngroups;
    const int rt = threadIdx.x - 64;
    const int grp = rt / nthr, gt = rt - grp * nthr;
    const bool peer_lane = gt < N && gt != r && gt != root;  // lanes that talk to the other non-root ranks
    const Region src = make_region(ub.recv, op.bytes);
    const bool al = region_aligned(src);
    for (u32 i = (u32)grp; i < g.rounds; i += (u32)ngroups) {
      const u64 tu = tile_units(g, i), first = tile_first(g, i), soff = slot_offset(g, i);
      const u64 su = (tu + nsh - 1) / nsh;
      if (r == root) {
        // slot free on every owner: D(i - depth)
        if (i >= g.depth && gt < N && gt != r) wait_flag(c, c.flags[r] + rowA + gt, E + 1 + (i - g.depth));
        named_bar_sync(1 + grp, nthr);
        for (int jj = 0; jj < nsh; ++jj) {
          int j = jj + (int)((blockIdx.x + i) % (u32)nsh);  // CTAs start on different owners
          if (j >= nsh) j -= nsh;
          u64 lo = (u64)j * su;
          if (lo > tu) lo = tu;
          u64 hi = lo + su;
          if (hi > tu) hi = tu;
          char* dst = c.stage[owner_of(j)] + soff;
          for (u64 b0 = lo + gt; b0 < hi; b0 += 4ull * nthr) {
            uint4 v[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const u64 u = b0 + (u64)q * nthr;
              if (u < hi) v[q] = region_ld(src, al, first + u);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const u64 u = b0 + (u64)q * nthr;
              if (u < hi) st16(dst + u * kUnit, v[q]);
            }
          }
        }
      } else {
        if (gt == 0) wait_flag(c, c.flags[r] + rowA + root, E + 1 + i);                             // own sub-shard is here
        if (i >= g.depth && peer_lane) wait_flag(c, c.flags[r] + rowA + gt, E + 1 + (i - g.depth));  // peers' slots are free
        named_bar_sync(1 + grp, nthr);
        int j = r - root - 1;
        if (j < 0) j += N;
        u64 lo = (u64)j * su;
        if (lo > tu) lo = tu;
        u64 hi = lo + su;
        if (hi > tu) hi = tu;
        const char* from = pb.local + soff;
        for (u64 b0 = lo + gt; b0 < hi; b0 += 4ull * nthr) {
          uint4 v[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const u64 u = b0 + (u64)q * nthr;
            if (u < hi) v[q] = ld16(from + u * kUnit);
          }
#pragma unroll
          for (int k = 1; k < kMaxRanks; ++k) {
            if (k < N) {
              int p = r + k;
              if (p >= N) p -= N;
              if (p != root) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                  const u64 u = b0 + (u64)q * nthr;
                  if (u < hi) st16(c.stage[p] + soff + u * kUnit, v[q]);
                }
              }
            }
          }
        }
      }
      named_bar_sync(1 + grp, nthr);
      if (gt < N) {
        // flags only grow and a signal for tile i must imply the one for i-1: groups signal in tile order
        while (sig_next != i) {
        }
        if (r == root) {
          if (gt != r) st_release_sys(c.flags[gt] + rowA + r, E + 1 + i);  // A(i)
        } else if (peer_lane) {
          st_release_sys(c.flags[gt] + rowB + r, E + 1 + i);  // B(i)
        }
        __syncwarp((1u << N) - 1u);
        if (gt == 0) {
          __threadfence_block();
          sig_next = i + 1;
        }
      }
    }
  } else if (warp == 1 && r != root) {
    TmaRing R = {pipe_smem, bars, 0};
    const Region dst = make_region(ub.recv, op.bytes);
    for (u32 i = 0; i < g.rounds; ++i) {
      if (lane == root) wait_flag(c, c.flags[r] + rowA + root, E + 1 + i);
      else if (lane < N && lane != r) wait_flag(c, c.flags[r] + rowB + lane, E + 1 + i);
      __syncwarp();
      if (lane == 0) fence_proxy_async();  // acquired data is read by the async proxy next
      __syncwarp();
      const u64 tu = tile_units(g, i);
      if (tu) warp_tile_copy(R, dst, tile_first(g, i), tu, pb.local + slot_offset(g, i), false, lane);
      if (lane == 0) {
        bulk_wait_all0();  // slot fully read, user memory written
        while (sig_next < i + 1) {  // ... and this rank's PUSH warps are done reading it too
        }
      }
      __syncwarp();
      if (lane < N && lane != r) st_release_sys(c.flags[lane] + rowA + r, E + 1 + i);  // D(i)
    }
  }
  pipe_close(c, E, g.rounds, op);
}

// all_reduce on symmetric memory from the multicast pool (nbd_mem_alloc): no staging at all.
// Barrier, rank r sums shard r inside the switch and writes it to every rank, barrier.
// send_mc / recv_mc are multicast aliases; op.bytes is a multiple of 16 (the host checks).
struct McPtrs {
  const char* send_mc;
  char* recv_mc;
};
template <typename T>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_allreduce_mcz(const __grid_constant__ Pack<1> pk, const McPtrs m, const OpParams op) {
  const RankCtx& c = pk.ctx[0];
  const int N = c.world, r = c.rank;
  KernelClock clk = kernel_begin(c);
  flag_barrier(c, ++clk.epoch);
  const u64 units = op.bytes / kUnit;
  const u64 su = (units + N - 1) / N;
  const u64 first = (u64)r * su;
  const u64 mine = first >= units ? 0 : (units - first < su ? units - first : su);
  u64 lo, hi;
  block_span(su, mine, lo, hi);
  const char* src = m.send_mc + first * kUnit;
  char* dst = m.recv_mc + first * kUnit;
  for (u64 b0 = lo + threadIdx.x; b0 < hi; b0 += 8ull * kThreads) {
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const u64 u = b0 + (u64)j * kThreads;
      if (u < hi) v[j] = McSum<T>::ld(src + u * kUnit);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const u64 u = b0 + (u64)j * kThreads;
      if (u < hi) mc_st(dst + u * kUnit, op.avg ? unit_div<T>(v[j], N) : v[j]);
    }
  }
  flag_barrier(c, ++clk.epoch);
  kernel_end(c, 2, 0);
}

}  // namespace nbd
