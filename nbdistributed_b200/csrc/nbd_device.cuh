// nbd_device.cuh — device-side building blocks shared by every collective
// kernel: arena layout, release/acquire flag barriers between the same-index
// CTA of every rank, 128-bit unit accessors for (possibly misaligned) user
// memory, typed reducers with a fixed ascending-rank order.
//
// Written for sm_100a only.  No counterpart in the reference (which has no
// native code); the semantics implemented are those of c10d's collectives as
// called from ref:src/nbdistributed/worker.py:162 (see include/nbdist_b200.h).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace nbd {

typedef unsigned long long u64;
typedef unsigned int u32;

constexpr int kMaxRanks = 8;      // NBD_MAX_RANKS
constexpr int kMaxBlocks = 1024;  // flag rows per arena
constexpr int kThreads = 512;     // CTA size of every collective kernel
constexpr int kMinBlocksPerSM = 1;  // guaranteed co-resident CTAs per SM (more if registers allow)
constexpr int kUnit = 16;         // bytes moved per 128-bit access

// Arena (one per rank, IPC-mapped by every peer):
//   [0, kFlagBytes)            u64 flag[kMaxBlocks][kMaxRanks]; flag[b][src] is
//                              written by rank `src`'s CTA b, read by our CTA b
//   [kFlagBytes, +stage_bytes) staging, two halves of half_bytes each
constexpr u64 kFlagBytes = (u64)kMaxBlocks * kMaxRanks * sizeof(u64);

// Rank-private device state (not mapped by peers).
struct LocalState {
  u64 epoch;      // flag-barrier epochs consumed by completed launches
  u64 chunk_seq;  // staging steps consumed; parity picks the staging half
  u32 done;       // CTAs of the running launch that have finished
  u32 failed;     // set (never cleared) when a flag wait of this rank gave up: later waits drain
  u64 ll_seq;     // LL (flag-in-data) operations completed; parity picks the LL buffer
  u64 ll_hist;    // bit 0: the last LL operation was a broadcast; bit 1: the one before it was
};

// Pinned host memory, visible to the device: async error reporting + abort.
constexpr int kMissRing = 24;
struct HostMailbox {
  u32 error;  // 0, or -nbd_result set by a kernel (timeout / abort)
  u32 abort;  // host sets to 1 to make flag waits give up
  u32 miss_count;        // auto-registration: number of (peer, export slot) pairs reported so far
  u32 pad;
  u64 miss[kMissRing];   // ring of (peer << 32 | slot) this rank has not imported yet
};

struct RankCtx {
  int rank;
  int world;
  u64* flags[kMaxRanks];  // flags[p]: rank p's flag array in OUR address space
  char* stage[kMaxRanks]; // stage[p]: rank p's staging base in OUR address space
  LocalState* state;
  HostMailbox* mbox;
  u64 half_bytes;   // bytes in one staging half (multiple of 16)
  u64 timeout_ns;   // 0 = never
  u64 ll_off;       // offset of the LL area inside every rank's staging (0: no LL area)
  u64 ll_slot;      // bytes of one source rank's slot in the LL area (2 x the LL payload limit)
  u64 p2p_off;      // offset of the point-to-point area (nbd_p2p.cuh) from every rank's staging base (0: none)
};

// LL ("low latency") area at the tail of the staging region: for each parity, one slot per
// source rank; a slot (RankCtx::ll_slot bytes) holds the LL payload limit of the communicator as
// 8-byte {data32, flag32} packets.  The limit is kLLBigMax when the staging is large enough to
// spare 8 MiB for it, else kLLMaxBytes (small test arenas).
constexpr u64 kLLMaxBytes = 32ull << 10;
constexpr u64 kLLBigMax = 256ull << 10;
constexpr u64 kLLBigStage = 64ull << 20;  // staging sizes from here on carry the large LL area
__host__ __device__ constexpr u64 ll_cap_of(u64 stage_bytes) { return stage_bytes >= kLLBigStage ? kLLBigMax : kLLMaxBytes; }
__host__ __device__ constexpr u64 ll_area_of(u64 stage_bytes) { return 2 * (u64)kMaxRanks * 2 * ll_cap_of(stage_bytes); }  // 1 or 8 MiB
// Row of the flag array where rank `src` publishes "LL broadcasts I have finished reading" (word
// = src, value = sequence number + 1).  A rank that WRITES packets for LL operation s must know
// that every receiver has finished operation s-2, the previous user of the same parity buffer.
// After an all_reduce that is implied (one has received every peer's packets of s-1, which they
// sent after finishing s-2); a broadcast's root hears nothing from its peers, so broadcasts are
// acknowledged, and only an operation that directly follows a broadcast has to look (ll_hist).
constexpr int kLLAckRow = 480;

struct RankBufs {
  const char* send;
  char* recv;
};

// NCTX = 1: a real rank (one process per GPU).  NCTX = kMaxRanks: emulated
// ranks, blockIdx.y selects the rank, launched cooperatively so that every
// rank's CTAs are co-resident.
template <int NCTX>
struct Pack {
  RankCtx ctx[NCTX];
  RankBufs buf[NCTX];
  // reduce_scatter, list form: shard[r][p] = rank r's contribution to rank p's result when the N
  // contributions are separate buffers; all null = one contiguous send buffer (buf[r].send).
  const char* shard[NCTX][kMaxRanks];
};

// Zero-copy collectives: every rank's user buffers are directly addressable (registered
// symmetric memory between processes, or simply local for emulated ranks).  send[p] / recv[p]
// are rank p's buffers of THIS operation in our address space (null where not needed).
struct ZcPtrs {
  const char* send[kMaxRanks];
  char* recv[kMaxRanks];
  u64 desc[6];  // {send registration, send offset, recv registration, recv offset, bytes, op tag}
};

template <int NCTX>
struct PackZc {
  RankCtx ctx[NCTX];
  ZcPtrs z;
};

// Auto-registration (ordinary, unregistered tensors): each rank exports the allocations its
// tensors live in (CUDA IPC) into EXPORT SLOTS of its arena, peers import them lazily.  Before
// an eligible collective a one-CTA kernel (k_zc_agree) exchanges {slot, offset} descriptors,
// resolves them through the local import table and agrees, across ranks, whether everybody
// can address everybody's buffers.  The zero-copy kernel and the staged kernel are then both
// enqueued; each returns at once if the decision is not for it.
constexpr int kExportRow0 = 512;                       // export slots use flag rows [512, 1008)
constexpr int kExportSlotRows = 2;                     // 2 rows x 64 B = 128 B per slot
constexpr int kMaxExports = (kMaxBlocks - 2 * kMaxRanks - kExportRow0) / kExportSlotRows;  // 248
constexpr u64 kNoSlot = ~0ull;

struct ExportSlot {   // 128 B, written by the exporting rank's host, read by peers' hosts
  u64 buffer_id;      // CU_POINTER_ATTRIBUTE_BUFFER_ID of the allocation (unique per allocation)
  u64 base;
  u64 size;
  u64 valid;          // 1 once the record is complete
  unsigned char ipc[64];
  u64 pad[4];
};

struct ZcDecision {
  u32 ok;             // 1: every rank can address every rank's buffers of the pending operation
  u32 pad;
  ZcPtrs z;
};

// Rank-private device block; LocalState stays first so that RankCtx::state needs no change.
struct LocalBlock {
  LocalState st;
  ZcDecision dec;
  u64 imp[kMaxRanks][kMaxExports];  // imp[p][slot]: rank p's exported allocation mapped here, or 0
  u64 prev_desc[kMaxRanks][6];      // every rank's descriptor of the previous agreement
  u64 p2p_tx[kMaxRanks];            // point-to-point chunks sent to rank p so far (nbd_p2p.cuh)
  u64 p2p_rx[kMaxRanks];            // ... and received from rank p
};

// Operation descriptors for the symmetric-use check live in flag rows no grid ever reaches
// (grids are capped well below kDescRow CTAs): row kDescRow + src holds rank src's 8 words.
constexpr int kDescRow = kMaxBlocks - 2 * kMaxRanks;

struct OpParams {
  u64 bytes;        // all_reduce/broadcast: message bytes; all_gather /
                    // reduce_scatter: bytes of ONE rank's shard
  u64 chunk_bytes;  // bytes handled per staging step (multiple of 16)
  u32 nchunks;
  int root;
  int op;           // nbd_redop
  int algo;         // kernel-specific variant
  int world;
  int avg;          // 1: divide the sum by world (NBD_AVG)
  int zc_mode;      // 0: unconditional; 1: zero-copy kernel, run only if the agreed decision is ok
                    // (buffers come from the decision); 2: staged kernel, run only if it is not
  int depth;        // tile pipeline (nbd_pipe.cuh): staging slots per CTA
};

// ---------------------------------------------------------------------------
// raw memory operations
// ---------------------------------------------------------------------------

__device__ __forceinline__ uint4 ld16(const void* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

__device__ __forceinline__ void st16(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w));
}

__device__ __forceinline__ void st_release_sys(u64* p, u64 v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v)
               : "memory");
}

__device__ __forceinline__ u64 ld_relaxed_sys(const u64* p) {
  u64 v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void fence_acq_rel_sys() {
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}

__device__ __forceinline__ u64 globaltimer_ns() {
  u64 t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// ---------------------------------------------------------------------------
// flag barrier between CTA b of every rank
// ---------------------------------------------------------------------------

// Spin until *f >= epoch (flags only ever grow).  Gives up on abort/timeout and
// records why in the host mailbox; the communicator is dead after that.
static __device__ __noinline__ void wait_flag_slow(LocalState* st, HostMailbox* mbox, u64 timeout_ns, const u64* f,
                                                   u64 epoch) {
  u64 t0 = 0;
  u32 spins = 0;
  // A launch that already gave up once (timeout/abort) must drain, not wait a full timeout at
  // every later flag: its results are void anyway and the communicator is dead.  The marker is
  // device memory on purpose: the mailbox is pinned HOST memory, and reading it on every entry
  // from hundreds of CTAs cost ~75 us per flag (PCIe reads serialise; measured on 2xB200).
  if (reinterpret_cast<volatile LocalState*>(st)->failed != 0) return;
  while (ld_relaxed_sys(f) < epoch) {
    if ((++spins & 0x3FFFu) == 0) {
      const u64 now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      volatile HostMailbox* mb = mbox;
      if (mb->abort) {
        mb->error = 5;  // -NBD_ERR_ABORTED
        reinterpret_cast<volatile LocalState*>(st)->failed = 1;
        break;
      }
      if (timeout_ns != 0 && now - t0 > timeout_ns) {
        mb->error = 4;  // -NBD_ERR_TIMEOUT
        reinterpret_cast<volatile LocalState*>(st)->failed = 1;
        break;
      }
      if (reinterpret_cast<volatile LocalState*>(st)->failed != 0) break;  // another thread of this rank gave up
    }
  }
}

__device__ __forceinline__ void wait_flag(const RankCtx& c, const u64* f,
                                          u64 epoch) {
  if (ld_relaxed_sys(f) < epoch) wait_flag_slow(c.state, c.mbox, c.timeout_ns, f, epoch);
  fence_acq_rel_sys();  // relaxed load + fence = acquire; also drops stale L1 lines
}

// Full barrier: on return, everything CTA b of every rank wrote before ITS
// call (to any arena or to its own memory) is visible to this CTA.
// bar.sync orders the CTA's writes before thread t's release store
// (cumulativity), and the acquiring thread's fence before the CTA's reads.
__device__ __forceinline__ void flag_barrier(const RankCtx& c, u64 epoch) {
  __syncthreads();
  const int t = threadIdx.x;
  if (t < c.world && t != c.rank) {
    const u64 row = (u64)blockIdx.x * kMaxRanks;
    st_release_sys(c.flags[t] + row + c.rank, epoch);
    wait_flag(c, c.flags[c.rank] + row + t, epoch);
  }
  __syncthreads();
}

struct KernelClock {
  u64 epoch;
  u64 seq;
};

__device__ __forceinline__ KernelClock kernel_begin(const RankCtx& c) {
  KernelClock k;
  const volatile LocalState* s = c.state;
  k.epoch = s->epoch;
  k.seq = s->chunk_seq;
  return k;
}

// The last CTA of this rank to finish advances the rank-private clocks.  Every
// CTA read them in kernel_begin() before it could be counted here.
__device__ __forceinline__ void kernel_end(const RankCtx& c, u64 nbarriers,
                                           u64 nsteps) {
  __syncthreads();
  if (threadIdx.x == 0) {
    LocalState* s = c.state;
    const u32 prev = atomicAdd(&s->done, 1u);
    if (prev == gridDim.x - 1) {
      s->done = 0;
      s->epoch += nbarriers;
      s->chunk_seq += nsteps;
      __threadfence();
    }
  }
}

// ---------------------------------------------------------------------------
// user memory as 16-byte units
// ---------------------------------------------------------------------------

// A span of user memory.  Unit i = bytes [16i, 16i+16) clipped to n.  When p is
// 16-byte aligned, full units take the 128-bit path; a misaligned base or the
// ragged last unit go byte by byte (slow, correct, rare: storage_offset views
// and counts that are not a multiple of 16 bytes).
struct Region {
  char* p;
  u64 n;
};

__device__ __forceinline__ Region make_region(const void* p, u64 n) {
  Region r;
  r.p = const_cast<char*>(static_cast<const char*>(p));
  r.n = n;
  return r;
}

__device__ __forceinline__ u64 units_of(u64 bytes) { return (bytes + kUnit - 1) / kUnit; }

static __device__ __noinline__ uint4 region_ld_slow(const char* rp, u64 rn, u64 i) {
  uint4 v = make_uint4(0, 0, 0, 0);
  const u64 off = i * kUnit;
  const volatile unsigned char* s = reinterpret_cast<const volatile unsigned char*>(rp) + off;
  u32 w[4] = {0, 0, 0, 0};
#pragma unroll
  for (int j = 0; j < kUnit; ++j) {
    if (off + j < rn) w[j >> 2] |= (u32)s[j] << (8 * (j & 3));
  }
  v.x = w[0]; v.y = w[1]; v.z = w[2]; v.w = w[3];
  return v;
}

static __device__ __noinline__ void region_st_slow(char* rp, u64 rn, u64 i, uint4 v) {
  const u64 off = i * kUnit;
  volatile unsigned char* d = reinterpret_cast<volatile unsigned char*>(rp) + off;
  const u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < kUnit; ++j) {
    if (off + j < rn) d[j] = (unsigned char)(w[j >> 2] >> (8 * (j & 3)));
  }
}

__device__ __forceinline__ bool region_aligned(const Region& r) {
  return (reinterpret_cast<uintptr_t>(r.p) & (kUnit - 1)) == 0;
}

__device__ __forceinline__ uint4 region_ld(const Region& r, bool aligned, u64 i) {
  if (aligned && (i + 1) * kUnit <= r.n) return ld16(r.p + i * kUnit);
  return region_ld_slow(r.p, r.n, i);
}

__device__ __forceinline__ void region_st(const Region& r, bool aligned, u64 i, const uint4& v) {
  if (aligned && (i + 1) * kUnit <= r.n) {
    st16(r.p + i * kUnit, v);
  } else {
    region_st_slow(r.p, r.n, i, v);
  }
}

// CTA b of nb owns units [lo, hi) of a span.  The cut is made on `full_units` (the span
// of a full-size chunk) and clipped to `units` (this chunk), so that a CTA touches the same
// staging offsets in every chunk: the flag barriers only order same-index CTAs.
__device__ __forceinline__ void block_span(u64 full_units, u64 units, u64& lo, u64& hi) {
  const u64 nb = gridDim.x;
  u64 per = (full_units + nb - 1) / nb;
  per = (per + kThreads - 1) / kThreads * kThreads;
  lo = (u64)blockIdx.x * per;
  if (lo > units) lo = units;
  hi = lo + per;
  if (hi > units) hi = units;
}

// Units [0, fast_units(r)) of a region can be moved with plain 128-bit accesses; the rest
// (a ragged last unit, or everything if the base is misaligned) go through the checked path.
// Loops are split accordingly so that the hot loop carries no per-access branches.
__device__ __forceinline__ u64 fast_units(const Region& r) { return region_aligned(r) ? r.n / kUnit : 0; }

template <int U, typename LD, typename ST>
__device__ __forceinline__ void copy_units(u64 lo, u64 hi, LD ld, ST st) {
  for (u64 base = lo + threadIdx.x; base < hi; base += (u64)U * kThreads) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) v[j] = ld(i);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) st(i, v[j]);
    }
  }
}

// NOTE: the three helpers below keep the per-access checked form on purpose.  Splitting them
// into an unchecked hot loop plus a checked tail (as the zero-copy kernels do with copy_units)
// raised the register pressure of k_allreduce_twoshot (stack 200 -> 328 B) and cost the staged
// two-shot all_reduce 18 % at 8xB200 (256 MiB: 574 -> 471 GB/s, profiles/r01_bench_n8_*.json).
// Copy units [lo,hi): user region -> staging (dst may be a peer's arena).
template <int U>
__device__ __forceinline__ void copy_region_to_stage(char* dst, const Region& src, u64 lo, u64 hi) {
  const bool al = region_aligned(src);
  for (u64 base = lo + threadIdx.x; base < hi; base += (u64)U * kThreads) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) v[j] = region_ld(src, al, i);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) st16(dst + i * kUnit, v[j]);
    }
  }
}

// Copy units [lo,hi): staging (maybe a peer's) -> user region.
template <int U>
__device__ __forceinline__ void copy_stage_to_region(const Region& dst, const char* src, u64 lo, u64 hi) {
  const bool al = region_aligned(dst);
  for (u64 base = lo + threadIdx.x; base < hi; base += (u64)U * kThreads) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) v[j] = ld16(src + i * kUnit);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) region_st(dst, al, i, v[j]);
    }
  }
}

// Copy units [lo,hi): user region -> user region (both local).
template <int U>
__device__ __forceinline__ void copy_region_to_region(const Region& dst, const Region& src, u64 lo, u64 hi) {
  const bool als = region_aligned(src), ald = region_aligned(dst);
  for (u64 base = lo + threadIdx.x; base < hi; base += (u64)U * kThreads) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) v[j] = region_ld(src, als, i);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * kThreads;
      if (i < hi) region_st(dst, ald, i, v[j]);
    }
  }
}

// ---------------------------------------------------------------------------
// typed reduction of 16-byte units, ascending rank order
// ---------------------------------------------------------------------------

enum { OP_SUM = 0, OP_AVG = 1, OP_PROD = 2, OP_MIN = 3, OP_MAX = 4, OP_BAND = 5, OP_BOR = 6, OP_BXOR = 7 };

template <typename A>
__device__ __forceinline__ A fmin_nan(A a, A b) { return (a != a) ? a : ((b != b) ? b : (b < a ? b : a)); }
template <typename A>
__device__ __forceinline__ A fmax_nan(A a, A b) { return (a != a) ? a : ((b != b) ? b : (b > a ? b : a)); }

template <typename A, bool IS_FLOAT>
struct Combine;

template <typename A>
struct Combine<A, true> {
  static __device__ __forceinline__ A apply(A a, A b, int op) {
    switch (op) {
      case OP_PROD: return a * b;
      case OP_MIN: return fmin_nan(a, b);
      case OP_MAX: return fmax_nan(a, b);
      default: return a + b;  // SUM, AVG
    }
  }
};

template <typename A>
struct Combine<A, false> {
  // A is the unsigned carrier for SUM/PROD (wrap-around two's complement) and
  // the signed/unsigned element type itself decides MIN/MAX via S.
  template <typename S>
  static __device__ __forceinline__ A apply_s(A a, A b, int op) {
    switch (op) {
      case OP_PROD: return a * b;
      case OP_MIN: return ((S)b < (S)a) ? b : a;
      case OP_MAX: return ((S)b > (S)a) ? b : a;
      case OP_BAND: return a & b;
      case OP_BOR: return a | b;
      case OP_BXOR: return a ^ b;
      default: return a + b;
    }
  }
};

// Reducer<T>: acc = x_0; acc = op(acc, x_k) for k = 1..N-1; one rounding at the end.
template <typename T>
struct Reducer;

template <>
struct Reducer<float> {
  float a[4];
  __device__ __forceinline__ void init(const uint4& v) {
    a[0] = __uint_as_float(v.x); a[1] = __uint_as_float(v.y);
    a[2] = __uint_as_float(v.z); a[3] = __uint_as_float(v.w);
  }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    const float b[4] = {__uint_as_float(v.x), __uint_as_float(v.y), __uint_as_float(v.z), __uint_as_float(v.w)};
#pragma unroll
    for (int e = 0; e < 4; ++e) a[e] = SUM ? __fadd_rn(a[e], b[e]) : Combine<float, true>::apply(a[e], b[e], op);
  }
  __device__ __forceinline__ uint4 result(int avg, int world) {
    if (avg) {
#pragma unroll
      for (int e = 0; e < 4; ++e) a[e] = __fdiv_rn(a[e], (float)world);
    }
    return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]));
  }
};

template <>
struct Reducer<double> {
  double a[2];
  __device__ __forceinline__ void init(const uint4& v) {
    a[0] = __longlong_as_double(((long long)v.y << 32) | v.x);
    a[1] = __longlong_as_double(((long long)v.w << 32) | v.z);
  }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    const double b[2] = {__longlong_as_double(((long long)v.y << 32) | v.x),
                         __longlong_as_double(((long long)v.w << 32) | v.z)};
#pragma unroll
    for (int e = 0; e < 2; ++e) a[e] = SUM ? __dadd_rn(a[e], b[e]) : Combine<double, true>::apply(a[e], b[e], op);
  }
  __device__ __forceinline__ uint4 result(int avg, int world) {
    if (avg) { a[0] = __ddiv_rn(a[0], (double)world); a[1] = __ddiv_rn(a[1], (double)world); }
    const u64 x = (u64)__double_as_longlong(a[0]), y = (u64)__double_as_longlong(a[1]);
    return make_uint4((u32)x, (u32)(x >> 32), (u32)y, (u32)(y >> 32));
  }
};

// bf16 / f16: fp32 accumulator, single round-to-nearest-even at the end.
template <>
struct Reducer<__nv_bfloat16> {
  float a[8];
  static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
    const u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  }
  __device__ __forceinline__ void init(const uint4& v) { unpack(v, a); }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    float b[8];
    unpack(v, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] = SUM ? __fadd_rn(a[e], b[e]) : Combine<float, true>::apply(a[e], b[e], op);
  }
  __device__ __forceinline__ uint4 result(int avg, int world) {
    if (avg) {
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] = __fdiv_rn(a[e], (float)world);
    }
    u32 w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
      w[i] = *reinterpret_cast<const u32*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

template <>
struct Reducer<__half> {
  float a[8];
  static __device__ __forceinline__ void unpack(const uint4& v, float* f) {
    const u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 p = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      f[2 * i] = p.x;
      f[2 * i + 1] = p.y;
    }
  }
  __device__ __forceinline__ void init(const uint4& v) { unpack(v, a); }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    float b[8];
    unpack(v, b);
#pragma unroll
    for (int e = 0; e < 8; ++e) a[e] = SUM ? __fadd_rn(a[e], b[e]) : Combine<float, true>::apply(a[e], b[e], op);
  }
  __device__ __forceinline__ uint4 result(int avg, int world) {
    if (avg) {
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] = __fdiv_rn(a[e], (float)world);
    }
    u32 w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __half2 h = __floats2half2_rn(a[2 * i], a[2 * i + 1]);
      w[i] = *reinterpret_cast<const u32*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// 32-bit integers (S = int32_t): unsigned carrier, wrap-around.
template <>
struct Reducer<int32_t> {
  u32 a[4];
  __device__ __forceinline__ void init(const uint4& v) { a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w; }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    const u32 b[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) a[e] = SUM ? a[e] + b[e] : Combine<u32, false>::apply_s<int32_t>(a[e], b[e], op);
  }
  __device__ __forceinline__ uint4 result(int, int) { return make_uint4(a[0], a[1], a[2], a[3]); }
};

template <>
struct Reducer<int64_t> {
  u64 a[2];
  __device__ __forceinline__ void init(const uint4& v) { a[0] = ((u64)v.y << 32) | v.x; a[1] = ((u64)v.w << 32) | v.z; }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    const u64 b[2] = {((u64)v.y << 32) | v.x, ((u64)v.w << 32) | v.z};
#pragma unroll
    for (int e = 0; e < 2; ++e) a[e] = SUM ? a[e] + b[e] : Combine<u64, false>::apply_s<long long>(a[e], b[e], op);
  }
  __device__ __forceinline__ uint4 result(int, int) {
    return make_uint4((u32)a[0], (u32)(a[0] >> 32), (u32)a[1], (u32)(a[1] >> 32));
  }
};

// 8-bit integers: 16 lanes, 32-bit carriers truncated at the end (SUM/PROD are
// congruent mod 256; MIN/MAX/bitwise stay within 8 bits).
template <typename S8>
struct Reducer8 {
  u32 a[16];
  static __device__ __forceinline__ void unpack(const uint4& v, u32* o) {
    const u32 w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const u32 byte = (w[j >> 2] >> (8 * (j & 3))) & 0xFFu;
      // sign- or zero-extend so signed MIN/MAX compare correctly
      o[j] = (u32)(int)(S8)byte;
    }
  }
  __device__ __forceinline__ void init(const uint4& v) { unpack(v, a); }
  template <bool SUM>
  __device__ __forceinline__ void add(const uint4& v, int op) {
    u32 b[16];
    unpack(v, b);
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const u32 r = SUM ? a[e] + b[e] : Combine<u32, false>::apply_s<int>(a[e], b[e], op);
      a[e] = (u32)(int)(S8)(r & 0xFFu);  // renormalise to the element's range
    }
  }
  __device__ __forceinline__ uint4 result(int, int) {
    u32 w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int j = 0; j < 16; ++j) w[j >> 2] |= (a[j] & 0xFFu) << (8 * (j & 3));
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <>
struct Reducer<uint8_t> : Reducer8<uint8_t> {};
template <>
struct Reducer<int8_t> : Reducer8<int8_t> {};

}  // namespace nbd
