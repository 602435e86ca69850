// nbd_p2p.cuh — point-to-point messages between ranks: the primitive under dist.send / dist.recv
// ($TORCH/.../Backend.hpp:355,364), all_to_all, gather and scatter (Backend.hpp:299-353), which
// cell code may call on the process group of ref:src/nbdistributed/worker.py:151 and which backend
// "nccl" serves there today.  No counterpart in the reference itself.
//
// One launch serves a BATCH of sends and receives of one rank, grouped into LANES: a lane is one
// (direction, peer) pair with its messages in issue order, worked by its own CTAs, so that a
// rank's sends and receives progress concurrently (a batch {send->1, recv<-1} on rank 0 and the
// same on rank 1 cannot deadlock, whatever the order they were issued in).
//
// Data path: sender CTAs copy the user buffer into the RECEIVER's P2P area (part of its arena,
// one region of kP2PSlots chunks of 1 MiB per source rank), receiver CTAs copy it out.  Unlike
// the collectives, only two ranks take part, so the all-rank epoch clock cannot be used: every
// ordered pair keeps its own chunk counter (LocalBlock::p2p_tx / p2p_rx) and signals with
//   ready[v][src]  (row kP2PRowReady + v of the receiver's flags) = chunks written so far
//   ack[v][dst]    (row kP2PRowAck + v of the sender's flags)     = chunks consumed so far
// where v < 32 indexes a fixed virtual span of the chunk.  The two sides may run different
// numbers of CTAs (1..32 each, a power of two): a CTA owns 32/G consecutive virtual spans.
#pragma once

#include "nbd_kernels.cuh"

namespace nbd {

constexpr int kP2PVirt = 32;
constexpr u64 kP2PChunk = 1ull << 20;
constexpr int kP2PSlots = 4;
constexpr int kP2PRowReady = 416, kP2PRowAck = 448;  // rows 416..479: above every grid's flag rows, below the export slots
constexpr u64 kP2PAreaBytes = (u64)kMaxRanks * kP2PSlots * kP2PChunk;  // 32 MiB per rank
constexpr int kP2PMaxLanes = 2 * (kMaxRanks - 1);
constexpr int kP2PLaneMsgs = 4;

struct P2PMsg {
  char* buf;
  u64 bytes;
};
struct P2PLane {
  int peer, send, nmsgs, cta0, nctas, pad;
  P2PMsg msg[kP2PLaneMsgs];
};
struct P2PRank {
  int nlanes, pad;
  P2PLane lane[kP2PMaxLanes];
};
template <int NCTX>
struct PackP2P {
  RankCtx ctx[NCTX];
  P2PRank r[NCTX];
};

// One CTA's share of one lane: its 32/G virtual spans of every chunk of every message.
__device__ __forceinline__ void p2p_lane(const RankCtx& c, const P2PLane& L, int b) {
  const int t = threadIdx.x, me = c.rank;

  const int peer = L.peer, G = L.nctas, j = b - L.cta0;
  const int vper = kP2PVirt / G, v0 = j * vper;
  LocalBlock* blk = reinterpret_cast<LocalBlock*>(c.state);
  u64 seq = *(L.send ? (volatile u64*)&blk->p2p_tx[peer] : (volatile u64*)&blk->p2p_rx[peer]);

  for (int m = 0; m < L.nmsgs; ++m) {
    const u64 bytes = L.msg[m].bytes;
    const u64 nchunks = (bytes + kP2PChunk - 1) / kP2PChunk;
    for (u64 k = 0; k < nchunks; ++k, ++seq) {
      const u64 off = k * kP2PChunk;
      const u64 cb = bytes - off < kP2PChunk ? bytes - off : kP2PChunk;
      const Region user = make_region(L.msg[m].buf + off, cb);
      const bool al = region_aligned(user);
      const u64 cu = units_of(cb), su = (cu + kP2PVirt - 1) / kP2PVirt;
      u64 lo = (u64)v0 * su, hi = (u64)(v0 + vper) * su;
      if (lo > cu) lo = cu;
      if (hi > cu) hi = cu;
      const u64 slot = seq % kP2PSlots;
      if (L.send) {
        char* dst = c.stage[peer] + c.p2p_off + ((u64)me * kP2PSlots + slot) * kP2PChunk;
        if (seq >= (u64)kP2PSlots && t < vper)  // the chunk that used this slot has been copied out
          wait_flag(c, c.flags[me] + (u64)(kP2PRowAck + v0 + t) * kMaxRanks + peer, seq - kP2PSlots + 1);
        __syncthreads();
        for (u64 base = lo + t; base < hi; base += 4ull * kThreads) {
          uint4 v[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const u64 u = base + (u64)q * kThreads;
            if (u < hi) v[q] = region_ld(user, al, u);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const u64 u = base + (u64)q * kThreads;
            if (u < hi) st16(dst + u * kUnit, v[q]);
          }
        }
        __syncthreads();
        if (t < vper) st_release_sys(c.flags[peer] + (u64)(kP2PRowReady + v0 + t) * kMaxRanks + me, seq + 1);
      } else {
        const char* src = c.stage[me] + c.p2p_off + ((u64)peer * kP2PSlots + slot) * kP2PChunk;
        if (t < vper) wait_flag(c, c.flags[me] + (u64)(kP2PRowReady + v0 + t) * kMaxRanks + peer, seq + 1);
        __syncthreads();
        for (u64 base = lo + t; base < hi; base += 4ull * kThreads) {
          uint4 v[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const u64 u = base + (u64)q * kThreads;
            if (u < hi) v[q] = ld16(src + u * kUnit);
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const u64 u = base + (u64)q * kThreads;
            if (u < hi) region_st(user, al, u, v[q]);
          }
        }
        __syncthreads();
        if (t < vper) st_release_sys(c.flags[peer] + (u64)(kP2PRowAck + v0 + t) * kMaxRanks + me, seq + 1);
      }
    }
  }
}

template <int NCTX>
__global__ void __launch_bounds__(kThreads, kMinBlocksPerSM)
k_p2p_batch(const __grid_constant__ PackP2P<NCTX> pk) {
  const int ci = ctx_index<NCTX>();
  const RankCtx& c = pk.ctx[ci];
  const P2PRank& pr = pk.r[ci];
  const int b = blockIdx.x, t = threadIdx.x;
  int li = -1;
  for (int l = 0; l < pr.nlanes; ++l)
    if (b >= pr.lane[l].cta0 && b < pr.lane[l].cta0 + pr.lane[l].nctas) li = l;
  if (li >= 0) p2p_lane(c, pr.lane[li], b);
  // The pair counters advance once per launch, by the last CTA to finish: every CTA of a lane
  // read them when it started, and CTAs of one launch start at different times.
  __syncthreads();
  if (t == 0) {
    LocalBlock* blk = reinterpret_cast<LocalBlock*>(c.state);
    const u32 prev = atomicAdd(&blk->st.done, 1u);
    if (prev == gridDim.x - 1) {
      blk->st.done = 0;
      for (int l = 0; l < pr.nlanes; ++l) {
        u64 chunks = 0;
        for (int m = 0; m < pr.lane[l].nmsgs; ++m) chunks += (pr.lane[l].msg[m].bytes + kP2PChunk - 1) / kP2PChunk;
        if (pr.lane[l].send) blk->p2p_tx[pr.lane[l].peer] += chunks;
        else blk->p2p_rx[pr.lane[l].peer] += chunks;
      }
      __threadfence();
    }
  }
}

}  // namespace nbd
