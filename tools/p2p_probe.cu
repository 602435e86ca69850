// p2p_probe.cu — microbenchmark of GPU-to-GPU copy strategies over NVLink, used to choose how
// the collective kernels should move bytes (measure, don't guess).  One process drives two GPUs
// with peer access enabled.  Variants:
//   mode 0  pull : thread loads U x 16 B from the PEER, stores locally
//   mode 1  push : thread loads U x 16 B locally, stores to the PEER
//   mode 2  both : half of the CTAs pull, half push (traffic in both directions from one GPU)
//   mode 3  tma pull : one thread per CTA, cp.async.bulk peer -> smem ring -> local
//   mode 4  tma push : one thread per CTA, cp.async.bulk local -> smem ring -> peer
// Built by tools/p2p_probe.py with nvcc -gencode arch=compute_100a,code=sm_100a.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

typedef unsigned long long u64;

__device__ __forceinline__ uint4 ld16(const void* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st16(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w));
}

template <int U>
__global__ void k_copy(char* __restrict__ dst, const char* __restrict__ src, u64 units, int split,
                       char* dst2, const char* src2) {
  // split: CTAs with odd index copy src2 -> dst2 instead (mode "both")
  const bool second = split && (blockIdx.x & 1);
  char* d = second ? dst2 : dst;
  const char* s = second ? src2 : src;
  const u64 nb = split ? gridDim.x / 2 : gridDim.x;
  const u64 b = split ? blockIdx.x / 2 : blockIdx.x;
  const u64 per = (units + nb - 1) / nb;
  const u64 lo = b * per, hi = (lo + per < units) ? lo + per : units;
  for (u64 base = lo + threadIdx.x; base < hi; base += (u64)U * blockDim.x) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * blockDim.x;
      if (i < hi) v[j] = ld16(s + i * 16);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const u64 i = base + (u64)j * blockDim.x;
      if (i < hi) st16(d + i * 16, v[j]);
    }
  }
}

// ---- TMA (bulk async copy) ring: global -> shared -> global, driven by one thread per CTA ----
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

constexpr int kMaxStages = 16;

__global__ void k_tma_copy(char* dst, const char* src, u64 bytes, int tile, int stages) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t bars[kMaxStages];
  const u64 per = ((bytes / gridDim.x) + 127) / 128 * 128;
  const u64 lo = (u64)blockIdx.x * per;
  if (lo >= bytes) return;
  const u64 hi = (lo + per < bytes) ? lo + per : bytes;
  const u64 ntiles = (hi - lo + tile - 1) / tile;
  if (threadIdx.x != 0) return;
  for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  auto tile_bytes = [&](u64 i) { return (uint32_t)((lo + (i + 1) * tile <= hi) ? tile : (hi - lo - i * tile)); };
  for (u64 i = 0; i < ntiles && i < (u64)stages; ++i) {
    mbar_expect_tx(&bars[i], tile_bytes(i));
    bulk_load(smem + (size_t)i * tile, src + lo + i * tile, tile_bytes(i), &bars[i]);
  }
  for (u64 i = 0; i < ntiles; ++i) {
    const int slot = (int)(i % stages);
    mbar_wait(&bars[slot], (uint32_t)((i / stages) & 1));
    bulk_store(dst + lo + i * tile, smem + (size_t)slot * tile, tile_bytes(i));
    bulk_commit();
    const u64 nxt = i + stages;
    if (nxt < ntiles) {
      bulk_wait_read<0>();  // the store has drained this slot
      mbar_expect_tx(&bars[slot], tile_bytes(nxt));
      bulk_load(smem + (size_t)slot * tile, src + lo + nxt * tile, tile_bytes(nxt), &bars[slot]);
    }
  }
  bulk_wait_all<0>();
}

extern "C" {

int probe_enable_peer(int dev, int peer) {
  cudaSetDevice(dev);
  int can = 0;
  cudaDeviceCanAccessPeer(&can, dev, peer);
  if (!can) return -1;
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return -2;
  cudaGetLastError();
  return 0;
}

// Launches on the current device.  Returns 0 or a cudaError.
int probe_copy(void* dst, const void* src, u64 bytes, int blocks, int threads, int unroll, int split, void* dst2,
               const void* src2, void* stream) {
  const u64 units = bytes / 16;
  cudaStream_t s = (cudaStream_t)stream;
  switch (unroll) {
    case 1: k_copy<1><<<blocks, threads, 0, s>>>((char*)dst, (const char*)src, units, split, (char*)dst2, (const char*)src2); break;
    case 2: k_copy<2><<<blocks, threads, 0, s>>>((char*)dst, (const char*)src, units, split, (char*)dst2, (const char*)src2); break;
    case 4: k_copy<4><<<blocks, threads, 0, s>>>((char*)dst, (const char*)src, units, split, (char*)dst2, (const char*)src2); break;
    case 8: k_copy<8><<<blocks, threads, 0, s>>>((char*)dst, (const char*)src, units, split, (char*)dst2, (const char*)src2); break;
    default: return -1;
  }
  return (int)cudaGetLastError();
}

int probe_tma_copy(void* dst, const void* src, u64 bytes, int blocks, int tile, int stages, void* stream) {
  if (stages > kMaxStages || tile % 16) return -1;
  const size_t smem = (size_t)tile * stages;
  cudaError_t e = cudaFuncSetAttribute(k_tma_copy, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  k_tma_copy<<<blocks, 32, smem, (cudaStream_t)stream>>>((char*)dst, (const char*)src, bytes, tile, stages);
  return (int)cudaGetLastError();
}
}
