// nvls_probe.cu — groundwork for round 2: can this box do in-switch reduction
// (NVLink SHARP / "NVLS") from our own kernels, and how fast?  NOT part of the product path and
// NOT yet run on hardware (written after the round's GPU budget was spent; it compiles for
// sm_100a).  One process drives N GPUs, so no file-descriptor passing is needed:
//
//   cuMulticastCreate(N devices) -> cuMulticastAddDevice x N
//   per device: cuMemCreate -> cuMulticastBindMem -> map a unicast VA (own memory, all devices
//   may access it) and the shared multicast VA
//   kernel on device d, shard d:  v = multimem.ld_reduce.add [mc + i]   (switch pulls and sums
//                                 the N copies)            multimem.st [mc + i], v  (switch
//                                 writes all N copies)
//
// Phases are separated by host-side synchronisation of all devices (kernels on different GPUs
// never wait on each other here).  Prints one JSON line per message size.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/_nvls_probe \
//        tools/nvls_probe.cu -lcuda     (tools/nvls_probe.py does it)
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#define CU(call)                                                                    \
  do {                                                                              \
    CUresult r__ = (call);                                                          \
    if (r__ != CUDA_SUCCESS) {                                                      \
      const char* s__ = nullptr;                                                    \
      cuGetErrorString(r__, &s__);                                                  \
      printf("{\"error\": \"%s failed: %s\"}\n", #call, s__ ? s__ : "?");         \
      return 2;                                                                     \
    }                                                                               \
  } while (0)
#define RT(call)                                                                    \
  do {                                                                              \
    cudaError_t e__ = (call);                                                       \
    if (e__ != cudaSuccess) {                                                       \
      printf("{\"error\": \"%s failed: %s\"}\n", #call, cudaGetErrorString(e__)); \
      return 2;                                                                     \
    }                                                                               \
  } while (0)

__global__ void k_fill(float* p, size_t n, float v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

// all_reduce of `shard_units` 16-byte units starting at unit `first`, entirely in the switch
__global__ void k_nvls_allreduce(char* mc, size_t first, size_t shard_units) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < shard_units; i += (size_t)gridDim.x * blockDim.x) {
    char* p = mc + (first + i) * 16;
    float a, b, c, d;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(a), "=f"(b), "=f"(c), "=f"(d)
                 : "l"(p)
                 : "memory");
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
  }
}

__global__ void k_check(const float* p, size_t n, float want, unsigned long long* bad) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    if (p[i] != want) atomicAdd(bad, 1ull);
}

int main(int argc, char** argv) {
  int ndev = 0;
  RT(cudaGetDeviceCount(&ndev));
  int N = argc > 1 ? atoi(argv[1]) : ndev;
  if (N > ndev) N = ndev;
  if (N < 2) {
    printf("{\"error\": \"needs >= 2 GPUs\"}\n");
    return 1;
  }
  CU(cuInit(0));
  std::vector<CUdevice> dev(N);
  for (int d = 0; d < N; ++d) {
    RT(cudaSetDevice(d));
    RT(cudaFree(0));
    CU(cuDeviceGet(&dev[d], d));
    int mc_ok = 0;
    CU(cuDeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev[d]));
    if (!mc_ok) {
      printf("{\"error\": \"device %d does not support multicast\"}\n", d);
      return 1;
    }
    for (int p = 0; p < N; ++p)
      if (p != d) {
        cudaError_t e = cudaDeviceEnablePeerAccess(p, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) RT(e);
        cudaGetLastError();
      }
  }

  const size_t want_bytes = 1ull << 30;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)N;
  mp.size = want_bytes;
  mp.handleTypes = 0;
  size_t gran = 0;
  CU(cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  const size_t size = (want_bytes + gran - 1) / gran * gran;
  mp.size = size;
  CUmemGenericAllocationHandle mc;
  CU(cuMulticastCreate(&mc, &mp));
  for (int d = 0; d < N; ++d) CU(cuMulticastAddDevice(mc, dev[d]));

  std::vector<CUmemGenericAllocationHandle> mem(N);
  std::vector<CUdeviceptr> uc(N);
  std::vector<CUmemAccessDesc> all(N);
  for (int d = 0; d < N; ++d) {
    all[d].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    all[d].location.id = d;
    all[d].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  for (int d = 0; d < N; ++d) {
    CUmemAllocationProp ap;
    memset(&ap, 0, sizeof(ap));
    ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    ap.location.id = d;
    CU(cuMemCreate(&mem[d], size, &ap, 0));
    CU(cuMulticastBindMem(mc, 0, mem[d], 0, size, 0));
    CU(cuMemAddressReserve(&uc[d], size, gran, 0, 0));
    CU(cuMemMap(uc[d], size, 0, mem[d], 0));
    CU(cuMemSetAccess(uc[d], size, all.data(), (size_t)N));
  }
  CUdeviceptr mcva = 0;
  CU(cuMemAddressReserve(&mcva, size, gran, 0, 0));
  CU(cuMemMap(mcva, size, 0, mc, 0));
  CU(cuMemSetAccess(mcva, size, all.data(), (size_t)N));

  std::vector<cudaStream_t> st(N);
  std::vector<cudaEvent_t> e0(N), e1(N);
  std::vector<unsigned long long*> bad(N);
  for (int d = 0; d < N; ++d) {
    RT(cudaSetDevice(d));
    RT(cudaStreamCreate(&st[d]));
    RT(cudaEventCreate(&e0[d]));
    RT(cudaEventCreate(&e1[d]));
    RT(cudaMalloc(&bad[d], sizeof(unsigned long long)));
  }
  auto sync_all = [&]() -> int {
    for (int d = 0; d < N; ++d) {
      RT(cudaSetDevice(d));
      RT(cudaDeviceSynchronize());
    }
    return 0;
  };

  const size_t sizes[] = {1ull << 20, 16ull << 20, 64ull << 20, 256ull << 20, 1ull << 30};
  for (size_t bytes : sizes) {
    const size_t units = bytes / 16, su = (units + N - 1) / N;
    for (int blocks : {32, 148, 296}) {
      float worst_ms = 0.f;
      const int iters = bytes <= (64ull << 20) ? 10 : 3;
      unsigned long long nbad = 0;
      for (int it = 0; it <= iters; ++it) {  // iteration 0 is the warm-up and the checked one
        for (int d = 0; d < N; ++d) {
          RT(cudaSetDevice(d));
          k_fill<<<296, 512, 0, st[d]>>>((float*)uc[d], bytes / 4, (float)(d + 1));
        }
        if (sync_all()) return 2;
        for (int d = 0; d < N; ++d) {
          RT(cudaSetDevice(d));
          const size_t first = (size_t)d * su;
          const size_t mine = first >= units ? 0 : (units - first < su ? units - first : su);
          RT(cudaEventRecord(e0[d], st[d]));
          k_nvls_allreduce<<<blocks, 512, 0, st[d]>>>((char*)mcva, first, mine);
          RT(cudaEventRecord(e1[d], st[d]));
        }
        if (sync_all()) return 2;
        if (it == 0) {
          const float want = (float)(N * (N + 1) / 2);
          for (int d = 0; d < N; ++d) {
            RT(cudaSetDevice(d));
            RT(cudaMemsetAsync(bad[d], 0, sizeof(unsigned long long), st[d]));
            k_check<<<296, 512, 0, st[d]>>>((const float*)uc[d], bytes / 4, want, bad[d]);
            unsigned long long b = 0;
            RT(cudaMemcpyAsync(&b, bad[d], sizeof(b), cudaMemcpyDeviceToHost, st[d]));
            RT(cudaStreamSynchronize(st[d]));
            nbad += b;
          }
        } else {
          for (int d = 0; d < N; ++d) {
            float ms = 0.f;
            RT(cudaEventElapsedTime(&ms, e0[d], e1[d]));
            if (ms > worst_ms) worst_ms = ms;  // slowest device of this iteration
          }
        }
      }
      // worst_ms holds the maximum over devices AND iterations: a conservative figure
      printf("{\"ranks\": %d, \"bytes\": %zu, \"blocks\": %d, \"us\": %.2f, \"busbw_GBps\": %.1f, \"wrong_elements\": %llu, "
             "\"granularity\": %zu}\n",
             N, bytes, blocks, worst_ms * 1e3, bytes * 2.0 * (N - 1) / N / (worst_ms * 1e-3) / 1e9, nbad, gran);
      fflush(stdout);
    }
  }
  return 0;
}
