// Tile-pipeline kernels (nbd_pipe.cuh): all_reduce for every dtype over the peer-to-peer
// transport (real and emulated ranks), all_reduce of f32/bf16/f16 sums and broadcast over the
// NVSwitch multicast transport (real ranks only).
#include "../../include/nbdist_b200.h"
#include "nbd_kernel_table.h"
#include "nbd_pipe.cuh"

namespace nbd {
namespace {
template <typename T>
const void* p2p(bool sum, bool emu) {
  if (emu) return sum ? (const void*)k_allreduce_pipe<T, true, kMaxRanks, false> : (const void*)k_allreduce_pipe<T, false, kMaxRanks, false>;
  return sum ? (const void*)k_allreduce_pipe<T, true, 1, false> : (const void*)k_allreduce_pipe<T, false, 1, false>;
}
}  // namespace

const void* pipe_allreduce_kernel(int dtype, bool sum, bool emu, bool mc) {
  if (mc) {
    if (emu || !sum) return nullptr;  // the switch only adds, and only between real ranks
    switch (dtype) {
      case NBD_F32: return (const void*)k_allreduce_pipe<float, true, 1, true>;
      case NBD_BF16: return (const void*)k_allreduce_pipe<__nv_bfloat16, true, 1, true>;
      case NBD_F16: return (const void*)k_allreduce_pipe<__half, true, 1, true>;
      default: return nullptr;
    }
  }
  switch (dtype) {
    case NBD_F32: return p2p<float>(sum, emu);
    case NBD_BF16: return p2p<__nv_bfloat16>(sum, emu);
    case NBD_F16: return p2p<__half>(sum, emu);
    case NBD_F64: return p2p<double>(sum, emu);
    case NBD_I32: return p2p<int32_t>(sum, emu);
    case NBD_I64: return p2p<int64_t>(sum, emu);
    case NBD_U8: return p2p<uint8_t>(sum, emu);
    case NBD_I8: return p2p<int8_t>(sum, emu);
    default: return nullptr;
  }
}

const void* pipe_broadcast_kernel(bool emu, bool mc) {
  if (mc) return emu ? nullptr : (const void*)k_broadcast_pipe<1, true>;
  return emu ? (const void*)k_broadcast_pipe<kMaxRanks, false> : (const void*)k_broadcast_pipe<1, false>;
}

const void* relay_broadcast_kernel(bool emu) {
  return emu ? (const void*)k_broadcast_relay<kMaxRanks> : (const void*)k_broadcast_relay<1>;
}

size_t pipe_smem_bytes() { return kPipeSmemBytes; }

const void* mcz_allreduce_kernel(int dtype) {
  switch (dtype) {
    case NBD_F32: return (const void*)k_allreduce_mcz<float>;
    case NBD_BF16: return (const void*)k_allreduce_mcz<__nv_bfloat16>;
    case NBD_F16: return (const void*)k_allreduce_mcz<__half>;
    default: return nullptr;
  }
}
}  // namespace nbd
