// Zero-copy reduction kernels (all_reduce, reduce_scatter) for every dtype.
#include "../../include/nbdist_b200.h"
#include "nbd_kernel_table.h"
#include "nbd_kernels.cuh"

namespace nbd {
namespace {
template <typename T, bool SUM, int NCTX>
const void* pick(int zkind) {
  return zkind == Z_AR ? (const void*)k_allreduce_zc<T, SUM, NCTX> : (const void*)k_reduce_scatter_zc<T, SUM, NCTX>;
}
template <typename T>
const void* pick_t(int zkind, bool sum, bool emu) {
  if (emu) return sum ? pick<T, true, kMaxRanks>(zkind) : pick<T, false, kMaxRanks>(zkind);
  return sum ? pick<T, true, 1>(zkind) : pick<T, false, 1>(zkind);
}
}  // namespace

const void* zc_reduce_kernel(int zkind, int dtype, bool sum, bool emu) {
  switch (dtype) {
    case NBD_F32: return pick_t<float>(zkind, sum, emu);
    case NBD_BF16: return pick_t<__nv_bfloat16>(zkind, sum, emu);
    case NBD_F16: return pick_t<__half>(zkind, sum, emu);
    case NBD_F64: return pick_t<double>(zkind, sum, emu);
    case NBD_I32: return pick_t<int32_t>(zkind, sum, emu);
    case NBD_I64: return pick_t<int64_t>(zkind, sum, emu);
    case NBD_U8: return pick_t<uint8_t>(zkind, sum, emu);
    case NBD_I8: return pick_t<int8_t>(zkind, sum, emu);
    default: return nullptr;
  }
}
}  // namespace nbd
