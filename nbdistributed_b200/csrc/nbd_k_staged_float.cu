// Staged reduction kernels (one-shot, two-shot, reduce_scatter, LL) for the floating-point dtypes.
#define NBD_TABLE_NAME staged_reduce_kernel_float
#define NBD_TABLE_TYPES                                           \
  case NBD_F32: return pick_t<float>(kind, sum, emu);             \
  case NBD_BF16: return pick_t<__nv_bfloat16>(kind, sum, emu);    \
  case NBD_F16: return pick_t<__half>(kind, sum, emu);            \
  case NBD_F64: return pick_t<double>(kind, sum, emu);
#include "nbd_k_staged.inc"
