// Staged reduction kernels (one-shot, two-shot, reduce_scatter, LL) for the integer dtypes.
#define NBD_TABLE_NAME staged_reduce_kernel_int
#define NBD_TABLE_TYPES                                           \
  case NBD_I32: return pick_t<int32_t>(kind, sum, emu);           \
  case NBD_I64: return pick_t<int64_t>(kind, sum, emu);           \
  case NBD_U8: return pick_t<uint8_t>(kind, sum, emu);            \
  case NBD_I8: return pick_t<int8_t>(kind, sum, emu);
#include "nbd_k_staged.inc"
