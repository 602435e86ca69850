// nbd_kernel_table.h — the reduction kernels are instantiated in separate translation units
// (nbd_k_staged_float.cu, nbd_k_staged_int.cu, nbd_k_zc.cu) so that the library builds in
// parallel; the host side (nbd_lib.cu) looks their entry points up here and launches them with
// cudaLaunchKernel / cudaLaunchCooperativeKernel.
#pragma once

namespace nbd {

enum Kind { K_AR1, K_AR2, K_RS, K_LL };  // staged: one-shot, two-shot, reduce_scatter, LL
enum ZKind { Z_AR, Z_RS };               // zero-copy: all_reduce, reduce_scatter

// dtype: nbd_dtype; sum: SUM/AVG specialisation; emu: emulated-rank (cooperative) variant.
// Return nullptr for dtypes the translation unit does not hold.
const void* staged_reduce_kernel_float(int kind, int dtype, bool sum, bool emu);
const void* staged_reduce_kernel_int(int kind, int dtype, bool sum, bool emu);
const void* zc_reduce_kernel(int zkind, int dtype, bool sum, bool emu);

}  // namespace nbd
