// nbd_kernel_table.h — the reduction kernels are instantiated in separate translation units
// (nbd_k_staged_float.cu, nbd_k_staged_int.cu, nbd_k_zc.cu) so that the library builds in
// parallel; the host side (nbd_lib.cu) looks their entry points up here and launches them with
// cudaLaunchKernel / cudaLaunchCooperativeKernel.
#pragma once

#include <stddef.h>

namespace nbd {

enum Kind { K_AR1, K_AR2, K_RS, K_LL };  // staged: one-shot, two-shot, reduce_scatter, LL
enum ZKind { Z_AR, Z_RS };               // zero-copy: all_reduce, reduce_scatter

// dtype: nbd_dtype; sum: SUM/AVG specialisation; emu: emulated-rank (cooperative) variant.
// Return nullptr for dtypes the translation unit does not hold.
const void* staged_reduce_kernel_float(int kind, int dtype, bool sum, bool emu);
const void* staged_reduce_kernel_int(int kind, int dtype, bool sum, bool emu);
const void* zc_reduce_kernel(int zkind, int dtype, bool sum, bool emu);
// tile-pipeline kernels (nbd_k_pipe.cu); mc: NVSwitch multicast transport (nullptr where the
// switch cannot do the reduction); all need pipe_smem_bytes() of dynamic shared memory
const void* pipe_allreduce_kernel(int dtype, bool sum, bool emu, bool mc);
const void* pipe_broadcast_kernel(bool emu, bool mc);
const void* relay_broadcast_kernel(bool emu);  // k_broadcast_relay: scatter + relay, peer-to-peer transport
size_t pipe_smem_bytes();
// zero-copy in-switch all_reduce on multicast-pool memory (f32/bf16/f16 sums); args: Pack<1>, McPtrs, OpParams
const void* mcz_allreduce_kernel(int dtype);

}  // namespace nbd
