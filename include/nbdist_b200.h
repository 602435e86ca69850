/*
 * nbdist_b200.h — C ABI of the B200-native collective data plane.
 *
 * This is the drop-in boundary for the tensor data path that nbdistributed
 * worker processes drive (dist.all_reduce / broadcast / all_gather /
 * reduce_scatter / barrier issued by cell code after worker.py's
 * init_process_group).  The reference has no native code: every byte on this
 * path moves inside torch's c10d backend ("nccl" | "gloo", selected at
 * ref:src/nbdistributed/worker.py:135-151).  The entry points below are what a
 * c10d::Backend (ours: nbdistributed_b200/csrc/c10d_backend.cpp) or any other
 * host (ctypes, cgo, JNI) binds to replace that backend on one NVLink/NVSwitch
 * node.  Plain pointers and sizes only; no torch types.
 *
 * Model: one process per GPU ("rank").  Each rank owns one ARENA in its HBM
 * (signal flags + a double-buffered staging area) that every peer maps through
 * CUDA IPC.  Collectives are single kernels that read/write peer arenas
 * directly over NVLink and synchronise with per-block release/acquire flags.
 * User buffers are ordinary device pointers (they need not be registered).
 *
 * All functions return NBD_OK (0) or a negative nbd_result code;
 * nbd_error_string() and nbd_last_error() describe it.  Collectives are
 * asynchronous with respect to the host: they enqueue work on `stream`
 * (a cudaStream_t passed as void*; NULL = the legacy default stream).
 * Every rank of a communicator must issue the same sequence of collectives
 * with the same counts/dtypes/ops/roots.
 */
#ifndef NBDIST_B200_H_
#define NBDIST_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NBD_VERSION 100 /* 0.1.0 */
#define NBD_MAX_RANKS 8 /* one NVSwitch node */
#define NBD_HANDLE_BYTES 256

typedef enum {
  NBD_OK = 0,
  NBD_ERR_INVALID_ARGUMENT = -1,
  NBD_ERR_CUDA = -2,
  NBD_ERR_NO_PEER_ACCESS = -3,
  NBD_ERR_TIMEOUT = -4,   /* a peer never arrived at a flag barrier */
  NBD_ERR_ABORTED = -5,
  NBD_ERR_UNSUPPORTED = -6, /* dtype/op combination not implemented */
  NBD_ERR_INTERNAL = -7,
  NBD_ERR_MISMATCH = -8 /* ranks disagreed on the buffers of a zero-copy collective */
} nbd_result;

/* Element types.  Copy collectives (broadcast, all_gather) are byte moves and
 * accept any of them; reductions accept all but NBD_BYTES. */
typedef enum {
  NBD_F32 = 0,
  NBD_BF16 = 1,
  NBD_F16 = 2,
  NBD_F64 = 3,
  NBD_I32 = 4,
  NBD_I64 = 5,
  NBD_U8 = 6,
  NBD_I8 = 7,
  NBD_BYTES = 8 /* opaque 1-byte elements, copy collectives only */
} nbd_dtype;

/* Reduction operators; numbering follows c10d::ReduceOp::RedOpType
 * ($TORCH/include/torch/csrc/distributed/c10d/Types.hpp:37-48). */
typedef enum {
  NBD_SUM = 0,
  NBD_AVG = 1, /* floating point only (gloo rejects it; NCCL semantics) */
  NBD_PROD = 2,
  NBD_MIN = 3,
  NBD_MAX = 4,
  NBD_BAND = 5, /* integer only */
  NBD_BOR = 6,  /* integer only */
  NBD_BXOR = 7  /* integer only */
} nbd_redop;

typedef struct nbd_comm* nbd_comm_t;

/* ---- library ------------------------------------------------------------ */

/* Version of the loaded library (NBD_VERSION it was built with). */
int nbd_version(void);
/* Static description of a result code. */
const char* nbd_error_string(int result);
/* Thread-local detail of the last failure in this thread ("" if none). */
const char* nbd_last_error(void);
/* Size in bytes of one element of `dtype`, 0 if unknown. */
size_t nbd_dtype_size(int dtype);

/* ---- communicator: one process per GPU ----------------------------------
 * Replaces the transport set-up torch does inside init_process_group
 * (ref:src/nbdistributed/worker.py:151).  Bootstrap is three steps so the
 * caller can use whatever rendezvous it already has (the c10d Store that
 * worker.py's init_process_group creates):
 *   1. nbd_comm_create       allocate the arena on `device`
 *   2. nbd_comm_get_handle   serialise this rank's IPC handle (NBD_HANDLE_BYTES)
 *      ... caller all-gathers the handles over its store ...
 *   3. nbd_comm_connect      map every peer's arena
 * stage_bytes = size of the staging area (0 = default, 256 MiB; rounded up to
 * 2 MiB).  Larger messages are pipelined through it in chunks. */
int nbd_comm_create(nbd_comm_t* comm, int rank, int world_size, int device,
                    size_t stage_bytes);
int nbd_comm_get_handle(nbd_comm_t comm, void* handle_out /*NBD_HANDLE_BYTES*/);
/* all_handles: world_size consecutive handles, index = rank. */
int nbd_comm_connect(nbd_comm_t comm, const void* all_handles);
/* Replaces the transport teardown of dist.destroy_process_group
 * (ref:src/nbdistributed/worker.py:578).  Never blocks on peers. */
int nbd_comm_destroy(nbd_comm_t comm);
/* Ask in-flight kernels of this rank to leave their flag waits
 * (c10d::Backend::abort).  The communicator is unusable afterwards. */
int nbd_comm_abort(nbd_comm_t comm);
/* NBD_OK, or the asynchronous error a kernel reported (timeout/abort). */
int nbd_comm_async_error(nbd_comm_t comm);
int nbd_comm_rank(nbd_comm_t comm);
int nbd_comm_world_size(nbd_comm_t comm);
/* Flag-wait watchdog in milliseconds (default 600000; 0 = wait forever). */
int nbd_comm_set_timeout_ms(nbd_comm_t comm, uint64_t timeout_ms);
/* Tuning knobs: "blocks" (max CTAs per collective), "ar1_max_bytes"
 * (one-shot/two-shot all-reduce switch), "bcast_direct_max_bytes",
 * "chunk_bytes" (pipeline chunk), "zero_copy_off" (1 = always stage),
 * "zc_min_bytes" (smallest message that goes zero-copy), "nvls_reduce" (1 = float
 * sums may be reduced in the NVSwitch), "nvls_off", "nvls_min_bytes", "pipe" (1 = use
 * the tile-pipeline kernels over peer-to-peer transport), "pipe_tile_bytes".  value 0
 * restores the default. */
int nbd_comm_set_tunable(nbd_comm_t comm, const char* name, uint64_t value);
int nbd_comm_get_tunable(nbd_comm_t comm, const char* name, uint64_t* value);

/* NVSwitch multicast ("NVLS").  nbd_comm_connect also tries to bind a staging buffer of every
 * rank (NBDIST_B200_NVLS_BYTES, default 128 MiB) into one multicast object, passing the handle
 * between the processes over a unix socket that rank 0 names in its NBD handle; no extra
 * rendezvous is asked of the caller.  If any rank cannot (no NVSwitch, driver without
 * multicast, NBDIST_B200_NVLS=0) all ranks agree to run without it.  With it:
 *   - nbd_broadcast of >= 512 KiB writes each tile once (multimem.st; the switch replicates it);
 *   - nbd_all_reduce of f32/bf16/f16 SUM/AVG of >= 1 MiB is reduced INSIDE the switch
 *     (multimem.ld_reduce) when the tunable "nvls_reduce" is 1 (or NBDIST_B200_NVLS_REDUCE=1)
 *     and the group has more than four ranks (2 forces it for smaller groups as well, where
 *     the peer-to-peer kernels measured faster); while it is on, nbd_mem_alloc hands out memory from a
 *     pool bound into the multicast object (NBDIST_B200_NVLS_POOL_BYTES, default 1 GiB) and
 *     all_reduce on such memory runs without any staging.
 *     Opt-in because the switch adds in an order of its own: results are identical on all ranks
 *     and within the float tolerance of the rank-order result, but not bit-identical with it.
 * Returns 1 if multicast is set up on this communicator, 0 if not (`why`, if given, receives
 * a one-line reason), negative on error.  Replaces nothing in the reference; it is what
 * backend "nccl" (ref:src/nbdistributed/worker.py:145) uses on the same hardware. */
int nbd_comm_nvls_status(nbd_comm_t comm, char* why, size_t why_bytes);

/* ---- symmetric memory (zero-copy collectives) --------------------------------
 * Replaces nothing in the reference; it is the backend side of c10d's
 * Backend::allocateTensor / supportsTensorAlloc ($TORCH/.../Backend.hpp:492-503), which FSDP
 * uses for its all_gather / reduce_scatter buffers.  A buffer obtained here is mapped by every
 * peer, so collectives on it skip the staging arena: barrier, one streaming pass over peer
 * memory, barrier.  Allocation is collective: every rank calls nbd_mem_alloc in the same
 * order with the same size, all-gathers the handles over its rendezvous, then calls
 * nbd_mem_connect.  Ranks must pass the same offset within the same allocation to a
 * collective (checked on the device; a mismatch surfaces as NBD_ERR_MISMATCH).  Collectives on
 * ordinary (unregistered) pointers keep using the staged kernels. */
int nbd_mem_alloc(nbd_comm_t comm, size_t bytes, void** ptr, void* handle_out /*NBD_HANDLE_BYTES*/);
int nbd_mem_connect(nbd_comm_t comm, void* ptr, const void* all_handles);
int nbd_mem_free(nbd_comm_t comm, void* ptr);

/* ---- emulated ranks: R ranks on ONE device, one process -------------------
 * The same device code as above, with every "peer" arena living on the same
 * GPU and all R ranks executed by one cooperative launch (grid.y = rank).
 * This is how the single-GPU parity tests and the 1-GPU bench row (local
 * copy / k-way reduce against HBM) run; it has no counterpart in the
 * reference.  comms_out receives world_size communicators (index = rank)
 * that must only be used with the nbd_emu_* calls below and destroyed with
 * nbd_emu_destroy(comms_out[0]). */
int nbd_emu_create(nbd_comm_t* comms_out, int world_size, int device,
                   size_t stage_bytes);
int nbd_emu_destroy(nbd_comm_t any_rank_comm);

/* ---- collectives ---------------------------------------------------------
 * count is in elements of `dtype` unless stated otherwise. */

/* dist.all_reduce(tensor, op)  — $TORCH/distributed/distributed_c10d.py:2979
 * -> Backend::allreduce (Backend.hpp:148); called from cell code
 * (ref:README.md:109).  recv[i] = op over ranks 0..N-1 (ascending, fixed
 * order; bf16/f16 accumulate in fp32 and round once).  send may equal recv. */
int nbd_all_reduce(nbd_comm_t comm, const void* send, void* recv, size_t count,
                   int dtype, int op, void* stream);

/* dist.broadcast(tensor, src) — distributed_c10d.py:2927 -> Backend::broadcast
 * (Backend.hpp:140); ref:README.md:122-123.  In place: root's buf is read,
 * every other rank's buf is overwritten. */
int nbd_broadcast(nbd_comm_t comm, void* buf, size_t count, int dtype, int root,
                  void* stream);

/* dist.all_gather_into_tensor(out, in) — distributed_c10d.py:4091 ->
 * Backend::_allgather_base (Backend.hpp:200).  recv holds world_size*send_count
 * elements; recv[r*send_count ...] = rank r's send.  send may alias its own
 * slot of recv. */
int nbd_all_gather(nbd_comm_t comm, const void* send, void* recv,
                   size_t send_count, int dtype, void* stream);

/* dist.reduce_scatter_tensor(out, in, op) — distributed_c10d.py:4563 ->
 * Backend::_reduce_scatter_base (Backend.hpp:269).  send holds
 * world_size*recv_count elements; recv = op over ranks of send[rank slot]. */
int nbd_reduce_scatter(nbd_comm_t comm, const void* send, void* recv,
                       size_t recv_count, int dtype, int op, void* stream);
/* dist.reduce_scatter(out, [in_0 .. in_{N-1}], op) — distributed_c10d.py reduce_scatter ->
 * Backend::reduce_scatter (Backend.hpp:259): the N contributions are separate buffers of
 * recv_count elements, sends[p] going into rank p's result.  Same result as nbd_reduce_scatter
 * on their concatenation; nothing is concatenated. */
int nbd_reduce_scatter_list(nbd_comm_t comm, const void* const* sends, void* recv,
                            size_t recv_count, int dtype, int op, void* stream);

/* dist.barrier() — ref:src/nbdistributed/worker.py:214 (from %sync) ->
 * Backend::barrier (Backend.hpp:382).  Enqueues a flag barrier on `stream`;
 * the caller synchronises the stream if it wants host-side semantics. */
int nbd_barrier(nbd_comm_t comm, void* stream);

/* Emulated-rank forms: one call runs the collective for ALL ranks of an
 * nbd_emu_create() group.  sends/recvs/bufs are arrays of world_size device
 * pointers (index = rank). */
int nbd_emu_all_reduce(nbd_comm_t any, const void* const* sends,
                       void* const* recvs, size_t count, int dtype, int op,
                       void* stream);
int nbd_emu_broadcast(nbd_comm_t any, void* const* bufs, size_t count,
                      int dtype, int root, void* stream);
int nbd_emu_all_gather(nbd_comm_t any, const void* const* sends,
                       void* const* recvs, size_t send_count, int dtype,
                       void* stream);
int nbd_emu_reduce_scatter(nbd_comm_t any, const void* const* sends,
                           void* const* recvs, size_t recv_count, int dtype,
                           int op, void* stream);
/* sends[r * world_size + p] = rank r's contribution to rank p's result */
int nbd_emu_reduce_scatter_list(nbd_comm_t any, const void* const* sends,
                                void* const* recvs, size_t recv_count, int dtype,
                                int op, void* stream);
int nbd_emu_barrier(nbd_comm_t any, void* stream);

/* ---- many broadcasts at once ---------------------------------------------------------
 * The reference's broadcast idiom is one dist.broadcast per parameter (ref:README.md:122-123);
 * issued inside a c10d coalescing block (Backend::startCoalescing / endCoalescing,
 * $TORCH/.../Backend.hpp:109-123) they arrive here as ONE call.  Every rank passes the same list
 * of (bytes, root) in the same order.  The list is taken 256 items at a time:
 *   - items that share a root and exceed 256 KiB together run as one flat table: the packed
 *     list moves as one chunked scatter pipeline (any total size, one launch);
 *   - otherwise (small, or mixed roots): every rank packs the items it is root of into its peers'
 *     staging, one flag barrier, every rank unpacks the rest; an item too large to share such a
 *     table runs as an ordinary nbd_broadcast.
 * Result: bit-identical to nitems calls of nbd_broadcast in list order. */
typedef struct {
  void* buf; /* device pointer: read on the root, written elsewhere */
  size_t bytes;
  int root;
} nbd_bcast_item;
int nbd_broadcast_table(nbd_comm_t comm, const nbd_bcast_item* items, int nitems, void* stream);
/* Emulated ranks: items[r] = rank r's list (same bytes/root per index, its own buffers). */
int nbd_emu_broadcast_table(nbd_comm_t any, const nbd_bcast_item* const* items, int nitems, void* stream);

/* ---- point to point -------------------------------------------------------------
 * dist.send / dist.recv / dist.batch_isend_irecv ($TORCH/distributed/distributed_c10d.py ->
 * Backend::send / recv, Backend.hpp:355,364) and the primitive under all_to_all, gather and
 * scatter (Backend.hpp:299-353): calls cell code can make on the process group of
 * ref:src/nbdistributed/worker.py:151.  One call takes a BATCH of this rank's sends and
 * receives and runs them concurrently in one launch (per peer and direction in issue order),
 * so matching batches on two ranks never deadlock on their relative order.  The matching call
 * on the peer must post the same message sizes in the same per-pair order.  Messages travel
 * through the receiver's arena (1 MiB chunks, 4 in flight per ordered pair); bytes == 0 is a
 * no-op on both sides; peer must differ from this rank (copy locally instead). */
#define NBD_P2P_SEND 0
#define NBD_P2P_RECV 1
typedef struct {
  int kind; /* NBD_P2P_SEND | NBD_P2P_RECV */
  int peer;
  void* buf; /* device pointer: read for a send, written for a receive */
  size_t bytes;
} nbd_p2p_op;
int nbd_p2p_batch(nbd_comm_t comm, const nbd_p2p_op* ops, int nops, void* stream);
int nbd_send(nbd_comm_t comm, const void* buf, size_t bytes, int peer, void* stream);
int nbd_recv(nbd_comm_t comm, void* buf, size_t bytes, int peer, void* stream);
/* Emulated ranks: ops[r] / nops[r] = rank r's batch; all ranks run in one cooperative launch. */
int nbd_emu_p2p_batch(nbd_comm_t any, const nbd_p2p_op* const* ops, const int* nops, void* stream);

/* ---- host buffers -------------------------------------------------------------
 * all_reduce of data that lives in HOST memory (pinned memory recommended): the call a host
 * makes when its tensors are on the CPU side, e.g. the reference's gloo users
 * (ref:src/nbdistributed/worker.py:146-147) moving to GPUs without changing where their data
 * lives.  The library pipelines H2D copy, collective and D2H copy in "host_chunk_bytes"
 * (tunable, default 32 MiB) chunks over three internal streams and a ring of device bounce
 * buffers (allocated on first use), so PCIe runs in both directions at once.  Ordered after
 * prior work on `stream`; `stream` waits for the result.  host_send may equal host_recv. */
int nbd_all_reduce_host(nbd_comm_t comm, const void* host_send, void* host_recv, size_t count,
                        int dtype, int op, void* stream);
int nbd_emu_all_reduce_host(nbd_comm_t any, const void* const* host_sends, void* const* host_recvs,
                            size_t count, int dtype, int op, void* stream);

/* ---- allocation-free variant ------------------------------------------------
 * No counterpart in the reference.
 * The all-reduce of nbd_emu_all_reduce (world_size buffers that all live on the CURRENT
 * device, reduced in ascending rank order, every recvs[r] receiving the result) for callers
 * that cannot allocate device memory at the time of the call — e.g. a process whose HBM
 * is already owned by another allocator, or a latency-critical loop that cannot take the
 * device-wide stall of cudaMalloc/cudaFree.  It needs no communicator: flags, staging and
 * clocks live in a fixed array inside the library (about 8.5 MiB of device memory on every
 * GPU that loads the library, held until the process exits) and the call performs no
 * cudaMalloc/cudaFree or any other allocating routine.
 *  - at most NBD_NOALLOC_MAX_BYTES per rank; larger messages are rejected with
 *    NBD_ERR_INVALID_ARGUMENT (use nbd_emu_all_reduce);
 *  - the first call on a device looks up the array's device address; make that first call
 *    before starting any stream capture on that device;
 *  - all calls share the one array: order them on the GPU (same stream, or events).  Calls
 *    whose device work overlaps race on it and produce wrong results;
 *  - the first call on a device also makes the CUDA runtime load this library's module and the
 *    kernel (driver-side memory the caller does not see): treat it as a warm-up;
 *  - there is no communicator, hence no asynchronous error channel: the flag-wait watchdog (10 s)
 *    of this entry point writes its verdict into the scratch array, which the host never reads.
 *    All ranks run inside ONE cooperative launch, so a wait can only time out if the GPU itself
 *    is wedged; the call's return value reports argument and launch errors only;
 *  - sends/recvs must be device pointers of the CURRENT device (not checked). */
#define NBD_NOALLOC_MAX_BYTES (1u << 20)
int nbd_local_all_reduce_noalloc(const void* const* sends, void* const* recvs, int world_size,
                                 size_t count, int dtype, int op, void* stream);

/* ---- introspection for benches/tests ------------------------------------- */

/* Number of kernels this library has launched in this process so far. */
uint64_t nbd_launch_count(void);
/* Ordinary (unregistered) buffers of at least "auto_min_bytes" (tunable, default 8 MiB) are
 * registered on the fly: the allocation is exported through CUDA IPC on first use, peers import
 * it lazily, and a one-CTA kernel agrees per operation whether every rank can address every
 * rank's buffers; if so the zero-copy kernel runs, otherwise the staged one (both are
 * enqueued, the decision gates them).  Returns 1 if the most recent such collective ran
 * zero-copy, 0 if staged.  Synchronises the device: diagnostics and tests only.
 * "auto_register_off" = 1 (or NBDIST_B200_AUTO_REGISTER=0) disables the mechanism. */
int nbd_comm_last_zc_decision(nbd_comm_t comm);
/* What the planner would choose for a collective, computed on the host without touching a GPU
 * (tests of the host logic).  kind: 0 all_reduce, 1 reduce_scatter, 2 all_gather (bytes = one
 * rank's shard for 1 and 2), 3 broadcast, 4 flat broadcast table (bytes = packed size of the
 * table), 5 tile pipeline (chunk_bytes = tile bytes, nchunks = rounds per CTA, algorithm = slots
 * per CTA).  out[6] = {algorithm, chunk_bytes, nchunks, CTAs, bytes per staging half, offset of
 * the LL area}.  stage_bytes 0 = default. */
int nbd_plan_query(int world_size, size_t stage_bytes, int num_sms, int emulated, int kind, size_t bytes,
                   uint64_t* out);
/* Name of the algorithm the last collective on `comm` used
 * ("ar_oneshot", "ar_twoshot", "bcast_direct", "bcast_scatter", ...). */
const char* nbd_last_algo(nbd_comm_t comm);

#ifdef __cplusplus
}
#endif
#endif /* NBDIST_B200_H_ */
