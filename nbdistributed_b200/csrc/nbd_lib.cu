// nbd_lib.cu — host side of the C ABI in include/nbdist_b200.h: arena
// allocation, CUDA-IPC peer mapping, algorithm/grid/chunk selection, kernel
// launches for real ranks (one process per GPU) and emulated ranks (all ranks
// in one cooperative launch on one GPU).
//
// Replaces, for one NVLink/NVSwitch node, the transport that torch's "nccl"
// backend provides behind ref:src/nbdistributed/worker.py:151.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "nbd_host.h"
#include "nbd_kernel_table.h"
#include "nbd_kernels.cuh"
#include "nbd_pipe.cuh"
#include "nbd_p2p.cuh"
#include "nbd_table.cuh"

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------
static thread_local char g_last_error[512] = "";
namespace nbd {
std::atomic<uint64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return code;
}
}  // namespace nbd

static uint64_t host_id() {
  char name[256] = {0};
  gethostname(name, sizeof(name) - 1);
  uint64_t h = 1469598103934665603ull;
  for (const char* p = name; *p; ++p) h = (h ^ (unsigned char)*p) * 1099511628211ull;
  // boot id distinguishes containers that share a hostname but not /dev/shm
  FILE* f = fopen("/proc/sys/kernel/random/boot_id", "r");
  if (f) {
    char b[64] = {0};
    if (fgets(b, sizeof(b), f))
      for (const char* p = b; *p; ++p) h = (h ^ (unsigned char)*p) * 1099511628211ull;
    fclose(f);
  }
  return h;
}

static const size_t kDefaultStage = 256ull << 20;
static const size_t kStageAlign = 2ull << 20;

static int alloc_rank(nbd_comm* c, size_t stage_bytes) {
  if (stage_bytes == 0) {
    const char* e = getenv("NBDIST_B200_STAGE_BYTES");
    stage_bytes = e ? strtoull(e, nullptr, 0) : kDefaultStage;
    if (stage_bytes == 0) stage_bytes = kDefaultStage;
  }
  stage_bytes = (stage_bytes + kStageAlign - 1) / kStageAlign * kStageAlign;
  c->stage_bytes = stage_bytes;
  c->arena_bytes = kFlagBytes + stage_bytes + kP2PAreaBytes;  // flags | staging (halves + LL) | point-to-point area
  cudaDeviceProp prop;
  NBD_CUDA(cudaGetDeviceProperties(&prop, c->device));
  c->num_sms = prop.multiProcessorCount;
  NBD_CUDA(cudaMalloc((void**)&c->arena, c->arena_bytes));
  NBD_CUDA(cudaMemset(c->arena, 0, kFlagBytes));
  NBD_CUDA(cudaMemset(c->arena + kFlagBytes + stage_bytes - ll_area_of(stage_bytes), 0, ll_area_of(stage_bytes)));  // LL flags start at 0
  NBD_CUDA(cudaMalloc((void**)&c->state, sizeof(LocalBlock)));  // LocalState is its first member
  NBD_CUDA(cudaMemset(c->state, 0, sizeof(LocalBlock)));
  if (!c->emu) {
    NBD_CUDA(cudaHostAlloc((void**)&c->export_mirror, sizeof(ExportSlot) * kMaxExports, cudaHostAllocPortable));
    memset(c->export_mirror, 0, sizeof(ExportSlot) * kMaxExports);
    NBD_CUDA(cudaHostAlloc((void**)&c->import_tmp, sizeof(ExportSlot), cudaHostAllocPortable));
    NBD_CUDA(cudaHostAlloc((void**)&c->imp_tmp, sizeof(uint64_t) * 64, cudaHostAllocPortable));
    NBD_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
  }
  NBD_CUDA(cudaHostAlloc((void**)&c->mbox, sizeof(HostMailbox), cudaHostAllocMapped | cudaHostAllocPortable));
  memset(c->mbox, 0, sizeof(HostMailbox));
  NBD_CUDA(cudaHostGetDevicePointer((void**)&c->mbox_dev, c->mbox, 0));
  NBD_CUDA(cudaDeviceSynchronize());
  const char* t = getenv("NBDIST_B200_TIMEOUT_MS");
  if (t) c->timeout_ms = strtoull(t, nullptr, 0);
  const char* b = getenv("NBDIST_B200_BLOCKS");
  if (b) c->t_blocks = strtoull(b, nullptr, 0);
  const char* ck = getenv("NBDIST_B200_CHUNK_BYTES");
  if (ck) c->t_chunk = strtoull(ck, nullptr, 0);
  const char* a1 = getenv("NBDIST_B200_AR1_MAX_BYTES");
  if (a1) c->t_ar1_max = strtoull(a1, nullptr, 0);
  const char* bd = getenv("NBDIST_B200_BCAST_DIRECT_MAX_BYTES");
  if (bd) c->t_bcast_direct_max = strtoull(bd, nullptr, 0);
  const char* ao = getenv("NBDIST_B200_AUTO_REGISTER");
  if (ao && ao[0] == '0') c->t_auto_off = 1;
  const char* am = getenv("NBDIST_B200_AUTO_MIN_BYTES");
  if (am) c->t_auto_min = strtoull(am, nullptr, 0);
  const char* nr = getenv("NBDIST_B200_NVLS_REDUCE");
  if (nr && nr[0] != '0') c->t_nvls_reduce = 1;
  return NBD_OK;
}

static void fill_ctx(nbd_comm* c) {
  RankCtx& x = c->ctx;
  x.rank = c->rank;
  x.world = c->world;
  for (int p = 0; p < kMaxRanks; ++p) {
    char* base = (char*)c->peer_base[p];
    x.flags[p] = base ? (u64*)base : nullptr;
    x.stage[p] = base ? base + kFlagBytes : nullptr;
  }
  x.state = c->state;
  x.mbox = c->mbox_dev;
  // the tail of the staging (1 or 8 MiB) is the LL area; the two pipeline halves share the rest
  x.ll_off = c->stage_bytes - ll_area_of(c->stage_bytes);
  x.ll_slot = 2 * ll_cap_of(c->stage_bytes);
  x.p2p_off = c->stage_bytes;
  x.half_bytes = ((c->stage_bytes - ll_area_of(c->stage_bytes)) / 2) / kUnit * kUnit;
  x.timeout_ns = c->timeout_ms * 1000000ull;
}

static void hostpipe_free(HostPipe& hp) {
  if (!hp.ready) return;
  for (int r = 0; r < kMaxRanks; ++r)
    for (int k = 0; k < kPipeSlots; ++k)
      if (hp.buf[r][k]) cudaFree(hp.buf[r][k]);
  for (int k = 0; k < kPipeSlots; ++k) {
    cudaEventDestroy(hp.ev_h2d[k]);
    cudaEventDestroy(hp.ev_comp[k]);
    cudaEventDestroy(hp.ev_d2h[k]);
  }
  cudaEventDestroy(hp.ev_start);
  cudaStreamDestroy(hp.h2d);
  cudaStreamDestroy(hp.comp);
  cudaStreamDestroy(hp.d2h);
  hp = HostPipe();
}

static void free_rank(nbd_comm* c) {
  if (c->mbox) c->mbox->abort = 1;  // let any stuck flag wait leave before cudaFree's implicit sync
  if (!c->emu) {
    for (int p = 0; p < c->world; ++p)
      if (p != c->rank && c->peer_base[p]) cudaIpcCloseMemHandle(c->peer_base[p]);
  }
  hostpipe_free(c->hp);
  nvls_teardown(c);
  for (auto& g : c->regs) {
    if (!c->emu)
      for (int p = 0; p < c->world; ++p)
        if (p != c->rank && g.peer[p]) cudaIpcCloseMemHandle(g.peer[p]);
    if (g.base && !g.mc) cudaFree(g.base);
  }
  c->regs.clear();
  for (void* m : c->import_ptrs) cudaIpcCloseMemHandle(m);
  c->import_ptrs.clear();
  if (c->side) cudaStreamDestroy(c->side);
  if (c->export_mirror) cudaFreeHost(c->export_mirror);
  if (c->import_tmp) cudaFreeHost(c->import_tmp);
  if (c->imp_tmp) cudaFreeHost(c->imp_tmp);
  if (c->arena) cudaFree(c->arena);
  if (c->state) cudaFree(c->state);
  if (c->mbox) cudaFreeHost(c->mbox);
}

extern "C" {

int nbd_version(void) { return NBD_VERSION; }

const char* nbd_error_string(int r) {
  switch (r) {
    case NBD_OK: return "ok";
    case NBD_ERR_INVALID_ARGUMENT: return "invalid argument";
    case NBD_ERR_CUDA: return "CUDA runtime error";
    case NBD_ERR_NO_PEER_ACCESS: return "peer GPU is not reachable over NVLink/PCIe P2P";
    case NBD_ERR_TIMEOUT: return "timed out waiting for a peer at a flag barrier";
    case NBD_ERR_ABORTED: return "communicator aborted";
    case NBD_ERR_UNSUPPORTED: return "unsupported dtype/op combination";
    case NBD_ERR_INTERNAL: return "internal error";
    case NBD_ERR_MISMATCH: return "ranks passed different registered buffers/offsets/sizes to a zero-copy collective";
    default: return "unknown error";
  }
}

const char* nbd_last_error(void) { return g_last_error; }

size_t nbd_dtype_size(int dtype) {
  switch (dtype) {
    case NBD_F32: case NBD_I32: return 4;
    case NBD_BF16: case NBD_F16: return 2;
    case NBD_F64: case NBD_I64: return 8;
    case NBD_U8: case NBD_I8: case NBD_BYTES: return 1;
    default: return 0;
  }
}

uint64_t nbd_launch_count(void) { return g_launches.load(); }

int nbd_comm_create(nbd_comm_t* out, int rank, int world, int device, size_t stage_bytes) {
  if (!out) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (world < 1 || world > kMaxRanks) return fail(NBD_ERR_INVALID_ARGUMENT, "world_size %d not in [1,%d]", world, kMaxRanks);
  if (rank < 0 || rank >= world) return fail(NBD_ERR_INVALID_ARGUMENT, "rank %d not in [0,%d)", rank, world);
  int ndev = 0;
  NBD_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail(NBD_ERR_INVALID_ARGUMENT, "device %d not in [0,%d)", device, ndev);
  DeviceGuard g(device);
  nbd_comm* c = new nbd_comm();
  c->rank = rank;
  c->world = world;
  c->device = device;
  int rc = alloc_rank(c, stage_bytes);
  if (rc != NBD_OK) {
    free_rank(c);
    delete c;
    return rc;
  }
  c->peer_base[rank] = c->arena;
  if (world == 1) {
    fill_ctx(c);
    c->connected = true;
  }
  nvls_prepare(c);  // rank 0 of a multi-rank group opens the multicast rendezvous socket
  *out = c;
  return NBD_OK;
}

int nbd_comm_get_handle(nbd_comm_t c, void* handle_out) {
  if (!c || !handle_out) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL argument");
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "emulated communicators have no IPC handle");
  DeviceGuard g(c->device);
  Handle h;
  memset(&h, 0, sizeof(h));
  h.magic = kMagic;
  h.version = NBD_VERSION;
  h.rank = c->rank;
  h.world = c->world;
  h.pid = (int64_t)getpid();
  h.hostid = host_id();
  h.arena_bytes = c->arena_bytes;
  h.stage_bytes = c->stage_bytes;
  NBD_CUDA(cudaDeviceGetPCIBusId(h.pci, sizeof(h.pci), c->device));
  NBD_CUDA(cudaIpcGetMemHandle(&h.ipc, c->arena));
  h.nvls_nonce = c->nvls.nonce;
  memset(handle_out, 0, NBD_HANDLE_BYTES);
  memcpy(handle_out, &h, sizeof(h));
  return NBD_OK;
}

int nbd_comm_connect(nbd_comm_t c, const void* all_handles) {
  if (!c || !all_handles) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL argument");
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "emulated communicators are connected at creation");
  if (c->connected) return NBD_OK;
  DeviceGuard g(c->device);
  const uint64_t me = host_id();
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank) continue;
    Handle h;
    memcpy(&h, (const char*)all_handles + (size_t)p * NBD_HANDLE_BYTES, sizeof(h));
    if (h.magic != kMagic || h.version != NBD_VERSION)
      return fail(NBD_ERR_INVALID_ARGUMENT, "handle %d: bad magic/version (%#x/%u)", p, h.magic, h.version);
    if (h.rank != p || h.world != c->world)
      return fail(NBD_ERR_INVALID_ARGUMENT, "handle %d describes rank %d of %d", p, h.rank, h.world);
    if (h.hostid != me)
      return fail(NBD_ERR_NO_PEER_ACCESS, "rank %d is on another host; this backend is single-node (NVLink/NVSwitch)", p);
    if (h.stage_bytes != c->stage_bytes)
      return fail(NBD_ERR_INVALID_ARGUMENT, "rank %d has stage_bytes %llu, this rank %llu", p,
                  (unsigned long long)h.stage_bytes, (unsigned long long)c->stage_bytes);
    if (h.pid == (int64_t)getpid())
      return fail(NBD_ERR_INVALID_ARGUMENT, "rank %d lives in this process; use one process per GPU (or nbd_emu_*)", p);
    int pdev = -1;
    if (cudaDeviceGetByPCIBusId(&pdev, h.pci) == cudaSuccess && pdev >= 0) {
      if (pdev == c->device)
        return fail(NBD_ERR_INVALID_ARGUMENT, "rank %d shares GPU %s with this rank; ranks must own distinct GPUs", p, h.pci);
      int can = 0;
      NBD_CUDA(cudaDeviceCanAccessPeer(&can, c->device, pdev));
      if (!can) return fail(NBD_ERR_NO_PEER_ACCESS, "GPU %d cannot access peer GPU %d (%s)", c->device, pdev, h.pci);
    } else {
      cudaGetLastError();  // peer not visible in this process (CUDA_VISIBLE_DEVICES); let IPC decide
    }
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h.ipc, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(NBD_ERR_NO_PEER_ACCESS, "cudaIpcOpenMemHandle(rank %d, %s): %s", p, h.pci, cudaGetErrorString(e));
    }
    c->peer_base[p] = ptr;
  }
  fill_ctx(c);
  {
    // NVSwitch multicast staging (nbd_nvls.cu): collective among the ranks, never fatal
    Handle h0;
    memcpy(&h0, (const char*)all_handles, sizeof(h0));
    nvls_setup(c, c->rank == 0 ? c->nvls.nonce : h0.nvls_nonce);
  }
  c->connected = true;
  return NBD_OK;
}

int nbd_comm_destroy(nbd_comm_t c) {
  if (!c) return NBD_OK;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_destroy for emulated communicators");
  DeviceGuard g(c->device);
  free_rank(c);
  delete c;
  return NBD_OK;
}

int nbd_comm_abort(nbd_comm_t c) {
  if (!c) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  c->mbox->abort = 1;
  c->dead = true;
  return NBD_OK;
}

int nbd_comm_async_error(nbd_comm_t c) {
  if (!c) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  const uint32_t e = ((volatile HostMailbox*)c->mbox)->error;
  if (e == 0) return NBD_OK;
  c->dead = true;
  return fail(-(int)e, "rank %d: a collective kernel gave up at a flag barrier (%s)", c->rank, nbd_error_string(-(int)e));
}

int nbd_comm_rank(nbd_comm_t c) { return c ? c->rank : -1; }
int nbd_comm_world_size(nbd_comm_t c) { return c ? c->world : -1; }

int nbd_comm_set_timeout_ms(nbd_comm_t c, uint64_t ms) {
  if (!c) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  c->timeout_ms = ms;
  c->ctx.timeout_ns = ms * 1000000ull;
  return NBD_OK;
}

static uint64_t* tunable_slot(nbd_comm* c, const char* name) {
  if (!strcmp(name, "blocks")) return &c->t_blocks;
  if (!strcmp(name, "ar1_max_bytes")) return &c->t_ar1_max;
  if (!strcmp(name, "bcast_direct_max_bytes")) return &c->t_bcast_direct_max;
  if (!strcmp(name, "chunk_bytes")) return &c->t_chunk;
  if (!strcmp(name, "zero_copy_off")) return &c->t_zc_off;
  if (!strcmp(name, "zc_min_bytes")) return &c->t_zc_min;
  if (!strcmp(name, "zc_ar_mode")) return &c->t_zc_fused;
  if (!strcmp(name, "zc_tile_bytes")) return &c->t_zc_tile;
  if (!strcmp(name, "host_chunk_bytes")) return &c->t_host_chunk;
  if (!strcmp(name, "ll_max_bytes")) return &c->t_ll_max;
  if (!strcmp(name, "auto_register_off")) return &c->t_auto_off;
  if (!strcmp(name, "auto_min_bytes")) return &c->t_auto_min;
  if (!strcmp(name, "nvls_reduce")) return &c->t_nvls_reduce;
  if (!strcmp(name, "nvls_off")) return &c->t_nvls_off;
  if (!strcmp(name, "nvls_min_bytes")) return &c->t_nvls_min;
  if (!strcmp(name, "pipe")) return &c->t_pipe;
  if (!strcmp(name, "pipe_tile_bytes")) return &c->t_pipe_tile;
  if (!strcmp(name, "debug_mute_rank")) return &c->t_debug_mute;
  if (!strcmp(name, "pipe_depth")) return &c->t_pipe_depth;
  if (!strcmp(name, "pipe_mode")) return &c->t_pipe_mode;
  if (!strcmp(name, "relay_min_bytes")) return &c->t_relay_min;
  if (!strcmp(name, "nvls_bcast")) return &c->t_nvls_bcast;
  return nullptr;
}

int nbd_comm_set_tunable(nbd_comm_t c, const char* name, uint64_t value) {
  if (!c || !name) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL argument");
  uint64_t* s = tunable_slot(c, name);
  if (!s) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown tunable '%s'", name);
  *s = value;
  return NBD_OK;
}

int nbd_comm_get_tunable(nbd_comm_t c, const char* name, uint64_t* value) {
  if (!c || !name || !value) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL argument");
  uint64_t* s = tunable_slot(c, name);
  if (!s) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown tunable '%s'", name);
  *value = *s;
  return NBD_OK;
}

const char* nbd_last_algo(nbd_comm_t c) { return c ? c->last_algo : "none"; }

// ---- emulated ranks -------------------------------------------------------

int nbd_emu_create(nbd_comm_t* out, int world, int device, size_t stage_bytes) {
  if (!out) return fail(NBD_ERR_INVALID_ARGUMENT, "comms_out is NULL");
  if (world < 1 || world > kMaxRanks) return fail(NBD_ERR_INVALID_ARGUMENT, "world_size %d not in [1,%d]", world, kMaxRanks);
  int ndev = 0;
  NBD_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail(NBD_ERR_INVALID_ARGUMENT, "device %d not in [0,%d)", device, ndev);
  DeviceGuard g(device);
  int coop = 0;
  NBD_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
  if (!coop) return fail(NBD_ERR_UNSUPPORTED, "device %d lacks cooperative launch", device);
  EmuGroup* grp = new EmuGroup();
  grp->world = world;
  int rc = NBD_OK;
  for (int r = 0; r < world && rc == NBD_OK; ++r) {
    nbd_comm* c = new nbd_comm();
    c->rank = r;
    c->world = world;
    c->device = device;
    c->emu = true;
    c->group = grp;
    grp->comms[r] = c;
    rc = alloc_rank(c, stage_bytes);
  }
  if (rc != NBD_OK) {
    for (int r = 0; r < world; ++r)
      if (grp->comms[r]) {
        free_rank(grp->comms[r]);
        delete grp->comms[r];
      }
    delete grp;
    return rc;
  }
  for (int r = 0; r < world; ++r) {
    nbd_comm* c = grp->comms[r];
    for (int p = 0; p < world; ++p) c->peer_base[p] = grp->comms[p]->arena;
    fill_ctx(c);
    c->connected = true;
    out[r] = c;
  }
  return NBD_OK;
}

int nbd_emu_destroy(nbd_comm_t any) {
  if (!any) return NBD_OK;
  if (!any->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "not an emulated communicator");
  EmuGroup* grp = any->group;
  DeviceGuard g(any->device);
  hostpipe_free(grp->hp);
  for (int r = 0; r < grp->world; ++r) {
    free_rank(grp->comms[r]);
    delete grp->comms[r];
  }
  delete grp;
  return NBD_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// planning: algorithm, chunking, grid
// ---------------------------------------------------------------------------
static inline u64 floor_to(u64 x, u64 m) { return x / m * m; }
static inline u64 ceil_div(u64 a, u64 b) { return (a + b - 1) / b; }

struct Plan {
  OpParams op;
  int nblocks;
};

// CTAs: enough that each handles >= 4 units per thread, at most the number that
// is guaranteed co-resident (flag waits need every CTA of every rank running).
// max_blocks is that bound, computed per kernel by occupancy_cap().
static int pick_blocks(const nbd_comm* c, u64 units_per_block_span, int emu_world) {
  u64 cap = (u64)c->num_sms * kMinBlocksPerSM;
  if (emu_world > 1) cap /= (u64)emu_world;
  if (c->t_blocks && c->t_blocks < cap) cap = c->t_blocks;
  if (cap > (u64)kExportRow0) cap = kExportRow0;  // rows above hold export slots and descriptors
  if (cap < 1) cap = 1;
  u64 want = ceil_div(units_per_block_span, (u64)kThreads * 4);
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

// A message that fits one chunk gets a chunk of exactly its (aligned) size: CTA spans are
// cut from chunk_bytes (see block_span), so an oversized chunk would idle most CTAs.  A larger
// message is cut into EQUAL chunks: a short ragged last chunk leaves most CTAs idle for a whole
// pipeline step (8xB200, 256 MiB: 3 chunks of 113+113+30 MiB 518 GB/s, 4 x 64 MiB 579 GB/s).
static inline u64 fit_chunk(u64 chunk, u64 bytes, u64 align) {
  if (bytes == 0) return chunk;
  const u64 n = ceil_div(bytes, chunk);
  const u64 even = ceil_div(ceil_div(bytes, n), align) * align;
  return even < chunk ? even : chunk;
}

// Default pipeline chunk: 64 MiB of message (or what the staging half holds, if less).  Every
// chunk costs one or two flag barriers: 8 MiB chunks measured 187 GB/s, 64 MiB 579 GB/s
// (8xB200, 256 MiB; 2xB200: 387 vs 555; profiles/r01_tune_n{2,8}.jsonl).  "chunk_bytes" overrides.
static const u64 kDefChunkAR = 64ull << 20;
static const u64 kDefChunkShard = 64ull << 20;
static const u64 kDefChunkBcast = 64ull << 20;

// LL threshold ("ll_max_bytes" overrides it, up to the communicator's LL limit of 32 or 256 KiB).
// Every rank sends 2(N-1) bytes per payload byte, so the hand-over to the barrier kernels comes
// earlier with more ranks.  Device-side us, LL against one-shot / direct (tools/tune_ll.py,
// profiles/r02_tune_ll_n{2,4,8}.jsonl):
//   all_reduce  64 KiB   N=2 6.5 / 10.7   N=4  9.7 / 14.1   N=8 13.3 / 19.6
//               128 KiB      7.6 / 11.3       10.7 / 15.0       16.9 / 19.9
//               256 KiB      7.4 / 10.3       13.5 / 14.8       23.5 / 20.4
//   broadcast   128 KiB      6.5 /  8.4        8.5 / 11.5       12.6 / 12.9
//               256 KiB      7.8 /  8.4       10.4 / 11.2       18.9 / 16.4
static u64 ll_default_max(const nbd_comm* c) { return c->world <= 4 ? (256ull << 10) : (128ull << 10); }

static Plan plan_all_reduce(nbd_comm* c, u64 bytes, int op, int emu_world) {
  Plan p;
  memset(&p, 0, sizeof(p));
  const int N = c->world;
  const u64 half = c->ctx.half_bytes;
  // one-shot moves N/2 times the bytes of two-shot but has one barrier: 8xB200 1 MiB 31 us vs 54 us,
  // 4 MiB 64 us vs 59 us (profiles/r01_tune_n8.jsonl); with two ranks it moves the same bytes
  const u64 ar1_max = c->t_ar1_max ? c->t_ar1_max : (N == 2 ? ~0ull : (2ull << 20));
  p.op.bytes = bytes;
  p.op.op = op;
  p.op.avg = (op == NBD_AVG);
  p.op.world = N;
  const u64 ll_max = c->t_ll_max ? c->t_ll_max : ll_default_max(c);
  if (c->ctx.ll_off != 0 && bytes <= ll_max && bytes <= c->ctx.ll_slot / 2) {
    p.op.algo = 3;  // LL one-shot: packets carry their own flags, no barrier
    p.op.chunk_bytes = bytes;
    p.op.nchunks = 1;
    u64 want = ceil_div(ceil_div(bytes, kUnit), (u64)kThreads);
    const u64 cap_real = bytes > kLLMaxBytes ? 32 : 8;  // one unit per thread up to 256 KiB
    u64 cap = emu_world > 1 ? (u64)c->num_sms / emu_world : cap_real;
    if (cap > cap_real) cap = cap_real;
    if (cap < 1) cap = 1;
    p.nblocks = (int)(want < 1 ? 1 : (want < cap ? want : cap));
    c->last_algo = "ar_ll";
    return p;
  }
  if (bytes <= ar1_max) {
    p.op.algo = 1;
    u64 cap = floor_to(half / N, kUnit);
    u64 chunk = c->t_chunk ? floor_to(c->t_chunk, kUnit) : kDefChunkAR;
    if (chunk < kUnit) chunk = kUnit;
    if (chunk > cap) chunk = cap;
    chunk = fit_chunk(chunk, bytes, kUnit);
    p.op.chunk_bytes = chunk;
    p.op.nchunks = (u32)ceil_div(bytes, chunk);
    p.nblocks = pick_blocks(c, ceil_div(chunk, kUnit), emu_world);
    c->last_algo = "ar_oneshot";
  } else {
    p.op.algo = 2;
    // (N+1) slots of ceil(units/N) units must fit in one half
    u64 cap = floor_to((half / (N + 1) - kUnit) * N, (u64)kUnit * N);
    u64 chunk = c->t_chunk ? floor_to(c->t_chunk, (u64)kUnit * N) : kDefChunkAR;
    if (chunk < (u64)kUnit * N) chunk = (u64)kUnit * N;
    if (chunk > cap) chunk = cap;
    chunk = fit_chunk(chunk, bytes, (u64)kUnit * N);
    p.op.chunk_bytes = chunk;
    p.op.nchunks = (u32)ceil_div(bytes, chunk);
    p.nblocks = pick_blocks(c, ceil_div(ceil_div(chunk, kUnit), N), emu_world);
    c->last_algo = "ar_twoshot";
  }
  return p;
}

static Plan plan_reduce_scatter(nbd_comm* c, u64 shard_bytes, int op, int emu_world) {
  Plan p;
  memset(&p, 0, sizeof(p));
  const int N = c->world;
  const u64 half = c->ctx.half_bytes;
  p.op.bytes = shard_bytes;
  p.op.op = op;
  p.op.avg = (op == NBD_AVG);
  p.op.world = N;
  u64 cap = floor_to(half / N, kUnit);
  u64 chunk = c->t_chunk ? floor_to(c->t_chunk, kUnit) : kDefChunkShard;
  if (chunk < kUnit) chunk = kUnit;
  if (chunk > cap) chunk = cap;
  chunk = fit_chunk(chunk, shard_bytes, kUnit);
  p.op.chunk_bytes = chunk;
  p.op.nchunks = (u32)ceil_div(shard_bytes, chunk);
  p.nblocks = pick_blocks(c, ceil_div(chunk, kUnit), emu_world);
  c->last_algo = "reduce_scatter_push";
  return p;
}

static Plan plan_all_gather(nbd_comm* c, u64 shard_bytes, int emu_world) {
  Plan p;
  memset(&p, 0, sizeof(p));
  const u64 half = c->ctx.half_bytes;
  p.op.bytes = shard_bytes;
  p.op.world = c->world;
  u64 cap = floor_to(half, kUnit);
  u64 chunk = c->t_chunk ? floor_to(c->t_chunk, kUnit) : kDefChunkShard;
  if (chunk < kUnit) chunk = kUnit;
  if (chunk > cap) chunk = cap;
  chunk = fit_chunk(chunk, shard_bytes, kUnit);
  p.op.chunk_bytes = chunk;
  p.op.nchunks = (u32)ceil_div(shard_bytes, chunk);
  p.nblocks = pick_blocks(c, ceil_div(chunk, kUnit), emu_world);
  c->last_algo = "all_gather_pull";
  return p;
}

static Plan plan_broadcast(nbd_comm* c, u64 bytes, int root, int emu_world) {
  Plan p;
  memset(&p, 0, sizeof(p));
  const int N = c->world;
  const u64 half = c->ctx.half_bytes;
  const u64 direct_max = c->t_bcast_direct_max ? c->t_bcast_direct_max : (N <= 2 ? ~0ull : (256ull << 10));
  p.op.bytes = bytes;
  p.op.root = root;
  p.op.world = N;
  const u64 ll_max = c->t_ll_max ? c->t_ll_max : ll_default_max(c);
  if (c->ctx.ll_off != 0 && bytes <= ll_max && bytes <= c->ctx.ll_slot / 2) {
    p.op.algo = 2;  // LL: packets carry their own flags, no barrier (4 KiB: ~half the latency of "direct")
    p.op.chunk_bytes = bytes;
    p.op.nchunks = 1;
    u64 want = ceil_div(ceil_div(bytes, kUnit), (u64)kThreads);
    const u64 cap_real = bytes > kLLMaxBytes ? 32 : 8;
    u64 cap = emu_world > 1 ? (u64)c->num_sms / emu_world : cap_real;
    if (cap > cap_real) cap = cap_real;
    if (cap < 1) cap = 1;
    p.nblocks = (int)(want < 1 ? 1 : (want < cap ? want : cap));
    c->last_algo = "bcast_ll";
    return p;
  }
  if (bytes <= direct_max) {
    p.op.algo = 0;
    u64 cap = floor_to(half, kUnit);
    u64 chunk = c->t_chunk ? floor_to(c->t_chunk, kUnit) : kDefChunkBcast;
    if (chunk < kUnit) chunk = kUnit;
    if (chunk > cap) chunk = cap;
    chunk = fit_chunk(chunk, bytes, kUnit);
    p.op.chunk_bytes = chunk;
    p.op.nchunks = (u32)ceil_div(bytes, chunk);
    p.nblocks = pick_blocks(c, ceil_div(chunk, kUnit), emu_world);
    c->last_algo = "bcast_direct";
  } else {
    p.op.algo = 1;
    const u64 nsh = (u64)(N - 1);
    u64 cap = floor_to((half - kUnit) * nsh, (u64)kUnit * nsh);
    u64 chunk = c->t_chunk ? floor_to(c->t_chunk, (u64)kUnit * nsh) : floor_to(kDefChunkBcast, (u64)kUnit * nsh);
    if (chunk < (u64)kUnit * nsh) chunk = (u64)kUnit * nsh;
    if (chunk > cap) chunk = cap;
    chunk = fit_chunk(chunk, bytes, (u64)kUnit * nsh);
    p.op.chunk_bytes = chunk;
    p.op.nchunks = (u32)ceil_div(bytes, chunk);
    p.nblocks = pick_blocks(c, ceil_div(ceil_div(chunk, kUnit), nsh), emu_world);
    c->last_algo = "bcast_scatter";
  }
  return p;
}

// Tile pipeline (nbd_pipe.cuh): G CTAs, each streaming `rounds` tiles of `tile` bytes through
// kPipeDepth staging slots.  Tiles are 16..256 KiB: small enough that copy-in, reduction and
// copy-out of neighbouring tiles overlap and a flag costs one tile of latency, large enough that
// the TMA rings and the 448 reducing threads have work.  capacity = staging bytes per rank.
static const u64 kPipeTileMin = 16ull << 10, kPipeTileMax = 1ull << 20, kPipeTileDef = 256ull << 10;

static int plan_pipe(nbd_comm* c, u64 bytes, u64 capacity, int emu_world, Plan* out) {
  Plan p;
  memset(&p, 0, sizeof(p));
  u64 gmax = (u64)c->num_sms * kMinBlocksPerSM;
  if (emu_world > 1) gmax /= (u64)emu_world;
  if (c->t_blocks && c->t_blocks < gmax) gmax = c->t_blocks;
  if (gmax > (u64)kPipeMaxBlocks) gmax = kPipeMaxBlocks;
  if (gmax < 1) gmax = 1;
  u64 G = ceil_div(bytes, kPipeTileMin);
  if (G < 1) G = 1;
  if (G > gmax) G = gmax;
  const u64 depth = c->t_pipe_depth ? (c->t_pipe_depth < 2 ? 2 : (c->t_pipe_depth > 8 ? 8 : c->t_pipe_depth)) : (u64)kPipeDepth;
  u64 tile = c->t_pipe_tile ? c->t_pipe_tile : ceil_div(bytes, G * 4);
  if (!c->t_pipe_tile) {
    if (tile < kPipeTileMin) tile = kPipeTileMin;
    if (tile > kPipeTileDef) tile = kPipeTileDef;
  }
  if (tile > kPipeTileMax) tile = kPipeTileMax;
  tile = ceil_div(tile, 128) * 128;
  const u64 cap_tile = floor_to(capacity / (G * depth), 128);
  if (tile > cap_tile) tile = cap_tile;
  if (tile < 128) return fail(NBD_ERR_INVALID_ARGUMENT, "staging of %llu bytes is too small for the tile pipeline (%llu CTAs)",
                             (unsigned long long)capacity, (unsigned long long)G);
  p.op.bytes = bytes;
  p.op.chunk_bytes = tile;
  p.op.nchunks = (u32)ceil_div(ceil_div(bytes, kUnit) * kUnit, G * tile);
  p.op.world = c->world;
  p.op.depth = (int)depth;
  p.op.algo = (int)(c->t_pipe_mode & 62);  // bits 3-5: timing diagnostics (skip copy-in / copy-out / reduction)  // callers add bit 0 (open with a barrier) for arena staging
  p.nblocks = (int)G;
  *out = p;
  return NBD_OK;
}

// The pipeline kernels need 64 KiB of dynamic shared memory (two TMA rings): opt in once per
// device and kernel.
static int pipe_launch(const void* fn, int device, int nblocks, int world, bool coop, cudaStream_t s, void* pk, OpParams* op) {
  static std::mutex mu;
  static std::set<std::pair<int, const void*>> prepared;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (!prepared.count({device, fn})) {
      NBD_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pipe_smem_bytes()));
      prepared.insert({device, fn});
    }
  }
  void* args[2] = {pk, (void*)op};
  if (!coop) {
    NBD_CUDA(cudaLaunchKernel(fn, dim3(nblocks, 1, 1), dim3(kThreads, 1, 1), args, pipe_smem_bytes(), s));
  } else {
    NBD_CUDA(cudaLaunchCooperativeKernel(fn, dim3(nblocks, world, 1), dim3(kThreads, 1, 1), args, pipe_smem_bytes(), s));
  }
  g_launches.fetch_add(1);
  return NBD_OK;
}

// In-switch reduction is opt-in ("nvls_reduce" / NBDIST_B200_NVLS_REDUCE=1): the switch adds the N
// copies in an order of its own, so float sums are not bit-identical with the ascending-rank
// order of every other kernel (they stay within the tolerance of tests/helpers.py and are
// identical on all ranks).  Broadcast over multicast is a plain copy and needs no opt-in.
// Measured at 8xB200 (profiles/r02_bench_n8_mid.json), busbw per rank in GB/s:
//   staged in-switch pipeline (ordinary buffers): 16 MiB 172, 64 MiB 465, 256 MiB 650, 1 GiB 739
//   peer-to-peer kernels, same buffers:           16 MiB 327, 64 MiB 508, 256 MiB 577-630, 1 GiB 658
//   zero-copy in-switch (multicast-pool memory):   1 MiB 76, 16 MiB 445, 64 MiB 632, 256 MiB 772
//   peer-to-peer zero-copy (symmetric memory):     1 MiB 61, 16 MiB 322, 64 MiB 565, 256 MiB 645
static const u64 kDefNvlsReduceMin = 128ull << 20;  // staged pipeline
static const u64 kDefNvlsZcMin = 256ull << 10;      // zero-copy on pool memory
static const u64 kDefNvlsBcastMin = 512ull << 10;

// In-switch float sums pay from more than four ranks on.  Wire bytes per GPU and direction are
// S(1 + 1/N) through the switch against 2(N-1)/N S peer to peer: a loss at two ranks (326 against
// 602 GB/s), and at four the smaller saving is eaten by the switch path's lower efficiency
// (4xB200, 256 MiB: staged 562 / pool 645 against 636 / 657 peer to peer; profiles/r02_bench_n4.json);
// at eight: 654-660 / 772 against 630 / 645.  "nvls_reduce" = 2 forces it at any world size.
static bool nvls_reduce_policy(const nbd_comm* c) {
  return c->nvls.enabled && !c->t_nvls_off && c->t_nvls_reduce && (c->world > 4 || c->t_nvls_reduce == 2);
}

static bool nvls_reduce_wanted(const nbd_comm* c, u64 bytes, int dtype, int op) {
  return nvls_reduce_policy(c) && (op == NBD_SUM || op == NBD_AVG) &&
         (dtype == NBD_F32 || dtype == NBD_BF16 || dtype == NBD_F16) &&
         bytes >= (c->t_nvls_min ? c->t_nvls_min : kDefNvlsZcMin);
}

// Broadcast over multimem.st is correct but, measured at 8xB200, slower than the scatter kernel at
// every size (256 MiB: 366 vs 535 GB/s; one source feeding the switch's replication), so it only
// runs on request ("nvls_bcast" = 1).
static bool nvls_bcast_wanted(const nbd_comm* c, u64 bytes) {
  return c->nvls.enabled && !c->t_nvls_off && c->t_nvls_bcast && bytes >= (c->t_nvls_min ? c->t_nvls_min : kDefNvlsBcastMin);
}

// P2P tile pipeline: only on request ("pipe" = 1) until measurements say where it wins
static bool pipe_wanted(const nbd_comm* c, u64 bytes) { return c->t_pipe == 1 && bytes > kLLMaxBytes; }

// Relay broadcast (k_broadcast_relay: the scatter algorithm with per-tile flags).  "pipe" = 3
// forces it; otherwise from "relay_min_bytes" on (0 = the default below).
static const u64 kDefRelayMin = ~0ull;
static bool relay_wanted(const nbd_comm* c, u64 bytes) {
  if (c->t_pipe == 3) return bytes > kLLMaxBytes;
  if (c->t_pipe != 0 || c->world <= 2) return false;
  return bytes >= (c->t_relay_min ? c->t_relay_min : kDefRelayMin);
}

static void fill_pipe(PackPipe<1>& pk, nbd_comm* c, const void* send, void* recv, bool mc) {
  memset(&pk, 0, sizeof(pk));
  pk.ctx[0] = c->ctx;
  pk.buf[0].send = (const char*)send;
  pk.buf[0].recv = (char*)recv;
  pk.pb[0].local = mc ? c->nvls.local : c->ctx.stage[c->rank];
  pk.pb[0].mc = mc ? c->nvls.mcva : nullptr;
}

static int pipe_all_reduce(nbd_comm* c, const void* send, void* recv, u64 bytes, int dtype, int op, bool mc, cudaStream_t s) {
  const bool sum = (op == NBD_SUM || op == NBD_AVG);
  const void* fn = pipe_allreduce_kernel(dtype, sum, false, mc);
  if (!fn) return fail(NBD_ERR_UNSUPPORTED, "dtype %d / op %d has no tile-pipeline kernel", dtype, op);
  Plan p;
  int rc = plan_pipe(c, bytes, mc ? (u64)c->nvls.stage_bytes : 2 * c->ctx.half_bytes, 0, &p);
  if (rc != NBD_OK) return rc;
  p.op.op = op;
  p.op.avg = (op == NBD_AVG);
  p.op.algo |= mc ? 0 : 1;  // arena staging is shared with the chunked kernels: open with a barrier
  if (mc && !c->t_pipe_mode) p.op.algo |= 2;  // two reducing groups: 8xB200 256 MiB 712 vs 735 us (profiles/r02_tune2_nvls_diag_n8.jsonl)
  PackPipe<1> pk;
  fill_pipe(pk, c, send, recv, mc);
  c->last_algo = mc ? "ar_nvls" : "ar_pipe";
  return pipe_launch(fn, c->device, p.nblocks, 1, false, s, &pk, &p.op);
}

// two pushing groups unless "pipe_mode" says otherwise (any non-zero value without bit 1: one group)
static int relay_mode(const nbd_comm* c) { return c->t_pipe_mode ? (int)(c->t_pipe_mode & 2) : 2; }

static int pipe_broadcast(nbd_comm* c, void* buf, u64 bytes, int root, bool mc, cudaStream_t s, bool relay = false) {
  const void* fn = relay ? relay_broadcast_kernel(false) : pipe_broadcast_kernel(false, mc);
  Plan p;
  int rc = plan_pipe(c, bytes, mc ? (u64)c->nvls.stage_bytes : 2 * c->ctx.half_bytes, 0, &p);
  if (rc != NBD_OK) return rc;
  p.op.root = root;
  p.op.algo |= mc ? 0 : 1;
  if (relay) p.op.algo = 1 | relay_mode(c);
  PackPipe<1> pk;
  fill_pipe(pk, c, buf, buf, mc);
  c->last_algo = relay ? "bcast_relay" : (mc ? "bcast_nvls" : "bcast_pipe");
  return pipe_launch(fn, c->device, p.nblocks, 1, false, s, &pk, &p.op);
}

// ---------------------------------------------------------------------------
// launch
// ---------------------------------------------------------------------------
template <int NCTX>
static int launch_any(const void* kernel, int nblocks, int world, cudaStream_t s, Pack<NCTX>& pk, OpParams* op) {
  void* args[2] = {(void*)&pk, (void*)op};
  if (NCTX == 1) {
    NBD_CUDA(cudaLaunchKernel(kernel, dim3(nblocks, 1, 1), dim3(kThreads, 1, 1), args, 0, s));
  } else {
    NBD_CUDA(cudaLaunchCooperativeKernel(kernel, dim3(nblocks, world, 1), dim3(kThreads, 1, 1), args, 0, s));
  }
  g_launches.fetch_add(1);
  return NBD_OK;
}

template <int NCTX>
static int launch_reduce(Kind k, int dtype, int op, int nblocks, int world, cudaStream_t s, Pack<NCTX>& pk,
                         OpParams* prm) {
  const bool sum = (op == NBD_SUM || op == NBD_AVG);
  const void* fn = staged_reduce_kernel_float((int)k, dtype, sum, NCTX != 1);
  if (!fn) fn = staged_reduce_kernel_int((int)k, dtype, sum, NCTX != 1);
  if (!fn) return fail(NBD_ERR_UNSUPPORTED, "dtype %d cannot be reduced", dtype);
  return launch_any<NCTX>(fn, nblocks, world, s, pk, prm);
}

static int check_usable(nbd_comm* c) {
  if (!c) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (!c->connected) return fail(NBD_ERR_INVALID_ARGUMENT, "communicator is not connected (call nbd_comm_connect)");
  if (c->dead) return fail(NBD_ERR_ABORTED, "communicator was aborted or hit an asynchronous error");
  return nbd_comm_async_error(c);
}

static int check_reduce_args(int dtype, int op) {
  const size_t es = nbd_dtype_size(dtype);
  if (es == 0 || dtype == NBD_BYTES) return fail(NBD_ERR_UNSUPPORTED, "dtype %d cannot be reduced", dtype);
  const bool is_float = (dtype == NBD_F32 || dtype == NBD_BF16 || dtype == NBD_F16 || dtype == NBD_F64);
  if (op < NBD_SUM || op > NBD_BXOR) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown reduce op %d", op);
  if (is_float && (op == NBD_BAND || op == NBD_BOR || op == NBD_BXOR))
    return fail(NBD_ERR_UNSUPPORTED, "bitwise reduce ops need an integer dtype");
  if (!is_float && op == NBD_AVG) return fail(NBD_ERR_UNSUPPORTED, "AVG needs a floating-point dtype");
  return NBD_OK;
}

// Fault injection for the watchdog tests ("debug_mute_rank" = rank + 1 on an emulated group): the
// muted rank's signals are redirected into its own LL area, so its peers never see it arrive and
// run into the flag-wait watchdog, exactly as if that rank's process had died.
static void mute_rank(const EmuGroup* g, RankCtx* ctx) {
  const uint64_t m = g->comms[0]->t_debug_mute;
  if (m == 0 || m > (uint64_t)g->world) return;
  RankCtx& x = ctx[m - 1];
  for (int p = 0; p < g->world; ++p)
    if (p != x.rank) x.flags[p] = reinterpret_cast<u64*>(x.stage[x.rank] + x.ll_off);
}

// Fills pk for a real rank (NCTX == 1) or every emulated rank.
static void fill_pack(Pack<1>& pk, nbd_comm* c, const void* send, void* recv) {
  memset(pk.shard, 0, sizeof(pk.shard));
  pk.ctx[0] = c->ctx;
  pk.buf[0].send = (const char*)send;
  pk.buf[0].recv = (char*)recv;
}

static int fill_pack_emu(Pack<kMaxRanks>& pk, EmuGroup* g, const void* const* sends, void* const* recvs) {
  memset(&pk, 0, sizeof(pk));
  for (int r = 0; r < g->world; ++r) {
    int rc = check_usable(g->comms[r]);
    if (rc != NBD_OK) return rc;
    pk.ctx[r] = g->comms[r]->ctx;
    pk.buf[r].send = sends ? (const char*)sends[r] : nullptr;
    pk.buf[r].recv = recvs ? (char*)recvs[r] : nullptr;
  }
  mute_rank(g, pk.ctx);
  return NBD_OK;
}


// ---------------------------------------------------------------------------
// zero-copy: registration lookup, planning, launch
// ---------------------------------------------------------------------------
// Registered buffers: below ~32 MiB the staged kernels are as fast or faster (2xB200: 16 MiB staged
// 40 us vs zero-copy 46 us, 64 MiB 143 vs 123; 8xB200: 16 MiB 89 vs 96, 64 MiB 234 vs 222).
static const u64 kDefZcMin = 32ull << 20;

static bool zc_wanted(const nbd_comm* c, u64 bytes) {
  // emulated ranks have no NVLink round trips to amortise: zero-copy halves the HBM traffic at any size
  return !c->t_zc_off && bytes >= (c->t_zc_min ? c->t_zc_min : (c->emu ? (256ull << 10) : kDefZcMin));
}

// index of the connected registration that contains [ptr, ptr+bytes), or -1
static int find_reg(const nbd_comm* c, const void* ptr, u64 bytes, u64* off) {
  const char* p = static_cast<const char*>(ptr);
  for (size_t k = 0; k < c->regs.size(); ++k) {
    const auto& g = c->regs[k];
    if (g.connected && g.base && !g.mc && p >= g.base && p + bytes <= g.base + g.bytes) {
      *off = (u64)(p - g.base);
      return (int)k;
    }
  }
  return -1;
}

// multicast-pool registration (nbd_mem_alloc with in-switch reduction on) that contains [ptr, ptr+bytes), or -1
static int find_mc_reg(const nbd_comm* c, const void* ptr, u64 bytes, u64* off) {
  const char* p = static_cast<const char*>(ptr);
  for (size_t k = 0; k < c->regs.size(); ++k) {
    const auto& g = c->regs[k];
    if (g.base && g.mc && p >= g.base && p + bytes <= g.base + g.bytes) {
      *off = (u64)(p - g.base);
      return (int)k;
    }
  }
  return -1;
}

static int launch_ptr(const void* kernel, int nblocks, int world, bool coop, cudaStream_t s, void* pk, OpParams* op) {
  void* args[2] = {pk, (void*)op};
  if (!coop) {
    NBD_CUDA(cudaLaunchKernel(kernel, dim3(nblocks, 1, 1), dim3(kThreads, 1, 1), args, 0, s));
  } else {
    NBD_CUDA(cudaLaunchCooperativeKernel(kernel, dim3(nblocks, world, 1), dim3(kThreads, 1, 1), args, 0, s));
  }
  g_launches.fetch_add(1);
  return NBD_OK;
}

template <int NCTX>
static int launch_zc_reduce(ZKind k, int dtype, int op, int nblocks, int world, cudaStream_t s, PackZc<NCTX>& pk,
                            OpParams* prm) {
  const bool sum = (op == NBD_SUM || op == NBD_AVG);
  const void* fn = zc_reduce_kernel((int)k, dtype, sum, NCTX != 1);
  if (!fn) return fail(NBD_ERR_UNSUPPORTED, "dtype %d cannot be reduced", dtype);
  return launch_ptr(fn, nblocks, world, NCTX != 1, s, &pk, prm);
}

static OpParams zc_params(const nbd_comm* c, u64 bytes, int op, int algo) {
  OpParams p;
  memset(&p, 0, sizeof(p));
  p.bytes = bytes;
  p.op = op;
  p.avg = (op == NBD_AVG);
  p.world = c->world;
  p.algo = algo;
  return p;
}

// all_reduce zero-copy variant: fused loop for two ranks, tiled read-then-push beyond
static void zc_ar_variant(const nbd_comm* c, OpParams* p) {
  // Emulated ranks (every "peer" is local HBM) always take the fused loop: the tile's re-read
  // buys nothing without NVLink (1 GPU, 8 x 256 MiB: 0.731 ms fused against 0.789 tiled).
  const bool tiled = c->t_zc_fused ? (c->t_zc_fused == 2) : (c->world > 2 && !c->emu);
  p->algo = tiled ? 1 : 0;
  p->chunk_bytes = tiled ? c->t_zc_tile : 0;
}

static int fill_zc_emu(PackZc<kMaxRanks>& pk, EmuGroup* g, const void* const* sends, void* const* recvs) {
  memset(&pk, 0, sizeof(pk));
  for (int r = 0; r < g->world; ++r) {
    int rc = check_usable(g->comms[r]);
    if (rc != NBD_OK) return rc;
    pk.ctx[r] = g->comms[r]->ctx;
    pk.z.send[r] = (const char*)sends[r];
    pk.z.recv[r] = (char*)recvs[r];
  }
  mute_rank(g, pk.ctx);
  return NBD_OK;
}

static void set_algo(nbd_comm* c, const char* name) {
  c->last_algo = name;
  if (c->emu)
    for (int r = 0; r < c->group->world; ++r) c->group->comms[r]->last_algo = name;
}


// ---------------------------------------------------------------------------
// auto-registration of ordinary (unregistered) buffers — see nbd_device.cuh
// ---------------------------------------------------------------------------
// Below this the agreement kernel (~15 us) costs more than zero-copy saves (2xB200: 16 MiB staged
// 40 us vs auto 62 us; 64 MiB 143 vs 139; 256 MiB 498 vs 444; profiles/r01_tune_n2.jsonl).
static const u64 kDefAutoMin = 64ull << 20;

typedef CUresult (*PFN_cuPointerGetAttribute)(void*, CUpointer_attribute, CUdeviceptr);
typedef CUresult (*PFN_cuMemGetAddressRange)(CUdeviceptr*, size_t*, CUdeviceptr);
static PFN_cuPointerGetAttribute p_cuPointerGetAttribute = nullptr;
static PFN_cuMemGetAddressRange p_cuMemGetAddressRange = nullptr;
static std::once_flag g_drv_once;

static void load_driver_entry_points() {
  std::call_once(g_drv_once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuPointerGetAttribute", &f, cudaEnableDefault, &q) == cudaSuccess && f)
      p_cuPointerGetAttribute = (PFN_cuPointerGetAttribute)f;
    f = nullptr;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q) == cudaSuccess && f)
      p_cuMemGetAddressRange = (PFN_cuMemGetAddressRange)f;
    cudaGetLastError();
  });
}

static bool auto_wanted(const nbd_comm* c, u64 bytes) {
  return !c->emu && c->world > 1 && !c->t_zc_off && !c->t_auto_off &&
         bytes >= (c->t_auto_min ? c->t_auto_min : kDefAutoMin);
}

// all_reduce takes the same threshold at every world size.  (Round 1 kept 8 ranks staged below
// 512 MiB on a stale measurement; 8xB200 round 2: 64 MiB staged 493 / auto 507, 256 MiB 577 / 630.)
static bool auto_wanted_ar(const nbd_comm* c, u64 bytes) { return auto_wanted(c, bytes); }

struct ExportRef {
  u64 slot = kNoSlot;
  u64 off = 0;
};

// The export slot of the allocation that holds [ptr, ptr+bytes), exporting it on first use.
// kNoSlot when the allocation cannot be shared (not a cudaMalloc allocation, table full, ...).
static ExportRef export_ref(nbd_comm* c, const void* ptr, u64 bytes, cudaStream_t s) {
  ExportRef ref;
  load_driver_entry_points();
  if (!p_cuPointerGetAttribute || !p_cuMemGetAddressRange) return ref;
  unsigned long long id = 0;
  CUdeviceptr base = 0;
  size_t size = 0;
  if (p_cuPointerGetAttribute(&id, CU_POINTER_ATTRIBUTE_BUFFER_ID, (CUdeviceptr)ptr) != CUDA_SUCCESS) return ref;
  if (p_cuMemGetAddressRange(&base, &size, (CUdeviceptr)ptr) != CUDA_SUCCESS) return ref;
  if ((CUdeviceptr)ptr + bytes > base + size) return ref;
  auto it = c->export_of.find(id);
  if (it == c->export_of.end()) {
    if (c->not_exportable.count(id) || c->n_exports >= kMaxExports) return ref;
    ExportSlot& e = c->export_mirror[c->n_exports];
    memset(&e, 0, sizeof(e));
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, (void*)base) != cudaSuccess) {
      cudaGetLastError();
      c->not_exportable.insert(id);  // e.g. expandable (VMM) segments of torch's allocator
      return ref;
    }
    e.buffer_id = id;
    e.base = (u64)base;
    e.size = (u64)size;
    memcpy(e.ipc, &h, sizeof(h));
    // The record first, its `valid` word last and in a copy of its own: a peer's host reads the
    // slot with a device-to-host copy that may run while ours is in flight, and must never see
    // valid == 1 on a half-written record.  (Both copies are ordered on `s`; the 1 comes from a
    // pinned cell of its own because the mirror is read when the copy executes, not when it is
    // enqueued.)
    e.valid = 0;
    char* row = c->arena + (size_t)(kExportRow0 + c->n_exports * kExportSlotRows) * kMaxRanks * sizeof(u64);
    uint64_t* one = &c->imp_tmp[c->imp_tmp_n++ % 64];
    *one = 1;
    if (cudaMemcpyAsync(row, &e, sizeof(e), cudaMemcpyHostToDevice, s) != cudaSuccess ||
        cudaMemcpyAsync(row + offsetof(ExportSlot, valid), one, sizeof(uint64_t), cudaMemcpyHostToDevice, s) != cudaSuccess) {
      cudaGetLastError();
      return ref;
    }
    it = c->export_of.emplace(id, c->n_exports++).first;
  }
  ref.slot = (u64)it->second;
  ref.off = (u64)((CUdeviceptr)ptr - (CUdeviceptr)c->export_mirror[it->second].base);
  return ref;
}

// Import the (peer, slot) pairs kernels reported as unresolved; later operations on those
// allocations run zero-copy.  Failures are remembered and simply keep the staged path.
static void process_misses(nbd_comm* c, cudaStream_t s) {
  const uint32_t count = ((volatile HostMailbox*)c->mbox)->miss_count;
  if (count == c->miss_seen) return;
  uint32_t first = c->miss_seen;
  if (count - first > (uint32_t)kMissRing) first = count - kMissRing;
  c->miss_seen = count;
  LocalBlock* blk = reinterpret_cast<LocalBlock*>(c->state);
  for (uint32_t i = first; i != count; ++i) {
    const u64 key = ((volatile HostMailbox*)c->mbox)->miss[i % kMissRing];
    const int peer = (int)(key >> 32);
    const u64 slot = key & 0xffffffffull;
    if (peer < 0 || peer >= c->world || peer == c->rank || slot >= (u64)kMaxExports) continue;
    if (!c->imported.insert(key).second) continue;
    const char* row = (const char*)c->peer_base[peer] + (size_t)(kExportRow0 + slot * kExportSlotRows) * kMaxRanks * sizeof(u64);
    if (cudaMemcpyAsync(c->import_tmp, row, sizeof(ExportSlot), cudaMemcpyDeviceToHost, c->side) != cudaSuccess ||
        cudaStreamSynchronize(c->side) != cudaSuccess) {
      cudaGetLastError();
      continue;
    }
    if (c->import_tmp->valid != 1) {
      c->imported.erase(key);  // not published yet: try again when it is reported again
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, c->import_tmp->ipc, sizeof(h));
    void* mapped = nullptr;
    if (cudaIpcOpenMemHandle(&mapped, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      continue;
    }
    c->import_ptrs.push_back(mapped);
    uint64_t* cell = &c->imp_tmp[c->imp_tmp_n++ % 64];
    *cell = (uint64_t)mapped;
    if (cudaMemcpyAsync(&blk->imp[peer][slot], cell, sizeof(uint64_t), cudaMemcpyHostToDevice, s) != cudaSuccess)
      cudaGetLastError();
  }
}

static int launch_agree(nbd_comm* c, const ExportRef& es, const ExportRef& er, u64 bytes, u64 optag, int need,
                        const void* send, void* recv, cudaStream_t s) {
  Pack<1> pk;
  fill_pack(pk, c, send, recv);
  AgreeArgs a;
  memset(&a, 0, sizeof(a));
  const u64 d[6] = {es.slot, es.off, er.slot, er.off, bytes, optag};
  memcpy(a.desc, d, sizeof(d));
  a.local_send = (const char*)send;
  a.local_recv = (char*)recv;
  a.need = need;
  void* args[2] = {(void*)&pk, (void*)&a};
  NBD_CUDA(cudaLaunchKernel((const void*)k_zc_agree, dim3(1, 1, 1), dim3(64, 1, 1), args, 0, s));
  g_launches.fetch_add(1);
  return NBD_OK;
}

extern "C" {

// ---- real ranks -----------------------------------------------------------

int nbd_all_reduce(nbd_comm_t c, const void* send, void* recv, size_t count, int dtype, int op, void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_all_reduce with emulated communicators");
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (count == 0) return NBD_OK;
  if (!send || !recv) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer");
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)count * nbd_dtype_size(dtype);
  if (c->world == 1) {
    c->last_algo = "local";
    if (send != recv) NBD_CUDA(cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, s));
    return NBD_OK;
  }
  {
    u64 offs = 0, offr = 0;
    const int ks = zc_wanted(c, bytes) ? find_reg(c, send, bytes, &offs) : -1;
    const int kr = ks >= 0 ? find_reg(c, recv, bytes, &offr) : -1;
    if (ks >= 0 && kr >= 0) {
      PackZc<1> zk;
      memset(&zk, 0, sizeof(zk));
      zk.ctx[0] = c->ctx;
      for (int q = 0; q < c->world; ++q) {
        zk.z.send[q] = (const char*)c->regs[ks].peer[q] + offs;
        zk.z.recv[q] = (char*)c->regs[kr].peer[q] + offr;
      }
      const u64 d[6] = {(u64)ks, offs, (u64)kr, offr, bytes, 1ull | ((u64)dtype << 8) | ((u64)op << 16)};
      memcpy(zk.z.desc, d, sizeof(d));
      OpParams prm = zc_params(c, bytes, op, 0);
      zc_ar_variant(c, &prm);
      c->last_algo = "ar_zerocopy";
      return launch_zc_reduce<1>(Z_AR, dtype, op, pick_blocks(c, ceil_div(ceil_div(bytes, kUnit), c->world), 0), 1, s, zk,
                                 &prm);
    }
  }
  if (nvls_reduce_wanted(c, bytes, dtype, op)) {
    // both buffers in the multicast pool, whole 16-byte units: no staging, the switch does it all
    u64 offs = 0, offr = 0;
    const int ks = (bytes % kUnit == 0) ? find_mc_reg(c, send, bytes, &offs) : -1;
    const int kr = ks >= 0 ? find_mc_reg(c, recv, bytes, &offr) : -1;
    const void* fn = (ks >= 0 && kr >= 0 && offs % kUnit == 0 && offr % kUnit == 0) ? mcz_allreduce_kernel(dtype) : nullptr;
    if (fn) {
      Pack<1> pk;
      fill_pack(pk, c, send, recv);
      McPtrs m = {c->regs[ks].mc + offs, c->regs[kr].mc + offr};
      OpParams prm = zc_params(c, bytes, op, 0);
      // 32 CTAs saturate the switch path and leave fewer flags to exchange at the two barriers:
      // 8xB200, 256 MiB: 800 GB/s with 32, 789 with 64, 764 with 148 (profiles/r02_tune3_n8.jsonl)
      int nb = pick_blocks(c, ceil_div(bytes / kUnit, (u64)c->world), 0);
      if (!c->t_blocks && nb > 32) nb = 32;
      void* args[3] = {(void*)&pk, (void*)&m, (void*)&prm};
      NBD_CUDA(cudaLaunchKernel(fn, dim3(nb, 1, 1), dim3(kThreads, 1, 1), args, 0, s));
      g_launches.fetch_add(1);
      c->last_algo = "ar_nvls_zerocopy";
      return NBD_OK;
    }
    if (bytes >= (c->t_nvls_min ? c->t_nvls_min : kDefNvlsReduceMin)) return pipe_all_reduce(c, send, recv, bytes, dtype, op, true, s);
  }
  if (pipe_wanted(c, bytes)) return pipe_all_reduce(c, send, recv, bytes, dtype, op, false, s);
  if (auto_wanted_ar(c, bytes)) {
    // ordinary tensors: agree on the device whether every rank can address every rank's
    // buffers; enqueue both kernels, the decision gates them (nbd_device.cuh)
    const ExportRef es = export_ref(c, send, bytes, s);
    const ExportRef er = (recv == send) ? es : export_ref(c, recv, bytes, s);
    process_misses(c, s);
    if ((rc = launch_agree(c, es, er, bytes, 1ull | ((u64)dtype << 8) | ((u64)op << 16), 3, send, recv, s)) != NBD_OK)
      return rc;
    PackZc<1> zk;
    memset(&zk, 0, sizeof(zk));
    zk.ctx[0] = c->ctx;
    OpParams prm = zc_params(c, bytes, op, 0);
    zc_ar_variant(c, &prm);
    prm.zc_mode = 1;
    if ((rc = launch_zc_reduce<1>(Z_AR, dtype, op, pick_blocks(c, ceil_div(ceil_div(bytes, kUnit), c->world), 0), 1, s,
                                  zk, &prm)) != NBD_OK)
      return rc;
    Plan p = plan_all_reduce(c, bytes, op, 0);
    p.op.zc_mode = 2;
    Pack<1> pk;
    fill_pack(pk, c, send, recv);
    c->last_algo = "ar_auto";
    return launch_reduce<1>(p.op.algo == 1 ? K_AR1 : K_AR2, dtype, op, p.nblocks, 1, s, pk, &p.op);
  }
  Plan p = plan_all_reduce(c, bytes, op, 0);
  Pack<1> pk;
  fill_pack(pk, c, send, recv);
  return launch_reduce<1>(p.op.algo == 3 ? K_LL : (p.op.algo == 1 ? K_AR1 : K_AR2), dtype, op, p.nblocks, 1, s, pk, &p.op);
}

int nbd_reduce_scatter(nbd_comm_t c, const void* send, void* recv, size_t recv_count, int dtype, int op,
                       void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_reduce_scatter with emulated communicators");
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (recv_count == 0) return NBD_OK;
  if (!send || !recv) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer");
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)recv_count * nbd_dtype_size(dtype);
  if (c->world == 1) {
    c->last_algo = "local";
    if (send != recv) NBD_CUDA(cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, s));
    return NBD_OK;
  }
  {
    u64 offs = 0;
    const int ks = zc_wanted(c, bytes) ? find_reg(c, send, bytes * c->world, &offs) : -1;
    if (ks >= 0) {
      PackZc<1> zk;
      memset(&zk, 0, sizeof(zk));
      zk.ctx[0] = c->ctx;
      for (int q = 0; q < c->world; ++q) zk.z.send[q] = (const char*)c->regs[ks].peer[q] + offs;
      zk.z.recv[c->rank] = (char*)recv;
      const u64 d[6] = {(u64)ks, offs, 0, 0, bytes, 2ull | ((u64)dtype << 8) | ((u64)op << 16)};
      memcpy(zk.z.desc, d, sizeof(d));
      OpParams prm = zc_params(c, bytes, op, 0);
      c->last_algo = "reduce_scatter_zerocopy";
      return launch_zc_reduce<1>(Z_RS, dtype, op, pick_blocks(c, ceil_div(bytes, kUnit), 0), 1, s, zk, &prm);
    }
  }
  if (auto_wanted(c, bytes * c->world)) {
    const ExportRef es = export_ref(c, send, bytes * c->world, s);
    process_misses(c, s);
    if ((rc = launch_agree(c, es, ExportRef(), bytes, 2ull | ((u64)dtype << 8) | ((u64)op << 16), 1, send, recv, s)) != NBD_OK)
      return rc;
    PackZc<1> zk;
    memset(&zk, 0, sizeof(zk));
    zk.ctx[0] = c->ctx;
    OpParams prm = zc_params(c, bytes, op, 0);
    prm.zc_mode = 1;
    if ((rc = launch_zc_reduce<1>(Z_RS, dtype, op, pick_blocks(c, ceil_div(bytes, kUnit), 0), 1, s, zk, &prm)) != NBD_OK)
      return rc;
    Plan p = plan_reduce_scatter(c, bytes, op, 0);
    p.op.zc_mode = 2;
    Pack<1> pk;
    fill_pack(pk, c, send, recv);
    c->last_algo = "reduce_scatter_auto";
    return launch_reduce<1>(K_RS, dtype, op, p.nblocks, 1, s, pk, &p.op);
  }
  Plan p = plan_reduce_scatter(c, bytes, op, 0);
  Pack<1> pk;
  fill_pack(pk, c, send, recv);
  return launch_reduce<1>(K_RS, dtype, op, p.nblocks, 1, s, pk, &p.op);
}

int nbd_all_gather(nbd_comm_t c, const void* send, void* recv, size_t send_count, int dtype, void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_all_gather with emulated communicators");
  const size_t es = nbd_dtype_size(dtype);
  if (es == 0) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown dtype %d", dtype);
  if (send_count == 0) return NBD_OK;
  if (!send || !recv) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer");
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)send_count * es;
  if (c->world == 1) {
    c->last_algo = "local";
    if (send != recv) NBD_CUDA(cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, s));
    return NBD_OK;
  }
  if (zc_wanted(c, bytes)) {
    u64 offs = 0, offr = 0;
    const int ks = find_reg(c, send, bytes, &offs);
    const int kr = ks >= 0 ? -1 : find_reg(c, recv, bytes * c->world, &offr);
    if (ks >= 0 || kr >= 0) {
      PackZc<1> zk;
      memset(&zk, 0, sizeof(zk));
      zk.ctx[0] = c->ctx;
      for (int q = 0; q < c->world; ++q) {
        if (ks >= 0) zk.z.send[q] = (const char*)c->regs[ks].peer[q] + offs;
        else zk.z.recv[q] = (char*)c->regs[kr].peer[q] + offr;
      }
      if (ks >= 0) zk.z.recv[c->rank] = (char*)recv;
      else zk.z.send[c->rank] = (const char*)send;
      const u64 d[6] = {(u64)(ks >= 0 ? ks : 0), offs, (u64)(kr >= 0 ? kr : 0), offr, bytes, 3ull | ((u64)(ks >= 0) << 8)};
      memcpy(zk.z.desc, d, sizeof(d));
      OpParams prm = zc_params(c, bytes, 0, ks >= 0 ? 0 : 1);
      c->last_algo = ks >= 0 ? "all_gather_zerocopy_pull" : "all_gather_zerocopy_push";
      return launch_ptr((const void*)k_all_gather_zc<1>, pick_blocks(c, ceil_div(bytes, kUnit), 0), 1, false, s, &zk, &prm);
    }
  }
  if (auto_wanted(c, bytes * c->world)) {
    const ExportRef es = export_ref(c, send, bytes, s);  // pull: peers read our send buffer
    process_misses(c, s);
    int rc2 = launch_agree(c, es, ExportRef(), bytes, 3ull | (1ull << 8), 1, send, recv, s);
    if (rc2 != NBD_OK) return rc2;
    PackZc<1> zk;
    memset(&zk, 0, sizeof(zk));
    zk.ctx[0] = c->ctx;
    OpParams prm = zc_params(c, bytes, 0, 0);
    prm.zc_mode = 1;
    if ((rc2 = launch_ptr((const void*)k_all_gather_zc<1>, pick_blocks(c, ceil_div(bytes, kUnit), 0), 1, false, s, &zk,
                          &prm)) != NBD_OK)
      return rc2;
    Plan p = plan_all_gather(c, bytes, 0);
    p.op.zc_mode = 2;
    Pack<1> pk;
    fill_pack(pk, c, send, recv);
    c->last_algo = "all_gather_auto";
    return launch_any<1>((const void*)k_all_gather<1>, p.nblocks, 1, s, pk, &p.op);
  }
  Plan p = plan_all_gather(c, bytes, 0);
  Pack<1> pk;
  fill_pack(pk, c, send, recv);
  return launch_any<1>((const void*)k_all_gather<1>, p.nblocks, 1, s, pk, &p.op);
}

int nbd_broadcast(nbd_comm_t c, void* buf, size_t count, int dtype, int root, void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_broadcast with emulated communicators");
  const size_t es = nbd_dtype_size(dtype);
  if (es == 0) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown dtype %d", dtype);
  if (root < 0 || root >= c->world) return fail(NBD_ERR_INVALID_ARGUMENT, "root %d not in [0,%d)", root, c->world);
  if (count == 0) return NBD_OK;
  if (!buf) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer");
  if (c->world == 1) {
    c->last_algo = "local";
    return NBD_OK;
  }
  DeviceGuard g(c->device);
  cudaStream_t s = (cudaStream_t)stream;
  if (nvls_bcast_wanted(c, (u64)count * es)) return pipe_broadcast(c, buf, (u64)count * es, root, true, s);
  if (pipe_wanted(c, (u64)count * es)) return pipe_broadcast(c, buf, (u64)count * es, root, false, s);
  if (relay_wanted(c, (u64)count * es)) return pipe_broadcast(c, buf, (u64)count * es, root, false, s, true);
  Plan p = plan_broadcast(c, (u64)count * es, root, 0);
  Pack<1> pk;
  fill_pack(pk, c, buf, buf);
  return launch_any<1>(p.op.algo == 2 ? (const void*)k_broadcast_ll<1> : (const void*)k_broadcast<1>, p.nblocks, 1, s, pk, &p.op);
}

int nbd_barrier(nbd_comm_t c, void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_barrier with emulated communicators");
  if (c->world == 1) return NBD_OK;
  DeviceGuard g(c->device);
  Pack<1> pk;
  fill_pack(pk, c, nullptr, nullptr);
  void* args[1] = {(void*)&pk};
  NBD_CUDA(cudaLaunchKernel((const void*)k_barrier<1>, dim3(1, 1, 1), dim3(kThreads, 1, 1), args, 0,
                            (cudaStream_t)stream));
  g_launches.fetch_add(1);
  c->last_algo = "barrier";
  return NBD_OK;
}

// ---- emulated ranks -------------------------------------------------------

static int emu_local_copies(EmuGroup* g, const void* const* sends, void* const* recvs, u64 bytes, cudaStream_t s) {
  for (int r = 0; r < g->world; ++r)
    if (sends[r] != recvs[r]) NBD_CUDA(cudaMemcpyAsync(recvs[r], sends[r], bytes, cudaMemcpyDeviceToDevice, s));
  return NBD_OK;
}

static int emu_check(nbd_comm_t any, const void* const* a, const void* const* b, bool need_b) {
  if (!any) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (!any->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "not an emulated communicator");
  if (!a || (need_b && !b)) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL pointer array");
  for (int r = 0; r < any->group->world; ++r)
    if (!a[r] || (need_b && !b[r])) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer for rank %d", r);
  return NBD_OK;
}

int nbd_emu_all_reduce(nbd_comm_t any, const void* const* sends, void* const* recvs, size_t count, int dtype,
                       int op, void* stream) {
  int rc;
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (any && any->emu && count == 0) return NBD_OK;
  if ((rc = emu_check(any, sends, (const void* const*)recvs, true)) != NBD_OK) return rc;
  EmuGroup* g = any->group;
  nbd_comm* c0 = g->comms[0];
  DeviceGuard dg(c0->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)count * nbd_dtype_size(dtype);
  if (g->world == 1) return emu_local_copies(g, sends, recvs, bytes, s);
  if (pipe_wanted(c0, bytes)) {  // the tile pipeline over the emulated arenas (P2P transport)
    const void* fn = pipe_allreduce_kernel(dtype, op == NBD_SUM || op == NBD_AVG, true, false);
    if (!fn) return fail(NBD_ERR_UNSUPPORTED, "dtype %d cannot be reduced", dtype);
    Plan p;
    if ((rc = plan_pipe(c0, bytes, 2 * c0->ctx.half_bytes, g->world, &p)) != NBD_OK) return rc;
    p.op.op = op;
    p.op.avg = (op == NBD_AVG);
    p.op.algo |= 1;
    PackPipe<kMaxRanks> pk;
    memset(&pk, 0, sizeof(pk));
    for (int r = 0; r < g->world; ++r) {
      if ((rc = check_usable(g->comms[r])) != NBD_OK) return rc;
      pk.ctx[r] = g->comms[r]->ctx;
      pk.buf[r].send = (const char*)sends[r];
      pk.buf[r].recv = (char*)recvs[r];
      pk.pb[r].local = g->comms[r]->ctx.stage[r];
    }
    mute_rank(g, pk.ctx);
    set_algo(c0, "ar_pipe");
    return pipe_launch(fn, c0->device, p.nblocks, g->world, true, s, &pk, &p.op);
  }
  if (zc_wanted(c0, bytes)) {  // emulated ranks share one address space: always zero-copy capable
    PackZc<kMaxRanks> zk;
    if ((rc = fill_zc_emu(zk, g, sends, recvs)) != NBD_OK) return rc;
    OpParams prm = zc_params(c0, bytes, op, 0);
    zc_ar_variant(c0, &prm);
    set_algo(c0, "ar_zerocopy");
    // grid per rank: the largest power of two that fits (8 ranks on 148 SMs: 16 CTAs each, 0.697 ms
    // against 0.734 with 18; profiles/r02_tune_emu_n1.jsonl)
    int nb = pick_blocks(c0, ceil_div(ceil_div(bytes, kUnit), g->world), g->world);
    if (!c0->t_blocks) {
      int p2 = 1;
      while (p2 * 2 <= nb) p2 *= 2;
      nb = p2;
    }
    return launch_zc_reduce<kMaxRanks>(Z_AR, dtype, op, nb, g->world, s, zk, &prm);
  }
  Pack<kMaxRanks> pk;
  if ((rc = fill_pack_emu(pk, g, sends, recvs)) != NBD_OK) return rc;
  Plan p = plan_all_reduce(c0, bytes, op, g->world);
  for (int r = 1; r < g->world; ++r) g->comms[r]->last_algo = c0->last_algo;
  return launch_reduce<kMaxRanks>(p.op.algo == 3 ? K_LL : (p.op.algo == 1 ? K_AR1 : K_AR2), dtype, op, p.nblocks, g->world,
                                  s, pk, &p.op);
}

int nbd_emu_reduce_scatter(nbd_comm_t any, const void* const* sends, void* const* recvs, size_t recv_count,
                           int dtype, int op, void* stream) {
  int rc;
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (any && any->emu && recv_count == 0) return NBD_OK;
  if ((rc = emu_check(any, sends, (const void* const*)recvs, true)) != NBD_OK) return rc;
  EmuGroup* g = any->group;
  nbd_comm* c0 = g->comms[0];
  DeviceGuard dg(c0->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)recv_count * nbd_dtype_size(dtype);
  if (g->world == 1) return emu_local_copies(g, sends, recvs, bytes, s);
  if (zc_wanted(c0, bytes)) {
    PackZc<kMaxRanks> zk;
    if ((rc = fill_zc_emu(zk, g, sends, recvs)) != NBD_OK) return rc;
    OpParams prm = zc_params(c0, bytes, op, 0);
    set_algo(c0, "reduce_scatter_zerocopy");
    return launch_zc_reduce<kMaxRanks>(Z_RS, dtype, op, pick_blocks(c0, ceil_div(bytes, kUnit), g->world), g->world, s, zk,
                                       &prm);
  }
  Pack<kMaxRanks> pk;
  if ((rc = fill_pack_emu(pk, g, sends, recvs)) != NBD_OK) return rc;
  Plan p = plan_reduce_scatter(c0, bytes, op, g->world);
  for (int r = 1; r < g->world; ++r) g->comms[r]->last_algo = c0->last_algo;
  return launch_reduce<kMaxRanks>(K_RS, dtype, op, p.nblocks, g->world, s, pk, &p.op);
}

// reduce_scatter whose N contributions are separate buffers (c10d's list form,
// $TORCH/.../Backend.hpp reduce_scatter(outputTensors, inputTensors)): the staged kernel takes
// the pointers as they are, nothing is concatenated.  sends[p] = contribution to rank p's result.
int nbd_reduce_scatter_list(nbd_comm_t c, const void* const* sends, void* recv, size_t recv_count, int dtype, int op,
                            void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_reduce_scatter_list with emulated communicators");
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (recv_count == 0) return NBD_OK;
  if (!sends || !recv) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer");
  for (int p = 0; p < c->world; ++p)
    if (!sends[p]) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL contribution for rank %d", p);
  if (c->world == 1) return nbd_reduce_scatter(c, sends[0], recv, recv_count, dtype, op, stream);
  DeviceGuard g(c->device);
  const u64 bytes = (u64)recv_count * nbd_dtype_size(dtype);
  Plan p = plan_reduce_scatter(c, bytes, op, 0);
  Pack<1> pk;
  fill_pack(pk, c, sends[0], recv);
  for (int q = 0; q < c->world; ++q) pk.shard[0][q] = (const char*)sends[q];
  c->last_algo = "reduce_scatter_list";
  return launch_reduce<1>(K_RS, dtype, op, p.nblocks, 1, (cudaStream_t)stream, pk, &p.op);
}

// Emulated ranks: sends[r * world + p] = rank r's contribution to rank p's result.
int nbd_emu_reduce_scatter_list(nbd_comm_t any, const void* const* sends, void* const* recvs, size_t recv_count, int dtype,
                                int op, void* stream) {
  int rc;
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (any && any->emu && recv_count == 0) return NBD_OK;
  if (!any) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (!any->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "not an emulated communicator");
  if (!sends || !recvs) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL pointer array");
  EmuGroup* g = any->group;
  const int N = g->world;
  for (int r = 0; r < N; ++r) {
    if (!recvs[r]) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL result buffer for rank %d", r);
    for (int p = 0; p < N; ++p)
      if (!sends[r * N + p]) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL contribution of rank %d for rank %d", r, p);
  }
  nbd_comm* c0 = g->comms[0];
  DeviceGuard dg(c0->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)recv_count * nbd_dtype_size(dtype);
  if (N == 1) {
    if (sends[0] != recvs[0]) NBD_CUDA(cudaMemcpyAsync(recvs[0], sends[0], bytes, cudaMemcpyDeviceToDevice, s));
    return NBD_OK;
  }
  const void* first[kMaxRanks];
  for (int r = 0; r < N; ++r) first[r] = sends[r * N];
  Pack<kMaxRanks> pk;
  if ((rc = fill_pack_emu(pk, g, first, recvs)) != NBD_OK) return rc;
  for (int r = 0; r < N; ++r)
    for (int p = 0; p < N; ++p) pk.shard[r][p] = (const char*)sends[r * N + p];
  Plan p = plan_reduce_scatter(c0, bytes, op, N);
  set_algo(c0, "reduce_scatter_list");
  return launch_reduce<kMaxRanks>(K_RS, dtype, op, p.nblocks, N, s, pk, &p.op);
}

int nbd_emu_all_gather(nbd_comm_t any, const void* const* sends, void* const* recvs, size_t send_count, int dtype,
                       void* stream) {
  int rc;
  const size_t es = nbd_dtype_size(dtype);
  if (es == 0) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown dtype %d", dtype);
  if (any && any->emu && send_count == 0) return NBD_OK;
  if ((rc = emu_check(any, sends, (const void* const*)recvs, true)) != NBD_OK) return rc;
  EmuGroup* g = any->group;
  nbd_comm* c0 = g->comms[0];
  DeviceGuard dg(c0->device);
  cudaStream_t s = (cudaStream_t)stream;
  const u64 bytes = (u64)send_count * es;
  if (g->world == 1) return emu_local_copies(g, sends, recvs, bytes, s);
  if (zc_wanted(c0, bytes)) {
    PackZc<kMaxRanks> zk;
    if ((rc = fill_zc_emu(zk, g, sends, recvs)) != NBD_OK) return rc;
    OpParams prm = zc_params(c0, bytes, 0, 0);
    set_algo(c0, "all_gather_zerocopy_pull");
    return launch_ptr((const void*)k_all_gather_zc<kMaxRanks>, pick_blocks(c0, ceil_div(bytes, kUnit), g->world), g->world,
                      true, s, &zk, &prm);
  }
  Pack<kMaxRanks> pk;
  if ((rc = fill_pack_emu(pk, g, sends, recvs)) != NBD_OK) return rc;
  Plan p = plan_all_gather(c0, bytes, g->world);
  for (int r = 1; r < g->world; ++r) g->comms[r]->last_algo = c0->last_algo;
  return launch_any<kMaxRanks>((const void*)k_all_gather<kMaxRanks>, p.nblocks, g->world, s, pk, &p.op);
}

int nbd_emu_broadcast(nbd_comm_t any, void* const* bufs, size_t count, int dtype, int root, void* stream) {
  int rc;
  const size_t es = nbd_dtype_size(dtype);
  if (es == 0) return fail(NBD_ERR_INVALID_ARGUMENT, "unknown dtype %d", dtype);
  if (any && any->emu && count == 0) return NBD_OK;
  if ((rc = emu_check(any, (const void* const*)bufs, nullptr, false)) != NBD_OK) return rc;
  EmuGroup* g = any->group;
  if (root < 0 || root >= g->world) return fail(NBD_ERR_INVALID_ARGUMENT, "root %d not in [0,%d)", root, g->world);
  if (g->world == 1) return NBD_OK;
  nbd_comm* c0 = g->comms[0];
  DeviceGuard dg(c0->device);
  const bool relay = relay_wanted(c0, (u64)count * es);
  if (relay || pipe_wanted(c0, (u64)count * es)) {
    Plan p;
    if ((rc = plan_pipe(c0, (u64)count * es, 2 * c0->ctx.half_bytes, g->world, &p)) != NBD_OK) return rc;
    p.op.root = root;
    p.op.algo |= 1;
    if (relay) p.op.algo = 1 | relay_mode(c0);
    PackPipe<kMaxRanks> pk;
    memset(&pk, 0, sizeof(pk));
    for (int r = 0; r < g->world; ++r) {
      if ((rc = check_usable(g->comms[r])) != NBD_OK) return rc;
      pk.ctx[r] = g->comms[r]->ctx;
      pk.buf[r].send = (const char*)bufs[r];
      pk.buf[r].recv = (char*)bufs[r];
      pk.pb[r].local = g->comms[r]->ctx.stage[r];
    }
    mute_rank(g, pk.ctx);
    set_algo(c0, relay ? "bcast_relay" : "bcast_pipe");
    return pipe_launch(relay ? relay_broadcast_kernel(true) : pipe_broadcast_kernel(true, false), c0->device, p.nblocks, g->world, true, (cudaStream_t)stream, &pk,
                       &p.op);
  }
  Pack<kMaxRanks> pk;
  if ((rc = fill_pack_emu(pk, g, (const void* const*)bufs, bufs)) != NBD_OK) return rc;
  Plan p = plan_broadcast(c0, (u64)count * es, root, g->world);
  for (int r = 1; r < g->world; ++r) g->comms[r]->last_algo = c0->last_algo;
  return launch_any<kMaxRanks>(p.op.algo == 2 ? (const void*)k_broadcast_ll<kMaxRanks> : (const void*)k_broadcast<kMaxRanks>,
                               p.nblocks, g->world, (cudaStream_t)stream, pk, &p.op);
}

int nbd_emu_barrier(nbd_comm_t any, void* stream) {
  if (!any) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (!any->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "not an emulated communicator");
  EmuGroup* g = any->group;
  if (g->world == 1) return NBD_OK;
  DeviceGuard dg(g->comms[0]->device);
  Pack<kMaxRanks> pk;
  int rc = fill_pack_emu(pk, g, nullptr, nullptr);
  if (rc != NBD_OK) return rc;
  void* args[1] = {(void*)&pk};
  NBD_CUDA(cudaLaunchCooperativeKernel((const void*)k_barrier<kMaxRanks>, dim3(1, g->world, 1),
                                       dim3(kThreads, 1, 1), args, 0, (cudaStream_t)stream));
  g_launches.fetch_add(1);
  return NBD_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// broadcast tables (nbd_table.cuh): many small broadcasts, one launch, one flag barrier
// ---------------------------------------------------------------------------
namespace {

// Chunking of a flat table (k_bcast_flat): the packed table is one message of `bytes`, every
// chunk is cut into N-1 shards that must each fit one staging half.
static Plan plan_flat(nbd_comm* c, u64 bytes, int root, int emu_world) {
  Plan p;
  memset(&p, 0, sizeof(p));
  const u64 nsh = (u64)(c->world - 1);
  p.op.bytes = bytes;
  p.op.root = root;
  p.op.world = c->world;
  p.op.algo = 1;
  const u64 cap = floor_to((c->ctx.half_bytes - kUnit) * nsh, (u64)kUnit * nsh);
  u64 chunk = c->t_chunk ? floor_to(c->t_chunk, (u64)kUnit * nsh) : floor_to(kDefChunkBcast, (u64)kUnit * nsh);
  if (chunk < (u64)kUnit * nsh) chunk = (u64)kUnit * nsh;
  if (chunk > cap) chunk = cap;
  chunk = fit_chunk(chunk, bytes, (u64)kUnit * nsh);
  p.op.chunk_bytes = chunk;
  p.op.nchunks = (u32)ceil_div(bytes, chunk);
  p.nblocks = pick_blocks(c, ceil_div(ceil_div(chunk, kUnit), nsh), emu_world);
  return p;
}

// Launches items [i0, i1) as one table.  items[r][k]: rank r's view of item k (real rank: r = 0).
// flat: the scatter pipeline over the packed table (all items share one root), any size;
// otherwise the direct push, whose packed size must fit one staging half.
template <int NCTX>
int launch_table(nbd_comm* c0, const RankCtx* ctx, int world, const nbd_bcast_item* const* items, int i0, int i1, bool flat,
                 cudaStream_t s) {
  static thread_local PackTable<NCTX> pk;  // 5-22 KiB: keep it off the stack
  memset(&pk, 0, sizeof(pk));
  for (int r = 0; r < NCTX && r < world; ++r) pk.ctx[r] = ctx[r];
  u64 units = 0;
  for (int k = i0; k < i1; ++k) {
    const int j = k - i0;
    pk.d.first[j] = (u32)units;
    pk.d.bytes[j] = (u64)items[0][k].bytes;
    pk.d.root[j] = (unsigned char)items[0][k].root;
    for (int r = 0; r < NCTX && r < world; ++r) pk.buf[r][j] = (char*)items[r][k].buf;
    units += ceil_div((u64)items[0][k].bytes, kUnit);
  }
  if (units > 0xffffffffull) return fail(NBD_ERR_INVALID_ARGUMENT, "broadcast table of %llu bytes: more than 64 GiB", (unsigned long long)(units * kUnit));
  pk.d.nitems = i1 - i0;
  pk.d.first[pk.d.nitems] = (u32)units;
  const int emu_world = NCTX == 1 ? 0 : world;
  Plan p;
  memset(&p, 0, sizeof(p));
  const void* fn;
  void* args[2] = {(void*)&pk, (void*)&p.op};  // k_bcast_table takes the first only
  if (flat) {
    p = plan_flat(c0, units * kUnit, items[0][i0].root, emu_world);
    fn = (const void*)k_bcast_flat<NCTX>;
  } else {
    p.nblocks = pick_blocks(c0, units, emu_world);
    fn = (const void*)k_bcast_table<NCTX>;
  }
  if (NCTX == 1) {
    NBD_CUDA(cudaLaunchKernel(fn, dim3(p.nblocks, 1, 1), dim3(kThreads, 1, 1), args, 0, s));
  } else {
    NBD_CUDA(cudaLaunchCooperativeKernel(fn, dim3(p.nblocks, world, 1), dim3(kThreads, 1, 1), args, 0, s));
  }
  g_launches.fetch_add(1);
  return NBD_OK;
}

// A table beyond this size whose items share one root runs as a flat table (scatter pipeline);
// the direct push of the small table would make the root send (N-1) copies of it.
static const u64 kFlatMinBytes = 256ull << 10;

// Cuts the item list into launches.  The list is taken kTableMax items at a time; a run whose
// items share a root and together exceed kFlatMinBytes is one flat table.  Any other run is cut
// into direct-push tables (packed size <= one staging half), and an item too large for one of
// those is handed to `big` (the ordinary broadcast).  table(i0, i1, flat) launches items [i0, i1).
template <typename Table, typename Big>
int run_tables(const nbd_comm* c0, const nbd_bcast_item* items, int n, Table table, Big big) {
  const u64 cap_units = c0->ctx.half_bytes / kUnit;
  for (int a = 0; a < n; a += kTableMax) {
    const int b = a + kTableMax < n ? a + kTableMax : n;
    bool one_root = true;
    u64 total = 0;
    for (int k = a; k < b; ++k) {
      one_root = one_root && items[k].root == items[a].root;
      total += (u64)items[k].bytes;
    }
    if (one_root && total > kFlatMinBytes) {
      int rc = table(a, b, true);
      if (rc != NBD_OK) return rc;
      continue;
    }
    int i0 = a;
    u64 units = 0;
    for (int k = a; k <= b; ++k) {
      const u64 u = k < b ? ceil_div((u64)items[k].bytes, kUnit) : 0;
      const bool oversize = k < b && u > cap_units / 2;
      if (k == b || oversize || units + u > cap_units) {
        if (k > i0) {
          int rc = table(i0, k, false);
          if (rc != NBD_OK) return rc;
        }
        i0 = k;
        units = 0;
        if (oversize) {
          int rc = big(k);
          if (rc != NBD_OK) return rc;
          i0 = k + 1;
          continue;
        }
      }
      units += u;
    }
  }
  return NBD_OK;
}

int check_items(const nbd_bcast_item* items, int n, int world) {
  for (int k = 0; k < n; ++k) {
    if (items[k].root < 0 || items[k].root >= world) return fail(NBD_ERR_INVALID_ARGUMENT, "table item %d: root %d not in [0,%d)", k, items[k].root, world);
    if (items[k].bytes && !items[k].buf) return fail(NBD_ERR_INVALID_ARGUMENT, "table item %d: NULL buffer", k);
  }
  return NBD_OK;
}

}  // namespace

extern "C" int nbd_broadcast_table(nbd_comm_t c, const nbd_bcast_item* items, int n, void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_broadcast_table with emulated communicators");
  if (n < 0 || (n > 0 && !items)) return fail(NBD_ERR_INVALID_ARGUMENT, "bad item list");
  if ((rc = check_items(items, n, c->world)) != NBD_OK) return rc;
  if (c->world == 1 || n == 0) return NBD_OK;
  DeviceGuard g(c->device);
  std::vector<nbd_bcast_item> live;  // empty broadcasts are no-ops
  for (int k = 0; k < n; ++k)
    if (items[k].bytes) live.push_back(items[k]);
  const nbd_bcast_item* rows[1] = {live.data()};
  bool flat_seen = false;
  rc = run_tables(
      c, live.data(), (int)live.size(),
      [&](int i0, int i1, bool flat) {
        flat_seen = flat_seen || flat;
        return launch_table<1>(c, &c->ctx, c->world, rows, i0, i1, flat, (cudaStream_t)stream);
      },
      [&](int k) { return nbd_broadcast(c, live[(size_t)k].buf, live[(size_t)k].bytes, NBD_BYTES, live[(size_t)k].root, stream); });
  c->last_algo = flat_seen ? "bcast_flat" : "bcast_table";
  return rc;
}

extern "C" int nbd_emu_broadcast_table(nbd_comm_t any, const nbd_bcast_item* const* items, int n, void* stream) {
  if (!any) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (!any->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "not an emulated communicator");
  if (n < 0 || (n > 0 && !items)) return fail(NBD_ERR_INVALID_ARGUMENT, "bad item list");
  EmuGroup* g = any->group;
  nbd_comm* c0 = g->comms[0];
  int rc;
  RankCtx ctx[kMaxRanks];
  for (int r = 0; r < g->world; ++r) {
    if ((rc = check_usable(g->comms[r])) != NBD_OK) return rc;
    if (!items[r]) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL item list for rank %d", r);
    if ((rc = check_items(items[r], n, g->world)) != NBD_OK) return rc;
    ctx[r] = g->comms[r]->ctx;
    for (int k = 0; k < n; ++k)
      if (items[r][k].bytes != items[0][k].bytes || items[r][k].root != items[0][k].root)
        return fail(NBD_ERR_MISMATCH, "table item %d differs between rank 0 and rank %d", k, r);
  }
  if (g->world == 1 || n == 0) return NBD_OK;
  DeviceGuard dg(c0->device);
  mute_rank(g, ctx);
  // drop empty items from every rank's row alike
  std::vector<nbd_bcast_item> live[kMaxRanks];
  const nbd_bcast_item* rows[kMaxRanks] = {};
  for (int r = 0; r < g->world; ++r) {
    for (int k = 0; k < n; ++k)
      if (items[0][k].bytes) live[r].push_back(items[r][k]);
    rows[r] = live[r].data();
  }
  bool flat_seen = false;
  rc = run_tables(
      c0, live[0].data(), (int)live[0].size(),
      [&](int i0, int i1, bool flat) {
        flat_seen = flat_seen || flat;
        return launch_table<kMaxRanks>(c0, ctx, g->world, rows, i0, i1, flat, (cudaStream_t)stream);
      },
      [&](int k) {
        void* bufs[kMaxRanks];
        for (int r = 0; r < g->world; ++r) bufs[r] = live[r][(size_t)k].buf;
        return nbd_emu_broadcast(any, bufs, live[0][(size_t)k].bytes, NBD_BYTES, live[0][(size_t)k].root, stream);
      });
  set_algo(c0, flat_seen ? "bcast_flat" : "bcast_table");
  return rc;
}

// ---------------------------------------------------------------------------
// point-to-point batches (nbd_p2p.cuh)
// ---------------------------------------------------------------------------
namespace {

struct LaneList {
  int peer = 0, send = 0;
  std::vector<P2PMsg> msgs;
};

// Groups a rank's operations into lanes (direction, peer), messages in issue order.  Local
// copies (peer == rank) are not part of the C ABI: callers do those themselves.
int build_lanes(int rank, int world, const nbd_p2p_op* ops, int nops, std::vector<LaneList>* lanes) {
  for (int i = 0; i < nops; ++i) {
    const nbd_p2p_op& o = ops[i];
    if (o.kind != NBD_P2P_SEND && o.kind != NBD_P2P_RECV) return fail(NBD_ERR_INVALID_ARGUMENT, "p2p op %d: kind %d is neither send nor recv", i, o.kind);
    if (o.peer < 0 || o.peer >= world || o.peer == rank)
      return fail(NBD_ERR_INVALID_ARGUMENT, "p2p op %d: peer %d is not another rank of this %d-rank group", i, o.peer, world);
    if (o.bytes == 0) continue;  // both sides skip empty messages
    if (!o.buf) return fail(NBD_ERR_INVALID_ARGUMENT, "p2p op %d: NULL buffer", i);
    LaneList* L = nullptr;
    for (auto& l : *lanes)
      if (l.peer == o.peer && l.send == (o.kind == NBD_P2P_SEND)) L = &l;
    if (!L) {
      lanes->emplace_back();
      L = &lanes->back();
      L->peer = o.peer;
      L->send = (o.kind == NBD_P2P_SEND);
    }
    L->msgs.push_back(P2PMsg{(char*)o.buf, (u64)o.bytes});
  }
  return NBD_OK;
}

// CTAs per lane: a power of two (it must divide the 32 virtual spans), as many as stay co-resident.
int lane_ctas(int cap, int nlanes) {
  int g = 1;
  while (g * 2 <= kP2PVirt && g * 2 * nlanes <= cap) g *= 2;
  return g;
}

// Fills launch `round` (messages [round*kP2PLaneMsgs, +kP2PLaneMsgs) of every lane); returns CTAs used.
int fill_p2p_rank(P2PRank* pr, const std::vector<LaneList>& lanes, size_t round, int g) {
  memset(pr, 0, sizeof(*pr));
  int cta = 0;
  for (const auto& l : lanes) {
    const size_t first = round * kP2PLaneMsgs;
    if (first >= l.msgs.size()) continue;
    P2PLane& L = pr->lane[pr->nlanes++];
    L.peer = l.peer;
    L.send = l.send;
    L.cta0 = cta;
    L.nctas = g;
    cta += g;
    for (size_t m = first; m < l.msgs.size() && m < first + kP2PLaneMsgs; ++m) L.msg[L.nmsgs++] = l.msgs[m];
  }
  return cta;
}

size_t p2p_rounds(const std::vector<LaneList>& lanes) {
  size_t r = 0;
  for (const auto& l : lanes) r = std::max(r, (l.msgs.size() + kP2PLaneMsgs - 1) / kP2PLaneMsgs);
  return r;
}

}  // namespace

extern "C" int nbd_p2p_batch(nbd_comm_t c, const nbd_p2p_op* ops, int nops, void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_p2p_batch with emulated communicators");
  if (nops < 0 || (nops > 0 && !ops)) return fail(NBD_ERR_INVALID_ARGUMENT, "bad p2p op list");
  std::vector<LaneList> lanes;
  if ((rc = build_lanes(c->rank, c->world, ops, nops, &lanes)) != NBD_OK) return rc;
  if (lanes.empty()) return NBD_OK;
  if (lanes.size() > (size_t)kP2PMaxLanes) return fail(NBD_ERR_INTERNAL, "too many p2p lanes");
  DeviceGuard g(c->device);
  int cap = c->num_sms * kMinBlocksPerSM;
  if (c->t_blocks && (int)c->t_blocks < cap) cap = (int)c->t_blocks;
  const int gl = lane_ctas(cap, (int)lanes.size());
  const size_t rounds = p2p_rounds(lanes);
  for (size_t r = 0; r < rounds; ++r) {
    PackP2P<1> pk;
    memset(&pk, 0, sizeof(pk));
    pk.ctx[0] = c->ctx;
    const int ctas = fill_p2p_rank(&pk.r[0], lanes, r, gl);
    if (ctas == 0) continue;
    void* args[1] = {(void*)&pk};
    NBD_CUDA(cudaLaunchKernel((const void*)k_p2p_batch<1>, dim3(ctas, 1, 1), dim3(kThreads, 1, 1), args, 0, (cudaStream_t)stream));
    g_launches.fetch_add(1);
  }
  c->last_algo = "p2p_batch";
  return NBD_OK;
}

extern "C" int nbd_send(nbd_comm_t c, const void* buf, size_t bytes, int peer, void* stream) {
  nbd_p2p_op op = {NBD_P2P_SEND, peer, const_cast<void*>(buf), bytes};
  return nbd_p2p_batch(c, &op, 1, stream);
}

extern "C" int nbd_recv(nbd_comm_t c, void* buf, size_t bytes, int peer, void* stream) {
  nbd_p2p_op op = {NBD_P2P_RECV, peer, buf, bytes};
  return nbd_p2p_batch(c, &op, 1, stream);
}

extern "C" int nbd_emu_p2p_batch(nbd_comm_t any, const nbd_p2p_op* const* ops, const int* nops, void* stream) {
  if (!any) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (!any->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "not an emulated communicator");
  if (!ops || !nops) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL op lists");
  EmuGroup* g = any->group;
  nbd_comm* c0 = g->comms[0];
  DeviceGuard dg(c0->device);
  std::vector<LaneList> lanes[kMaxRanks];
  size_t maxl = 0, rounds = 0;
  int rc;
  for (int r = 0; r < g->world; ++r) {
    if ((rc = check_usable(g->comms[r])) != NBD_OK) return rc;
    if ((rc = build_lanes(r, g->world, ops[r], nops[r], &lanes[r])) != NBD_OK) return rc;
    maxl = std::max(maxl, lanes[r].size());
    rounds = std::max(rounds, p2p_rounds(lanes[r]));
  }
  if (maxl == 0) return NBD_OK;
  int cap = c0->num_sms * kMinBlocksPerSM / g->world;
  if (c0->t_blocks && (int)c0->t_blocks < cap) cap = (int)c0->t_blocks;
  if (cap < 1) cap = 1;
  if ((int)maxl > cap) return fail(NBD_ERR_UNSUPPORTED, "%zu lanes per emulated rank do not fit %d co-resident CTAs", maxl, cap);
  const int gl = lane_ctas(cap, (int)maxl);
  for (size_t rd = 0; rd < rounds; ++rd) {
    PackP2P<kMaxRanks> pk;
    memset(&pk, 0, sizeof(pk));
    int ctas = 0;
    for (int r = 0; r < g->world; ++r) {
      pk.ctx[r] = g->comms[r]->ctx;
      ctas = std::max(ctas, fill_p2p_rank(&pk.r[r], lanes[r], rd, gl));
    }
    if (ctas == 0) continue;
    void* args[1] = {(void*)&pk};
    NBD_CUDA(cudaLaunchCooperativeKernel((const void*)k_p2p_batch<kMaxRanks>, dim3(ctas, g->world, 1), dim3(kThreads, 1, 1), args, 0,
                                         (cudaStream_t)stream));
    g_launches.fetch_add(1);
  }
  set_algo(c0, "p2p_batch");
  return NBD_OK;
}

// ---------------------------------------------------------------------------
// nbd_local_all_reduce_noalloc — all-reduce of R buffers on ONE GPU without any allocation
// ---------------------------------------------------------------------------
// For hosts that must not allocate when they call us: a process whose device memory is
// already parcelled out to another allocator (a serving stack, torch's caching allocator
// holding nearly all of HBM) where a driver allocation may fail or evict that allocator's
// cache, and latency-sensitive loops for which the device-wide stall of cudaMalloc/cudaFree
// is not acceptable.  The emulated-rank all-reduce normally takes its flags, staging and
// clocks from arenas that nbd_emu_create() cudaMallocs; here they live in the file-scope
// array below, which every GPU that loads this library carries until the process exits.
//
// Layout of g_noalloc_scratch (offsets in bytes):
//   [0, R*kFlagBytes)                     flag arrays, one per rank
//   [.., + R*256)                         LocalState per rank
//   [.., + R*256)                         mailbox per rank (device-resident stand-in)
//   [kNaStageOff, + R*kNaStageBytes)      staging, one per rank (two halves each)
// The control part (everything before kNaStageOff) is zeroed on the caller's stream at the
// start of every call: flags and clocks must start equal on all ranks, and a previous call
// with another world size leaves them unequal.  Staging needs no clearing (always written
// before it is read).  All calls share the one array, so two calls whose GPU work overlaps
// (different streams, no ordering between them) corrupt each other.
namespace {

constexpr size_t kNaStageBytes = 1u << 20;  // per rank; half of it per pipeline step
constexpr size_t kNaCtlBytes = (size_t)kMaxRanks * (kFlagBytes + 256 + 256);
constexpr size_t kNaStageOff = (kNaCtlBytes + 255) / 256 * 256;
constexpr size_t kNaTotalBytes = kNaStageOff + (size_t)kMaxRanks * kNaStageBytes;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaTotalBytes];

constexpr int kNaMaxDevices = 64;
struct NoAllocDevice {
  bool ready = false;
  unsigned char* base = nullptr;  // DEVICE address of g_noalloc_scratch on this device
  nbd_comm plan;                  // host-only planning record (tunables, SM count); owns nothing
};
static NoAllocDevice g_na_dev[kNaMaxDevices];
static std::mutex g_na_mu;

// Once per device, with that device current.  g_noalloc_scratch named in host code is only a
// host-side handle of the symbol; the address kernels can use comes from the runtime.
static int noalloc_setup(int device, NoAllocDevice** out) {
  std::lock_guard<std::mutex> lk(g_na_mu);
  NoAllocDevice& d = g_na_dev[device];
  if (!d.ready) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      cudaGetLastError();
      return fail(NBD_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", device,
                  cudaGetErrorString(e));
    }
    int sms = 0, coop = 0;
    NBD_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    NBD_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device));
    if (!coop) return fail(NBD_ERR_UNSUPPORTED, "device %d lacks cooperative launch", device);
    d.base = static_cast<unsigned char*>(ptr);
    d.plan.device = device;
    d.plan.num_sms = sms;
    d.plan.emu = true;
    d.plan.connected = true;
    d.ready = true;
  }
  *out = &d;
  return NBD_OK;
}

}  // namespace

extern "C" int nbd_local_all_reduce_noalloc(const void* const* sends, void* const* recvs, int world_size, size_t count,
                                            int dtype, int op, void* stream) {
  int rc = check_reduce_args(dtype, op);
  if (rc != NBD_OK) return rc;
  if (world_size < 1 || world_size > kMaxRanks)
    return fail(NBD_ERR_INVALID_ARGUMENT, "world_size %d not in [1,%d]", world_size, kMaxRanks);
  const u64 bytes = (u64)count * nbd_dtype_size(dtype);
  if (bytes > (u64)NBD_NOALLOC_MAX_BYTES)
    return fail(NBD_ERR_INVALID_ARGUMENT,
                "nbd_local_all_reduce_noalloc: %llu bytes per rank exceeds the fixed scratch limit of %llu bytes "
                "(NBD_NOALLOC_MAX_BYTES); use nbd_emu_all_reduce for larger messages",
                (unsigned long long)bytes, (unsigned long long)NBD_NOALLOC_MAX_BYTES);
  if (count == 0) return NBD_OK;
  if (!sends || !recvs) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL pointer array");
  for (int r = 0; r < world_size; ++r)
    if (!sends[r] || !recvs[r]) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer for rank %d", r);
  int device = 0;
  NBD_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= kNaMaxDevices) return fail(NBD_ERR_INVALID_ARGUMENT, "device %d out of range", device);
  NoAllocDevice* d = nullptr;
  if ((rc = noalloc_setup(device, &d)) != NBD_OK) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (world_size == 1) {
    if (sends[0] != recvs[0]) NBD_CUDA(cudaMemcpyAsync(recvs[0], sends[0], bytes, cudaMemcpyDeviceToDevice, s));
    return NBD_OK;
  }

  std::lock_guard<std::mutex> lk(g_na_mu);  // d->plan is shared host state
  unsigned char* base = d->base;
  Pack<kMaxRanks> pk;
  memset(&pk, 0, sizeof(pk));
  unsigned char* states = base + (size_t)kMaxRanks * kFlagBytes;
  unsigned char* mboxes = states + (size_t)kMaxRanks * 256;
  for (int r = 0; r < world_size; ++r) {
    RankCtx& x = pk.ctx[r];
    x.rank = r;
    x.world = world_size;
    for (int p = 0; p < world_size; ++p) {
      x.flags[p] = reinterpret_cast<u64*>(base + (size_t)p * kFlagBytes);
      x.stage[p] = reinterpret_cast<char*>(base + kNaStageOff + (size_t)p * kNaStageBytes);
    }
    x.state = reinterpret_cast<LocalState*>(states + (size_t)r * 256);
    x.mbox = reinterpret_cast<HostMailbox*>(mboxes + (size_t)r * 256);
    x.half_bytes = (kNaStageBytes / 2) / kUnit * kUnit;
    x.timeout_ns = 10ull * 1000000000ull;
    pk.buf[r].send = static_cast<const char*>(sends[r]);
    pk.buf[r].recv = static_cast<char*>(recvs[r]);
  }
  d->plan.world = world_size;
  d->plan.ctx.half_bytes = pk.ctx[0].half_bytes;
  Plan p = plan_all_reduce(&d->plan, bytes, op, world_size);
  // flags, clocks and mailboxes start from zero on every call (see the layout note above)
  NBD_CUDA(cudaMemsetAsync(base, 0, kNaCtlBytes, s));
  return launch_reduce<kMaxRanks>(p.op.algo == 1 ? K_AR1 : K_AR2, dtype, op, p.nblocks, world_size, s, pk, &p.op);
}

// ---------------------------------------------------------------------------
// symmetric memory (see include/nbdist_b200.h)
// ---------------------------------------------------------------------------
extern "C" int nbd_mem_alloc(nbd_comm_t c, size_t bytes, void** ptr, void* handle_out) {
  if (!c || !ptr || !handle_out) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL argument");
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "emulated ranks share one address space; no registration needed");
  if (bytes == 0) return fail(NBD_ERR_INVALID_ARGUMENT, "cannot allocate 0 bytes");
  DeviceGuard g(c->device);
  const size_t rounded = (bytes + kStageAlign - 1) / kStageAlign * kStageAlign;
  if (nvls_reduce_policy(c)) {
    // In-switch reduction is on: hand out memory that is bound into the multicast object, so
    // that all_reduce on it needs no staging.  Peers do not map it (multimem.* reaches every
    // copy through the switch); every other collective treats it as an ordinary buffer.
    char* mc = nullptr;
    char* p = nvls_pool_alloc(c, rounded, &mc);
    if (p) {
      Handle h;
      memset(&h, 0, sizeof(h));
      h.magic = kMagic;
      h.version = NBD_VERSION;
      h.rank = c->rank;
      h.world = c->world;
      h.pid = (int64_t)getpid();
      h.hostid = host_id();
      h.arena_bytes = rounded;
      h.stage_bytes = (uint64_t)c->regs.size();
      h.nvls_nonce = (uint64_t)(p - c->nvls.local) | (1ull << 63);  // pool offset: must match on every rank
      nbd_comm::Registration reg;
      reg.base = p;
      reg.bytes = rounded;
      reg.peer[c->rank] = p;
      reg.mc = mc;
      reg.connected = false;
      c->regs.push_back(reg);
      memset(handle_out, 0, NBD_HANDLE_BYTES);
      memcpy(handle_out, &h, sizeof(h));
      *ptr = p;
      return NBD_OK;
    }
  }
  char* base = nullptr;
  NBD_CUDA(cudaMalloc((void**)&base, rounded));
  Handle h;
  memset(&h, 0, sizeof(h));
  h.magic = kMagic;
  h.version = NBD_VERSION;
  h.rank = c->rank;
  h.world = c->world;
  h.pid = (int64_t)getpid();
  h.hostid = host_id();
  h.arena_bytes = rounded;
  h.stage_bytes = (uint64_t)c->regs.size();  // allocation index: must match on every rank
  if (c->world > 1) {
    cudaError_t e = cudaIpcGetMemHandle(&h.ipc, base);
    if (e != cudaSuccess) {
      cudaFree(base);
      cudaGetLastError();
      return fail(NBD_ERR_CUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
    }
  }
  nbd_comm::Registration reg;
  reg.base = base;
  reg.bytes = rounded;
  reg.peer[c->rank] = base;
  reg.connected = (c->world == 1);
  c->regs.push_back(reg);
  memset(handle_out, 0, NBD_HANDLE_BYTES);
  memcpy(handle_out, &h, sizeof(h));
  *ptr = base;
  return NBD_OK;
}

extern "C" int nbd_mem_connect(nbd_comm_t c, void* ptr, const void* all_handles) {
  if (!c || !ptr || !all_handles) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL argument");
  int k = -1;
  for (size_t i = 0; i < c->regs.size(); ++i)
    if (c->regs[i].base == ptr) k = (int)i;
  if (k < 0) return fail(NBD_ERR_INVALID_ARGUMENT, "pointer was not returned by nbd_mem_alloc on this communicator");
  auto& reg = c->regs[(size_t)k];
  if (reg.connected) return NBD_OK;
  if (reg.mc) {  // multicast pool: nothing to map, only check that every rank got the same piece
    const uint64_t mine = (uint64_t)(reg.base - c->nvls.local) | (1ull << 63);
    for (int p = 0; p < c->world; ++p) {
      Handle h;
      memcpy(&h, (const char*)all_handles + (size_t)p * NBD_HANDLE_BYTES, sizeof(h));
      if (h.magic != kMagic || h.rank != p || h.world != c->world) return fail(NBD_ERR_INVALID_ARGUMENT, "memory handle %d is malformed", p);
      if (h.nvls_nonce != mine || h.arena_bytes != reg.bytes)
        return fail(NBD_ERR_MISMATCH, "rank %d holds a different piece of the multicast pool: nbd_mem_alloc/free must be called in the "
                    "same order with the same sizes on every rank", p);
    }
    reg.connected = true;
    return NBD_OK;
  }
  DeviceGuard g(c->device);
  for (int p = 0; p < c->world; ++p) {
    if (p == c->rank) continue;
    Handle h;
    memcpy(&h, (const char*)all_handles + (size_t)p * NBD_HANDLE_BYTES, sizeof(h));
    if (h.magic != kMagic || h.rank != p || h.world != c->world)
      return fail(NBD_ERR_INVALID_ARGUMENT, "memory handle %d is malformed", p);
    if (h.stage_bytes != (uint64_t)k || h.arena_bytes != reg.bytes)
      return fail(NBD_ERR_MISMATCH, "rank %d allocated buffer #%llu of %llu bytes, this rank #%d of %llu bytes: "
                  "nbd_mem_alloc must be called in the same order with the same size on every rank", p,
                  (unsigned long long)h.stage_bytes, (unsigned long long)h.arena_bytes, k, (unsigned long long)reg.bytes);
    void* mapped = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&mapped, h.ipc, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(NBD_ERR_NO_PEER_ACCESS, "cudaIpcOpenMemHandle(buffer of rank %d): %s", p, cudaGetErrorString(e));
    }
    reg.peer[p] = mapped;
  }
  reg.connected = true;
  return NBD_OK;
}

extern "C" int nbd_mem_free(nbd_comm_t c, void* ptr) {
  if (!c || !ptr) return NBD_OK;
  DeviceGuard g(c->device);
  for (auto& reg : c->regs) {
    if (reg.base != ptr) continue;
    if (reg.mc) {
      nvls_pool_free(c, reg.base);
      reg.base = nullptr;
      reg.mc = nullptr;
      reg.bytes = 0;
      reg.connected = false;
      return NBD_OK;
    }
    for (int p = 0; p < c->world; ++p)
      if (p != c->rank && reg.peer[p]) {
        cudaIpcCloseMemHandle(reg.peer[p]);
        reg.peer[p] = nullptr;
      }
    cudaFree(reg.base);
    reg.base = nullptr;  // the slot stays so that allocation indices remain aligned across ranks
    reg.bytes = 0;
    reg.connected = false;
    return NBD_OK;
  }
  return fail(NBD_ERR_INVALID_ARGUMENT, "pointer was not returned by nbd_mem_alloc on this communicator");
}

extern "C" int nbd_comm_nvls_status(nbd_comm_t c, char* why, size_t why_bytes) {
  if (!c) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  if (why && why_bytes) snprintf(why, why_bytes, "%s", c->nvls.why);
  return c->nvls.enabled ? 1 : 0;
}

// 1 if the most recent auto-registered collective on `comm` ran zero-copy, 0 if it ran staged,
// negative on error.  Synchronises the device (tests and diagnostics only).
extern "C" int nbd_comm_last_zc_decision(nbd_comm_t c) {
  if (!c) return fail(NBD_ERR_INVALID_ARGUMENT, "comm is NULL");
  DeviceGuard g(c->device);
  NBD_CUDA(cudaDeviceSynchronize());
  unsigned int ok = 0;
  NBD_CUDA(cudaMemcpy(&ok, &reinterpret_cast<LocalBlock*>(c->state)->dec.ok, sizeof(ok), cudaMemcpyDeviceToHost));
  return (int)ok;
}

// ---------------------------------------------------------------------------
// host-buffer all_reduce (see include/nbdist_b200.h): chunked H2D -> collective -> D2H pipeline
// ---------------------------------------------------------------------------
static const size_t kDefHostChunk = 32ull << 20;

static int hostpipe_init(HostPipe& hp, int nranks, size_t chunk) {
  if (hp.ready && hp.nranks == nranks && hp.chunk == chunk) return NBD_OK;
  hostpipe_free(hp);
  hp.nranks = nranks;
  hp.chunk = chunk;
  NBD_CUDA(cudaStreamCreateWithFlags(&hp.h2d, cudaStreamNonBlocking));
  NBD_CUDA(cudaStreamCreateWithFlags(&hp.comp, cudaStreamNonBlocking));
  NBD_CUDA(cudaStreamCreateWithFlags(&hp.d2h, cudaStreamNonBlocking));
  NBD_CUDA(cudaEventCreateWithFlags(&hp.ev_start, cudaEventDisableTiming));
  for (int k = 0; k < kPipeSlots; ++k) {
    NBD_CUDA(cudaEventCreateWithFlags(&hp.ev_h2d[k], cudaEventDisableTiming));
    NBD_CUDA(cudaEventCreateWithFlags(&hp.ev_comp[k], cudaEventDisableTiming));
    NBD_CUDA(cudaEventCreateWithFlags(&hp.ev_d2h[k], cudaEventDisableTiming));
    for (int r = 0; r < nranks; ++r) NBD_CUDA(cudaMalloc((void**)&hp.buf[r][k], chunk));
  }
  hp.ready = true;
  return NBD_OK;
}

// One pipelined pass.  `reduce(slot, bytes, stream)` enqueues the collective on the bounce
// buffers of `slot`.
template <typename F>
static int hostpipe_run(HostPipe& hp, const void* const* hsend, void* const* hrecv, u64 bytes, cudaStream_t user, F reduce) {
  NBD_CUDA(cudaEventRecord(hp.ev_start, user));
  NBD_CUDA(cudaStreamWaitEvent(hp.h2d, hp.ev_start, 0));
  NBD_CUDA(cudaStreamWaitEvent(hp.comp, hp.ev_start, 0));
  const u64 nchunks = ceil_div(bytes, (u64)hp.chunk);
  int last_slot = 0;
  for (u64 k = 0; k < nchunks; ++k) {
    const int slot = (int)(k % kPipeSlots);
    const u64 off = k * hp.chunk;
    const u64 n = bytes - off < hp.chunk ? bytes - off : hp.chunk;
    if (hp.used[slot]) NBD_CUDA(cudaStreamWaitEvent(hp.h2d, hp.ev_d2h[slot], 0));  // bounce buffer drained
    for (int r = 0; r < hp.nranks; ++r)
      NBD_CUDA(cudaMemcpyAsync(hp.buf[r][slot], (const char*)hsend[r] + off, n, cudaMemcpyHostToDevice, hp.h2d));
    NBD_CUDA(cudaEventRecord(hp.ev_h2d[slot], hp.h2d));
    NBD_CUDA(cudaStreamWaitEvent(hp.comp, hp.ev_h2d[slot], 0));
    int rc = reduce(slot, n, hp.comp);
    if (rc != NBD_OK) return rc;
    NBD_CUDA(cudaEventRecord(hp.ev_comp[slot], hp.comp));
    NBD_CUDA(cudaStreamWaitEvent(hp.d2h, hp.ev_comp[slot], 0));
    for (int r = 0; r < hp.nranks; ++r)
      NBD_CUDA(cudaMemcpyAsync((char*)hrecv[r] + off, hp.buf[r][slot], n, cudaMemcpyDeviceToHost, hp.d2h));
    NBD_CUDA(cudaEventRecord(hp.ev_d2h[slot], hp.d2h));
    hp.used[slot] = true;
    last_slot = slot;
  }
  NBD_CUDA(cudaStreamWaitEvent(user, hp.ev_d2h[last_slot], 0));  // d2h is in order: all chunks are back
  return NBD_OK;
}

static size_t host_chunk_bytes(const nbd_comm* c) {
  size_t chunk = c->t_host_chunk ? (size_t)c->t_host_chunk : kDefHostChunk;
  chunk = (chunk + 255) / 256 * 256;
  return chunk;
}

extern "C" int nbd_all_reduce_host(nbd_comm_t c, const void* host_send, void* host_recv, size_t count, int dtype, int op,
                                   void* stream) {
  int rc = check_usable(c);
  if (rc != NBD_OK) return rc;
  if (c->emu) return fail(NBD_ERR_INVALID_ARGUMENT, "use nbd_emu_all_reduce_host with emulated communicators");
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (count == 0) return NBD_OK;
  if (!host_send || !host_recv) return fail(NBD_ERR_INVALID_ARGUMENT, "NULL buffer");
  DeviceGuard g(c->device);
  const size_t es = nbd_dtype_size(dtype);
  if ((rc = hostpipe_init(c->hp, 1, host_chunk_bytes(c))) != NBD_OK) return rc;
  const void* hs[1] = {host_send};
  void* hr[1] = {host_recv};
  HostPipe& hp = c->hp;
  return hostpipe_run(hp, hs, hr, (u64)count * es, (cudaStream_t)stream, [&](int slot, u64 n, cudaStream_t s) {
    return nbd_all_reduce(c, hp.buf[0][slot], hp.buf[0][slot], (size_t)(n / es), dtype, op, s);
  });
}

extern "C" int nbd_emu_all_reduce_host(nbd_comm_t any, const void* const* host_sends, void* const* host_recvs, size_t count,
                                       int dtype, int op, void* stream) {
  int rc;
  if ((rc = check_reduce_args(dtype, op)) != NBD_OK) return rc;
  if (any && any->emu && count == 0) return NBD_OK;
  if ((rc = emu_check(any, host_sends, (const void* const*)host_recvs, true)) != NBD_OK) return rc;
  EmuGroup* g = any->group;
  DeviceGuard dg(g->comms[0]->device);
  const size_t es = nbd_dtype_size(dtype);
  if ((rc = hostpipe_init(g->hp, g->world, host_chunk_bytes(g->comms[0]))) != NBD_OK) return rc;
  HostPipe& hp = g->hp;
  return hostpipe_run(hp, host_sends, host_recvs, (u64)count * es, (cudaStream_t)stream, [&](int slot, u64 n, cudaStream_t s) {
    const void* sends[kMaxRanks];
    void* recvs[kMaxRanks];
    for (int r = 0; r < g->world; ++r) {
      sends[r] = hp.buf[r][slot];
      recvs[r] = hp.buf[r][slot];
    }
    return nbd_emu_all_reduce(any, sends, recvs, (size_t)(n / es), dtype, op, s);
  });
}

// ---------------------------------------------------------------------------
// nbd_plan_query — which algorithm / chunking / grid the planner picks, without a GPU
// ---------------------------------------------------------------------------
// kind: 0 all_reduce, 1 reduce_scatter (bytes = one shard), 2 all_gather (bytes = one shard),
// 3 broadcast, 4 flat broadcast table (bytes = packed size), 5 tile pipeline (all_reduce_pipe,
// broadcast_pipe, broadcast_relay: chunk_bytes = tile, nchunks = rounds, algo = slots per CTA).
// out: {algo, chunk_bytes, nchunks, nblocks, half_bytes, ll_off}.
extern "C" int nbd_plan_query(int world, size_t stage_bytes, int num_sms, int emu, int kind, size_t bytes, uint64_t* out) {
  if (!out || world < 2 || world > kMaxRanks || num_sms < 1) return fail(NBD_ERR_INVALID_ARGUMENT, "bad plan query");
  static std::mutex mu;
  std::lock_guard<std::mutex> lk(mu);
  nbd_comm c;  // planning record only: no device memory is touched
  c.world = world;
  c.num_sms = num_sms;
  c.emu = emu != 0;
  if (stage_bytes == 0) stage_bytes = kDefaultStage;
  c.stage_bytes = (stage_bytes + kStageAlign - 1) / kStageAlign * kStageAlign;
  c.ctx.ll_off = c.stage_bytes - ll_area_of(c.stage_bytes);
  c.ctx.ll_slot = 2 * ll_cap_of(c.stage_bytes);
  c.ctx.half_bytes = ((c.stage_bytes - ll_area_of(c.stage_bytes)) / 2) / kUnit * kUnit;
  Plan p;
  const int ew = emu ? world : 0;
  switch (kind) {
    case 0: p = plan_all_reduce(&c, bytes, NBD_SUM, ew); break;
    case 1: p = plan_reduce_scatter(&c, bytes, NBD_SUM, ew); break;
    case 2: p = plan_all_gather(&c, bytes, ew); break;
    case 3: p = plan_broadcast(&c, bytes, 0, ew); break;
    case 4: p = plan_flat(&c, bytes, 0, ew); break;
    case 5: {
      int rc = plan_pipe(&c, bytes, 2 * c.ctx.half_bytes, ew, &p);
      if (rc != NBD_OK) return rc;
      p.op.algo = p.op.depth;
      break;
    }
    default: return fail(NBD_ERR_INVALID_ARGUMENT, "unknown collective kind %d", kind);
  }
  out[0] = (uint64_t)p.op.algo;
  out[1] = p.op.chunk_bytes;
  out[2] = p.op.nchunks;
  out[3] = (uint64_t)p.nblocks;
  out[4] = c.ctx.half_bytes;
  out[5] = c.ctx.ll_off;
  return NBD_OK;
}
