// nbd_host.h — host-side records shared by the translation units of libnbdist_b200.so
// (nbd_lib.cu: communicators, planning, launches; nbd_nvls.cu: NVSwitch multicast set-up).
// No counterpart in the reference (which has no native code).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <atomic>
#include <map>
#include <set>
#include <vector>

#include "../../include/nbdist_b200.h"
#include "nbd_device.cuh"

namespace nbd {
int fail(int code, const char* fmt, ...);  // records the thread-local detail, returns code
extern std::atomic<uint64_t> g_launches;
}  // namespace nbd
using nbd::fail;
using nbd::g_launches;

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------

#define NBD_CUDA(call)                                                                      \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess)                                                                 \
      return fail(NBD_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                  __FILE__, __LINE__);                                                      \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool changed = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) changed = (cudaSetDevice(dev) == cudaSuccess);
  }
  ~DeviceGuard() {
    if (changed) cudaSetDevice(prev);
  }
};

// ---------------------------------------------------------------------------
// communicator
// ---------------------------------------------------------------------------
using namespace nbd;
struct EmuGroup;

// Host-buffer collectives (nbd_*_host): three streams and a ring of device bounce buffers so
// that the H2D copy of chunk k+1, the collective on chunk k and the D2H copy of chunk k-1
// overlap; PCIe is full duplex, so a step costs ~max(H2D, D2H) instead of their sum.
constexpr int kPipeSlots = 3;
struct HostPipe {
  bool ready = false;
  int nranks = 0;        // 1 for a real rank, R for an emulated group
  size_t chunk = 0;      // bytes per bounce buffer
  cudaStream_t h2d = nullptr, comp = nullptr, d2h = nullptr;
  cudaEvent_t ev_start = nullptr, ev_h2d[kPipeSlots] = {}, ev_comp[kPipeSlots] = {}, ev_d2h[kPipeSlots] = {};
  bool used[kPipeSlots] = {};
  char* buf[kMaxRanks][kPipeSlots] = {};
};

// NVSwitch multicast ("NVLS") staging of one rank: a VMM allocation bound into a multicast
// object shared by all ranks (nbd_nvls.cu).  `local` addresses this rank's own copy, `mcva`
// addresses every rank's copy at once and is only valid for multimem.* instructions.
struct NvlsState {
  bool enabled = false;     // set only when every rank of the communicator finished the set-up
  int listen_fd = -1;       // rank 0: the socket its peers connect to during nbd_comm_connect
  uint64_t nonce = 0;       // names that socket (abstract unix namespace); 0 = multicast not wanted
  size_t bytes = 0;         // bytes bound into the multicast object per rank: staging, then the symmetric pool
  size_t stage_bytes = 0;   // [0, stage_bytes): staging of the tile-pipeline kernels
  size_t pool_bytes = 0;    // [stage_bytes, +pool_bytes): symmetric memory handed out by nbd_mem_alloc
  struct PoolBlock {
    size_t off, len;
    bool used;
  };
  std::vector<PoolBlock> pool;  // address-ordered blocks of the symmetric pool
  size_t granularity = 0;
  unsigned long long mem = 0, mc = 0;  // CUmemGenericAllocationHandle of the staging / the multicast object
  bool have_mem = false, have_mc = false, added = false, bound = false;
  char* local = nullptr;
  char* mcva = nullptr;
  char why[192] = "not set up";  // why multicast is off (diagnostics: nbd_comm_nvls_status)
};

struct nbd_comm {
  int rank = 0, world = 1, device = 0;
  int num_sms = 0;
  bool emu = false, connected = false, dead = false;
  EmuGroup* group = nullptr;
  char* arena = nullptr;
  size_t arena_bytes = 0, stage_bytes = 0;
  LocalState* state = nullptr;
  HostMailbox* mbox = nullptr;      // host pointer (pinned, mapped)
  HostMailbox* mbox_dev = nullptr;  // device alias
  void* peer_base[kMaxRanks] = {};
  RankCtx ctx = {};
  // tunables (0 = default)
  uint64_t t_blocks = 0, t_ar1_max = 0, t_bcast_direct_max = 0, t_chunk = 0;
  uint64_t t_zc_off = 0, t_zc_min = 0;  // zero-copy switch / size threshold
  uint64_t t_zc_fused = 0, t_zc_tile = 0;  // all_reduce: 1 = fused loop, 2 = tiled; tile bytes (0 = whole span)
  struct Registration {                 // symmetric memory: same index on every rank
    char* base = nullptr;
    size_t bytes = 0;
    void* peer[kMaxRanks] = {};
    bool connected = false;
    char* mc = nullptr;                 // multicast alias (memory from the NVSwitch-multicast pool): multimem.* only
  };
  std::vector<Registration> regs;
  // auto-registration of ordinary allocations (exports: ours; imports: peers')
  uint64_t t_auto_off = 0, t_auto_min = 0;
  ExportSlot* export_mirror = nullptr;          // pinned host copy of our export slots
  int n_exports = 0;
  std::map<uint64_t, int> export_of;            // buffer id -> slot
  std::set<uint64_t> not_exportable;            // buffer ids cudaIpcGetMemHandle refused
  std::set<uint64_t> imported;                  // (peer << 32 | slot) already mapped
  std::vector<void*> import_ptrs;               // mappings to close at destroy
  uint32_t miss_seen = 0;
  cudaStream_t side = nullptr;                  // for reading peers' export slots
  ExportSlot* import_tmp = nullptr;             // pinned scratch for one slot
  uint64_t* imp_tmp = nullptr;                  // pinned ring of pointers handed to cudaMemcpyAsync
  int imp_tmp_n = 0;
  uint64_t t_host_chunk = 0;
  uint64_t t_ll_max = 0;  // LL all_reduce up to this many bytes (tunable "ll_max_bytes"; 1 = off)
  HostPipe hp;
  uint64_t timeout_ms = 600000;
  const char* last_algo = "none";
  NvlsState nvls;
  uint64_t t_nvls_reduce = 0;   // 1: float SUM/AVG all_reduce may use in-switch reduction (order != rank order)
  uint64_t t_nvls_off = 0;      // 1: never use the multicast kernels
  uint64_t t_nvls_min = 0;      // smallest message for the multicast kernels
  uint64_t t_pipe = 0;          // 1: force the tile-pipeline kernels (P2P transport) where they apply; 2: never;
                                // 3: force the relay broadcast (all_reduce as with 0)
  uint64_t t_relay_min = 0;     // broadcasts from this size on use the relay kernel (0 = default)
  uint64_t t_pipe_tile = 0;     // tile bytes per CTA per round (0 = planner's choice)
  uint64_t t_pipe_depth = 0;    // staging slots per CTA (0 = default)
  uint64_t t_pipe_mode = 0;     // bit 1: two reducing groups; bit 2: 8 loads in flight per thread (0 = default)
  uint64_t t_nvls_bcast = 0;    // 1: broadcast through multimem.st where multicast is set up
  uint64_t t_debug_mute = 0;    // tests: rank (value - 1) of an emulated group never signals its peers
};

// nbd_nvls.cu
int nvls_prepare(nbd_comm* c);                       // nbd_comm_create: rank 0 opens the rendezvous socket
void nvls_setup(nbd_comm* c, uint64_t nonce_rank0);  // nbd_comm_connect: collective; leaves nvls.enabled
void nvls_teardown(nbd_comm* c);
char* nvls_pool_alloc(nbd_comm* c, size_t bytes, char** mc_alias);  // nullptr: pool absent or exhausted
bool nvls_pool_free(nbd_comm* c, char* ptr);                        // false: not a pool pointer

struct EmuGroup {
  int world = 0;
  nbd_comm* comms[kMaxRanks] = {};
  HostPipe hp;
};

struct Handle {  // serialised into NBD_HANDLE_BYTES
  uint32_t magic, version;
  int32_t rank, world;
  int64_t pid;
  uint64_t hostid, arena_bytes, stage_bytes;
  char pci[32];
  cudaIpcMemHandle_t ipc;
  uint64_t nvls_nonce;  // rank 0: name of its multicast rendezvous socket (0 = none)
};
static_assert(sizeof(Handle) <= NBD_HANDLE_BYTES, "handle too large");
static const uint32_t kMagic = 0x4e424432u;  // "NBD2"

