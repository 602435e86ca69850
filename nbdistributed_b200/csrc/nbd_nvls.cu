// nbd_nvls.cu — NVSwitch multicast ("NVLS") set-up for one communicator.
//
// What it gives the kernels: per rank a staging buffer in its own HBM (`nvls.local`) that is
// also bound, at the same offset, into ONE multicast object shared by every rank
// (`nvls.mcva`).  `multimem.ld_reduce [mcva + x]` makes the switch fetch and sum the N copies
// of x; `multimem.st [mcva + x]` makes the switch write all N copies.  That is the transport of
// the tile-pipeline kernels in nbd_pipe.cuh; what the reference's users get
// today from backend "nccl" (ref:src/nbdistributed/worker.py:145) on an NVSwitch box.
//
// How ranks (one process each) come to share the object without a new rendezvous service: a
// multicast handle can only travel between processes as a POSIX file descriptor, and an fd can
// only travel over a unix socket.  Rank 0 therefore listens on an ABSTRACT unix socket whose
// name is a random nonce carried in its NBD handle (which the caller already all-gathers over
// its store for nbd_comm_connect).  Every peer connects and the steps below run in lock-step;
// after each step the peers report a status and rank 0 answers with the AND, so either all
// ranks end up with multicast enabled or none does:
//
//   HELLO   every rank allocated its staging (cuMemCreate) and supports multicast
//   MC      rank 0: cuMulticastCreate, export, send the fd (SCM_RIGHTS); peers import it
//   ADDED   every rank: cuMulticastAddDevice
//   BOUND   every rank: cuMulticastBindMem + map the multicast address range
//
// Nothing here can fail the communicator: on any error the rank reports "not ok", the group
// agrees to run without multicast, and collectives keep using the peer-to-peer kernels.
#include <errno.h>
#include <fcntl.h>
#include <poll.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <sys/socket.h>
#include <sys/un.h>
#include <time.h>
#include <unistd.h>

#include <mutex>

#include "nbd_host.h"

namespace {

// ---- driver entry points (the library links no libcuda; the runtime resolves them) ----------
struct Drv {
  bool ok = false;
  CUresult (*DeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags) = nullptr;
  CUresult (*GetErrorString)(CUresult, const char**) = nullptr;
};
Drv g_drv;
std::once_flag g_drv_once;

template <typename F>
bool entry(const char* name, F* out) {
  void* f = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &f, cudaEnableDefault, &q) != cudaSuccess || !f) {
    cudaGetLastError();
    return false;
  }
  *out = reinterpret_cast<F>(f);
  return true;
}

const Drv& drv() {
  std::call_once(g_drv_once, [] {
    Drv& d = g_drv;
    bool ok = true;
    ok &= entry("cuDeviceGet", &d.DeviceGet);
    ok &= entry("cuDeviceGetAttribute", &d.DeviceGetAttribute);
    ok &= entry("cuMemCreate", &d.MemCreate);
    ok &= entry("cuMemRelease", &d.MemRelease);
    ok &= entry("cuMemAddressReserve", &d.MemAddressReserve);
    ok &= entry("cuMemAddressFree", &d.MemAddressFree);
    ok &= entry("cuMemMap", &d.MemMap);
    ok &= entry("cuMemUnmap", &d.MemUnmap);
    ok &= entry("cuMemSetAccess", &d.MemSetAccess);
    ok &= entry("cuMemExportToShareableHandle", &d.MemExportToShareableHandle);
    ok &= entry("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
    ok &= entry("cuMulticastCreate", &d.MulticastCreate);
    ok &= entry("cuMulticastAddDevice", &d.MulticastAddDevice);
    ok &= entry("cuMulticastBindMem", &d.MulticastBindMem);
    ok &= entry("cuMulticastUnbind", &d.MulticastUnbind);
    ok &= entry("cuMulticastGetGranularity", &d.MulticastGetGranularity);
    ok &= entry("cuGetErrorString", &d.GetErrorString);
    d.ok = ok;
  });
  return g_drv;
}

void why(nbd_comm* c, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(c->nvls.why, sizeof(c->nvls.why), fmt, ap);
  va_end(ap);
  const char* dbg = getenv("NBDIST_B200_DEBUG");
  if (dbg && dbg[0] != '0') fprintf(stderr, "[b200 nvls] rank %d: %s\n", c->rank, c->nvls.why);
}

const char* cu_str(CUresult r) {
  const char* s = nullptr;
  if (g_drv.GetErrorString) g_drv.GetErrorString(r, &s);
  return s ? s : "unknown CUDA driver error";
}

int64_t now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (int64_t)ts.tv_sec * 1000 + ts.tv_nsec / 1000000;
}

int64_t step_timeout_ms() {
  const char* e = getenv("NBDIST_B200_NVLS_TIMEOUT_MS");
  const int64_t v = e ? atoll(e) : 20000;
  return v > 0 ? v : 20000;
}

// ---- messages over the unix socket ------------------------------------------------------------
enum MsgType : uint32_t { M_HELLO = 1, M_MC = 2, M_ADDED = 3, M_BOUND = 4, M_VERDICT = 5 };
struct Msg {
  uint32_t type;
  int32_t rank;
  int32_t ok;
  uint32_t pad;
  uint64_t a, b;  // HELLO: staging bytes, granularity
};

socklen_t abstract_addr(sockaddr_un* sa, uint64_t nonce) {
  memset(sa, 0, sizeof(*sa));
  sa->sun_family = AF_UNIX;
  char name[64];
  const int n = snprintf(name, sizeof(name), "nbd-b200-mc-%016llx", (unsigned long long)nonce);
  memcpy(sa->sun_path + 1, name, (size_t)n);  // sun_path[0] == 0: abstract namespace, no file, gone with the fd
  return (socklen_t)(offsetof(sockaddr_un, sun_path) + 1 + n);
}

bool wait_io(int fd, short ev, int64_t deadline) {
  for (;;) {
    const int64_t left = deadline - now_ms();
    if (left <= 0) return false;
    pollfd p = {fd, ev, 0};
    const int r = poll(&p, 1, (int)(left > 1000 ? 1000 : left));
    if (r > 0) return (p.revents & (ev | POLLHUP | POLLERR)) != 0;
    if (r < 0 && errno != EINTR) return false;
  }
}

// Sends msg, with `pass_fd` attached as SCM_RIGHTS when >= 0.
bool send_msg(int fd, const Msg& m, int pass_fd, int64_t deadline) {
  if (!wait_io(fd, POLLOUT, deadline)) return false;
  iovec iov = {const_cast<Msg*>(&m), sizeof(m)};
  msghdr h;
  memset(&h, 0, sizeof(h));
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
  if (pass_fd >= 0) {
    memset(ctl, 0, sizeof(ctl));
    h.msg_control = ctl;
    h.msg_controllen = sizeof(ctl);
    cmsghdr* cm = CMSG_FIRSTHDR(&h);
    cm->cmsg_level = SOL_SOCKET;
    cm->cmsg_type = SCM_RIGHTS;
    cm->cmsg_len = CMSG_LEN(sizeof(int));
    memcpy(CMSG_DATA(cm), &pass_fd, sizeof(int));
  }
  for (;;) {
    const ssize_t n = sendmsg(fd, &h, MSG_NOSIGNAL);
    if (n == (ssize_t)sizeof(m)) return true;
    if (n < 0 && (errno == EINTR || errno == EAGAIN)) {
      if (!wait_io(fd, POLLOUT, deadline)) return false;
      continue;
    }
    return false;
  }
}

// Receives one message; *got_fd receives an attached descriptor (or -1).
bool recv_msg(int fd, Msg* m, int* got_fd, int64_t deadline) {
  if (got_fd) *got_fd = -1;
  if (!wait_io(fd, POLLIN, deadline)) return false;
  iovec iov = {m, sizeof(*m)};
  msghdr h;
  memset(&h, 0, sizeof(h));
  h.msg_iov = &iov;
  h.msg_iovlen = 1;
  alignas(cmsghdr) char ctl[CMSG_SPACE(sizeof(int))];
  memset(ctl, 0, sizeof(ctl));
  h.msg_control = ctl;
  h.msg_controllen = sizeof(ctl);
  for (;;) {
    const ssize_t n = recvmsg(fd, &h, MSG_WAITALL | MSG_CMSG_CLOEXEC);
    if (n == (ssize_t)sizeof(*m)) break;
    if (n < 0 && (errno == EINTR || errno == EAGAIN)) {
      if (!wait_io(fd, POLLIN, deadline)) return false;
      continue;
    }
    return false;
  }
  for (cmsghdr* cm = CMSG_FIRSTHDR(&h); cm; cm = CMSG_NXTHDR(&h, cm)) {
    if (cm->cmsg_level == SOL_SOCKET && cm->cmsg_type == SCM_RIGHTS && cm->cmsg_len >= CMSG_LEN(sizeof(int))) {
      int f = -1;
      memcpy(&f, CMSG_DATA(cm), sizeof(int));
      if (got_fd && *got_fd < 0) *got_fd = f;
      else if (f >= 0) close(f);
    }
  }
  return true;
}

uint64_t random_nonce() {
  uint64_t v = 0;
  const int fd = open("/dev/urandom", O_RDONLY | O_CLOEXEC);
  if (fd >= 0) {
    if (read(fd, &v, sizeof(v)) != (ssize_t)sizeof(v)) v = 0;
    close(fd);
  }
  if (v == 0) v = ((uint64_t)getpid() << 32) ^ (uint64_t)now_ms() ^ 0x9e3779b97f4a7c15ull;
  return v | 1ull;  // never 0
}

bool env_off() {
  const char* e = getenv("NBDIST_B200_NVLS");
  return e && e[0] == '0';
}

size_t env_bytes(const char* name, size_t dflt) {
  const char* e = getenv(name);
  if (!e) return dflt;
  return (size_t)strtoull(e, nullptr, 0);
}

size_t round_up(size_t v, size_t g) { return (v + g - 1) / g * g; }

// ---- the steps every rank performs locally ------------------------------------------------------
struct Local {
  CUdevice dev = 0;
  CUmulticastObjectProp prop;
};

// HELLO: capability + staging allocation.  Returns false with nvls.why set.
bool step_alloc(nbd_comm* c, Local* L) {
  const Drv& d = drv();
  if (env_off()) return why(c, "disabled by NBDIST_B200_NVLS=0"), false;
  if (!d.ok) return why(c, "the CUDA driver lacks the multicast entry points"), false;
  CUresult r = d.DeviceGet(&L->dev, c->device);
  if (r != CUDA_SUCCESS) return why(c, "cuDeviceGet: %s", cu_str(r)), false;
  int mc_ok = 0;
  r = d.DeviceGetAttribute(&mc_ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, L->dev);
  if (r != CUDA_SUCCESS || !mc_ok) return why(c, "device %d does not support multicast (no NVSwitch / fabric manager?)", c->device), false;
  memset(&L->prop, 0, sizeof(L->prop));
  L->prop.numDevices = (unsigned)c->world;
  L->prop.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  L->prop.size = 128ull << 20;
  size_t gran = 0;
  r = d.MulticastGetGranularity(&gran, &L->prop, CU_MULTICAST_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS || gran == 0) return why(c, "cuMulticastGetGranularity: %s", cu_str(r)), false;
  NvlsState& s = c->nvls;
  s.granularity = gran;
  // staging for the tile-pipeline kernels, then a pool that nbd_mem_alloc hands out as symmetric
  // memory (zero-copy in-switch all_reduce); both in units of the recommended granularity
  size_t stage = env_bytes("NBDIST_B200_NVLS_BYTES", 128ull << 20);
  if (stage == 0) stage = 128ull << 20;
  s.stage_bytes = round_up(stage, gran);
  const size_t pool = env_bytes("NBDIST_B200_NVLS_POOL_BYTES", 1ull << 30);
  s.pool_bytes = pool ? round_up(pool, gran) : 0;
  s.bytes = s.stage_bytes + s.pool_bytes;
  s.pool.clear();
  if (s.pool_bytes) s.pool.push_back({s.stage_bytes, s.pool_bytes, false});
  L->prop.size = s.bytes;
  CUmemAllocationProp ap;
  memset(&ap, 0, sizeof(ap));
  ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  ap.location.id = c->device;
  // "externally shareable as well as imported multicast objects can be bound only to externally
  // shareable memory" (cuMulticastBindMem): the staging is never exported, but must be exportable
  ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle mem = 0;
  r = d.MemCreate(&mem, s.bytes, &ap, 0);
  if (r != CUDA_SUCCESS) return why(c, "cuMemCreate(%zu bytes): %s", s.bytes, cu_str(r)), false;
  s.mem = mem;
  s.have_mem = true;
  CUdeviceptr va = 0;
  r = d.MemAddressReserve(&va, s.bytes, gran, 0, 0);
  if (r != CUDA_SUCCESS) return why(c, "cuMemAddressReserve: %s", cu_str(r)), false;
  s.local = reinterpret_cast<char*>(va);
  r = d.MemMap(va, s.bytes, 0, mem, 0);
  if (r != CUDA_SUCCESS) {
    d.MemAddressFree(va, s.bytes);
    s.local = nullptr;
    return why(c, "cuMemMap(local): %s", cu_str(r)), false;
  }
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(va, s.bytes, &acc, 1);
  if (r != CUDA_SUCCESS) return why(c, "cuMemSetAccess(local): %s", cu_str(r)), false;
  return true;
}

bool step_add(nbd_comm* c, Local* L) {
  const CUresult r = drv().MulticastAddDevice(c->nvls.mc, L->dev);
  if (r != CUDA_SUCCESS) return why(c, "cuMulticastAddDevice: %s", cu_str(r)), false;
  c->nvls.added = true;
  return true;
}

bool step_bind(nbd_comm* c, Local* L) {
  const Drv& d = drv();
  NvlsState& s = c->nvls;
  CUresult r = d.MulticastBindMem(s.mc, 0, s.mem, 0, s.bytes, 0);
  if (r != CUDA_SUCCESS) return why(c, "cuMulticastBindMem: %s", cu_str(r)), false;
  s.bound = true;
  CUdeviceptr va = 0;
  r = d.MemAddressReserve(&va, s.bytes, s.granularity, 0, 0);
  if (r != CUDA_SUCCESS) return why(c, "cuMemAddressReserve(multicast): %s", cu_str(r)), false;
  r = d.MemMap(va, s.bytes, 0, s.mc, 0);
  if (r != CUDA_SUCCESS) {
    d.MemAddressFree(va, s.bytes);
    return why(c, "cuMemMap(multicast): %s", cu_str(r)), false;
  }
  s.mcva = reinterpret_cast<char*>(va);
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = c->device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.MemSetAccess(va, s.bytes, &acc, 1);
  if (r != CUDA_SUCCESS) return why(c, "cuMemSetAccess(multicast): %s", cu_str(r)), false;
  (void)L;
  return true;
}

// ---- rank 0 -----------------------------------------------------------------------------------
// Collects one message of `type` from every peer; returns the AND of their `ok`.
bool collect(int* fds, int world, MsgType type, int64_t deadline) {
  bool all = true;
  for (int p = 1; p < world; ++p) {
    if (fds[p] < 0) {
      all = false;
      continue;
    }
    Msg m;
    if (!recv_msg(fds[p], &m, nullptr, deadline) || m.type != (uint32_t)type || !m.ok) all = false;
  }
  return all;
}

bool verdict(int* fds, int world, MsgType about, bool ok, int64_t deadline) {
  Msg v = {M_VERDICT, 0, ok ? 1 : 0, 0, (uint64_t)about, 0};
  bool sent = true;
  for (int p = 1; p < world; ++p)
    if (fds[p] >= 0 && !send_msg(fds[p], v, -1, deadline)) sent = false;
  return ok && sent;
}

void server(nbd_comm* c) {
  NvlsState& s = c->nvls;
  const Drv& d = drv();
  const int64_t T = step_timeout_ms();
  int fds[kMaxRanks];
  for (int p = 0; p < kMaxRanks; ++p) fds[p] = -1;
  Local L;
  bool ok = step_alloc(c, &L);
  // HELLO from every peer (each on its own connection)
  int64_t deadline = now_ms() + T;
  int connected = 0;
  while (connected < c->world - 1) {
    if (!wait_io(s.listen_fd, POLLIN, deadline)) break;
    const int fd = accept4(s.listen_fd, nullptr, nullptr, SOCK_CLOEXEC);
    if (fd < 0) {
      if (errno == EINTR || errno == EAGAIN) continue;
      break;
    }
    Msg m;
    if (!recv_msg(fd, &m, nullptr, deadline) || m.type != M_HELLO || m.rank <= 0 || m.rank >= c->world || fds[m.rank] >= 0) {
      close(fd);
      continue;
    }
    fds[m.rank] = fd;
    ++connected;
    if (!m.ok) {
      if (ok) why(c, "rank %d cannot use multicast", m.rank);
      ok = false;
    } else if (ok && (m.a != (uint64_t)s.bytes || m.b != (uint64_t)s.granularity)) {
      why(c, "rank %d sized its staging %llu/%llu, rank 0 %zu/%zu", m.rank, (unsigned long long)m.a, (unsigned long long)m.b,
          s.bytes, s.granularity);
      ok = false;
    }
  }
  close(s.listen_fd);
  s.listen_fd = -1;
  if (connected < c->world - 1) {
    if (ok) why(c, "only %d of %d peers reached the multicast rendezvous", connected, c->world - 1);
    ok = false;
  }
  // MC: create the object and hand every peer a descriptor of it
  int mcfd = -1;
  if (ok) {
    CUmemGenericAllocationHandle mc = 0;
    CUresult r = d.MulticastCreate(&mc, &L.prop);
    if (r != CUDA_SUCCESS) {
      why(c, "cuMulticastCreate(%d devices, %zu bytes): %s", c->world, s.bytes, cu_str(r));
      ok = false;
    } else {
      s.mc = mc;
      s.have_mc = true;
      r = d.MemExportToShareableHandle(&mcfd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
      if (r != CUDA_SUCCESS) {
        why(c, "cuMemExportToShareableHandle(multicast): %s", cu_str(r));
        ok = false;
        mcfd = -1;
      }
    }
  }
  deadline = now_ms() + T;
  {
    Msg m = {M_MC, 0, ok ? 1 : 0, 0, 0, 0};
    for (int p = 1; p < c->world; ++p)
      if (fds[p] >= 0 && !send_msg(fds[p], m, ok ? mcfd : -1, deadline)) ok = false;
  }
  if (mcfd >= 0) close(mcfd);
  if (ok) {
    // ADDED
    bool mine = step_add(c, &L);
    bool theirs = collect(fds, c->world, M_ADDED, now_ms() + T);
    if (mine && !theirs) why(c, "a peer failed cuMulticastAddDevice");
    ok = verdict(fds, c->world, M_ADDED, mine && theirs, now_ms() + T);
  }
  if (ok) {
    // BOUND
    bool mine = step_bind(c, &L);
    bool theirs = collect(fds, c->world, M_BOUND, now_ms() + T);
    if (mine && !theirs) why(c, "a peer failed cuMulticastBindMem");
    ok = verdict(fds, c->world, M_BOUND, mine && theirs, now_ms() + T);
  }
  for (int p = 1; p < c->world; ++p)
    if (fds[p] >= 0) close(fds[p]);
  s.enabled = ok;
}

// ---- ranks 1..N-1 -----------------------------------------------------------------------------
void client(nbd_comm* c, uint64_t nonce) {
  NvlsState& s = c->nvls;
  const Drv& d = drv();
  const int64_t T = step_timeout_ms();
  Local L;
  bool ok = step_alloc(c, &L);
  const int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) return why(c, "socket(AF_UNIX): %s", strerror(errno));
  sockaddr_un sa;
  const socklen_t len = abstract_addr(&sa, nonce);
  int64_t deadline = now_ms() + T;
  for (;;) {  // rank 0 listens from nbd_comm_create on; retry only covers a full backlog
    if (connect(fd, (sockaddr*)&sa, len) == 0) break;
    if ((errno == EAGAIN || errno == ECONNREFUSED || errno == EINTR) && now_ms() < deadline) {
      usleep(2000);
      continue;
    }
    why(c, "connect to rank 0's multicast rendezvous: %s", strerror(errno));
    close(fd);
    return;
  }
  Msg hello = {M_HELLO, c->rank, ok ? 1 : 0, 0, (uint64_t)s.bytes, (uint64_t)s.granularity};
  Msg m;
  int mcfd = -1;
  bool alive = send_msg(fd, hello, -1, deadline);
  // the MC message arrives after rank 0 has heard from everybody: allow two step timeouts
  alive = alive && recv_msg(fd, &m, &mcfd, now_ms() + 2 * T) && m.type == M_MC;
  if (!alive) {
    if (ok) why(c, "lost the multicast rendezvous with rank 0");
    ok = false;
  } else if (!m.ok) {
    if (ok) why(c, "the group agreed to run without multicast (see rank 0 or NBDIST_B200_DEBUG=1)");
    ok = false;
  }
  if (ok) {
    if (mcfd < 0) {
      why(c, "rank 0 sent no multicast descriptor");
      ok = false;
    } else {
      CUmemGenericAllocationHandle mc = 0;
      const CUresult r = d.MemImportFromShareableHandle(&mc, (void*)(uintptr_t)mcfd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
      if (r != CUDA_SUCCESS) {
        why(c, "cuMemImportFromShareableHandle(multicast): %s", cu_str(r));
        ok = false;
      } else {
        s.mc = mc;
        s.have_mc = true;
      }
    }
  }
  if (mcfd >= 0) close(mcfd);
  if (alive && m.ok) {
    // we take part in the ADDED / BOUND rounds whenever rank 0 started them (m.ok), reporting
    // our own failure if we had one, so that nobody waits for a rank that has given up
    const bool added = ok && step_add(c, &L);
    Msg a = {M_ADDED, c->rank, added ? 1 : 0, 0, 0, 0};
    Msg v;
    bool go = send_msg(fd, a, -1, now_ms() + T) && recv_msg(fd, &v, nullptr, now_ms() + 2 * T) && v.type == M_VERDICT && v.ok;
    if (!go && ok) why(c, "the group gave up after cuMulticastAddDevice");
    ok = ok && go;
    if (go) {
      const bool bound = ok && step_bind(c, &L);
      Msg b = {M_BOUND, c->rank, bound ? 1 : 0, 0, 0, 0};
      go = send_msg(fd, b, -1, now_ms() + T) && recv_msg(fd, &v, nullptr, now_ms() + 2 * T) && v.type == M_VERDICT && v.ok;
      if (!go && ok) why(c, "the group gave up after cuMulticastBindMem");
      ok = ok && go;
    }
  }
  close(fd);
  s.enabled = ok;
}

}  // namespace

// Rank 0 opens the rendezvous socket when its communicator is created, i.e. before any peer can
// hold its handle, so a peer's connect() never races the listener.
int nvls_prepare(nbd_comm* c) {
  NvlsState& s = c->nvls;
  if (c->emu || c->world < 2 || c->rank != 0) return NBD_OK;
  if (env_off()) {
    why(c, "disabled by NBDIST_B200_NVLS=0");
    return NBD_OK;
  }
  const int fd = socket(AF_UNIX, SOCK_STREAM | SOCK_CLOEXEC, 0);
  if (fd < 0) {
    why(c, "socket(AF_UNIX): %s", strerror(errno));
    return NBD_OK;
  }
  for (int attempt = 0; attempt < 8; ++attempt) {
    const uint64_t nonce = random_nonce();
    sockaddr_un sa;
    const socklen_t len = abstract_addr(&sa, nonce);
    if (bind(fd, (sockaddr*)&sa, len) == 0 && listen(fd, kMaxRanks) == 0) {
      s.listen_fd = fd;
      s.nonce = nonce;
      return NBD_OK;
    }
    if (errno != EADDRINUSE) break;
  }
  why(c, "bind/listen on the multicast rendezvous socket: %s", strerror(errno));
  close(fd);
  return NBD_OK;
}

void nvls_setup(nbd_comm* c, uint64_t nonce_rank0) {
  NvlsState& s = c->nvls;
  s.enabled = false;
  if (c->emu || c->world < 2) return;
  if (nonce_rank0 == 0) {
    if (c->rank != 0) why(c, "rank 0 does not want multicast (NBDIST_B200_NVLS=0 there, or its socket failed)");
    return;
  }
  DeviceGuard g(c->device);
  if (c->rank == 0) server(c);
  else client(c, nonce_rank0);
  if (s.enabled) snprintf(s.why, sizeof(s.why), "enabled: %zu MiB staging + %zu MiB symmetric pool per rank, granularity %zu KiB",
                          s.stage_bytes >> 20, s.pool_bytes >> 20, s.granularity >> 10);
  else nvls_teardown(c);
  const char* dbg = getenv("NBDIST_B200_DEBUG");
  if (dbg && dbg[0] != '0') fprintf(stderr, "[b200 nvls] rank %d: %s\n", c->rank, s.why);
}

void nvls_teardown(nbd_comm* c) {
  NvlsState& s = c->nvls;
  const Drv& d = g_drv;
  if (s.listen_fd >= 0) {
    close(s.listen_fd);
    s.listen_fd = -1;
  }
  if (!d.ok) return;
  s.enabled = false;
  if (s.mcva) {
    d.MemUnmap((CUdeviceptr)s.mcva, s.bytes);
    d.MemAddressFree((CUdeviceptr)s.mcva, s.bytes);
    s.mcva = nullptr;
  }
  if (s.bound) {
    CUdevice dev = 0;
    if (d.DeviceGet(&dev, c->device) == CUDA_SUCCESS) d.MulticastUnbind(s.mc, dev, 0, s.bytes);
    s.bound = false;
  }
  if (s.have_mc) {
    d.MemRelease(s.mc);
    s.have_mc = false;
  }
  if (s.local) {
    d.MemUnmap((CUdeviceptr)s.local, s.bytes);
    d.MemAddressFree((CUdeviceptr)s.local, s.bytes);
    s.local = nullptr;
  }
  if (s.have_mem) {
    d.MemRelease(s.mem);
    s.have_mem = false;
  }
}

// Symmetric pool: first fit over address-ordered blocks, 2 MiB granules.  Every rank makes the
// same calls in the same order (nbd_mem_alloc is collective), so every rank gets the same
// offset; the descriptor check of the zero-copy kernels catches a rank that did not.
char* nvls_pool_alloc(nbd_comm* c, size_t bytes, char** mc_alias) {
  NvlsState& s = c->nvls;
  if (!s.enabled || s.pool_bytes == 0 || bytes == 0) return nullptr;
  const size_t need = round_up(bytes, 2ull << 20);
  for (size_t k = 0; k < s.pool.size(); ++k) {
    NvlsState::PoolBlock& b = s.pool[k];
    if (b.used || b.len < need) continue;
    const size_t off = b.off;
    if (b.len > need) {
      const NvlsState::PoolBlock rest = {b.off + need, b.len - need, false};
      b.len = need;
      b.used = true;
      s.pool.insert(s.pool.begin() + (long)k + 1, rest);
    } else {
      b.used = true;
    }
    *mc_alias = s.mcva + off;
    return s.local + off;
  }
  return nullptr;
}

bool nvls_pool_free(nbd_comm* c, char* ptr) {
  NvlsState& s = c->nvls;
  if (!s.local || ptr < s.local + s.stage_bytes || ptr >= s.local + s.bytes) return false;
  const size_t off = (size_t)(ptr - s.local);
  for (size_t k = 0; k < s.pool.size(); ++k) {
    if (s.pool[k].off != off || !s.pool[k].used) continue;
    s.pool[k].used = false;
    if (k + 1 < s.pool.size() && !s.pool[k + 1].used) {
      s.pool[k].len += s.pool[k + 1].len;
      s.pool.erase(s.pool.begin() + (long)k + 1);
    }
    if (k > 0 && !s.pool[k - 1].used) {
      s.pool[k - 1].len += s.pool[k].len;
      s.pool.erase(s.pool.begin() + (long)k);
    }
    return true;
  }
  return false;
}
