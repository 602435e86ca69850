// c10d_backend.cpp — the c10d::Backend that torch.distributed loads for
// backend string "b200".  It is glue only: tensor checks, stream/event
// bookkeeping, IPC-handle exchange over the store torch hands us, and calls
// into the C ABI (include/nbdist_b200.h) where the sm_100a kernels live.
//
// It replaces, for CUDA tensors, the ProcessGroupNCCL instance that
// ref:src/nbdistributed/worker.py:145-151 selects with backend="nccl":
// the same user calls (dist.all_reduce / broadcast / all_gather /
// reduce_scatter / barrier from cell code, ref:src/nbdistributed/worker.py:162,
// :214) arrive here through c10d::Backend's virtuals
// ($TORCH/include/torch/csrc/distributed/c10d/Backend.hpp:140-382).
#include <ATen/cuda/CUDAContext.h>
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDACachingAllocator.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <pybind11/chrono.h>
#include <torch/csrc/distributed/c10d/Backend.hpp>
#include <torch/csrc/distributed/c10d/Store.hpp>
#include <torch/csrc/distributed/c10d/Types.hpp>
#include <torch/csrc/distributed/c10d/Work.hpp>
#include <torch/extension.h>
#include <ATen/ops/from_blob.h>

#include <execinfo.h>
#include <signal.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

#include "nbdist_b200.h"

namespace {

constexpr const char* kBackendName = "b200";

bool debug_on() {
  static const bool on = [] { const char* e = std::getenv("NBDIST_B200_DEBUG"); return e && e[0] != '0'; }();
  return on;
}
#define NBD_DBG(...)                                 \
  do {                                               \
    if (debug_on()) {                                \
      std::fprintf(stderr, "[b200 glue] " __VA_ARGS__); \
      std::fprintf(stderr, "\n");                    \
      std::fflush(stderr);                           \
    }                                                \
  } while (0)

void segv_backtrace(int sig) {
  void* frames[64];
  const int n = backtrace(frames, 64);
  const char msg[] = "[b200 glue] fatal signal, native backtrace:\n";
  (void)!write(2, msg, sizeof(msg) - 1);
  backtrace_symbols_fd(frames, n, 2);
  signal(sig, SIG_DFL);
  raise(sig);
}

[[noreturn]] void throw_nbd(int rc, const char* what) {
  std::string msg = c10::str("[", kBackendName, "] ", what, ": ", nbd_error_string(rc));
  const char* detail = nbd_last_error();
  if (detail && *detail) msg += c10::str(" — ", detail);
  C10_THROW_ERROR(DistBackendError, msg);
}

#define NBD_CHECK(call, what)       \
  do {                              \
    int rc__ = (call);              \
    if (rc__ != NBD_OK) throw_nbd(rc__, what); \
  } while (0)

int to_nbd_dtype(at::ScalarType t, bool for_reduce) {
  switch (t) {
    case at::kFloat: return NBD_F32;
    case at::kBFloat16: return NBD_BF16;
    case at::kHalf: return NBD_F16;
    case at::kDouble: return NBD_F64;
    case at::kInt: return NBD_I32;
    case at::kLong: return NBD_I64;
    case at::kByte: return NBD_U8;
    case at::kChar: return NBD_I8;
    default:
      TORCH_CHECK(!for_reduce, "[b200] dtype ", t, " cannot be reduced (supported: float32, bfloat16, float16, "
                  "float64, int32, int64, uint8, int8)");
      return NBD_BYTES;
  }
}

int to_nbd_op(const c10d::ReduceOp& op) {
  switch (op.op_) {
    case c10d::ReduceOp::SUM: return NBD_SUM;
    case c10d::ReduceOp::AVG: return NBD_AVG;
    case c10d::ReduceOp::PRODUCT: return NBD_PROD;
    case c10d::ReduceOp::MIN: return NBD_MIN;
    case c10d::ReduceOp::MAX: return NBD_MAX;
    case c10d::ReduceOp::BAND: return NBD_BAND;
    case c10d::ReduceOp::BOR: return NBD_BOR;
    case c10d::ReduceOp::BXOR: return NBD_BXOR;
    default: TORCH_CHECK(false, "[b200] unsupported ReduceOp (PREMUL_SUM is NCCL-specific)");
  }
}

void check_tensor(const at::Tensor& t) {
  TORCH_CHECK(t.is_cuda(), "[b200] tensors must be CUDA tensors (this backend is registered for devices=['cuda'])");
  TORCH_CHECK(!t.is_sparse(), "[b200] sparse tensors are not supported");
  TORCH_CHECK_VALUE(t.is_contiguous(t.suggest_memory_format()), "Tensors must be contiguous");
}

// ---------------------------------------------------------------------------
class WorkB200 : public c10d::Work {
 public:
  WorkB200(int rank, c10d::OpType type, const char* title, c10::DeviceIndex dev, nbd_comm_t comm,
           std::vector<at::Tensor> outputs, bool is_barrier)
      : c10d::Work(rank, type, title), dev_(dev), comm_(comm), outputs_(std::move(outputs)), is_barrier_(is_barrier) {}

  // record completion on `stream` and publish a CUDA-aware future bound to it
  void mark_enqueued(c10::cuda::CUDAStream stream) {
    done_.record(stream);
    future_ = c10::make_intrusive<at::ivalue::Future>(c10::ListType::create(c10::TensorType::get()),
                                                      std::vector<c10::Device>{c10::Device(c10::kCUDA, dev_)});
    c10::cuda::CUDAStreamGuard g(stream);
    future_->markCompleted(at::IValue(outputs_));
  }

  bool isCompleted() override { return done_.query(); }
  bool isSuccess() const override { return !comm_ || nbd_comm_async_error(comm_) == NBD_OK; }

  void synchronize() override {
    auto cur = c10::cuda::getCurrentCUDAStream(dev_);
    done_.block(cur);
    if (is_barrier_) {
      // dist.barrier() is a host-side rendezvous (as ProcessGroupNCCL does for barrier works)
      c10::cuda::CUDAGuard g(dev_);
      C10_CUDA_CHECK(cudaStreamSynchronize(cur));
      int rc = nbd_comm_async_error(comm_);
      if (rc != NBD_OK) throw_nbd(rc, "barrier");
    }
  }

  // wait() without a timeout is a stream wait (what cell code gets after a synchronous call).
  // With a timeout it blocks the HOST until the work's event fires or the deadline passes; on
  // expiry the communicator is aborted (in-flight kernels leave their flag waits) and the call
  // throws, instead of the reference's hang-forever (ref:src/nbdistributed/communication.py:263-269).
  bool wait(std::chrono::milliseconds timeout = kNoTimeout) override {
    if (timeout != kNoTimeout && timeout.count() > 0) {
      const auto deadline = std::chrono::steady_clock::now() + timeout;
      while (!done_.query()) {
        if (std::chrono::steady_clock::now() >= deadline) {
          if (comm_) nbd_comm_abort(comm_);
          C10_THROW_ERROR(DistBackendError, c10::str("[", kBackendName, "] wait(timeout=", timeout.count(),
                                                     " ms) expired; the communicator was aborted"));
        }
        std::this_thread::sleep_for(std::chrono::microseconds(50));
      }
      const int rc = comm_ ? nbd_comm_async_error(comm_) : (int)NBD_OK;
      if (rc != NBD_OK) throw_nbd(rc, "wait");
    }
    synchronize();
    return true;
  }

  std::vector<at::Tensor> result() override { return outputs_; }
  c10::intrusive_ptr<c10::ivalue::Future> getFuture() override { return future_; }

 private:
  c10::DeviceIndex dev_;
  nbd_comm_t comm_;
  std::vector<at::Tensor> outputs_;
  bool is_barrier_;
  at::cuda::CUDAEvent done_;
  c10::intrusive_ptr<at::ivalue::Future> future_;
};

// ---------------------------------------------------------------------------
class BackendB200 : public c10d::Backend {
 public:
  BackendB200(c10::intrusive_ptr<c10d::Store> store, int rank, int size, std::chrono::milliseconds timeout)
      : c10d::Backend(rank, size), store_(std::move(store)), timeout_(timeout) {
    TORCH_CHECK(size >= 1 && size <= NBD_MAX_RANKS, "[b200] world_size ", size, " unsupported: this backend drives one "
                "NVLink/NVSwitch node of at most ", NBD_MAX_RANKS, " GPUs");
    TORCH_CHECK(at::cuda::is_available(), "[b200] no CUDA device: the b200 backend has no CPU fallback "
                "(use 'gloo' for CPU tensors, e.g. backend='cpu:gloo,cuda:b200')");
    if (debug_on()) signal(SIGSEGV, segv_backtrace);
    const char* inl = std::getenv("NBDIST_B200_INLINE_SYNC_OPS");
    inline_sync_ops_ = !(inl && inl[0] == '0');
  }

  ~BackendB200() override {
    try {
      shutdown();
    } catch (...) {
    }
  }

  const std::string getBackendName() const override { return kBackendName; }

  void setTimeout(std::chrono::milliseconds timeout) override {
    timeout_ = timeout;
    if (comm_) nbd_comm_set_timeout_ms(comm_, (uint64_t)timeout.count());
  }

  void abort() override {
    std::lock_guard<std::mutex> lk(mu_);
    if (comm_) nbd_comm_abort(comm_);
  }

  // Teardown never blocks on peers for long: a bounded store handshake (so that
  // nobody unmaps an arena a slow peer is still pulling from), then free.
  void shutdown() override {
    std::lock_guard<std::mutex> lk(mu_);
    if (!comm_) return;
    {
      c10::cuda::CUDAGuard g(device_);
      if (comm_stream_) cudaStreamSynchronize(comm_stream_->stream());
      cudaStreamSynchronize(c10::cuda::getCurrentCUDAStream(device_));
    }
    if (getSize() > 1) {
      try {
        const std::string key = "nbd_b200/shutdown";
        store_->add(key, 1);
        const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(shutdown_wait_ms());
        while (store_->add(key, 0) < getSize() && std::chrono::steady_clock::now() < deadline)
          std::this_thread::sleep_for(std::chrono::milliseconds(5));
      } catch (...) {
        // store already gone (rank 0 exited or peers were killed): nothing to wait for
      }
    }
    alive_->store(false);
    nbd_comm_destroy(comm_);
    comm_ = nullptr;
    comm_stream_.reset();
  }

  c10d::ErrorType getError() override {
    if (!comm_) return c10d::ErrorType::SUCCESS;
    const int rc = nbd_comm_async_error(comm_);
    if (rc == NBD_OK) return c10d::ErrorType::SUCCESS;
    return rc == NBD_ERR_TIMEOUT ? c10d::ErrorType::TIMEOUT : c10d::ErrorType::COMM_ERROR;
  }

  // ---- collectives ---------------------------------------------------------
  c10::intrusive_ptr<c10d::Work> allreduce(std::vector<at::Tensor>& tensors,
                                           const c10d::AllreduceOptions& opts) override {
    TORCH_CHECK(tensors.size() == 1, "[b200] allreduce expects exactly one tensor");
    auto& t = tensors[0];
    check_tensor(t);
    const int dtype = to_nbd_dtype(t.scalar_type(), true), op = to_nbd_op(opts.reduceOp);
    return run(c10d::OpType::ALLREDUCE, "b200:all_reduce", t.get_device(), {t}, {t}, opts.asyncOp, false,
               [&](cudaStream_t s) {
                 return nbd_all_reduce(comm_, t.data_ptr(), t.data_ptr(), (size_t)t.numel(), dtype, op, s);
               });
  }

  c10::intrusive_ptr<c10d::Work> allreduce_coalesced(std::vector<at::Tensor>& tensors,
                                                     const c10d::AllreduceCoalescedOptions& opts) override {
    TORCH_CHECK(!tensors.empty(), "[b200] allreduce_coalesced needs at least one tensor");
    for (auto& t : tensors) check_tensor(t);
    const int op = to_nbd_op(opts.reduceOp);
    return run(c10d::OpType::ALLREDUCE_COALESCED, "b200:all_reduce_coalesced", tensors[0].get_device(), tensors,
               tensors, opts.asyncOp, false, [&](cudaStream_t s) {
                 for (auto& t : tensors) {
                   int rc = nbd_all_reduce(comm_, t.data_ptr(), t.data_ptr(), (size_t)t.numel(),
                                           to_nbd_dtype(t.scalar_type(), true), op, s);
                   if (rc != NBD_OK) return rc;
                 }
                 return (int)NBD_OK;
               });
  }

  c10::intrusive_ptr<c10d::Work> broadcast(std::vector<at::Tensor>& tensors,
                                           const c10d::BroadcastOptions& opts) override {
    TORCH_CHECK(tensors.size() == 1, "[b200] broadcast expects exactly one tensor");
    auto& t = tensors[0];
    check_tensor(t);
    TORCH_CHECK(opts.rootRank >= 0 && opts.rootRank < getSize(), "[b200] invalid root rank ", opts.rootRank);
    const int root = (int)opts.rootRank;
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (coalescing_) {  // collected: endCoalescing runs the block's broadcasts as tables (nbd_broadcast_table)
        co_bcast_.push_back({t, root});
        co_tensors_.push_back(t);
        return collected_work(t);
      }
    }
    return run(c10d::OpType::BROADCAST, "b200:broadcast", t.get_device(), {t}, {t}, opts.asyncOp, false,
               [&](cudaStream_t s) {
                 return nbd_broadcast(comm_, t.data_ptr(), (size_t)t.nbytes(), NBD_BYTES, root, s);
               });
  }

  c10::intrusive_ptr<c10d::Work> _allgather_base(at::Tensor& out, at::Tensor& in,
                                                 const c10d::AllgatherOptions& opts) override {
    check_tensor(out);
    check_tensor(in);
    TORCH_CHECK(out.scalar_type() == in.scalar_type(), "[b200] all_gather: output and input dtypes differ");
    TORCH_CHECK(out.numel() == in.numel() * getSize(), "[b200] all_gather: output must hold world_size * input elements");
    return run(c10d::OpType::_ALLGATHER_BASE, "b200:all_gather_into_tensor", in.get_device(), {in}, {out},
               opts.asyncOp, false, [&](cudaStream_t s) {
                 return nbd_all_gather(comm_, in.data_ptr(), out.data_ptr(), (size_t)in.nbytes(), NBD_BYTES, s);
               });
  }

  c10::intrusive_ptr<c10d::Work> allgather(std::vector<std::vector<at::Tensor>>& outs, std::vector<at::Tensor>& ins,
                                           const c10d::AllgatherOptions& opts) override {
    TORCH_CHECK(ins.size() == 1 && outs.size() == 1, "[b200] allgather expects one input tensor and one output list");
    auto& in = ins[0];
    auto& list = outs[0];
    check_tensor(in);
    TORCH_CHECK((int)list.size() == getSize(), "[b200] all_gather: output list must have world_size tensors");
    for (auto& o : list) {
      check_tensor(o);
      TORCH_CHECK(o.scalar_type() == in.scalar_type() && o.numel() == in.numel(),
                  "[b200] all_gather: every output tensor must match the input's dtype and size");
    }
    std::vector<at::Tensor> all_out(list.begin(), list.end());
    all_out.push_back(in);
    const int N = getSize(), me = getRank();
    return run(c10d::OpType::ALLGATHER, "b200:all_gather", in.get_device(), {in}, all_out, opts.asyncOp, false,
               [&](cudaStream_t s) {
                 // straight into the caller's list: one point-to-point batch (a send to and a
                 // receive from every peer, all concurrent) plus the local copy - no flat
                 // temporary, no scatter copies
                 std::vector<nbd_p2p_op> ops;
                 for (int k = 1; k < N; ++k) ops.push_back({NBD_P2P_SEND, (me + k) % N, in.data_ptr(), (size_t)in.nbytes()});
                 for (int k = 1; k < N; ++k) {
                   const int p = (me - k + N) % N;
                   ops.push_back({NBD_P2P_RECV, p, list[p].data_ptr(), (size_t)list[p].nbytes()});
                 }
                 if (in.nbytes() > 0 && list[me].data_ptr() != in.data_ptr())
                   if (cudaMemcpyAsync(list[me].data_ptr(), in.data_ptr(), (size_t)in.nbytes(), cudaMemcpyDeviceToDevice, s) != cudaSuccess)
                     return (int)NBD_ERR_CUDA;
                 return nbd_p2p_batch(comm_, ops.data(), (int)ops.size(), s);
               });
  }

  c10::intrusive_ptr<c10d::Work> allgather_into_tensor_coalesced(std::vector<at::Tensor>& outs,
                                                                 std::vector<at::Tensor>& ins,
                                                                 const c10d::AllgatherOptions& opts) override {
    TORCH_CHECK(outs.size() == ins.size() && !ins.empty(), "[b200] all_gather coalesced: list sizes differ");
    for (size_t i = 0; i < ins.size(); ++i) {
      check_tensor(ins[i]);
      check_tensor(outs[i]);
      TORCH_CHECK(outs[i].numel() == ins[i].numel() * getSize() && outs[i].scalar_type() == ins[i].scalar_type(),
                  "[b200] all_gather coalesced: bad output size/dtype at index ", i);
    }
    return run(c10d::OpType::COALESCED, "b200:all_gather_into_tensor_coalesced", ins[0].get_device(), ins, outs,
               opts.asyncOp, false, [&](cudaStream_t s) {
                 for (size_t i = 0; i < ins.size(); ++i) {
                   int rc = nbd_all_gather(comm_, ins[i].data_ptr(), outs[i].data_ptr(), (size_t)ins[i].nbytes(),
                                           NBD_BYTES, s);
                   if (rc != NBD_OK) return rc;
                 }
                 return (int)NBD_OK;
               });
  }

  c10::intrusive_ptr<c10d::Work> _reduce_scatter_base(at::Tensor& out, at::Tensor& in,
                                                      const c10d::ReduceScatterOptions& opts) override {
    check_tensor(out);
    check_tensor(in);
    TORCH_CHECK(out.scalar_type() == in.scalar_type(), "[b200] reduce_scatter: output and input dtypes differ");
    TORCH_CHECK(in.numel() == out.numel() * getSize(), "[b200] reduce_scatter: input must hold world_size * output elements");
    const int dtype = to_nbd_dtype(in.scalar_type(), true), op = to_nbd_op(opts.reduceOp);
    return run(c10d::OpType::_REDUCE_SCATTER_BASE, "b200:reduce_scatter_tensor", in.get_device(), {in}, {out},
               opts.asyncOp, false, [&](cudaStream_t s) {
                 return nbd_reduce_scatter(comm_, in.data_ptr(), out.data_ptr(), (size_t)out.numel(), dtype, op, s);
               });
  }

  c10::intrusive_ptr<c10d::Work> reduce_scatter(std::vector<at::Tensor>& outs,
                                                std::vector<std::vector<at::Tensor>>& ins,
                                                const c10d::ReduceScatterOptions& opts) override {
    TORCH_CHECK(outs.size() == 1 && ins.size() == 1, "[b200] reduce_scatter expects one output tensor and one input list");
    auto& out = outs[0];
    auto& list = ins[0];
    check_tensor(out);
    TORCH_CHECK((int)list.size() == getSize(), "[b200] reduce_scatter: input list must have world_size tensors");
    for (auto& t : list) {
      check_tensor(t);
      TORCH_CHECK(t.scalar_type() == out.scalar_type() && t.numel() == out.numel(),
                  "[b200] reduce_scatter: every input tensor must match the output's dtype and size");
    }
    const int dtype = to_nbd_dtype(out.scalar_type(), true), op = to_nbd_op(opts.reduceOp);
    std::vector<at::Tensor> all_in(list.begin(), list.end());
    return run(c10d::OpType::REDUCE_SCATTER, "b200:reduce_scatter", out.get_device(), all_in, {out}, opts.asyncOp,
               false, [&](cudaStream_t s) {
                 // the kernel takes the N contributions where they lie: no concatenation
                 std::vector<const void*> parts;
                 parts.reserve(list.size());
                 for (auto& t : list) parts.push_back(t.data_ptr());
                 return nbd_reduce_scatter_list(comm_, parts.data(), out.data_ptr(), (size_t)out.numel(), dtype, op, s);
               });
  }

  c10::intrusive_ptr<c10d::Work> reduce_scatter_tensor_coalesced(std::vector<at::Tensor>& outs,
                                                                 std::vector<at::Tensor>& ins,
                                                                 const c10d::ReduceScatterOptions& opts) override {
    TORCH_CHECK(outs.size() == ins.size() && !ins.empty(), "[b200] reduce_scatter coalesced: list sizes differ");
    for (size_t i = 0; i < ins.size(); ++i) {
      check_tensor(ins[i]);
      check_tensor(outs[i]);
      TORCH_CHECK(ins[i].numel() == outs[i].numel() * getSize() && outs[i].scalar_type() == ins[i].scalar_type(),
                  "[b200] reduce_scatter coalesced: bad input size/dtype at index ", i);
    }
    const int op = to_nbd_op(opts.reduceOp);
    return run(c10d::OpType::COALESCED, "b200:reduce_scatter_tensor_coalesced", ins[0].get_device(), ins, outs,
               opts.asyncOp, false, [&](cudaStream_t s) {
                 for (size_t i = 0; i < ins.size(); ++i) {
                   int rc = nbd_reduce_scatter(comm_, ins[i].data_ptr(), outs[i].data_ptr(), (size_t)outs[i].numel(),
                                               to_nbd_dtype(ins[i].scalar_type(), true), op, s);
                   if (rc != NBD_OK) return rc;
                 }
                 return (int)NBD_OK;
               });
  }

  // ---- point to point and the collectives built on it (csrc/nbd_p2p.cuh) ---------------------
  // What cell code may also call on the group worker.py creates (ref:src/nbdistributed/worker.py:151,
  // :162): dist.send / recv / isend / irecv / batch_isend_irecv, all_to_all[_single], gather,
  // scatter, reduce.  One nbd_p2p_batch launch runs all of a rank's sends and receives of a call
  // concurrently.  Tags are not matched (as with ProcessGroupNCCL): messages between two ranks
  // pair up in issue order.
  c10::intrusive_ptr<c10d::Work> send(std::vector<at::Tensor>& tensors, int dstRank, int /*tag*/) override {
    return p2p_one(tensors, dstRank, NBD_P2P_SEND, c10d::OpType::SEND, "b200:send");
  }

  c10::intrusive_ptr<c10d::Work> recv(std::vector<at::Tensor>& tensors, int srcRank, int /*tag*/) override {
    return p2p_one(tensors, srcRank, NBD_P2P_RECV, c10d::OpType::RECV, "b200:recv");
  }

  c10::intrusive_ptr<c10d::Work> recvAnysource(std::vector<at::Tensor>& /*tensors*/, int /*tag*/) override {
    TORCH_CHECK(false, "[b200] recv from any source is not supported (neither does backend 'nccl'): name the source rank");
  }

  bool supportsCoalescing() const override { return true; }

  // dist.batch_isend_irecv and `with dist._coalescing_manager(group, device)`: point-to-point
  // operations issued in between are collected and run as ONE batch (so their relative order
  // on different ranks cannot deadlock); broadcasts are collected into table launches (small ones
  // one direct push, a state_dict's worth of same-root tensors one scatter pipeline).
  void startCoalescing() override {
    std::lock_guard<std::mutex> lk(mu_);
    TORCH_CHECK(!coalescing_, "[b200] startCoalescing called twice");
    coalescing_ = true;
    co_ops_.clear();
    co_tensors_.clear();
    co_bcast_.clear();
  }

  c10::intrusive_ptr<c10d::Work> endCoalescing() override {
    std::vector<nbd_p2p_op> ops;
    std::vector<at::Tensor> tensors;
    std::vector<CoBcast> bc;
    {
      std::lock_guard<std::mutex> lk(mu_);
      TORCH_CHECK(coalescing_, "[b200] endCoalescing without startCoalescing");
      coalescing_ = false;
      ops.swap(co_ops_);
      tensors.swap(co_tensors_);
      bc.swap(co_bcast_);
    }
    c10::DeviceIndex dev = comm_ ? device_ : c10::cuda::current_device();
    if (!tensors.empty()) dev = tensors[0].get_device();
    return run(c10d::OpType::COALESCED, "b200:coalesced", dev, tensors, tensors, /*async_op=*/true, false,
               [&](cudaStream_t s) {
                 int rc = NBD_OK;
                 if (!bc.empty()) {
                   std::vector<nbd_bcast_item> items;
                   items.reserve(bc.size());
                   for (auto& b : bc) items.push_back(nbd_bcast_item{b.t.data_ptr(), (size_t)b.t.nbytes(), b.root});
                   rc = nbd_broadcast_table(comm_, items.data(), (int)items.size(), s);
                 }
                 if (rc == NBD_OK && !ops.empty()) rc = nbd_p2p_batch(comm_, ops.data(), (int)ops.size(), s);
                 return rc;
               });
  }

  c10::intrusive_ptr<c10d::Work> alltoall_base(at::Tensor& out, at::Tensor& in, std::vector<int64_t>& outSplits,
                                               std::vector<int64_t>& inSplits, const c10d::AllToAllOptions& opts) override {
    check_tensor(out);
    check_tensor(in);
    TORCH_CHECK(out.scalar_type() == in.scalar_type(), "[b200] all_to_all: output and input dtypes differ");
    const int N = getSize(), me = getRank();
    auto offsets = [&](const at::Tensor& t, const std::vector<int64_t>& splits, std::vector<int64_t>* off, std::vector<int64_t>* len) {
      const int64_t row = t.dim() > 0 && t.size(0) > 0 ? t.numel() / t.size(0) : 1;  // elements per index of dim 0
      off->assign(N, 0);
      len->assign(N, 0);
      if (splits.empty()) {
        TORCH_CHECK(t.dim() > 0 && t.size(0) % N == 0, "[b200] all_to_all_single: dim 0 must be divisible by world_size");
        for (int r = 0; r < N; ++r) {
          (*len)[r] = t.size(0) / N * row;
          (*off)[r] = r * (*len)[r];
        }
      } else {
        TORCH_CHECK((int)splits.size() == N, "[b200] all_to_all_single: split sizes must have world_size entries");
        int64_t acc = 0;
        for (int r = 0; r < N; ++r) {
          (*off)[r] = acc * row;
          (*len)[r] = splits[r] * row;
          acc += splits[r];
        }
        TORCH_CHECK(acc == (t.dim() > 0 ? t.size(0) : 0), "[b200] all_to_all_single: split sizes do not add up to dim 0");
      }
    };
    std::vector<int64_t> ooff, olen, ioff, ilen;
    offsets(out, outSplits, &ooff, &olen);
    offsets(in, inSplits, &ioff, &ilen);
    TORCH_CHECK(olen[me] == ilen[me], "[b200] all_to_all_single: this rank's own input and output splits differ");
    const size_t es = in.element_size();
    return run(c10d::OpType::ALLTOALL_BASE, "b200:all_to_all_single", in.get_device(), {in}, {out}, opts.asyncOp, false,
               [&](cudaStream_t s) {
                 std::vector<nbd_p2p_op> ops;
                 for (int k = 1; k < N; ++k) {  // rotated start: no peer is hit by every sender at once
                   const int p = (me + k) % N;
                   ops.push_back({NBD_P2P_SEND, p, (char*)in.data_ptr() + ioff[p] * es, (size_t)ilen[p] * es});
                 }
                 for (int k = 1; k < N; ++k) {
                   const int p = (me - k + N) % N;
                   ops.push_back({NBD_P2P_RECV, p, (char*)out.data_ptr() + ooff[p] * es, (size_t)olen[p] * es});
                 }
                 if (ilen[me] > 0 && (char*)out.data_ptr() + ooff[me] * es != (char*)in.data_ptr() + ioff[me] * es)
                   if (cudaMemcpyAsync((char*)out.data_ptr() + ooff[me] * es, (char*)in.data_ptr() + ioff[me] * es,
                                       (size_t)ilen[me] * es, cudaMemcpyDeviceToDevice, s) != cudaSuccess)
                     return (int)NBD_ERR_CUDA;
                 return nbd_p2p_batch(comm_, ops.data(), (int)ops.size(), s);
               });
  }

  c10::intrusive_ptr<c10d::Work> alltoall(std::vector<at::Tensor>& outs, std::vector<at::Tensor>& ins,
                                          const c10d::AllToAllOptions& opts) override {
    const int N = getSize(), me = getRank();
    TORCH_CHECK((int)outs.size() == N && (int)ins.size() == N, "[b200] all_to_all: tensor lists must have world_size entries");
    for (int r = 0; r < N; ++r) {
      check_tensor(outs[r]);
      check_tensor(ins[r]);
    }
    TORCH_CHECK(outs[me].nbytes() == ins[me].nbytes(), "[b200] all_to_all: this rank's own input and output sizes differ");
    std::vector<at::Tensor> all(ins.begin(), ins.end());
    all.insert(all.end(), outs.begin(), outs.end());
    return run(c10d::OpType::ALLTOALL, "b200:all_to_all", ins[0].get_device(), all, all, opts.asyncOp, false,
               [&](cudaStream_t s) {
                 std::vector<nbd_p2p_op> ops;
                 for (int k = 1; k < N; ++k) {
                   const int p = (me + k) % N;
                   ops.push_back({NBD_P2P_SEND, p, ins[p].data_ptr(), (size_t)ins[p].nbytes()});
                 }
                 for (int k = 1; k < N; ++k) {
                   const int p = (me - k + N) % N;
                   ops.push_back({NBD_P2P_RECV, p, outs[p].data_ptr(), (size_t)outs[p].nbytes()});
                 }
                 if (ins[me].nbytes() > 0 && outs[me].data_ptr() != ins[me].data_ptr())
                   if (cudaMemcpyAsync(outs[me].data_ptr(), ins[me].data_ptr(), (size_t)ins[me].nbytes(), cudaMemcpyDeviceToDevice, s) !=
                       cudaSuccess)
                     return (int)NBD_ERR_CUDA;
                 return nbd_p2p_batch(comm_, ops.data(), (int)ops.size(), s);
               });
  }

  c10::intrusive_ptr<c10d::Work> gather(std::vector<std::vector<at::Tensor>>& outs, std::vector<at::Tensor>& ins,
                                        const c10d::GatherOptions& opts) override {
    const int N = getSize(), me = getRank(), root = (int)opts.rootRank;
    TORCH_CHECK(ins.size() == 1, "[b200] gather expects one input tensor");
    TORCH_CHECK(root >= 0 && root < N, "[b200] invalid root rank ", root);
    check_tensor(ins[0]);
    std::vector<at::Tensor> all{ins[0]};
    if (me == root) {
      TORCH_CHECK(outs.size() == 1 && (int)outs[0].size() == N, "[b200] gather: the root needs world_size output tensors");
      for (auto& o : outs[0]) {
        check_tensor(o);
        TORCH_CHECK(o.nbytes() == ins[0].nbytes() && o.scalar_type() == ins[0].scalar_type(),
                    "[b200] gather: every output tensor must match the input's dtype and size");
        all.push_back(o);
      }
    } else {
      TORCH_CHECK(outs.empty() || outs[0].empty(), "[b200] gather: only the root passes output tensors");
    }
    return run(c10d::OpType::GATHER, "b200:gather", ins[0].get_device(), all, all, opts.asyncOp, false, [&](cudaStream_t s) {
      std::vector<nbd_p2p_op> ops;
      if (me == root) {
        for (int k = 1; k < N; ++k) {
          const int p = (me + k) % N;
          ops.push_back({NBD_P2P_RECV, p, outs[0][p].data_ptr(), (size_t)outs[0][p].nbytes()});
        }
        if (ins[0].nbytes() > 0 && outs[0][me].data_ptr() != ins[0].data_ptr())
          if (cudaMemcpyAsync(outs[0][me].data_ptr(), ins[0].data_ptr(), (size_t)ins[0].nbytes(), cudaMemcpyDeviceToDevice, s) !=
              cudaSuccess)
            return (int)NBD_ERR_CUDA;
      } else {
        ops.push_back({NBD_P2P_SEND, root, ins[0].data_ptr(), (size_t)ins[0].nbytes()});
      }
      return nbd_p2p_batch(comm_, ops.data(), (int)ops.size(), s);
    });
  }

  c10::intrusive_ptr<c10d::Work> scatter(std::vector<at::Tensor>& outs, std::vector<std::vector<at::Tensor>>& ins,
                                         const c10d::ScatterOptions& opts) override {
    const int N = getSize(), me = getRank(), root = (int)opts.rootRank;
    TORCH_CHECK(outs.size() == 1, "[b200] scatter expects one output tensor");
    TORCH_CHECK(root >= 0 && root < N, "[b200] invalid root rank ", root);
    check_tensor(outs[0]);
    std::vector<at::Tensor> all{outs[0]};
    if (me == root) {
      TORCH_CHECK(ins.size() == 1 && (int)ins[0].size() == N, "[b200] scatter: the root needs world_size input tensors");
      for (auto& t : ins[0]) {
        check_tensor(t);
        TORCH_CHECK(t.nbytes() == outs[0].nbytes() && t.scalar_type() == outs[0].scalar_type(),
                    "[b200] scatter: every input tensor must match the output's dtype and size");
        all.push_back(t);
      }
    } else {
      TORCH_CHECK(ins.empty() || ins[0].empty(), "[b200] scatter: only the root passes input tensors");
    }
    return run(c10d::OpType::SCATTER, "b200:scatter", outs[0].get_device(), all, all, opts.asyncOp, false, [&](cudaStream_t s) {
      std::vector<nbd_p2p_op> ops;
      if (me == root) {
        for (int k = 1; k < N; ++k) {
          const int p = (me + k) % N;
          ops.push_back({NBD_P2P_SEND, p, ins[0][p].data_ptr(), (size_t)ins[0][p].nbytes()});
        }
        if (outs[0].nbytes() > 0 && outs[0].data_ptr() != ins[0][me].data_ptr())
          if (cudaMemcpyAsync(outs[0].data_ptr(), ins[0][me].data_ptr(), (size_t)outs[0].nbytes(), cudaMemcpyDeviceToDevice, s) !=
              cudaSuccess)
            return (int)NBD_ERR_CUDA;
      } else {
        ops.push_back({NBD_P2P_RECV, root, outs[0].data_ptr(), (size_t)outs[0].nbytes()});
      }
      return nbd_p2p_batch(comm_, ops.data(), (int)ops.size(), s);
    });
  }

  // dist.reduce: the root's tensor receives the reduction; the other ranks' tensors are left
  // untouched (they reduce into a scratch tensor), as with gloo.
  c10::intrusive_ptr<c10d::Work> reduce(std::vector<at::Tensor>& tensors, const c10d::ReduceOptions& opts) override {
    TORCH_CHECK(tensors.size() == 1, "[b200] reduce expects exactly one tensor");
    auto& t = tensors[0];
    check_tensor(t);
    const int root = (int)opts.rootRank;
    TORCH_CHECK(root >= 0 && root < getSize(), "[b200] invalid root rank ", root);
    const int dtype = to_nbd_dtype(t.scalar_type(), true), op = to_nbd_op(opts.reduceOp);
    return run(c10d::OpType::REDUCE, "b200:reduce", t.get_device(), {t}, {t}, opts.asyncOp, false, [&](cudaStream_t s) {
      if (getRank() == root) return nbd_all_reduce(comm_, t.data_ptr(), t.data_ptr(), (size_t)t.numel(), dtype, op, s);
      at::Tensor scratch = at::empty_like(t);
      note_stream(scratch);
      return nbd_all_reduce(comm_, t.data_ptr(), scratch.data_ptr(), (size_t)t.numel(), dtype, op, s);
    });
  }

  c10::intrusive_ptr<c10d::Work> barrier(const c10d::BarrierOptions& opts) override {
    c10::DeviceIndex dev = device_;
    if (!comm_) {
      if (opts.device.has_value() && opts.device->has_index()) dev = opts.device->index();
      else if (!opts.device_ids.empty()) dev = (c10::DeviceIndex)opts.device_ids[0];
      else if (getBoundDeviceId().has_value()) dev = getBoundDeviceId()->index();
      else dev = c10::cuda::current_device();
    }
    return run(c10d::OpType::BARRIER, "b200:barrier", dev, {}, {}, opts.asyncOp, true,
               [&](cudaStream_t s) { return nbd_barrier(comm_, s); });
  }

  // ---- symmetric memory: Backend::allocateTensor ($TORCH/.../Backend.hpp:492-503) -------
  // Collective (every rank, same order, same size), like ncclMemAlloc + window registration.
  // Tensors obtained here are mapped by every peer; collectives on them skip the staging arena.
  bool supportsTensorAlloc(c10::DeviceIndex /*deviceIdx*/) override { return true; }

  at::Tensor allocateTensor(long size, at::TensorOptions options = {}) override {
    TORCH_CHECK(size > 0, "[b200] allocateTensor needs a positive element count");
    c10::DeviceIndex dev = options.has_device() && options.device().has_index() ? options.device().index()
                                                                               : (comm_ ? device_ : c10::cuda::current_device());
    std::lock_guard<std::mutex> lk(mu_);
    ensure_comm(dev);
    c10::cuda::CUDAGuard g(dev);
    const auto dtype = options.has_dtype() ? c10::typeMetaToScalarType(options.dtype()) : at::kFloat;
    const size_t bytes = (size_t)size * c10::elementSize(dtype);
    void* ptr = nullptr;
    std::vector<uint8_t> mine(NBD_HANDLE_BYTES);
    NBD_CHECK(nbd_mem_alloc(comm_, bytes, &ptr, mine.data()), "nbd_mem_alloc");
    const int idx = mem_seq_++;
    if (getSize() > 1) {
      store_->set(c10::str("nbd_b200/mem/", idx, "/", getRank()), mine);
      std::vector<uint8_t> all((size_t)getSize() * NBD_HANDLE_BYTES);
      for (int r = 0; r < getSize(); ++r) {
        std::vector<uint8_t> h = (r == getRank()) ? mine : store_->get(c10::str("nbd_b200/mem/", idx, "/", r));
        TORCH_CHECK(h.size() == NBD_HANDLE_BYTES, "[b200] malformed memory handle from rank ", r);
        std::memcpy(all.data() + (size_t)r * NBD_HANDLE_BYTES, h.data(), NBD_HANDLE_BYTES);
      }
      NBD_CHECK(nbd_mem_connect(comm_, ptr, all.data()), "nbd_mem_connect");
    }
    auto alive = alive_;
    nbd_comm_t comm = comm_;
    auto deleter = [alive, comm](void* p) {
      if (alive->load()) nbd_mem_free(comm, p);  // after shutdown the communicator already freed it
    };
    return at::from_blob(ptr, {(int64_t)size}, deleter,
                         at::TensorOptions().dtype(dtype).device(c10::Device(c10::kCUDA, dev)));
  }

  // Enqueue `iters` in-place SUM all_reduces of `t` back to back on the caller's stream without
  // returning to Python in between: what bench.py times for the small-message latency, where
  // the ~15 us of Python/c10d dispatch per dist.all_reduce call would otherwise hide the kernel.
  void repeat_all_reduce(at::Tensor t, int64_t iters) {
    check_tensor(t);
    const int dtype = to_nbd_dtype(t.scalar_type(), true);
    std::lock_guard<std::mutex> lk(mu_);
    ensure_comm(t.get_device());
    c10::cuda::CUDAGuard g(t.get_device());
    auto stream = c10::cuda::getCurrentCUDAStream(t.get_device());
    if (has_last_ && *last_stream_ != stream) last_done_.block(stream);
    for (int64_t i = 0; i < iters; ++i)
      NBD_CHECK(nbd_all_reduce(comm_, t.data_ptr(), t.data_ptr(), (size_t)t.numel(), dtype, NBD_SUM, stream.stream()),
                "repeat_all_reduce");
    last_done_.record(stream);
    last_stream_ = stream;
    has_last_ = true;
  }

  // Same for broadcast (root's tensor to everybody), `iters` times back to back.
  void repeat_broadcast(at::Tensor t, int64_t root, int64_t iters) {
    check_tensor(t);
    std::lock_guard<std::mutex> lk(mu_);
    ensure_comm(t.get_device());
    c10::cuda::CUDAGuard g(t.get_device());
    auto stream = c10::cuda::getCurrentCUDAStream(t.get_device());
    if (has_last_ && *last_stream_ != stream) last_done_.block(stream);
    for (int64_t i = 0; i < iters; ++i)
      NBD_CHECK(nbd_broadcast(comm_, t.data_ptr(), (size_t)t.nbytes(), NBD_BYTES, (int)root, stream.stream()), "repeat_broadcast");
    last_done_.record(stream);
    last_stream_ = stream;
    has_last_ = true;
  }

  // A list of tensors from one root in one call (a state_dict, a module's parameters): what the
  // per-parameter broadcast loop of ref:README.md:122-123 does, without one dispatcher trip per
  // tensor.  Stream-ordered on the current stream, like the loop it replaces.
  void broadcast_many(std::vector<at::Tensor> tensors, int64_t root) {
    TORCH_CHECK(root >= 0 && root < getSize(), "[b200] invalid root rank ", root);
    if (tensors.empty()) return;
    std::vector<nbd_bcast_item> items;
    items.reserve(tensors.size());
    const c10::DeviceIndex dev = tensors[0].get_device();
    for (auto& t : tensors) {
      check_tensor(t);
      TORCH_CHECK(t.get_device() == dev, "[b200] broadcast_many: tensors on different devices");
      items.push_back(nbd_bcast_item{t.data_ptr(), (size_t)t.nbytes(), (int)root});
    }
    std::lock_guard<std::mutex> lk(mu_);
    ensure_comm(dev);
    c10::cuda::CUDAGuard g(dev);
    auto stream = c10::cuda::getCurrentCUDAStream(dev);
    if (has_last_ && *last_stream_ != stream) last_done_.block(stream);
    NBD_CHECK(nbd_broadcast_table(comm_, items.data(), (int)items.size(), stream.stream()), "broadcast_many");
    last_done_.record(stream);
    last_stream_ = stream;
    has_last_ = true;
  }

  // all_reduce of CPU tensors through the GPUs: nbd_all_reduce_host pipelines H2D, the
  // collective and D2H in chunks.  Returns when `recv` holds the result (CPU-tensor semantics).
  void all_reduce_host(at::Tensor send, at::Tensor recv, int64_t op) {
    TORCH_CHECK(send.device().is_cpu() && recv.device().is_cpu(), "[b200] all_reduce_host takes CPU tensors");
    TORCH_CHECK(send.is_contiguous() && recv.is_contiguous(), "Tensors must be contiguous");
    TORCH_CHECK(send.scalar_type() == recv.scalar_type() && send.numel() == recv.numel(),
                "[b200] all_reduce_host: send and recv must match");
    const int dtype = to_nbd_dtype(send.scalar_type(), true);
    std::lock_guard<std::mutex> lk(mu_);
    const c10::DeviceIndex dev = comm_ ? device_ : c10::cuda::current_device();
    ensure_comm(dev);
    c10::cuda::CUDAGuard g(dev);
    auto stream = c10::cuda::getCurrentCUDAStream(dev);
    if (has_last_ && *last_stream_ != stream) last_done_.block(stream);
    NBD_CHECK(nbd_all_reduce_host(comm_, send.data_ptr(), recv.data_ptr(), (size_t)send.numel(), dtype, (int)op,
                                  stream.stream()),
              "all_reduce_host");
    last_done_.record(stream);
    last_stream_ = stream;
    has_last_ = true;
    C10_CUDA_CHECK(cudaStreamSynchronize(stream));
    int rc = nbd_comm_async_error(comm_);
    if (rc != NBD_OK) throw_nbd(rc, "all_reduce_host");
  }

  // ---- extras exposed to Python ------------------------------------------------
  void set_tunable(const std::string& name, uint64_t value) {
    tunables_.emplace_back(name, value);
    if (comm_) NBD_CHECK(nbd_comm_set_tunable(comm_, name.c_str(), value), "set_tunable");
  }
  std::string last_algo() { return comm_ ? nbd_last_algo(comm_) : "none"; }
  int last_zc_decision() { return comm_ ? nbd_comm_last_zc_decision(comm_) : -1; }
  // (enabled, one-line reason): whether NVSwitch multicast was set up for this group
  std::pair<bool, std::string> nvls_status() {
    if (!comm_) return {false, "no communicator yet (created by the first collective)"};
    char why[256] = "";
    const int rc = nbd_comm_nvls_status(comm_, why, sizeof(why));
    return {rc == 1, std::string(why)};
  }
  int64_t device_index() const { return comm_ ? (int64_t)device_ : -1; }
  void set_inline_sync_ops(bool v) { inline_sync_ops_ = v; }

 private:
  struct CoBcast {
    at::Tensor t;
    int root;
  };

  // What an operation collected by a coalescing block returns: a Work that is already complete
  // (the real one comes from endCoalescing).  Not null: the call travels through the dispatcher.
  c10::intrusive_ptr<c10d::Work> collected_work(const at::Tensor& t) {
    auto work = c10::make_intrusive<WorkB200>(getRank(), c10d::OpType::COALESCED, "b200:collected", t.get_device(), nullptr,
                                              std::vector<at::Tensor>{t}, false);
    c10::cuda::CUDAGuard g(t.get_device());
    work->mark_enqueued(c10::cuda::getCurrentCUDAStream(t.get_device()));
    return work;
  }

  c10::intrusive_ptr<c10d::Work> p2p_one(std::vector<at::Tensor>& tensors, int peer, int kind, c10d::OpType type, const char* title) {
    TORCH_CHECK(tensors.size() == 1, "[b200] send/recv expect exactly one tensor");
    auto& t = tensors[0];
    check_tensor(t);
    TORCH_CHECK(peer >= 0 && peer < getSize() && peer != getRank(), "[b200] send/recv: peer ", peer,
                " is not another rank of this group");
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (coalescing_) {  // collected; endCoalescing launches the whole batch and returns its Work
        co_ops_.push_back({kind, peer, t.data_ptr(), (size_t)t.nbytes()});
        co_tensors_.push_back(t);
        return collected_work(t);
      }
    }
    return run(type, title, t.get_device(), {t}, {t}, /*async_op=*/true, false, [&](cudaStream_t s) {
      nbd_p2p_op op = {kind, peer, t.data_ptr(), (size_t)t.nbytes()};
      return nbd_p2p_batch(comm_, &op, 1, s);
    });
  }

  static int64_t shutdown_wait_ms() {
    const char* e = std::getenv("NBDIST_B200_SHUTDOWN_WAIT_MS");
    return e ? std::atoll(e) : 2000;
  }

  // First collective binds the communicator to the tensor's device (not to the
  // rank: worker.py sets LOCAL_RANK=rank regardless of --gpu-ids)
  // and exchanges CUDA-IPC handles over the store torch created at
  // ref:src/nbdistributed/worker.py:151.
  void ensure_comm(c10::DeviceIndex dev) {
    if (comm_) {
      TORCH_CHECK(dev == device_, "[b200] this process group is bound to cuda:", (int)device_, " but got a tensor on cuda:",
                  (int)dev, " (one GPU per rank)");
      return;
    }
    device_ = dev;
    c10::cuda::CUDAGuard g(dev);
    size_t stage = 0;
    NBD_DBG("rank %d: creating communicator on cuda:%d", getRank(), (int)dev);
    NBD_CHECK(nbd_comm_create(&comm_, getRank(), getSize(), (int)dev, stage), "nbd_comm_create");
    nbd_comm_set_timeout_ms(comm_, (uint64_t)timeout_.count());
    for (auto& kv : tunables_) nbd_comm_set_tunable(comm_, kv.first.c_str(), kv.second);
    if (getSize() > 1) {
      std::vector<uint8_t> mine(NBD_HANDLE_BYTES);
      NBD_CHECK(nbd_comm_get_handle(comm_, mine.data()), "nbd_comm_get_handle");
      NBD_DBG("rank %d: publishing IPC handle", getRank());
      store_->set(c10::str("nbd_b200/handle/", getRank()), mine);
      NBD_DBG("rank %d: published", getRank());
      std::vector<uint8_t> all((size_t)getSize() * NBD_HANDLE_BYTES);
      for (int r = 0; r < getSize(); ++r) {
        std::vector<uint8_t> h = (r == getRank()) ? mine : store_->get(c10::str("nbd_b200/handle/", r));
        TORCH_CHECK(h.size() == NBD_HANDLE_BYTES, "[b200] malformed IPC handle from rank ", r);
        std::memcpy(all.data() + (size_t)r * NBD_HANDLE_BYTES, h.data(), NBD_HANDLE_BYTES);
      }
      NBD_DBG("rank %d: got all handles, connecting", getRank());
      int rc = nbd_comm_connect(comm_, all.data());
      NBD_DBG("rank %d: connect rc=%d", getRank(), rc);
      if (rc != NBD_OK) {
        nbd_comm_destroy(comm_);
        comm_ = nullptr;
        throw_nbd(rc, "nbd_comm_connect");
      }
    }
    comm_stream_ = c10::cuda::getStreamFromPool(/*isHighPriority=*/true, dev);
  }

  // temporaries created inside an enqueue lambda are used on cur_stream_
  void note_stream(const at::Tensor& t) {
    if (cur_stream_ && *cur_stream_ != c10::cuda::getCurrentCUDAStream(device_))
      c10::cuda::CUDACachingAllocator::recordStream(t.storage().data_ptr(), *cur_stream_);
  }

  template <typename F>
  c10::intrusive_ptr<c10d::Work> run(c10d::OpType type, const char* title, c10::DeviceIndex dev,
                                     std::vector<at::Tensor> inputs, std::vector<at::Tensor> outputs, bool async_op,
                                     bool is_barrier, F&& enqueue) {
    std::lock_guard<std::mutex> lk(mu_);
    ensure_comm(dev);
    c10::cuda::CUDAGuard g(dev);
    auto caller = c10::cuda::getCurrentCUDAStream(dev);
    // Synchronous calls (dist.all_reduce(x) from a cell: async_op=False, wait() follows
    // immediately) run in line on the caller's stream: same semantics, no event round trip.
    // Asynchronous calls (DDP buckets) run on the backend's stream so they overlap compute.
    const bool inline_op = inline_sync_ops_ && !async_op;
    c10::cuda::CUDAStream stream = inline_op ? caller : *comm_stream_;
    if (!inline_op) {
      at::cuda::CUDAEvent ready;
      ready.record(caller);
      ready.block(stream);
      for (auto& t : inputs) c10::cuda::CUDACachingAllocator::recordStream(t.storage().data_ptr(), stream);
      for (auto& t : outputs) c10::cuda::CUDACachingAllocator::recordStream(t.storage().data_ptr(), stream);
    }
    // collectives of one communicator share its staging arena and flags: never let two
    // of them overlap, whichever streams they were issued on
    if (has_last_ && *last_stream_ != stream) last_done_.block(stream);
    int rc;
    {
      c10::cuda::CUDAStreamGuard sg(stream);
      cur_stream_ = stream;
      NBD_DBG("rank %d: enqueue %s (inline=%d)", getRank(), title, (int)inline_op);
      rc = enqueue(stream.stream());
      NBD_DBG("rank %d: enqueue %s rc=%d", getRank(), title, rc);
      cur_stream_.reset();
    }
    if (rc != NBD_OK) throw_nbd(rc, title);
    auto work = c10::make_intrusive<WorkB200>(getRank(), type, title, dev, comm_, std::move(outputs), is_barrier);
    work->mark_enqueued(stream);
    last_done_.record(stream);
    last_stream_ = stream;
    has_last_ = true;
    return work;
  }

  c10::intrusive_ptr<c10d::Store> store_;
  std::chrono::milliseconds timeout_;
  std::mutex mu_;
  nbd_comm_t comm_ = nullptr;
  c10::DeviceIndex device_ = -1;
  std::optional<c10::cuda::CUDAStream> comm_stream_;
  std::optional<c10::cuda::CUDAStream> cur_stream_;
  at::cuda::CUDAEvent last_done_;
  std::optional<c10::cuda::CUDAStream> last_stream_;
  bool has_last_ = false;
  bool inline_sync_ops_ = true;
  int mem_seq_ = 0;
  std::shared_ptr<std::atomic<bool>> alive_ = std::make_shared<std::atomic<bool>>(true);
  std::vector<std::pair<std::string, uint64_t>> tunables_;
  bool coalescing_ = false;
  std::vector<nbd_p2p_op> co_ops_;
  std::vector<at::Tensor> co_tensors_;
  std::vector<CoBcast> co_bcast_;
};

c10::intrusive_ptr<c10d::Backend> create_backend(const c10::intrusive_ptr<c10d::Store>& store, int rank, int size,
                                                 const std::chrono::duration<float>& timeout) {
  return c10::make_intrusive<BackendB200>(store, rank, size,
                                          std::chrono::duration_cast<std::chrono::milliseconds>(timeout));
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "c10d backend 'b200': NVLink/NVSwitch P2P collectives for one B200 node";
  py::class_<BackendB200, c10::intrusive_ptr<BackendB200>, c10d::Backend>(m, "BackendB200")
      .def("set_tunable", &BackendB200::set_tunable)
      .def("last_algo", &BackendB200::last_algo)
      .def("last_zc_decision", &BackendB200::last_zc_decision)
      .def("nvls_status", &BackendB200::nvls_status)
      .def("all_reduce_host", &BackendB200::all_reduce_host, py::arg("send"), py::arg("recv"), py::arg("op") = 0,
           py::call_guard<py::gil_scoped_release>())
      .def("repeat_all_reduce", &BackendB200::repeat_all_reduce, py::call_guard<py::gil_scoped_release>())
      .def("repeat_broadcast", &BackendB200::repeat_broadcast, py::call_guard<py::gil_scoped_release>())
      .def("broadcast_many", &BackendB200::broadcast_many, py::arg("tensors"), py::arg("root"),
           py::call_guard<py::gil_scoped_release>())
      .def("device_index", &BackendB200::device_index)
      .def("set_inline_sync_ops", &BackendB200::set_inline_sync_ops);
  m.def("create_backend", &create_backend, py::arg("store"), py::arg("rank"), py::arg("size"), py::arg("timeout"),
        "creator for torch.distributed.Backend.register_backend");
  m.def("store_roundtrip", [](const c10::intrusive_ptr<c10d::Store>& store, const std::string& key, int nbytes) {
    // the same set/get sequence ensure_comm() uses for the IPC handles; lets the CPU test
    // suite exercise the store leg of the boundary without a GPU
    std::vector<uint8_t> v((size_t)nbytes);
    for (int i = 0; i < nbytes; ++i) v[(size_t)i] = (uint8_t)(i * 7 + 1);
    store->set(key, v);
    std::vector<uint8_t> back = store->get(key);
    return back == v;
  }, py::call_guard<py::gil_scoped_release>());
  m.def("launch_count", []() { return nbd_launch_count(); });
  m.def("nbd_version", []() { return nbd_version(); });
}
