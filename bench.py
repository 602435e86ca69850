#!/usr/bin/env python
"""bench.py — the headline benchmark of the b200 collective backend.

Metric (BASELINE.json): all_reduce bus GB/s (nccl-tests convention) — plus a message-size
sweep for all_reduce / broadcast and the small-message latency as extra keys.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

* N = 1 (plain ``python bench.py``): the 1-GPU row of SURVEY.md §8(d).  R = 8 ranks are
  emulated on the one GPU by a single cooperative launch per collective (same kernels, flag
  protocol and staging as the multi-process backend; "peers" are local buffers), so the
  bound is local HBM.  A step = one fp32 SUM all_reduce of 256 MiB per rank.
* N > 1 (under ``python -m torch.distributed.run --nproc-per-node N``): one process per
  GPU through ``dist.all_reduce`` on backend "b200".  A step = the same 256 MiB all_reduce.
  The bound is NVLink (900 GB/s per direction per GPU nominal).  NCCL on the same GPUs is
  timed beside it as a comparison point (not the oracle, not the baseline).
* ``--impl reference``: the reference's own CPU path — gloo over TCP loopback, which is
  what ref:src/nbdistributed/worker.py:146-147 selects without CUDA — on the box's host
  cores, same metric and config (R or N local processes), on a bounded sample.

``value`` is the whole-job figure: the sum over ranks of bus bytes (S * 2(R-1)/R each) per
second with inputs resident in HBM; ``busbw_per_rank`` = value / ranks is the nccl-tests
number that the 900 GB/s NVLink roofline applies to.  ``e2e`` is the same metric through the
public call with pinned HOST buffers, H2D and D2H copies inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "all_reduce bus GB/s"
UNIT = "GB/s"
MSG_BYTES = 256 << 20          # per-rank message of the headline step
EMU_RANKS = 8                  # ranks emulated on one GPU at N = 1
NVLINK_PEAK_GBS = 900.0        # per direction per GPU, nominal (BASELINE.json north_star)
KERNEL_OF_ALGO = {  # last_algo() of the timed step -> the kernel that ran (real ranks, fp32 SUM)
    "ar_nvls": "nbd::k_allreduce_pipe<float,true,1,true> (multimem.ld_reduce / multimem.st)",
    "ar_pipe": "nbd::k_allreduce_pipe<float,true,1,false>",
    "ar_auto": "nbd::k_zc_agree + nbd::k_allreduce_zc<float,true,1> (staged launch gated out on the device)",
    "ar_zerocopy": "nbd::k_allreduce_zc<float,true,1>",
    "ar_twoshot": "nbd::k_allreduce_twoshot<float,true,1>",
    "ar_oneshot": "nbd::k_allreduce_oneshot<float,true,1>",
}
HBM_FALLBACK_GBS = 6650.0      # /opt/skills/guides/B200_PROFILING.md fallback


def log(msg: str) -> None:
    if int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)


def bus_factor(ranks: int) -> float:
    return 2.0 * (ranks - 1) / ranks


def measured_hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


def recorded_traffic(key: str):
    """dram bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(key)
    except Exception:
        return None


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
        0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
        0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, device_index: int = 0, period_s: float = 0.005):
        self.device_index = device_index
        self.period_s = period_s
        self.samples = []
        self.reason_mask = 0
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self._nv = None

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            self._nv = nv
            self._h = nv.nvmlDeviceGetHandleByIndex(self.device_index)
            self.max_mhz = int(nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM))
        except Exception:
            self._nv = None
            return self
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def _run(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                try:
                    self.reason_mask |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                except Exception:
                    self.reason_mask |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        s = sorted(self.samples)
        reasons = [name for bit, name in self.REASONS.items() if self.reason_mask & bit and name != "gpu_idle"]
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(s)}


def physical_gpu_index(local_index: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_index])
        except Exception:
            return local_index
    return local_index


def nvlink_bytes(device_index: int):
    """Cumulative NVLink payload bytes (tx, rx) of one GPU, all links, from NVML's throughput
    counters (KiB granularity); None where the driver does not expose them."""
    try:
        import pynvml as nv

        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(device_index)
        vals = nv.nvmlDeviceGetFieldValues(h, [(nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF),
                                               (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF)])
        if all(v.nvmlReturn == 0 for v in vals):
            return tuple(int(v.value.ullVal) * 1024 for v in vals)
        # no aggregate scope on this driver: sum the 18 links
        tx = rx = 0
        got = False
        for link in range(18):
            vals = nv.nvmlDeviceGetFieldValues(h, [(nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, link),
                                                   (nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, link)])
            if all(v.nvmlReturn == 0 for v in vals):
                tx += int(vals[0].value.ullVal) * 1024
                rx += int(vals[1].value.ullVal) * 1024
                got = True
        return (tx, rx) if got else None
    except Exception:
        return None


# --------------------------------------------------------------------------------------
# CPU baseline: the reference's gloo path on this box's host cores (bounded sample)
# --------------------------------------------------------------------------------------
def gloo_cpu_baseline(ranks: int, msg_bytes: int, iters: int, warmup: int):
    """Seconds per all_reduce (max over ranks) of the reference's CPU path: `ranks` local
    processes, backend gloo, fp32 tensors of msg_bytes each (oracle/gloo_runner.py)."""
    from oracle.gloo_runner import time_gloo

    numel = msg_bytes // 4
    t = time_gloo(ranks, [{"kind": "all_reduce", "dtype": "f32", "numel": numel, "iters": iters, "warmup": warmup}],
                  timeout=240.0)[0]
    return t


def cpu_baseline_object(ranks: int, msg_bytes: int):
    sample_bytes = min(msg_bytes, 64 << 20)  # bounded: ~10-30 s of CPU work including start-up
    iters, warm = 5, 1
    try:
        t = gloo_cpu_baseline(ranks, sample_bytes, iters, warm)
    except Exception as e:  # noqa: BLE001 - the bench line must still be printed
        return {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference",
                "sample": f"gloo baseline failed: {str(e)[:160]}"}
    per_rank = sample_bytes * bus_factor(ranks) / t / 1e9
    return {
        "value": round(per_rank * ranks, 3), "unit": UNIT, "busbw_per_rank": round(per_rank, 3),
        "cores": os.cpu_count(), "processes": ranks, "kind": "reference",
        "sample": f"gloo (torch {_torch_version()}) all_reduce fp32 SUM, {sample_bytes >> 20} MiB per rank, {ranks} local "
                  f"processes over TCP loopback, {warm} warm-up + {iters} timed iterations, {t * 1e3:.1f} ms/iter; each gloo "
                  f"rank runs 1 reduction thread + 1 I/O thread",
    }


def configs_cpu_baseline(ranks: int, configs: dict) -> None:
    """The reference's gloo path on the same two workloads, one pass each on the host cores
    (config 3: 149 broadcasts; config 4: 77 bf16 all_reduces); fills `gloo_ms` into `configs`."""
    from oracle.gloo_runner import time_gloo

    try:
        jobs = [{"kind": "broadcast", "dtype": "f32", "numel": n, "iters": 1, "warmup": 0, "root": 0} for n in gpt2_small_numels()]
        t3 = sum(time_gloo(ranks, jobs, timeout=240.0))
        configs["cfg3_gpt2_state_dict_broadcast"]["gloo_ms"] = round(t3 * 1e3, 1)
        jobs = [{"kind": "all_reduce", "dtype": "bf16", "numel": n, "iters": 1, "warmup": 0} for n in ddp_bucket_numels()]
        t4 = sum(time_gloo(ranks, jobs, timeout=300.0))
        configs["cfg4_ddp_bf16_gradient_all_reduce"]["gloo_ms"] = round(t4 * 1e3, 1)
        configs["gloo_note"] = (f"gloo, {ranks} local processes on {os.cpu_count()} host cores, one pass per tensor "
                                "(barrier-fenced per tensor, summed)")
    except Exception as e:  # noqa: BLE001 - the bench line must still be printed
        configs["gloo_note"] = f"gloo baseline failed: {str(e)[:160]}"


def _torch_version():
    import torch

    return torch.__version__


# --------------------------------------------------------------------------------------
# our arm, N = 1: emulated ranks on one GPU (HBM-bound 1-GPU row)
# --------------------------------------------------------------------------------------
def run_ours_single(args):
    import torch

    from nbdistributed_b200 import capi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 backend has no CPU fallback "
                         "(use --impl reference for the gloo CPU arm)")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    R, S = EMU_RANKS, MSG_BYTES
    numel = S // 4
    grp = capi.EmuGroup(R, device=0)
    g = torch.Generator(device=dev)
    sends = []
    for r in range(R):
        g.manual_seed(1234 + r)
        sends.append(torch.randn(numel, generator=g, device=dev, dtype=torch.float32))
    recvs = [torch.empty_like(s) for s in sends]

    def step():
        grp.all_reduce(sends, recvs)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    grp.async_error()
    algo = grp.last_algo()

    sampler = ClockSampler(physical_gpu_index(0)).start()
    l0 = capi.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    launches = capi.launch_count() - l0
    clocks = sampler.stop()
    grp.async_error()
    ms = ev0.elapsed_time(ev1) / args.steps
    per_rank = S * bus_factor(R) / (ms * 1e-3) / 1e9
    value = per_rank * R

    # correctness of what was timed: a sample re-reduced with plain fp32 adds in the kernels' fixed
    # order (ascending rank) must match bit for bit on the first and the last rank.  (The oracle
    # under oracle/ is only used by the cpu_baseline / --impl reference legs of this file.)
    idx = torch.randint(0, numel, (8192,), device=dev, generator=g)
    want = sends[0][idx].clone()
    for r in range(1, R):
        want = want + sends[r][idx]
    for r in (0, R - 1):
        if not torch.equal(recvs[r][idx].view(torch.int32), want.view(torch.int32)):
            raise SystemExit("bench.py: the timed all_reduce does not match a rank-order fp32 re-reduction")

    # ---- end to end: pinned host buffers -> H2D -> all_reduce -> D2H, every step --------
    host_in = [torch.empty(numel, dtype=torch.float32).pin_memory() for _ in range(R)]
    host_out = [torch.empty(numel, dtype=torch.float32).pin_memory() for _ in range(R)]
    for r in range(R):
        host_in[r].copy_(sends[r])
    e2e_steps = max(2, min(args.steps, 5))

    def e2e_step():
        # the library's host-buffer call: H2D, collective and D2H pipelined in chunks inside it
        grp.all_reduce_host(host_in, host_out)

    e2e_step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ev0.record()
    for _ in range(e2e_steps):
        e2e_step()
    ev1.record()
    torch.cuda.synchronize()
    e2e_ms = max(ev0.elapsed_time(ev1), (time.perf_counter() - t0) * 1e3) / e2e_steps
    e2e_value = S * bus_factor(R) * R / (e2e_ms * 1e-3) / 1e9
    # ceiling of that leg: the same pinned H2D + D2H bytes with no collective in between
    d_in = [torch.empty_like(s) for s in sends[:2]]
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def copies_only():
        for r in range(R):
            with torch.cuda.stream(s1):
                d_in[r % 2].copy_(host_in[r], non_blocking=True)
            with torch.cuda.stream(s2):
                host_out[r].copy_(recvs[r], non_blocking=True)

    copies_only()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        copies_only()
    torch.cuda.synchronize()
    ceiling_ms = (time.perf_counter() - t0) * 1e3 / 3
    del d_in

    # ---- roofline of the dominant kernel (the all_reduce launch itself) ------------------
    peak, peak_src = measured_hbm_peak()
    alg_bytes = 2.0 * R * S  # every rank's input read once + every rank's output written once
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    traffic = recorded_traffic(f"emu{R}_{algo}_f32_{S >> 20}MiB")
    kernel = {"ar_zerocopy": "nbd::k_allreduce_zc<float,true,8>", "ar_twoshot": "nbd::k_allreduce_twoshot<float,true,8>",
              "ar_oneshot": "nbd::k_allreduce_oneshot<float,true,8>"}.get(algo, algo)

    extras = {}
    if not args.no_sweep:
        extras["sweep"] = emu_sweep(torch, grp, dev, R)
    del host_in, host_out
    grp.close()

    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "busbw_per_rank": round(per_rank, 2),
        "config": {"workload": f"all_reduce fp32 SUM, {S >> 20} MiB per rank, {R} ranks emulated on 1 GPU "
                               f"(one cooperative launch; peers = local buffers), out of place, algo={algo}",
                   "ranks": R, "message_bytes": S, "l2": "inputs (2 GiB) larger than L2; no reuse between steps",
                   "value_definition": "sum over ranks of S*2(R-1)/R bytes per second"},
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": R * S, "d2h_bytes_per_step": R * S,
                "ms_per_step": round(e2e_ms, 3), "host_copy_only_ms": round(ceiling_ms, 3),
                "frac_of_host_copy_ceiling": round(ceiling_ms / e2e_ms, 3),
                "api": "nbd_emu_all_reduce_host (pinned host buffers in, pinned host buffers out; chunked H2D/collective/D2H "
                       "pipeline inside the call)"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": alg_bytes,
                     "kernel": kernel,
                     "note": "algorithmic bytes = 2*R*S: every rank's input read once, every rank's output written once"},
    }
    line.update(extras)
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_object(R, S)
    return line


def emu_sweep(torch, grp, dev, R):
    """Message-size sweep on the emulated group (extra information, outside the timed region)."""
    out = []
    for kind in ("all_reduce", "broadcast"):
        for dt, es in ((torch.float32, 4), (torch.bfloat16, 2)):
            for nbytes in (4 << 10, 64 << 10, 1 << 20, 16 << 20, 64 << 20, 256 << 20):
                n = nbytes // es
                bufs = [torch.ones(n, device=dev, dtype=dt) for _ in range(R)]
                outs = [torch.empty_like(b) for b in bufs]
                iters = 50 if nbytes <= (1 << 20) else 10
                fn = (lambda: grp.all_reduce(bufs, outs)) if kind == "all_reduce" else (lambda: grp.broadcast(bufs, root=0))
                for _ in range(3):
                    fn()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record()
                for _ in range(iters):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / iters
                factor = bus_factor(R) if kind == "all_reduce" else 1.0
                out.append({"op": kind, "dtype": str(dt).replace("torch.", ""), "bytes": nbytes, "us": round(us, 2),
                            "busbw_per_rank": round(nbytes * factor / us / 1e3, 2), "algo": grp.last_algo()})
                del bufs, outs
    return out


# --------------------------------------------------------------------------------------
# our arm, N > 1: one process per GPU through torch.distributed backend "b200"
# --------------------------------------------------------------------------------------
def _time_collective(torch, dist, fn, iters, warmup, cpu_group):
    """ms per call: CUDA events on the launching stream, barrier + synchronize on both sides,
    max over ranks."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu_group)
    return float(t.item())


# --------------------------------------------------------------------------------------
# parity of what is about to be timed, at the actual world size (driver-visible)
# --------------------------------------------------------------------------------------
def _seeded(torch, dev, rank, n, dtype, seed):
    """Rank `rank`'s synthetic input: every process can regenerate every rank's tensor."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed + rank)
    if dtype in (torch.int32, torch.int64):
        return torch.randint(-(2 ** 30), 2 ** 30, (n,), generator=g, device=dev, dtype=dtype)
    if dtype == torch.uint8:
        return torch.randint(0, 256, (n,), generator=g, device=dev, dtype=torch.uint8)
    return torch.randn(n, generator=g, device=dev, dtype=torch.float32).to(dtype)


def _rank_order_sum(torch, dev, world, n, dtype, seed):
    """Self-contained restatement of the kernels' fixed order (ascending rank; bf16 accumulates
    in fp32 and is rounded once; integers wrap) - oracle/ stays out of the product arm."""
    acc_t = torch.float32 if dtype in (torch.bfloat16, torch.float16) else dtype
    acc = _seeded(torch, dev, 0, n, dtype, seed).to(acc_t)
    for r in range(1, world):
        acc = acc + _seeded(torch, dev, r, n, dtype, seed).to(acc_t)
    return acc.to(dtype)


def parity_check(torch, dist, dev, rank, world, cpu_group, be, B):
    """All four collectives + barrier through dist.* on backend b200, every kernel family, checked
    bit for bit against a rank-order re-reduction (integers and copies also against gloo through
    the cpu group; float sums against gloo within the north_star tolerance)."""
    cases, failed, algos, notes = 0, [], set(), []

    def bits(t):
        return t.contiguous().view(torch.uint8)

    def check(name, ok):
        nonlocal cases
        cases += 1
        algos.add(be.last_algo())
        if not ok:
            failed.append(f"{name} [{be.last_algo()}]")

    def gloo_sum(x):
        c = x.detach().cpu()
        if c.dtype == torch.bfloat16:
            c = c.float()  # compare magnitudes only; gloo's bf16 rounding differs by design
        dist.all_reduce(c, group=cpu_group)
        return c

    eps = {torch.float32: 1e-6, torch.bfloat16: world * 2.0 ** -8}
    # all_reduce: LL, one-shot, two-shot, auto-registered zero-copy
    be.set_tunable("auto_min_bytes", 8 << 20)
    for dtype, n in ((torch.float32, 1000), (torch.float32, 64 << 10), (torch.float32, (1 << 20) + 3),
                     (torch.float32, 4 << 20), (torch.bfloat16, 4099), (torch.bfloat16, (2 << 20) + 8),
                     (torch.int32, 70001), (torch.int64, 1 << 20)):
        x = _seeded(torch, dev, rank, n, dtype, 4321)
        want = _rank_order_sum(torch, dev, world, n, dtype, 4321)
        reps = 3 if n * x.element_size() >= (8 << 20) else 1  # zero-copy takes over once peers imported the segment
        for _ in range(reps):
            y = x.clone()
            dist.all_reduce(y)
            check(f"all_reduce {dtype} n={n}", torch.equal(bits(y), bits(want)))
        if dtype in (torch.int32, torch.int64):
            check(f"all_reduce {dtype} n={n} vs gloo", torch.equal(y.cpu(), gloo_sum(x)))
        elif n <= (1 << 20) + 3:
            ref = gloo_sum(x).double()
            mag = gloo_sum(x.abs()).double()
            check(f"all_reduce {dtype} n={n} vs gloo (tolerance)", bool(((y.cpu().double() - ref).abs() <= eps[dtype] * mag + 1e-30).all()))
    be.set_tunable("auto_min_bytes", 0)
    # symmetric memory: zero-copy kernels
    sym = B.empty_symmetric(2 << 20, torch.float32)
    x = _seeded(torch, dev, rank, 2 << 20, torch.float32, 4400)
    sym.copy_(x)
    be.set_tunable("zc_min_bytes", 1 << 20)
    dist.all_reduce(sym)
    check("all_reduce on symmetric memory", torch.equal(bits(sym), bits(_rank_order_sum(torch, dev, world, 2 << 20, torch.float32, 4400))))
    be.set_tunable("zc_min_bytes", 0)
    del sym
    # the tile pipeline over peer-to-peer transport
    be.set_tunable("pipe", 1)
    y = _seeded(torch, dev, rank, (3 << 20) + 5, torch.bfloat16, 4500)
    dist.all_reduce(y)
    check("all_reduce tile pipeline", torch.equal(bits(y), bits(_rank_order_sum(torch, dev, world, (3 << 20) + 5, torch.bfloat16, 4500))))
    be.set_tunable("pipe", 0)
    # broadcast: direct, scatter / multicast; first and last root
    for n in (400, (1 << 20) + 17, (40 << 20) + 1):
        for root in sorted({0, world - 1}):
            be.set_tunable("nvls_bcast", 1 if (root == 0 and n > 400) else 0)  # multimem.st path where multicast is on
            y = _seeded(torch, dev, rank, n, torch.uint8, 4600 + root)
            dist.broadcast(y, src=root)
            check(f"broadcast {n}B root {root}", torch.equal(y, _seeded(torch, dev, root, n, torch.uint8, 4600 + root)))
    be.set_tunable("nvls_bcast", 0)
    # all_gather / reduce_scatter (tensor forms)
    n = (1 << 20) + 8
    x = _seeded(torch, dev, rank, n, torch.bfloat16, 4700)
    out = torch.empty(world * n, dtype=torch.bfloat16, device=dev)
    dist.all_gather_into_tensor(out, x)
    check("all_gather_into_tensor", torch.equal(bits(out), bits(torch.cat([_seeded(torch, dev, r, n, torch.bfloat16, 4700) for r in range(world)]))))
    n = 1 << 18
    x = _seeded(torch, dev, rank, world * n, torch.float32, 4800)
    o = torch.empty(n, dtype=torch.float32, device=dev)
    dist.reduce_scatter_tensor(o, x)
    want = _rank_order_sum(torch, dev, world, world * n, torch.float32, 4800)[rank * n:(rank + 1) * n]
    check("reduce_scatter_tensor", torch.equal(bits(o), bits(want)))
    dist.barrier(device_ids=[dev.index])
    cases += 1
    # NVSwitch multicast, where the box has it: copy exact, in-switch sums within tolerance,
    # identical on all ranks, exact on integer-valued data
    nv_on, nv_why = be.nvls_status()
    notes.append(f"multicast: {'on' if nv_on else 'off'} ({nv_why})")
    if nv_on:
        be.set_tunable("nvls_reduce", 2)  # 2: also between two ranks
        be.set_tunable("nvls_min_bytes", 1)  # the staged in-switch pipeline is only the default from 128 MiB on
        n = (4 << 20) + 3
        x = _seeded(torch, dev, rank, n, torch.float32, 4900)
        y = x.clone()
        dist.all_reduce(y)
        want = _rank_order_sum(torch, dev, world, n, torch.float32, 4900)
        mag = x.abs()
        dist.all_reduce(mag)
        check("in-switch all_reduce f32 (tolerance 1e-6 * sum|x|)", be.last_algo() == "ar_nvls" and
              bool(((y.double() - want.double()).abs() <= 1e-6 * mag.double() + 1e-30).all()))
        chk = y.view(torch.int32).sum(dtype=torch.int64).cpu().reshape(1)
        allchk = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allchk, chk, group=cpu_group)
        check("in-switch all_reduce identical on all ranks", all(int(a) == int(allchk[0]) for a in allchk))
        xi = _seeded(torch, dev, rank, n, torch.int32, 4950) % 1000
        yi = xi.to(torch.float32)
        dist.all_reduce(yi)
        wi = xi.clone()
        dist.all_reduce(wi)
        check("in-switch all_reduce of integer-valued fp32 (exact)", torch.equal(yi, wi.to(torch.float32)))
        be.set_tunable("nvls_reduce", 0)
        be.set_tunable("nvls_min_bytes", 0)
    torch.cuda.synchronize()
    res = {"world": world, "cases": cases, "failed": len(failed), "algos": sorted(algos), "notes": notes}
    if failed:
        res["failures"] = failed[:8]
    # every rank must agree that everything passed
    flag = torch.tensor([len(failed)], dtype=torch.int64)
    dist.all_reduce(flag, group=cpu_group)
    res["failed_any_rank"] = int(flag.item())
    return res



def bind_near_gpu(physical_index: int):
    """Run this process on the CPUs NVML names as closest to its GPU, so that the pinned host
    buffers of the end-to-end leg are allocated on that NUMA node.  Returns how many CPUs."""
    try:
        import pynvml as nv

        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(physical_index)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def host_copy_ceiling(torch, dist, dev, cpu_group, nbytes, iters=4):
    """What the box gives every GPU at once for pinned H2D + D2H of `nbytes` each way with NO
    collective in between: the ceiling of the end-to-end leg (one cudaMemcpyAsync per copy, two
    streams per rank, all ranks concurrently; ms per step, max over ranks)."""
    h_in = torch.empty(nbytes // 4, dtype=torch.float32).pin_memory()
    h_out = torch.empty(nbytes // 4, dtype=torch.float32).pin_memory()
    d_in = torch.empty(nbytes // 4, dtype=torch.float32, device=dev)
    d_out = torch.zeros(nbytes // 4, dtype=torch.float32, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def step():
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)

    step()
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)
    t0 = time.perf_counter()
    for _ in range(iters):
        step()
    torch.cuda.synchronize()
    t = torch.tensor([(time.perf_counter() - t0) * 1e3 / iters], dtype=torch.float64)
    dist.barrier(group=cpu_group)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu_group)
    del h_in, h_out, d_in, d_out
    return float(t.item())


def run_ours_multi(args):
    import torch
    import torch.distributed as dist

    from nbdistributed_b200 import backend as B
    from nbdistributed_b200 import worker_init as W

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    bound_cpus = bind_near_gpu(physical_gpu_index(local))
    ns = W.init_worker(rank, world, os.environ.get("MASTER_ADDR", "127.0.0.1"), os.environ.get("MASTER_PORT", "29500"),
                       gpu_id=local, backend="b200")
    os.environ["LOCAL_RANK"] = str(local)
    dev = torch.device("cuda", local)
    cpu_group = dist.new_group(backend="gloo")
    S = MSG_BYTES
    numel = S // 4
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    src = torch.randn(numel, generator=g, device=dev, dtype=torch.float32) * 0.01
    x = src.clone()
    be = B.get_backend()
    # first collective = lazy communicator set-up (arena, CUDA-IPC exchange over the store, NVSwitch
    # multicast rendezvous): what lengthens the reference's fixed 2 s start-up wait
    t_init = time.perf_counter()
    dist.all_reduce(torch.zeros(256, device=dev))
    torch.cuda.synchronize()
    init_ms = (time.perf_counter() - t_init) * 1e3

    def step():
        dist.all_reduce(x)

    W0 = max(args.warmup, 3)
    # parity first: every kernel family at this world size, bit for bit (see parity_check)
    log("parity check")
    pc = parity_check(torch, dist, dev, rank, world, cpu_group, be, B)
    if pc["failed_any_rank"]:
        print(f"[bench] rank {rank}: PARITY FAILURE {pc}", file=sys.stderr, flush=True)
        dist.barrier(group=cpu_group)
        raise SystemExit(f"bench.py: parity check failed on {pc['failed_any_rank']} case(s); nothing was timed")
    # in-switch reduction (NVSwitch multicast) is an opt-in of the backend; the headline uses it
    # where the box has it and says so (config.workload, parity_check.notes)
    nvls_on = bool(be.nvls_status()[0]) and os.environ.get("NBD_BENCH_NVLS", "1") != "0"
    if nvls_on:
        be.set_tunable("nvls_reduce", 1)
    for _ in range(W0):
        step()
    torch.cuda.synchronize()
    algo = be.last_algo()
    # correctness of the timed op itself: one fresh all_reduce against a rank-order re-reduction of
    # the same seeded inputs - bit for bit on the peer-to-peer paths, 1e-6 * sum|x| in-switch
    x.copy_(src)
    dist.all_reduce(x)
    want = None
    mag = None
    for r in range(world):
        gr = torch.Generator(device=dev)
        gr.manual_seed(1234 + r)
        xr = torch.randn(numel, generator=gr, device=dev, dtype=torch.float32) * 0.01
        want = xr if want is None else want + xr
        mag = xr.abs() if mag is None else mag + xr.abs()
        del xr
    if algo == "ar_nvls":
        ok = bool(((x.double() - want.double()).abs() <= 1e-6 * mag.double() + 1e-30).all())
    else:
        ok = torch.equal(x.view(torch.int32), want.view(torch.int32))
    del want, mag
    if not ok:
        raise SystemExit(f"bench.py: the timed all_reduce ({algo}) does not match a rank-order fp32 re-reduction")
    x.copy_(src)

    sampler = ClockSampler(physical_gpu_index(local)).start() if rank == 0 else None
    l0 = B.launch_count()
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)
    launches = B.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    sampler = None
    # NVLink payload bytes of the same step, counted in a SEPARATE untimed pass (NVML queries
    # stay out of the timed region)
    nvlink = None
    nvl0 = nvlink_bytes(physical_gpu_index(local)) if rank == 0 else None
    dist.barrier(group=cpu_group)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    dist.barrier(group=cpu_group)
    nvl1 = nvlink_bytes(physical_gpu_index(local)) if rank == 0 else None
    if nvl0 and nvl1:
        nvlink = {"tx_bytes_per_step": (nvl1[0] - nvl0[0]) / 5, "rx_bytes_per_step": (nvl1[1] - nvl0[1]) / 5,
                  "source": "NVML NVLINK_THROUGHPUT_DATA_TX/RX of rank 0's GPU, all links, over 5 untimed repeats of the step"}
    t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu_group)
    ms = float(t.item())
    per_rank = S * bus_factor(world) / (ms * 1e-3) / 1e9

    # ---- end to end: pinned host -> H2D -> dist.all_reduce -> D2H, each step -------------
    h_in = torch.empty(numel, dtype=torch.float32).pin_memory()
    h_out = torch.empty(numel, dtype=torch.float32).pin_memory()
    h_in.copy_(src)

    def e2e_step():
        B.all_reduce_host(h_in, out=h_out)  # pinned host in -> pinned host out, pipelined inside the backend

    log(f"timed region done: {ms:.3f} ms/step, busbw/rank {per_rank:.1f} GB/s; measuring e2e")
    e2e_ms = _time_collective(torch, dist, e2e_step, max(2, min(args.steps, 5)), 1, cpu_group)
    e2e_per_rank = S * bus_factor(world) / (e2e_ms * 1e-3) / 1e9
    ceiling_ms = host_copy_ceiling(torch, dist, dev, cpu_group, S)

    extras = {}
    if nvls_on:
        # the same step on the exact (ascending-rank) peer-to-peer path, for comparison
        be.set_tunable("nvls_reduce", 0)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        ms_p2p = _time_collective(torch, dist, step, max(3, args.steps // 2), 2, cpu_group)
        extras["p2p_exact_order"] = {"ms_per_step": round(ms_p2p, 4), "algo": be.last_algo(),
                                     "busbw_per_rank": round(S * bus_factor(world) / (ms_p2p * 1e-3) / 1e9, 2)}
    if not args.no_sweep:
        log("message-size sweep")
        extras["sweep"] = multi_sweep(torch, dist, dev, world, cpu_group, be)
        if nvls_on:
            log("in-switch reduction sweep")
            be.set_tunable("nvls_reduce", 2)  # 2: measure the switch path at any world size (the default policy uses it beyond 4 ranks)
            extras["nvls_sweep"] = multi_sweep(torch, dist, dev, world, cpu_group, be, only_all_reduce=True)
            log("in-switch reduction on multicast-pool memory (zero copy)")
            extras["nvls_registered"] = nvls_registered_sweep(torch, dist, dev, world, cpu_group, be)
            be.set_tunable("nvls_reduce", 0)
        log("device-side small-message latency")
        extras["latency"] = kernel_latency(torch, dist, dev, world, cpu_group, be)
        log("registered-memory sweep")
        extras["registered"] = registered_sweep(torch, dist, dev, world, cpu_group, be)
    if not args.no_sweep:
        log("BASELINE configs 3 and 4")
        extras["configs"] = model_configs(torch, dist, dev, rank, world, cpu_group)
    if not args.no_nccl:
        log("NCCL comparison point")
        extras["nccl"] = nccl_comparison(torch, dist, dev, world, cpu_group, S)
    log("collectives done; tearing down")

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(per_rank * world, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": W0, "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "busbw_per_rank": round(per_rank, 2),
            "config": {"workload": f"all_reduce fp32 SUM, {S >> 20} MiB per rank, {world} ranks (one process per GPU) "
                                   f"through dist.all_reduce on backend '{ns['__backend__']}', in place, algo={algo}"
                                   + (" (in-switch reduction: backend tunable nvls_reduce=1, an opt-in because the NVSwitch sums in "
                                      "its own order; checked within 1e-6*sum|x| of the rank-order result)" if algo == "ar_nvls" else ""),
                       "ranks": world, "message_bytes": S, "l2": "message (256 MiB) larger than L2",
                       "value_definition": "sum over ranks of S*2(N-1)/N bytes per second"},
            "clocks": clocks,
            "e2e": {"value": round(e2e_per_rank * world, 2), "unit": UNIT, "h2d_bytes_per_step": S * world,
                    "d2h_bytes_per_step": S * world, "ms_per_step": round(e2e_ms, 3),
                    "host_copy_only_ms": round(ceiling_ms, 3), "frac_of_host_copy_ceiling": round(ceiling_ms / e2e_ms, 3),
                    "ceiling_note": "host_copy_only_ms = the same pinned H2D + D2H on every GPU at once with no collective "
                                    "(what this box's PCIe/host memory gives); the end-to-end step cannot beat it",
                    "cpus_bound_near_gpu": bound_cpus,
                    "api": "nbdistributed_b200.backend.all_reduce_host (BackendB200 -> nbd_all_reduce_host): pinned host "
                           "buffers in and out, chunked H2D/collective/D2H pipeline inside the call"},
            "gpu_launches": int(launches),
            "parity_check": pc,
            "communicator_init_ms": round(init_ms, 1),
            "roofline": {"bound": "nvlink", "achieved": round(per_rank, 1), "peak": NVLINK_PEAK_GBS, "unit": "GB/s",
                         "frac": round(per_rank / NVLINK_PEAK_GBS, 4),
                         "traffic": (nvlink["tx_bytes_per_step"] + nvlink["rx_bytes_per_step"]) if nvlink else None,
                         "nvlink": nvlink,
                         "traffic_note": "NVLink payload bytes per launch, tx + rx of one GPU (not DRAM: ncu must not wrap a "
                                         "multi-rank run); algorithmic_bytes_per_launch counts ONE direction",
                         "peak_source": "nominal NVLink 5 per direction per GPU (BASELINE.json); measured peer copy on this "
                                        "pool: 770 GB/s (B200_PROFILING.md)",
                         "algorithmic_bytes_per_launch": S * bus_factor(world),
                         "kernel": KERNEL_OF_ALGO.get(algo, algo)},
        }
        line.update(extras)
    dist.barrier(group=cpu_group)
    W.shutdown_worker()
    if rank == 0 and not args.no_cpu_baseline:
        log("gloo CPU baseline")
        line["cpu_baseline"] = cpu_baseline_object(world, S)
        if "configs" in line:
            configs_cpu_baseline(world, line["configs"])
    return line


def registered_sweep(torch, dist, dev, world, cpu_group, be):
    """The same collectives on symmetric memory (Backend.allocate_tensor): zero-copy kernels."""
    from nbdistributed_b200 import backend as B

    out = []
    big = B.empty_symmetric((1 << 30) // 4, torch.float32)
    for nbytes in (1 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30):
        buf = big[: nbytes // 4]
        buf.fill_(1.0)
        iters = 50 if nbytes <= (1 << 20) else (20 if nbytes <= (64 << 20) else 5)
        ms = _time_collective(torch, dist, lambda: dist.all_reduce(buf), iters, 3, cpu_group)
        out.append({"op": "all_reduce", "dtype": "float32", "bytes": nbytes, "us": round(ms * 1e3, 2),
                    "busbw_per_rank": round(nbytes * bus_factor(world) / (ms * 1e-3) / 1e9, 2), "algo": be.last_algo()})
    # ZeRO-style pair on a flat bf16 buffer (BASELINE config 5, scaled to 1 GiB total)
    flat = big.view(torch.bfloat16)[: (1 << 30) // 2]
    shard = torch.empty(flat.numel() // world, dtype=torch.bfloat16, device=dev)
    flat.fill_(0.01)
    for name, fn in (("reduce_scatter", lambda: dist.reduce_scatter_tensor(shard, flat)),
                     ("all_gather", lambda: dist.all_gather_into_tensor(flat, shard))):
        ms = _time_collective(torch, dist, fn, 5, 2, cpu_group)
        out.append({"op": name, "dtype": "bfloat16", "bytes": 1 << 30, "us": round(ms * 1e3, 2),
                    "busbw_per_rank": round((1 << 30) * (world - 1) / world / (ms * 1e-3) / 1e9, 2), "algo": be.last_algo()})
    del flat, big
    return out


def gpt2_small_numels():
    """Element counts of GPT2LMHeadModel(GPT2Config()).state_dict() incl. the tied lm_head alias
    (149 tensors, BASELINE config 3; pinned by tests/test_configs_shapes.py)."""
    d, v, p, layers = 768, 50257, 1024, 12
    n = [v * d, p * d]
    for _ in range(layers):
        n += [d, d, d * 3 * d, 3 * d, d * d, d, d, d, d * 4 * d, 4 * d, 4 * d * d, d]
    return n + [d, d, v * d]


def ddp_bucket_numels():
    """bf16 gradient buckets of the 1.236 G-parameter Llama-style model of BASELINE config 4:
    the tied embedding, DDP's 1 MiB first bucket, then 25 MiB buckets."""
    params, emb = 1_235_814_400, 128256 * 2048
    sizes = [emb, (1 << 20) // 2]
    rest, cap = params - sum(sizes), (25 << 20) // 2
    while rest > 0:
        sizes.append(min(cap, rest))
        rest -= sizes[-1]
    return sizes


def model_configs(torch, dist, dev, rank, world, cpu_group):
    """BASELINE configs 3 and 4 through dist.* on backend b200 (checked, then timed).  The gloo
    figures beside them are added on rank 0 after the GPUs are released (configs_cpu_baseline)."""
    out = {}
    tensors = [(torch.randn(n, device=dev) if rank == 0 else torch.zeros(n, device=dev)) for n in gpt2_small_numels()]
    total = sum(t.numel() * 4 for t in tensors)

    def per_tensor():
        for t in tensors:
            dist.broadcast(t, src=0)

    def coalesced():
        with dist._coalescing_manager(device=dev):
            for t in tensors:
                dist.broadcast(t, src=0)

    per_tensor()
    torch.cuda.synchronize()
    chk = torch.stack([t.view(torch.int32).sum(dtype=torch.int64) for t in tensors]).cpu()
    allc = [torch.empty_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk, group=cpu_group)
    ok = all(torch.equal(a, allc[0]) for a in allc)
    ms_a = _time_collective(torch, dist, per_tensor, 5, 1, cpu_group)
    ms_b = _time_collective(torch, dist, coalesced, 5, 1, cpu_group)
    from nbdistributed_b200 import backend as B

    for t in tensors[1:]:
        if rank != 0:
            t.zero_()
    B.broadcast_tensors(tensors, src=0)  # the same 149 broadcasts as one call: one flat-table launch
    torch.cuda.synchronize()
    chk_c = torch.stack([t.view(torch.int32).sum(dtype=torch.int64) for t in tensors]).cpu()
    ok = ok and bool(torch.equal(chk_c, allc[0]))
    ms_c = _time_collective(torch, dist, lambda: B.broadcast_tensors(tensors, src=0), 5, 1, cpu_group)
    out["cfg3_gpt2_state_dict_broadcast"] = {
        "tensors": len(tensors), "bytes": total, "identical_on_all_ranks": ok,
        "ms_per_tensor_loop": round(ms_a, 3), "ms_in_coalescing_block": round(ms_b, 3), "ms_one_call": round(ms_c, 3),
        "effective_GBps": round(total / min(ms_a, ms_b, ms_c) / 1e6, 1)}
    del tensors
    buckets = [torch.full((n,), float(rank + 1), device=dev, dtype=torch.bfloat16) for n in ddp_bucket_numels()]

    def ddp_step():
        works = [dist.all_reduce(b, async_op=True) for b in reversed(buckets)]
        for w in works:
            w.wait()

    ddp_step()
    torch.cuda.synchronize()
    want = float(world * (world + 1) // 2)
    ok4 = all(bool((b[:: max(1, b.numel() // 1000)] == want).all()) for b in buckets)
    for b in buckets:
        b.fill_(0.001)
    ms4 = _time_collective(torch, dist, ddp_step, 5, 2, cpu_group)
    nbytes4 = sum(b.numel() for b in buckets) * 2
    out["cfg4_ddp_bf16_gradient_all_reduce"] = {
        "buckets": len(buckets), "bytes": nbytes4, "exact": ok4, "ms_per_step": round(ms4, 3),
        "busbw_per_rank": round(nbytes4 * bus_factor(world) / ms4 / 1e6, 1)}
    del buckets
    torch.cuda.empty_cache()
    return out


def nvls_registered_sweep(torch, dist, dev, world, cpu_group, be):
    """all_reduce on symmetric memory allocated while in-switch reduction is on: the buffer is
    bound into the multicast object, so the kernel is barrier, multimem.ld_reduce + multimem.st
    of the rank's shard, barrier - no staging."""
    from nbdistributed_b200 import backend as B

    out = []
    try:
        big = B.empty_symmetric((1 << 30) // 4, torch.float32)
    except Exception as e:  # noqa: BLE001
        return [{"unavailable": str(e)[:160]}]
    for nbytes in (1 << 20, 16 << 20, 64 << 20, 256 << 20, 1 << 30):
        buf = big[: nbytes // 4]
        buf.fill_(1.0)
        dist.all_reduce(buf)
        ok = bool((buf == float(world)).all())
        buf.zero_()
        iters = 50 if nbytes <= (1 << 20) else (20 if nbytes <= (64 << 20) else 5)
        ms = _time_collective(torch, dist, lambda: dist.all_reduce(buf), iters, 3, cpu_group)
        out.append({"op": "all_reduce", "dtype": "float32", "bytes": nbytes, "us": round(ms * 1e3, 2), "ok": ok,
                    "busbw_per_rank": round(nbytes * bus_factor(world) / (ms * 1e-3) / 1e9, 2), "algo": be.last_algo()})
    del big
    return out


def multi_sweep(torch, dist, dev, world, cpu_group, be, only_all_reduce=False):
    out = []
    if not only_all_reduce:
        # ZeRO-style pair on ordinary tensors (BASELINE config 5, scaled to 1 GiB total)
        flat = torch.full(((1 << 30) // 2,), 0.01, device=dev, dtype=torch.bfloat16)
        shard = torch.empty(flat.numel() // world, dtype=torch.bfloat16, device=dev)
        for name, fn in (("reduce_scatter", lambda: dist.reduce_scatter_tensor(shard, flat)),
                         ("all_gather", lambda: dist.all_gather_into_tensor(flat, shard))):
            for _ in range(3):  # first touches export/import the allocations (auto-registration)
                fn()
                torch.cuda.synchronize()
            ms = _time_collective(torch, dist, fn, 5, 2, cpu_group)
            out.append({"op": name, "dtype": "bfloat16", "bytes": 1 << 30, "us": round(ms * 1e3, 2),
                        "busbw_per_rank": round((1 << 30) * (world - 1) / world / (ms * 1e-3) / 1e9, 2), "algo": be.last_algo()})
        del flat, shard
    for kind in (("all_reduce",) if only_all_reduce else ("all_reduce", "broadcast")):
        for dt, es in ((torch.float32, 4), (torch.bfloat16, 2)):
            for nbytes in [1 << k for k in range(12, 31)]:  # BASELINE config 2: 4 KB ... 1 GB, 19 sizes
                if only_all_reduce and nbytes < (1 << 20):
                    continue
                if dt == torch.bfloat16 and nbytes not in (4 << 10, 1 << 20, 64 << 20, 1 << 30):
                    continue
                buf = torch.ones(nbytes // es, device=dev, dtype=dt)
                iters = 100 if nbytes <= (1 << 20) else (20 if nbytes <= (64 << 20) else 5)
                fn = (lambda: dist.all_reduce(buf)) if kind == "all_reduce" else (lambda: dist.broadcast(buf, src=0))
                if nbytes >= (8 << 20):
                    for _ in range(3):  # first touches export/import the allocation
                        fn()
                        torch.cuda.synchronize()
                ms = _time_collective(torch, dist, fn, iters, 5 if nbytes <= (1 << 20) else 2, cpu_group)
                factor = bus_factor(world) if kind == "all_reduce" else 1.0
                out.append({"op": kind, "dtype": str(dt).replace("torch.", ""), "bytes": nbytes, "us": round(ms * 1e3, 2),
                            "busbw_per_rank": round(nbytes * factor / (ms * 1e-3) / 1e9, 2), "algo": be.last_algo()})
                del buf
    return out


def kernel_latency(torch, dist, dev, world, cpu_group, be):
    """Per-operation time of small all_reduces issued back to back from C++ (no Python between
    launches): the device-side latency the 'single-digit-us' target is about."""
    out = []
    for nbytes in (8, 4 << 10, 64 << 10, 256 << 10, 1 << 20):
        buf = torch.zeros(max(1, nbytes // 4), device=dev, dtype=torch.float32)
        iters = 200
        be.repeat_all_reduce(buf, 20)
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        be.repeat_all_reduce(buf, iters)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu_group)
        out.append({"op": "all_reduce", "bytes": nbytes, "us": round(float(t.item()), 2), "algo": be.last_algo()})
    for nbytes in (8, 4 << 10, 32 << 10, 256 << 10):
        buf = torch.zeros(max(1, nbytes // 4), device=dev, dtype=torch.float32)
        iters = 200
        be.repeat_broadcast(buf, 0, 20)
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        be.repeat_broadcast(buf, 0, iters)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
        t = torch.tensor([e0.elapsed_time(e1) * 1e3 / iters], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=cpu_group)
        out.append({"op": "broadcast", "bytes": nbytes, "us": round(float(t.item()), 2), "algo": be.last_algo()})
    return out


def nccl_comparison(torch, dist, dev, world, cpu_group, S):
    """NCCL (torch-bundled) on the same GPUs, same sizes: comparison point only."""
    try:
        pg = dist.new_group(backend="nccl")
        out = []
        for nbytes in (4 << 10, 1 << 20, 64 << 20, S, 1 << 30):
            buf = torch.ones(nbytes // 4, device=dev, dtype=torch.float32)
            iters = 100 if nbytes <= (1 << 20) else (20 if nbytes <= (64 << 20) else 5)
            ms = _time_collective(torch, dist, lambda: dist.all_reduce(buf, group=pg), iters, 5, cpu_group)
            out.append({"op": "all_reduce", "dtype": "float32", "bytes": nbytes, "us": round(ms * 1e3, 2),
                        "busbw_per_rank": round(nbytes * bus_factor(world) / (ms * 1e-3) / 1e9, 2)})
            del buf
        return {"version": ".".join(map(str, torch.cuda.nccl.version())), "results": out}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": str(e)[:200]}


# --------------------------------------------------------------------------------------
# reference arm: the reference's CPU path (gloo) on the host cores
# --------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None  # under torchrun only rank 0 works; the others exit 0
    n = max(1, args.gpus)
    ranks = EMU_RANKS if n == 1 else n
    sample_bytes = MSG_BYTES  # the same 256 MiB step as our arm; ~0.15 s per iteration on gloo
    warm, iters = max(1, args.warmup), max(1, args.steps)
    t = gloo_cpu_baseline(ranks, sample_bytes, iters, warm)
    per_rank = sample_bytes * bus_factor(ranks) / t / 1e9
    value = round(per_rank * ranks, 3)
    sample = (f"gloo all_reduce fp32 SUM, {sample_bytes >> 20} MiB per rank (the full step of our arm), "
              f"{ranks} local processes over TCP loopback, {warm} warm-up + {iters} timed iterations")
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n, "steps": iters, "warmup": warm,
        "ms_per_step": round(t * 1e3, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "busbw_per_rank": round(per_rank, 3),
        "config": {"workload": f"all_reduce fp32 SUM, {ranks} ranks, reference CPU path (backend gloo, "
                               f"ref:src/nbdistributed/worker.py:146-147)", "ranks": ranks, "message_bytes": sample_bytes,
                   "value_definition": "sum over ranks of S*2(R-1)/R bytes per second"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": os.cpu_count(), "processes": ranks, "kind": "reference",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-sweep", action="store_true", help="skip the message-size sweep (extra keys)")
    ap.add_argument("--no-nccl", action="store_true", help="skip the NCCL comparison point (N > 1)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip timing gloo on the host cores")
    args = ap.parse_args()
    args.steps = max(1, args.steps)

    # stdout carries exactly ONE JSON line: libraries that print to fd 1 (NCCL's version
    # banner, for one) are sent to stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    if args.impl == "reference":
        line = run_reference(args)
    elif int(os.environ.get("WORLD_SIZE", "1")) > 1:
        line = run_ours_multi(args)
    else:
        if args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with python -m torch.distributed.run "
                             "--nproc-per-node N (one process per GPU)")
        line = run_ours_single(args)
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    os.close(real_stdout)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
