#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 convolution engine (BASELINE.json metric).

Workload (BASELINE.json configs[1]): batched 2-D LINEAR_SAME, 65536 images 100x100 (*) one shared
15x15 kernel, FP64, per GPU.  A "step" is one pass of the hot path over that batch:
kernel-spectrum kernel + fused convolution kernel (b2c_convolve_batch_device).

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this engine
  python bench.py --impl reference [...]                         # reference CPU path on host cores
  torchrun ... bench.py --gpus N ...                             # one rank per GPU, weak scaling

Prints ONE JSON line (rank 0).  Numbers:
  value     conv/s, inputs resident in HBM, CUDA events on the launching stream, max over ranks
  e2e       conv/s through b2c_convolve_batch_host with pinned HOST buffers (H2D + D2H inside)
  roofline  dominant kernel vs the measured HBM copy bandwidth (MEASURED_PEAKS.json)
  cpu_baseline  the reference's convolution_fftw.h (unmodified, over the stand-in FFT) on all host cores
The oracle/reference libraries are used here ONLY as the timed CPU baseline, never on the GPU path.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

H = W = 100
HK = WK = 15
MODE_LINEAR_SAME = 2
BATCH = 65536
SEED = 0x5EED0002
ALGO_BYTES_PER_CONV = (H * W + H * W) * 8  # 80 000 read + 80 000 written
METRIC = "2D conv/s (f64, 100x100 (*) 15x15 LINEAR_SAME batch)"
UNIT = "conv/s"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def ncu_traffic():
    """dram bytes read+written per launch of the dominant kernel from the committed ncu capture."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))["traffic_bytes_per_launch"]
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU during the timed region (NVML, ~5 ms period)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.reasons = set()
        self.max_mhz = None
        self.stop_flag = False
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.nv = None

    def sample(self):
        if not self.ok:
            return
        nv = self.nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
                     0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}
            for bit, name in names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def run(self):
        while not self.stop_flag:
            self.sample()
            time.sleep(0.005)

    def result(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------
def reference_batch_fn():
    """(callable, kind): the reference CPU implementation of the path for a batch of images."""
    import _oracle
    ref = _oracle.load_ref()
    if ref is not None:
        return ref, "reference"
    return _oracle.load_oracle(), "port"


def time_cpu_reference(n_sample, threads, reps=1):
    import numpy as np
    import _oracle
    lib, kind = reference_batch_fn()
    src = _oracle.splitmix_uniform(SEED, 0, n_sample * H * W).reshape(n_sample, H, W)
    ker = np.random.default_rng(SEED).random((HK, WK))
    best = None
    for _ in range(reps):
        t0 = time.perf_counter()
        lib.batch_fftw(MODE_LINEAR_SAME, src, ker, threads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n_sample / best, kind, best


def time_cpu_scipy_port(n_sample, threads):
    """A modern-CPU restatement of convolution_fftw.h:169-284 on scipy.fft (pocketfft, SIMD, `threads`
    workers) at the reference's own period (find_closest_factor -> 110 x 110).  Extra information only."""
    import numpy as np
    import scipy.fft as sfft
    import _oracle
    P, off = 110, HK // 2
    src = _oracle.splitmix_uniform(SEED, 0, n_sample * H * W).reshape(n_sample, H, W)
    ker = np.random.default_rng(SEED).random((HK, WK))
    t0 = time.perf_counter()
    ks = sfft.rfft2(ker, s=(P, P))
    out = sfft.irfft2(sfft.rfft2(src, s=(P, P), workers=threads) * ks, s=(P, P), workers=threads)[:, off:off + H, off:off + W]
    out = np.ascontiguousarray(out)
    dt = time.perf_counter() - t0
    return n_sample / dt, dt


def cpu_sample_size(threads, target_s):
    """Calibrate a bounded sample (about target_s seconds of CPU work on `threads` threads)."""
    rate, _, _ = time_cpu_reference(4 * threads, threads)
    n = int(max(4 * threads, min(BATCH, rate * target_s)))
    return n


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    n_sample = cpu_sample_size(cores, 2.0)
    for _ in range(min(args.warmup, 1)):
        time_cpu_reference(n_sample, cores)
    times = []
    kind = None
    for _ in range(args.steps):
        rate, kind, dt = time_cpu_reference(n_sample, cores)
        times.append(dt)
    total = sum(times)
    value = n_sample * args.steps / total
    sample = "%d of %d images per step, convolution_fftw.h %s, %d threads each with its own Workspace" % (
        n_sample, BATCH, "unmodified over stand-in FFT (no libfftw3 on box)" if kind == "reference" else "restated (oracle port)",
        cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "configs[1]: 65536 x 100x100 (*) shared 15x15 LINEAR_SAME f64 (bounded sample per step)",
                   "batch_per_step": n_sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)
    return 0


# ------------------------------------------------------------------------------------------------
def pin_to_gpu_numa_node(torch, local_rank):
    """Bind this rank's host threads (and, by first touch, its pinned staging buffers) to the NUMA node
    its GPU hangs off -- otherwise the end-to-end leg of N ranks fights over one socket's memory.
    Best effort: returns a description for the JSON line, or None when the topology cannot be read."""
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return {"numa_node": node, "cpus": len(cpus)}
    except Exception:
        return None


def allreduce_u64_sum(torch, dist, world, word):
    """Wrapping (mod 2^64) sum over ranks of one unsigned 64-bit word: the two 32-bit halves travel as
    int64 (NCCL has no unsigned 64-bit sum in torch) and are recombined, so the result equals the checksum a
    single GPU would compute over the concatenated shards."""
    t = torch.tensor([word & 0xFFFFFFFF, (word >> 32) & 0xFFFFFFFF], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    lo, hi = int(t[0].item()), int(t[1].item())
    return (lo + (hi << 32)) & 0xFFFFFFFFFFFFFFFF


# ---- secondary workloads (BASELINE.json configs 3, 4, 1 and a strong-scaling split of config 2) ------------
SEED3, SEED4 = 0x5EED0003, 0x5EED0004
N3 = 16384          # config 3: images per GPU
N4 = 16384          # config 4: signals per GPU


class Ctx:
    """What the timed regions share: torch handles, rank layout, the launching stream."""

    def __init__(self, torch, dist, b2c, rank, world, local_rank):
        self.torch, self.dist, self.b2c = torch, dist, b2c
        self.rank, self.world, self.local_rank = rank, world, local_rank
        self.stream = torch.cuda.current_stream().cuda_stream
        self.peak, self.peak_kind = measured_peak()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps, warmup=3):
        """ms per step of fn() on the launching stream (CUDA events, barrier + synchronize on both sides, max
        over ranks) and the clocks sampled during that region."""
        torch = self.torch
        for _ in range(warmup):
            fn()
        self.barrier()
        sampler = ClockSampler(self.local_rank)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler.start()
        self.barrier()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        sampler.sample()
        self.barrier()
        sampler.sample()
        sampler.stop_flag = True
        return self.max_over_ranks(ev0.elapsed_time(ev1)) / steps, sampler.result()


def device_batch_record(cx, name, mode, hs, ws, hk, wk, n, shared, seed, steps, bytes_per_conv):
    """One secondary record: `n` problems per GPU, inputs generated on the device, kernel(s) included in the step."""
    torch, b2c = cx.torch, cx.b2c
    plan = b2c.Plan(mode, hs, ws, hk, wk)
    first, _ = b2c.partition(n * cx.world, cx.world, cx.rank)
    d_src = torch.empty((n, hs, ws), dtype=torch.float64, device="cuda")
    d_ker = torch.empty((1 if shared else n, hk, wk), dtype=torch.float64, device="cuda")
    d_dst = torch.empty((n, plan.h_dst, plan.w_dst), dtype=torch.float64, device="cuda")
    b2c.fill_uniform_device(d_src.data_ptr(), d_src.numel(), seed, first * hs * ws, cx.stream)
    b2c.fill_uniform_device(d_ker.data_ptr(), d_ker.numel(), seed ^ 0xABCDEF0000, 0 if shared else first * hk * wk, cx.stream)

    def step():
        plan.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1 if shared else 0, d_dst.data_ptr(), n, cx.stream)

    ms, clocks = cx.timed(step, steps)
    cs = allreduce_u64_sum(torch, cx.dist, cx.world, b2c.checksum_device(d_dst.data_ptr(), d_dst.numel(), cx.stream)[0])
    value = n * cx.world / (ms * 1e-3)
    gbs = n * bytes_per_conv / (ms * 1e-3) / 1e9
    rec = {"name": name, "value": value, "unit": UNIT, "ms_per_step": ms, "steps": steps, "per_gpu": n,
           "fft_period": [plan.h_fft, plan.w_fft], "path": plan.info()["path"],
           "roofline": {"bound": "hbm", "achieved": gbs, "peak": cx.peak, "unit": "GB/s", "frac": gbs / cx.peak,
                        "algorithmic_bytes_per_conv": bytes_per_conv, "peak_kind": cx.peak_kind},
           "clocks": clocks, "checksum_u64": cs}
    plan.destroy()
    del d_src, d_ker, d_dst
    torch.cuda.empty_cache()
    return rec


def single_call_record(cx):
    """configs[0]: one 50x50 (*) 10x10 LINEAR_SAME convolution per call through b2c_convolve_host, which is what
    the drop-in header's convolve(ws, src, kernel) calls (pageable src/kernel, page-locked ws.dst): H2D, kernel,
    D2H and the synchronisation are all inside the timed call.  Inputs as convolution_benchmark.cc:137-145."""
    import numpy as np
    b2c = cx.b2c
    L = b2c.lib()
    rng = np.random.default_rng(1)
    src, ker = rng.random((50, 50)), rng.random((10, 10))
    plan = b2c.Plan(b2c.LINEAR_SAME, 50, 50, 10, 10)
    pinned = ctypes.c_void_p()
    b2c.capi.check(L.b2c_host_alloc(ctypes.byref(pinned), 50 * 50 * 8))
    for _ in range(20):
        b2c.capi.check(L.b2c_convolve_host(plan._h, src.ctypes.data, ker.ctypes.data, pinned))
    cx.barrier()
    sampler = ClockSampler(cx.local_rank)
    sampler.start()
    calls = 300
    t0 = time.perf_counter()
    for _ in range(calls):
        L.b2c_convolve_host(plan._h, src.ctypes.data, ker.ctypes.data, pinned)
    dt = time.perf_counter() - t0
    sampler.sample()
    sampler.stop_flag = True
    dt = cx.max_over_ranks(dt)
    got = np.ctypeslib.as_array(ctypes.cast(pinned, ctypes.POINTER(ctypes.c_double)), shape=(50, 50)).copy()
    L.b2c_host_free(pinned)
    plan.destroy()
    return {"name": "configs[0]: single 50x50 (*) 10x10 LINEAR_SAME per convolve() call, host buffers", "value": calls / dt,
            "unit": "conv/s", "us_per_call": 1e6 * dt / calls, "calls": calls, "result_norm": float(np.linalg.norm(got)),
            "clocks": sampler.result()}


def pcie_probe(cx, mb=512):
    """Raw ceiling of the end-to-end leg: one page-locked H2D copy and one D2H copy of `mb` MiB running
    concurrently on two streams, all ranks at once; GB/s each way per GPU (the slower direction)."""
    torch = cx.torch
    n = mb << 20
    h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
    h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    best = None
    for it in range(3):
        cx.barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        dt = cx.max_over_ranks(time.perf_counter() - t0)
        if it:
            best = dt if best is None else min(best, dt)
    del h_in, h_out, d_in, d_out
    return n / best / 1e9


def run_b2c_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import fftconvolution_b200 as b2c

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU fallback")
    pinned = pin_to_gpu_numa_node(torch, local_rank) if (world > 1 and not os.environ.get("B2C_BENCH_NO_PIN")) else None
    torch.cuda.set_device(local_rank)
    b2c.lib().b2c_set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n_local = args.batch                      # weak scaling: every GPU owns a full 65536-image shard
    n_global = n_local * world
    first, _ = b2c.partition(n_global, world, rank)
    stream = torch.cuda.current_stream().cuda_stream

    d_src = torch.empty((n_local, H, W), dtype=torch.float64, device="cuda")
    d_dst = torch.empty((n_local, H, W), dtype=torch.float64, device="cuda")
    b2c.fill_uniform_device(d_src.data_ptr(), d_src.numel(), SEED, first * H * W, stream)
    ker = np.random.default_rng(SEED).random((HK, WK))
    d_ker = torch.from_numpy(ker).cuda()
    plan = b2c.Plan(b2c.LINEAR_SAME, H, W, HK, WK)

    def step():
        plan.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), n_local, stream)

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    sampler = ClockSampler(local_rank)
    L = b2c.lib()
    L.b2c_kernel_timer(1)
    launches0 = L.b2c_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    barrier()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    sampler.sample()
    barrier()
    sampler.sample()
    sampler.stop_flag = True
    elapsed_ms = ev0.elapsed_time(ev1)
    launches = L.b2c_launch_count() - launches0
    avg_ms, nl = ctypes.c_double(), ctypes.c_long()
    L.b2c_kernel_timer_read(ctypes.byref(avg_ms), ctypes.byref(nl))
    L.b2c_kernel_timer(0)
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    elapsed_ms = float(t.item())
    value = n_global * args.steps / (elapsed_ms * 1e-3)

    # verification checksum of the sharded result: wrapping sum of bit patterns, all-reduced
    cs = b2c.checksum_device(d_dst.data_ptr(), d_dst.numel(), stream)
    checksum = allreduce_u64_sum(torch, dist, world, cs[0])

    # ---- end to end: pinned host buffers through the C-ABI host entry point
    e2e = None
    try:
        if args.no_e2e:
            raise RuntimeError("skipped (--no-e2e)")
        n_e2e = args.e2e_batch
        src_bytes, dst_bytes = n_e2e * H * W * 8, n_e2e * H * W * 8
        h_src = torch.empty((n_e2e, H, W), dtype=torch.float64).pin_memory()
        h_dst = torch.empty((n_e2e, H, W), dtype=torch.float64).pin_memory()
        h_src.copy_(d_src[:n_e2e])
        h_ker = torch.from_numpy(ker).pin_memory()
        e_steps = max(1, min(args.steps, 3))
        plan.convolve_batch_host_ptr(h_src.data_ptr(), h_ker.data_ptr(), 1, h_dst.data_ptr(), n_e2e)  # warm
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            plan.convolve_batch_host_ptr(h_src.data_ptr(), h_ker.data_ptr(), 1, h_dst.data_ptr(), n_e2e)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        ok = bool(torch.equal(h_dst[:64], d_dst[:64].cpu()))
        e2e = {"value": n_e2e * world * e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": src_bytes + HK * WK * 8,
               "d2h_bytes_per_step": dst_bytes, "batch_per_gpu": n_e2e, "steps": e_steps,
               "matches_device_path": ok, "host_affinity": pinned}
        del h_src, h_dst
    except Exception as ex:  # pinned allocation can fail on a small host; say so instead of faking a number
        e2e = {"value": None, "unit": UNIT, "error": str(ex)[:200]}

    cx = Ctx(torch, dist, b2c, rank, world, local_rank)
    if e2e and e2e.get("value"):
        try:   # name the ceiling of the end-to-end leg: raw concurrent page-locked copies on this box
            gbs = pcie_probe(cx)
            e2e["pcie_gbs_each_way_per_gpu"] = gbs
            e2e["roof"] = world * gbs * 1e9 / (H * W * 8)
            e2e["frac"] = e2e["value"] / e2e["roof"]
        except Exception as ex:
            e2e["pcie_probe_error"] = str(ex)[:120]

    # ---- strong scaling: the SAME 65536 images split over the ranks (b2c_partition); the all-reduced checksum of
    # the shards must equal the checksum of rank 0's own full batch from the weak run above (images [0, 65536))
    strong = None
    try:
        ref_cs = torch.tensor([cs[0] & 0xFFFFFFFF, (cs[0] >> 32) & 0xFFFFFFFF], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.broadcast(ref_cs, src=0)
        ref_word = (int(ref_cs[0].item()) + (int(ref_cs[1].item()) << 32)) & 0xFFFFFFFFFFFFFFFF
        g_begin, g_end = b2c.partition(args.batch, world, rank)
        m = g_end - g_begin
        if world > 1:
            b2c.fill_uniform_device(d_src.data_ptr(), m * H * W, SEED, g_begin * H * W, stream)

        def strong_step():
            plan.convolve_batch_device(d_src.data_ptr(), d_ker.data_ptr(), 1, d_dst.data_ptr(), m, stream)

        s_ms, s_clocks = cx.timed(strong_step, args.steps)
        s_cs = allreduce_u64_sum(torch, dist, world, b2c.checksum_device(d_dst.data_ptr(), m * H * W, stream)[0])
        strong = {"name": "configs[1] split over the ranks: 65536 images in total", "value": args.batch / (s_ms * 1e-3),
                  "unit": UNIT, "ms_per_step": s_ms, "steps": args.steps, "images_total": args.batch, "scaling": "strong",
                  "checksum_u64": s_cs, "checksum_n1": ref_word, "checksum_matches_n1": bool(s_cs == ref_word),
                  "clocks": s_clocks}
    except Exception as ex:
        strong = {"error": str(ex)[:200]}

    secondary = []
    if not args.no_secondary:
        del d_src, d_dst
        torch.cuda.empty_cache()
        jobs = [
            ("configs[2]: 100x100 (*) 100x100 CIRCULAR_FULL, per-image kernels, %d per GPU" % N3,
             b2c.CIRCULAR_FULL, 100, 100, 100, 100, N3, False, SEED3, 5, 8 * (2 * 100 * 100 + 199 * 199)),
            ("configs[1] shape with per-image kernels: 100x100 (*) 15x15 LINEAR_SAME, %d per GPU" % (2 * N3),
             b2c.LINEAR_SAME, 100, 100, 15, 15, 2 * N3, False, SEED3, 10, 8 * (2 * 100 * 100 + 15 * 15)),
            ("configs[4] small-kernel point: 100x100 (*) 3x3 LINEAR_SAME, shared kernel (direct kernel), %d per GPU" % (2 * N3),
             b2c.LINEAR_SAME, 100, 100, 3, 3, 2 * N3, True, SEED3, 20, 8 * (2 * 100 * 100)),
            ("configs[4] odd-width small-kernel point: 99x99 (*) 3x3 LINEAR_SAME, shared kernel (direct kernel), %d per GPU" % (2 * N3),
             b2c.LINEAR_SAME, 99, 99, 3, 3, 2 * N3, True, SEED3, 20, 8 * (2 * 99 * 99)),
            ("configs[4] wrapped small-kernel point: 100x100 (*) 3x3 CIRCULAR_SAME, shared kernel (direct kernel), %d per GPU" % (2 * N3),
             b2c.CIRCULAR_SAME, 100, 100, 3, 3, 2 * N3, True, SEED3, 20, 8 * (2 * 100 * 100)),
            ("configs[4] large-kernel point: 100x100 (*) 80x80 LINEAR_FULL, shared kernel (period 200 strip kernel), %d per GPU" % (N3 // 2),
             b2c.LINEAR_FULL, 100, 100, 80, 80, N3 // 2, True, SEED3, 3, 8 * (100 * 100 + 179 * 179)),
            ("configs[3] short-kernel point: 1x4096 (*) 1x3 LINEAR_SAME rows (direct kernel), %d per GPU" % (4 * N4),
             b2c.LINEAR_SAME, 1, 4096, 1, 3, 4 * N4, True, SEED4, 20, 8 * (4096 + 4096)),
            ("configs[3]: 1x4096 (*) 1x255 LINEAR_SAME rows, shared kernel, %d per GPU" % N4,
             b2c.LINEAR_SAME, 1, 4096, 1, 255, N4, True, SEED4, 20, 8 * (4096 + 4096)),
            ("configs[3]: 4096x1 (*) 255x1 LINEAR_SAME columns, shared kernel, %d per GPU" % N4,
             b2c.LINEAR_SAME, 4096, 1, 255, 1, N4, True, SEED4, 20, 8 * (4096 + 4096)),
            ("configs[3]: 1x4095 (*) 1x256 LINEAR_FULL rows (odd source, even kernel), shared kernel, %d per GPU" % N4,
             b2c.LINEAR_FULL, 1, 4095, 1, 256, N4, True, SEED4, 20, 8 * (4095 + 4350)),
        ]
        for job in jobs:
            try:
                secondary.append(device_batch_record(cx, *job))
            except Exception as ex:
                secondary.append({"name": job[0], "error": str(ex)[:200]})
        try:
            secondary.append(single_call_record(cx))
        except Exception as ex:
            secondary.append({"name": "configs[0]", "error": str(ex)[:200]})

    peak, peak_kind = measured_peak()
    kernel_s = avg_ms.value * 1e-3
    achieved = (n_local * ALGO_BYTES_PER_CONV / kernel_s) / 1e9 if kernel_s > 0 else None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": ncu_traffic(),
                "peak_kind": peak_kind,
                "kernel": ("b2c::pfa::conv_kernel" if os.environ.get("B2C_P108") == "pfa" else "b2c::t2::conv_kernel<12,9>") if (plan.h_fft, plan.w_fft) == (108, 108) else "b2c::conv_shared_kernel", "kernel_ms": avg_ms.value,
                "kernel_launches": nl.value, "algorithmic_bytes_per_launch": n_local * ALGO_BYTES_PER_CONV}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = host_cores()
        n_sample = cpu_sample_size(cores, 12.0)
        rate, kind, dt = time_cpu_reference(n_sample, cores)
        cpu_baseline = {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
                        "sample": "%d of %d images, convolution_fftw.h %s, %d threads (%.1f s)" % (
                            n_sample, BATCH,
                            "unmodified over stand-in FFT (no libfftw3 on box)" if kind == "reference" else "restated (oracle port)",
                            cores, dt)}
        try:   # one core beside all cores, and the reference's nested loops (convolution_std.h)
            n1 = max(64, int(rate / cores * 2.0))
            r1, _, d1 = time_cpu_reference(n1, 1)
            cpu_baseline["one_core"] = {"value": r1, "unit": UNIT, "cores": 1, "sample": "%d images (%.1f s)" % (n1, d1)}
            import _oracle
            ref = _oracle.load_ref()
            if ref is not None:
                import numpy as np
                ns = 64 * cores
                s_src = _oracle.splitmix_uniform(SEED, 0, ns * H * W).reshape(ns, H, W)
                t0 = time.perf_counter()
                out_std = ref.batch_std(1, s_src, ker, threads=cores)   # STD_Convolution::LINEAR_SAME
                d_std = time.perf_counter() - t0
                out_fft = ref.batch_fftw(MODE_LINEAR_SAME, s_src[:4], ker, threads=1)
                cpu_baseline["nested_loops"] = {"value": ns / d_std, "unit": UNIT, "cores": cores,
                                                "sample": "%d images, convolution_std.h unmodified (%.2f s)" % (ns, d_std),
                                                "agrees_with_fft_path": bool(np.allclose(out_std[:4], out_fft, rtol=1e-10, atol=1e-10))}
        except Exception as ex:
            cpu_baseline["one_core"] = {"error": str(ex)[:100]}
        try:
            n2 = 4096
            r2, d2 = time_cpu_scipy_port(n2, cores)
            cpu_baseline["scipy_pocketfft_port"] = {"value": r2, "unit": UNIT, "cores": cores,
                                                    "sample": "%d images, rfft2/irfft2 at period 110 (%.2f s)" % (n2, d2)}
        except Exception as ex:
            cpu_baseline["scipy_pocketfft_port"] = {"error": str(ex)[:100]}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": elapsed_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[1]: 65536 x 100x100 (*) shared 15x15 LINEAR_SAME f64, per GPU",
                       "batch_per_gpu": n_local, "fft_period": [plan.h_fft, plan.w_fft],
                       "l2": "inputs 5.24 GB + outputs 5.24 GB per GPU, far larger than the 126 MB L2 (no flush needed)",
                       "rng": "splitmix64 seed 0x5EED0002 on device"},
            "roofline": roofline, "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": sampler.result(), "checksum_u64": checksum,
            "strong": strong, "secondary": secondary,
        }
        emit_line(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


_RESULT_FD = None


def claim_stdout():
    """The contract is ONE JSON line on stdout, but libraries write there too (NCCL prints its version
    banner on stdout).  Keep a private handle on the real stdout for the result line and point fd 1 at
    stderr for everything else."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    os.write(_RESULT_FD if _RESULT_FD is not None else 1, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b2c", choices=["b2c", "reference"])
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--e2e-batch", type=int, default=BATCH)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: skip the host-buffer leg")
    ap.add_argument("--no-secondary", action="store_true", help="profiling runs: skip the secondary workloads")
    args = ap.parse_args()
    claim_stdout()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b2c_arm(args)


if __name__ == "__main__":
    sys.exit(main())
