#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 all-pairs N-body step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--bodies BODIES] [--ic uniform|plummer] [--variant ID]

Metric (BASELINE.json): Giga body-interactions/s, device-timed, plus the fraction
of B200 FP32 peak at 20 FLOP/interaction.  A step = one full all-pairs step
(N^2 interactions incl. the self term) with the fused integrator.  Workload at
every N: N = 1 048 576 bodies (BASELINE configs[3]), uniform-cube ICs, sharded
by target bodies across the N ranks (strong scaling), NCCL all-gather of the
packed positions overlapped with the own-slice interactions.

`--impl reference` times the reference's own CPU path (CPU_compute of
TestProject/validation.cpp:28-52, compiled unchanged into oracle/_ref with its
OpenMP pragma enabled) on the host cores; when oracle/_ref is absent it falls
back to the C restatement (oracle/nbody_oracle.c).  This and the cpu_baseline
leg are the only places bench.py touches oracle/.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Giga body-interactions/s (all-pairs step, 20 FLOP/interaction, device-timed)"
UNIT = "G interactions/s"
FLOP_PER_INTERACTION = 20.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bodies", "--n", dest="n", type=int, default=1 << 20)   # torchrun swallows a bare --n: use --bodies there
    ap.add_argument("--ic", default="uniform", choices=["uniform", "plummer"])
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--exchange", default="p2p", choices=["nccl", "p2p"],
                    help="multi-GPU position exchange: P2P stores fused into the step's epilogue (default; r1: +0.3%% at N=1M, "
                         "+4.4%% at N=256k on 8 GPUs), or an NCCL all-gather overlapped with the own-slice kernel")
    ap.add_argument("--barrier", default="auto", choices=["auto", "device", "collective"],
                    help="fused exchange only: cross-rank barrier inside the step kernel (flags in peer memory) or a 4-byte all-reduce")
    ap.add_argument("--cpu-series", action="store_true",
                    help="with --impl reference: also print baseline B1 of BASELINE.md (reference CPU path at N=1024x10, 16384x3, "
                         "131072x1, -O2 and -O3 builds) as extra JSON lines BEFORE the contract line")
    ap.add_argument("--ref-kernel-series", action="store_true",
                    help="with --impl reference: also print baseline B2 of BASELINE.md (the reference's kernel.cu rebuilt for sm_100, "
                         "oracle/_ref/libref_kernel_sm100*.so, as shipped and re-tuned to 256 threads) as extra JSON lines; needs a GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def config_of(args, world):
    return {"workload": f"N={args.n} bodies, {args.ic} ICs, EPS2=0.002 DT=0.1 (constants.h), one all-pairs step + integrator per step",
            "n_bodies": args.n, "ic": args.ic, "sharding": f"targets/{world}" if world > 1 else "none",
            "l2": "flushed between timed steps (256 MiB memset, untimed)"}


def ncu_traffic(n, kernel):
    """DRAM bytes per launch of the force kernel from the committed `ncu --set full` capture
    (profiles/r1_ncu_full_v105_n1m_raw.csv: dram__bytes_read.sum + dram__bytes_write.sum); None when
    no capture exists for this (N, kernel)."""
    # the candidates the engine chooses among at large N differ in source order / unit size only: same buffers, same traffic
    if n != (1 << 20) or not kernel.startswith(("T6_NT256", "T8_NT256")):
        return None
    try:
        tot = 0.0
        for line in open(os.path.join(ROOT, "profiles", "r1_ncu_full_v105_n1m_raw.csv")):
            f = line.strip().split(",")
            if f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                tot += float(f[2]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[f[1]]
        return int(tot) if tot > 0 else None
    except Exception:  # noqa: BLE001
        return None


def peaks():
    p = {"sm_max_mhz": 1965.0, "src": "fallback"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p.update(sm_max_mhz=float(m["sm_max_mhz"]), src="MEASURED_PEAKS.json sm_max_mhz")
    except Exception:  # noqa: BLE001
        pass
    return p


# ---- clocks during the timed region ----------------------------------------------------------
class ClockSampler:
    def __init__(self, index=0, period=0.1):
        self.samples, self.reasons, self.max_mhz, self.power = [], set(), None, []
        self._stop = threading.Event()
        self._t = None
        self.index, self.period = index, period
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    self.power.append(nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                except Exception:  # noqa: BLE001
                    pass
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        s = sorted(self.samples)
        out = {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "samples": len(s), "reasons": sorted(self.reasons)}
        if self.power:
            out["power_w_max"] = round(max(self.power), 1)
        return out


# ---- reference arm / cpu baseline ------------------------------------------------------------------
def cpu_reference_rate(n, steps, warmup):
    """reference CPU_compute (all host threads) on n bodies; returns (G inter/s, kind, cores, seconds/step)"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle as o
    x = o.ic_reference(n)
    v, a = np.zeros_like(x), np.zeros_like(x)
    if o.ref_lib(omp=True) is not None:
        kind, fn = "reference", (lambda: o.ref_cpu_compute(x, a, v, omp=True))
    else:
        kind, fn = "port", (lambda: o.step_snapshot(x, a, v))
    for _ in range(warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = (time.perf_counter() - t0) / steps
    return n * float(n) / dt / 1e9, kind, o.num_threads(), dt


def cpu_series():
    """baseline B1: the reference's CPU_compute at its three CPU-runnable configs"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle as o
    for build, flags in ((True, "-O2 -fopenmp"), ("o3", "-O3 -march=x86-64-v3 -fopenmp")):
        if o.ref_lib(build) is None:
            continue
        for n, steps in ((1024, 10), (16384, 3), (131072, 1)):
            x = o.ic_reference(n)
            v, a = np.zeros_like(x), np.zeros_like(x)
            o.ref_cpu_compute(x, a, v, omp=build)            # warm-up (thread pool, page faults)
            t0 = time.perf_counter()
            for _ in range(steps):
                o.ref_cpu_compute(x, a, v, omp=build)
            dt = (time.perf_counter() - t0) / steps
            print(json.dumps({"baseline": "B1 reference CPU_compute (validation.cpp:28-52)", "build": flags, "n": n, "steps": steps,
                              "threads": o.num_threads(), "ms_per_step": round(dt * 1e3, 3),
                              "g_inter_per_s": round(n * float(n) / dt / 1e9, 3)}), flush=True)


def ref_kernel_series():
    """baseline B2: the reference's own kernel.cu rebuilt for sm_100 (one-token fix at
    kernel.cu:99, nvcc -O3, no fast-math; see oracle/Makefile), timed with CUDA events around its own simulate()
    (kernel.cu:628-645).  A reported baseline; none of our kernels run here."""
    import ctypes as C
    import torch
    import nbody_b200 as nb   # seeded ICs only
    for so, label, sizes in (
            ("libref_kernel_sm100.so", "B2 reference kernel.cu rebuilt for sm_100 (THREADS_PER_BLOCK 32 as shipped)",
             ((16384, 20), (131072, 3), (1 << 20, 1))),
            ("libref_kernel_sm100_tpb256.so", "B2' reference kernel.cu, RE-TUNED to THREADS_PER_BLOCK 256 (not as shipped)",
             ((16384, 20), (131072, 3), (1 << 20, 1)))):
        path = os.path.join(ROOT, "oracle", "_ref", so)
        if not os.path.exists(path):
            continue
        sim = getattr(C.CDLL(path), "_Z8simulateP6float4S0_S0_i")
        sim.argtypes, sim.restype = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int], None
        for n, steps in sizes:
            x = torch.from_numpy(nb.ic_uniform(n, seed=12345)).cuda()
            v, a = torch.zeros_like(x), torch.zeros_like(x)
            if n <= 131072:
                sim(x.data_ptr(), a.data_ptr(), v.data_ptr(), n)   # warm-up
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                sim(x.data_ptr(), a.data_ptr(), v.data_ptr(), n)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            print(json.dumps({"baseline": label, "n": n, "ms_per_step": round(ms, 3), "g_inter_per_s": round(n * float(n) / ms / 1e6, 2)}), flush=True)
    print(json.dumps({"baseline": "B2 reference kernel.cu rebuilt for sm_100", "n": 8388608,
                      "error": "the reference's launcher throws \"Shared memory request too large.\" (kernel.cu:639-641): 16*N B > 228 KiB x 148 SMs; "
                               "not called here (a C++ exception through ctypes aborts the process); a guard-bypassed run would take ~100 s per step"}), flush=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.cpu_series:
        cpu_series()
    if args.ref_kernel_series:
        ref_kernel_series()
    n = 32768   # bounded sample: ~1 s per step on 8 cores; throughput is N-independent at this size
    rate, kind, cores, dt = cpu_reference_rate(n, args.steps, min(args.warmup, 1))
    sample = f"CPU_compute (validation.cpp:28-52, OpenMP pragma on) over {n} bodies per step ({n * n:.3g} interactions), {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(rate, 4), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": config_of(args, args.gpus),
        "cpu_baseline": {"value": round(rate, 4), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(rate, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


# ---- our arm ----------------------------------------------------------------------------------------
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import nbody_b200 as nb
    from importlib import import_module
    par = import_module("n-bodysimulation_b200.parallel")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (the N-body step has no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n, K, W = args.n, args.steps, max(args.warmup, 0)

    if args.ic == "plummer":
        x0, v0 = nb.ic_plummer(n, seed=12345)
        eps2, dt = 0.002, 0.01
    else:
        x0, v0 = nb.ic_uniform(n, seed=12345), None
        eps2, dt = 0.002, 0.1
    sim = nb.Simulation(n, eps2=eps2, dt=dt, rank=rank, world=world, device=local, variant=args.variant)
    sim.upload(x0, v0)
    exchange = args.exchange if world > 1 else "none"
    if world > 1 and exchange == "p2p":
        # peer mapping needs CUDA IPC + P2P access between all GPUs of the box; agree on it collectively
        try:
            runner = par.ShardedSimulation(sim, rank, world, device=dev, exchange="p2p", barrier=args.barrier)
            ok = torch.ones(1, device=dev)
        except Exception as e:  # noqa: BLE001
            print(f"[bench] rank {rank}: fused exchange unavailable ({e}); using the NCCL all-gather", file=sys.stderr, flush=True)
            ok = torch.zeros(1, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) < 1:
            exchange = "nccl"
            sim.close()
            sim = nb.Simulation(n, eps2=eps2, dt=dt, rank=rank, world=world, device=local, variant=args.variant)
            sim.upload(x0, v0)
    if world == 1 or exchange == "nccl":
        runner = par.ShardedSimulation(sim, rank, world, device=dev, exchange="nccl")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def timed_step():
        """one step, device-timed on the streams the kernels and the gather run on"""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(runner.compute):
            e0.record()
        runner.step()
        with torch.cuda.stream(runner.comm):
            if world == 1 or runner.exchange == "p2p":   # nothing runs on the comm stream then
                runner.comm.wait_stream(runner.compute)
            e1.record()
        runner.synchronize()
        return e0.elapsed_time(e1)

    for _ in range(W):
        timed_step()
    barrier()
    launches0 = nb.launch_count()
    total_ms = 0.0
    with ClockSampler(local) as clk:
        for _ in range(K):
            flush.zero_()
            torch.cuda.synchronize(dev)
            total_ms += timed_step()
    barrier()
    launches = nb.launch_count() - launches0
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / K
    value = n * float(n) * K / (total_ms * 1e-3) / 1e9

    # ---- e2e: host buffers in, host buffers out, every step (pinned memory) -------------------------
    e2e = None
    if not args.no_e2e:
        hx = torch.from_numpy(x0).pin_memory()
        hv = torch.zeros(n, 4).pin_memory() if v0 is None else torch.from_numpy(v0).pin_memory()
        ox = torch.empty(sim.count, 4).pin_memory()
        ov = torch.empty(sim.count, 4).pin_memory()
        oa = torch.empty(sim.count, 4).pin_memory()

        def e2e_step():
            nb._check(nb.lib.nbody_upload(sim._h, hx.data_ptr(), hv.data_ptr()))
            if world > 1:
                runner.step()
                runner.synchronize()
            else:
                sim.step(1)
            nb._check(nb.lib.nbody_download(sim._h, ox.data_ptr(), ov.data_ptr(), oa.data_ptr()))

        for _ in range(2):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            e2e_step()
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = {"value": round(n * float(n) * K / float(te.item()) / 1e9, 3), "unit": UNIT,
               "h2d_bytes_per_step": int(hx.numel() * 4 + hv.numel() * 4), "d2h_bytes_per_step": int(3 * ox.numel() * 4),
               "api": "nbody_upload + step + nbody_download per step (host float4 buffers, pinned)"}

    info = sim.info()
    pk = peaks()
    clocks = clk.summary()
    sms = info["sm_count"]
    peak_tflops = sms * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12
    achieved = FLOP_PER_INTERACTION * value * 1e9 / 1e12 / world          # per GPU, TFLOP/s
    roof = {"bound": "fp32", "achieved": round(achieved, 3), "peak": round(peak_tflops, 3), "unit": "TFLOP/s",
            "frac": round(achieved / peak_tflops, 4), "traffic": ncu_traffic(n, info["variant_name"]) if world == 1 else None,
            "traffic_unit": "DRAM bytes per launch (ncu --set full, profiles/r1_ncu_full_v105_n1m_raw.csv); algorithmic 112*N = 117 MB",
            "peak_src": f"nominal FP32: {sms} SMs x 128 lanes x 2 x {pk['sm_max_mhz']:.0f} MHz ({pk['src']}); "
                        "MEASURED_PEAKS.json holds HBM and bf16-tensor peaks only, neither bounds this kernel",
            "kernel": info["variant_name"], "launches_per_step": launches / K}
    if clocks["sm_mhz"]:
        roof["frac_at_sampled_clock"] = round(achieved / (sms * 128 * 2 * clocks["sm_mhz"] * 1e6 / 1e12), 4)
    try:   # clock64 / globaltimer inside the last timed kernel: the clock the SMs really delivered
        mhz = sim.kernel_clock_mhz()
        clocks["sm_mhz_in_kernel"] = round(mhz, 1)
        roof["frac_at_in_kernel_clock"] = round(achieved / (sms * 128 * 2 * mhz * 1e6 / 1e12), 4)
    except Exception:  # noqa: BLE001  (small-N family kernels do not record it)
        pass

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ncpu = 49152   # bounded sample sized for ~10-20 s of CPU work on 8 cores (2 timed steps)
        rate, kind, cores, dtc = cpu_reference_rate(ncpu, 2, 1)
        cpu = {"value": round(rate, 4), "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"CPU_compute (validation.cpp:28-52) x2 steps over {ncpu} bodies ({2 * ncpu * ncpu:.3g} interactions), {cores} threads, {dtc:.2f} s/step"}

    if rank == 0:
        line = {"metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": round(ms_per_step, 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config_of(args, world), "clocks": clocks,
                "e2e": e2e, "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
                "pct_fp32_peak": round(100 * achieved / peak_tflops, 2),
                "engine": {"variant": info["variant_name"], "grid_ctas": info["grid_ctas"], "ctas_per_sm": info["ctas_per_sm"],
                           "exchange": exchange, "barrier": getattr(runner, "barrier", None) if exchange == "p2p" else None,
                           "autotune": nb.autotune_report()}}
        print(json.dumps(line), flush=True)
    sim.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
