#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 all-pairs N-body step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--bodies BODIES] [--ic uniform|plummer] [--variant ID]
                    [--exchange p2p|nccl] [--driver torchrun|capi]

Metric (BASELINE.json): Giga body-interactions/s, device-timed, plus the fraction
of B200 FP32 peak at 20 FLOP/interaction.  A step = one full all-pairs step
(N^2 interactions incl. the self term) with the fused integrator.  Workload at
every N: N = 1 048 576 bodies (BASELINE configs[3]), uniform-cube ICs, sharded
by target bodies across the N ranks (strong scaling).  At N > 1 the one JSON line
also carries `series` (BASELINE configs[4]: N = 8 388 608 on 8 GPUs = the weak
series N = G x 2^20 at every G), a `parity` block (the sharded path checked against
the fp64 checker and against the same engine on one GPU, barrier time-outs) and
`capi` (the same step driven by ONE process through the multi-device C ABI).

`--impl reference` times the reference's own CPU path (CPU_compute of
TestProject/validation.cpp:28-52, compiled unchanged into oracle/_ref with its
OpenMP pragma enabled) on all host cores; when oracle/_ref is absent it falls
back to the C restatement (oracle/nbody_oracle.c).  That arm, the cpu_baseline /
ref_gpu_baseline legs and the parity checks (all AFTER the timed regions) are the
only places bench.py touches oracle/.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Giga body-interactions/s (all-pairs step, 20 FLOP/interaction, device-timed)"
UNIT = "G interactions/s"
FLOP_PER_INTERACTION = 20.0
CPU_SAMPLE_BODIES = 32768   # bounded sample of the reference arm: ~0.15-1 s per step on 8-32 host threads


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
                    help="multi-GPU position exchange: P2P stores fused into the step's epilogue (default), or an NCCL "
                         "all-gather overlapped with the own-slice kernel")
    ap.add_argument("--barrier", default="auto", choices=["auto", "device", "collective"],
                    help="fused exchange only: cross-rank barrier inside the step kernel (flags in peer memory) or a 4-byte all-reduce")
    ap.add_argument("--driver", default="torchrun", choices=["torchrun", "capi"],
                    help="capi: ONE process drives all --gpus devices through the multi-device C ABI (nbody_params.n_devices); "
                         "prints its own JSON line (the torchrun form embeds it under `capi`)")
    ap.add_argument("--cpu-series", action="store_true",
                    help="with --impl reference: also print baseline B1 of BASELINE.md (reference CPU path at N=1024x10, 16384x3, "
                         "131072x1, -O2 and -O3 builds) as extra JSON lines BEFORE the contract line")
    ap.add_argument("--ref-kernel-noguard", type=int, default=0, metavar="N",
                    help="--impl reference: also time ONE step of the reference kernel.cu at N bodies with its launcher's "
                         "total-shared-memory guard dropped (oracle/_ref/libref_kernel_sm100_noguard.so; N = 8388608 takes ~100 s); needs a GPU")
    ap.add_argument("--ref-kernel-series", action="store_true",
                    help="with --impl reference: also print baseline B2 of BASELINE.md (the reference's kernel.cu rebuilt for sm_100, "
                         "oracle/_ref/libref_kernel_sm100*.so, as shipped and re-tuned to 256 threads) as extra JSON lines; needs a GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-series", action="store_true", help="skip the weak / config-5 series at N > 1")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity block (checker legs after the timed regions)")
    ap.add_argument("--no-capi", action="store_true", help="skip the one-process C-ABI leg at N > 1")
    ap.add_argument("--no-extras", action="store_true", help="skip small_n and ref_gpu_baseline")
    return ap.parse_args()


def config_of(args, world, sample=None):
    w = f"N={args.n} bodies, {args.ic} ICs, EPS2=0.002 DT=0.1 (constants.h), one all-pairs step + integrator per step"
    if sample:
        w += f"; {sample}"
    return {"workload": w, "n_bodies": args.n, "ic": args.ic, "sharding": f"targets/{world}" if world > 1 else "none",
            "l2": "flushed between timed steps (256 MiB memset on the step's stream, outside the per-step event pairs)"}


def peaks():
    p = {"sm_max_mhz": 1965.0, "src": "fallback"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p.update(sm_max_mhz=float(m["sm_max_mhz"]), src="MEASURED_PEAKS.json sm_max_mhz")
    except Exception:  # noqa: BLE001
        pass
    return p


def ncu_traffic(n, kernel):
    """DRAM bytes per launch of the force kernel from the committed `ncu --set full` capture; None when no capture
    exists for this (N, kernel family)."""
    if n != (1 << 20) or not kernel.startswith("T6_NT256"):
        return None, None
    for name in ("r2_ncu_full_n1m_raw.csv", "r1_ncu_full_v105_n1m_raw.csv"):
        try:
            tot = 0.0
            for line in open(os.path.join(ROOT, "profiles", name)):
                f = line.strip().split(",")
                if f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    tot += float(f[2]) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[f[1]]
            if tot > 0:
                return int(tot), name
        except Exception:  # noqa: BLE001
            continue
    return None, None


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
def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as o
    o.set_num_threads()   # all host cores, whatever OMP_NUM_THREADS says (torchrun exports 1 to its workers)
    return o


def cpu_reference_rate(n, steps, warmup):
    """reference CPU_compute (all host threads) on n bodies; returns (G inter/s, kind, cores, seconds/step)"""
    import numpy as np
    o = _oracle()
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
    import numpy as np
    o = _oracle()
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


REF_KERNEL_SYM = "_Z8simulateP6float4S0_S0_i"   # void simulate(float4*, float4*, float4*, int), TestProject/kernel.cu:628


def ref_kernel_rate(so, n, steps, warm):
    """the reference's own kernel.cu (rebuilt for sm_100, oracle/Makefile) through its own simulate(); cudaEvent-timed"""
    import ctypes as C
    import torch
    import nbody_b200 as nb   # seeded ICs only
    path = os.path.join(ROOT, "oracle", "_ref", so)
    if not os.path.exists(path):
        return None
    sim = getattr(C.CDLL(path), REF_KERNEL_SYM)
    sim.argtypes, sim.restype = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int], None
    x = torch.from_numpy(nb.ic_uniform(n, seed=12345)).cuda()
    v, a = torch.zeros_like(x), torch.zeros_like(x)
    for _ in range(warm):
        sim(x.data_ptr(), a.data_ptr(), v.data_ptr(), n)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sim(x.data_ptr(), a.data_ptr(), v.data_ptr(), n)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return ms, n * float(n) / ms / 1e6


def ref_kernel_series():
    """baseline B2: reference kernel.cu, one-token fix at :99, nvcc -O3, no fast-math.
    A reported baseline; none of our kernels run here."""
    for so, label in (("libref_kernel_sm100.so", "B2 reference kernel.cu rebuilt for sm_100 (THREADS_PER_BLOCK 32 as shipped)"),
                      ("libref_kernel_sm100_tpb256.so", "B2' reference kernel.cu, RE-TUNED to THREADS_PER_BLOCK 256 (not as shipped)")):
        for n, steps in ((16384, 20), (131072, 3), (1 << 20, 1)):
            r = ref_kernel_rate(so, n, steps, 1 if n <= 131072 else 0)
            if r:
                print(json.dumps({"baseline": label, "n": n, "ms_per_step": round(r[0], 3), "g_inter_per_s": round(r[1], 2)}), flush=True)
    print(json.dumps({"baseline": "B2 reference kernel.cu rebuilt for sm_100", "n": 8388608,
                      "error": "the reference's launcher throws \"Shared memory request too large.\" (kernel.cu:639-641): 16*N B > 228 KiB x 148 SMs; "
                               "not called here (a C++ exception through ctypes aborts the process); guard bypassed: --ref-kernel-noguard 8388608 (96.5 s per step = 729 G/s, profiles/r2_reference_kernel_noguard_8m.json)"}), flush=True)


def ref_kernel_noguard(n):
    """BASELINE.md section 3, B2 at N = 8 388 608: the reference launcher refuses this size ("Shared memory request too
    large.", kernel.cu:639-641); the same kernel.cu with that half of the guard dropped, one step, labelled as such"""
    ref_kernel_rate("libref_kernel_sm100_noguard.so", 16384, 1, 0)     # context + module load outside the timed step
    r = ref_kernel_rate("libref_kernel_sm100_noguard.so", n, 1, 0)
    if r:
        print(json.dumps({"baseline": "B2 reference kernel.cu rebuilt for sm_100, launcher guard BYPASSED (kernel.cu:639-641, second half of the condition dropped in a /tmp copy; not the reference as shipped)",
                          "n": n, "steps": 1, "ms_per_step": round(r[0], 1), "g_inter_per_s": round(r[1], 2)}), flush=True)
    else:
        print(json.dumps({"baseline": "B2 guard bypassed", "error": "oracle/_ref/libref_kernel_sm100_noguard.so not built"}), flush=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.ref_kernel_noguard:
        ref_kernel_noguard(args.ref_kernel_noguard)
    if args.cpu_series:
        cpu_series()
    if args.ref_kernel_series:
        ref_kernel_series()
    n = CPU_SAMPLE_BODIES
    rate, kind, cores, dt = cpu_reference_rate(n, args.steps, min(args.warmup, 1))
    sample = (f"CPU_compute (validation.cpp:28-52, OpenMP pragma on) over {n} bodies per step ({n * n:.3g} interactions), "
              f"{cores} threads (all host cores, set explicitly)")
    note = (f"REFERENCE ARM: CPU path sampled at N={n} per step (the all-pairs rate is N-independent at this size; "
            f"a full N={args.n} step would take ~{args.n * float(args.n) / rate / 1e9:.0f} s); GPU arm runs N={args.n}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(rate, 4), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": config_of(args, args.gpus, note),
        "cpu_baseline": {"value": round(rate, 4), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(rate, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "sampled_n_bodies": n, "gpu_launches": 0}), flush=True)


# ---- helpers of our arm ------------------------------------------------------------------------------
def norm_rel(p, q):
    import numpy as np
    p, q = np.asarray(p, np.float64)[:, :3], np.asarray(q, np.float64)[:, :3]
    d = np.linalg.norm(p - q, axis=1)
    den = np.linalg.norm(q, axis=1)
    return d / np.where(den > 0, den, 1.0)


def stats(e):
    import numpy as np
    return {"max": float(e.max()), "p99.9": float(np.quantile(e, 0.999)), "median": float(np.median(e))}


class Timed:
    """K steps enqueued back to back on the runner's streams -- no host synchronisation inside the region, so the
    ranks stay in the lock step the exchange itself imposes (a host sync per step lets them drift apart by host
    jitter, which the next kernel's flag wait then absorbs INSIDE its event pair: the 0.90 of r1's SCALE record).
    Every step has its own event pair; the 256 MiB L2 flush runs on the same stream between the pairs."""

    def __init__(self, torch, runner, world, flush):
        self.torch, self.runner, self.world, self.flush = torch, runner, world, flush

    def run(self, K):
        torch, r = self.torch, self.runner
        nccl = self.world > 1 and r.exchange == "nccl"
        pairs = []
        eb, ee = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eb.record(r.compute)
        for _ in range(K):
            if self.flush is not None:
                with torch.cuda.stream(r.compute):
                    self.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(r.compute)
            r.step()
            if nccl:
                e1.record(r.comm)          # the step ends when the gathered positions have landed
            else:
                e1.record(r.compute)
            pairs.append((e0, e1))
        if nccl:
            r.compute.wait_stream(r.comm)
        ee.record(r.compute)
        r.synchronize()
        per_step = [a.elapsed_time(b) for a, b in pairs]
        region = eb.elapsed_time(ee)
        # fused exchange / one GPU: the sum of the per-step pairs (flushes excluded).  NCCL exchange: the own-slice
        # kernel of step s+1 overlaps the gather of step s, so the pairs overlap -- the whole region counts there.
        return (region if nccl else sum(per_step)), region, per_step


def make_sim(nb, par, torch, dist, args, n, x0, v0, eps2, dt, rank, world, local, dev, exchange, barrier):
    """this rank's context + runner; rank 0 resolves the kernel variant first and everybody runs that one"""
    variant = args.variant
    sim = None
    if world > 1:
        vt = torch.zeros(1, dtype=torch.int32, device=dev)
        if rank == 0:
            sim = nb.Simulation(n, eps2=eps2, dt=dt, rank=rank, world=world, device=local, variant=variant)
            vt[0] = sim.info()["variant"]
        dist.broadcast(vt, 0)
        variant = int(vt.item())
    if sim is None:
        sim = nb.Simulation(n, eps2=eps2, dt=dt, rank=rank, world=world, device=local, variant=variant)
    sim.upload(x0, v0)
    if world > 1 and exchange == "p2p":
        # peer mapping needs CUDA IPC + P2P access between all GPUs of the box; agree on it collectively
        try:
            runner = par.ShardedSimulation(sim, rank, world, device=dev, exchange="p2p", barrier=barrier)
            ok = torch.ones(1, device=dev)
        except Exception as e:  # noqa: BLE001
            print(f"[bench] rank {rank}: fused exchange unavailable ({e}); using the NCCL all-gather", file=sys.stderr, flush=True)
            ok = torch.zeros(1, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok.item()) < 1:
            exchange = "nccl"
            sim.close()
            sim = nb.Simulation(n, eps2=eps2, dt=dt, rank=rank, world=world, device=local, variant=variant)
            sim.upload(x0, v0)
    if world == 1 or exchange == "nccl":
        runner = par.ShardedSimulation(sim, rank, world, device=dev, exchange="nccl")
    return sim, runner, (exchange if world > 1 else "none")


def gather_rows(dist, world, begin, arr):
    """rows of every rank in body order (host arrays; outside every timed region)"""
    import numpy as np
    if world == 1:
        return arr
    parts = [None] * world
    dist.all_gather_object(parts, (begin, arr))
    parts.sort(key=lambda p: p[0])
    return np.concatenate([p[1] for p in parts])


def check_last_step(o, nb, dist, torch, dev, sim, runner, world, n, eps2, k_samples=64):
    """(i) of the parity block: one more step from the CURRENT state; K sampled owned targets of every rank against
    the checker's fp64 sum over the gathered positions, with the reference's own fp32 arithmetic beside it."""
    import numpy as np
    runner.synchronize()
    x_pre_own, _, _ = sim.download()
    x_pre = np.ascontiguousarray(gather_rows(dist, world, sim.begin, x_pre_own))
    runner.step()
    runner.synchronize()
    x1, v1, a1 = sim.download()
    loc = np.unique(np.linspace(0, sim.count - 1, min(k_samples, sim.count)).astype(np.int64))
    idx = (sim.begin + loc).astype(np.int32)
    a64 = o.accel_f64(x_pre, idx, eps2)
    err = float(norm_rel(a1[loc], a64).max())
    ref = float(norm_rel(o.accel_f32(x_pre, idx, eps2), a64).max())
    finite = bool(np.isfinite(a1).all() and np.isfinite(x1).all() and np.isfinite(v1).all())
    mass_ok = bool(np.array_equal(x1[:, 3], x_pre_own[:, 3]) and np.all(a1[:, 3] == 0))
    t = torch.tensor([err, ref, 0.0 if finite and mass_ok else 1.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    err, ref, bad = (float(v) for v in t.tolist())
    return {"sampled_targets_per_rank": int(len(loc)), "err_a_vs_fp64": err, "ref_fp32_err_vs_fp64": ref,
            "finite_and_mass_untouched": bad == 0.0, "tol": "err_a_vs_fp64 <= max(1e-5, 2 x ref_fp32_err_vs_fp64) (SURVEY section 8c)",
            "ok": bool(bad == 0.0 and err <= max(1e-5, 2 * ref))}


def barrier_health(torch, dist, dev, sim, world):
    t = torch.tensor([sim.shard_barrier_timeouts() if world > 1 else 0], dtype=torch.int32, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return int(t.item())


def versus_one_gpu(o, nb, dist, torch, dev, local, sim, runner, rank, world, n, x0, v0, eps2, dt, variant):
    """(ii) of the parity block: ONE step of the sharded run from the initial conditions against the same engine,
    unsharded, on rank 0's GPU.  Same input positions, so the accelerations differ by summation order only; the worst
    bodies are listed with their fp64 conditioning (sum |a_ij| / |sum a_ij|) and both engines' error against fp64."""
    import numpy as np
    runner.quiesce()
    sim.upload(x0, v0)
    runner.quiesce()
    runner.step()
    runner.synchronize()
    x, v, a = sim.download()
    X, V, A = (gather_rows(dist, world, sim.begin, t) for t in (x, v, a))
    out = None
    if rank == 0:
        with nb.Simulation(n, eps2=eps2, dt=dt, device=local) as one:
            one.upload(x0, v0).step(1)
            xr, vr, ar = one.download()
            one_variant = one.info()["variant_name"]
        ea, ev, ex = norm_rel(A, ar), norm_rel(V, vr), norm_rel(X, xr)
        worst = np.argsort(ea)[-4:][::-1].astype(np.int32)
        a64 = o.accel_f64(x0, worst, eps2)
        rows = []
        x64 = x0.astype(np.float64)
        for k, i in enumerate(worst):
            d = x64[:, :3] - x64[i, :3]
            r2 = (d * d).sum(1) + eps2
            terms = d * (x64[:, 3] / (r2 * np.sqrt(r2)))[:, None]
            cond = float(np.linalg.norm(terms, axis=1).sum() / max(np.linalg.norm(terms.sum(0)), 1e-300))
            rows.append({"body": int(i), "err_sharded_vs_1gpu": float(ea[i]), "cond": round(cond, 1),
                         "sharded_vs_fp64": float(norm_rel(A[i:i + 1], a64[k:k + 1])[0]),
                         "one_gpu_vs_fp64": float(norm_rel(ar[i:i + 1], a64[k:k + 1])[0]),
                         "ref_fp32_vs_fp64": float(norm_rel(o.accel_f32(x0, worst[k:k + 1], eps2), a64[k:k + 1])[0])})
        # the bulk must agree to fp32 rounding; the tail is bounded by what fp32 itself can deliver on those bodies:
        # neither engine may be further from fp64 than 4 x the other one or the reference's own fp32 sum
        tail_ok = all(r["sharded_vs_fp64"] <= max(1e-5, 4 * max(r["one_gpu_vs_fp64"], r["ref_fp32_vs_fp64"])) for r in rows)
        sa, sx = stats(ea), stats(ex)
        out = {"steps": 1, "one_gpu_kernel": one_variant, "a": sa, "v": stats(ev), "x": sx, "worst_bodies": rows,
               "tol": "a: median <= 1e-6, p99.9 <= 1e-5; x: p99.9 <= 1e-6; worst bodies: sharded_vs_fp64 <= max(1e-5, 4 x max(one_gpu_vs_fp64, ref_fp32_vs_fp64))",
               "ok": bool(len(A) == n and sa["median"] <= 1e-6 and sa["p99.9"] <= 1e-5 and sx["p99.9"] <= 1e-6 and tail_ok)}
    if world > 1:
        box = [out]
        dist.broadcast_object_list(box, 0)
        out = box[0]
    return out


def measure(nb, par, torch, dist, args, n, ic, K, W, rank, world, local, dev, flush, want_e2e, want_one_gpu, sampler=None):
    """one workload: warm-up, K timed steps, (optionally) the end-to-end loop, the parity block"""
    import numpy as np
    if ic == "plummer":
        x0, v0 = nb.ic_plummer(n, seed=12345)
        eps2, dt = 0.002, 0.01
    else:
        x0, v0 = nb.ic_uniform(n, seed=12345), None
        eps2, dt = 0.002, 0.1
    sim, runner, exchange = make_sim(nb, par, torch, dist, args, n, x0, v0, eps2, dt, rank, world, local, dev, args.exchange, args.barrier)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    timed = Timed(torch, runner, world, flush)
    if W > 0:
        timed.run(W)
    barrier()
    if world > 1:
        sim.shard_barrier_wait_ns(reset=True)
    launches0 = nb.launch_count()
    if sampler is not None:
        sampler.__enter__()
    total_ms, region_ms, per_step = timed.run(K)
    if sampler is not None:
        sampler.__exit__(None, None, None)
    barrier()
    launches = nb.launch_count() - launches0
    wait = sim.shard_barrier_wait_ns() if world > 1 else {"total_ns": 0, "max_ns": 0, "waits": 0}
    mine = {"rank": rank, "ms_per_step": round(total_ms / K, 4), "variant": sim.info()["variant_name"],
            "barrier_wait_ms_per_step": round(wait["total_ns"] / 1e6 / K, 4), "barrier_wait_ms_max": round(wait["max_ns"] / 1e6, 4)}
    per_rank = [mine]
    if world > 1:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)
    t = torch.tensor([total_ms, region_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, region_ms = (float(v) for v in t.tolist())
    res = {"n": n, "ic": ic, "steps": K, "ms_per_step": total_ms / K, "region_ms_per_step": region_ms / K,
           "value": n * float(n) * K / (total_ms * 1e-3) / 1e9, "launches": int(launches), "exchange": exchange,
           "barrier": getattr(runner, "barrier", None) if exchange == "p2p" else None, "per_rank": per_rank,
           "info": sim.info(), "eps2": eps2}
    try:   # clock64 / globaltimer inside the last timed kernel: the clock the SMs really delivered
        res["sm_mhz_in_kernel"] = round(sim.kernel_clock_mhz(), 1)
    except Exception:  # noqa: BLE001  (small-N family kernels do not record it)
        pass

    # ---- e2e: host buffers in, host buffers out, every step (pinned memory) -------------------------
    if want_e2e:
        hx = torch.from_numpy(x0).pin_memory()
        hv = torch.zeros(n, 4).pin_memory() if v0 is None else torch.from_numpy(v0).pin_memory()
        ox = torch.empty(sim.count, 4).pin_memory()
        ov = torch.empty(sim.count, 4).pin_memory()
        oa = torch.empty(sim.count, 4).pin_memory()

        def e2e_step():
            # upload rewrites both source buffers, which the peers' epilogues store into: nobody may still be stepping
            # (a cross-rank barrier on each side -- part of what this API costs at N > 1, so inside the timed region)
            if world > 1:
                runner.quiesce()
            nb._check(nb.lib.nbody_upload(sim._h, hx.data_ptr(), hv.data_ptr()))
            if world > 1:
                runner.quiesce()
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
        res["e2e"] = {"value": round(n * float(n) * K / float(te.item()) / 1e9, 3), "unit": UNIT,
                      "h2d_bytes_per_step": int(hx.numel() * 4 + hv.numel() * 4), "d2h_bytes_per_step": int(3 * ox.numel() * 4),
                      "api": "nbody_upload + step + nbody_download per step (host float4 buffers, pinned)"
                             + ("; a cross-rank barrier on each side of the upload" if world > 1 else "")}
        res["_e2e_out"] = (ox.numpy().copy(), oa.numpy().copy())

    # ---- parity block: checker legs, after every timed region ------------------------------------------
    if not args.no_parity:
        o = _oracle()
        par_blk = {}
        if want_e2e:   # the state the e2e loop produced: one step from x0
            _, oa_h = res.pop("_e2e_out")
            loc = np.unique(np.linspace(0, sim.count - 1, 32).astype(np.int64))
            idx = (sim.begin + loc).astype(np.int32)
            a64 = o.accel_f64(x0, idx, eps2)
            te = torch.tensor([float(norm_rel(oa_h[loc], a64).max())], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
            par_blk["e2e_err_a_vs_fp64"] = float(te.item())
        par_blk.update(check_last_step(o, nb, dist, torch, dev, sim, runner, world, n, eps2))
        par_blk["barrier_timeouts"] = barrier_health(torch, dist, dev, sim, world)
        if world > 1 and want_one_gpu:
            par_blk["err_vs_1gpu"] = versus_one_gpu(o, nb, dist, torch, dev, local, sim, runner, rank, world, n, x0, v0, eps2, dt, args.variant)
            par_blk["barrier_timeouts"] += barrier_health(torch, dist, dev, sim, world)
        par_blk["ok"] = bool(par_blk["ok"] and par_blk["barrier_timeouts"] == 0 and par_blk.get("err_vs_1gpu", {"ok": True})["ok"]
                             and par_blk.get("e2e_err_a_vs_fp64", 0.0) <= max(1e-5, 2 * par_blk["ref_fp32_err_vs_fp64"]))
        res["parity"] = par_blk
    res.pop("_e2e_out", None)
    runner.synchronize()
    barrier()
    sim.close()
    return res


def roofline_of(value, world, info, n, sm_mhz_sampled, sm_mhz_kernel, launches_per_step):
    pk = peaks()
    sms = info["sm_count"]
    peak_tflops = sms * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12
    achieved = FLOP_PER_INTERACTION * value * 1e9 / 1e12 / world          # per GPU, TFLOP/s
    traffic, src = ncu_traffic(n, info["variant_name"]) if world == 1 else (None, None)
    roof = {"bound": "fp32", "achieved": round(achieved, 3), "peak": round(peak_tflops, 3), "unit": "TFLOP/s",
            "frac": round(achieved / peak_tflops, 4), "traffic": traffic,
            "traffic_unit": f"DRAM bytes per launch (ncu --set full, profiles/{src}); algorithmic 112*N = 117 MB" if src else None,
            "peak_src": f"nominal FP32: {sms} SMs x 128 lanes x 2 x {pk['sm_max_mhz']:.0f} MHz ({pk['src']}); "
                        "MEASURED_PEAKS.json holds HBM and bf16-tensor peaks only, neither bounds this kernel",
            "kernel": info["variant_name"], "launches_per_step": launches_per_step}
    if sm_mhz_sampled:
        roof["frac_at_sampled_clock"] = round(achieved / (sms * 128 * 2 * sm_mhz_sampled * 1e6 / 1e12), 4)
    if sm_mhz_kernel:
        roof["frac_at_in_kernel_clock"] = round(achieved / (sms * 128 * 2 * sm_mhz_kernel * 1e6 / 1e12), 4)
    return roof, peak_tflops, achieved


# ---- our arm ----------------------------------------------------------------------------------------
def run_ours(args):
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
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    clk = ClockSampler(local)
    main = measure(nb, par, torch, dist, args, n, args.ic, K, W, rank, world, local, dev, flush,
                   want_e2e=not args.no_e2e, want_one_gpu=True, sampler=clk)
    clocks = clk.summary()
    if "sm_mhz_in_kernel" in main:
        clocks["sm_mhz_in_kernel"] = main["sm_mhz_in_kernel"]
    roof, peak_tflops, achieved = roofline_of(main["value"], world, main["info"], n, clocks["sm_mhz"], main.get("sm_mhz_in_kernel"),
                                              main["launches"] / K)

    # ---- series: BASELINE configs[4] -- N = 8 388 608 on 8 GPUs is the G = 8 point of the weak series N = G x 2^20 ----
    series = None
    if world > 1 and not args.no_series:
        series = []
        one = None
        if rank == 0:   # the 1-GPU rate of THIS process and box, the denominator of the efficiencies below
            with nb.Simulation(1 << 20, device=local) as s1:
                s1.upload(nb.ic_uniform(1 << 20, seed=12345)).step(1)
                s1.step(3)
                one = (1 << 40) * 3 / (s1.elapsed_ms() * 1e-3) / 1e9
        box = [one]
        dist.broadcast_object_list(box, 0)
        one = box[0]
        nw = world << 20
        r = measure(nb, par, torch, dist, args, nw, "uniform", 3, 1, rank, world, local, dev, flush, want_e2e=False, want_one_gpu=False)
        series.append({"name": f"weak N = G x 2^20 = {nw}" + (" = BASELINE configs[4] strong N = 8 388 608 on 8 GPUs" if world == 8 else ""),
                       "n_bodies": nw, "n_gpus": world, "steps": 3, "warmup": 1, "ms_per_step": round(r["ms_per_step"], 3),
                       "value": round(r["value"], 2), "unit": UNIT, "pct_fp32_peak_per_gpu": round(100 * FLOP_PER_INTERACTION * r["value"] * 1e9 / 1e12 / world / peak_tflops, 2),
                       "one_gpu_rate_same_process": round(one, 2), "efficiency_vs_one_gpu": round(r["value"] / (world * one), 4),
                       "kernel": r["info"]["variant_name"], "exchange": r["exchange"], "barrier": r["barrier"],
                       "per_rank": r["per_rank"], "parity": r.get("parity"),
                       "reference": "the reference's launcher refuses N > 2.16 M on a B200 (kernel.cu:637-641: \"Shared memory request too large.\")"})
        series.append({"name": f"strong N = {n} (the main line)", "n_bodies": n, "n_gpus": world, "value": round(main["value"], 2),
                       "one_gpu_rate_same_process": round(one, 2), "efficiency_vs_one_gpu": round(main["value"] / (world * one), 4)})

    # ---- mid sizes on one GPU (BASELINE configs[1], [2]) -----------------------------------------------
    small_n = None
    ref_gpu = None
    single_level = None
    if rank == 0 and not args.no_extras:
        if world == 1:
            small_n = []
            for nn, st in ((8192, 800), (16384, 400), (131072, 20)):   # 8 192 = the reference's shipped N_BODIES (constants.h:13), 16 384 = config 2, 131 072 = config 3
                with nb.Simulation(nn, device=local) as s:
                    s.upload(nb.ic_uniform(nn, seed=12345)).step(max(st // 10, 2))
                    s.step(st)
                    ms = s.elapsed_ms() / st
                    small_n.append({"n_bodies": nn, "steps": st, "us_per_step": round(ms * 1e3, 2), "value": round(nn * float(nn) / ms / 1e6, 1),
                                    "pct_fp32_peak": round(100 * FLOP_PER_INTERACTION * nn * float(nn) / ms / 1e9 / peak_tflops, 2),
                                    "kernel": s.info()["variant_name"], "split_tiles_finished_by": "no tile is split (row-split kernel: whole target tiles per CTA)" if s.info()["variant_name"].startswith("R_") else "all their CTAs (cooperative launch)" if s.info()["spread_fixup"] else "the last CTA to arrive",
                                    "timing": "nbody_step(ctx, steps), one CUDA-event pair around the loop"})
        if world == 1:   # the same kernel with ONE running sum per target (nbody_params.summation = NBODY_SUM_SINGLE): speed vs accuracy
            import numpy as np
            o = _oracle()
            xs = nb.ic_uniform(n, seed=12345)
            idx = np.random.default_rng(3).choice(n, 128, replace=False).astype(np.int32)
            a64 = o.accel_f64(xs, idx)
            single_level = {}
            for name, mode in (("two_level (default)", nb.NBODY_SUM_TWO_LEVEL), ("single_level", nb.NBODY_SUM_SINGLE)):
                with nb.Simulation(n, device=local, summation=mode) as s:
                    s.upload(xs).step(1)
                    a1 = s.download()[2]
                    s.step(3)
                    ms = s.elapsed_ms() / 3
                single_level[name] = {"value": round(n * float(n) / ms / 1e6, 1), "ms_per_step": round(ms, 3),
                                      "err_a_vs_fp64_128_targets": float(norm_rel(a1[idx], a64).max())}
            single_level["reference_fp32_err_vs_fp64"] = float(norm_rel(o.accel_f32(xs, idx), a64).max())
        try:   # B2: the reference's own kernel.cu rebuilt for sm_100, through its own simulate(); reported baseline
            ref_kernel_rate("libref_kernel_sm100.so", 16384, 1, 1)
            r = ref_kernel_rate("libref_kernel_sm100.so", 1 << 20, 1, 0)
            if r:
                ref_gpu = {"value": round(r[1], 2), "unit": UNIT, "ms_per_step": round(r[0], 2), "n_bodies": 1 << 20, "steps": 1, "n_gpus": 1,
                           "what": "reference TestProject/kernel.cu rebuilt for sm_100 (one-token fix at :99, nvcc -O3, THREADS_PER_BLOCK 32 as shipped) "
                                   "through its own simulate() (kernel.cu:628-645), cudaEvent-timed; oracle/_ref/libref_kernel_sm100.so"}
        except Exception as e:  # noqa: BLE001
            ref_gpu = {"error": str(e)[:200]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ncpu = 49152   # bounded sample sized for ~10-20 s of CPU work on 8 cores (2 timed steps)
        rate, kind, cores, dtc = cpu_reference_rate(ncpu, 2, 1)
        cpu = {"value": round(rate, 4), "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"CPU_compute (validation.cpp:28-52) x2 steps over {ncpu} bodies ({2 * ncpu * ncpu:.3g} interactions), {cores} threads, {dtc:.2f} s/step"}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    del flush
    torch.cuda.empty_cache()

    # ---- the same step driven by ONE process through the multi-device C ABI (after the ranks have left) ----
    capi = None
    if rank == 0 and world > 1 and not args.no_capi:
        env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS")}
        time.sleep(1.0)   # let the other ranks' contexts go away
        try:
            p = subprocess.run([sys.executable, os.path.abspath(__file__), "--driver", "capi", "--gpus", str(world), "--steps", str(K), "--warmup", str(W),
                                "--bodies", str(n), "--ic", args.ic], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
            lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
            capi = json.loads(lines[-1]) if lines else {"error": (p.stderr or "no output")[-300:]}
        except Exception as e:  # noqa: BLE001
            capi = {"error": str(e)[:300]}

    if rank == 0:
        info = main["info"]
        line = {"metric": METRIC, "value": round(main["value"], 3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": round(main["ms_per_step"], 4), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config_of(args, world), "clocks": clocks,
                "e2e": main.get("e2e"), "gpu_launches": int(main["launches"]), "roofline": roof, "cpu_baseline": cpu,
                "pct_fp32_peak": round(100 * achieved / peak_tflops, 2), "parity": main.get("parity"),
                "engine": {"variant": info["variant_name"], "grid_ctas": info["grid_ctas"], "ctas_per_sm": info["ctas_per_sm"],
                           "exchange": main["exchange"], "barrier": main["barrier"], "autotune": nb.autotune_report(),
                           "timing": "K steps enqueued back to back, one CUDA-event pair per step on the step's stream, no host sync inside the region; "
                                     "value = sum of the pairs, max over ranks" + (" (NCCL exchange: the whole region, steps overlap)" if main["exchange"] == "nccl" else ""),
                           "region_ms_per_step_incl_flush": round(main["region_ms_per_step"], 4), "per_rank": main["per_rank"]}}
        if series is not None:
            line["series"] = series
        if small_n is not None:
            line["small_n"] = small_n
        if single_level is not None:
            line["summation"] = single_level
        if ref_gpu is not None:
            line["ref_gpu_baseline"] = ref_gpu
        if capi is not None:
            line["capi"] = capi
        print(json.dumps(line), flush=True)
        if main.get("parity") is not None and not main["parity"]["ok"]:
            print("[bench] PARITY FAILED: " + json.dumps(main["parity"]), file=sys.stderr, flush=True)
            sys.exit(3)


def run_capi(args):
    """ONE process, ONE nbody_ctx spanning --gpus devices (nbody_params.n_devices): what the reference's caller loop
    (main.cpp:142-160) becomes on a multi-GPU box without new host code.  Prints one JSON line."""
    import numpy as np
    import torch
    import nbody_b200 as nb
    assert torch.cuda.is_available(), "bench.py needs a GPU (the N-body step has no CPU fallback)"
    G, n, K, W = args.gpus, args.n, args.steps, max(args.warmup, 1)
    assert torch.cuda.device_count() >= G, f"{G} devices asked for, {torch.cuda.device_count()} visible"
    x0 = nb.ic_uniform(n, seed=12345)
    ex = nb.NBODY_EXCHANGE_NCCL if args.exchange == "nccl" else nb.NBODY_EXCHANGE_P2P
    out = {"driver": "capi", "n_gpus": G, "n_bodies": n, "steps": K, "warmup": W}
    with nb.Simulation(n, devices=list(range(G)), exchange=ex, variant=args.variant) as sim:
        sim.upload(x0)
        sim.step(W)
        sim.shard_barrier_wait_ns(reset=True)
        sim.step(K)
        ms = sim.elapsed_ms() / K
        gi = sim.group_info() if G > 1 else {"n_devices": 1, "exchange": "none", "in_kernel_barrier": False}
        wait = sim.shard_barrier_wait_ns() if G > 1 else {"total_ns": 0}
        out.update({"ms_per_step": round(ms, 4), "value": round(n * float(n) / ms / 1e6, 2), "unit": UNIT, "kernel": sim.info()["variant_name"],
                    "exchange": gi["exchange"], "in_kernel_barrier": gi["in_kernel_barrier"], "barrier_wait_ms_per_step": round(wait["total_ns"] / 1e6 / K, 4),
                    "timing": "nbody_step(ctx, K): K steps enqueued back to back on every device, one CUDA-event pair per device, max over devices"})
        # e2e through the same context: host arrays in and out every step
        hx = torch.from_numpy(x0).pin_memory()
        hv = torch.zeros(n, 4).pin_memory()
        ox, ov, oa = (torch.empty(n, 4).pin_memory() for _ in range(3))
        t0 = None
        for i in range(2 + K):
            if i == 2:
                t0 = time.perf_counter()
            nb._check(nb.lib.nbody_upload(sim._h, hx.data_ptr(), hv.data_ptr()))
            nb._check(nb.lib.nbody_step(sim._h, 1))
            nb._check(nb.lib.nbody_download(sim._h, ox.data_ptr(), ov.data_ptr(), oa.data_ptr()))
        out["e2e"] = {"value": round(n * float(n) * K / (time.perf_counter() - t0) / 1e9, 2), "unit": UNIT,
                      "h2d_bytes_per_step": int(G * hx.numel() * 4 + hv.numel() * 4), "d2h_bytes_per_step": int(3 * n * 16)}
        if not args.no_parity:   # checker leg: the e2e state is one step from x0
            o = _oracle()
            idx = np.unique(np.linspace(0, n - 1, 64 * G).astype(np.int32))
            a64 = o.accel_f64(x0, idx, 0.002)
            err = float(norm_rel(oa.numpy()[idx], a64).max())
            ref = float(norm_rel(o.accel_f32(x0, idx, 0.002), a64).max())
            tmo = sim.shard_barrier_timeouts() if G > 1 else 0
            out["parity"] = {"sampled_targets": int(len(idx)), "err_a_vs_fp64": err, "ref_fp32_err_vs_fp64": ref, "barrier_timeouts": tmo,
                             "ok": bool(err <= max(1e-5, 2 * ref) and tmo == 0 and np.isfinite(oa.numpy()).all())}
    print(json.dumps(out), flush=True)
    if "parity" in out and not out["parity"]["ok"]:
        sys.exit(3)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.driver == "capi":
        run_capi(a)
    else:
        run_ours(a)
