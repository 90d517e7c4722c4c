#!/usr/bin/env python
"""Benchmark of the KMeans Lloyd hot path (BASELINE.json metric: rows*iter/sec, % of HBM roofline).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg4|cfg5] [--impl reference]

One "step" = one Lloyd iteration over the resident batch: hkm_assign_reduce (fused assign + per-cluster
reduce) -> all-reduce of k*(d+1)+1 doubles when N > 1 -> hkm_update_centers.  N > 1 runs under torchrun,
one rank per GPU, rows sharded (weak scaling: per-GPU rows fixed).  Rank 0 prints ONE JSON line.
`--impl reference` times the CPU restatement of the MLlib row solver (oracle/, all host threads) on a
bounded sample of the same workload; Spark itself cannot run here (no JVM / pyspark).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (rows per GPU, d, k, description)
    "cfg2": (10_000_000, 32, 8, "KMeans k=8, 10M rows x 32 features fp64, maxIter=20 (BASELINE configs[1])"),
    "cfg3": (100_000_000, 64, 64, "KMeans k=64, 100M rows x 64 features fp64 (BASELINE configs[2]; tcgen05 screening variant); rows are the 1-GPU total, divided by N"),
    "cfg4": (125_000_000, 16, 16, "KMeans k=16, 125M rows x 16 features fp64 per GPU (BASELINE configs[3] shard)"),
    "cfg5": (5_000_000, 24, 16, "KMeans k=16, 5M rows x 24 features fp64 (BASELINE configs[4], mid sweep)"),
}
SEED = 20250331


class NvmlSampler(threading.Thread):
    """Samples SM clock and clock-event (throttle) reasons through NVML every few ms while the
    timed region runs -- the region is too short for nvidia-smi's process start-up."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index: int, period_s: float = 0.004):
        super().__init__(daemon=True)
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
        self.period, self.sm, self.bits, self._stop_evt = period_s, [], 0, threading.Event()
        self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))

    def run(self):
        nv = self.nv
        getr = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop_evt.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.bits |= int(getr(self.h))
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        busy = [s for s in self.sm if s > 0.5 * self.mx] or self.sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": self.mx,
                "reasons": sorted(v for b, v in self.REASONS.items() if self.bits & b), "samples": len(self.sm), "source": "nvml"}


def make_sampler(index: int):
    try:
        return NvmlSampler(index)
    except Exception:
        return ClockSampler(index)


class ClockSampler(threading.Thread):
    """Fallback: samples nvidia-smi clocks / throttle reasons every 200 ms while the timed region runs."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s in sm if s > 0.5 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


def ncu_traffic(workload: str):
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get(workload)
        except Exception:
            return None
    return None


def host_threads() -> int:
    """Cores this process may use.  torchrun exports OMP_NUM_THREADS=1 to its workers, which would silently make the
    CPU arm single-threaded; the oracle takes an explicit thread count instead."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def run_reference(args):
    """CPU arm: the oracle's Lloyd pass (C, OpenMP over Spark-style partitions) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    from oracle import hkm_oracle as O
    n_full, d, k, desc = WORKLOADS[args.workload]
    n = min(n_full, args.ref_rows)
    rng = np.random.default_rng(SEED)
    prof = rng.normal(size=(k, d)) * 2.0
    X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d)) * 0.7
    C = X[:k].copy()
    threads = host_threads()
    P = max(8, 4 * threads)
    for _ in range(max(args.warmup, 1)):
        O.lloyd_pass(X, C, P=P, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sums, wsum, cost, _ = O.lloyd_pass(X, C, P=P, threads=threads)
        C, _ = O.update_centers(sums, wsum, C, 0.0)
    dt = time.perf_counter() - t0
    val = n * args.steps / dt
    line = {
        "impl": "reference", "metric": "kmeans_rows_iter_per_sec", "value": val, "unit": "rows*iter/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "rows_per_step": n, "d": d, "k": k},
        "cpu_baseline": {"value": val, "unit": "rows*iter/s", "cores": threads, "kind": "port",
                         "sample": f"{n} of {n_full} rows x {d}, {args.steps} Lloyd passes; CPU restatement of the MLlib row solver (not Spark: no JVM here)"},
        "e2e": {"value": val, "unit": "rows*iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


_JSON_FD = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else any library prints (NCCL's version banner, warnings)
    was sent to stderr by main(), so stdout holds exactly one line."""
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU")
    ap.add_argument("--ref-rows", type=int, default=10_000_000, help="rows per step of the CPU arm (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import hkm_b200
    from hkm_b200 import datagen
    N = hkm_b200._native
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n, d, k, desc = WORKLOADS[args.workload]
    if args.workload == "cfg3":
        n //= world                       # configs[2] is a fixed 100M-row table sharded over the GPUs (strong scaling)
    if args.rows:
        n = args.rows

    X = datagen.hospital_matrix_device(n, d, k, SEED + rank, dev)
    C0 = X[:k].clone()
    if world > 1:
        dist.broadcast(C0, 0)
    P = k * d + k + 1
    partial = torch.zeros(P, dtype=torch.float64, device=dev)
    scratch = N.new_scratch(d, k, dev)
    state = N.new_state(dev)
    xsum = torch.zeros(1, dtype=torch.float64, device=dev)
    C = C0.clone()

    xchg = hkm_b200._runtime.PeerExchange.get(dev) if world > 1 else None

    def step(mode, ev=None):
        if ev is not None:
            ev[0].record()
        # tol < 0: never "converged", every step does full work (rows*iter well defined)
        if world == 1:
            N.lloyd_step(X, C, partial, -1.0, state, scratch, xnorm2_sum=xsum, xnorm2_mode=mode)   # pass + centre update, one launch
            if ev is not None:
                ev[1].record()
            return
        if xchg is not None:     # ranks > 1: still one launch, the k*(d+1)+1 sums cross NVLink inside the kernel tail
            N.lloyd_step_peer(X, C, partial, -1.0, state, scratch, xchg, xnorm2_sum=xsum, xnorm2_mode=mode)
            if ev is not None:
                ev[1].record()
            return
        N.assign_reduce(X, C, partial, state=state, scratch=scratch, xnorm2_sum=xsum, xnorm2_mode=mode)
        if ev is not None:
            ev[1].record()
        dist.all_reduce(partial)
        N.update_centers(partial, C, -1.0, state)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step(1)                                  # STORE pass: sum |x|^2 (outside the timed region, like the first Lloyd iteration)
    for _ in range(args.warmup):
        step(2)
    # N == 1: a step IS one launch of the kernel, so the region's own events time it; N > 1: bracket each launch
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)] if world > 1 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = make_sampler(local_rank) if rank == 0 else None
    barrier()
    if sampler:
        sampler.start()
    e0.record()
    for i in range(args.steps):
        step(2, kev[i] if kev else None)
    e1.record()
    barrier()
    clocks = sampler.stop() if sampler else None
    ms_total = e0.elapsed_time(e1)
    kernel_ms = (sum(a.elapsed_time(b) for a, b in kev) / args.steps) if kev else ms_total / args.steps
    t = torch.tensor([ms_total, kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, kernel_ms = float(t[0]), float(t[1])
    st = N.read_state(state)
    assert st.iter == 1 + args.warmup + args.steps, "device state says a step was skipped"
    value = n * world * args.steps / (ms_total * 1e-3)

    # ---------------- e2e: the public estimator call on HOST (pinned) features
    e2e = None
    if not args.no_e2e:
        from hkm_b200.ml.clustering import KMeans
        e2e_iters = 20
        host = torch.empty((n, d), dtype=torch.float64, pin_memory=True)
        host.copy_(X)
        init = C0.cpu().numpy()
        km = KMeans(k=k, seed=42, maxIter=e2e_iters, tol=0.0).setInitialCenters(init)
        del X
        torch.cuda.empty_cache()
        fits = 1
        km.fit({"features": host})          # warm-up fit
        times = []
        for _ in range(5):                   # median of 5 single fits: the H2D copy shares the host's PCIe with other tenants
            barrier()
            t0 = time.perf_counter()
            model = km.fit({"features": host})   # H2D of the features, 20 Lloyd steps, summary pass, D2H of centres/cost/sizes
            _ = model.summary.trainingCost
            torch.cuda.synchronize()
            tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            times.append(float(tt[0]))
        dt = statistics.median(times)
        e2e = {"value": n * world * e2e_iters * fits / dt, "unit": "rows*iter/s",
               "h2d_bytes_per_step": int(n * d * 8 / e2e_iters), "d2h_bytes_per_step": int((k * d * 8 + 64 + n * 0) / e2e_iters),
               "call": f"KMeans(k={k}, maxIter={e2e_iters}, tol=0).fit(host table): one H2D of the pinned features per fit "
                       f"({n * d * 8} B), {e2e_iters} Lloyd steps + summary pass on device, D2H of centres/cost/sizes",
               "ms_per_fit": dt / fits * 1e3, "ms_per_fit_all": [round(t * 1e3, 2) for t in times]}
        X = None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_kind = measured_peak_gbs()
    alg_bytes = n * d * 8                       # 8*d bytes per row*iteration, one launch = n rows
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    line = {
        "metric": "kmeans_rows_iter_per_sec", "value": value, "unit": "rows*iter/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "strong" if args.workload == "cfg3" else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "rows_per_gpu": n, "d": d, "k": k,
                   "cache": f"inputs ({alg_bytes / 1e9:.2f} GB per GPU) are larger than the 126 MB L2; no flush needed",
                   "step": ("hkm_lloyd_step (assign + reduce + centre update in one launch, xnorm2 REUSE)" if world == 1 else ("hkm_lloyd_step_peer (one launch: pass + NVLink peer-memory sum of k*(d+1)+1 f64 + centre update)" if xchg is not None else "hkm_assign_reduce (xnorm2 REUSE) + NCCL allreduce(k*(d+1)+1 f64) + hkm_update_centers")) + ", convergence test off"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})", "kernel": "lloyd_tc_kernel (hkm_lloyd_tc.cuh)" if (k > 32 and d > 32) else ("lloyd_mma_kernel (hkm_lloyd_mma.cuh)" if k <= 16 else "lloyd_pass_kernel (hkm_lloyd.cuh)"),
                     "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes, "traffic": ncu_traffic(args.workload)},
        "gpu_launches": (1 if (world == 1 or xchg is not None) else 2) * args.steps,
        "clocks": clocks,
    }
    if e2e:
        line["e2e"] = e2e
    if not args.no_cpu_baseline and world == 1:
        from oracle import hkm_oracle as O
        rng = np.random.default_rng(SEED)
        ns = min(n, 10_000_000)
        prof = rng.normal(size=(k, d)) * 2.0
        Xh = prof[rng.integers(0, k, size=ns)] + rng.normal(size=(ns, d)) * 0.7
        Ch = Xh[:k].copy()
        threads = host_threads()
        O.lloyd_pass(Xh, Ch, P=4 * threads, threads=threads)
        reps, t0 = 0, time.perf_counter()
        while reps < 3 or (time.perf_counter() - t0 < 10 and reps < 200):
            O.lloyd_pass(Xh, Ch, P=4 * threads, threads=threads)
            reps += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": ns * reps / dt, "unit": "rows*iter/s", "cores": threads, "kind": "port",
                                "sample": f"{reps} Lloyd passes over {ns} x {d} rows (k={k}); C/OpenMP restatement of the MLlib row solver "
                                          "(Spark local[N] cannot run here: no JVM/pyspark)"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
