#!/usr/bin/env python
"""Benchmark of the KMeans Lloyd hot path (BASELINE.json metric: rows*iter/sec, % of HBM roofline).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg4|cfg5] [--impl reference]

One "step" = one Lloyd iteration over the resident batch: one launch of the fused pass (assign + per-cluster
reduce + cross-GPU sum over NVLink peer memory when N > 1 + centre update).  N > 1 runs under torchrun, one
rank per GPU, rows sharded (weak scaling: per-GPU rows fixed).  Rank 0 prints ONE JSON line.  The headline
workload is BASELINE configs[1] (cfg2); the same line carries, under "extra", 20 timed steps each of the
configs[3] shard (cfg4) and of configs[2] (cfg3, its 100 M rows divided by N), at N > 1 a "parity_check" of the
in-kernel peer reduction against a rank-ordered sum of the per-rank partials, and two end-to-end numbers
through the public estimator (given initial centres, and the default k-means|| initialiser).
`--impl reference` times the CPU restatement of the MLlib row solver (oracle/, all host threads) on the SAME
host-built hospital table (same generator, seed and rows); Spark itself cannot run here (no JVM / pyspark).
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

    # One sample when the region starts, then one every 20 ms.  Measured at N = 2 (profiles/r02_sampler_skew.txt): with a 4 ms period
    # the sampled rank arrived 7 us later at every step's exchange (0.4361 ms per step against 0.4282 without a sampler); at N = 1
    # the period makes no difference.  HKM_BENCH_NVML_PERIOD / HKM_BENCH_SAMPLER_RANK are developer switches for that measurement.
    def __init__(self, index: int, period_s: float = float(os.environ.get("HKM_BENCH_NVML_PERIOD", "0.02"))):
        super().__init__(daemon=True)
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
        self.period, self.sm, self.bits, self._stop_evt, self._go = period_s, [], 0, threading.Event(), threading.Event()
        self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))

    def run(self):
        nv = self.nv
        getr = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        self._go.wait()        # the thread is started before the ranks' barrier (its start-up must not delay this rank's first step)
        self._stop_evt.wait(0.002)   # first sample two milliseconds into the region: the GPU is under load by then
        while not self._stop_evt.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.bits |= int(getr(self.h))
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def go(self):
        self._go.set()

    def stop(self):
        self._go.set()
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
        self.index, self.rows, self._stop_evt, self._go = index, [], threading.Event(), threading.Event()

    def go(self):
        self._go.set()

    def run(self):
        self._go.wait()
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
        self._go.set()
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


PARTITIONS = 8          # Spark partition layout both arms emulate


def config_for(args, n, d, k, desc):
    """The `config` object of the JSON line -- identical, key for key, in both arms."""
    return {"workload": args.workload, "description": desc, "rows_per_gpu": n, "d": d, "k": k,
            "table": "datagen.hospital_matrix(rows_per_gpu, d, profiles=k, seed=20250331 + rank): synthetic hospital table, StandardScaler defaults",
            "partitions": PARTITIONS, "init": "first k rows of rank 0's shard", "iterations_forced": True}


def host_table(n, d, k, rank):
    from hkm_b200 import datagen
    return datagen.hospital_matrix(n, d, profiles=k, seed=SEED + rank)


def cpu_lloyd_throughput(Xh, Ch, threads, min_seconds, max_reps):
    """rows*iter/s of the oracle's Lloyd pass (C, OpenMP over P Spark-style partitions) on all host threads."""
    from oracle import hkm_oracle as O
    O.lloyd_pass(Xh, Ch, P=PARTITIONS, threads=threads)
    C, reps, t0 = Ch.copy(), 0, time.perf_counter()
    while reps < 3 or (time.perf_counter() - t0 < min_seconds and reps < max_reps):
        sums, wsum, cost, _ = O.lloyd_pass(Xh, C, P=PARTITIONS, threads=threads)
        C, _ = O.update_centers(sums, wsum, C, 0.0)
        reps += 1
    dt = time.perf_counter() - t0
    return Xh.shape[0] * reps / dt, reps, dt


def run_reference(args):
    """CPU arm: the oracle's Lloyd pass (C, OpenMP over Spark-style partitions) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import hkm_b200  # noqa: F401
    n_full, d, k, desc = WORKLOADS[args.workload]
    n = min(n_full, args.ref_rows)
    X = host_table(n, d, k, 0)
    C = X[:k].copy()
    threads = host_threads()
    from oracle import hkm_oracle as O
    for _ in range(max(args.warmup, 1)):
        O.lloyd_pass(X, C, P=PARTITIONS, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sums, wsum, cost, _ = O.lloyd_pass(X, C, P=PARTITIONS, threads=threads)
        C, _ = O.update_centers(sums, wsum, C, 0.0)
    dt = time.perf_counter() - t0
    val = n * args.steps / dt
    line = {
        "impl": "reference", "metric": "kmeans_rows_iter_per_sec", "value": val, "unit": "rows*iter/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_for(args, n, d, k, desc),
        "cpu_baseline": {"value": val, "unit": "rows*iter/s", "cores": threads, "kind": "port",
                         "sample": f"{n} of {n_full} rows x {d}, {args.steps} Lloyd passes, P = {PARTITIONS} partitions; CPU restatement of the MLlib row solver "
                                   "(not Spark: no JVM here); the same host-built hospital table as the GPU arm's rank 0"},
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
    ap.add_argument("--no-extra", action="store_true", help="skip the cfg4 / cfg3 extras of the default (cfg2) run")
    ap.add_argument("--extra-steps", type=int, default=20)
    ap.add_argument("--extra-rows", type=int, default=0, help="cap the rows per GPU of the extra workloads (smoke runs)")
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
    xchg_all = hkm_b200._runtime.PeerExchange.get(dev) if world > 1 else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def limb_shape(d_, k_):
        """REUSE passes of this shape run the limb-reduction variant (the dispatch rule of hkm_lloyd_wl.cu)."""
        kp = ((k_ + 15) // 16) * 16
        return 8 < k_ <= 64 and d_ % 2 == 0 and d_ <= 64 and not (kp == 16 and d_ > 32)

    def kernel_name(d_, k_):
        if limb_shape(d_, k_):
            return "lloyd_wl_kernel (hkm_lloyd_wl.cuh)"
        if k_ <= 8 or (k_ <= 16 and d_ > 32):
            return "lloyd_mma_kernel (hkm_lloyd_mma.cuh)"
        if k_ <= 16 and 8 < d_ <= 24:
            return "lloyd_wg_kernel (hkm_lloyd_wg.cuh)"
        return "lloyd_tp_kernel (hkm_lloyd_tp.cuh)"

    def timed_loop(X, C0, steps, warmup, sample_clocks):
        """`steps` Lloyd iterations over the resident shard X, device-timed; returns (ms_total, kernel_ms, clocks, state, C, partial)."""
        k_, d_ = C0.shape
        nloc = X.shape[0]
        partial = torch.zeros(k_ * d_ + k_ + 1, dtype=torch.float64, device=dev)
        scratch = N.new_scratch(d_, k_, dev)
        state = N.new_state(dev)
        xsum = torch.zeros(1, dtype=torch.float64, device=dev)
        C = C0.clone()
        xchg = xchg_all if (xchg_all is not None and xchg_all.fits(k_, d_)) else None

        def step(mode, ev=None):
            if ev is not None:
                ev[0].record()
            # tol < 0: never "converged", every step does full work (rows*iter well defined)
            if world == 1:
                N.lloyd_step(X, C, partial, -1.0, state, scratch, xnorm2_sum=xsum, xnorm2_mode=mode)   # pass + centre update, one launch
            elif xchg is not None:    # ranks > 1: still one launch, the k*(d+1)+1 sums cross NVLink inside the kernel tail
                N.lloyd_step_peer(X, C, partial, -1.0, state, scratch, xchg, xnorm2_sum=xsum, xnorm2_mode=mode)
            else:
                N.assign_reduce(X, C, partial, state=state, scratch=scratch, xnorm2_sum=xsum, xnorm2_mode=mode)
                if ev is not None:
                    ev[1].record()
                dist.all_reduce(partial)
                N.update_centers(partial, C, -1.0, state)
                return
            if ev is not None:
                ev[1].record()

        step(1)                                  # STORE pass: sum |x|^2 (outside the timed region, like the first Lloyd iteration)
        for _ in range(warmup):
            step(2)
        torch.cuda.synchronize()
        st_warm = N.read_state(state)             # exchange timing so far (start-up skew of the first exchange included): subtracted below
        # N == 1: a step IS one launch of the kernel, so the region's own events time it; N > 1: bracket each launch
        kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)] if world > 1 else None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sampler = make_sampler(local_rank) if (rank == int(os.environ.get("HKM_BENCH_SAMPLER_RANK", "0")) and sample_clocks) else None
        if sampler:
            sampler.start()              # armed here, released after the barrier
        barrier()
        e0.record()
        if sampler:
            sampler.go()
        for i in range(steps):
            step(2, kev[i] if kev else None)
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms_total = e0.elapsed_time(e1)
        kernel_ms = (sum(a.elapsed_time(b) for a, b in kev) / steps) if kev else ms_total / steps
        t = torch.tensor([ms_total, kernel_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        st = N.read_state(state)
        assert st.iter == 1 + warmup + steps, "device state says a step was skipped"
        peer_t = None
        if xchg is not None:
            xchg.check()
            # per-rank exchange timing the kernel tail accumulated over the timed steps: store into the peers / wait for their flags
            lo = ((st.reserved & 0xffffffff) - (st_warm.reserved & 0xffffffff)) & 0xffffffff
            hi = (((st.reserved >> 32) & 0xffffffff) - ((st_warm.reserved >> 32) & 0xffffffff)) & 0xffffffff
            mine = torch.tensor([lo / steps / 1e3, hi / steps / 1e3], dtype=torch.float64, device=dev)
            allr = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(allr, mine)
            peer_t = {"store_us_per_step_by_rank": [round(float(a[0]), 2) for a in allr], "wait_us_per_step_by_rank": [round(float(a[1]), 2) for a in allr],
                      "note": "inside the last CTA of every rank's pass: time to store the k*(d+1)+1 sums into every peer, then time until all peers' flags arrived (arrival skew between ranks + NVLink flight)"}
        return float(t[0]), float(t[1]), clocks, (C, partial, scratch, state, xsum, xchg, peer_t)

    # ---------------- the table: built on the host by the same generator the CPU arm uses, then made resident
    # (cfg2, the headline).  The larger workloads do not fit host RAM and are generated in HBM by the device twin of the generator.
    on_host = n * d * 8 <= 4 << 30
    Xh = host_table(n, d, k, rank) if on_host else None
    X = torch.from_numpy(Xh).to(dev) if on_host else datagen.hospital_matrix_device(n, d, k, SEED + rank, dev)
    C0 = X[:k].clone()
    if world > 1:
        dist.broadcast(C0, 0)
    ms_total, kernel_ms, clocks, (C, partial, scratch, state, xsum, xchg, peer_t) = timed_loop(X, C0, args.steps, args.warmup, True)
    value = n * world * args.steps / (ms_total * 1e-3)

    # ---------------- N > 1: the in-kernel peer reduction against a rank-ordered sum of the per-rank partials
    parity = None
    if world > 1 and xchg is not None:
        Ca, Cb = C.clone(), C.clone()
        pa = torch.zeros_like(partial); pb = torch.zeros_like(partial)
        sa, sb = N.new_state(dev), N.new_state(dev)
        N.lloyd_step_peer(X, Ca, pa, -1.0, sa, scratch, xchg, xnorm2_sum=xsum, xnorm2_mode=2)
        N.assign_reduce(X, Cb, pb, state=sb, scratch=scratch, xnorm2_sum=xsum, xnorm2_mode=2)
        gathered = [torch.empty_like(pb) for _ in range(world)]
        dist.all_gather(gathered, pb)
        tot = gathered[0].clone()
        for g in gathered[1:]:
            tot = tot + g                                   # rank order, the order the kernel's tail uses
        N.update_centers(tot, Cb, -1.0, sb)
        torch.cuda.synchronize()
        xchg.check()
        same_sum = torch.tensor([1 if torch.equal(pa, tot) else 0, 1 if torch.equal(Ca, Cb) else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(same_sum, op=dist.ReduceOp.MIN)
        cmin, cmax = Ca.clone(), Ca.clone()
        dist.all_reduce(cmin, op=dist.ReduceOp.MIN); dist.all_reduce(cmax, op=dist.ReduceOp.MAX)
        parity = {"peer_vs_rank_ordered_sum_bitequal": bool(same_sum[0].item()), "centres_peer_vs_rank_ordered_bitequal": bool(same_sum[1].item()),
                  "centres_identical_on_all_ranks": bool(torch.equal(cmin, cmax)), "ranks": world,
                  "how": "one extra step from the same centres: hkm_lloyd_step_peer vs hkm_assign_reduce + all_gather + sum in rank order + hkm_update_centers"}
    del C, partial

    # ---------------- e2e: the public estimator call on HOST (pinned) features
    e2e = e2e_default = None
    if not args.no_e2e and on_host:
        from hkm_b200.ml.clustering import KMeans
        e2e_iters = 20
        host = torch.from_numpy(Xh).pin_memory()
        init = C0.cpu().numpy()
        del X
        torch.cuda.empty_cache()
        # the host -> device copy alone (same pinned buffer), for the breakdown
        hcopy = []
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            tmp = host.to(dev, non_blocking=True)
            torch.cuda.synchronize()
            hcopy.append(time.perf_counter() - t0)
            del tmp
        ms_h2d = statistics.median(hcopy) * 1e3

        def time_fits(km, reps):
            km.fit({"features": host})          # warm-up fit
            times = []
            for _ in range(reps):                # median of single fits: the H2D copy shares the host's PCIe with other tenants
                barrier()
                t0 = time.perf_counter()
                model = km.fit({"features": host})   # H2D of the features, init, Lloyd steps, summary pass, D2H of centres/cost/sizes
                _ = model.summary.trainingCost
                torch.cuda.synchronize()
                tt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
                if world > 1:
                    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                times.append(float(tt[0]))
            return statistics.median(times), times, model

        km = KMeans(k=k, seed=42, maxIter=e2e_iters, tol=0.0).setNumPartitions(PARTITIONS).setInitialCenters(init)
        dt, times, _ = time_fits(km, 5)
        e2e = {"value": n * world * e2e_iters / dt, "unit": "rows*iter/s",
               "h2d_bytes_per_step": int(n * d * 8 / e2e_iters), "d2h_bytes_per_step": int((k * d * 8 + 64) / e2e_iters),
               "call": f"KMeans(k={k}, maxIter={e2e_iters}, tol=0).setInitialCenters(first k rows).fit(host table): one H2D of the pinned features "
                       f"per fit ({n * d * 8} B), {e2e_iters} Lloyd steps + summary pass on device, D2H of centres/cost/sizes",
               "ms_per_fit": dt * 1e3, "ms_per_fit_all": [round(t * 1e3, 2) for t in times], "ms_h2d_alone": round(ms_h2d, 2),
               "h2d_note": "every rank copies its shard from host memory of the same NUMA node; the copy, not the kernels, sets the fit time"}
        kmd = KMeans(k=k, seed=42, maxIter=e2e_iters, tol=0.0).setNumPartitions(PARTITIONS)      # pyspark defaults: initMode k-means||, initSteps 2
        dtd, timesd, md = time_fits(kmd, 3)
        e2e_default = {"value": n * world * md.summary.numIter / dtd, "unit": "rows*iter/s", "h2d_bytes_per_step": int(n * d * 8 / max(md.summary.numIter, 1)),
                       "d2h_bytes_per_step": int((k * d * 8 + 64) / max(md.summary.numIter, 1)),
                       "call": f"KMeans(k={k}, maxIter={e2e_iters}, tol=0, seed=42).fit(host table) with the default k-means|| initialiser (initSteps 2, "
                               f"P = {PARTITIONS}): H2D + k-means|| passes + LocalKMeans + {md.summary.numIter} Lloyd steps + summary pass",
                       "ms_per_fit": dtd * 1e3, "ms_per_fit_all": [round(t * 1e3, 2) for t in timesd], "numIter": md.summary.numIter}
        del host
    else:
        del X
    torch.cuda.empty_cache()

    # ---------------- extra workloads in the same line: configs[3] shard and configs[2], generated on device (too large for host RAM)
    extra = {}
    if args.workload == "cfg2" and not args.no_extra:
        for name in ("cfg4", "cfg3"):
            en, ed, ek, edesc = WORKLOADS[name]
            if name == "cfg3":
                en //= world
            if args.extra_rows:
                en = min(en, args.extra_rows)
            try:
                Xe = datagen.hospital_matrix_device(en, ed, ek, SEED + rank, dev)
                Ce = Xe[:ek].clone()
                if world > 1:
                    dist.broadcast(Ce, 0)
                ems, ekms, eclk, _st = timed_loop(Xe, Ce, args.extra_steps, 3, True)
                epeer = _st[-1]
                del Xe, _st
                torch.cuda.empty_cache()
                ebytes = en * ed * 8
                extra[name] = {"description": edesc, "rows_per_gpu": en, "d": ed, "k": ek, "steps": args.extra_steps, "ms_per_step": ems / args.extra_steps,
                               "value": en * world * args.extra_steps / (ems * 1e-3), "unit": "rows*iter/s", "scaling": "strong" if name == "cfg3" else "weak",
                               "kernel": kernel_name(ed, ek), "kernel_ms": ekms, "achieved_gbs": ebytes / (ekms * 1e-3) / 1e9,
                               "frac": ebytes / (ekms * 1e-3) / 1e9 / measured_peak_gbs()[0], "clocks": eclk,
                               "table": "datagen.hospital_matrix_device (generated in HBM, seed 20250331 + rank)"}
                if epeer:
                    extra[name]["peer_exchange"] = epeer
            except Exception as ex:      # noqa: BLE001  (e.g. out of memory on a smaller GPU): report, do not lose the headline
                extra[name] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
                torch.cuda.empty_cache()

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
        "config": config_for(args, n, d, k, desc) if on_host else dict(config_for(args, n, d, k, desc), table="datagen.hospital_matrix_device (generated in HBM)"),
        "notes": {"cache": f"inputs ({alg_bytes / 1e9:.2f} GB per GPU) are larger than the 126 MB L2; no flush needed",
                  "step": ("hkm_lloyd_step (assign + reduce + centre update in one launch, xnorm2 REUSE)" if world == 1 else ("hkm_lloyd_step_peer (one launch: pass + NVLink peer-memory sum of k*(d+1)+1 f64 + centre update)" if xchg is not None else "hkm_assign_reduce (xnorm2 REUSE) + NCCL allreduce(k*(d+1)+1 f64) + hkm_update_centers")) + ", convergence test off"},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_kind})", "kernel": kernel_name(d, k),
                     "kernel_ms": kernel_ms, "algorithmic_bytes_per_launch": alg_bytes, "traffic": ncu_traffic(args.workload)},
        # per step: the fused Lloyd kernel (+ the update kernel on the NCCL path); shapes on the limb-reduction variant launch the
        # variant behind it as well, which returns at once when the limb pass has run
        "gpu_launches": ((1 if (world == 1 or xchg is not None) else 2) + (1 if limb_shape(d, k) else 0)) * args.steps,
        "clocks": clocks,
    }
    if e2e:
        line["e2e"] = e2e
    if e2e_default:
        line["e2e_default_init"] = e2e_default
    if extra:
        line["extra"] = extra
    if parity:
        line["parity_check"] = parity
    if peer_t:
        line["peer_exchange"] = peer_t
    if not args.no_cpu_baseline and world == 1 and on_host:
        ns = min(n, 10_000_000)
        threads = host_threads()
        v, reps, dt = cpu_lloyd_throughput(Xh[:ns], Xh[:k].copy(), threads, 10.0, 200)
        line["cpu_baseline"] = {"value": v, "unit": "rows*iter/s", "cores": threads, "kind": "port",
                                "sample": f"{reps} Lloyd passes over the first {ns} rows x {d} of the same host-built table (k={k}, P = {PARTITIONS}); C/OpenMP "
                                          "restatement of the MLlib row solver (Spark local[N] cannot run here: no JVM/pyspark)"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
