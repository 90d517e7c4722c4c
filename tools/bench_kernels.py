"""Per-kernel roofline numbers for every hot-path kernel besides the cfg2 Lloyd pass that bench.py reports:
algorithmic bytes / CUDA-event time, as a fraction of the measured HBM peak.  One JSON line each."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
N = hkm_b200._native
from hkm_b200.ml import _kmeans_init as KI
from hkm_b200 import _sparkrand as SR

PEAK = 6537.0
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def report(name, shape, bytes_alg, ms, note=""):
    gbs = bytes_alg / ms * 1e-6
    print(json.dumps({"kernel": name, "shape": shape, "ms": round(ms, 4), "algorithmic_GB": round(bytes_alg / 1e9, 4), "GBps": round(gbs, 1),
                      "frac_of_measured_hbm": round(gbs / PEAK, 3), "note": note}), flush=True)


def data(n, d, k, seed=1, spread=2.0):
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * spread
    lab = torch.randint(0, k, (n,), device="cuda", generator=g)
    X = prof[lab] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    return X, prof + 0.1 * torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g), lab.int()


def lloyd(n, d, k, tag):
    X, C, _ = data(n, d, k)
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    ms = timeit(lambda: N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2))
    report("lloyd_pass (hkm_assign_reduce, REUSE)", f"{n}x{d} k={k}", n * d * 8, ms, tag)
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    ms = timeit(lambda: N.assign_reduce(X, C, labels=labels, do_reduce=False))
    report("predict (hkm_assign_reduce, labels only)", f"{n}x{d} k={k}", n * (d * 8 + 4), ms, tag)


def silhouette(n, d, k):
    X, C, lab = data(n, d, k)
    stats = torch.zeros(k * d + 2 * k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
    work = torch.empty(k * d + 2 * k, dtype=torch.float64, device="cuda"); out2 = torch.zeros(2, dtype=torch.float64, device="cuda")
    ms1 = timeit(lambda: N.silhouette_stats(X, lab, k, stats, scratch))
    report("silhouette_stats", f"{n}x{d} k={k}", n * (d * 8 + 4), ms1)
    ms2 = timeit(lambda: N.silhouette_score(X, lab, k, stats, work, out2, scratch))
    report("silhouette_score", f"{n}x{d} k={k}", n * (d * 8 + 4), ms2)


def loader(n, d):
    g = torch.Generator(device="cuda"); g.manual_seed(2)
    cols = [torch.randint(0, 1000, (n,), device="cuda", generator=g, dtype=torch.int32) if i % 2 == 0 else
            torch.randn(n, dtype=torch.float64, device="cuda", generator=g) for i in range(d)]
    valids = [None] * d
    in_bytes = sum(c.element_size() for c in cols) * n
    shift = torch.randn(d, dtype=torch.float64, device="cuda"); scale = torch.rand(d, dtype=torch.float64, device="cuda") + 0.5
    ms = timeit(lambda: N.assemble(cols, valids, n, shift, scale), iters=5, warm=2)
    report("assemble (+fused standardise)", f"{n} rows, {d} cols (int32/f64 mix)", in_bytes + n * d * 8, ms, "includes torch.empty of the output")
    pivot = torch.stack([c[0].double() for c in cols])
    ms = timeit(lambda: N.column_moments(cols, valids, n, pivot), iters=5, warm=2)
    report("column_moments", f"{n} rows, {d} cols", in_bytes, ms)
    X = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    ms = timeit(lambda: N.moments(X, X[0].clone()), iters=5, warm=2)
    report("moments (row-major)", f"{n}x{d}", n * d * 8, ms)
    ms = timeit(lambda: N.scale_rows(X, shift, scale), iters=5, warm=2)
    report("scale_rows", f"{n}x{d}", 2 * n * d * 8, ms, "includes torch.empty_like of the output")


def kmpar(n, d, k):
    X, C, _ = data(n, d, k)
    costs = torch.full((n,), float("inf"), dtype=torch.float64, device="cuda"); s = torch.zeros(1, dtype=torch.float64, device="cuda")
    scratch = N.kmpar_new_scratch(d, "cuda")
    newc = C[: min(2 * k, 64)].contiguous()
    ms = timeit(lambda: N.kmpar_cost_update(X, newc, costs, s, scratch), iters=5, warm=2)
    report("kmpar_cost_update", f"{n}x{d}, {newc.shape[0]} new centres", n * (d * 8 + 16), ms)
    lay = KI.PartitionLayout.build([n], 0, 8)
    strip = N.kmpar_strip_rows(); lsz = lay.local_part_sizes
    part_rows = torch.tensor(SR.partition_bounds(lsz), dtype=torch.int64, device="cuda")
    strips = [(x + strip - 1) // strip for x in lsz]
    strip_start = torch.tensor(SR.partition_bounds(strips), dtype=torch.int64, device="cuda")
    idx = torch.empty(1 << 16, dtype=torch.int64, device="cuda"); cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    ms = timeit(lambda: N.kmpar_sample(costs, part_rows, strip_start, sum(strips), 0, 12345, 1, k, s, idx, cnt), iters=5, warm=2)
    report("kmpar_sample (XORShift jump-ahead)", f"{n} rows, 8 partitions", n * 8, ms)


if __name__ == "__main__":
    if "sil" in sys.argv:
        silhouette(5_000_000, 24, 8); silhouette(10_000_000, 32, 8); silhouette(5_000_000, 24, 16); silhouette(5_000_000, 24, 32); sys.exit(0)
    if "loader" in sys.argv:
        loader(10_000_000, 32); loader(10_000_000, 24); sys.exit(0)
    lloyd(10_000_000, 32, 8, "cfg2")
    lloyd(20_000_000, 16, 16, "cfg4 shard shape")
    lloyd(5_000_000, 24, 16, "cfg5 mid sweep")
    lloyd(5_000_000, 24, 32, "cfg5 top of sweep")
    lloyd(20_000_000, 64, 64, "cfg3 shape (tcgen05 variant)")
    silhouette(5_000_000, 24, 8)
    silhouette(5_000_000, 24, 32)
    loader(10_000_000, 32)
    kmpar(10_000_000, 32, 8)
