import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib, json, torch, numpy as np
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
def bench(n, d, k, iters=10, spread=2.0):
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * spread
    lab = torch.randint(0, k, (n,), device="cuda", generator=g)
    X = prof[lab] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    C = prof + 0.1 * torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g)
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    scratch = N.new_scratch(d, k, "cuda"); xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    for _ in range(2): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(json.dumps(dict(n=n, d=d, k=k, spread=spread, ms=round(ms, 4), GBps=round(n * d * 8 / ms * 1e-6, 1), rows_per_s=round(n / ms * 1e3))), flush=True)
for args in [(5_000_000, 64, 64, 10, 2.0), (5_000_000, 64, 64, 10, 0.5), (5_000_000, 64, 32, 10, 2.0), (5_000_000, 48, 24, 10, 2.0), (20_000_000, 64, 64, 5, 2.0)]:
    bench(*args)
