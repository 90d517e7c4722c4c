"""Lloyd pass (REUSE mode) over a grid of shapes: which variant runs, ms per pass, fraction of the measured HBM peak."""
import importlib, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
PEAK = 6537.0
rows = []
for d in (8, 16, 24, 32, 48, 64, 96):
    for k in (4, 8, 16, 24, 32, 48, 64):
        n = 2_400_000_000 // (d * 8)
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
        X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
        C = prof + 0.1
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        try:
            scratch = N.new_scratch(d, k, "cuda")
            for _ in range(2): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for _ in range(5): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            rows.append(dict(d=d, k=k, n=n, ms=round(ms, 4), GBps=round(n * d * 8 / ms * 1e-6, 1), frac=round(n * d * 8 / ms * 1e-6 / PEAK, 3)))
        except Exception as e:
            rows.append(dict(d=d, k=k, n=n, error=str(e)[-80:]))
        print(json.dumps(rows[-1]), flush=True)
        del X
