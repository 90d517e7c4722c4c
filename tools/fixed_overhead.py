"""Fixed cost of one Lloyd pass (set-up + last-CTA fold): time of a pass over a few rows, per kernel family."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
for n, d, k in [(1000, 16, 16), (1000, 32, 8), (1000, 24, 32), (1000, 64, 64), (125_000_000, 16, 16)]:
    X = torch.randn(n, d, dtype=torch.float64, device="cuda"); C = X[:k].clone()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    row = dict(n=n, d=d, k=k)
    for v, name in ((5, "round1"), (4, "tp"), (6, "wg"), (0, "default")):
        N.set_lloyd_variant(v)
        for _ in range(3):
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        it = 50 if n < 10**6 else 10
        for _ in range(it):
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
        e1.record(); torch.cuda.synchronize()
        row[name + "_us"] = round(e0.elapsed_time(e1) / it * 1000, 1)
    N.set_lloyd_variant(0)
    del X
    print(json.dumps(row), flush=True)
