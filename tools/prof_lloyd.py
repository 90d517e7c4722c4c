"""Minimal driver for ncu: a few Lloyd passes on one shape (default cfg2: 10M x 32, k=8)."""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
n, d, k = (int(a) for a in (sys.argv[1:4] if len(sys.argv) >= 4 else (10_000_000, 32, 8)))
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 5
g = torch.Generator(device="cuda"); g.manual_seed(1)
X = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
C = X[:k].clone()
partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
scratch = N.new_scratch(d, k, "cuda")
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
xs = torch.zeros(1, dtype=torch.float64, device="cuda")
mode = int(os.environ.get("HKM_XMODE", "2"))
for _ in range(2): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
torch.cuda.synchronize(); e0.record()
for _ in range(reps): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=mode)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
print(f"n={n} d={d} k={k} ms={ms:.4f} GB/s={n*d*8/ms*1e-6:.1f}")
