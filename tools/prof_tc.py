import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib, torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
n, d, k = 5_000_000, 64, 64
g = torch.Generator(device="cuda"); g.manual_seed(1)
prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2.0
X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
C = prof + 0.1 * torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g)
partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
scratch = N.new_scratch(d, k, "cuda"); xs = torch.zeros(1, dtype=torch.float64, device="cuda")
N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
for _ in range(3): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
torch.cuda.synchronize()
print("ok")
