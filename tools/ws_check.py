"""Warp-specialised Lloyd variant (hkm_lloyd_ws.cuh): parity vs the oracle and timing against the other variants (run with
HKM_LLOYD_VARIANT=ws|gather|mma or unset)."""
import importlib, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
from oracle import hkm_oracle as O

def check(n, d, k, seed=0):
    rng = np.random.default_rng(seed)
    Xh = rng.normal(size=(n, d)) + rng.integers(0, k, size=(n, 1)) * 2.0
    Ch = Xh[rng.choice(n, k, replace=False)].copy()
    X = torch.from_numpy(Xh).cuda(); C = torch.from_numpy(Ch).cuda()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda"); mind2 = torch.empty(n, dtype=torch.float64, device="cuda")
    scratch = N.new_scratch(d, k, "cuda")
    for rep in range(2): N.assign_reduce(X, C, partial, labels, mind2, scratch=scratch)
    xs = torch.zeros(1, dtype=torch.float64, device="cuda"); p1 = torch.empty_like(partial); p2 = torch.empty_like(partial)
    N.assign_reduce(X, C, p1, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    N.assign_reduce(X, C, p2, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
    torch.cuda.synchronize()
    assert torch.equal(p1, partial) and torch.equal(p1[:-1], p2[:-1])
    sums, wsum, cost, lab = O.lloyd_pass(Xh, Ch, P=1, want_labels=True)
    p = partial.cpu().numpy()
    agree = (labels.cpu().numpy() == lab).mean()
    es = np.abs(p[:k * d].reshape(k, d) - sums).max() / max(1.0, np.abs(sums).max())
    ew = np.abs(p[k * d:k * d + k] - wsum).max(); ec = abs(p[-1] - cost) / cost; ec2 = abs(p2[-1].item() - cost) / cost
    ok = agree == 1.0 and es < 1e-12 and ew == 0 and ec < 1e-12 and ec2 < 1e-10
    print(f"n={n} d={d} k={k}: agree={agree:.6f} sums_rel={es:.2e} wsum_abs={ew:.2e} cost_rel={ec:.2e}/{ec2:.2e} {'OK' if ok else 'FAIL'}", flush=True)
    return ok

def bench(n, d, k, iters=10):
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
    X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    C = prof + 0.1
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    for _ in range(3): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(json.dumps(dict(variant=os.environ.get("HKM_LLOYD_VARIANT", "default"), n=n, d=d, k=k, ms=round(ms, 4), GBps=round(n * d * 8 / ms * 1e-6, 1))), flush=True)

if __name__ == "__main__":
    if "check" in sys.argv:
        ok = all([check(*s) for s in [(30000, 24, 32), (30011, 24, 17), (5000, 16, 16), (128, 32, 24), (100, 24, 9), (50000, 32, 20), (40000, 16, 40), (70000, 24, 12), (60000, 32, 64), (45000, 20, 50), (300000, 7, 33), (1000003, 24, 32)]])
        print("ALL OK" if ok else "FAILED"); sys.exit(0 if ok else 1)
    shapes = [(5_000_000, 24, 32), (5_000_000, 24, 17), (5_000_000, 24, 16)] if "short" in sys.argv else None
    if "wide" in sys.argv:
        shapes = [(5_000_000, 24, 32), (5_000_000, 24, 24), (5_000_000, 32, 32), (5_000_000, 32, 20), (5_000_000, 16, 48), (5_000_000, 16, 24),
                  (5_000_000, 24, 64), (5_000_000, 32, 64), (5_000_000, 8, 40), (5_000_000, 12, 64)]
    for s in shapes or [(5_000_000, 24, 32), (5_000_000, 24, 24), (5_000_000, 24, 17), (5_000_000, 24, 16), (5_000_000, 24, 12), (20_000_000, 16, 16), (5_000_000, 32, 20), (5_000_000, 16, 48), (5_000_000, 32, 32), (5_000_000, 24, 64), (5_000_000, 32, 64)]:
        bench(*s)
