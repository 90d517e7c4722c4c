"""Developer GPU check: Lloyd pass vs the CPU oracle on a few shapes, plus a quick timing."""
import importlib, sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
from oracle import hkm_oracle as O

def check(n, d, k, seed=0, weights=False):
    rng = np.random.default_rng(seed)
    Xh = rng.normal(size=(n, d)) + rng.integers(0, k, size=(n, 1)) * 2.0
    Ch = Xh[rng.choice(n, k, replace=False)].copy()
    wh = rng.uniform(0.5, 2.0, size=n) if weights else None
    X = torch.from_numpy(Xh).cuda(); C = torch.from_numpy(Ch).cuda()
    w = torch.from_numpy(wh).cuda() if weights else None
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda"); mind2 = torch.empty(n, dtype=torch.float64, device="cuda")
    scratch = N.new_scratch(d, k, "cuda")
    for rep in range(2):
        N.assign_reduce(X, C, partial, labels, mind2, weights=w, scratch=scratch)
    torch.cuda.synchronize()
    # STORE then REUSE must reproduce the same partials (cost to rounding)
    xs = torch.zeros(1, dtype=torch.float64, device="cuda"); p1 = torch.empty_like(partial); p2 = torch.empty_like(partial)
    N.assign_reduce(X, C, p1, weights=w, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    N.assign_reduce(X, C, p2, weights=w, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
    torch.cuda.synchronize()
    assert torch.equal(p1, partial), "STORE pass differs from PER_ROW pass"
    assert torch.equal(p1[:-1], p2[:-1]), "REUSE sums differ"
    rel = abs(p2[-1].item() - p1[-1].item()) / max(p1[-1].item(), 1e-300)
    assert rel < 1e-10, f"REUSE cost rel err {rel}"
    sums, wsum, cost, lab = O.lloyd_pass(Xh, Ch, w=wh, P=1, want_labels=True)
    _, dist = O.predict(Xh, Ch, want_dist=True)
    p = partial.cpu().numpy()
    agree = (labels.cpu().numpy() == lab).mean()
    es = np.abs(p[:k * d].reshape(k, d) - sums).max() / max(1.0, np.abs(sums).max())
    ew = np.abs(p[k * d:k * d + k] - wsum).max()
    ec = abs(p[-1] - cost) / max(cost, 1e-300)
    ed = np.abs(mind2.cpu().numpy() - dist).max()
    ok = agree == 1.0 and es < 1e-12 and ew < 1e-9 and ec < 1e-12 and ed < 1e-9
    print(f"n={n} d={d} k={k} w={weights}: agree={agree:.6f} sums_rel={es:.2e} wsum_abs={ew:.2e} cost_rel={ec:.2e} mind2_abs={ed:.2e} {'OK' if ok else 'FAIL'}", flush=True)
    return ok

def bench(n, d, k, iters=20):
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    X = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    C = X[:k].clone()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    scratch = N.new_scratch(d, k, "cuda")
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    for _ in range(3): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(json.dumps(dict(n=n, d=d, k=k, ms=round(ms, 4), GBps=round(n * d * 8 / ms * 1e-6, 1), rows_per_s=round(n / ms * 1e3))), flush=True)

if __name__ == "__main__":
    ok = True
    for (n, d, k) in [(1000, 8, 5), (128, 32, 8), (4096, 32, 8), (100003, 32, 8), (50000, 16, 16), (30000, 24, 7), (30000, 24, 32),
                      (20000, 5, 3), (20000, 7, 2), (20000, 64, 16), (20000, 33, 9), (77, 3, 2), (20000, 64, 64)]:
        ok &= check(n, d, k)
    ok &= check(50000, 32, 8, weights=True)
    ok &= check(50000, 16, 16, weights=True)
    print("ALL OK" if ok else "SOME FAILED")
    if "--bench" in sys.argv:
        bench(10_000_000, 32, 8); bench(20_000_000, 16, 16); bench(5_000_000, 24, 8); bench(5_000_000, 24, 16); bench(5_000_000, 24, 32); bench(5_000_000, 64, 16); bench(5_000_000, 64, 64); bench(20_000_000, 8, 5)
