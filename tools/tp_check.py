"""Pipelined tcgen05 Lloyd variant (hkm_lloyd_tp.cuh): parity against the oracle and the round-1 kernels, and timing
of both families on the same shapes.

  python tools/tp_check.py check            # every shape in its own process (a protocol bug traps after ~3 s, it cannot hang the call)
  python tools/tp_check.py bench [wide]     # ms per pass and fraction of the HBM peak, variant 4 (tp) next to variant 5 (round 1)
  python tools/tp_check.py one n d k        # a single parity case (what `check` spawns)
"""
import importlib, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHECK_SHAPES = [(5000, 16, 16), (128, 16, 16), (100, 24, 9), (30011, 24, 17), (50000, 16, 16), (30000, 24, 32), (128, 32, 24), (50000, 32, 20),
                (40000, 16, 40), (70000, 24, 12), (60000, 32, 64), (45000, 20, 50), (300000, 7, 33), (20000, 64, 16), (20000, 64, 64),
                (6000, 48, 24), (5000, 40, 40), (4000, 62, 64), (9000, 12, 20), (20000, 33, 9), (3000, 31, 33), (70003, 5, 12), (1000003, 24, 32),
                (200000, 64, 64), (200000, 16, 16), (33333, 2, 10), (2000, 1, 9), (65, 64, 64)]
BENCH_SHAPES = [(20_000_000, 16, 16), (5_000_000, 24, 16), (20_000_000, 64, 64), (5_000_000, 24, 32), (5_000_000, 32, 64), (5_000_000, 24, 64),
                (10_000_000, 32, 16), (10_000_000, 8, 16)]
WIDE = [(5_000_000, 24, 12), (5_000_000, 24, 24), (5_000_000, 32, 32), (5_000_000, 32, 20), (5_000_000, 16, 48), (5_000_000, 16, 24), (5_000_000, 8, 40),
        (5_000_000, 12, 64), (5_000_000, 48, 24), (5_000_000, 64, 16), (5_000_000, 64, 32), (5_000_000, 48, 48), (5_000_000, 16, 9), (5_000_000, 32, 9)]


def _mods():
    import numpy as np, torch
    N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
    from oracle import hkm_oracle as O
    return np, torch, N, O


def one(n, d, k, seed=0, tpv=4):
    np, torch, N, O = _mods()
    if os.environ.get("HKM_WG_GROUPS"):
        import ctypes
        N.load().hkm_debug_wg_groups(ctypes.c_int(int(os.environ["HKM_WG_GROUPS"])))
    rng = np.random.default_rng(seed + n + 31 * d + 977 * k)
    prof = rng.normal(size=(k, d)) * 1.5
    Xh = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    Ch = prof + 0.05 * rng.normal(size=(k, d))
    X = torch.from_numpy(Xh).cuda(); C = torch.from_numpy(Ch).cuda()
    P = k * d + k + 1
    scratch = N.new_scratch(d, k, "cuda")
    res = {}
    for variant in (5, tpv):
        N.set_lloyd_variant(variant)
        partial = torch.full((P,), -7.0, dtype=torch.float64, device="cuda")
        labels = torch.full((n,), -9, dtype=torch.int32, device="cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        p1 = torch.empty_like(partial); p2 = torch.empty_like(partial); lab2 = torch.full((n,), -9, dtype=torch.int32, device="cuda")
        N.assign_reduce(X, C, partial, labels, scratch=scratch)                                   # labels + sums, per-row norms in-pass
        N.assign_reduce(X, C, p1, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)                  # STORE
        N.assign_reduce(X, C, p2, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)                  # REUSE
        N.assign_reduce(X, C, None, lab2, do_reduce=False)                                        # predict
        torch.cuda.synchronize()
        res[4 if variant == tpv else 5] = (partial.cpu().numpy(), labels.cpu().numpy(), p1.cpu().numpy(), p2.cpu().numpy(), lab2.cpu().numpy(), xs.item())
    N.set_lloyd_variant(0)
    sums, wsum, cost, lab = O.lloyd_pass(Xh, Ch, P=1, want_labels=True)
    p, labels, p1, p2, lab2, xs = res[4]
    po = res[5][0]
    agree = float((labels == lab).mean()); agree_old = float((labels == res[5][1]).mean())
    es = float(np.abs(p[:k * d].reshape(k, d) - sums).max() / max(1.0, np.abs(sums).max()))
    ew = float(np.abs(p[k * d:k * d + k] - wsum).max())
    ec = abs(p[-1] - cost) / max(cost, 1e-300); ec2 = abs(p2[-1] - cost) / max(cost, 1e-300)
    same_modes = bool(np.array_equal(p, p1) and np.array_equal(p[:-1], p2[:-1]) and np.array_equal(labels, lab2))
    xs_ok = abs(xs - float((Xh * Xh).sum())) <= 1e-12 * float((Xh * Xh).sum())
    ok = agree == 1.0 and es < 1e-11 and ew == 0 and ec < 1e-9 and ec2 < 1e-9 and same_modes and xs_ok
    print(json.dumps(dict(n=n, d=d, k=k, v=tpv, agree=agree, agree_round1=agree_old, sums_rel=es, wsum_abs=ew, cost_rel=[ec, ec2], same_modes=same_modes,
                          xs_ok=bool(xs_ok), cost_vs_round1=abs(p[-1] - po[-1]) / max(po[-1], 1e-300), ok=bool(ok))), flush=True)
    return ok


def bench(shapes, iters=10, extra=()):
    np, torch, N, O = _mods()
    peak = 6554.6
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    for n, d, k in shapes:
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
        X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
        C = prof + 0.1
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        row = dict(n=n, d=d, k=k)
        for variant, name in ((5, "round1"), (4, "tp")) + tuple(extra):
            N.set_lloyd_variant(variant)
            try:
                for _ in range(3):
                    N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize(); e0.record()
                for _ in range(iters):
                    N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
                e1.record(); torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / iters
                row[name + "_ms"] = round(ms, 4); row[name + "_frac"] = round(n * d * 8 / ms * 1e-6 / peak, 3)
            except Exception as ex:      # noqa: BLE001
                row[name + "_err"] = str(ex)[:120]
        N.set_lloyd_variant(0)
        del X
        print(json.dumps(row), flush=True)


ROLE = ["screen"] * 8 + ["reduce"] * 4 + ["convert"] * 4 + ["producer", "mma"]


def prof(shapes, iters=3, hospital=False, variant=4, hook="hkm_debug_tp_profile", nw=18):
    """Per-warp cycle breakdown of CTA 0: where each role waits (site 0/1), refine cycles, listed rows."""
    import ctypes
    np, torch, N, O = _mods()
    N.set_lloyd_variant(variant)
    for n, d, k in shapes:
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        if hospital:        # bench.py's table and centres: first k rows, then 10 Lloyd steps
            import hkm_b200
            from hkm_b200 import datagen
            X = datagen.hospital_matrix_device(n, d, k, 20250331, torch.device("cuda"))
            C = X[:k].clone()
            pp = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); sc = N.new_scratch(d, k, "cuda"); st = N.new_state("cuda")
            for _ in range(10):
                N.lloyd_step(X, C, pp, -1.0, st, sc)
        else:
            prof_c = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
            X = prof_c[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
            C = prof_c + 0.1
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        buf = torch.zeros(nw * 8, dtype=torch.int64, device="cuda")
        N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
        fn = getattr(N.load(), hook); fn.argtypes = [ctypes.c_void_p]; fn.restype = None
        fn(ctypes.c_void_p(buf.data_ptr()))
        for _ in range(iters):
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
        torch.cuda.synchronize()
        fn(None)
        b = buf.cpu().numpy().reshape(nw, 8)
        tiles = max(int(b[0, 6]), 1)
        print(json.dumps(dict(n=n, d=d, k=k, tiles_cta0=tiles, clk_per_tile=round(float(b[0, 0]) / tiles, 1))))
        for w in range(nw):
            if b[w, 0] == 0:
                continue
            print(f"  warp {w:2d} total/tile {b[w, 0] / tiles:8.1f}  wait0 {b[w, 1] / tiles:8.1f}  wait1 {b[w, 2] / tiles:8.1f}  refine {b[w, 3] / tiles:7.1f}  refined rows/tile {b[w, 5] / tiles:6.2f}")
        sys.stdout.flush()
        del X
    N.set_lloyd_variant(0)


if __name__ == "__main__":
    cmd = sys.argv[1] if len(sys.argv) > 1 else "check"
    if cmd == "single":     # python tools/tp_check.py single n d k [variant]: three passes, for ncu
        np, torch, N, O = _mods()
        n, d, k = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
        N.set_lloyd_variant(int(sys.argv[5]) if len(sys.argv) > 5 else 4)
        if os.environ.get("HKM_WG_GROUPS"):
            import ctypes
            N.load().hkm_debug_wg_groups(ctypes.c_int(int(os.environ["HKM_WG_GROUPS"])))
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        pc = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
        X = pc[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
        C = pc + 0.1
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
        for _ in range(2):
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
        torch.cuda.synchronize()
        sys.exit(0)
    if cmd == "prof":
        prof([(20_000_000, 16, 16), (5_000_000, 24, 32), (20_000_000, 64, 64), (5_000_000, 24, 64)] if "hospital" not in sys.argv else
             [(20_000_000, 16, 16), (20_000_000, 64, 64), (5_000_000, 24, 16)], hospital="hospital" in sys.argv)
        sys.exit(0)
    if cmd == "one":
        sys.exit(0 if one(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), tpv=int(sys.argv[5]) if len(sys.argv) > 5 else 4) else 1)
    if cmd == "check":
        bad = 0
        # default scheme for every shape, then the other reduce schemes pinned (42 register-owned, 40/41 private sets) on a subset
        cases = [(n, d, k, 4) for n, d, k in CHECK_SHAPES]
        cases += [(n, d, k, 42) for n, d, k in CHECK_SHAPES if d % 2 == 0 and (k <= 16 or d <= 32)]
        cases += [(n, d, k, 40) for n, d, k in CHECK_SHAPES if d <= 32 and k <= 16]
        for n, d, k, v in cases:
            try:
                r = subprocess.run([sys.executable, __file__, "one", str(n), str(d), str(k), str(v)], capture_output=True, text=True, timeout=180)
                out = (r.stdout.strip().splitlines() or ["(no output)"])[-1]
                if r.returncode != 0:
                    bad += 1
                    out += "  rc=%d %s" % (r.returncode, r.stderr.strip()[-400:].replace("\n", " | "))
                print(out, flush=True)
            except subprocess.TimeoutExpired:
                bad += 1
                print(json.dumps(dict(n=n, d=d, k=k, v=v, timeout=True)), flush=True)
        print("ALL OK" if bad == 0 else f"FAILED {bad}")
        sys.exit(0 if bad == 0 else 1)
    if cmd == "wgcheck":    # the warp-group variant (6) on every shape it takes
        bad = 0
        shapes = [(n, d, k) for n, d, k in CHECK_SHAPES if k <= 32 and d <= 32 and k > 8] + [(1, 16, 16), (129, 13, 9), (4097, 8, 31), (555, 32, 32), (100001, 17, 16), (3000000, 16, 16)]
        for n, d, k in shapes:
            for g in ([0] if n != 50000 else [0, 2, 3, 5]):
                env = dict(os.environ, HKM_WG_GROUPS=str(g))
                r = subprocess.run([sys.executable, __file__, "one", str(n), str(d), str(k), "6"], capture_output=True, text=True, timeout=180, env=env)
                out = (r.stdout.strip().splitlines() or ["(no output)"])[-1]
                if r.returncode != 0:
                    bad += 1
                    out += "  rc=%d %s" % (r.returncode, r.stderr.strip()[-400:].replace("\n", " | "))
                print(("g=%d " % g) + out, flush=True)
        print("ALL OK" if bad == 0 else f"FAILED {bad}")
        sys.exit(0 if bad == 0 else 1)
    if cmd == "wgprof":
        import ctypes
        np, torch, N, O = _mods()
        for g in [int(a) for a in sys.argv[2:] if a.isdigit()] or [0]:
            N.load().hkm_debug_wg_groups(ctypes.c_int(g))
            print("groups", g, flush=True)
            prof([(20_000_000, 16, 16), (10_000_000, 32, 16), (5_000_000, 24, 32)], hospital="hospital" in sys.argv, variant=6, hook="hkm_debug_wg_profile", nw=24)
        sys.exit(0)
    if cmd == "wgbench":
        shapes = [(20_000_000, 16, 16), (5_000_000, 24, 16), (10_000_000, 32, 16), (10_000_000, 8, 16), (5_000_000, 24, 32), (5_000_000, 32, 32), (5_000_000, 16, 24),
                  (5_000_000, 24, 12), (5_000_000, 16, 9), (5_000_000, 32, 20)]
        import ctypes
        np, torch, N, O = _mods()
        if "short" in sys.argv:
            shapes = shapes[:3]
        for g in [int(a) for a in sys.argv[2:] if a.isdigit()] or [0]:
            N.load().hkm_debug_wg_groups(ctypes.c_int(g))
            print("groups", g, flush=True)
            bench(shapes, extra=[(6, "wg")])
        sys.exit(0)
    if cmd == "bench":
        extra = []
        if "modes" in sys.argv:
            extra = [(42, "tp_own"), (40, "tp_priv")]
        if "budgets" in sys.argv:
            extra = [(50, "w0"), (51, "w1"), (53, "w3"), (54, "w4_2cta")]
        bench(BENCH_SHAPES + (WIDE if "wide" in sys.argv else []), extra=extra)
