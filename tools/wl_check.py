"""Limb-reduction Lloyd pass (hkm_lloyd_wl.cuh): parity against the oracle and the other variants, and timing.
python tools/wl_check.py check | bench | prof"""
import ctypes, importlib, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _mods():
    import numpy as np, torch
    N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
    from oracle import hkm_oracle as O
    return np, torch, N, O


def one(n, d, k, kind="blobs"):
    np, torch, N, O = _mods()
    if os.environ.get("WL_PW"):
        N.load().hkm_debug_wl_pw(ctypes.c_int(int(os.environ["WL_PW"])))
    rng = np.random.default_rng(n + 31 * d + 977 * k)
    prof = rng.normal(size=(k, d)) * (0.2 if kind == "close" else 1.5)     # "close": overlapping clusters, many rows re-scored exactly
    Xh = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    if kind == "offset":
        Xh += 300.0
        prof = prof + 300.0
    if kind == "scales":
        Xh *= np.logspace(-6, 6, d)[None, :]
        prof = prof * np.logspace(-6, 6, d)[None, :]
    if kind == "outlier":                      # one huge value in a column of small ones: the range pass must decline
        Xh[n // 2, 3] = 1e15
    Ch = prof + 0.05 * rng.normal(size=(k, d)) * (np.abs(prof).mean())
    X = torch.from_numpy(np.ascontiguousarray(Xh)).cuda(); C = torch.from_numpy(np.ascontiguousarray(Ch)).cuda()
    P = k * d + k + 1
    scratch = N.new_scratch(d, k, "cuda")
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    p1 = torch.empty(P, dtype=torch.float64, device="cuda"); p2 = torch.full((P,), -7.0, dtype=torch.float64, device="cuda"); p3 = torch.empty_like(p2)
    N.assign_reduce(X, C, p1, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)       # STORE (measures the column ranges)
    N.assign_reduce(X, C, p2, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)       # REUSE: the limb variant when the record admits the table
    N.load().hkm_debug_wl_off(ctypes.c_int(1))
    N.assign_reduce(X, C, p3, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)       # REUSE without it
    N.load().hkm_debug_wl_off(ctypes.c_int(0))
    torch.cuda.synchronize()
    tail = scratch.view(torch.uint8)
    p1, p2, p3 = p1.cpu().numpy(), p2.cpu().numpy(), p3.cpu().numpy()
    sums, wsum, cost, lab = O.lloyd_pass(Xh, Ch, P=1, want_labels=True)
    scale = max(1.0, np.abs(sums).max())
    es = float(np.abs(p2[:k * d].reshape(k, d) - sums).max() / scale)
    es_col = float((np.abs(p2[:k * d].reshape(k, d) - sums) / np.maximum(np.abs(sums).max(axis=0), 1e-300)[None, :]).max())
    e23 = float(np.abs(p2[:k * d] - p3[:k * d]).max() / scale)
    ew = float(np.abs(p2[k * d:k * d + k] - wsum).max())
    ec = abs(p2[-1] - cost) / max(cost, 1e-300)
    used = not np.array_equal(p2[:k * d], p3[:k * d])            # bit-different sums: the limb variant produced p2
    ok = es < 1e-11 and es_col < 1e-9 and ew == 0 and ec < 1e-9 and e23 < 1e-11
    print(json.dumps(dict(n=n, d=d, k=k, kind=kind, limb_used=bool(used), sums_rel=es, sums_rel_per_column=es_col, vs_other_variant=e23, wsum_abs=ew, cost_rel=ec, ok=bool(ok))), flush=True)
    return ok


def bench(shapes, iters=10):
    np, torch, N, O = _mods()
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    for n, d, k in shapes:
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
        X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
        C = prof + 0.1
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        row = dict(n=n, d=d, k=k)
        for off, name in ((1, "other"), (0, "limb")):
            N.load().hkm_debug_wl_off(ctypes.c_int(off))
            for _ in range(2):
                N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for _ in range(iters):
                N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / iters
            row[name + "_ms"] = round(ms, 4); row[name + "_frac"] = round(n * d * 8 / ms * 1e-6 / peak, 3)
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(iters):
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
        e1.record(); torch.cuda.synchronize()
        row["store_pass_ms"] = round(e0.elapsed_time(e1) / iters, 4)
        N.load().hkm_debug_wl_off(ctypes.c_int(0))
        del X
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    cmd = sys.argv[1] if len(sys.argv) > 1 else "check"
    if cmd == "one":
        sys.exit(0 if one(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5] if len(sys.argv) > 5 else "blobs") else 1)
    if cmd == "check":
        cases = [(5000, 16, 16, "blobs"), (128, 16, 16, "blobs"), (100, 16, 9, "blobs"), (1, 16, 12, "blobs"), (200000, 16, 16, "blobs"), (30011, 16, 33, "blobs"),
                 (40000, 16, 40, "blobs"), (45000, 16, 64, "blobs"), (50000, 32, 20, "blobs"), (60000, 32, 64, "blobs"), (129, 32, 17, "blobs"),
                 (30000, 24, 17, "blobs"), (30011, 24, 32, "blobs"), (9000, 12, 20, "blobs"), (33333, 2, 10, "blobs"), (70003, 8, 12, "blobs"), (3000, 30, 33, "blobs"), (4097, 22, 64, "blobs"),
                 (20000, 64, 64, "blobs"), (20000, 64, 16, "blobs"), (6000, 48, 24, "blobs"), (5000, 40, 40, "blobs"), (4000, 62, 64, "blobs"), (65, 64, 64, "blobs"),
                 (20011, 64, 64, "close"), (20000, 64, 33, "close"), (9000, 64, 17, "close"), (7000, 58, 12, "close"), (30000, 64, 48, "blobs"), (5000, 32, 32, "close"),
                 (70000, 16, 16, "offset"), (70000, 16, 24, "scales"), (50000, 16, 16, "outlier"), (3000000, 16, 16, "blobs")]
        bad = 0
        for n, d, k, kind in cases:
            try:
                r = subprocess.run([sys.executable, __file__, "one", str(n), str(d), str(k), kind], capture_output=True, text=True, timeout=180)
                out = (r.stdout.strip().splitlines() or ["(no output)"])[-1]
                if r.returncode != 0:
                    bad += 1
                    out += "  rc=%d %s" % (r.returncode, r.stderr.strip()[-300:].replace("\n", " | "))
            except subprocess.TimeoutExpired:
                bad += 1; out = json.dumps(dict(n=n, d=d, k=k, timeout=True))
            print(out, flush=True)
        print("ALL OK" if bad == 0 else f"FAILED {bad}")
        sys.exit(0 if bad == 0 else 1)
    if cmd == "soak":        # python tools/wl_check.py soak [cases] [seed]: random limb shapes against the oracle, in one process
        import random
        cases = int(sys.argv[2]) if len(sys.argv) > 2 else 150
        rnd = random.Random(int(sys.argv[3]) if len(sys.argv) > 3 else 1)
        bad = 0
        for it in range(cases):
            d = 2 * rnd.randint(1, 32); k = rnd.randint(9, 64)
            n = rnd.choice([1, 31, 127, 128, 129, 1000, rnd.randint(2, 70000), rnd.randint(2, 300000)])
            kind = rnd.choice(["blobs", "blobs", "close", "close", "offset", "scales"])
            try:
                ok = one(n, d, k, kind)
            except Exception as e:                      # noqa: BLE001
                ok = False; print("EXC", n, d, k, kind, repr(e)[:200], flush=True)
            bad += 0 if ok else 1
        print("SOAK OK" if bad == 0 else f"SOAK FAILED {bad}")
        sys.exit(0 if bad == 0 else 1)
    if cmd == "single":     # python tools/wl_check.py single n d k: STORE pass, then three REUSE passes (for ncu)
        np, torch, N, O = _mods()
        n, d, k = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
        g = torch.Generator(device="cuda"); g.manual_seed(1)
        prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
        X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
        C = prof + 0.1
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
        for _ in range(3):
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
        torch.cuda.synchronize()
        sys.exit(0)
    if cmd == "prof":
        np, torch, N, O = _mods()
        for n, d, k in ([(20_000_000, 64, 64)] if "64" in sys.argv else [(5_000_000, 24, 32), (5_000_000, 24, 16)] if "cfg5" in sys.argv else [(20_000_000, 16, 16), (10_000_000, 32, 16), (20_000_000, 16, 32)]):
            g = torch.Generator(device="cuda"); g.manual_seed(1)
            partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda"); scratch = N.new_scratch(d, k, "cuda")
            xs = torch.zeros(1, dtype=torch.float64, device="cuda")
            N.load().hkm_debug_wl_pw(ctypes.c_int(0 if "nopw" in sys.argv else 1))
            if "hospital" in sys.argv:      # bench.py's table and centres: first k rows, then 10 Lloyd steps
                import hkm_b200
                from hkm_b200 import datagen
                X = datagen.hospital_matrix_device(n, d, 8 if "cfg5" in sys.argv else k, 20250331, torch.device("cuda"))
                C = X[:k].clone(); st = N.new_state("cuda")
                for it in range(10):
                    N.lloyd_step(X, C, partial, -1.0, st, scratch, xnorm2_sum=xs, xnorm2_mode=1 if it == 0 else 2)
            else:
                prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
                X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
                C = prof + 0.1
            N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
            buf = torch.zeros(24 * 8 + 8 + 32, dtype=torch.int64, device="cuda")
            fn = N.load().hkm_debug_wl_profile; fn.argtypes = [ctypes.c_void_p]; fn.restype = None
            fn(ctypes.c_void_p(buf.data_ptr()))
            for _ in range(3):
                N.assign_reduce(X, C, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
            torch.cuda.synchronize(); fn(None)
            bx = buf.cpu().numpy(); b = bx[:192].reshape(24, 8); tiles = max(int(b[0, 6]), 1)
            print(json.dumps(dict(n=n, d=d, k=k, tiles_cta0=tiles, clk_per_tile=round(float(b[0, 0]) / tiles, 1))))
            for w in range(24):
                if b[w, 0]:
                    print(f"  warp {w:2d} total/tile {b[w, 0] / tiles:8.1f}  wait_tma {b[w, 1] / tiles:8.1f}  wait_mma {b[w, 2] / tiles:8.1f}  refine {b[w, 3] / tiles:7.1f}  refined rows/tile {b[w, 5] / tiles:6.3f}  convert {b[w, 4] / tiles:7.1f}  group_barrier {b[w, 7] / tiles:7.1f}  sweep+exchange {(bx[192 + w] / tiles if w < 8 else -1):7.1f}" + (f"  top {bx[200 + 4 * w] / tiles:7.1f}  issue+close(incl. its wait) {bx[201 + 4 * w] / tiles:7.1f}  tail(incl. refine) {bx[202 + 4 * w] / tiles:7.1f}" if w < 8 else ""))
            del X
        sys.exit(0)
    if cmd == "fbench":
        np, torch, N, O = _mods()
        for f in (0, 1):
            N.load().hkm_debug_wl_flags(ctypes.c_int(f)); print("flags", f, flush=True)
            bench([(20_000_000, 64, 64), (5_000_000, 64, 32), (10_000_000, 32, 16), (5_000_000, 48, 24)] if "wide" in sys.argv else
                  [(20_000_000, 16, 16), (20_000_000, 16, 32), (10_000_000, 32, 16), (10_000_000, 32, 32), (5_000_000, 16, 12)])
        sys.exit(0)
    if cmd == "bench64":
        bench([(20_000_000, 64, 64), (5_000_000, 64, 64), (5_000_000, 64, 16), (5_000_000, 64, 32), (5_000_000, 48, 24), (5_000_000, 48, 48), (5_000_000, 40, 40)])
        sys.exit(0)
    if cmd == "qbench":      # with and without the loading warp
        np, torch, N, O = _mods()
        for w in (0, 1):
            N.load().hkm_debug_wl_pw(ctypes.c_int(w)); print("loading warp", w, flush=True)
            bench([(20_000_000, 64, 64), (5_000_000, 64, 32), (5_000_000, 50, 20), (10_000_000, 32, 16), (5_000_000, 48, 48)])
        sys.exit(0)
    if cmd == "hbench":      # rows of 16 ns + 8 features: half slab against padding to whole slabs
        np, torch, N, O = _mods()
        for f in (0, 1):
            N.load().hkm_debug_wl_nohalf(ctypes.c_int(f)); print("nohalf", f, flush=True)
            bench([(5_000_000, 24, 12), (5_000_000, 24, 16), (5_000_000, 24, 32), (5_000_000, 24, 64), (5_000_000, 20, 24), (5_000_000, 40, 40), (20_000_000, 24, 16)])
        sys.exit(0)
    if cmd == "gbench":
        np, torch, N, O = _mods()
        for g in (6, 5, 4):
            N.load().hkm_debug_wl_groups(ctypes.c_int(g)); print("groups", g, flush=True)
            bench([(20_000_000, 16, 16), (5_000_000, 16, 12)])
        sys.exit(0)
    if cmd == "bench":
        bench([(20_000_000, 16, 16), (20_000_000, 16, 32), (20_000_000, 16, 64), (10_000_000, 32, 16), (10_000_000, 32, 32), (10_000_000, 32, 64), (5_000_000, 16, 9), (5_000_000, 16, 48),
               (5_000_000, 24, 12), (5_000_000, 24, 16), (5_000_000, 24, 32), (5_000_000, 24, 64), (10_000_000, 8, 16), (5_000_000, 12, 64), (5_000_000, 20, 24)])
