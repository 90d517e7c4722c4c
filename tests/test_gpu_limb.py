"""GPU parity of the limb-reduction Lloyd pass (hkm_lloyd_wl.cuh): REUSE passes whose per-cluster sums run on the integer tensor
cores in fixed point.  Against the CPU oracle (labels exact, counts exact, sums / cost far inside north_star's 1e-6), against the
variant it replaces, and the cases in which it must decline (columns with an outlier, non-finite data, a scratch without the
column range record)."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _three(N, Xh, Ch, scratch=None, want_labels=True):
    """STORE pass, REUSE pass (limb variant where the record admits the table), REUSE pass with the variant switched off."""
    X = torch.from_numpy(np.ascontiguousarray(Xh)).cuda(); C = torch.from_numpy(np.ascontiguousarray(Ch)).cuda()
    n, d = Xh.shape; k = Ch.shape[0]
    P = k * d + k + 1
    scratch = N.new_scratch(d, k, "cuda") if scratch is None else scratch
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    outs = []
    lib = N.load()
    for mode, off in ((1, 0), (2, 0), (2, 1)):
        lib.hkm_debug_wl_off(ctypes.c_int(off))
        partial = torch.full((P,), -7.0, dtype=torch.float64, device="cuda")
        labels = torch.full((n,), -9, dtype=torch.int32, device="cuda") if want_labels else None
        try:
            N.assign_reduce(X, C, partial, labels, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=mode)
            torch.cuda.synchronize()
        finally:
            lib.hkm_debug_wl_off(ctypes.c_int(0))
        p = partial.cpu().numpy()
        outs.append(dict(sums=p[:k * d].reshape(k, d), wsum=p[k * d:k * d + k], cost=p[-1], labels=None if labels is None else labels.cpu().numpy()))
    return outs


def _blobs(n, d, k, seed, spread=1.5):
    rng = np.random.default_rng(seed)
    prof = rng.normal(size=(k, d)) * spread
    X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    C = prof + 0.05 * rng.normal(size=(k, d))
    return X, C


SHAPES = [(5000, 16, 16), (128, 16, 16), (127, 16, 9), (1, 16, 12), (200_000, 16, 16), (30011, 16, 33), (40000, 16, 40), (45000, 16, 64),
          (50000, 32, 20), (60000, 32, 64), (129, 32, 17), (30011, 24, 32), (9000, 12, 20), (33333, 2, 10), (70003, 8, 12), (3000, 30, 33),
          (4097, 22, 64), (20000, 64, 64), (6000, 48, 24), (5000, 40, 40), (4000, 62, 64), (65, 64, 64),
          # rows of 16 ns + 1..8 features: ns slabs of 16 and a half slab of 8 (64-byte rows, M = 64 limb product)
          (30000, 24, 12), (20000, 18, 9), (129, 24, 16), (50000, 36, 48), (200_000, 24, 32), (70001, 20, 64), (33000, 34, 20),
          # shapes served with a loading warp (64-feature rows; d = 32 with k <= 16)
          (70001, 32, 16), (100, 32, 9), (9000, 64, 17), (20011, 64, 48), (300_000, 64, 64), (40000, 50, 20)]


@pytest.mark.parametrize("n,d,k", SHAPES)
def test_limb_pass_matches_oracle(hkm, oracle, n, d, k):
    X, C = _blobs(n, d, k, n + 31 * d + 977 * k)
    store, limb, other = _three(hkm._native, X, C)
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    assert np.array_equal(limb["labels"], lab) and np.array_equal(limb["wsum"], wsum)
    scale = max(1.0, np.abs(sums).max())
    assert np.abs(limb["sums"] - sums).max() <= 1e-11 * scale
    # per column too: the fixed-point quantum is relative to the column's largest magnitude
    col = np.maximum(np.abs(X).max(axis=0), 1e-300) * max(n, 1)
    assert (np.abs(limb["sums"] - sums) / col[None, :]).max() <= 1e-13
    assert limb["cost"] == pytest.approx(cost, rel=1e-9, abs=1e-9 * float((X * X).sum() + 1))
    # the variant really ran (its sums differ in the last bits from the fp64 accumulation) and agrees with what it replaces
    assert np.abs(limb["sums"] - other["sums"]).max() <= 1e-11 * scale
    assert np.array_equal(limb["labels"], other["labels"]) and np.array_equal(store["labels"], other["labels"])
    if n >= 1000:
        assert not np.array_equal(limb["sums"], other["sums"])


@pytest.mark.parametrize("kind", ["offset", "scales", "integers", "sparse"])
def test_limb_pass_data_kinds(hkm, oracle, kind):
    n, d, k = 70000, 16, 24
    X, C = _blobs(n, d, k, 5)
    if kind == "offset":
        X += 300.0; C += 300.0
    if kind == "scales":
        s = np.logspace(-6, 6, d); X *= s; C *= s
    if kind == "integers":
        X = np.round(X * 3); C = np.round(C * 3)
    if kind == "sparse":
        X[np.random.default_rng(1).random(X.shape) < 0.7] = 0.0
    store, limb, other = _three(hkm._native, X, C)
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    assert np.array_equal(limb["labels"], lab) and np.array_equal(limb["wsum"], wsum)
    col = np.maximum(np.abs(X).max(axis=0), 1e-300) * n
    assert (np.abs(limb["sums"] - sums) / col[None, :]).max() <= 1e-13
    if kind == "integers":
        assert np.array_equal(limb["sums"], sums)                  # integer data: both the fixed point and the fp64 sums are exact
    assert limb["cost"] == pytest.approx(cost, rel=1e-9, abs=1e-9 * float((X * X).sum() + 1))


@pytest.mark.parametrize("bad", ["outlier", "inf", "nan", "huge"])
def test_limb_pass_declines(hkm, oracle, bad):
    """Columns the fixed point cannot carry leave the pass to the other variant: bit-identical results with the switch on and off."""
    n, d, k = 50000, 16, 16
    X, C = _blobs(n, d, k, 9)
    if bad == "outlier":
        X[n // 2, 3] = 1e15                                        # quantum of column 3 would be ~1, its typical values are ~1
    if bad == "inf":
        X[7, 0] = np.inf
    if bad == "nan":
        X[7, 0] = np.nan
    if bad == "huge":
        X[:, 5] *= 1e307
    store, limb, other = _three(hkm._native, X, C, want_labels=False)
    assert np.array_equal(limb["sums"], other["sums"], equal_nan=True) and np.array_equal(limb["wsum"], other["wsum"])
    assert (limb["cost"] == other["cost"]) or (np.isnan(limb["cost"]) and np.isnan(other["cost"]))


def test_limb_pass_needs_the_record(hkm, oracle):
    """A scratch of only the core size (no room for the column range record), or a REUSE pass without a STORE pass on this scratch,
    runs the other variant."""
    N = hkm._native
    n, d, k = 40000, 16, 16
    X, C = _blobs(n, d, k, 11)
    Xd = torch.from_numpy(X).cuda(); Cd = torch.from_numpy(C).cuda()
    P = k * d + k + 1
    xs = torch.tensor([float((X * X).sum())], dtype=torch.float64, device="cuda")
    full = N.new_scratch(d, k, "cuda")
    ref = torch.empty(P, dtype=torch.float64, device="cuda")
    N.load().hkm_debug_wl_off(ctypes.c_int(1))
    try:
        N.assign_reduce(Xd, Cd, ref, scratch=full, xnorm2_sum=xs, xnorm2_mode=2)
    finally:
        N.load().hkm_debug_wl_off(ctypes.c_int(0))
    fresh = N.new_scratch(d, k, "cuda")                            # no STORE pass has written a record here
    a = torch.empty(P, dtype=torch.float64, device="cuda")
    N.assign_reduce(Xd, Cd, a, scratch=fresh, xnorm2_sum=xs, xnorm2_mode=2)
    torch.cuda.synchronize()
    assert torch.equal(a, ref)
    # the record of ANOTHER table (same shape) must not be used either
    X2 = torch.from_numpy(X + 1.0).cuda()
    xs2 = torch.zeros(1, dtype=torch.float64, device="cuda")
    N.assign_reduce(X2, Cd, a, scratch=fresh, xnorm2_sum=xs2, xnorm2_mode=1)
    N.assign_reduce(Xd, Cd, a, scratch=fresh, xnorm2_sum=xs, xnorm2_mode=2)
    torch.cuda.synchronize()
    assert torch.equal(a, ref)


def test_limb_pass_is_bit_stable_and_order_free(hkm):
    """Integer sums: repeated passes are bit-identical, and so are passes over the same rows in another order."""
    N = hkm._native
    n, d, k = 300_000, 16, 16
    X, C = _blobs(n, d, k, 13)
    perm = np.random.default_rng(2).permutation(n)
    res = []
    for Xh in (X, X, X[perm]):
        store, limb, other = _three(N, Xh, C, want_labels=False)
        res.append(limb)
    assert np.array_equal(res[0]["sums"], res[1]["sums"]) and res[0]["cost"] == res[1]["cost"]
    # another row order: other CTAs own the rows, so the per-CTA roundings differ, but only there (one rounding per CTA and entry)
    assert np.abs(res[0]["sums"] - res[2]["sums"]).max() <= 1e-12 * np.abs(res[0]["sums"]).max()
    assert np.array_equal(res[0]["wsum"], res[2]["wsum"])


def test_fit_with_limb_passes_matches_oracle(hkm, oracle):
    """A whole fit (fused driver steps: STORE pass, then REUSE passes on the limb variant) against the oracle's Lloyd loop."""
    from hkm_b200.ml.clustering import KMeans
    n, d, k = 120_000, 16, 16
    X, C = _blobs(n, d, k, 17, spread=1.0)
    model = KMeans(k=k, maxIter=8, tol=0.0, seed=1).setInitialCenters(C).fit({"features": X})
    Co, it, cost = oracle.lloyd(X, C, max_iter=8, tol=0.0, P=1)
    G = np.array(model.clusterCenters())
    assert model.summary.numIter == it
    assert np.allclose(G, Co, rtol=1e-9, atol=1e-11)
    assert model.summary.trainingCost == pytest.approx(cost, rel=1e-9)


@pytest.mark.parametrize("n,d,m", [(50000, 16, 12), (30011, 24, 33), (40000, 32, 64), (20000, 48, 20), (129, 16, 16), (70003, 8, 40), (60000, 16, 100),
                                   (50000, 20, 9), (33001, 18, 17), (20000, 40, 64), (9000, 34, 48), (5, 24, 12)])
def test_distance_variant_kmpar_passes(hkm, n, d, m):
    """k-means|| data passes on the distance-returning variant of the same kernel (running minimum of the squared distance to the
    new centres, its sum, nearest candidate per row and the counts): against NumPy in the oracle's arithmetic, and against the
    round-1 kernels with the variant switched off."""
    N = hkm._native
    X, C = _blobs(n, d, m, 3 * n + d + m)
    Xd = torch.from_numpy(X).cuda(); Cd = torch.from_numpy(C).cuda()
    rng = np.random.default_rng(5)
    start = rng.random(n) * 40.0                                          # costs so far: some rows improve, some do not
    # reference: dist = max(|x|^2 - 2 (x.c - 0.5|c|^2), 0), nearest = lowest index among equals
    xx = (X * X).sum(axis=1)
    sc = X @ C.T - 0.5 * (C * C).sum(axis=1)[None, :]
    dist = np.maximum(xx[:, None] - 2.0 * sc, 0.0)
    want = np.minimum(start, dist.min(axis=1))
    scale = float(xx.max()) + 1.0
    outs = []
    for off in (0, 1):
        N.load().hkm_debug_wl_off(ctypes.c_int(off))
        try:
            costs = torch.from_numpy(start.copy()).cuda(); tot = torch.zeros(1, dtype=torch.float64, device="cuda")
            N.kmpar_cost_update(Xd, Cd, costs, tot, N.new_scratch(d, min(m, 64), "cuda"))
            labels = torch.empty(n, dtype=torch.int32, device="cuda"); mind2 = torch.empty(n, dtype=torch.float64, device="cuda")
            counts = torch.zeros(m, dtype=torch.int64, device="cuda")
            N.closest_counts(Xd, Cd, labels, mind2, counts)
            torch.cuda.synchronize()
        finally:
            N.load().hkm_debug_wl_off(ctypes.c_int(0))
        outs.append((costs.cpu().numpy(), tot.item(), labels.cpu().numpy(), mind2.cpu().numpy(), counts.cpu().numpy()))
    for costs, tot, labels, mind2, counts in outs:
        assert np.abs(costs - want).max() <= 1e-9 * scale
        assert tot == pytest.approx(want.sum(), rel=1e-12)
        near = dist.argmin(axis=1)
        same = labels == near
        assert same.mean() >= 0.9999
        assert np.abs(mind2 - dist.min(axis=1)).max() <= 1e-9 * scale
        assert counts.sum() == n and np.array_equal(counts, np.bincount(labels, minlength=m))
    assert (outs[0][2] == outs[1][2]).mean() >= 0.9999 and np.abs(outs[0][0] - outs[1][0]).max() <= 1e-9 * scale


@pytest.mark.parametrize("n,d,k,spread", [(30011, 24, 32, 1.5), (1, 16, 17, 1.5), (129, 48, 64, 1.5), (70001, 16, 40, 0.2), (50000, 30, 33, 0.3),
                                          (20000, 2, 20, 1.0), (9000, 48, 17, 0.2), (40001, 20, 64, 0.3), (7000, 40, 24, 1.0)])
def test_labels_only_pass_k_over_16(hkm, oracle, n, d, k, spread):
    """transform / predict with 16 < k <= 64, even d <= 48: the labels-only form of the distance-returning kernel -- labels
    array_equal to the oracle (also on overlapping clusters, where many rows are re-scored exactly), equal to those of the
    kernel it replaces, and nothing else written."""
    N = hkm._native
    Xh, Ch = _blobs(n, d, k, seed=n + d + k, spread=spread)
    _, _, _, olab = oracle.lloyd_pass(Xh, Ch, P=1, want_labels=True)
    X = torch.from_numpy(Xh).cuda(); C = torch.from_numpy(Ch).cuda()
    lib = N.load()
    got = []
    for off in (0, 1):
        lib.hkm_debug_wl_off(ctypes.c_int(off))
        try:
            labels = torch.full((n,), -9, dtype=torch.int32, device="cuda")
            N.assign_reduce(X, C, None, labels=labels, scratch=N.new_scratch(d, k, "cuda"), do_reduce=False)
            torch.cuda.synchronize()
        finally:
            lib.hkm_debug_wl_off(ctypes.c_int(0))
        got.append(labels.cpu().numpy())
    assert np.array_equal(got[0], olab)
    assert np.array_equal(got[1], olab)
