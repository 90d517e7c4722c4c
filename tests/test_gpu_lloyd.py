"""GPU parity of the Lloyd data pass (hkm_assign_reduce / hkm_update_centers, through the C ABI)
against the CPU oracle on seeded inputs, plus size-independent properties at larger sizes."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

# north_star tolerances: labels >= 99.99 % identical (fp near-ties documented), sums / centres / cost 1e-6 relative.
# The kernels do far better on well-conditioned data; the asserts below use tighter bounds where exactness is expected.
LABEL_AGREE = 0.9999
REL = 1e-6


def _pass(N, Xh, Ch, wh=None, want=("labels", "mind2"), mode=0, xs=None):
    X = torch.from_numpy(Xh).cuda(); C = torch.from_numpy(Ch).cuda()
    n, d = Xh.shape; k = Ch.shape[0]
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda") if "labels" in want else None
    mind2 = torch.empty(n, dtype=torch.float64, device="cuda") if "mind2" in want else None
    w = torch.from_numpy(wh).cuda() if wh is not None else None
    scratch = N.new_scratch(d, k, "cuda")
    N.assign_reduce(X, C, partial, labels, mind2, weights=w, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=mode)
    torch.cuda.synchronize()
    p = partial.cpu().numpy()
    return dict(sums=p[:k * d].reshape(k, d), wsum=p[k * d:k * d + k], cost=p[-1],
                labels=None if labels is None else labels.cpu().numpy(), mind2=None if mind2 is None else mind2.cpu().numpy())


def _check(g, o_sums, o_wsum, o_cost, o_lab, o_dist, scale):
    assert (g["labels"] == o_lab).mean() >= LABEL_AGREE
    same = g["labels"] == o_lab
    assert np.abs(g["mind2"][same] - o_dist[same]).max() <= 1e-9 * scale
    if same.all():
        assert np.abs(g["sums"] - o_sums).max() <= 1e-11 * max(1.0, np.abs(o_sums).max())
        assert np.abs(g["wsum"] - o_wsum).max() <= 1e-9 * max(1.0, o_wsum.max())
    assert g["cost"] == pytest.approx(o_cost, rel=REL, abs=1e-9 * scale)


SHAPES = [(1000, 8, 5), (128, 32, 8), (127, 32, 8), (129, 32, 8), (4096, 32, 8), (100003, 32, 8), (50000, 16, 16), (30000, 24, 7),
          (30000, 24, 32), (20000, 5, 3), (20000, 7, 2), (20000, 64, 16), (20000, 33, 9), (77, 3, 2), (1, 4, 2), (20000, 64, 64),
          (5000, 1, 3), (5000, 2, 8), (9000, 12, 20), (6000, 48, 24), (3000, 63, 33), (3000, 96, 5), (5000, 80, 40), (4000, 66, 64),
          # 16 < k, d <= 32: the warp-specialised variant (score warps + reduce warps), all NT, ragged tails, one tile, < one tile
          (40000, 32, 64), (45001, 20, 50), (70003, 7, 33), (33333, 32, 17), (50000, 29, 41), (128, 24, 32), (100, 24, 24), (60000, 3, 64),
          (200000, 24, 32)]


@pytest.mark.parametrize("n,d,k", SHAPES)
def test_pass_matches_oracle(hkm, oracle, n, d, k):
    rng = np.random.default_rng(n * 131 + d * 17 + k)
    X = rng.normal(size=(n, d)) + rng.integers(0, k, size=(n, 1)) * 2.0
    C = X[rng.choice(n, k, replace=n < k)].copy() if n >= k else rng.normal(size=(k, d))
    g = _pass(hkm._native, X, C)
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    _, dist = oracle.predict(X, C, want_dist=True)
    _check(g, sums, wsum, cost, lab, dist, scale=float((X * X).sum(1).max()))


def test_weighted_pass(hkm, oracle):
    rng = np.random.default_rng(5)
    for n, d, k in [(50000, 32, 8), (50000, 16, 16), (3000, 9, 40)]:
        X = rng.normal(size=(n, d)) * 2; C = X[:k].copy(); w = rng.uniform(0.0, 3.0, size=n); w[::97] = 0.0
        g = _pass(hkm._native, X, C, wh=w)
        sums, wsum, cost, lab = oracle.lloyd_pass(X, C, w=w, P=1, want_labels=True)
        _, dist = oracle.predict(X, C, want_dist=True)
        _check(g, sums, wsum, cost, lab, dist, scale=float((X * X).sum(1).max()))


def test_exact_ties_pick_lowest_index(hkm, oracle):
    # integer lattice data with duplicated centres and midpoints: every tie must go to the lower index
    rng = np.random.default_rng(1)
    X = rng.integers(-3, 4, size=(20000, 6)).astype(np.float64)
    C = np.array([[0, 0, 0, 0, 0, 0], [2, 0, 0, 0, 0, 0], [0, 0, 0, 0, 0, 0], [-2, 0, 0, 0, 0, 0], [2, 0, 0, 0, 0, 0],
                  [0, 2, 0, 0, 0, 0], [0, -2, 0, 0, 0, 0], [1, 1, 1, 1, 1, 1], [1, 1, 1, 1, 1, 1]], dtype=np.float64)
    g = _pass(hkm._native, X, C)
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    assert np.array_equal(g["labels"], lab)            # exact arithmetic on small integers: no near-ties, only true ties
    assert wsum[2] == 0 and wsum[4] == 0 and wsum[8] == 0 and np.array_equal(g["wsum"], wsum)
    assert np.array_equal(g["sums"], sums) and g["cost"] == cost


def test_duplicate_rows_and_empty_clusters(hkm, oracle):
    X = np.repeat(np.array([[0.0, 0, 0], [1, 1, 1], [5, 5, 5]]), 1000, axis=0)
    C = np.array([[0.0, 0, 0], [1, 1, 1], [5, 5, 5], [100, 100, 100], [-50, 0, 0]])
    g = _pass(hkm._native, X, C)
    assert g["wsum"].tolist() == [1000, 1000, 1000, 0, 0] and g["cost"] == 0.0 and np.all(g["mind2"] == 0.0)
    assert np.array_equal(g["sums"][:3], C[:3] * 1000) and np.all(g["sums"][3:] == 0)


def test_store_then_reuse_modes(hkm, oracle):
    rng = np.random.default_rng(8)
    X = rng.normal(size=(70001, 24)) + 1.5; C = X[:12].copy()
    N = hkm._native
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    a = _pass(N, X, C, want=(), mode=0)
    b = _pass(N, X, C, want=(), mode=1, xs=xs)
    c = _pass(N, X, C, want=(), mode=2, xs=xs)
    assert np.array_equal(a["sums"], b["sums"]) and a["cost"] == b["cost"] and np.array_equal(a["sums"], c["sums"])
    assert xs.item() == pytest.approx(float((X * X).sum()), rel=1e-13)
    assert c["cost"] == pytest.approx(a["cost"], rel=1e-11)


def test_update_centers_and_state(hkm, oracle):
    rng = np.random.default_rng(4)
    N = hkm._native
    X = rng.normal(size=(30000, 10)) + rng.integers(0, 4, size=(30000, 1)) * 3
    C0 = np.vstack([X[:5], [[500.0] * 10]])          # last centre attracts nothing and must stay put
    Xd = torch.from_numpy(X).cuda(); C = torch.from_numpy(C0).cuda()
    k, d = C0.shape
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    scratch = N.new_scratch(d, k, "cuda"); state = N.new_state("cuda")
    Co = C0.copy()
    for it in range(1, 4):
        N.assign_reduce(Xd, C, partial, state=state, scratch=scratch)
        N.update_centers(partial, C, 1e-4, state)
        sums, wsum, cost, _ = oracle.lloyd_pass(X, Co, P=1)
        Co, conv = oracle.update_centers(sums, wsum, Co, 1e-4)
        st = N.read_state(state)
        assert st.iter == it and st.cost == pytest.approx(cost, rel=1e-12) and bool(st.converged) == conv
        assert np.allclose(C.cpu().numpy(), Co, rtol=1e-12, atol=1e-12) and C.cpu().numpy()[5, 0] == 500.0
    # once converged, further passes are no-ops on the device state
    C2, it_o, cost_o = oracle.lloyd(X, C0, max_iter=100, tol=1e-4, P=1)
    for _ in range(100):
        N.assign_reduce(Xd, C, partial, state=state, scratch=scratch)
        N.update_centers(partial, C, 1e-4, state)
    st = N.read_state(state)
    assert st.converged == 1 and st.iter == it_o and st.cost == pytest.approx(cost_o, rel=1e-10)
    assert np.allclose(C.cpu().numpy(), C2, rtol=1e-9, atol=1e-9)


def test_argument_errors(hkm):
    N = hkm._native
    X = torch.zeros(10, 4, dtype=torch.float64, device="cuda"); C = torch.zeros(2, 4, dtype=torch.float64, device="cuda")
    with pytest.raises(N.HkmError, match="CUDA tensor"):
        N.assign_reduce(X.cpu(), C, do_reduce=False)
    with pytest.raises(N.HkmError, match="reduce needs"):
        N.assign_reduce(X, C, do_reduce=True)
    with pytest.raises(N.HkmError, match="d <= 4096"):
        N.assign_reduce(torch.zeros(10, 5000, dtype=torch.float64, device="cuda"), torch.zeros(2, 5000, dtype=torch.float64, device="cuda"), do_reduce=False)
    with pytest.raises(N.HkmError, match="no fused step|no general-shape path"):
        Xw = torch.zeros(10, 500, dtype=torch.float64, device="cuda"); Cw = torch.zeros(2, 500, dtype=torch.float64, device="cuda")
        N.lloyd_step(Xw, Cw, torch.zeros(2 * 500 + 3, dtype=torch.float64, device="cuda"), 0.0, N.new_state("cuda"), N.new_scratch(500, 2, "cuda"))
    with pytest.raises(N.HkmError, match="16-byte aligned"):
        N.assign_reduce(torch.zeros(41, dtype=torch.float64, device="cuda")[1:].view(10, 4), C, do_reduce=False)


def test_full_size_properties(hkm):
    """cfg2 size (10 M x 32, k = 8): properties that need no oracle -- counts add up, sums equal
    the column totals, shard-by-shard partials add up to the whole (the multi-GPU reduction identity),
    and the pass is bit-stable run to run."""
    N = hkm._native
    n, d, k = 10_000_000, 32, 8
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    X = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    C = X[:k].clone()
    P = k * d + k + 1
    scratch = N.new_scratch(d, k, "cuda")
    p1 = torch.empty(P, dtype=torch.float64, device="cuda"); p2 = torch.empty_like(p1)
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    N.assign_reduce(X, C, p1, labels=labels, scratch=scratch)
    N.assign_reduce(X, C, p2, scratch=scratch)
    assert torch.equal(p1, p2)
    assert p1[k * d:k * d + k].sum().item() == n
    assert torch.equal(p1[k * d:k * d + k], torch.bincount(labels, minlength=k).double())
    col = X.sum(0)
    assert torch.allclose(p1[:k * d].view(k, d).sum(0), col, rtol=1e-9, atol=1e-6)
    acc = torch.zeros_like(p1)
    bounds = [0, 1_234_567, 5_000_064, 9_999_999, n]
    for lo, hi in zip(bounds[:-1], bounds[1:]):
        lo16 = lo  # rows are 256 B, so any row offset keeps 16-byte alignment
        N.assign_reduce(X[lo16:hi], C, p2, scratch=scratch)
        acc += p2
    assert torch.equal(acc[k * d:k * d + k], p1[k * d:k * d + k])
    assert torch.allclose(acc, p1, rtol=1e-10, atol=1e-8)


# ---------------------------------------------------------------------------------------------
# tcgen05 screening variant (k > 32, even 32 < d <= 64): the approximate tensor-core scores only
# pre-select; labels, sums and cost must still match the fp64 oracle.
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,d,k,offset,spread", [
    (40_000, 64, 64, 0.0, 2.0),      # well separated: almost every row has one candidate
    (40_000, 64, 64, 0.0, 0.05),     # centres nearly coincide: most rows are ambiguous and re-scored exactly
    (30_000, 64, 40, 25.0, 1.0),     # large common offset (withMean=False data): bound is loose, still exact
    (30_000, 34, 33, 0.0, 1.0),      # smallest eligible shape, padded K and N
    (20_011, 50, 64, 3.0, 0.5),      # ragged tail tile
])
def test_tensor_core_variant_matches_oracle(hkm, oracle, n, d, k, offset, spread):
    rng = np.random.default_rng(n + d + k)
    prof = rng.normal(size=(k, d)) * spread + offset
    X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    C = prof + 0.05 * rng.normal(size=(k, d))
    g = _pass(hkm._native, X, C)
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    _, dist = oracle.predict(X, C, want_dist=True)
    assert np.array_equal(g["labels"], lab)
    _check(g, sums, wsum, cost, lab, dist, scale=float((X * X).sum(1).max()))
    # steady-state mode (xnorm2 REUSE: winner scores come from the gather) gives the same sums and cost
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    a = _pass(hkm._native, X, C, want=(), mode=1, xs=xs)
    b = _pass(hkm._native, X, C, want=(), mode=2, xs=xs)
    assert np.array_equal(a["sums"], b["sums"]) and np.array_equal(a["wsum"], b["wsum"]) and np.array_equal(a["wsum"], g["wsum"])
    # the pass that also returns per-row distances runs another kernel family: same rows per cluster, another summation order
    assert np.abs(a["sums"] - g["sums"]).max() <= 1e-11 * max(1.0, np.abs(g["sums"]).max())
    assert b["cost"] == pytest.approx(cost, rel=1e-9, abs=1e-9 * float((X * X).sum()))


def test_tensor_core_variant_exact_ties(hkm, oracle):
    rng = np.random.default_rng(3)
    d, k = 40, 48
    X = rng.integers(-2, 3, size=(30_000, d)).astype(np.float64)
    C = rng.integers(-2, 3, size=(k, d)).astype(np.float64)
    C[7] = C[3]; C[40] = C[3]; C[20] = C[11]           # duplicated centres: ties must go to the lowest index
    g = _pass(hkm._native, X, C)
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    assert np.array_equal(g["labels"], lab) and wsum[7] == 0 and wsum[40] == 0 and wsum[20] == 0
    assert np.array_equal(g["sums"], sums) and np.array_equal(g["wsum"], wsum) and g["cost"] == cost


# ---------------------------------------------------------------------------------------------
# Pipelined tcgen05 variant (hkm_lloyd_tp.cuh): plain passes with k > 8 (labels and / or sums, no per-row distances).
# Every reduce scheme it has (one-hot DMMA, private shared-memory sets, owned shared set) must give the oracle's labels,
# sums, counts and cost, and the same results as the fp64 kernel families.
# ---------------------------------------------------------------------------------------------
TP_SHAPES = [(5000, 16, 16), (128, 16, 16), (100, 24, 9), (30011, 24, 17), (30000, 24, 32), (60000, 32, 64), (45000, 20, 50), (300000, 7, 33),
             (20000, 64, 16), (20000, 64, 64), (6000, 48, 24), (4000, 62, 64), (9000, 12, 20), (3000, 31, 33), (70003, 5, 12), (33333, 2, 10),
             (2000, 1, 9), (65, 64, 64), (200000, 16, 16)]


@pytest.mark.parametrize("variant", [4, 40, 42, 43])
@pytest.mark.parametrize("n,d,k", TP_SHAPES)
def test_pipelined_variant_matches_oracle(hkm, oracle, n, d, k, variant):
    if variant == 40 and d > 32 or variant == 42 and d % 2 or variant == 43 and (k > 16 or (d > 32 and d % 2)) or variant == 4 and d > 32 and d % 2:
        pytest.skip("scheme not built for this shape")
    N = hkm._native
    rng = np.random.default_rng(n + 31 * d + 977 * k)
    prof = rng.normal(size=(k, d)) * 1.5 + 3.0                       # a common offset, as withMean=False tables have
    X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    C = prof + 0.05 * rng.normal(size=(k, d))
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    N.set_lloyd_variant(variant)
    try:
        xs = torch.zeros(1, dtype=torch.float64, device="cuda")
        a = _pass(N, X, C, want=("labels",))
        b = _pass(N, X, C, want=(), mode=1, xs=xs)
        c = _pass(N, X, C, want=(), mode=2, xs=xs)
        Xd = torch.from_numpy(X).cuda(); Cd = torch.from_numpy(C).cuda()
        pred = torch.full((n,), -5, dtype=torch.int32, device="cuda")
        N.assign_reduce(Xd, Cd, None, pred, do_reduce=False)
        N.set_lloyd_variant(5)
        ref = _pass(N, X, C, want=("labels",))
    finally:
        N.set_lloyd_variant(0)
    assert np.array_equal(a["labels"], lab) and np.array_equal(pred.cpu().numpy(), lab) and np.array_equal(ref["labels"], lab)
    assert np.array_equal(a["wsum"], wsum)
    assert np.abs(a["sums"] - sums).max() <= 1e-11 * max(1.0, np.abs(sums).max())
    assert a["cost"] == pytest.approx(cost, rel=1e-9) and c["cost"] == pytest.approx(cost, rel=1e-9)
    assert np.array_equal(a["sums"], b["sums"]) and a["cost"] == b["cost"] and np.array_equal(a["sums"], c["sums"])
    assert xs.item() == pytest.approx(float((X * X).sum()), rel=1e-12)
    assert a["cost"] == pytest.approx(ref["cost"], rel=1e-10)


def test_pipelined_variant_near_ties_and_exact_ties(hkm, oracle):
    """Centres that nearly coincide (most rows undecided by the tf32 screen -> exact fp64 re-scoring), a large common offset (loose
    bound), and exact ties on an integer lattice (lowest index must win)."""
    N = hkm._native
    rng = np.random.default_rng(12)
    for n, d, k, offset, spread in [(40_000, 16, 16, 0.0, 0.02), (30_000, 24, 32, 40.0, 1.0), (20_011, 32, 64, 3.0, 0.05), (30_000, 64, 40, 25.0, 1.0)]:
        prof = rng.normal(size=(k, d)) * spread + offset
        X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
        C = prof + 0.05 * rng.normal(size=(k, d))
        g = _pass(N, X, C, want=("labels",))
        sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
        assert np.array_equal(g["labels"], lab) and np.array_equal(g["wsum"], wsum)
        assert g["cost"] == pytest.approx(cost, rel=1e-9, abs=1e-9 * float((X * X).sum()))
    X = rng.integers(-2, 3, size=(30_000, 12)).astype(np.float64)
    C = rng.integers(-2, 3, size=(14, 12)).astype(np.float64)
    C[7] = C[3]; C[12] = C[3]; C[9] = C[1]
    g = _pass(N, X, C, want=("labels",))
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    assert np.array_equal(g["labels"], lab) and wsum[7] == 0 and wsum[12] == 0 and wsum[9] == 0
    assert np.array_equal(g["sums"], sums) and np.array_equal(g["wsum"], wsum) and g["cost"] == pytest.approx(cost, rel=1e-12)


@pytest.mark.parametrize("n,d,k", [(400_000, 24, 32), (300_001, 16, 48), (250_000, 32, 64), (300_000, 32, 8), (300_000, 16, 16)])
def test_passes_are_bit_stable(hkm, n, d, k):
    """Fixed summation order in every variant (no atomics on the data path): repeated passes over the same input give
    bit-identical sums, counts, cost and labels -- which is also what a shared-memory race would break."""
    N = hkm._native
    g = torch.Generator(device="cuda"); g.manual_seed(n + d + k)
    prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
    X = prof[torch.randint(0, k, (n,), device="cuda", generator=g)] + torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g)
    C = prof + 0.05
    scratch = N.new_scratch(d, k, "cuda")
    xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    ref = None
    for rep in range(6):
        partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
        labels = torch.empty(n, dtype=torch.int32, device="cuda")
        N.assign_reduce(X, C, partial, labels, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1 if rep == 0 else 2)
        if rep == 0:
            continue                       # STORE pass: its cost is formed per row, the REUSE passes below from the sums
        cur = (partial.clone(), labels.clone())
        if ref is None:
            ref = cur
        else:
            assert torch.equal(ref[0], cur[0]) and torch.equal(ref[1], cur[1])
    assert int(ref[0][k * d:k * d + k].sum().item()) == n


@pytest.mark.parametrize("n,d,k", [(36_000_000, 64, 64), (70_000_000, 32, 8), (90_000_000, 24, 32), (46_000_000, 48, 24)])
def test_more_than_2_31_elements(hkm, n, d, k):
    """n * d > 2^31 (cfg3 / cfg4 sized shards): 64-bit row addressing in every variant.  Properties that need no oracle:
    counts add up and equal the histogram of the labels, per-cluster sums add up to the column totals, and the labels
    of the LAST rows (beyond element 2^31) are the fp64 argmin computed by torch."""
    assert n * d > 2 ** 31
    N = hkm._native
    g = torch.Generator(device="cuda"); g.manual_seed(n % 1000 + d)
    prof = torch.randn(k, d, dtype=torch.float64, device="cuda", generator=g) * 2
    X = torch.empty(n, d, dtype=torch.float64, device="cuda")
    step = 4_000_000
    for lo in range(0, n, step):                     # generated in slabs: no second full-size temporary
        hi = min(n, lo + step)
        X[lo:hi] = prof[torch.randint(0, k, (hi - lo,), device="cuda", generator=g)] + torch.randn(hi - lo, d, dtype=torch.float64, device="cuda", generator=g)
    C = prof + 0.05
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    N.assign_reduce(X, C, partial, labels, scratch=N.new_scratch(d, k, "cuda"))
    counts = partial[k * d:k * d + k]
    assert counts.sum().item() == n
    assert torch.equal(counts, torch.bincount(labels, minlength=k).double())
    col = torch.zeros(d, dtype=torch.float64, device="cuda")
    for lo in range(0, n, step):
        col += X[lo:lo + step].sum(0)
    assert torch.allclose(partial[:k * d].view(k, d).sum(0), col, rtol=1e-9, atol=1e-5)
    tail = X[n - 200_000:]
    d2 = (tail * tail).sum(1, keepdim=True) - 2.0 * tail @ C.T + (C * C).sum(1)[None, :]
    assert (d2.argmin(1).int() == labels[n - 200_000:]).float().mean().item() >= 0.99999
