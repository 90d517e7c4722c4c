"""Property-based GPU-vs-oracle parity of the Lloyd data pass and of a short fit: hypothesis draws the shape
(n in [1, 1e5], d in [1, 96], k in [2, 64]) and the pathologies -- duplicated rows, constant columns, centres nothing is
assigned to, duplicated centres (exact ties), integer-valued data (more exact ties), a large common offset.
Tolerances are north_star's: labels >= 99.99 % identical, sums / centres / cost 1e-6 relative."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
hypothesis = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings, strategies as st   # noqa: E402

pytestmark = pytest.mark.gpu


@st.composite
def problems(draw):
    d = draw(st.one_of(st.integers(1, 96), st.sampled_from([1, 2, 8, 16, 24, 31, 32, 33, 48, 63, 64, 65, 96])))
    k = draw(st.one_of(st.integers(2, 64), st.sampled_from([2, 8, 9, 16, 17, 32, 33, 48, 64])))
    n = draw(st.one_of(st.integers(1, 2000), st.integers(1, 100_000), st.sampled_from([1, 127, 128, 129, 4096])))
    seed = draw(st.integers(0, 2**31 - 1))
    flags = draw(st.fixed_dictionaries({"dups": st.booleans(), "const_col": st.booleans(), "far_centre": st.booleans(), "dup_centre": st.booleans(),
                                        "integers": st.booleans(), "offset": st.sampled_from([0.0, 0.0, 5.0, 300.0])}))
    return n, d, k, seed, flags


def build(n, d, k, seed, flags):
    rng = np.random.default_rng(seed)
    prof = rng.normal(size=(k, d)) * 2.0
    X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d))
    if flags["integers"]:
        X = np.round(X)
    X = X + flags["offset"]
    if flags["dups"] and n > 3:
        X[n // 2:] = X[: n - n // 2]                       # every row of the second half repeats one of the first
    if flags["const_col"]:
        X[:, rng.integers(0, d)] = 7.0
    C = X[rng.integers(0, n, size=k)].copy() if n >= 1 else prof
    if n < k:                                              # fewer rows than centres: the rest are synthetic
        C = np.vstack([X, prof[: k - n] + flags["offset"]])[:k]
    if flags["far_centre"]:
        C[k - 1] = 1e4                                     # attracts nothing: an empty cluster
    if flags["dup_centre"] and k > 2:
        C[k - 2] = C[0]                                    # exact tie: the lower index must win
    return np.ascontiguousarray(X), np.ascontiguousarray(C)


@settings(max_examples=60, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large, HealthCheck.function_scoped_fixture])
@given(problems())
def test_lloyd_pass_property(hkm, oracle, prob):
    n, d, k, seed, flags = prob
    X, C = build(n, d, k, seed, flags)
    N = hkm._native
    Xd = torch.from_numpy(X).cuda(); Cd = torch.from_numpy(C).cuda()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    N.assign_reduce(Xd, Cd, partial, labels, scratch=N.new_scratch(d, k, "cuda"))
    p = partial.cpu().numpy(); lab = labels.cpu().numpy()
    sums, wsum, cost, olab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    agree = (lab == olab).mean()
    assert agree >= 0.9999 or (lab != olab).sum() <= 1, (prob, agree)
    assert p[k * d:k * d + k].sum() == n
    if (lab == olab).all():
        assert np.array_equal(p[k * d:k * d + k], wsum)
        assert np.abs(p[:k * d].reshape(k, d) - sums).max() <= 1e-6 * max(1.0, np.abs(sums).max())
        if flags["dup_centre"] and k > 2:
            assert wsum[k - 2] == 0                        # the duplicate never wins a tie against centre 0
    scale = float((X * X).sum()) + 1.0
    assert p[-1] == pytest.approx(cost, rel=1e-6, abs=1e-9 * scale)


@settings(max_examples=15, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large, HealthCheck.function_scoped_fixture])
@given(problems())
def test_short_fit_property(hkm, oracle, prob):
    """Three Lloyd iterations from the same initial centres through the estimator: centres, cost, iteration count."""
    n, d, k, seed, flags = prob
    if n < 2:
        n = 2
    X, C = build(n, d, k, seed, flags)
    from hkm_b200.ml.clustering import KMeans
    C0 = np.unique(C, axis=0)                              # Spark's initial model holds distinct centres
    if C0.shape[0] < 2:
        return
    model = KMeans(k=max(2, C0.shape[0]), maxIter=3, tol=0.0, seed=1).setInitialCenters(C0).fit({"features": X})
    Co, it, cost = oracle.lloyd(X, C0, max_iter=3, tol=0.0, P=1)
    G = np.array(model.clusterCenters())
    assert model.summary.numIter == it
    scale = float((X * X).sum()) + 1.0
    # a label flip on an exact near-tie moves one row between two clusters; everything else is reassociation error
    lab_g = model.summary.predictions.to_numpy("prediction"); lab_o = oracle.predict(X, Co)
    if (lab_g == lab_o).all():
        assert np.allclose(G, Co, rtol=1e-6, atol=1e-9 * max(1.0, np.abs(Co).max()))
        assert model.summary.trainingCost == pytest.approx(cost, rel=1e-6, abs=1e-9 * scale)


@st.composite
def limb_problems(draw):
    """Shapes the limb-reduction variant takes (even d <= 64, 8 < k <= 64), same pathologies."""
    d = 2 * draw(st.one_of(st.integers(1, 32), st.sampled_from([1, 4, 8, 12, 15, 16, 17, 24, 31, 32])))
    k = draw(st.one_of(st.integers(9, 64), st.sampled_from([9, 16, 17, 32, 33, 48, 49, 64])))
    n = draw(st.one_of(st.integers(1, 2000), st.integers(1, 100_000), st.sampled_from([1, 127, 128, 129, 4096])))
    seed = draw(st.integers(0, 2**31 - 1))
    flags = draw(st.fixed_dictionaries({"dups": st.booleans(), "const_col": st.booleans(), "far_centre": st.booleans(), "dup_centre": st.booleans(),
                                        "integers": st.booleans(), "offset": st.sampled_from([0.0, 0.0, 5.0, 300.0])}))
    return n, d, k, seed, flags


@settings(max_examples=60, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow, HealthCheck.data_too_large, HealthCheck.function_scoped_fixture])
@given(limb_problems())
def test_reuse_pass_property(hkm, oracle, prob):
    """STORE pass, then a REUSE pass on the same scratch (the limb-reduction variant wherever the column ranges admit the table)."""
    n, d, k, seed, flags = prob
    X, C = build(n, d, k, seed, flags)
    N = hkm._native
    Xd = torch.from_numpy(X).cuda(); Cd = torch.from_numpy(C).cuda()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    scratch = N.new_scratch(d, k, "cuda"); xs = torch.zeros(1, dtype=torch.float64, device="cuda")
    N.assign_reduce(Xd, Cd, partial, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=1)
    N.assign_reduce(Xd, Cd, partial, labels, scratch=scratch, xnorm2_sum=xs, xnorm2_mode=2)
    p = partial.cpu().numpy(); lab = labels.cpu().numpy()
    sums, wsum, cost, olab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    agree = (lab == olab).mean()
    assert agree >= 0.9999 or (lab != olab).sum() <= 1, (prob, agree)
    assert p[k * d:k * d + k].sum() == n
    if (lab == olab).all():
        assert np.array_equal(p[k * d:k * d + k], wsum)
        assert np.abs(p[:k * d].reshape(k, d) - sums).max() <= 1e-6 * max(1.0, np.abs(sums).max())
        if flags["dup_centre"] and k > 2:
            assert wsum[k - 2] == 0
    scale = float((X * X).sum()) + 1.0
    # REUSE forms the cost as sum|x|^2 - 2 sum score: absolute accuracy ~1e-16 sum|x|^2 (include/hkm.h)
    assert p[-1] == pytest.approx(cost, rel=1e-6, abs=1e-9 * scale)
