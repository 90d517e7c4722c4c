"""Pins the CPU oracle (and the product's host-side RNG code) to every known answer that can be
reproduced offline.  The reference itself holds no KMeans code, tests or fixtures, so
these upstream constants are what "parity" is anchored to."""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")
KAT = json.load(open(os.path.join(GOLD, "kat.json")))


def iris():
    a = np.loadtxt(os.path.join(GOLD, "iris_uci.csv"), delimiter=",")
    return np.ascontiguousarray(a[:, :4]), a[:, 4].astype(np.int32)


def test_java_random_kat(oracle):
    k = KAT["java_random"]
    assert oracle.JavaRandom(0).next_int() == k["seed0_nextInt"]
    assert oracle.JavaRandom(42).next_int() == k["seed42_nextInt"]
    assert oracle.JavaRandom(0).next_double() == k["seed0_nextDouble"]
    r = oracle.JavaRandom(42)
    assert [r.next_int(10) for _ in range(5)] == k["seed42_nextInt10_x5"]


def test_xorshift_matches_pyspark_rand_docs(oracle):
    k = KAT["xorshift"]
    r = oracle.XORShiftRandom(42)
    assert [r.next_double() * 3, r.next_double() * 3] == k["seed42_rand_times3"]
    r = oracle.XORShiftRandom(42)
    assert [r.next_gaussian(), r.next_gaussian()] == k["seed42_randn"]


def test_hash_seed_bits(oracle):
    # XORShiftRandomSuite "hashSeed has random bits throughout"
    total = 0
    for seed in range(10):
        bits = bin(oracle.hash_seed(seed)).count("1")
        assert 20 < bits < 44
        total += bits
    assert 32 * 10 - 30 < total < 32 * 10 + 30
    assert oracle.XORShiftRandom(0).next_int() != 0  # "XORShift with zero seed"


def test_iris_silhouette_constants(oracle):
    X, y = iris()
    k = KAT["iris"]
    assert oracle.silhouette(X, y) == pytest.approx(k["silhouette_sqeuclidean"], rel=k["rel_tol"])
    assert oracle.silhouette(X, y, distance="cosine") == pytest.approx(k["silhouette_cosine"], rel=k["rel_tol"])
    assert oracle.np_silhouette(X, y) == pytest.approx(oracle.silhouette(X, y), rel=1e-12)


def test_clustering_evaluator_doctest(oracle):
    """pyspark.ml.evaluation.ClusteringEvaluator doctest: 0.9079... (unweighted: an upstream constant; weighted: the published
    formulas applied to the doctest's weights, see make_golden.py)."""
    k = KAT["clustering_evaluator_doctest"]
    X = np.array(k["points"]); lab = np.array(k["prediction"], dtype=np.int32)
    s = oracle.silhouette(X, lab, P=2)
    assert abs(s - k["silhouette_prefix"]) < 1e-4 and f"{s:.10f}".startswith("0.9079")
    assert oracle.np_silhouette(X, lab) == pytest.approx(s, rel=1e-13)
    sw = oracle.silhouette(X, lab, w=np.array(k["weights"]), P=3)
    assert sw == pytest.approx(k["weighted_by_formula"], rel=1e-12)
    assert oracle.np_silhouette(X, lab, np.array(k["weights"])) == pytest.approx(sw, rel=1e-13)
    # uniform weights change nothing
    assert oracle.silhouette(X, lab, w=np.full(6, 2.5), P=2) == pytest.approx(s, rel=1e-13)


def test_standard_scaler_doctest(oracle):
    k = KAT["standard_scaler_doctest"]
    X = np.array(k["rows"])
    mean, std = oracle.scaler_fit(X, P=2)
    assert mean == pytest.approx(k["mean"], abs=1e-15) and std == pytest.approx(k["std"], abs=k["print_tol"]) and std[0] == pytest.approx(np.sqrt(2.0), rel=1e-15)
    Y = oracle.scaler_transform(X, mean, std, with_mean=False, with_std=True)      # pyspark defaults
    assert Y == pytest.approx(np.array(k["scaled"]), abs=k["print_tol"]) and Y[0, 0] == 0.0


def test_silhouette_equals_sklearn_sqeuclidean(oracle):
    from sklearn.metrics import silhouette_score
    rng = np.random.default_rng(3)
    X = rng.normal(size=(500, 6)) + rng.integers(0, 3, size=(500, 1)) * 2.5
    lab = rng.integers(0, 3, size=500).astype(np.int32)
    assert oracle.silhouette(X, lab, P=4) == pytest.approx(silhouette_score(X, lab, metric="sqeuclidean"), abs=1e-12)


def test_silhouette_singleton_and_weights(oracle):
    X = np.array([[0.0, 0], [0, 1], [5, 5], [9, 9]])
    lab = np.array([0, 0, 1, 2], dtype=np.int32)  # clusters 1 and 2 are singletons -> s = 0
    s = oracle.silhouette(X, lab, P=1)
    assert s == pytest.approx(oracle.np_silhouette(X, lab), rel=1e-13)
    w = np.array([1.0, 2.0, 3.0, 0.5])
    assert oracle.silhouette(X, lab, w=w, P=2) == pytest.approx(oracle.np_silhouette(X, lab, w), rel=1e-13)
    assert np.isnan(oracle.silhouette(X, np.zeros(4, dtype=np.int32)))  # one cluster -> Spark requires > 1


@pytest.mark.parametrize("mode", ["random", "k-means||"])
def test_kmeans_single_cluster(oracle, mode):
    k = KAT["kmeans_single_cluster"]
    pts = np.array(k["points"], dtype=np.float64)
    for max_iter in (1, 2, 5):
        r = oracle.kmeans_fit(pts, 1, seed=1, init_mode=mode, max_iter=max_iter, P=2)
        if mode == "random" and max_iter == 1:
            continue  # one Lloyd step from a data point already lands on the mean; keep the loop simple
        assert np.allclose(r["centers"], [k["center"]], atol=1e-12)


def test_pyspark_kmeans_doctest_cost(oracle):
    k = KAT["pyspark_doctest"]
    pts = np.array(k["points"], dtype=np.float64)
    w = np.array(k["weights"], dtype=np.float64)
    for seed in (1, 7, 42):
        r = oracle.kmeans_fit(pts, k["k"], seed=seed, max_iter=10, P=2, w=w)
        assert r["training_cost"] == pytest.approx(k["trainingCost"], abs=1e-12)
        assert np.allclose(sorted(map(list, r["centers"])), k["centers_sorted"])


def test_lloyd_matches_numpy_twin_and_sklearn(oracle):
    from sklearn.cluster import KMeans as SK
    rng = np.random.default_rng(0)
    X = rng.normal(size=(4000, 6)) + rng.integers(0, 4, size=(4000, 1)) * 3
    C0 = oracle.init_kmeans_parallel(X, 4, seed=42, P=8)
    C, it, cost = oracle.lloyd(X, C0, max_iter=50, tol=1e-9, P=8)
    Cn, itn, costn = oracle.np_lloyd(X, C0, max_iter=50, tol=1e-9)
    assert it == itn and cost == pytest.approx(costn, rel=1e-12) and np.allclose(C, Cn, atol=1e-10)
    sk = SK(n_clusters=4, init=C0, n_init=1, algorithm="lloyd", tol=0, max_iter=300).fit(X)
    assert np.allclose(np.sort(sk.cluster_centers_, axis=0), np.sort(C, axis=0), atol=1e-6)


def test_pruned_find_closest_equals_plain(oracle):
    rng = np.random.default_rng(5)
    X = rng.normal(size=(3000, 7)); C = X[:11].copy()
    a = oracle.lloyd_pass(X, C, P=3, use_stats=False, want_labels=True)
    b = oracle.lloyd_pass(X, C, P=3, use_stats=True, want_labels=True)
    assert np.array_equal(a[3], b[3]) and a[2] == b[2] and np.array_equal(a[0], b[0])
    # the pre-3.1 norm-expansion form agrees to rounding
    c = oracle.lloyd_pass(X, C, P=3, form=1, want_labels=True)
    assert (c[3] == a[3]).mean() == 1.0 and c[2] == pytest.approx(a[2], rel=1e-12)


def test_training_cost_is_cost_before_last_update(oracle):
    rng = np.random.default_rng(9)
    X = rng.normal(size=(500, 3)); C0 = X[:3].copy()
    C, it, cost = oracle.lloyd(X, C0, max_iter=1, tol=0.0, P=2)
    assert it == 1
    assert cost == pytest.approx(oracle.lloyd_pass(X, C0, P=2)[2], rel=1e-15)      # w.r.t. centres BEFORE the update
    assert cost > oracle.lloyd_pass(X, C, P=2)[2]
    C2, it0, cost0 = oracle.lloyd(X, C0, max_iter=0, P=2)
    assert it0 == 0 and cost0 == 0.0 and np.array_equal(C2, C0)


def test_empty_cluster_keeps_centre(oracle):
    X = np.array([[0.0], [0.1], [0.2], [10.0]])
    C0 = np.array([[0.1], [10.0], [100.0]])  # third centre attracts nothing
    C, it, cost = oracle.lloyd(X, C0, max_iter=5, tol=1e-4, P=2)
    assert C[2, 0] == 100.0 and C[1, 0] == 10.0 and C[0, 0] == pytest.approx(0.1)


def test_scaler_matches_numpy(oracle):
    rng = np.random.default_rng(2)
    X = rng.normal(size=(1000, 5)) * [1, 10, 100, 0.1, 5] + [0, 5, -3, 100, 0]
    X[:, 2] = np.round(X[:, 2]); X[::3, 2] = 0.0   # exercise the zero-skipping path
    X = np.column_stack([X, np.full(1000, 7.0)])    # constant column -> std 0 -> scaled to 0
    mean, std = oracle.scaler_fit(X, P=8)
    assert np.allclose(mean, X.mean(0), rtol=1e-12)
    assert np.allclose(std[:5], X[:, :5].std(0, ddof=1), rtol=1e-12) and std[5] == pytest.approx(0.0, abs=1e-9)
    Y = oracle.scaler_transform(X, mean, np.where(std < 1e-9, 0.0, std), with_mean=True, with_std=True)
    assert np.allclose(Y[:, :5].std(0, ddof=1), 1.0) and np.all(Y[:, 5] == 0.0)


def test_take_sample_properties(oracle):
    for n, P, num, seed in [(1000, 4, 3, 7), (100000, 8, 1, -5), (37, 3, 5, 99), (5, 2, 9, 1)]:
        idx = oracle.take_sample(n, P, num, seed)
        assert len(idx) == min(num, n) and len(set(idx.tolist())) == len(idx) and idx.min() >= 0 and idx.max() < n
        assert np.array_equal(idx, oracle.take_sample(n, P, num, seed))
    assert not np.array_equal(oracle.take_sample(1000, 4, 3, 7), oracle.take_sample(1000, 5, 3, 7))  # layout dependent


def test_kmeans_parallel_structure(oracle):
    rng = np.random.default_rng(11)
    X = rng.normal(size=(20000, 4)) + rng.integers(0, 6, size=(20000, 1)) * 4
    C, dbg = oracle.init_kmeans_parallel(X, 6, seed=42, P=8, debug=True)
    assert C.shape == (6, 4) and dbg["step_counts"].shape == (2,)
    assert 2 <= dbg["step_counts"][1] <= 60       # ~2k expected per step
    assert len(dbg["candidates"]) <= 1 + dbg["step_counts"].sum()
    # explicit first centre makes the run independent of takeSample
    C2 = oracle.init_kmeans_parallel(X, 6, seed=42, P=8, first_idx=int(dbg["chosen_idx"][0]))
    assert np.array_equal(C, C2)
    # fewer distinct points than k -> fewer centres, no error
    Xd = np.repeat(np.array([[0.0, 0], [1, 1], [2, 2]]), 50, axis=0)
    r = oracle.kmeans_fit(Xd, 5, seed=3, P=2)
    assert r["centers"].shape[0] <= 3 and r["training_cost"] == pytest.approx(0.0, abs=1e-20)


def test_cosine_kmeans_spark_suite_case_and_numpy_twin(oracle):
    """ml KMeansSuite "KMeans with cosine distance": 6 points in 3 directions, 2 partitions, k=3, random init, seed 42,
    tol 1e-6 -> three clusters, same-direction points together, unit-length centres.  Plus an independent NumPy Lloyd."""
    X = np.array([[1.0, 1.0], [10.0, 10.0], [1.0, 0.5], [10.0, 4.4], [-1.0, 1.0], [-100.0, 90.0]])
    r = oracle.kmeans_fit(X, 3, 42, tol=1e-6, init_mode="random", P=2, distance="cosine")
    lab = r["labels"]
    assert len(set(lab.tolist())) == 3 and lab[0] == lab[1] and lab[2] == lab[3] and lab[4] == lab[5]
    assert np.allclose(np.linalg.norm(r["centers"], axis=1), 1.0, atol=1e-15)

    rng = np.random.default_rng(0)
    Xr = rng.normal(size=(5000, 6)) + rng.integers(0, 4, size=(5000, 1)) * np.array([1, -1, 2, 0, 1, 3.0])
    C = Xr[:5].copy()
    Co, it_o, cost_o = oracle.lloyd(Xr, C, max_iter=20, tol=1e-4, P=4, form=oracle.FORM_COSINE)
    cn, xn = np.linalg.norm(C, axis=1), np.linalg.norm(Xr, axis=1)
    for it in range(20):
        D = 1 - (Xr @ C.T) / cn[None, :] / xn[:, None]
        lab = D.argmin(1)
        cost = D[np.arange(len(Xr)), lab].sum()
        conv = True
        for j in range(len(C)):
            m = lab == j
            if m.any():
                s = (Xr[m] / xn[m, None]).sum(0) / m.sum()
                s /= np.linalg.norm(s)
                conv &= bool(1 - C[j] @ s / cn[j] <= 1e-4)
                C[j], cn[j] = s, 1.0
        if conv:
            break
    assert it_o == it + 1 and cost_o == pytest.approx(cost, rel=1e-12) and np.abs(Co - C).max() < 1e-12
