"""GPU parity of ClusteringEvaluator (squared-Euclidean silhouette) against the oracle and against
Spark's own iris constant.  Tolerance from north_star: 1e-5 (the kernels are ~1e-12)."""
import json
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
KAT = json.load(open(os.path.join(GOLD, "kat.json")))
TOL = 1e-5


def test_iris_constant_from_spark_suite(hkm):
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    a = np.loadtxt(os.path.join(GOLD, "iris_uci.csv"), delimiter=",")
    ev = ClusteringEvaluator(featuresCol="features", predictionCol="prediction")
    s = ev.evaluate({"features": np.ascontiguousarray(a[:, :4]), "prediction": a[:, 4].astype(np.int32)})
    assert s == pytest.approx(KAT["iris"]["silhouette_sqeuclidean"], rel=KAT["iris"]["rel_tol"])
    assert ev.isLargerBetter() and ev.getMetricName() == "silhouette" and ev.getDistanceMeasure() == "squaredEuclidean"


def test_pyspark_doctest_constant(hkm, oracle):
    """pyspark.ml.evaluation.ClusteringEvaluator doctest: 0.9079... on six points; with the doctest's weights the value the
    published formulas give (tests/golden/make_golden.py)."""
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    k = KAT["clustering_evaluator_doctest"]
    data = {"features": np.array(k["points"]), "prediction": np.array(k["prediction"], dtype=np.float64), "weight": np.array(k["weights"])}
    ev = ClusteringEvaluator()
    s = ev.evaluate(data)
    assert f"{s:.10f}".startswith("0.9079") and s == pytest.approx(oracle.silhouette(data["features"], np.array(k["prediction"], dtype=np.int32)), abs=1e-12)
    ev.setWeightCol("weight")
    assert ev.evaluate(data) == pytest.approx(k["weighted_by_formula"], abs=1e-12)


@pytest.mark.parametrize("n,d,k", [(500, 6, 3), (20000, 8, 5), (100_003, 24, 16), (30000, 32, 32), (5000, 3, 2), (40000, 64, 9), (129, 5, 64)])
def test_matches_oracle(hkm, oracle, n, d, k):
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    rng = np.random.default_rng(n + d + k)
    X = rng.normal(size=(n, d)) + rng.integers(0, k, size=(n, 1)) * 2.0
    lab = oracle.predict(X, X[rng.choice(n, k, replace=False)])
    s = ClusteringEvaluator().evaluate({"features": X, "prediction": lab})
    assert s == pytest.approx(oracle.silhouette(X, lab, kmax=k), abs=TOL)
    assert s == pytest.approx(oracle.silhouette(X, lab, kmax=k), abs=1e-10)


def test_weights_singletons_gaps_and_errors(hkm, oracle):
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    rng = np.random.default_rng(1)
    X = rng.normal(size=(3000, 4)) * 2
    lab = rng.integers(0, 4, size=3000).astype(np.int32) * 3        # ids 0,3,6,9: gaps never occur in `prediction`
    lab[0] = 11                                                      # a singleton cluster -> s = 0 for that row
    w = rng.uniform(0.2, 3.0, size=3000)
    s = ClusteringEvaluator(weightCol="w").evaluate({"features": X, "prediction": lab, "w": w})
    assert s == pytest.approx(oracle.silhouette(X, lab, w=w), abs=1e-10)
    s2 = ClusteringEvaluator().evaluate({"features": X, "prediction": lab.astype(np.float64)})   # Spark casts prediction to double
    assert s2 == pytest.approx(oracle.silhouette(X, lab), abs=1e-10)
    with pytest.raises(AssertionError, match="greater than one"):
        ClusteringEvaluator().evaluate({"features": X, "prediction": np.zeros(3000, dtype=np.int32)})
    with pytest.raises(hkm.IllegalArgumentException):
        ClusteringEvaluator(metricName="rmse")
    with pytest.raises(hkm.IllegalArgumentException):
        ClusteringEvaluator(distanceMeasure="manhattan")


def test_model_selection_sweep_cfg5_shape(hkm, oracle):
    """BASELINE configs[4] idiom at reduced n: fit k = 2..6 and score each model."""
    from hkm_b200 import datagen
    from hkm_b200.ml.clustering import KMeans
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    X = datagen.hospital_matrix(20_000, d=24, profiles=4, seed=7)
    ev = ClusteringEvaluator()
    for k in range(2, 7):
        model = KMeans(k=k, seed=42, maxIter=20).fit({"features": X})
        s = ev.evaluate(model.summary.predictions)                  # device-resident features + prediction columns
        ref = oracle.kmeans_fit(X, k, 42)
        assert s == pytest.approx(oracle.silhouette(X, ref["labels"]), abs=TOL)


def test_full_size_invariants(hkm):
    """5 M x 24 (cfg5 size): label-permutation invariance and the [-1, 1] range."""
    from hkm_b200.ml.evaluation import silhouette
    g = torch.Generator(device="cuda"); g.manual_seed(5)
    n, d, k = 5_000_000, 24, 8
    lab = torch.randint(0, k, (n,), device="cuda", generator=g, dtype=torch.int32)
    X = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g) + lab.double()[:, None] * 1.5
    s = silhouette(X, lab)
    perm = torch.tensor([3, 0, 7, 1, 6, 2, 5, 4], device="cuda", dtype=torch.int32)
    assert -1.0 <= s <= 1.0 and s > 0.3
    assert silhouette(X, perm[lab.long()]) == pytest.approx(s, abs=1e-12)
    assert silhouette(X, lab) == s                                   # bit-stable


def test_cosine_silhouette_iris_constant_and_oracle(hkm, oracle):
    """distanceMeasure='cosine' (SURVEY 8f-4): Spark's ClusteringEvaluatorSuite constant on iris, and the oracle."""
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    a = np.loadtxt(os.path.join(GOLD, "iris_uci.csv"), delimiter=",")
    ev = ClusteringEvaluator(distanceMeasure="cosine")
    s = ev.evaluate({"features": np.ascontiguousarray(a[:, :4]), "prediction": a[:, 4].astype(np.int32)})
    assert s == pytest.approx(KAT["iris"]["silhouette_cosine"], rel=KAT["iris"]["rel_tol"])
    rng = np.random.default_rng(4)
    X = rng.normal(size=(30_000, 24)) + rng.integers(0, 6, size=(30_000, 1)) * 1.5 + 0.5
    lab = rng.integers(0, 6, size=30_000).astype(np.int32)
    w = rng.uniform(0.5, 2.0, size=30_000)
    assert ev.evaluate({"features": X, "prediction": lab}) == pytest.approx(oracle.silhouette(X, lab, distance="cosine"), abs=1e-10)
    assert ClusteringEvaluator(distanceMeasure="cosine", weightCol="w").evaluate({"features": X, "prediction": lab, "w": w}) == \
        pytest.approx(oracle.silhouette(X, lab, w=w, distance="cosine"), abs=1e-10)
    X[7] = 0.0
    with pytest.raises(hkm.IllegalArgumentException, match="zero-length"):
        ev.evaluate({"features": X, "prediction": lab})
