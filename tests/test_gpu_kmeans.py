"""GPU parity of the pyspark.ml-shaped KMeans estimator against the CPU oracle (same inputs, seed
and partition layout): initialisation (random, k-means||), Lloyd loop, summary, transform."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

REL = 1e-6           # north_star: centroids and trainingCost within 1e-6 relative
LABEL_AGREE = 0.9999  # north_star: assignments >= 99.99 % identical


def _fit_both(hkm, oracle, X, k, seed, init_mode="k-means||", P=8, max_iter=20, tol=1e-4, w=None, init_steps=2):
    from hkm_b200.ml.clustering import KMeans
    km = KMeans(k=k, seed=seed, initMode=init_mode, maxIter=max_iter, tol=tol, initSteps=init_steps).setNumPartitions(P)
    data = {"features": X}
    if w is not None:
        data["w"] = w
        km.setWeightCol("w")
    model = km.fit(data)
    ref = oracle.kmeans_fit(X, k, seed, max_iter=max_iter, tol=tol, init_mode=init_mode, init_steps=init_steps, P=P, w=w)
    return model, ref


def _assert_model_matches(model, ref):
    C = np.array(model.clusterCenters())
    assert C.shape == ref["centers"].shape
    assert np.allclose(C, ref["centers"], rtol=REL, atol=1e-9)
    s = model.summary
    assert s.numIter == ref["num_iter"]
    assert s.trainingCost == pytest.approx(ref["training_cost"], rel=REL, abs=1e-12)
    pred = s.predictions.to_numpy("prediction")
    assert pred.dtype == np.int32 and (pred == ref["labels"]).mean() >= LABEL_AGREE
    if (pred == ref["labels"]).all():
        assert s.clusterSizes == ref["cluster_sizes"].tolist()


@pytest.mark.parametrize("init_mode", ["k-means||", "random"])
def test_cfg1_hospital_table(hkm, oracle, init_mode):
    """BASELINE configs[0]: k=5 on the synthetic hospital table, 10k rows x 8 features."""
    from hkm_b200 import datagen
    X = datagen.hospital_matrix(10_000, d=8, profiles=5, seed=20250331)
    model, ref = _fit_both(hkm, oracle, X, k=5, seed=42, init_mode=init_mode)
    _assert_model_matches(model, ref)
    assert model.summary.k == 5 and model.hasSummary and sum(model.summary.clusterSizes) == 10_000


@pytest.mark.parametrize("n,d,k,P,seed", [(20000, 4, 6, 8, 42), (50000, 16, 16, 8, 7), (30011, 24, 3, 5, 1), (8000, 32, 8, 1, 123), (40000, 8, 2, 16, -9),
                                          # k > 16: the warp-specialised, gather and tcgen05 variants behind the fused driver step
                                          (60000, 24, 32, 8, 5), (50000, 32, 20, 4, 11), (40000, 64, 40, 8, 3), (30000, 16, 48, 8, 2),
                                          # d = 28..30 with k near 20: each k-means|| step scores chunks of ~2k new centres, whose pass runs
                                          # more CTAs per SM than a k = 64 chunk -- the k-means|| scratch must cover that
                                          (30000, 28, 20, 8, 4), (20000, 30, 16, 4, 9), (20000, 29, 18, 8, 6)])
def test_fit_matches_oracle(hkm, oracle, n, d, k, P, seed):
    rng = np.random.default_rng(abs(seed) + n)
    X = rng.normal(size=(n, d)) + rng.integers(0, k, size=(n, 1)) * 3.0
    model, ref = _fit_both(hkm, oracle, X, k, seed, P=P)
    _assert_model_matches(model, ref)


def test_kmeans_parallel_candidates_match_oracle(hkm, oracle):
    """The k-means|| sampling is partition-exact: same chosen rows, same candidate weights."""
    from hkm_b200.ml._kmeans_init import PartitionLayout, init_kmeans_parallel
    rng = np.random.default_rng(21)
    for n, d, k, P, seed in [(20000, 4, 6, 8, 42), (100_000, 8, 20, 8, 5), (5000, 3, 40, 3, 77), (300_000, 6, 4, 16, 2024)]:
        X = rng.normal(size=(n, d)) + rng.integers(0, max(k // 2, 2), size=(n, 1)) * 4.0
        dbg = {}
        C = init_kmeans_parallel(torch.from_numpy(X).cuda(), PartitionLayout.build([n], 0, P), k, 2, seed, debug=dbg)
        Co, odbg = oracle.init_kmeans_parallel(X, k, seed, 2, P, debug=True)
        assert dbg["first_idx"] == int(odbg["chosen_idx"][0])
        assert dbg["step_counts"] == odbg["step_counts"].tolist()
        assert np.array_equal(dbg["candidates"], odbg["candidates"])
        assert np.allclose(C, Co, rtol=1e-12, atol=1e-12)


def test_weighted_fit_and_pyspark_doctest(hkm, oracle):
    from hkm_b200.ml.clustering import KMeans
    pts = np.array([[0.0, 0.0], [1.0, 1.0], [9.0, 8.0], [8.0, 9.0]])
    model = KMeans(k=2, seed=1, maxIter=10, weightCol="weighCol").fit({"features": pts, "weighCol": np.full(4, 2.0)})
    centers = sorted(map(list, model.clusterCenters()))
    assert np.allclose(centers, [[0.5, 0.5], [8.5, 8.5]])
    assert model.summary.trainingCost == pytest.approx(4.0, abs=1e-12) and model.summary.clusterSizes == [2, 2]
    t = model.transform({"features": pts})
    p = t.to_numpy("prediction")
    assert p[0] == p[1] and p[2] == p[3] and p[0] != p[2]
    assert model.predict(pts[0]) == p[0] and model.predict(np.array([8.6, 8.4])) == p[2]
    rng = np.random.default_rng(3)
    X = rng.normal(size=(20000, 5)) + rng.integers(0, 3, size=(20000, 1)) * 3; w = rng.uniform(0.1, 4, size=20000)
    m, ref = _fit_both(hkm, oracle, X, 3, 11, w=w)
    _assert_model_matches(m, ref)


def test_single_cluster_kat_and_few_distinct_points(hkm, oracle):
    from hkm_b200.ml.clustering import KMeans
    pts = np.array([[1.0, 2, 6], [1, 3, 0], [1, 4, 6]])
    for mode in ("random", "k-means||"):
        km = KMeans(k=2, seed=1, initMode=mode, maxIter=5).setInitialCenters(pts[:1])  # k=1 is not a legal Param; one initial centre is
        m = km.fit({"features": pts})
        assert np.allclose(m.clusterCenters(), [[1, 3, 4]])
    Xd = np.repeat(np.array([[0.0, 0], [1, 1], [2, 2]]), 50, axis=0)
    m, ref = _fit_both(hkm, oracle, Xd, 5, 3, P=2)
    assert len(m.clusterCenters()) == ref["centers"].shape[0] <= 3 and m.summary.trainingCost == pytest.approx(0.0, abs=1e-18)


def test_max_iter_zero_and_tol_zero(hkm, oracle):
    rng = np.random.default_rng(0)
    X = rng.normal(size=(5000, 6)) + rng.integers(0, 4, size=(5000, 1)) * 3
    m0, r0 = _fit_both(hkm, oracle, X, 4, 42, max_iter=0)
    assert m0.summary.numIter == 0 and m0.summary.trainingCost == 0.0 and np.allclose(m0.clusterCenters(), r0["init_centers"])
    m, r = _fit_both(hkm, oracle, X, 4, 42, max_iter=20, tol=0.0)
    _assert_model_matches(m, r)


def test_transform_save_load_and_errors(hkm, oracle, tmp_path):
    from hkm_b200.ml.clustering import KMeans, KMeansModel
    rng = np.random.default_rng(2)
    X = rng.normal(size=(3000, 4)) + rng.integers(0, 3, size=(3000, 1)) * 5
    model = KMeans(k=3, seed=5, featuresCol="f", predictionCol="cluster").fit({"f": X, "id": np.arange(3000)})
    out = model.transform({"f": X[:100], "id": np.arange(100)})
    assert out.columns == ["f", "id", "cluster"] and np.array_equal(out.to_numpy("cluster"), oracle.predict(X[:100], np.array(model.clusterCenters())))
    path = str(tmp_path / "model")
    model.write().overwrite().save(path)            # ref:L241
    with pytest.raises(IOError):
        model.save(path)
    m2 = KMeansModel.load(path)
    assert np.array_equal(m2.clusterCenters(), model.clusterCenters()) and m2.getPredictionCol() == "cluster" and not m2.hasSummary
    assert np.array_equal(m2.transform({"f": X[:100]}).to_numpy("cluster"), out.to_numpy("cluster"))
    with pytest.raises(hkm.IllegalArgumentException):
        model.transform({"f": X[:, :3]})
    with pytest.raises(hkm.IllegalArgumentException):
        KMeans(k=2).fit({"features": np.zeros((0, 3))})
    with pytest.raises(KeyError):
        KMeans(k=2).fit({"feats": X})


def test_cfg2_fit_end_to_end(hkm):
    """BASELINE configs[1] at full size through the public API: 10 M x 32, k = 8, maxIter = 20."""
    from hkm_b200 import datagen
    from hkm_b200.ml.clustering import KMeans
    X = datagen.hospital_matrix_device(10_000_000, 32, 8, 20250331, torch.device("cuda"))
    t = hkm.Table({"features": hkm.DeviceMatrix(X)})
    model = KMeans(k=8, seed=42, maxIter=20, tol=1e-4).fit(t)
    s = model.summary
    assert 1 <= s.numIter <= 20 and sum(s.clusterSizes) == 10_000_000 and len(model.clusterCenters()) == 8
    # idempotence: refitting from the fitted centres converges at once to the same model and cost
    m2 = KMeans(k=8, seed=42, maxIter=20, tol=1e-4).setInitialCenters(np.array(model.clusterCenters())).fit(t)
    if s.numIter < 20:
        assert m2.summary.numIter == 1
        assert np.allclose(m2.clusterCenters(), model.clusterCenters(), rtol=1e-6, atol=1e-6)
    # the cost of a fixed point equals the sum of squared distances to the assigned centres
    C = torch.from_numpy(np.array(m2.clusterCenters())).cuda()
    lab = m2.summary.predictions["prediction"].values.long()
    direct = 0.0
    for lo in range(0, 10_000_000, 1_000_000):
        direct += float(((X[lo:lo + 1_000_000] - C[lab[lo:lo + 1_000_000]]) ** 2).sum())
    assert direct <= m2.summary.trainingCost * (1 + 1e-6)


@pytest.mark.parametrize("init_mode", ["random", "k-means||"])
def test_cosine_distance_measure(hkm, oracle, init_mode):
    """distanceMeasure='cosine' (SURVEY 8f-4): unit rows + unit centres through the same kernels == CosineDistanceMeasure."""
    from hkm_b200.ml.clustering import KMeans
    rng = np.random.default_rng(5)
    n, d, k, P = 40_000, 12, 6, 8
    dirs = rng.normal(size=(k, d))
    X = dirs[rng.integers(0, k, size=n)] * rng.uniform(0.5, 20.0, size=(n, 1)) + 0.3 * rng.normal(size=(n, d))
    w = rng.uniform(0.5, 2.0, size=n) if init_mode == "random" else None
    km = KMeans(k=k, seed=11, initMode=init_mode, distanceMeasure="cosine", maxIter=20, tol=1e-5).setNumPartitions(P)
    t = {"features": X}
    if w is not None:
        km.setWeightCol("w")
        t["w"] = w
    model = km.fit(t)
    ref = oracle.kmeans_fit(X, k, 11, tol=1e-5, init_mode=init_mode, P=P, w=w, distance="cosine")
    C = np.array(model.clusterCenters())
    assert model.summary.numIter == ref["num_iter"]
    assert np.abs(C - ref["centers"]).max() < 1e-9 and np.allclose(np.linalg.norm(C, axis=1), 1.0, atol=1e-14)
    assert model.summary.trainingCost == pytest.approx(ref["training_cost"], rel=1e-8)
    pred = model.transform({"features": X}).to_numpy("prediction")
    assert (pred == ref["labels"]).mean() >= 0.9999
    assert model.summary.clusterSizes == np.bincount(pred, minlength=k).tolist()
    assert model.predict(X[3]) == ref["labels"][3]
    # ml KMeansSuite "KMeans with cosine distance"
    S = np.array([[1.0, 1.0], [10.0, 10.0], [1.0, 0.5], [10.0, 4.4], [-1.0, 1.0], [-100.0, 90.0]])
    m2 = KMeans(k=3, seed=42, initMode="random", tol=1e-6, distanceMeasure="cosine").setNumPartitions(2).fit({"features": S})
    p2 = m2.transform({"features": S}).to_numpy("prediction")
    assert len(set(p2.tolist())) == 3 and p2[0] == p2[1] and p2[2] == p2[3] and p2[4] == p2[5]
    S[2] = 0.0
    with pytest.raises(hkm.IllegalArgumentException, match="zero-length"):
        KMeans(k=2, distanceMeasure="cosine").fit({"features": S})


def test_fit_and_evaluate_beyond_2_31_elements(hkm):
    """One whole fit (k-means||, 3 Lloyd steps, summary pass) + silhouette on a 70 M x 32 device-resident table
    (2.24e9 elements, 17.9 GB): every kernel on the path addresses rows in 64 bits."""
    from hkm_b200 import datagen
    from hkm_b200.table import Table, DeviceMatrix
    from hkm_b200.ml.clustering import KMeans
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    n, d, k = 70_000_000, 32, 6
    X = datagen.hospital_matrix_device(n, d, k, 20250331, torch.device("cuda"))
    t = Table({"features": DeviceMatrix(X)})
    model = KMeans(k=k, seed=3, maxIter=3).fit(t)
    s = model.summary
    assert sum(s.clusterSizes) == n and s.numIter >= 1 and s.trainingCost > 0
    pred = s.predictions["prediction"].values
    assert int(pred.min().item()) >= 0 and int(pred.max().item()) < k
    assert torch.equal(torch.bincount(pred, minlength=k).cpu(), torch.tensor(s.clusterSizes))
    sil = ClusteringEvaluator().evaluate(s.predictions)
    assert -1.0 <= sil <= 1.0


@pytest.mark.parametrize("n,d,k", [(30000, 16, 100), (20000, 8, 256), (5000, 24, 65),
                                   # wide rows and (d, k) pairs whose tile ring does not fit: the general-shape kernels
                                   (20000, 128, 8), (6000, 200, 70), (8000, 96, 64), (5000, 94, 25), (3000, 1000, 5)])
def test_more_than_64_clusters(hkm, oracle, n, d, k):
    """KMeans(k=100) and beyond, d = 128 and beyond (the pyspark estimator caps neither): the data pass runs in chunks of centres,
    through the tiled kernels where the chunk fits them and through the general-shape kernels (hkm_generic.cu) where not."""
    from hkm_b200.ml.clustering import KMeans
    rng = np.random.default_rng(n + d + k)
    prof = rng.normal(size=(k, d)) * 3.0
    X = prof[rng.integers(0, k, size=n)] + rng.normal(size=(n, d)) * 0.5
    N = hkm._native
    C = X[rng.choice(n, k, replace=False)].copy()
    Xd = torch.from_numpy(X).cuda(); Cd = torch.from_numpy(C).cuda()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda"); mind2 = torch.empty(n, dtype=torch.float64, device="cuda")
    N.assign_reduce(Xd, Cd, partial, labels, mind2, scratch=N.new_scratch(d, k, "cuda"))
    sums, wsum, cost, lab = oracle.lloyd_pass(X, C, P=1, want_labels=True)
    p = partial.cpu().numpy()
    assert np.array_equal(labels.cpu().numpy(), lab) and np.array_equal(p[k * d:k * d + k], wsum)
    assert np.abs(p[:k * d].reshape(k, d) - sums).max() <= 1e-11 * max(1.0, np.abs(sums).max())
    assert p[-1] == pytest.approx(cost, rel=1e-9)
    pred = torch.empty(n, dtype=torch.int32, device="cuda")
    N.assign_reduce(Xd, Cd, None, pred, do_reduce=False)
    assert np.array_equal(pred.cpu().numpy(), lab)
    # the estimator end to end: given initial centres, and the default k-means|| initialiser
    model = KMeans(k=k, maxIter=5, tol=0.0, seed=3).setInitialCenters(C).fit({"features": X})
    Co, it, cost5 = oracle.lloyd(X, C, max_iter=5, tol=0.0, P=1)
    assert model.summary.numIter == it and np.allclose(model.clusterCenters(), Co, rtol=1e-6, atol=1e-9)
    assert model.summary.trainingCost == pytest.approx(cost5, rel=1e-6) and sum(model.summary.clusterSizes) == n
    m2, ref = _fit_both(hkm, oracle, X, k, 11, P=4, max_iter=3)
    _assert_model_matches(m2, ref)
    assert model.predict(X[0]) == int(model.summary.predictions.to_numpy("prediction")[0])


@pytest.mark.gpu
def test_label_counts_matches_bincount(hkm):
    """hkm_label_counts (cluster sizes of a summary from the labels of a transform pass): every label in [0, m) counted once,
    labels outside the range ignored, empty input -> zeros."""
    import torch
    from hkm_b200 import _native as N
    rng = np.random.default_rng(5)
    for n, m in ((0, 3), (1, 1), (1000, 7), (300_001, 64), (50_000, 1000)):
        lab = rng.integers(-2, m + 2, size=n).astype(np.int32)
        got = N.label_counts(torch.from_numpy(lab).cuda(), m).cpu().numpy()
        ok = lab[(lab >= 0) & (lab < m)]
        assert np.array_equal(got, np.bincount(ok, minlength=m).astype(np.int64))
