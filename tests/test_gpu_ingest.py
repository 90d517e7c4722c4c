"""GPU test of the ingest step right before the hot path: CSV with the hospital schema (ref:L64-78) -> training
time window (ref:L123-127) -> na.drop (ref:L128) -> VectorAssembler (ref:L134-136) -> KMeans.  The window and the null masks
are applied on the device, inside the loader's compaction pass (hkm_assemble_filtered); the oracle sees the NumPy-filtered rows."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

FEATURES = ["admission_count", "current_occupancy", "emergency_visits", "seasonality_index"]   # ref:L134


def _write_csv(path, n, seed):
    rng = np.random.default_rng(seed)
    t0 = np.datetime64("2025-03-31T20:00:00", "s")
    ts = t0 + rng.integers(0, 5 * 3600, size=n).astype("timedelta64[s]")
    prof = rng.integers(0, 4, size=n)
    adm = rng.poisson(10 + 15 * prof); occ = rng.poisson(100 + 200 * prof); ed = rng.poisson(3 + 4 * prof)
    season = 1 + 0.1 * prof + rng.normal(0, 0.02, size=n); los = rng.gamma(2.0, 1.0 + prof)
    null_at = rng.random((n, 7)) < 0.01
    with open(path, "w") as f:
        f.write("hospital_id,event_time,admission_count,current_occupancy,emergency_visits,seasonality_index,length_of_stay\n")
        for i in range(n):
            vals = [f"h{prof[i]}", str(ts[i]).replace("T", " "), str(adm[i]), str(occ[i]), str(ed[i]), f"{season[i]:.6f}", f"{los[i]:.6f}"]
            f.write(",".join("" if null_at[i, c] else v for c, v in enumerate(vals)) + "\n")


def test_csv_window_nadrop_assemble_fit(hkm, oracle, tmp_path):
    from hkm_b200 import io
    from hkm_b200.ml.clustering import KMeans
    from hkm_b200.ml.feature import StandardScaler, VectorAssembler
    p = tmp_path / "uploads"
    p.mkdir()
    _write_csv(p / "a.csv", 6000, 1); _write_csv(p / "b.csv", 5000, 2)
    lo, hi = "2025-03-31 22:00:00", "2025-03-31 23:59:59"
    t = io.read_csv(str(p))
    assert t.count() == 11000
    # host path (NumPy masks): the rows the oracle sees
    host = io.between(t, "event_time", lo, hi).na.drop()
    Xh = np.column_stack([np.asarray(host.to_numpy(c), dtype=np.float64) for c in FEATURES])
    assert 0 < host.count() < 11000
    # device path: the window and the null masks go into the loader's compaction pass
    lazy = io.between(t, "event_time", lo, hi, lazy=True).na.drop()
    assert lazy._pending is not None
    out = VectorAssembler(inputCols=FEATURES, outputCol="raw").transform(lazy)
    assert out._pending is None and isinstance(out["raw"], hkm.DeviceMatrix)
    X = out.to_numpy("raw")
    assert X.shape == Xh.shape and np.array_equal(X, Xh)
    # the table's other columns were filtered by the same keep mask
    assert out.to_numpy("hospital_id").tolist() == host.to_numpy("hospital_id").tolist()
    assert np.array_equal(np.asarray(out.to_numpy("length_of_stay")), np.asarray(host.to_numpy("length_of_stay")))
    # ... and the hot path on top of it: scale, fit, against the oracle on the NumPy-filtered rows
    t2 = StandardScaler(inputCol="raw", outputCol="features").fit(out).transform(out)
    model = KMeans(k=4, seed=42, maxIter=20).setNumPartitions(4).fit(t2)
    mean, std = oracle.scaler_fit(Xh, P=4)
    ref = oracle.kmeans_fit(oracle.scaler_transform(Xh, mean, std), 4, 42, max_iter=20, P=4)
    assert np.allclose(model.clusterCenters(), ref["centers"], rtol=1e-6, atol=1e-9)
    assert model.summary.trainingCost == pytest.approx(ref["training_cost"], rel=1e-6)
    assert (model.summary.predictions.to_numpy("prediction") == ref["labels"]).mean() >= 0.9999


def test_filtered_loader_counts_and_edges(hkm):
    """hkm_assemble_filtered through the binding: counts (outside the window / null inside it), null timestamps never match,
    inclusive bounds, extra validity masks, no window at all, empty result."""
    N = hkm._native
    n = 1000
    rng = np.random.default_rng(0)
    a = torch.from_numpy(rng.normal(size=n)).cuda(); b = torch.arange(n, dtype=torch.int32, device="cuda")
    va = torch.ones(n, dtype=torch.uint8, device="cuda"); va[::50] = 0                      # 20 nulls in column a
    ts = torch.arange(n, dtype=torch.int64, device="cuda") * 10
    tsv = torch.ones(n, dtype=torch.uint8, device="cuda"); tsv[105] = 0                     # a null timestamp inside the window
    extra = torch.ones(n, dtype=torch.uint8, device="cuda"); extra[107] = 0                 # a null in a non-feature column
    X, keep, outside, bad = N.assemble_filtered([a, b], [va, None], n, window=(ts, tsv, 1000, 1990), extra_valid=[extra])
    k = keep.cpu().numpy().astype(bool)
    exp = np.zeros(n, dtype=bool); exp[100:200] = True; exp[105] = False; exp[107] = False; exp[100:200:50] = False
    assert np.array_equal(k, exp) and X.shape[0] == exp.sum()
    assert outside == n - 100 + 1 and bad == 3                                             # rows 100, 150 (column a) and 107 (extra mask)
    assert np.array_equal(X[:, 1].cpu().numpy(), np.flatnonzero(exp).astype(np.float64))
    X2, keep2, outside2, bad2 = N.assemble_filtered([a, b], [va, None], n)                   # no window: plain na.drop on the features
    assert outside2 == 0 and bad2 == 20 and X2.shape[0] == n - 20
    X3, keep3, outside3, bad3 = N.assemble_filtered([a, b], [va, None], n, window=(ts, None, 5, 6))
    assert X3.shape[0] == 0 and outside3 == n and bad3 == 0 and not keep3.any()
