"""GPU parity of the loader path: VectorAssembler (columnar -> row-major fp64, null/NaN handling) and
StandardScaler (moments, fused standardisation) against NumPy and the oracle's Summarizer restatement."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _hospital(hkm, n, d=8, seed=3, null_fraction=0.0):
    from hkm_b200 import datagen
    return datagen.hospital_columns(n, d=d, profiles=5, seed=seed, null_fraction=null_fraction)


def test_assembler_matches_numpy_stack(hkm):
    from hkm_b200.ml.feature import VectorAssembler
    for n in (1, 127, 128, 129, 10_000, 100_003):
        cols = _hospital(hkm, n)
        names = list(cols)
        t = VectorAssembler(inputCols=names, outputCol="features").transform(dict(cols))   # ref:L134-136 idiom
        assert t.columns == names + ["features"]
        X = t.to_numpy("features")
        ref = np.stack([np.asarray(v, dtype=np.float64) for v in cols.values()], axis=1)
        assert X.dtype == np.float64 and X.shape == (n, 8) and np.array_equal(X, ref)      # int32 -> fp64 is exact


def test_assembler_dtypes_and_vector_inputs(hkm):
    from hkm_b200.ml.feature import VectorAssembler
    rng = np.random.default_rng(0)
    n = 5000
    data = {"f64": rng.normal(size=n), "f32": rng.normal(size=n).astype(np.float32), "i32": rng.integers(-1000, 1000, n).astype(np.int32),
            "i64": rng.integers(-2**40, 2**40, n), "i16": rng.integers(-300, 300, n).astype(np.int16), "i8": rng.integers(-100, 100, n).astype(np.int8),
            "b": rng.random(n) < 0.5, "u8": rng.integers(0, 255, n).astype(np.uint8), "u32": rng.integers(0, 2**31, n).astype(np.uint32),
            "vec": rng.normal(size=(n, 3))}
    t = VectorAssembler(inputCols=list(data), outputCol="v").transform(data)
    ref = np.column_stack([np.asarray(data[k], dtype=np.float64) for k in data])
    assert np.array_equal(t.to_numpy("v"), ref)
    dev = {"a": torch.arange(n, dtype=torch.int32, device="cuda"), "b": torch.randn(n, dtype=torch.float64, device="cuda")}
    X = VectorAssembler(inputCols=["b", "a"], outputCol="v").transform(dev).to_numpy("v")
    assert np.array_equal(X[:, 1], np.arange(n)) and np.array_equal(X[:, 0], dev["b"].cpu().numpy())


def test_handle_invalid_error_skip_keep(hkm):
    import pyarrow as pa
    from hkm_b200.ml.feature import VectorAssembler
    cols = _hospital(hkm, 20_000, null_fraction=0.01)
    cols["length_of_stay"] = np.asarray(cols["length_of_stay"]).copy()
    cols["length_of_stay"][5] = np.nan                                       # NaN counts as invalid, like null
    names = list(cols)
    with pytest.raises(hkm.SparkException, match="handleInvalid"):
        VectorAssembler(inputCols=names, outputCol="features").transform(dict(cols))
    bad = np.zeros(20_000, dtype=bool)
    for v in cols.values():
        if isinstance(v, np.ma.MaskedArray):
            bad |= np.ma.getmaskarray(v)
    bad[5] = True
    ref = np.stack([np.ma.filled(np.ma.asarray(v).astype(np.float64), np.nan) for v in cols.values()], axis=1)
    t = VectorAssembler(inputCols=names, outputCol="features", handleInvalid="skip").transform(dict(cols, row=np.arange(20_000)))
    assert t.count() == int((~bad).sum()) and np.array_equal(t.to_numpy("features"), ref[~bad])
    assert np.array_equal(t.to_numpy("row"), np.arange(20_000)[~bad])        # the other columns are filtered alike
    k = VectorAssembler(inputCols=names, outputCol="features", handleInvalid="keep").transform(dict(cols)).to_numpy("features")
    assert np.array_equal(np.isnan(k).any(1), bad) and np.array_equal(k[~bad], ref[~bad])
    # the reference's own flow: na.drop() first (ref:L128), then the default handleInvalid='error'
    arrow = pa.table({nm: pa.array(np.ma.filled(np.ma.asarray(v), 0), mask=np.ma.getmaskarray(np.ma.asarray(v))) for nm, v in cols.items()})
    clean = hkm.Table.wrap(arrow).na.drop()
    t2 = VectorAssembler(inputCols=names, outputCol="features").transform(clean)
    assert np.array_equal(t2.to_numpy("features"), ref[~bad])


@pytest.mark.parametrize("with_mean,with_std", [(False, True), (True, True), (True, False)])
def test_standard_scaler_fused_and_unfused(hkm, oracle, with_mean, with_std):
    from hkm_b200.ml.feature import StandardScaler, VectorAssembler
    cols = _hospital(hkm, 50_000, d=12, seed=9)
    cols["const"] = np.full(50_000, 7.0)                                     # zero variance -> scaled to 0.0
    names = list(cols)
    ref = np.stack([np.asarray(v, dtype=np.float64) for v in cols.values()], axis=1)
    omean, ostd = oracle.scaler_fit(ref, P=8)
    assembled = VectorAssembler(inputCols=names, outputCol="features").transform(dict(cols))
    sc = StandardScaler(inputCol="features", outputCol="scaled", withMean=with_mean, withStd=with_std)
    model = sc.fit(assembled)                                                # lazy column: moments straight from the typed columns
    assert np.allclose(model.mean, ref.mean(0), rtol=1e-12, atol=1e-12) and np.allclose(model.mean, omean, rtol=1e-12, atol=1e-12)
    assert np.allclose(model.std[:-1], ref[:, :-1].std(0, ddof=1), rtol=1e-12) and model.std[-1] == 0.0
    assert np.allclose(model.std[:-1], ostd[:-1], rtol=1e-11)
    fused = model.transform(assembled)
    assert isinstance(fused["scaled"], hkm.AssembledColumn)                  # nothing written yet: one fused launch on demand
    Y = fused.to_numpy("scaled")
    Yo = oracle.scaler_transform(ref, model.mean, model.std, with_mean, with_std)
    assert np.allclose(Y, Yo, rtol=1e-14, atol=1e-14) and np.all(Y[:, -1] == 0.0) == (with_std or with_mean)
    # same answer from an already materialised matrix (moments over rows + hkm_scale_rows)
    dense = {"features": ref}
    m2 = sc.fit(dense)
    assert np.allclose(m2.mean, model.mean, rtol=1e-13, atol=1e-13) and np.allclose(m2.std, model.std, rtol=1e-12, atol=1e-15)
    Yo2 = oracle.scaler_transform(ref, m2.mean, m2.std, with_mean, with_std)
    assert np.allclose(m2.transform(dense).to_numpy("scaled"), Yo2, rtol=1e-14, atol=1e-14)


def test_pyspark_feature_doctests(hkm):
    """pyspark.ml.feature doctests (tests/golden/make_golden.py cites the API-doc pages): VectorAssembler (1, 0, 3) -> [1.0, 0.0, 3.0];
    StandardScaler rows [0.0], [2.0] -> mean 1.0, std 1.4142, scaled [0.0], [1.4142]."""
    import json, os
    from hkm_b200.ml.feature import StandardScaler, VectorAssembler
    kat = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat.json")))
    va = kat["vector_assembler_doctest"]
    t = VectorAssembler(inputCols=["a", "b", "c"], outputCol="features").transform(
        {"a": np.array(va["a"], dtype=np.int64), "b": np.array(va["b"], dtype=np.int64), "c": np.array(va["c"], dtype=np.int64)})
    assert np.array_equal(t.to_numpy("features"), np.array(va["features"]))
    ss = kat["standard_scaler_doctest"]
    data = {"a": np.array(ss["rows"], dtype=np.float64)}
    model = StandardScaler(inputCol="a", outputCol="scaled").fit(data)
    assert model.mean == pytest.approx(ss["mean"], abs=1e-15) and model.std == pytest.approx(ss["std"], abs=ss["print_tol"])
    assert model.std[0] == pytest.approx(np.sqrt(2.0), rel=1e-15)
    Y = model.transform(data).to_numpy("scaled")
    assert Y == pytest.approx(np.array(ss["scaled"]), abs=ss["print_tol"]) and Y[0, 0] == 0.0


def test_pipeline_assemble_scale_fit_evaluate(hkm, oracle):
    """The whole path on the hospital table: VectorAssembler -> StandardScaler -> KMeans -> ClusteringEvaluator."""
    from hkm_b200.ml.clustering import KMeans
    from hkm_b200.ml.evaluation import ClusteringEvaluator
    from hkm_b200.ml.feature import StandardScaler, VectorAssembler
    cols = _hospital(hkm, 10_000, d=8, seed=20250331)
    raw = np.stack([np.asarray(v, dtype=np.float64) for v in cols.values()], axis=1)
    t = VectorAssembler(inputCols=list(cols), outputCol="raw").transform(dict(cols))
    t = StandardScaler(inputCol="raw", outputCol="features").fit(t).transform(t)
    model = KMeans(k=5, seed=42).fit(t)
    sil = ClusteringEvaluator().evaluate(model.transform(t))
    mean, std = oracle.scaler_fit(raw, P=8)
    Xo = oracle.scaler_transform(raw, mean, std)
    ref = oracle.kmeans_fit(Xo, 5, 42)
    assert np.allclose(model.clusterCenters(), ref["centers"], rtol=1e-6, atol=1e-9)
    assert model.summary.trainingCost == pytest.approx(ref["training_cost"], rel=1e-6)
    assert sil == pytest.approx(oracle.silhouette(Xo, ref["labels"]), abs=1e-5)
