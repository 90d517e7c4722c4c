"""CPU-only checks of the host side: the pyspark.ml-shaped Param surface, the driver-side Spark
RNG/sampling restatement shipped in the product (cross-checked against the independent C oracle),
the Table stand-in, and that libhkm.so loads and exports every symbol include/hkm.h declares."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KAT = json.load(open(os.path.join(ROOT, "tests", "golden", "kat.json")))


def test_library_exports_every_declared_symbol(hkm):
    hdr = open(os.path.join(ROOT, "include", "hkm.h")).read()
    declared = set(re.findall(r"HKM_API\s+[\w\s\*]+?\b(hkm_\w+)\s*\(", hdr))
    assert len(declared) >= 10
    so = hkm._native.SO_PATH
    assert os.path.exists(so), "build the library first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(so)
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, f"declared in hkm.h but not exported: {missing}"
    assert set(hkm._native.EXPORTS) <= declared
    lib.hkm_last_error.restype = ctypes.c_char_p
    assert lib.hkm_version() >= 100 and lib.hkm_max_k() >= 32 and isinstance(lib.hkm_last_error(), bytes)


def test_no_cpu_fallback(hkm):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from hkm_b200.ml.clustering import KMeans
    with pytest.raises(hkm.HkmError, match="no CPU fallback"):
        KMeans(k=2, seed=1).fit({"features": np.zeros((10, 2))})


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "clustermachinelearningforhospitalnetworks-apache-spark_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "hkm_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_kmeans_params_surface(hkm):
    from hkm_b200.ml.clustering import KMeans
    km = KMeans()
    assert (km.getK(), km.getInitMode(), km.getInitSteps(), km.getTol(), km.getMaxIter()) == (2, "k-means||", 2, 1e-4, 20)
    assert km.getFeaturesCol() == "features" and km.getPredictionCol() == "prediction" and km.getDistanceMeasure() == "euclidean"
    assert km.getSeed() == KAT["kmeans_default_seed"] and km.getSolver() == "auto" and km.getMaxBlockSizeInMB() == 0.0
    assert not km.isDefined("weightCol")
    km2 = KMeans(featuresCol="f", predictionCol="p", k=5, initMode="random", initSteps=3, tol=0.0, maxIter=7, seed=42, weightCol="w")
    assert (km2.getK(), km2.getInitMode(), km2.getSeed(), km2.getWeightCol(), km2.getMaxIter()) == (5, "random", 42, "w", 7)
    assert km.setK(3).setSeed(7).setMaxIter(0).setTol(0.5) is km and km.getK() == 3
    c = km.copy({"k": 9})
    assert c.getK() == 9 and km.getK() == 3
    assert "k: The number of clusters to create" in km.explainParams()
    for bad in (dict(k=1), dict(k=0), dict(initMode="kmeans++"), dict(initSteps=0), dict(maxIter=-1), dict(tol=-1e-3),
                dict(distanceMeasure="manhattan"), dict(solver="gpu")):
        with pytest.raises(hkm.IllegalArgumentException):
            KMeans(**bad)
    with pytest.raises(TypeError):
        KMeans(k="three")
    with pytest.raises(TypeError):
        KMeans(5)  # keyword-only, as pyspark's @keyword_only


def test_sparkrand_matches_oracle(hkm, oracle):
    SR = hkm._sparkrand
    r = SR.XORShiftRandom(42)
    assert [r.next_double() * 3, r.next_double() * 3] == KAT["xorshift"]["seed42_rand_times3"]
    assert SR.JavaRandom(0).next_int() == KAT["java_random"]["seed0_nextInt"]
    j = SR.JavaRandom(42)
    assert [j.next_int(10) for _ in range(5)] == KAT["java_random"]["seed42_nextInt10_x5"]
    for seed in (0, 1, -1, 42, 2**40 + 3, -(2**62)):
        assert SR.hash_seed(seed) == oracle.hash_seed(seed)
        a, b = SR.XORShiftRandom(seed), oracle.XORShiftRandom(seed)
        assert [a.next_int() for _ in range(4)] == [b.next_int() for _ in range(4)]
        a, b = SR.JavaRandom(seed), oracle.JavaRandom(seed)
        assert [a.next_long() for _ in range(3)] == [b.next_long() for _ in range(3)]
        assert [a.next_int(7) for _ in range(20)] == [b.next_int(7) for _ in range(20)]
    for n, P, num, seed in [(1000, 4, 3, 7), (5, 2, 9, 7), (100000, 8, 1, -123456), (37, 3, 5, 99), (10, 1, 4, 5), (3, 2, 2, 8), (2_000_000, 8, 64, 11)]:
        assert SR.take_sample(SR.equal_partition_sizes(n, P), num, seed) == list(oracle.take_sample(n, P, num, seed))


def test_local_kmeanspp_matches_oracle(hkm, oracle):
    rng = np.random.default_rng(0)
    for m, d, k in [(40, 3, 5), (257, 8, 64), (9, 2, 8)]:
        pts = rng.normal(size=(m, d)) * 3
        w = rng.integers(0, 10, size=m).astype(float)
        w[0] = max(w[0], 1.0)
        assert np.allclose(hkm._sparkrand.local_kmeans_pp(0, pts, w, k), oracle.local_kmeanspp(0, pts, w, k), atol=1e-12)
        # the product's driver step (host C++ inside libhkm; runs without a GPU) -- same summation order as Vectors.sqdist
        assert np.array_equal(hkm._native.local_kmeanspp(0, pts, w, k), oracle.local_kmeanspp(0, pts, w, k))


def test_xorshift_jump_table(hkm):
    SR = hkm._sparkrand
    J = SR.xorshift_jump_table(20)

    def apply(j, x):
        y = 0
        for i in range(64):
            if (x >> i) & 1:
                y ^= int(J[j, i])
        return y

    for seed, steps in [(3, 5), (42, 1000), (-7, 123457)]:
        r = SR.XORShiftRandom(seed)
        x = r.s
        for _ in range(steps):
            r.next(32)
        for j in range(20):
            if (steps >> j) & 1:
                x = apply(j, x)
        assert x == r.s


def test_distinct_rows_keeps_first_occurrence(hkm):
    V = np.array([[1.0, 2], [3, 4], [1, 2], [0.0, -0.0], [-0.0, 0.0], [3, 4]])
    out = hkm._sparkrand.distinct_rows(V)
    assert out.tolist() == [[1, 2], [3, 4], [0, 0]]


def test_partition_layout(hkm):
    from hkm_b200.ml._kmeans_init import PartitionLayout
    lay = PartitionLayout.build([10], 0, 4)
    assert lay.part_sizes == [2, 3, 2, 3] and lay.part_index0 == 0 and lay.n_global == 10
    lay = PartitionLayout.build([100, 60], 1, 8)
    assert lay.parts_per_rank == [4, 4] and lay.part_index0 == 4 and lay.row_offset == 100
    assert lay.local_part_sizes == [15, 15, 15, 15] and sum(lay.part_sizes) == 160


def test_table_surface(hkm):
    import pyarrow as pa
    T = hkm.Table
    t = T.wrap(pa.table({"a": pa.array([1, 2, None, 4], type=pa.int32()), "b": [0.5, float("nan"), 1.5, 2.5], "id": ["x", "y", "z", "w"]}))
    assert t.columns == ["a", "b", "id"] and t.count() == 4
    clean = t.na.drop()                          # ref:L128
    assert clean.count() == 2 and clean.to_numpy("a").tolist() == [1, 4] and clean.to_numpy("id").tolist() == ["x", "w"]
    assert t.na.drop(subset=["a"]).count() == 3
    s = t.select("a", "b")
    assert s.columns == ["a", "b"]
    t2 = t.withColumn("features", np.arange(8.0).reshape(4, 2)).drop("id")
    assert t2.columns == ["a", "b", "features"] and t2.to_numpy("features").shape == (4, 2)
    with pytest.raises(KeyError):
        t["nope"]
    with pytest.raises(ValueError):
        T({"a": [1, 2], "b": [1, 2, 3]})
    import pandas as pd
    tp = T.wrap(pd.DataFrame({"a": [1.0, None, 3.0], "b": [1, 2, 3]}))
    assert tp.na.drop().count() == 2
    assert list(t.toPandas().columns) == ["a", "b", "id"]


def test_datagen_schema(hkm):
    from hkm_b200 import datagen
    cols = datagen.hospital_columns(1000, d=8, profiles=5, seed=1, null_fraction=0.01)
    assert list(cols)[:5] == datagen.BASE_COLUMNS and len(cols) == 8
    assert np.asarray(cols["admission_count"]).dtype == np.int32 and np.asarray(cols["length_of_stay"]).dtype == np.float64
    assert any(isinstance(v, np.ma.MaskedArray) for v in cols.values())
    X = datagen.hospital_matrix(2000, d=8, profiles=5, seed=1)
    assert X.shape == (2000, 8) and np.allclose(X.std(0, ddof=1), 1.0)
    assert np.array_equal(X, datagen.hospital_matrix(2000, d=8, profiles=5, seed=1))


def test_ingest_csv_window_and_na_drop(hkm, tmp_path):
    """The steps right before the hot path in the reference: CSV with the hospital schema (ref:L64-78), the
    training time window (ref:L123-127) and na.drop (ref:L128)."""
    from hkm_b200 import io
    p = tmp_path / "hospitals"
    p.mkdir()
    (p / "a.csv").write_text(
        "hospital_id,event_time,admission_count,current_occupancy,emergency_visits,seasonality_index,length_of_stay\n"
        "h1,2025-03-31 21:59:59,10,100,3,1.0,4.5\n"
        "h1,2025-03-31 22:00:00,11,101,4,1.1,5.5\n"
        "h2,2025-03-31 22:30:00,12,,5,1.2,6.5\n"
        "h2,2025-03-31 23:00:00,13,103,6,,7.5\n")
    (p / "b.csv").write_text(
        "hospital_id,event_time,admission_count,current_occupancy,emergency_visits,seasonality_index,length_of_stay\n"
        "h3,2025-03-31 22:45:00,14,104,7,1.4,8.5\n"
        "h3,2025-03-31 23:00:01,15,105,8,1.5,9.5\n")
    t = io.read_csv(str(p))
    assert t.columns == [n for n, _ in io.HOSPITAL_SCHEMA] and t.count() == 6
    assert np.asarray(t["admission_count"].values).dtype == np.int32
    w = io.between(t, "event_time", "2025-03-31 22:00:00", "2025-03-31 23:00:00")   # BETWEEN is inclusive on both ends
    assert w.count() == 4 and w.to_numpy("hospital_id").tolist() == ["h1", "h2", "h2", "h3"]
    clean = w.na.drop()
    assert clean.count() == 2 and clean.to_numpy("admission_count").tolist() == [11, 14]
    assert clean.to_numpy("length_of_stay").tolist() == [5.5, 8.5]
    # the lazy form records the selection (the device loader fuses it); every other use resolves it on the host to the same rows
    lz = io.between(t, "event_time", "2025-03-31 22:00:00", "2025-03-31 23:00:00", lazy=True).na.drop()
    assert lz._pending == {"window": ("event_time", "2025-03-31 22:00:00", "2025-03-31 23:00:00"), "dropna": True}
    assert lz.count() == 2 and lz.to_numpy("admission_count").tolist() == [11, 14] and lz["length_of_stay"].values.tolist() == [5.5, 8.5]
