"""KMeans / KMeansModel / KMeansSummary with the pyspark.ml.clustering surface.

Drop-in for the Estimator/Transformer idiom of the reference (ref:L146-148: `Est(featuresCol=...)`,
`.fit(df)`, `.transform(df)`; ref:L241 `model.write().overwrite().save(path)`).  Semantics follow
Spark MLlib 3.5 "row" solver (mllib/clustering/KMeans.scala): k-means|| or random init, Lloyd
iterations until every updated centre moves <= tol, trainingCost = cost of the last executed
iteration, summary built from one extra transform pass.  All per-row work runs in libhkm's CUDA
kernels; rows may be sharded over torch.distributed ranks (one allreduce of k*(d+1)+1 doubles
per iteration).
"""
from __future__ import annotations

import json
import os
from typing import List, Optional

import numpy as np
import torch

from .. import _native as N
from .. import _runtime as RT
from .. import _sparkrand as SR
from ..table import DeviceMatrix, Table, features_to_device, Column
from .param import IllegalArgumentException, Param, Params, TypeConverters, add_accessors, java_string_hash

_INIT_MODES = ("random", "k-means||")
_DISTANCES = ("euclidean", "cosine")
_SOLVERS = ("auto", "row", "block")


class _KMeansParams(Params):
    _params = Params._declare(
        Param("featuresCol", "features column name.", converter=TypeConverters.toString),
        Param("predictionCol", "prediction column name.", converter=TypeConverters.toString),
        Param("k", "The number of clusters to create. Must be > 1.", lambda v: v > 1, TypeConverters.toInt),
        Param("initMode", 'The initialization algorithm. Supported options: "random" and "k-means||".',
              lambda v: v in _INIT_MODES, TypeConverters.toString),
        Param("initSteps", "The number of steps for k-means|| initialization mode. Must be > 0.", lambda v: v > 0, TypeConverters.toInt),
        Param("tol", "the convergence tolerance for iterative algorithms (>= 0).", lambda v: v >= 0, TypeConverters.toFloat),
        Param("maxIter", "max number of iterations (>= 0).", lambda v: v >= 0, TypeConverters.toInt),
        Param("seed", "random seed.", converter=TypeConverters.toInt),
        Param("distanceMeasure", "the distance measure. Supported options: 'euclidean' and 'cosine'.",
              lambda v: v in _DISTANCES, TypeConverters.toString),
        Param("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
              converter=TypeConverters.toString),
        Param("solver", "The solver algorithm for optimization. Supported options: auto, row, block.",
              lambda v: v in _SOLVERS, TypeConverters.toString),
        Param("maxBlockSizeInMB", "maximum memory in MB for stacking input data into blocks (>= 0).", lambda v: v >= 0, TypeConverters.toFloat),
    )

    def __init__(self):
        super().__init__()
        self._setDefault(featuresCol="features", predictionCol="prediction", k=2, initMode="k-means||", initSteps=2,
                         tol=1e-4, maxIter=20, distanceMeasure="euclidean", solver="auto", maxBlockSizeInMB=0.0,
                         seed=java_string_hash("org.apache.spark.ml.clustering.KMeans"))


add_accessors(_KMeansParams, ["featuresCol", "predictionCol", "k", "initMode", "initSteps", "tol", "maxIter", "seed",
                              "distanceMeasure", "weightCol", "solver", "maxBlockSizeInMB"], setters=False)


def _weights_to_device(table: Table, name: Optional[str], dev) -> Optional[torch.Tensor]:
    if not name:
        return None
    c = table[name]
    if not isinstance(c, Column):
        raise IllegalArgumentException(f"weightCol {name!r} must be a scalar numeric column")
    v = c.values
    t = v if isinstance(v, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64))
    t = t.to(device=dev, dtype=torch.float64).contiguous()
    if bool((t < 0).any()):
        raise IllegalArgumentException("weights must be non-negative")
    return t


class KMeansSummary:
    """pyspark.ml.clustering.KMeansSummary."""

    def __init__(self, predictions: Table, predictionCol: str, featuresCol: str, k: int, numIter: int, trainingCost: float,
                 clusterSizes: List[int]):
        self.predictions = predictions
        self.predictionCol = predictionCol
        self.featuresCol = featuresCol
        self.k = k
        self.numIter = numIter
        self.trainingCost = trainingCost
        self.clusterSizes = clusterSizes

    @property
    def cluster(self) -> Table:
        return self.predictions.select(self.predictionCol)


class _ModelWriter:
    def __init__(self, model: "KMeansModel"):
        self._m, self._overwrite = model, False

    def overwrite(self) -> "_ModelWriter":
        self._overwrite = True
        return self

    def save(self, path: str) -> None:
        self._m._save(path, self._overwrite)


class KMeansModel(_KMeansParams):
    """Model fitted by KMeans (pyspark.ml.clustering.KMeansModel)."""

    def __init__(self, centers: np.ndarray):
        super().__init__()
        self._centers = np.ascontiguousarray(centers, dtype=np.float64)
        self._summary: Optional[KMeansSummary] = None
        self._dev_centers: Optional[torch.Tensor] = None

    # ---- pyspark surface
    def clusterCenters(self) -> List[np.ndarray]:
        return [c.copy() for c in self._centers]

    @property
    def numFeatures(self) -> int:
        return int(self._centers.shape[1])

    @property
    def hasSummary(self) -> bool:
        return self._summary is not None

    @property
    def summary(self) -> KMeansSummary:
        if self._summary is None:
            raise RuntimeError(f"No training summary available for this {type(self).__name__}")
        return self._summary

    def setFeaturesCol(self, value):
        return self._set(featuresCol=value)

    def setPredictionCol(self, value):
        return self._set(predictionCol=value)

    def _centers_on(self, dev) -> torch.Tensor:
        if self._dev_centers is None or self._dev_centers.device != dev:
            self._dev_centers = torch.from_numpy(self._centers).to(dev)
        return self._dev_centers

    def _pass_centers(self, dev) -> torch.Tensor:
        """Centres as the data passes take them: unit length under the cosine measure."""
        if self.getDistanceMeasure() != "cosine":
            return self._centers_on(dev)
        return torch.from_numpy(_unit_centers(self._centers)).to(dev)

    def _predict_matrix(self, X: torch.Tensor) -> torch.Tensor:
        labels = torch.empty(X.shape[0], dtype=torch.int32, device=X.device)
        if X.shape[0]:
            if self.getDistanceMeasure() == "cosine":
                X = _unit_rows(X)
            N.assign_reduce(X, self._pass_centers(X.device), labels=labels, do_reduce=False)
        return labels

    def transform(self, dataset, params=None) -> Table:
        """Adds predictionCol (int32, device-resident) = index of the closest centre (8a-9)."""
        if params:
            return self.copy(params).transform(dataset)
        t = Table.wrap(dataset)
        fm = features_to_device(t[self.getFeaturesCol()], RT.device())
        if fm.shape[1] != self.numFeatures:
            raise IllegalArgumentException(f"features have {fm.shape[1]} columns, the model was fit on {self.numFeatures}")
        labels = self._predict_matrix(fm.X)
        return t.withColumn(self.getFeaturesCol(), fm).withColumn(self.getPredictionCol(), Column(labels))

    def predict(self, value) -> int:
        x = np.asarray(value, dtype=np.float64).reshape(1, -1)
        if x.shape[1] != self.numFeatures:
            raise IllegalArgumentException(f"vector has {x.shape[1]} entries, the model was fit on {self.numFeatures}")
        X = torch.from_numpy(x).to(RT.device())
        return int(self._predict_matrix(X).item())

    # ---- persistence (ref:L241-243 `model.write().overwrite().save(path)`)
    def write(self) -> _ModelWriter:
        return _ModelWriter(self)

    def save(self, path: str) -> None:
        self._save(path, False)

    def _save(self, path: str, overwrite: bool) -> None:
        if os.path.exists(path) and not overwrite:
            raise IOError(f"Path {path} already exists. To overwrite it, please use write.overwrite().save(path) for Scala and use write().overwrite().save(path) for Java and Python.")
        os.makedirs(os.path.join(path, "metadata"), exist_ok=True)
        os.makedirs(os.path.join(path, "data"), exist_ok=True)
        meta = {"class": "org.apache.spark.ml.clustering.KMeansModel", "uid": self.uid,
                "paramMap": {k: v for k, v in self._paramMap.items()}, "defaultParamMap": {k: v for k, v in self._defaultParamMap.items()}}
        with open(os.path.join(path, "metadata", "part-00000"), "w") as f:
            json.dump(meta, f)
        np.save(os.path.join(path, "data", "clusterCenters.npy"), self._centers)

    @classmethod
    def load(cls, path: str) -> "KMeansModel":
        with open(os.path.join(path, "metadata", "part-00000")) as f:
            meta = json.load(f)
        m = cls(np.load(os.path.join(path, "data", "clusterCenters.npy")))
        m.uid = meta["uid"]
        m._defaultParamMap.update(meta.get("defaultParamMap", {}))
        m._paramMap.update(meta.get("paramMap", {}))
        return m

    @classmethod
    def read(cls):
        class _R:
            @staticmethod
            def load(path):
                return cls.load(path)
        return _R()


class KMeans(_KMeansParams):
    """pyspark.ml.clustering.KMeans on B200 (fit/transform, k, maxIter, tol, seed, initMode, ...)."""

    def __init__(self, *, featuresCol="features", predictionCol="prediction", k=2, initMode="k-means||", initSteps=2, tol=1e-4,
                 maxIter=20, seed=None, distanceMeasure="euclidean", weightCol=None, solver="auto", maxBlockSizeInMB=0.0):
        super().__init__()
        self.setParams(featuresCol=featuresCol, predictionCol=predictionCol, k=k, initMode=initMode, initSteps=initSteps, tol=tol,
                       maxIter=maxIter, seed=seed, distanceMeasure=distanceMeasure, weightCol=weightCol, solver=solver,
                       maxBlockSizeInMB=maxBlockSizeInMB)
        # Not pyspark Params: the Spark partition layout this run emulates (k-means|| and takeSample
        # depend on it by construction) and optional explicit initial centres (mllib setInitialModel).
        self.numPartitions = 8
        self.initialCenters: Optional[np.ndarray] = None

    def setParams(self, **kw):
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def setNumPartitions(self, p: int):
        if p < 1:
            raise IllegalArgumentException("numPartitions must be >= 1")
        self.numPartitions = int(p)
        return self

    def setInitialCenters(self, centers):
        self.initialCenters = None if centers is None else np.ascontiguousarray(centers, dtype=np.float64)
        return self

    # ---- fit
    def fit(self, dataset, params=None) -> KMeansModel:
        if params:
            return self.copy(params).fit(dataset)
        return self._fit(Table.wrap(dataset))

    def _fit(self, table: Table) -> KMeansModel:
        from . import _kmeans_init as KI
        cosine = self.getDistanceMeasure() == "cosine"
        dev = RT.device()
        fm = features_to_device(table[self.getFeaturesCol()], dev)
        X = fm.X
        n_local, d = X.shape
        w = _weights_to_device(table, self.getOrDefault("weightCol") if self.isDefined("weightCol") else None, dev)
        rank, world, _ = RT.dist_info()
        sizes = RT.all_gather_object(int(n_local))
        n_global = sum(sizes)
        if n_global == 0:
            raise IllegalArgumentException("requirement failed: KMeans requires a non-empty dataset")
        k = self.getK()
        if d > N.load().hkm_max_d() or k > N.load().hkm_max_k():
            raise N.HkmError(f"d={d}, k={k} is outside the built CUDA kernels (d<={N.load().hkm_max_d()}, k<={N.load().hkm_max_k()})")

        layout = KI.PartitionLayout.build(sizes, rank, self.numPartitions)
        # cosine: every data pass runs on the unit rows (one extra resident copy); the initialisers still pick RAW rows
        Xp = _unit_rows(X) if cosine else X
        if self.initialCenters is not None:
            C0 = SR.distinct_rows(self.initialCenters)
        elif self.getInitMode() == "random":
            C0 = KI.init_random(X, layout, k, self.getSeed())
        else:
            C0 = KI.init_kmeans_parallel(X, layout, k, self.getInitSteps(), self.getSeed(), X_unit=Xp if cosine else None)
        centers, num_iter, cost = lloyd(Xp, _unit_centers(C0) if cosine else C0, w, self.getMaxIter(), self.getTol(), cosine=cosine)
        if num_iter == 0:
            centers = C0          # maxIter = 0: Spark hands back the initial centres as drawn (raw rows)

        model = KMeansModel(centers)
        self._copyValues(model)
        # summary: one more pass, as ml.clustering.KMeans.fit does with model.transform(dataset)
        pred, sizes_k = _predict_and_count(Xp, model._pass_centers(dev))
        del Xp
        pred_table = table.withColumn(self.getFeaturesCol(), fm).withColumn(self.getPredictionCol(), Column(pred))
        model._summary = KMeansSummary(pred_table, self.getPredictionCol(), self.getFeaturesCol(), self.getK(), num_iter, cost, sizes_k)
        return model


add_accessors(KMeans, ["featuresCol", "predictionCol", "k", "initMode", "initSteps", "tol", "maxIter", "seed",
                       "distanceMeasure", "weightCol", "solver", "maxBlockSizeInMB"])


def _unit_rows(X: torch.Tensor) -> torch.Tensor:
    """x / |x|_2 per row; Spark's CosineDistanceMeasure refuses zero-length vectors (global over ranks)."""
    Xu, zero_rows = N.normalize_rows(X)
    tot = torch.tensor([zero_rows], dtype=torch.int64, device=X.device)
    RT.all_reduce_sum(tot)
    if int(tot.item()):
        raise IllegalArgumentException("requirement failed: Cosine distance is not defined for zero-length vectors.")
    return Xu


def _unit_centers(C: np.ndarray) -> np.ndarray:
    nrm = np.sqrt((C * C).sum(axis=1))
    if (nrm == 0).any():
        raise IllegalArgumentException("requirement failed: Cosine distance is not defined for zero-length vectors.")
    return C * (1.0 / nrm)[:, None]


def lloyd(X: torch.Tensor, C0: np.ndarray, w: Optional[torch.Tensor], max_iter: int, tol: float, check_every: int = 4,
          cosine: bool = False):
    """KMeans.runAlgorithmWithWeight loop on the current device / process group.

    Every iteration is hkm_assign_reduce -> (all-reduce of k*(d+1)+1 doubles) -> hkm_update_centers,
    enqueued without host round trips; the device-side state makes passes after convergence no-ops
    and the host only reads it back every `check_every` iterations.
    Returns (centers [k, d] ndarray, numIter, trainingCost).
    """
    dev = X.device
    k, d = C0.shape
    C = torch.from_numpy(np.ascontiguousarray(C0, dtype=np.float64)).to(dev)
    if max_iter == 0:
        return C0.copy(), 0, 0.0
    P = k * d + k + 1
    partial = torch.zeros(P, dtype=torch.float64, device=dev)
    scratch = N.new_scratch(d, k, dev)
    state = N.new_state(dev)
    xsum = torch.zeros(1, dtype=torch.float64, device=dev)
    world = RT.dist_info()[1]

    big_k = bool(N.load().hkm_needs_workspace(d, k))     # k > 64 or a shape outside the tiled kernels: chunked pass + separate update
    lab_ws = torch.empty(X.shape[0], dtype=torch.int32, device=dev) if big_k else None
    d2_ws = torch.empty(X.shape[0], dtype=torch.float64, device=dev) if big_k else None
    xchg = RT.PeerExchange.get(dev) if world > 1 and not cosine and not big_k else None
    if xchg is not None and not xchg.fits(k, d):
        xchg = None          # partial sums larger than an exchange slot: NCCL all-reduce + hkm_update_centers instead

    def one_iter(mode):
        if cosine or big_k:   # cosine: X holds unit rows, C unit centres, CosineDistanceMeasure.centroid as the driver step
            N.assign_reduce(X, C, partial, lab_ws, d2_ws, weights=w, state=state, scratch=scratch, xnorm2_sum=xsum, xnorm2_mode=mode)
            if world > 1:
                RT.all_reduce_sum(partial)
            N.update_centers(partial, C, tol, state, cosine=cosine)
            return
        if world == 1:   # one launch: the last CTA of the pass also updates the centres
            N.lloyd_step(X, C, partial, tol, state, scratch, weights=w, xnorm2_sum=xsum, xnorm2_mode=mode)
            return
        if xchg is not None:   # one launch too: the cross-GPU sum runs over NVLink peer memory inside the kernel tail
            N.lloyd_step_peer(X, C, partial, tol, state, scratch, xchg, weights=w, xnorm2_sum=xsum, xnorm2_mode=mode)
            return
        N.assign_reduce(X, C, partial, weights=w, state=state, scratch=scratch, xnorm2_sum=xsum, xnorm2_mode=mode)
        RT.all_reduce_sum(partial)
        N.update_centers(partial, C, tol, state)

    one_iter(1)  # HKM_XNORM_STORE: exact per-row cost, and sum w|x|^2 for the later passes
    if xchg is not None:
        xchg.check()
    st = N.read_state(state)
    # REUSE needs cost / sum|x|^2 well above fp64 rounding; fall back to per-row norms otherwise
    xs_total = xsum.clone()
    if world > 1:
        RT.all_reduce_sum(xs_total)
    ratio_ok = st.cost > 1e-6 * float(xs_total.item())
    mode = 2 if ratio_ok else 0
    it = 1
    while it < max_iter and not st.converged:
        burst = min(check_every, max_iter - it)
        for _ in range(burst):
            one_iter(mode)
        if xchg is not None:
            xchg.check()     # a timed-out exchange stops here, on every rank, before more iterations run on partial sums
        st = N.read_state(state)
        it = st.iter
    return C.cpu().numpy(), int(st.iter), float(st.cost)


def _predict_and_count(X: torch.Tensor, C: torch.Tensor):
    """Final transform pass: labels plus cluster sizes (global over ranks)."""
    n, d = X.shape
    k = C.shape[0]
    labels = torch.empty(n, dtype=torch.int32, device=X.device)
    scratch = N.new_scratch(d, k, X.device)
    # a labels-only pass (no per-cluster sums: it runs at the rate of transform) and a histogram of the labels
    N.assign_reduce(X, C, None, labels=labels, scratch=scratch, do_reduce=False)
    counts = N.label_counts(labels, k)
    RT.all_reduce_sum(counts)
    return labels, [int(v) for v in counts.cpu().tolist()]
