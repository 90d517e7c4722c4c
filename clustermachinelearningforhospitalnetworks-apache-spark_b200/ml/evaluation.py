"""ClusteringEvaluator with the pyspark.ml.evaluation surface.

Idiom kept from the reference: `Evaluator(predictionCol=..., metricName=...).evaluate(predictions)`
(ref:L162-163).  Semantics: Spark's closed-form squared-Euclidean silhouette
(ml/evaluation/ClusteringMetrics.scala) -- two streaming passes, O(n k d), not the pairwise O(n^2) one.
"""
from __future__ import annotations

import torch

from .. import _native as N
from .. import _runtime as RT
from ..table import Column, Table, features_to_device
from .param import IllegalArgumentException, Param, Params, TypeConverters, add_accessors


class ClusteringEvaluator(Params):
    _params = Params._declare(
        Param("predictionCol", "prediction column name.", converter=TypeConverters.toString),
        Param("featuresCol", "features column name.", converter=TypeConverters.toString),
        Param("metricName", "metric name in evaluation (silhouette)", lambda v: v in ("silhouette",), TypeConverters.toString),
        Param("distanceMeasure", "The distance measure. Supported options: 'squaredEuclidean' and 'cosine'.",
              lambda v: v in ("squaredEuclidean", "cosine"), TypeConverters.toString),
        Param("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
              converter=TypeConverters.toString),
    )

    def __init__(self, *, predictionCol="prediction", featuresCol="features", metricName="silhouette",
                 distanceMeasure="squaredEuclidean", weightCol=None):
        super().__init__()
        self._setDefault(predictionCol="prediction", featuresCol="features", metricName="silhouette", distanceMeasure="squaredEuclidean")
        self.setParams(predictionCol=predictionCol, featuresCol=featuresCol, metricName=metricName,
                       distanceMeasure=distanceMeasure, weightCol=weightCol)

    def setParams(self, **kw):
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def isLargerBetter(self) -> bool:
        return True

    def evaluate(self, dataset, params=None) -> float:
        if params:
            return self.copy(params).evaluate(dataset)
        t = Table.wrap(dataset)
        dev = RT.device()
        X = features_to_device(t[self.getFeaturesCol()], dev).X
        pc = t[self.getPredictionCol()]
        if not isinstance(pc, Column):
            raise IllegalArgumentException(f"predictionCol {self.getPredictionCol()!r} must be a scalar column")
        lab = pc.values if isinstance(pc.values, torch.Tensor) else torch.as_tensor(pc.values)
        lab = lab.to(device=dev, dtype=torch.int32).contiguous()      # Spark casts the prediction column to a number as well
        w = None
        if self.isDefined("weightCol") and self.getWeightCol():
            wc = t[self.getWeightCol()]
            wv = wc.values if isinstance(wc.values, torch.Tensor) else torch.as_tensor(wc.values)
            w = wv.to(device=dev, dtype=torch.float64).contiguous()
        return silhouette(X, lab, w, cosine=self.getDistanceMeasure() == "cosine")


add_accessors(ClusteringEvaluator, ["predictionCol", "featuresCol", "metricName", "distanceMeasure", "weightCol"])


def silhouette(X: torch.Tensor, labels: torch.Tensor, w=None, cosine: bool = False) -> float:
    """Silhouette of device-resident rows/labels (global over torch.distributed ranks).

    squared Euclidean: Spark's closed form D(x,G) = |x|^2 + Psi_G/N_G - 2 x.Y_G/N_G.
    cosine (CosineSilhouette): D(x,G) = 1 - x_hat . Omega_G / N_G on normalised rows x_hat -- the same two passes run on
    x_hat with Y := Omega/2 and Psi := 0, since |x_hat|^2 - 2 x_hat.(Omega/2N) = 1 - x_hat.Omega/N."""
    n, d = X.shape
    if cosine:
        X, zero_rows = N.normalize_rows(X)
        tot = torch.tensor([zero_rows], dtype=torch.int64, device=X.device)
        RT.all_reduce_sum(tot)
        if int(tot.item()):
            raise IllegalArgumentException("requirement failed: Cosine distance is not defined for zero-length vectors.")
    kmax = torch.zeros(1, dtype=torch.int64, device=X.device)
    if n:
        lo, hi = N.label_range(labels)                     # one pass over the ids (hkm_label_range)
        if lo < 0:
            raise IllegalArgumentException("prediction column holds negative cluster ids")
        kmax[0] = hi + 1
    import torch.distributed as dist
    if RT.dist_info()[1] > 1:
        dist.all_reduce(kmax, op=dist.ReduceOp.MAX)
    k = int(kmax.item())
    if k > N.load().hkm_max_k_single_pass() or N.load().hkm_needs_workspace(d, max(k, 1)):
        raise N.HkmError(f"silhouette: {k} cluster ids / d={d} outside the silhouette kernels (ids < {N.load().hkm_max_k_single_pass()}, d <= 96)")
    k = max(k, 1)
    stats = torch.zeros(k * d + 2 * k + 1, dtype=torch.float64, device=X.device)
    scratch = N.new_scratch(d, k, X.device)
    N.silhouette_stats(X, labels, k, stats, scratch, weights=w)
    RT.all_reduce_sum(stats)
    present = int((stats[k * d:k * d + k] > 0).sum().item())
    if present < 2:
        raise AssertionError("requirement failed: Number of clusters must be greater than one.")
    if cosine:
        stats[:k * d] *= 0.5
        stats[k * d + k + 1:] = 0.0
    work = torch.empty(k * d + 2 * k, dtype=torch.float64, device=X.device)
    out2 = torch.zeros(2, dtype=torch.float64, device=X.device)
    N.silhouette_score(X, labels, k, stats, work, out2, scratch, weights=w)
    RT.all_reduce_sum(out2)
    num, den = out2.cpu().tolist()
    return num / den
