"""VectorAssembler / StandardScaler / StandardScalerModel with the pyspark.ml.feature surface
(SURVEY 8a-1, 8a-2, 8b).  The reference uses VectorAssembler exactly like this (ref:L134-136):

    assembler = VectorAssembler(inputCols=feature_cols, outputCol="features")
    final_data = assembler.transform(training_df).select("features", ...)

Here transform() only uploads the typed source columns and returns a lazy vector column; the
row-major fp64 matrix is produced by one hkm_assemble launch when something needs it, with
StandardScalerModel's (shift, scale) folded in when a scaler sits in between.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from .. import _native as N
from .. import _runtime as RT
from ..table import AssembledColumn, Column, DeviceMatrix, Table, features_to_device
from .param import IllegalArgumentException, Param, Params, TypeConverters, add_accessors


class SparkException(RuntimeError):
    """Stand-in for org.apache.spark.SparkException (VectorAssembler with handleInvalid='error')."""


_TORCH_OK = set(N.DTYPE_CODES)


def _to_device_scalar(c: Column, dev):
    v, valid = c.values, c.valid
    if not isinstance(v, torch.Tensor):
        a = np.asarray(v)
        if a.dtype == object:
            raise IllegalArgumentException("VectorAssembler only takes numeric / boolean / vector columns")
        if a.dtype.kind == "u" and a.dtype.itemsize > 1:
            a = a.astype(np.int64)
        elif a.dtype == np.float16:
            a = a.astype(np.float32)
        v = torch.from_numpy(np.ascontiguousarray(a))
    if v.dtype not in _TORCH_OK:
        v = v.to(torch.float64)
    v = v.to(dev, non_blocking=True).contiguous()
    if valid is not None:
        valid = valid if isinstance(valid, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(valid, dtype=np.uint8)))
        valid = valid.to(device=dev, dtype=torch.uint8).contiguous()
    return v, valid


class VectorAssembler(Params):
    _params = Params._declare(
        Param("inputCols", "input column names.", converter=TypeConverters.toListString),
        Param("outputCol", "output column name.", converter=TypeConverters.toString),
        Param("handleInvalid", "How to handle invalid data (NULL and NaN values). Options are 'skip' (filter out rows with "
              "invalid data), 'error' (throw an error), or 'keep' (return relevant number of NaN in the output).",
              lambda v: v in ("skip", "error", "keep"), TypeConverters.toString),
    )

    def __init__(self, *, inputCols: Optional[List[str]] = None, outputCol: Optional[str] = None, handleInvalid: str = "error"):
        super().__init__()
        self._setDefault(handleInvalid="error", outputCol=self.uid + "__output")
        self.setParams(inputCols=inputCols, outputCol=outputCol, handleInvalid=handleInvalid)

    def setParams(self, **kw):
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def transform(self, dataset, params=None) -> Table:
        if params:
            return self.copy(params).transform(dataset)
        t = Table.wrap(dataset)
        dev = RT.device()
        if not self.isDefined("inputCols"):
            raise IllegalArgumentException("VectorAssembler: inputCols is not set")
        fused = self._fused_selection(t, dev)
        if fused is not None:
            return fused
        cols, valids, names = [], [], []
        for name in self.getInputCols():
            c = t[name]
            if isinstance(c, (DeviceMatrix, AssembledColumn)) or (not isinstance(c, Column)):
                M = features_to_device(c, dev).X          # vector input: its components become columns
                for j in range(M.shape[1]):
                    cols.append(M[:, j].contiguous()); valids.append(None); names.append(f"{name}_{j}")
            else:
                v, valid = _to_device_scalar(c, dev)
                cols.append(v); valids.append(valid); names.append(name)
        n = t.count()
        if len(cols) > 64:
            raise N.HkmError("VectorAssembler: more than 64 assembled columns are not supported by hkm_assemble")
        mode = self.getHandleInvalid()
        if mode == "skip":
            # one compaction on the device; its keep mask (a byte per row) filters the table's other columns
            X, keep, _, bad = N.assemble_filtered(cols, valids, n)
            if bad:
                t = t.filter_rows(keep.cpu().numpy().astype(bool))
            return t.withColumn(self.getOutputCol(), DeviceMatrix(X.contiguous()))
        col = AssembledColumn(cols, valids, names, n)
        if mode == "error":
            # pyspark raises lazily when an action runs; we check now so the failure names this stage
            bad = _count_invalid(cols, valids, n)
            if bad:
                raise SparkException(f"Encountered null while assembling a row with handleInvalid = \"error\". Consider removing "
                                     f"nulls from dataset or using handleInvalid = \"keep\" or \"skip\". ({bad} invalid rows)")
        return t.withColumn(self.getOutputCol(), col)


def _host_validity(c: Column):
    """uint8 validity of a scalar host column for na.drop: not null and (floats) not NaN; None when every row is present."""
    vals = np.asarray(c.values) if not isinstance(c.values, torch.Tensor) else None
    ok = None if c.valid is None else np.asarray(c.valid.cpu() if isinstance(c.valid, torch.Tensor) else c.valid, dtype=bool)
    if vals is not None and vals.dtype.kind == "f":
        nn = ~np.isnan(vals)
        ok = nn if ok is None else ok & nn
    elif vals is not None and vals.dtype == object:
        nn = np.array([x is not None for x in vals], dtype=bool)
        ok = nn if ok is None else ok & nn
    if ok is None or ok.all():
        return None
    return np.ascontiguousarray(ok, dtype=np.uint8)


def _fused_selection_impl(self, t: Table, dev):
    """The reference's batch extraction (ref:L123-128: time window, then na.drop) fused with the assembly: the window predicate and the
    null masks ride in the loader's compaction pass on the device, and only the keep mask (one byte per row) comes back to filter
    the table's other columns."""
    pend = t._pending
    if pend is None or not (pend["dropna"] or self.getHandleInvalid() == "skip"):
        return None
    from .. import io
    raw = t._cols
    win = None
    if pend["window"] is not None:
        wname, lo, hi = pend["window"]
        wc = raw[wname]
        w = io.window_int64(wc, lo, hi) if isinstance(wc, Column) and not isinstance(wc.values, torch.Tensor) else None
        if w is None:
            return None                                          # not a timestamp / integer window: resolved on the host
        ts = torch.from_numpy(np.ascontiguousarray(w[0])).to(dev)
        tsv = None
        if wc.valid is not None:
            tsv = torch.from_numpy(np.ascontiguousarray(np.asarray(wc.valid, dtype=np.uint8))).to(dev)
        win = (ts, tsv, w[1], w[2])
    cols, valids = [], []
    for name in self.getInputCols():
        c = raw[name]
        if not isinstance(c, Column):
            return None                                          # vector inputs: the general path
        v, valid = _to_device_scalar(c, dev)
        cols.append(v); valids.append(valid)
    if len(cols) > 64:
        raise N.HkmError("VectorAssembler: more than 64 assembled columns are not supported by hkm_assemble")
    extra = []
    if pend["dropna"]:                                           # na.drop looks at every column, not only the features
        for name, c in raw.items():
            if name in self.getInputCols() or not isinstance(c, Column):
                continue
            ok = _host_validity(c)
            if ok is not None:
                extra.append(torch.from_numpy(ok).to(dev))
    if len(extra) > 8:
        return None
    n = _nrows_raw(raw)
    X, keep, outside, bad = N.assemble_filtered(cols, valids, n, window=win, extra_valid=extra)
    if bad and not pend["dropna"] and self.getHandleInvalid() == "error":
        raise SparkException(f"Encountered null while assembling a row with handleInvalid = \"error\". ({bad} invalid rows)")
    keep_np = keep.cpu().numpy().astype(bool)
    out = Table(raw).filter_rows(keep_np)
    return out.withColumn(self.getOutputCol(), DeviceMatrix(X.contiguous()))


def _nrows_raw(raw) -> int:
    for c in raw.values():
        return len(c.values) if isinstance(c, Column) else len(c)
    return 0


VectorAssembler._fused_selection = _fused_selection_impl


def _count_invalid(cols, valids, n) -> int:
    """Rows holding a null or NaN; skipped entirely when no column can hold one."""
    if n == 0 or (all(v is None for v in valids) and not any(c.dtype.is_floating_point for c in cols)):
        return 0
    return N.count_invalid_rows(cols, valids, n)


add_accessors(VectorAssembler, ["inputCols", "outputCol", "handleInvalid"])


class _ScalerParams(Params):
    _params = Params._declare(
        Param("inputCol", "input column name.", converter=TypeConverters.toString),
        Param("outputCol", "output column name.", converter=TypeConverters.toString),
        Param("withMean", "Center data with mean", converter=TypeConverters.toBoolean),
        Param("withStd", "Scale to unit standard deviation", converter=TypeConverters.toBoolean),
    )

    def __init__(self):
        super().__init__()
        self._setDefault(withMean=False, withStd=True, outputCol=self.uid + "__output")


add_accessors(_ScalerParams, ["inputCol", "outputCol", "withMean", "withStd"], setters=False)


class StandardScalerModel(_ScalerParams):
    def __init__(self, mean: np.ndarray, std: np.ndarray):
        super().__init__()
        self._mean = np.asarray(mean, dtype=np.float64)
        self._std = np.asarray(std, dtype=np.float64)

    @property
    def mean(self) -> np.ndarray:
        return self._mean.copy()

    @property
    def std(self) -> np.ndarray:
        return self._std.copy()

    def setInputCol(self, v):
        return self._set(inputCol=v)

    def setOutputCol(self, v):
        return self._set(outputCol=v)

    def transform(self, dataset, params=None) -> Table:
        """values = (values - shift) * scale with scale = (std == 0) ? 0 : 1/std -- a multiply by the
        reciprocal, as StandardScalerModel.getTransformFunc does."""
        if params:
            return self.copy(params).transform(dataset)
        t = Table.wrap(dataset)
        dev = RT.device()
        shift = torch.from_numpy(self._mean).to(dev) if self.getWithMean() else None
        scale = None
        if self.getWithStd():
            with np.errstate(divide="ignore"):
                scale = torch.from_numpy(np.where(self._std == 0, 0.0, 1.0 / np.where(self._std == 0, 1.0, self._std))).to(dev)
        c = t[self.getInputCol()]
        if isinstance(c, AssembledColumn) and c.shift is None and c.scale is None and c._mat is None:
            out = c.with_scaling(shift, scale)            # fused into the loader's single write
        else:
            X = features_to_device(c, dev).X
            if X.shape[1] != len(self._mean):
                raise IllegalArgumentException(f"vector has {X.shape[1]} entries, the scaler was fit on {len(self._mean)}")
            out = DeviceMatrix(N.scale_rows(X, shift, scale)) if (shift is not None or scale is not None) else DeviceMatrix(X)
        return t.withColumn(self.getOutputCol(), out)


class StandardScaler(_ScalerParams):
    """pyspark.ml.feature.StandardScaler: defaults withMean=False, withStd=True; std is the
    corrected sample standard deviation (Summarizer), zero-variance features scale to 0."""

    def __init__(self, *, withMean: bool = False, withStd: bool = True, inputCol: Optional[str] = None, outputCol: Optional[str] = None):
        super().__init__()
        self.setParams(withMean=withMean, withStd=withStd, inputCol=inputCol, outputCol=outputCol)

    def setParams(self, **kw):
        return self._set(**{k: v for k, v in kw.items() if v is not None})

    def fit(self, dataset, params=None) -> StandardScalerModel:
        if params:
            return self.copy(params).fit(dataset)
        t = Table.wrap(dataset)
        dev = RT.device()
        c = t[self.getInputCol()]
        rank, world, _ = RT.dist_info()
        lazy = isinstance(c, AssembledColumn) and c.shift is None and c.scale is None and c._mat is None
        if lazy:
            n, d = c.shape
            pivot = torch.stack([col[0].to(torch.float64) for col in c.cols]) if n else torch.zeros(d, dtype=torch.float64, device=dev)
        else:
            X = features_to_device(c, dev).X
            n, d = X.shape
            pivot = X[0].clone() if n else torch.zeros(d, dtype=torch.float64, device=dev)
        pivot = torch.nan_to_num(pivot)
        if world > 1:
            import torch.distributed as dist
            dist.broadcast(pivot, 0)
        s = N.column_moments(c.cols, c.valids, n, pivot) if lazy else N.moments(X, pivot)
        tot = torch.cat([s.reshape(-1), torch.tensor([float(n)], dtype=torch.float64, device=dev)])
        RT.all_reduce_sum(tot)
        ntot = float(tot[-1].item())
        if ntot == 0:
            raise IllegalArgumentException("requirement failed: StandardScaler requires a non-empty dataset")
        s = tot[:-1].view(d, 2).cpu().numpy()
        K = pivot.cpu().numpy()
        mean = K + s[:, 0] / ntot
        var = np.zeros(d) if ntot <= 1 else np.maximum((s[:, 1] - s[:, 0] * s[:, 0] / ntot) / (ntot - 1.0), 0.0)
        model = StandardScalerModel(mean, np.sqrt(var))
        self._copyValues(model)
        return model


add_accessors(StandardScaler, ["inputCol", "outputCol", "withMean", "withStd"])
