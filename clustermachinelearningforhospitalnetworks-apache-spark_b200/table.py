"""Columnar table standing in for the pyspark DataFrame at the reference's L4/L5 seam.

pyspark DataFrames do not exist on this box (no JVM), so Estimators/Transformers here accept a
`Table`: named columns that are NumPy arrays, torch tensors (host or device), pyarrow columns
(nulls kept as a validity mask, as Spark's nullable IntegerType/DoubleType of ref:L64-72) or a
`DeviceMatrix` -- the row-major fp64 feature matrix in HBM that VectorAssembler produces
(ref:L135-136) and everything downstream consumes without leaving the GPU.
Only what the reference's idiom uses is provided: select / withColumn / drop / count /
na.drop (ref:L128) / toPandas (ref:L205).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Any, Iterable, Optional

import numpy as np
import torch


class DeviceMatrix:
    """Row-major fp64 [n, d] feature matrix resident in HBM (a pyspark 'Vector' column)."""

    __slots__ = ("X",)

    def __init__(self, X: torch.Tensor):
        if X.dtype != torch.float64 or X.dim() != 2 or not X.is_contiguous():
            raise TypeError("DeviceMatrix needs a contiguous 2-D float64 tensor")
        self.X = X

    @property
    def shape(self):
        return tuple(self.X.shape)

    def __len__(self):
        return self.X.shape[0]

    def numpy(self) -> np.ndarray:
        return self.X.detach().cpu().numpy()


class AssembledColumn:
    """Lazy VectorAssembler output: references to the typed source columns (already on the device)
    plus an optional (shift, scale) from StandardScalerModel.transform.  Materialised by ONE
    hkm_assemble launch when a consumer needs the row-major matrix, so assemble -> scale -> fit writes
    the standardised matrix to HBM exactly once."""

    def __init__(self, cols, valids, names, n, shift=None, scale=None):
        self.cols, self.valids, self.names, self.n = cols, valids, names, n
        self.shift, self.scale = shift, scale
        self._mat = None

    @property
    def shape(self):
        return (self.n, len(self.cols))

    def __len__(self):
        return self.n

    def with_scaling(self, shift, scale) -> "AssembledColumn":
        if self.shift is not None or self.scale is not None:
            raise ValueError("column is already scaled; materialise it first")
        return AssembledColumn(self.cols, self.valids, self.names, self.n, shift, scale)

    def materialize(self) -> "DeviceMatrix":
        if self._mat is None:
            from . import _native as N
            X, _, _ = N.assemble(self.cols, self.valids, self.n, self.shift, self.scale, skip_invalid=False)
            self._mat = DeviceMatrix(X)
        return self._mat

    def numpy(self):
        return self.materialize().numpy()


class Column:
    """One scalar column: values plus an optional validity mask (True = present)."""

    __slots__ = ("values", "valid")

    def __init__(self, values, valid=None):
        self.values = values
        self.valid = valid

    def __len__(self):
        return len(self.values)


def _as_column(v) -> Any:
    if isinstance(v, (Column, DeviceMatrix, AssembledColumn)):
        return v
    try:
        import pyarrow as pa
        if isinstance(v, (pa.ChunkedArray, pa.Array)):
            arr = v.combine_chunks() if isinstance(v, pa.ChunkedArray) else v
            if pa.types.is_fixed_size_list(arr.type) or pa.types.is_list(arr.type):
                return np.asarray(arr.to_pylist(), dtype=np.float64)
            if arr.null_count:
                valid = np.asarray(arr.is_valid())
                if pa.types.is_floating(arr.type) or pa.types.is_integer(arr.type):
                    vals = arr.fill_null(0).to_numpy(zero_copy_only=False)
                elif pa.types.is_timestamp(arr.type):
                    vals = arr.to_numpy(zero_copy_only=False)
                else:
                    vals = np.asarray(arr.to_pylist(), dtype=object)
                return Column(vals, valid)
            if pa.types.is_floating(arr.type) or pa.types.is_integer(arr.type) or pa.types.is_boolean(arr.type) or pa.types.is_timestamp(arr.type):
                return Column(arr.to_numpy(zero_copy_only=False))
            return Column(np.asarray(arr.to_pylist(), dtype=object))
    except ImportError:
        pass
    if isinstance(v, np.ma.MaskedArray):
        return Column(np.asarray(v.filled(0)), ~np.ma.getmaskarray(v))
    if isinstance(v, torch.Tensor):
        if v.dim() == 2:
            return v
        return Column(v)
    a = np.asarray(v)
    if a.ndim == 2:
        return a
    if a.dtype == object and a.size and isinstance(a.flat[0], (list, tuple, np.ndarray)):
        return np.asarray([np.asarray(x, dtype=np.float64) for x in a], dtype=np.float64)
    return Column(a)


def _nrows(c) -> int:
    if isinstance(c, (DeviceMatrix, AssembledColumn)):
        return len(c)
    if isinstance(c, Column):
        return len(c.values)
    return c.shape[0]


def _take(c, mask_np: np.ndarray):
    """Row filter by boolean mask (host mask; device columns get a device copy of it)."""
    if isinstance(c, AssembledColumn):
        c = c.materialize()
    if isinstance(c, DeviceMatrix):
        m = torch.from_numpy(mask_np).to(c.X.device)
        return DeviceMatrix(c.X[m].contiguous())
    if isinstance(c, Column):
        if isinstance(c.values, torch.Tensor):
            m = torch.from_numpy(mask_np).to(c.values.device)
            return Column(c.values[m], None if c.valid is None else (c.valid[m] if isinstance(c.valid, torch.Tensor) else c.valid[mask_np]))
        return Column(c.values[mask_np], None if c.valid is None else np.asarray(c.valid)[mask_np])
    if isinstance(c, torch.Tensor):
        return c[torch.from_numpy(mask_np).to(c.device)]
    return c[mask_np]


class _NaFunctions:
    def __init__(self, table: "Table"):
        self._t = table

    def drop(self, how: str = "any", subset: Optional[Iterable[str]] = None) -> "Table":
        """DataFrame.na.drop (ref:L128): drop rows holding a null (or NaN) in `subset` columns.  On a table with a pending
        (lazy) time window the drop stays pending too, so the loader can fuse both."""
        t = self._t
        if t._pending is not None and how == "any" and subset is None:
            return t.with_pending(dropna=True)
        t = t.resolve()
        names = list(subset) if subset is not None else t.columns
        n = t.count()
        bad_any = np.zeros(n, dtype=bool)
        bad_all = np.ones(n, dtype=bool)
        for name in names:
            c = t._cols[name]
            if not isinstance(c, Column):
                continue
            bad = np.zeros(n, dtype=bool)
            if c.valid is not None:
                v = c.valid.cpu().numpy() if isinstance(c.valid, torch.Tensor) else np.asarray(c.valid)
                bad |= ~v
            vals = c.values.cpu().numpy() if isinstance(c.values, torch.Tensor) else c.values
            if vals.dtype.kind == "f":
                bad |= np.isnan(vals)
            elif vals.dtype == object:
                bad |= np.array([x is None for x in vals])
            bad_any |= bad
            bad_all &= bad
        keep = ~(bad_any if how == "any" else bad_all)
        if keep.all():
            return t
        return Table(OrderedDict((k, _take(c, keep)) for k, c in t._cols.items()))


class Table:
    """Columns by name.  A table may carry a PENDING row selection (`io.between(..., lazy=True)`, then `.na.drop()`): the
    reference's `WHERE event_time BETWEEN ... ` + `.na.drop()` (ref:L123-128).  VectorAssembler fuses a pending selection into the
    device loader's compaction pass; any other use of the table resolves it on the host first."""

    def __init__(self, columns=None):
        self._pending = None          # {"window": (column, lo, hi) | None, "dropna": bool}
        self._resolved = None
        self._cols: "OrderedDict[str, Any]" = OrderedDict()
        if columns is not None:
            items = columns.items() if hasattr(columns, "items") else columns
            for k, v in items:
                self._cols[k] = _as_column(v)
        ns = {_nrows(c) for c in self._cols.values()}
        if len(ns) > 1:
            raise ValueError(f"columns have different lengths: {sorted(ns)}")

    # ---- constructors
    @staticmethod
    def from_arrow(tbl) -> "Table":
        return Table(OrderedDict((name, tbl.column(name)) for name in tbl.column_names))

    @staticmethod
    def from_pandas(df) -> "Table":
        cols = OrderedDict()
        for name in df.columns:
            s = df[name]
            if s.isna().any() and s.dtype.kind in "fiu":
                cols[name] = Column(s.fillna(0).to_numpy(), (~s.isna()).to_numpy())
            elif s.isna().any():
                cols[name] = Column(s.to_numpy(dtype=object), (~s.isna()).to_numpy())
            else:
                cols[name] = s.to_numpy()
        return Table(cols)

    @staticmethod
    def wrap(obj) -> "Table":
        if isinstance(obj, Table):
            return obj
        try:
            import pyarrow as pa
            if isinstance(obj, pa.Table):
                return Table.from_arrow(obj)
        except ImportError:
            pass
        try:
            import pandas as pd
            if isinstance(obj, pd.DataFrame):
                return Table.from_pandas(obj)
        except ImportError:
            pass
        if isinstance(obj, dict):
            return Table(obj)
        raise TypeError(f"cannot interpret {type(obj).__name__} as a table (pass Table, pyarrow.Table, pandas.DataFrame or dict)")

    # ---- pending row selection
    def with_pending(self, window=None, dropna=None) -> "Table":
        t = Table(self._cols)
        cur = dict(self._pending or {"window": None, "dropna": False})
        if window is not None:
            if cur["window"] is not None:
                raise ValueError("a pending time window is already set; resolve() the table before adding another")
            cur["window"] = window
        if dropna is not None:
            cur["dropna"] = bool(dropna)
        t._pending = cur
        return t

    def resolve(self) -> "Table":
        """The table with its pending selection applied on the host (NumPy masks); a table without one is returned as is."""
        if self._pending is None:
            return self
        if self._resolved is None:
            t = Table(self._cols)
            if self._pending["window"] is not None:
                from . import io
                col, lo, hi = self._pending["window"]
                t = io.between(t, col, lo, hi)
            if self._pending["dropna"]:
                t = t.na.drop()
            self._resolved = t
        return self._resolved

    # ---- DataFrame-like surface
    @property
    def columns(self):
        return list(self._cols.keys())

    @property
    def na(self) -> _NaFunctions:
        return _NaFunctions(self)

    def count(self) -> int:
        if self._pending is not None:
            return self.resolve().count()
        for c in self._cols.values():
            return _nrows(c)
        return 0

    def __contains__(self, name):
        return name in self._cols

    def __getitem__(self, name):
        if name not in self._cols:
            raise KeyError(f"column {name!r} does not exist; available: {self.columns}")
        return self.resolve()._cols[name]

    def select(self, *names) -> "Table":
        if len(names) == 1 and isinstance(names[0], (list, tuple)):
            names = tuple(names[0])
        return Table(OrderedDict((n, self[n]) for n in names))

    def withColumn(self, name: str, col) -> "Table":
        cols = OrderedDict(self.resolve()._cols)
        cols[name] = _as_column(col)
        return Table(cols)

    def drop(self, *names) -> "Table":
        return Table(OrderedDict((k, v) for k, v in self.resolve()._cols.items() if k not in names))

    def filter_rows(self, keep: np.ndarray) -> "Table":
        return Table(OrderedDict((k, _take(c, keep)) for k, c in self.resolve()._cols.items()))

    def to_numpy(self, name: str) -> np.ndarray:
        c = self[name]
        if isinstance(c, (DeviceMatrix, AssembledColumn)):
            return c.numpy()
        if isinstance(c, Column):
            v = c.values.cpu().numpy() if isinstance(c.values, torch.Tensor) else np.asarray(c.values)
            if c.valid is not None:
                ok = c.valid.cpu().numpy() if isinstance(c.valid, torch.Tensor) else np.asarray(c.valid)
                v = np.ma.MaskedArray(v, mask=~ok)
            return v
        if isinstance(c, torch.Tensor):
            return c.cpu().numpy()
        return np.asarray(c)

    def toPandas(self):
        import pandas as pd
        data = OrderedDict()
        for k in self.columns:
            a = self.to_numpy(k)
            if isinstance(a, np.ma.MaskedArray):
                a = a.astype(object).filled(None)
            data[k] = list(a) if getattr(a, "ndim", 1) == 2 else a
        return pd.DataFrame(data)

    def __repr__(self):
        return f"Table[{self.count()} rows: {', '.join(self.columns)}]"


def features_to_device(col, device) -> DeviceMatrix:
    """A 'features' column in any accepted form -> DeviceMatrix (H2D copy if it lives on host)."""
    if isinstance(col, DeviceMatrix):
        return col
    if isinstance(col, AssembledColumn):
        return col.materialize()
    if isinstance(col, Column):
        raise TypeError("features column must be a vector column (2-D array or VectorAssembler output), got a scalar column")
    if isinstance(col, torch.Tensor):
        t = col
    else:
        t = torch.from_numpy(np.ascontiguousarray(col, dtype=np.float64))
    if t.dim() != 2:
        raise TypeError("features column must be 2-D [rows, features]")
    t = t.to(device=device, dtype=torch.float64, non_blocking=True).contiguous()
    return DeviceMatrix(t)
