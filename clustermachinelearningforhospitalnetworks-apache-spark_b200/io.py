"""Columnar ingest for the step right before the hot path: the reference reads hospital CSV uploads
with an explicit schema (ref:L64-78), keeps a time window of them (ref:L123-127: `event_time BETWEEN start AND end`)
and drops rows with nulls (ref:L128) before assembling features.  Here pyarrow does the file parsing on the host;
the resulting `Table` keeps nullable columns as values + validity, which is what `VectorAssembler` uploads.
"""
from __future__ import annotations

import datetime as _dt
from typing import Optional, Union

import numpy as np

from .table import Column, Table

# ref:L64-72 -- hospital_id:string, event_time:timestamp, three int32 counters, two doubles
HOSPITAL_SCHEMA = (
    ("hospital_id", "string"), ("event_time", "timestamp[us]"), ("admission_count", "int32"), ("current_occupancy", "int32"),
    ("emergency_visits", "int32"), ("seasonality_index", "float64"), ("length_of_stay", "float64"),
)


def _arrow_schema(schema):
    import pyarrow as pa
    return pa.schema([(n, pa.type_for_alias(t) if isinstance(t, str) and "[" not in t else (pa.timestamp("us") if isinstance(t, str) else t))
                      for n, t in schema])


def read_csv(path: str, schema=HOSPITAL_SCHEMA, header: bool = True) -> Table:
    """spark.read.option("header", True).schema(schema).csv(path) for one file or a directory of files."""
    import os
    import pyarrow as pa
    import pyarrow.csv as pacsv
    sch = _arrow_schema(schema)
    files = sorted(os.path.join(path, f) for f in os.listdir(path) if f.endswith(".csv")) if os.path.isdir(path) else [path]
    opts = pacsv.ConvertOptions(column_types=sch, strings_can_be_null=True, null_values=["", "null", "NULL", "NaN"])
    ropts = pacsv.ReadOptions(column_names=None if header else [n for n, _ in schema])
    tables = [pacsv.read_csv(f, read_options=ropts, convert_options=opts) for f in files]
    return Table.from_arrow(pa.concat_tables(tables).select([n for n, _ in schema]))


def read_parquet(path: str, columns=None) -> Table:
    import pyarrow.parquet as pq
    return Table.from_arrow(pq.read_table(path, columns=columns))


def _to_us(v: Union[str, _dt.datetime, np.datetime64, int]) -> int:
    if isinstance(v, (int, np.integer)):
        return int(v)
    return int(np.datetime64(v, "us").astype(np.int64))


def window_int64(c: Column, start, end):
    """A BETWEEN window in integer form: (values as int64, lo, hi), microseconds for timestamps; None when the column is
    neither a timestamp nor an integer column (such windows are applied on the host)."""
    vals = np.asarray(c.values)
    if vals.dtype.kind == "M":
        return vals.astype("datetime64[us]").astype(np.int64), _to_us(start), _to_us(end)
    if vals.dtype == object:   # python datetimes
        v = np.array([np.datetime64(x, "us").astype(np.int64) if x is not None else np.iinfo(np.int64).min for x in vals])
        return v, _to_us(start), _to_us(end)
    if vals.dtype.kind in "iu":
        return vals.astype(np.int64), int(start), int(end)
    return None


def between(table: Table, column: str, start, end, lazy: bool = False) -> Table:
    """WHERE column BETWEEN start AND end (inclusive on both sides, like Spark SQL; ref:L123-127).
    Null timestamps never match.  lazy=True only records the window: `VectorAssembler` then applies it (and a following
    `.na.drop()`) on the device, inside the loader's compaction pass; any other use of the table applies it on the host."""
    if lazy:
        if column not in table:
            raise KeyError(column)
        return table.with_pending(window=(column, start, end))
    c = table[column]
    if not isinstance(c, Column):
        raise TypeError(f"{column!r} is not a scalar column")
    w = window_int64(c, start, end)
    v, lo, hi = w if w is not None else (np.asarray(c.values), start, end)
    keep = (v >= lo) & (v <= hi)
    if c.valid is not None:
        keep &= np.asarray(c.valid, dtype=bool)
    return table.filter_rows(keep)
