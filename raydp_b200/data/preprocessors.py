"""``ray.data.preprocessors``: fit statistics on a dataset, transform datasets and batches.

Used between the Spark-to-dataset hand-off (reference ``python/raydp/spark/dataset.py``) and a trainer:
scale numeric columns, encode categories, impute, concatenate columns into one feature vector.
Statistics come from the dataset aggregates (:mod:`raydp_b200.data.extras`); transforms are
``map_batches`` over Arrow blocks, so they run as remote tasks like any other dataset op.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc


class PreprocessorNotFittedException(RuntimeError):
    pass


class Preprocessor:
    _is_fittable = True

    def __init__(self):
        self.stats_: Optional[Dict[str, Any]] = None

    def fit(self, ds) -> "Preprocessor":
        if self._is_fittable:
            self.stats_ = self._fit(ds)
        return self

    def fit_transform(self, ds):
        return self.fit(ds).transform(ds)

    def _check(self) -> None:
        if self._is_fittable and self.stats_ is None:
            raise PreprocessorNotFittedException("`fit` must be called before `transform`.")

    def transform(self, ds):
        self._check()
        return ds.map_batches(self._transform_table)

    def transform_batch(self, batch):
        """A pandas DataFrame, a dict of numpy arrays or an Arrow table; the same kind comes back."""
        self._check()
        import pandas as pd
        if isinstance(batch, pd.DataFrame):
            # without the pandas metadata: transformed columns keep the types the transform gave them
            table = pa.Table.from_pandas(batch, preserve_index=False).replace_schema_metadata(None)
            return self._transform_table(table).to_pandas()
        if isinstance(batch, dict):
            out = self._transform_table(pa.table({k: pa.array(list(v) if getattr(v, "ndim", 1) > 1 else v)
                                                  for k, v in batch.items()}))
            return {n: _to_numpy(out.column(n)) for n in out.column_names}
        return self._transform_table(batch)

    def _fit(self, ds) -> Dict[str, Any]:
        raise NotImplementedError

    def _transform_table(self, table: pa.Table) -> pa.Table:
        raise NotImplementedError

    def __repr__(self):
        return f"{type(self).__name__}({', '.join(f'{k}={v!r}' for k, v in vars(self).items() if k != 'stats_')})"


def _to_numpy(col) -> np.ndarray:
    arr = col.combine_chunks() if isinstance(col, pa.ChunkedArray) else col
    if pa.types.is_list(arr.type):
        return np.asarray(arr.to_pylist())
    return arr.to_numpy(zero_copy_only=False)


def _replace(table: pa.Table, name: str, values) -> pa.Table:
    values = values if isinstance(values, (pa.Array, pa.ChunkedArray)) else pa.array(values)
    if name in table.column_names:
        return table.set_column(table.column_names.index(name), name, values)
    return table.append_column(name, values)


class StandardScaler(Preprocessor):
    def __init__(self, columns: List[str], ddof: int = 0):
        super().__init__()
        self.columns, self.ddof = list(columns), ddof

    def _fit(self, ds):
        from raydp_b200.data.extras import Mean, Std
        res = ds.aggregate(*[Mean(c) for c in self.columns], *[Std(c, ddof=self.ddof) for c in self.columns])
        return {**{f"mean({c})": res[f"mean({c})"] for c in self.columns},
                **{f"std({c})": res[f"std({c})"] for c in self.columns}}

    def _transform_table(self, table):
        for c in self.columns:
            mean, std = self.stats_[f"mean({c})"], self.stats_[f"std({c})"] or 0.0
            centred = pc.subtract(pc.cast(table.column(c), pa.float64()), float(mean))
            table = _replace(table, c, pc.divide(centred, float(std)) if std else pc.multiply(centred, 0.0))
        return table


class MinMaxScaler(Preprocessor):
    def __init__(self, columns: List[str]):
        super().__init__()
        self.columns = list(columns)

    def _fit(self, ds):
        from raydp_b200.data.extras import Max, Min
        return ds.aggregate(*[Min(c) for c in self.columns], *[Max(c) for c in self.columns])

    def _transform_table(self, table):
        for c in self.columns:
            lo, hi = self.stats_[f"min({c})"], self.stats_[f"max({c})"]
            shifted = pc.subtract(pc.cast(table.column(c), pa.float64()), float(lo))
            span = float(hi) - float(lo)
            table = _replace(table, c, pc.divide(shifted, span) if span else pc.multiply(shifted, 0.0))
        return table


class MaxAbsScaler(Preprocessor):
    def __init__(self, columns: List[str]):
        super().__init__()
        self.columns = list(columns)

    def _fit(self, ds):
        from raydp_b200.data.extras import AbsMax
        return ds.aggregate(*[AbsMax(c) for c in self.columns])

    def _transform_table(self, table):
        for c in self.columns:
            top = float(self.stats_[f"absmax({c})"] or 1.0)
            table = _replace(table, c, pc.divide(pc.cast(table.column(c), pa.float64()), top))
        return table


class SimpleImputer(Preprocessor):
    def __init__(self, columns: List[str], strategy: str = "mean", fill_value=None):
        super().__init__()
        if strategy not in ("mean", "most_frequent", "constant"):
            raise ValueError(f"strategy {strategy!r} is not one of mean / most_frequent / constant")
        self.columns, self.strategy, self.fill_value = list(columns), strategy, fill_value
        self._is_fittable = strategy != "constant"
        if strategy == "constant":
            self.stats_ = {}

    def _fit(self, ds):
        from raydp_b200.data.extras import Mean
        if self.strategy == "mean":
            return ds.aggregate(*[Mean(c) for c in self.columns])
        table = ds.select_columns(self.columns).to_arrow()
        out = {}
        for c in self.columns:
            counts = pc.value_counts(table.column(c).drop_null().combine_chunks())
            best = int(np.argmax(np.asarray(pc.struct_field(counts, "counts")))) if len(counts) else None
            out[f"most_frequent({c})"] = None if best is None else pc.struct_field(counts, "values")[best].as_py()
        return out

    def _transform_table(self, table):
        for c in self.columns:
            fill = self.fill_value if self.strategy == "constant" else self.stats_[f"{self.strategy}({c})"]
            col = table.column(c)
            if pa.types.is_null(col.type):              # an all-missing batch carries no type
                col = pa.nulls(len(col), pa.scalar(fill).type)
            if pa.types.is_floating(col.type):
                col = pc.if_else(pc.fill_null(pc.is_nan(col), False), pa.nulls(len(col), col.type), col)
            if pa.types.is_integer(col.type) and isinstance(fill, float) and not float(fill).is_integer():
                col = pc.cast(col, pa.float64())
            table = _replace(table, c, pc.fill_null(col, pa.scalar(fill, type=col.type)))
        return table


class _CategoryEncoder(Preprocessor):
    def __init__(self, columns: List[str]):
        super().__init__()
        self.columns = list(columns)

    def _fit(self, ds):
        table = ds.select_columns(self.columns).to_arrow()
        out = {}
        for c in self.columns:
            values = sorted(v for v in pc.unique(table.column(c).combine_chunks()).to_pylist() if v is not None)
            out[f"unique_values({c})"] = {v: i for i, v in enumerate(values)}
        return out

    def _codes(self, table: pa.Table, c: str) -> pa.Array:
        mapping = self.stats_[f"unique_values({c})"]
        col = table.column(c).combine_chunks()
        return pc.index_in(col, value_set=pa.array(list(mapping), type=col.type))


class OrdinalEncoder(_CategoryEncoder):
    def _transform_table(self, table):
        for c in self.columns:
            table = _replace(table, c, pc.cast(self._codes(table, c), pa.int64()))
        return table


class LabelEncoder(_CategoryEncoder):
    def __init__(self, label_column: str):
        super().__init__([label_column])
        self.label_column = label_column

    def _transform_table(self, table):
        return _replace(table, self.label_column, pc.cast(self._codes(table, self.label_column), pa.int64()))

    def inverse_transform(self, ds):
        self._check()
        values = list(self.stats_[f"unique_values({self.label_column})"])

        def back(table: pa.Table) -> pa.Table:
            return _replace(table, self.label_column, pa.array(values).take(table.column(self.label_column).combine_chunks()))
        return ds.map_batches(back)


class OneHotEncoder(_CategoryEncoder):
    def _transform_table(self, table):
        for c in self.columns:
            codes = np.asarray(pc.fill_null(self._codes(table, c), -1))
            for value, i in self.stats_[f"unique_values({c})"].items():
                table = table.append_column(f"{c}_{value}", pa.array((codes == i).astype(np.int64)))
            table = table.drop_columns([c])
        return table


class Concatenator(Preprocessor):
    """Numeric columns -> one vector column (``array<double>``), what a torch model takes as input."""

    _is_fittable = False

    def __init__(self, output_column_name: str = "concat_out", include: Optional[List[str]] = None,
                 exclude: Optional[List[str]] = None, dtype=None, raise_if_missing: bool = False, columns=None):
        super().__init__()
        self.output_column_name, self.include, self.exclude = output_column_name, include or columns, list(exclude or [])
        self.dtype = dtype

    def _transform_table(self, table):
        names = [c for c in (self.include or table.column_names) if c not in self.exclude]
        width = len(names)
        m = np.stack([pc.cast(table.column(c), pa.float64()).to_numpy(zero_copy_only=False) for c in names], axis=1) \
            if table.num_rows else np.zeros((0, width))
        offsets = pa.array(np.arange(0, (len(m) + 1) * width, width, dtype=np.int32))
        vec = pa.ListArray.from_arrays(offsets, pa.array(m.reshape(-1)))
        return table.drop_columns(names).append_column(self.output_column_name, vec)


class BatchMapper(Preprocessor):
    _is_fittable = False

    def __init__(self, fn: Callable, batch_format: str = "pandas", batch_size=None):
        super().__init__()
        self.fn, self.batch_format = fn, batch_format

    def transform(self, ds):
        return ds.map_batches(self.fn, batch_format=self.batch_format)

    def _transform_table(self, table):
        if self.batch_format == "pandas":
            return pa.Table.from_pandas(self.fn(table.to_pandas()), preserve_index=False)
        if self.batch_format == "numpy":
            out = self.fn({n: _to_numpy(table.column(n)) for n in table.column_names})
            return pa.table({k: pa.array(list(v) if getattr(v, "ndim", 1) > 1 else v) for k, v in out.items()})
        return self.fn(table)


class Chain(Preprocessor):
    def __init__(self, *preprocessors: Preprocessor):
        super().__init__()
        self.preprocessors = list(preprocessors)

    def fit(self, ds) -> "Chain":
        for i, p in enumerate(self.preprocessors):
            p.fit(ds)
            if i < len(self.preprocessors) - 1:
                ds = p.transform(ds)
        self.stats_ = {}
        return self

    def transform(self, ds):
        for p in self.preprocessors:
            ds = p.transform(ds)
        return ds

    def _transform_table(self, table):
        for p in self.preprocessors:
            table = p._transform_table(table)
        return table
