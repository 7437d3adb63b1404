"""PySpark-shaped DataFrame over partitioned Arrow tables.

Parity: the DataFrame surface the reference exercises: ``withColumn,
filter, select, drop, randomSplit, repartition, sortWithinPartitions, cache/persist/unpersist,
count, take, collect, show, schema, rdd.getNumPartitions, groupBy().count(), fillna, join,
withColumnRenamed, agg, mapInPandas, write.parquet`` (reference ``examples/data_process.py``,
``python/raydp/utils.py:77``, ``spark/dataset.py:177-180,268``, ``torch/estimator.py:371``).
Narrow operators extend the plan; wide ones run a two-stage job through the session backend.
"""
from __future__ import annotations

import uuid
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple, Union

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql import types as T
from raydp_b200.sql.column import Column, as_column
from raydp_b200.sql.dataframe_ext import DataFrameExtras
from raydp_b200.sql.plan import LocalSource, Plan, RefSource


class StorageLevel:
    """Accepted for API compatibility; blocks always live in executor memory / shared memory."""

    def __init__(self, name: str):
        self.name = name

    def __repr__(self):
        return f"StorageLevel.{self.name}"


for _n in ("MEMORY_ONLY", "MEMORY_AND_DISK", "DISK_ONLY", "MEMORY_ONLY_2", "MEMORY_AND_DISK_2",
           "OFF_HEAP", "MEMORY_AND_DISK_DESER"):
    setattr(StorageLevel, _n, StorageLevel(_n))


class DataFrame(DataFrameExtras):
    def __init__(self, session, plan: Plan):
        self._session = session
        self._plan = plan
        self._cache_pending = False
        self._cached_refs: Optional[list] = None
        self._lineage: Optional[Plan] = None
        self._schema_cache: Optional[pa.Schema] = None

    # ------------------------------------------------------------------ plumbing
    @property
    def sparkSession(self):
        return self._session

    @property
    def sql_ctx(self):
        return _SQLContextShim(self._session)

    def _derive(self, op: tuple) -> "DataFrame":
        self._materialise_cache()
        return DataFrame(self._session, self._plan.with_op(op))

    def _materialise_cache(self) -> None:
        if not self._cache_pending:
            return
        self._cache_pending = False
        results = self._session._backend.run(self._plan, "put", {})
        schema = self._plan.schema()
        self._lineage = self._plan
        self._cached_refs = [[r[0]] for r in results]
        self._plan = Plan(RefSource(self._cached_refs, schema))

    def _run(self, action: str, args: Optional[dict] = None, plan: Optional[Plan] = None) -> list:
        self._materialise_cache()
        return self._session._backend.run(plan or self._plan, action, args or {})

    def _arrow_schema(self) -> pa.Schema:
        if self._schema_cache is None:
            self._schema_cache = self._plan.schema()
        return self._schema_cache

    # ------------------------------------------------------------------ metadata
    @property
    def schema(self) -> T.StructType:
        return T.from_arrow_schema(self._arrow_schema())

    @property
    def columns(self) -> List[str]:
        return list(self._arrow_schema().names)

    @property
    def dtypes(self) -> List[Tuple[str, str]]:
        return [(f.name, f.dataType.simpleString()) for f in self.schema]

    def printSchema(self) -> None:
        print("root")
        for f in self.schema:
            print(f" |-- {f.name}: {f.dataType.simpleString()} (nullable = {str(f.nullable).lower()})")

    @property
    def rdd(self):
        from raydp_b200.sql.rdd import DataFrameRDD
        return DataFrameRDD(self)

    def __getitem__(self, item):
        if isinstance(item, str):
            if item not in self.columns:
                raise KeyError(f"cannot resolve column '{item}' given input columns {self.columns}")
            return Column("col", (item,))
        if isinstance(item, Column):
            return self.filter(item)
        if isinstance(item, (list, tuple)):
            return self.select(*item)
        raise TypeError(f"unexpected item type: {type(item)}")

    def __getattr__(self, name: str):
        if name.startswith("_"):
            raise AttributeError(name)
        if name in self.columns:
            return Column("col", (name,))
        raise AttributeError(f"'DataFrame' object has no attribute '{name}'")

    def __repr__(self):
        return "DataFrame[" + ", ".join(f"{n}: {t}" for n, t in self.dtypes) + "]"

    # ------------------------------------------------------------------ narrow transforms
    def withColumn(self, name: str, col: Column) -> "DataFrame":
        if not isinstance(col, Column):
            raise TypeError("col should be Column")
        from raydp_b200.sql import windowing
        if windowing.has_window(col):
            found: list = []
            rewritten = windowing.extract_windows(col, "__win_", found)
            if rewritten._kind == "col" and len(found) == 1:          # the window expression IS the column
                return self._with_windows([(name, found[0][1])])
            return self._with_windows(found)._derive(("with_column", name, rewritten)).drop(*[t for t, _ in found])
        if col._kind in ("posexplode", "explode", "inline"):
            others = [Column("col", (c,)) for c in self.columns if c != name]
            return self.select(*others, col.alias(name) if col._kind == "explode" else col)
        return self._derive(("with_column", name, col))

    def withColumns(self, cols: Dict[str, Column]) -> "DataFrame":
        df = self
        for k, v in cols.items():
            df = df.withColumn(k, v)
        return df

    def withColumnRenamed(self, existing: str, new: str) -> "DataFrame":
        return self._derive(("rename", existing, new))

    def toDF(self, *names: str) -> "DataFrame":
        if len(names) != len(self.columns):
            raise ValueError("number of names does not match number of columns")
        return self._derive(("rename_all", tuple(names)))

    def filter(self, condition: Union[Column, str]) -> "DataFrame":
        if isinstance(condition, str):
            from raydp_b200.sql.sqlparser import parse_expression
            condition = parse_expression(condition)
        return self._derive(("filter", condition))

    where = filter

    def select(self, *cols) -> "DataFrame":
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        exprs = [as_column(c) for c in cols]
        exprs = self._expand_star(exprs)
        known = self.columns
        exprs = [e.alias(e._args[0].rsplit(".", 1)[1])          # select("s.a") / select("t.k") shows up as a / k
                 if e._kind == "col" and e._name is None and "." in e._args[0] and e._args[0] not in known else e
                 for e in exprs]
        gens = [i for i, e in enumerate(exprs) if e._kind in ("posexplode", "explode", "inline")]
        if len(gens) > 1:
            raise ValueError("only one generator (explode / posexplode) is allowed per select")
        from raydp_b200.sql import windowing
        if any(windowing.has_window(e) for e in exprs):
            found: list = []
            rewritten = [windowing.extract_windows(e, "__win_", found) for e in exprs]
            rewritten = [r.alias(e._output_name()) if r._name is None and windowing.has_window(e) else r
                         for r, e in zip(rewritten, exprs)]
            return self._with_windows(found).select(*rewritten)
        if gens:
            return self._derive(("generate", exprs, gens[0]))
        if any(_contains_agg(e) for e in exprs):
            return self.agg(*exprs)
        return self._derive(("select", exprs))

    def _expand_star(self, exprs: List[Column]) -> List[Column]:
        out: List[Column] = []
        for e in exprs:
            if e._kind == "col" and e._args[0] == "*":
                out.extend(Column("col", (c,)) for c in self.columns)
            elif e._kind == "col" and e._args[0].endswith(".*"):
                # struct expansion: select("s.*")
                base = e._args[0][:-2]
                field_type = self._arrow_schema().field(base).type
                out.extend(Column("col", (base,)).getField(f.name).alias(f.name) for f in field_type)
            else:
                out.append(e)
        return out

    def _with_windows(self, requests) -> "DataFrame":
        from raydp_b200.sql import windowing
        table = windowing.evaluate_windows(self._collect_table(), list(requests))
        return self._session._from_tables(_split_table(table, max(1, self.rdd.getNumPartitions())))

    def selectExpr(self, *exprs) -> "DataFrame":
        from raydp_b200.sql.sqlparser import parse_expression
        if len(exprs) == 1 and isinstance(exprs[0], (list, tuple)):
            exprs = tuple(exprs[0])
        cols: list = []
        for e in exprs:
            cols.extend(self.columns if e.strip() == "*" else [parse_expression(e)])
        return self.select(*cols)

    def createOrReplaceTempView(self, name: str) -> None:
        self._session.catalog.register_view(name, self)

    createTempView = createOrReplaceTempView
    registerTempTable = createOrReplaceTempView

    def drop(self, *cols) -> "DataFrame":
        names = [c if isinstance(c, str) else c._output_name() for c in cols]
        return self._derive(("drop", names))

    def fillna(self, value, subset: Optional[Union[str, Sequence[str]]] = None) -> "DataFrame":
        if isinstance(subset, str):
            subset = [subset]
        return self._derive(("fillna", value, list(subset) if subset is not None else None))

    @property
    def na(self):
        return _NAFunctions(self)

    def dropna(self, how: str = "any", thresh=None, subset=None) -> "DataFrame":
        if isinstance(subset, str):
            subset = [subset]
        return self._derive(("dropna", list(subset) if subset else None))

    def sortWithinPartitions(self, *cols, ascending: Union[bool, List[bool]] = True) -> "DataFrame":
        from raydp_b200.sql.windowing import sort_spec
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        asc = ascending if isinstance(ascending, (list, tuple)) else [ascending] * len(cols)
        return self._derive(("sort_within", [sort_spec(c, bool(a)) for c, a in zip(cols, asc)]))

    def sample(self, withReplacement=None, fraction=None, seed=None) -> "DataFrame":
        if isinstance(withReplacement, float) and fraction is None:
            fraction, withReplacement = withReplacement, False
        seed = int(seed) if seed is not None else int(np.random.SeedSequence().entropy % (2 ** 31))
        return self._derive(("sample", float(fraction), seed))

    def randomSplit(self, weights: List[float], seed: Optional[int] = None) -> List["DataFrame"]:
        if any(w < 0 for w in weights):
            raise ValueError("Weights must be positive. Found weight value: %s" % min(weights))
        seed = int(seed) if seed is not None else int(np.random.SeedSequence().entropy % (2 ** 31))
        total = float(sum(weights))
        bounds = np.concatenate([[0.0], np.cumsum([w / total for w in weights])])
        bounds[-1] = 1.0 + 1e-12
        return [self._derive(("split", float(bounds[i]), float(bounds[i + 1]), seed))
                for i in range(len(weights))]

    def mapInPandas(self, func, schema) -> "DataFrame":
        if isinstance(schema, str):
            schema = T.parse_ddl(schema)
        arrow_schema = schema.to_arrow_schema() if isinstance(schema, T.StructType) else schema
        return self._derive(("map_pandas", func, arrow_schema))

    def mapInArrow(self, func, schema=None) -> "DataFrame":
        return self._derive(("map_arrow", func))

    def distinct(self) -> "DataFrame":
        keys = self.columns
        return self.repartition(self.rdd.getNumPartitions(), *keys)._derive(("distinct",))

    def dropDuplicates(self, subset=None) -> "DataFrame":
        if subset:
            keys = [subset] if isinstance(subset, str) else list(subset)
            return self.repartition(self.rdd.getNumPartitions(), *keys)._derive(("distinct_subset", tuple(keys)))
        return self.distinct()

    drop_duplicates = dropDuplicates

    def unionByName(self, other: "DataFrame", allowMissingColumns: bool = False) -> "DataFrame":
        mine, theirs = self.columns, other.columns
        if allowMissingColumns:
            from raydp_b200.sql import functions as F
            for c in mine:
                if c not in theirs:
                    other = other.withColumn(c, F.lit(None))
            for c in theirs:
                if c not in mine:
                    self = self.withColumn(c, F.lit(None))     # noqa: PLW0642
            mine = self.columns
        elif set(mine) != set(theirs):
            raise ValueError(f"unionByName: column sets differ: {mine} vs {theirs}")
        return self.union(other.select(*mine))

    def limit(self, n: int) -> "DataFrame":
        table = self._collect_table(limit=n)
        return self._session._from_tables([table.slice(0, n)])

    # ------------------------------------------------------------------ wide transforms
    def repartition(self, numPartitions, *cols) -> "DataFrame":
        if not isinstance(numPartitions, int):
            cols = (numPartitions,) + cols
            numPartitions = self.rdd.getNumPartitions()
        keys = [c if isinstance(c, str) else c._output_name() for c in cols]
        schema = self._arrow_schema()
        import uuid
        pieces = self._run("shuffle_map", {"n_out": numPartitions, "keys": keys or None,
                                           "shuffle_id": uuid.uuid4().hex[:12]})
        refs = [[row[d] for row in pieces if row[d] is not None] for d in range(numPartitions)]
        out = DataFrame(self._session, Plan(RefSource(refs, schema)))
        out._cached_refs = refs
        self._session._track_refs([r for part in refs for r in part])
        return out

    def coalesce(self, numPartitions: int) -> "DataFrame":
        if numPartitions >= self.rdd.getNumPartitions():
            return self
        return self.repartition(numPartitions)

    def groupBy(self, *cols) -> "GroupedData":
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        df, keys = self, []
        for c in cols:
            if isinstance(c, str) or c._kind == "col":
                name = c if isinstance(c, str) else c._args[0]
                if name not in df.columns and "." in name and name.split(".", 1)[1] in df.columns:
                    name = name.split(".", 1)[1]            # alias.column of a SQL query
                keys.append(name)
            else:                                   # groupBy(F.year("ts"), (col("a") % 2).alias("odd"))
                name = c._output_name()
                df = df.withColumn(name, c)
                keys.append(name)
        return GroupedData(df, keys)

    groupby = groupBy

    def agg(self, *exprs) -> "DataFrame":
        return GroupedData(self, []).agg(*exprs)

    def join(self, other: "DataFrame", on=None, how: str = "inner") -> "DataFrame":
        """Broadcast-hash join: the right side is collected once and shipped with the plan.  ``on`` is a
        column name, a list of names, or an equality expression ``left.a == right.b`` (conjunctions
        with ``&`` allowed).  Right / full outer joins first gather the left side into one partition
        (unmatched right rows must be emitted exactly once)."""
        how = how.lower()
        if on is None or how == "cross":
            return self.crossJoin(other)
        if isinstance(on, (list, tuple)) and on and all(isinstance(c, Column) and c._kind != "col" for c in on):
            combined = on[0]
            for c in on[1:]:
                combined = combined & c
            on = combined
        if isinstance(on, Column):
            try:
                pairs = _equi_join_pairs(on)
            except NotImplementedError:
                return self._theta_join(other, on, how)
            la, ra = getattr(self, "_alias", None), getattr(other, "_alias", None)
            if (la or ra) and any("." in n for pair in pairs for n in pair):
                return self._join_aliased(other, pairs, how, la, ra)
            keys, right_keys = [p[0] for p in pairs], [p[1] for p in pairs]
        else:
            keys = [on] if isinstance(on, str) else [k if isinstance(k, str) else k._output_name() for k in on]
            right_keys = keys
        left = self
        if how in ("right", "right_outer", "rightouter", "outer", "full", "full_outer", "fullouter"):
            left = self.coalesce(1)
        right = other._collect_table()
        how = {"rightouter": "right_outer", "fullouter": "full_outer"}.get(how, how)
        return left._derive(("join", right, keys, how, right_keys))

    def _join_aliased(self, other: "DataFrame", pairs, how: str, la: Optional[str], ra: Optional[str]) -> "DataFrame":
        """``df.alias("a").join(df.alias("b"), col("a.k") == col("b.k"))`` (self-joins included): qualified key
        names are assigned to their side by alias; right-hand columns that collide with left-hand ones stay
        reachable as ``<right alias>.<column>``, left-hand ones as ``<left alias>.<column>`` or bare."""
        def side(name: str):
            if "." in name:
                qual, bare = name.split(".", 1)
                if qual == la and bare in self.columns:
                    return "l", bare
                if qual == ra and bare in other.columns:
                    return "r", bare
            if name in self.columns and name not in other.columns:
                return "l", name
            if name in other.columns and name not in self.columns:
                return "r", name
            raise ValueError(f"join key {name!r} is ambiguous or unknown: qualify it with "
                             f"{la or '<left alias>'}. / {ra or '<right alias>'}.")
        left_keys, right_keys = [], []
        for a, b in pairs:
            (sa, na), (sb, nb) = side(a), side(b)
            if sa == sb:
                raise ValueError(f"join condition {a} == {b} does not relate the two sides")
            left_keys.append(na if sa == "l" else nb)
            right_keys.append(nb if sa == "l" else na)
        tag = ra or "right"
        renamed = other
        semi = how in ("left_semi", "leftsemi", "semi", "left_anti", "leftanti", "anti")
        for c in other.columns:
            if c in self.columns and c not in right_keys and not semi:
                renamed = renamed.withColumnRenamed(c, f"{tag}.{c}")
        for lk, rk in zip(left_keys, right_keys):
            if lk != rk:
                if lk in renamed.columns:
                    renamed = renamed.withColumnRenamed(lk, f"{tag}.{lk}")
                renamed = renamed.withColumnRenamed(rk, lk)
        how = {"semi": "left_semi", "anti": "left_anti"}.get(how, how)
        out = self.join(renamed, left_keys, how)
        out._alias = None
        return out

    def _theta_join(self, other: "DataFrame", condition: Column, how: str) -> "DataFrame":
        """A join on an arbitrary predicate (``a.x > b.y``, ranges, mixed conditions): the pairs are formed
        by a cross join and filtered; outer variants add back the rows that found no partner."""
        from raydp_b200.sql import functions as F
        how = {"leftouter": "left", "left_outer": "left", "leftsemi": "left_semi", "leftanti": "left_anti",
               "semi": "left_semi", "anti": "left_anti"}.get(how, how)
        if how not in ("inner", "left", "left_semi", "left_anti"):
            raise NotImplementedError(f"{how} joins need equality conditions; a non-equality predicate is supported "
                                      "for inner, left, left_semi and left_anti joins")
        clash = [c for c in other.columns if c in self.columns]
        if clash:
            raise ValueError(f"both sides of a non-equality join have columns {clash}: rename one side first")
        tagged = self.withColumn("__lid", F.monotonically_increasing_id())
        matched = tagged.crossJoin(other).filter(condition)
        if how == "inner":
            return matched.drop("__lid")
        hit_ids = [r[0] for r in matched.select("__lid").distinct().collect()]
        if how == "left_semi":
            return tagged.filter(F.col("__lid").isin(hit_ids)).drop("__lid")
        lonely = tagged.filter(~F.col("__lid").isin(hit_ids)) if hit_ids else tagged
        if how == "left_anti":
            return lonely.drop("__lid")
        right_schema = other._arrow_schema()
        for f in right_schema:
            lonely = lonely.withColumn(f.name, F.lit(None).cast(T.from_arrow_type(f.type)))
        return matched.union(lonely.select(*matched.columns)).drop("__lid")

    def crossJoin(self, other: "DataFrame") -> "DataFrame":
        from raydp_b200.sql import functions as F
        key = "__cross_key"
        joined = self.withColumn(key, F.lit(1))._derive(
            ("join", other.withColumn(key, F.lit(1))._collect_table(), [key], "inner", [key]))
        return joined.drop(key)

    def union(self, other: "DataFrame") -> "DataFrame":
        tables = self._run("collect") + other._run("collect")
        schema = self._arrow_schema()
        theirs = other._arrow_schema()
        if len(theirs) != len(schema):
            raise ValueError(f"Union can only be performed on tables with the same number of columns, but the first "
                             f"table has {len(schema)} columns and the second table has {len(theirs)} columns")
        fields = []
        for mine, yours in zip(schema, theirs):              # positional; numeric columns widen like Spark's
            t = mine.type
            if mine.type != yours.type:
                try:
                    t = pa.unify_schemas([pa.schema([pa.field("c", mine.type)]), pa.schema([pa.field("c", yours.type)])],
                                         promote_options="permissive").field("c").type
                except (pa.ArrowInvalid, pa.ArrowTypeError):
                    t = mine.type
            fields.append(pa.field(mine.name, t))
        schema = pa.schema(fields)
        tables = [t.rename_columns(schema.names).cast(schema, safe=False) for t in tables]
        return self._session._from_tables(tables)

    unionAll = union

    def orderBy(self, *cols, ascending: Union[bool, List[bool]] = True) -> "DataFrame":
        from raydp_b200.sql.windowing import sort_table
        table = sort_table(self._collect_table(), cols, ascending)
        n = max(1, self.rdd.getNumPartitions())
        return self._session._from_tables(_split_table(table, n))

    sort = orderBy

    # ------------------------------------------------------------------ persistence
    def cache(self) -> "DataFrame":
        return self.persist()

    def persist(self, storageLevel: Optional[StorageLevel] = None) -> "DataFrame":
        if self._cached_refs is None:
            self._cache_pending = True
        return self

    def unpersist(self, blocking: bool = False) -> "DataFrame":
        self._cache_pending = False
        if self._cached_refs is not None and self._lineage is not None:
            from raydp_b200 import runtime
            runtime.free([r for part in self._cached_refs for r in part])
            self._plan = self._lineage
            self._cached_refs = None
            self._lineage = None
        return self

    @property
    def is_cached(self) -> bool:
        return self._cache_pending or self._cached_refs is not None

    @property
    def storageLevel(self) -> StorageLevel:
        return StorageLevel.MEMORY_AND_DISK if self.is_cached else StorageLevel("NONE")

    # ------------------------------------------------------------------ actions
    def count(self) -> int:
        return int(sum(self._run("count")))

    def _collect_table(self, limit: Optional[int] = None) -> pa.Table:
        schema = self._arrow_schema()
        if limit is not None:
            tables = self._run("take", {"n": limit})
        else:
            tables = self._run("collect")
        tables = [t for t in tables if t.num_rows] or [schema.empty_table()]
        return pa.concat_tables([t.cast(schema, safe=False) if t.schema != schema else t
                                 for t in tables]) if len(tables) > 1 else tables[0]

    def toArrow(self) -> pa.Table:
        return self._collect_table()

    def toPandas(self):
        return self._collect_table().to_pandas()

    def collect(self) -> List[T.Row]:
        return T.table_to_rows(self._collect_table())

    def take(self, num: int) -> List[T.Row]:
        return T.table_to_rows(self._collect_table(limit=num).slice(0, num))

    def head(self, n: Optional[int] = None):
        rows = self.take(1 if n is None else n)
        return (rows[0] if rows else None) if n is None else rows

    def first(self):
        return self.head()

    def isEmpty(self) -> bool:
        return len(self.take(1)) == 0

    def show(self, n: int = 20, truncate: bool = True, vertical: bool = False) -> None:
        table = self._collect_table(limit=n + 1)
        more = table.num_rows > n
        table = table.slice(0, n)
        names = table.column_names
        if isinstance(truncate, int) and not isinstance(truncate, bool) and truncate > 0:
            limit_chars = truncate
            cells = [[(lambda t: t[:limit_chars - 3] + "..." if len(t) > limit_chars > 3 else t[:limit_chars])(
                _fmt(v, False)) for v in c.to_pylist()] for c in table.columns]
        else:
            cells = [[_fmt(v, bool(truncate)) for v in c.to_pylist()] for c in table.columns]
        if vertical:
            width = max([len(nm) for nm in names] + [1])
            for i in range(table.num_rows):
                print(f"-RECORD {i}" + "-" * 20)
                for nm, col in zip(names, cells):
                    print(f" {nm.ljust(width)} | {col[i]}")
            if more:
                print(f"only showing top {n} row{'s' if n != 1 else ''}")
            return
        widths = [max([len(nm)] + [len(v) for v in col]) for nm, col in zip(names, cells)]
        line = "+" + "+".join("-" * w for w in widths) + "+"
        print(line)
        print("|" + "|".join(nm.rjust(w) for nm, w in zip(names, widths)) + "|")
        print(line)
        for i in range(table.num_rows):
            print("|" + "|".join(col[i].rjust(w) for col, w in zip(cells, widths)) + "|")
        print(line)
        if more:
            print(f"only showing top {n} row{'s' if n != 1 else ''}")

    def foreachPartition(self, f) -> None:
        def run(table, _f=f):
            _f(iter(T.table_to_rows(table)))
            return None
        self._run("py_partition", {"fn": run})

    def describe(self, *cols) -> "DataFrame":
        table = self._collect_table()
        names = list(cols) or [f.name for f in table.schema
                               if pa.types.is_integer(f.type) or pa.types.is_floating(f.type)]
        rows = {"summary": ["count", "mean", "stddev", "min", "max"]}
        for nme in names:
            c = table.column(nme)
            rows[nme] = [str(pc.count(c).as_py()), str(pc.mean(c).as_py()),
                         str(pc.stddev(c, ddof=1).as_py()), str(pc.min(c).as_py()),
                         str(pc.max(c).as_py())]
        return self._session._from_tables([pa.table(rows)])

    @property
    def write(self) -> "DataFrameWriter":
        return DataFrameWriter(self)

    def to_pandas_on_spark(self, index_col=None):
        from raydp_b200.sql import pandas_api
        return pandas_api.DataFrame(self)

    pandas_api = to_pandas_on_spark
    to_koalas = to_pandas_on_spark


class _SQLContextShim:
    def __init__(self, session):
        self.sparkSession = session


class _NAFunctions:
    def __init__(self, df: DataFrame):
        self._df = df

    def replace(self, to_replace, value=None, subset=None):
        return self._df.replace(to_replace, value, subset)

    def fill(self, value, subset=None):
        return self._df.fillna(value, subset)

    def drop(self, how="any", thresh=None, subset=None):
        return self._df.dropna(how, thresh, subset)


def _show_text(v) -> str:
    if v is None:
        return "null"
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, dict):                             # a struct (maps come out of Arrow as lists of pairs)
        return "{" + ", ".join(_show_text(x) for x in v.values()) + "}"
    if isinstance(v, list) and v and all(isinstance(x, tuple) and len(x) == 2 for x in v):
        return "{" + ", ".join(f"{_show_text(k)} -> {_show_text(x)}" for k, x in v) + "}"
    if isinstance(v, (list, tuple)):
        return "[" + ", ".join(_show_text(x) for x in v) + "]"
    return str(v)


def _fmt(v, truncate: bool) -> str:
    s = _show_text(v)
    return s[:17] + "..." if truncate and len(s) > 20 else s


def _equi_join_pairs(cond: Column) -> List[Tuple[str, str]]:
    """``(a == b) & (c == d)`` -> ``[(a, b), (c, d)]``; anything else is not an equi-join."""
    if cond._kind == "binop" and cond._args[0] == "&":
        return _equi_join_pairs(cond._args[1]) + _equi_join_pairs(cond._args[2])
    if cond._kind == "binop" and cond._args[0] == "==" and all(
            isinstance(a, Column) and a._kind == "col" for a in cond._args[1:3]):
        return [(cond._args[1]._args[0], cond._args[2]._args[0])]
    raise NotImplementedError("join conditions must be equalities between columns (optionally combined with &)")


def _sort_keys(cols, ascending) -> List[Tuple[str, str]]:
    if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
        cols = tuple(cols[0])
    asc = ascending if isinstance(ascending, (list, tuple)) else [ascending] * len(cols)
    keys = []
    for c, a in zip(cols, asc):
        if isinstance(c, Column) and c._kind == "sort":
            keys.append((c._args[0]._output_name(), c._args[1]))
        else:
            name = c if isinstance(c, str) else c._output_name()
            keys.append((name, "ascending" if a else "descending"))
    return keys


def _split_table(table: pa.Table, n: int) -> List[pa.Table]:
    n = max(1, min(n, max(1, table.num_rows)))
    per, rem = divmod(table.num_rows, n)
    out, off = [], 0
    for i in range(n):
        size = per + (1 if i < rem else 0)
        out.append(table.slice(off, size))
        off += size
    return out


def _contains_agg(e) -> bool:
    if isinstance(e, Column):
        if e._kind == "agg":
            return True
        if e._kind == "window":
            return False
        return any(_contains_agg(a) for a in e._args)
    if isinstance(e, (tuple, list)):
        return any(_contains_agg(a) for a in e)
    return False


class GroupedData:
    def __init__(self, df: DataFrame, keys: List[str]):
        self._df = df
        self._keys = keys

    def count(self) -> DataFrame:
        return self.agg(Column("agg", ("count", None), "count"))

    def _simple(self, fn: str, cols) -> DataFrame:
        if not cols:
            schema = self._df._arrow_schema()
            cols = [f.name for f in schema if f.name not in self._keys and
                    (pa.types.is_integer(f.type) or pa.types.is_floating(f.type))]
        spark_fn = {"mean": "avg"}.get(fn, fn)
        return self.agg(*[Column("agg", (fn, as_column(c)), f"{spark_fn}({c})") for c in cols])

    def sum(self, *cols): return self._simple("sum", cols)
    def avg(self, *cols): return self._simple("mean", cols)
    mean = avg
    def min(self, *cols): return self._simple("min", cols)
    def max(self, *cols): return self._simple("max", cols)

    def agg(self, *exprs) -> DataFrame:
        if len(exprs) == 1 and isinstance(exprs[0], dict):
            exprs = tuple(Column("agg", ({"avg": "mean"}.get(fn, fn), None if c == "*" else as_column(c)),
                                 f"{fn}({c})") for c, fn in exprs[0].items())
        if len(exprs) == 1 and isinstance(exprs[0], (list, tuple)):
            exprs = tuple(exprs[0])
        # struct keys (F.window(...) is one) are grouped by their fields and put back together afterwards:
        # Arrow's group-by hashes flat columns only
        key_schema = self._df._arrow_schema()
        struct_keys = {k: [f.name for f in key_schema.field(k).type] for k in self._keys
                       if pa.types.is_struct(key_schema.field(k).type)}
        if struct_keys:
            from raydp_b200.sql import functions as F
            flat_df, flat_keys = self._df, []
            for k in self._keys:
                if k in struct_keys:
                    for f in struct_keys[k]:
                        flat_df = flat_df.withColumn(f"__sk_{k}_{f}", Column("col", (k,)).getField(f))
                        flat_keys.append(f"__sk_{k}_{f}")
                else:
                    flat_keys.append(k)
            grouped = GroupedData(flat_df, flat_keys).agg(*exprs)
            rebuilt = []
            for k in self._keys:
                if k in struct_keys:
                    rebuilt.append(F.struct(*[Column("col", (f"__sk_{k}_{f}",)).alias(f) for f in struct_keys[k]]).alias(k))
                else:
                    rebuilt.append(Column("col", (k,)))
            rest = [Column("col", (c,)) for c in grouped.columns if c not in flat_keys]
            return grouped.select(*rebuilt, *rest)
        # an output may be an expression OVER aggregates (sum(a) / count(b), round(avg(x), 2)): the aggregate
        # nodes are computed first, the surrounding expression is applied to the aggregated table
        leaves: List[Column] = []
        outputs: List[Tuple[str, Column]] = []
        for e in exprs:
            if not isinstance(e, Column):
                raise TypeError("agg() expects aggregate expressions such as count('*')")
            rewritten = _extract_aggregates(e, leaves)
            if not leaves:
                raise TypeError("agg() expects aggregate expressions such as count('*')")
            outputs.append((e._name or _agg_default_name(e), rewritten))
        df = self._df
        aggs: List[Tuple[str, Optional[str]]] = []
        params: List[tuple] = []
        in_types: List[Optional[pa.DataType]] = []
        base_schema = df._arrow_schema()
        numeric_only = {"sum", "mean", "stddev", "variance", "stddev_pop", "var_pop", "skewness", "kurtosis", "median",
                        "product", "sum_distinct", "percentile"}
        for i, leaf in enumerate(leaves):
            fn, target = leaf._args[0], leaf._args[1]
            if target is not None and target._kind == "col" and fn in numeric_only and \
                    target._args[0] in base_schema.names and pa.types.is_string(base_schema.field(target._args[0]).type):
                target = target.cast("double")          # Spark's implicit cast: unparsable text counts as NULL
            if target is None:
                col_name = None
            elif target._kind == "col":
                col_name = target._args[0]
            else:
                col_name = f"__agg_in_{i}"
                df = df.withColumn(col_name, target)
            if fn == "count_distinct" and target is not None and target._kind == "call" and target._args[0] == "struct":
                fn = "count_distinct_rows"            # structs have no hash kernel: gathered instead
            aggs.append((fn, col_name))
            params.append(tuple(leaf._args[2:]))
        schema = df._arrow_schema()
        in_types = [None if c is None else schema.field(c).type for _, c in aggs]
        names = [f"__agg_out_{i}" for i in range(len(leaves))]
        partials = df._run("agg_partial", {"keys": self._keys, "aggs": list(dict.fromkeys(aggs))})
        merged = pa.concat_tables(partials, promote_options="default")
        result = _final_agg(merged, self._keys, aggs, names, params, in_types)
        out_cols = [result.column(k) for k in self._keys]
        out_names = list(self._keys)
        for name, expr in outputs:
            out_cols.append(expr.evaluate_array(result))
            out_names.append(name)
        return df._session._from_tables([pa.table(out_cols, names=out_names)])

    def pivot(self, pivot_col: str, values: Optional[Sequence[Any]] = None) -> "PivotedData":
        return PivotedData(self, pivot_col, values)

    def applyInPandas(self, func, schema) -> DataFrame:
        """Each group as one pandas DataFrame through ``func``; ``schema`` (DDL string or StructType) types
        the output."""
        import pandas as pd
        struct = T.parse_ddl(schema) if isinstance(schema, str) else schema
        arrow_schema = pa.schema([(f.name, f.dataType.arrow) for f in struct.fields])
        table = self._df._collect_table()
        pdf = table.to_pandas()
        pieces = []
        if self._keys:
            for _, group in pdf.groupby(self._keys, sort=True, dropna=False):
                pieces.append(func(group.reset_index(drop=True)))
        else:
            pieces.append(func(pdf))
        pieces = [p for p in pieces if p is not None and len(p)]
        if not pieces:
            return self._df._session._from_tables([arrow_schema.empty_table()])
        out = pd.concat(pieces, ignore_index=True)
        result = pa.Table.from_pandas(out[[f.name for f in struct.fields]], schema=arrow_schema, preserve_index=False)
        return self._df._session._from_tables(_split_table(result, max(1, self._df.rdd.getNumPartitions())))

    apply = applyInPandas

    def cogroup(self, other: "GroupedData") -> "CoGroupedData":
        return CoGroupedData(self, other)


class GroupingSets:
    """``rollup`` / ``cube``: the aggregation runs once per grouping set; keys outside a set are null."""

    def __init__(self, df: DataFrame, keys: List[str], sets: List[List[str]]):
        self._df, self._keys, self._sets = df, keys, sets

    def agg(self, *exprs) -> DataFrame:
        schema = self._df._arrow_schema()
        pieces = []
        for subset in self._sets:
            part = GroupedData(self._df, list(subset)).agg(*exprs)._collect_table()
            cols, names = [], []
            for k in self._keys:
                names.append(k)
                cols.append(part.column(k) if k in subset else pa.nulls(part.num_rows, schema.field(k).type))
            for nme in part.column_names:
                if nme not in subset:
                    names.append(nme)
                    cols.append(part.column(nme))
            pieces.append(pa.table(cols, names=names))
        return self._df._session._from_tables([pa.concat_tables(pieces, promote_options="default")])

    def count(self): return self.agg(Column("agg", ("count", None), "count"))
    def sum(self, *cols): return self.agg(*[Column("agg", ("sum", as_column(c)), f"sum({c})") for c in cols])
    def avg(self, *cols): return self.agg(*[Column("agg", ("mean", as_column(c)), f"avg({c})") for c in cols])
    mean = avg
    def min(self, *cols): return self.agg(*[Column("agg", ("min", as_column(c)), f"min({c})") for c in cols])
    def max(self, *cols): return self.agg(*[Column("agg", ("max", as_column(c)), f"max({c})") for c in cols])


class CoGroupedData:
    def __init__(self, left: GroupedData, right: GroupedData):
        self._left, self._right = left, right

    def applyInPandas(self, func, schema) -> DataFrame:
        import pandas as pd
        struct = T.parse_ddl(schema) if isinstance(schema, str) else schema
        arrow_schema = pa.schema([(f.name, f.dataType.arrow) for f in struct.fields])
        lp, rp = self._left._df.toPandas(), self._right._df.toPandas()
        lg = {k: g for k, g in lp.groupby(self._left._keys, sort=True, dropna=False)}
        rg = {k: g for k, g in rp.groupby(self._right._keys, sort=True, dropna=False)}
        pieces = []
        for key in sorted(set(lg) | set(rg), key=repr):
            a = lg.get(key, lp.iloc[0:0]).reset_index(drop=True)
            b = rg.get(key, rp.iloc[0:0]).reset_index(drop=True)
            pieces.append(func(a, b))
        pieces = [p for p in pieces if p is not None and len(p)]
        session = self._left._df._session
        if not pieces:
            return session._from_tables([arrow_schema.empty_table()])
        out = pd.concat(pieces, ignore_index=True)
        return session._from_tables([pa.Table.from_pandas(out[[f.name for f in struct.fields]], schema=arrow_schema,
                                                          preserve_index=False)])


class PivotedData:
    """``groupBy(keys).pivot(col[, values]).agg(...)``: one output column per pivot value (and per
    aggregate when several are given)."""

    def __init__(self, grouped: GroupedData, pivot_col: str, values):
        self._grouped, self._pivot_col, self._values = grouped, pivot_col, values

    def agg(self, *exprs) -> DataFrame:
        g, pcol = self._grouped, self._pivot_col
        values = self._values
        if values is None:
            found = g._df.select(pcol).distinct()._collect_table().column(0).to_pylist()
            values = sorted((v for v in found if v is not None), key=lambda v: (str(type(v)), v))
        base = GroupedData(g._df, g._keys + [pcol]).agg(*[e.alias(f"__pv_{i}") for i, e in enumerate(exprs)])
        table = base._collect_table()
        from raydp_b200.sql.windowing import sort_table
        if g._keys:
            table = sort_table(table, g._keys)
        key_rows = list(zip(*[table.column(k).to_pylist() for k in g._keys])) if g._keys else [()] * table.num_rows
        pivots = table.column(pcol).to_pylist()
        groups: Dict[tuple, int] = {}
        for kr in key_rows:
            groups.setdefault(kr, len(groups))
        if not g._keys:
            groups.setdefault((), 0)
        labels, cells, types = [], {}, {}
        for v in values:
            for i, e in enumerate(exprs):
                label = str(v) if len(exprs) == 1 else f"{v}_{e._name or _agg_default_name(e)}"
                labels.append(label)
                cells[label] = [None] * len(groups)
                types[label] = table.schema.field(f"__pv_{i}").type
        measures = [table.column(f"__pv_{i}").to_pylist() for i in range(len(exprs))]
        wanted = {_hashable(v) for v in values}
        for row, (kr, pv) in enumerate(zip(key_rows, pivots)):
            if _hashable(pv) not in wanted:
                continue
            for i, e in enumerate(exprs):
                label = str(pv) if len(exprs) == 1 else f"{pv}_{e._name or _agg_default_name(e)}"
                cells[label][groups[kr]] = measures[i][row]
        ordered = list(groups)
        cols = [pa.array([kr[j] for kr in ordered], type=table.schema.field(k).type) for j, k in enumerate(g._keys)]
        cols += [pa.array(cells[lb], type=types[lb]) for lb in labels]
        return g._df._session._from_tables([pa.table(cols, names=g._keys + labels)])

    def count(self): return self.agg(Column("agg", ("count", None), "count"))
    def sum(self, *cols): return self.agg(*[Column("agg", ("sum", as_column(c)), f"sum({c})") for c in cols])
    def avg(self, *cols): return self.agg(*[Column("agg", ("mean", as_column(c)), f"avg({c})") for c in cols])
    mean = avg
    def min(self, *cols): return self.agg(*[Column("agg", ("min", as_column(c)), f"min({c})") for c in cols])
    def max(self, *cols): return self.agg(*[Column("agg", ("max", as_column(c)), f"max({c})") for c in cols])


_AGG_DISPLAY = {"mean": "avg", "count_distinct": "count", "stddev": "stddev", "variance": "variance"}


def _agg_default_name(e: Column) -> str:
    if e._kind == "agg":
        fn, target = e._args[0], e._args[1]
        inner = "1" if target is None else target._output_name()
        if fn == "count_distinct":
            return f"count(DISTINCT {inner})".replace("DISTINCT ", "DISTINCT ")
        return f"{_AGG_DISPLAY.get(fn, fn)}({inner})".replace(" ", "")
    if e._kind == "binop":
        return f"({_agg_default_name(e._args[1])} {e._args[0]} {_agg_default_name(e._args[2])})"
    if e._kind == "call":
        inner = ", ".join(_agg_default_name(a) if isinstance(a, Column) else str(a) for a in e._args[1])
        return f"{e._args[0]}({inner})"
    return e._output_name()


def _extract_aggregates(expr: Column, leaves: List[Column]) -> Column:
    """Replace aggregate nodes inside ``expr`` by references to ``__agg_out_<i>`` columns."""
    if not isinstance(expr, Column):
        return expr
    if expr._kind == "agg":
        leaves.append(expr)
        return Column("col", (f"__agg_out_{len(leaves) - 1}",))
    if expr._kind in ("col", "lit"):
        return expr

    def walk(node):
        if isinstance(node, Column):
            return _extract_aggregates(node, leaves)
        if isinstance(node, tuple):
            return tuple(walk(x) for x in node)
        if isinstance(node, list):
            return [walk(x) for x in node]
        return node
    return Column(expr._kind, tuple(walk(a) for a in expr._args), expr._name)


def _final_agg(merged: pa.Table, keys: List[str], aggs, names, params=None, in_types=None) -> pa.Table:
    """Second aggregation stage: merge per-partition partial states."""
    from raydp_b200.sql.aggregates import MERGEABLE, reduce_groups
    params = params or [()] * len(aggs)
    in_types = in_types or [None] * len(aggs)

    def grouped(specs):
        if keys:
            return merged.group_by(keys, use_threads=False).aggregate(specs)
        if merged.num_rows == 0:
            return pa.TableGroupBy(merged, [], use_threads=False).aggregate(specs)
        keyed = merged.append_column("__all", pa.array(np.zeros(merged.num_rows, dtype=np.int8)))
        return keyed.group_by(["__all"], use_threads=False).aggregate(specs).drop_columns(["__all"])

    if not keys and merged.num_rows == 0:
        # a global aggregate over no rows still answers with one row: counts are 0, collections empty
        from raydp_b200.sql.aggregates import result_type
        cols = []
        for (fn, col), in_type in zip(aggs, in_types):
            in_type = in_type if in_type is not None else pa.null()
            if fn in ("count", "count_distinct", "count_distinct_rows", "approx_count_distinct", "count_if"):
                cols.append(pa.array([0], type=pa.int64()))
            elif fn in ("collect_list", "collect_set"):
                cols.append(pa.array([[]], type=pa.list_(in_type)))
            elif fn in ("sum", "min", "max", "first"):
                cols.append(pa.nulls(1, in_type))
            elif fn in MERGEABLE:
                cols.append(pa.nulls(1, pa.float64()))
            else:
                cols.append(pa.nulls(1, result_type(fn, in_type) or pa.float64()))
        return pa.table(cols, names=list(names))
    merge_specs = []
    gathered: List[str] = []
    for fn, col in aggs:
        if fn == "count":
            merge_specs.append(("count_all" if col is None else f"{col}_count", "sum"))
        elif fn == "sum":
            merge_specs.append((f"{col}_sum", "sum"))
        elif fn in ("min", "max"):
            merge_specs.append((f"{col}_{fn}", fn))
        elif fn == "mean":
            merge_specs.extend([(f"{col}_sum", "sum"), (f"{col}_count", "sum")])
        elif fn in ("stddev", "variance"):
            merge_specs.extend([(f"{col}_sum", "sum"), (f"__sq_{col}_sum", "sum"), (f"{col}_count", "sum")])
        elif fn == "first":
            merge_specs.append((f"{col}_first", "first"))
        elif fn == "count_distinct":
            if f"{col}_distinct" not in gathered:
                gathered.append(f"{col}_distinct")
        elif f"{col}_list" not in gathered:
            gathered.append(f"{col}_list")
    if gathered:
        merged = merged.append_column("__rowidx", pa.array(np.arange(merged.num_rows, dtype=np.int64)))
        merge_specs.append(("__rowidx", "list"))
    uniq = list(dict.fromkeys(merge_specs))
    g = grouped(uniq)
    if gathered:
        from raydp_b200.sql.executor import gather_by_index
        g = gather_by_index(g, "__rowidx_list", merged, gathered)       # -> <col>_list_list: list<list<T>>
    out_cols, out_names = [g.column(k) for k in keys], list(keys)
    for (fn, col), name, prm, in_type in zip(aggs, names, params, in_types):
        if fn == "count":
            src = "count_all_sum" if col is None else f"{col}_count_sum"
            out_cols.append(pc.cast(g.column(src), pa.int64()))
        elif fn == "sum":
            out_cols.append(g.column(f"{col}_sum_sum"))
        elif fn in ("min", "max"):
            out_cols.append(g.column(f"{col}_{fn}_{fn}"))
        elif fn == "mean":
            out_cols.append(pc.divide(pc.cast(g.column(f"{col}_sum_sum"), pa.float64()),
                                      pc.cast(g.column(f"{col}_count_sum"), pa.float64())))
        elif fn in ("stddev", "variance"):
            s = pc.cast(g.column(f"{col}_sum_sum"), pa.float64())
            sq = pc.cast(g.column(f"__sq_{col}_sum_sum"), pa.float64())
            cnt = pc.cast(g.column(f"{col}_count_sum"), pa.float64())
            var = pc.divide(pc.subtract(sq, pc.divide(pc.multiply(s, s), cnt)), pc.subtract(cnt, 1.0))
            out_cols.append(pc.sqrt(var) if fn == "stddev" else var)
        elif fn == "first":
            out_cols.append(g.column(f"{col}_first_first"))
        elif fn == "count_distinct":
            lists = g.column(f"{col}_distinct_list").to_pylist()
            out_cols.append(pa.array([len({_hashable(v) for sub in row for v in sub if v is not None})
                                      for row in lists], type=pa.int64()))
        else:
            assert fn not in MERGEABLE or fn == "count_distinct_rows"
            nested = g.column(f"{col}_list_list").combine_chunks()      # one inner list per partition
            inner = pc.list_flatten(nested)                             # list<T>, all partial lists in group order
            group_offsets = inner.offsets.take(nested.offsets) if len(inner) else \
                pa.array(np.zeros(len(nested) + 1, dtype=np.int32))
            flat = pa.ListArray.from_arrays(group_offsets, pc.list_flatten(inner))
            out_cols.append(reduce_groups(fn, flat, prm, flat.type.value_type))
        out_names.append(name)
    return pa.table(out_cols, names=out_names)


def _hashable(v):
    if isinstance(v, dict):
        return tuple((k, _hashable(x)) for k, x in v.items())
    if isinstance(v, list):
        return tuple(_hashable(x) for x in v)
    return v


class DataFrameWriter:
    def __init__(self, df: DataFrame):
        self._df = df
        self._mode = "errorifexists"
        self._options: Dict[str, Any] = {}

    def mode(self, saveMode: str) -> "DataFrameWriter":
        self._mode = (saveMode or "errorifexists").lower()
        return self

    def option(self, key: str, value) -> "DataFrameWriter":
        self._options[key.lower()] = value
        return self

    def format(self, fmt: str) -> "DataFrameWriter":
        self._options["__format"] = fmt
        return self

    def _prepare(self, path: str) -> Optional[str]:
        import os
        import shutil
        from raydp_b200.sql.plan import strip_scheme
        path = strip_scheme(path)
        if os.path.exists(path) and (os.path.isfile(path) or os.listdir(path)):
            if self._mode in ("overwrite",):
                shutil.rmtree(path, ignore_errors=True)
            elif self._mode in ("ignore",):
                return None
            elif self._mode in ("append",):
                pass
            else:
                raise FileExistsError(f"path {path} already exists.")
        os.makedirs(path, exist_ok=True)
        return path

    def partitionBy(self, *cols) -> "DataFrameWriter":
        cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
        self._partition_by = list(cols)
        return self

    def bucketBy(self, numBuckets: int, col, *cols) -> "DataFrameWriter":
        return self                                   # layout hint only

    def sortBy(self, col, *cols) -> "DataFrameWriter":
        return self

    def options(self, **opts) -> "DataFrameWriter":
        for k, v in opts.items():
            self.option(k, v)
        return self

    def _write(self, fmt: str, path: str, mode: Optional[str], **extra) -> None:
        import os
        if mode:
            self.mode(mode)
        target = self._prepare(path)
        if target is None:
            return
        args = {"path": target, "format": fmt, "job": uuid.uuid4().hex[:8],
                "partition_by": getattr(self, "_partition_by", None)}
        args.update({k: v for k, v in extra.items() if v is not None})
        self._df._run("write_files", args)
        open(os.path.join(target, "_SUCCESS"), "w").close()

    def parquet(self, path: str, mode: Optional[str] = None, partitionBy=None, compression: Optional[str] = None) -> None:
        if partitionBy is not None:
            self.partitionBy(partitionBy)
        self._write("parquet", path, mode, compression=compression or self._options.get("compression"))

    def orc(self, path: str, mode: Optional[str] = None, partitionBy=None, compression: Optional[str] = None) -> None:
        if partitionBy is not None:
            self.partitionBy(partitionBy)
        self._write("orc", path, mode)

    def json(self, path: str, mode: Optional[str] = None, **_kw) -> None:
        self._write("json", path, mode)

    def text(self, path: str, compression=None, lineSep=None) -> None:
        self._write("text", path, None)

    def saveAsTable(self, name: str, format: Optional[str] = None, mode: Optional[str] = None, **_kw) -> None:  # noqa: A002
        """Persist as a parquet table under the warehouse directory (see ``sqlparser.Catalog``)."""
        self._df._session.catalog.save_table(self._df, name, (mode or self._mode))

    def insertInto(self, name: str, overwrite: bool = False) -> None:
        self._df._session.catalog.save_table(self._df, name, "overwrite" if overwrite else "append")

    def csv(self, path: str, mode: Optional[str] = None, header=None, sep=None, **_kw) -> None:
        hdr = header if header is not None else self._options.get("header", False)
        hdr = str(hdr).lower() == "true" if not isinstance(hdr, bool) else hdr
        self._write("csv", path, mode, header=hdr, sep=sep or self._options.get("sep") or self._options.get("delimiter"))

    def save(self, path: Optional[str] = None, format: Optional[str] = None, mode: Optional[str] = None,  # noqa: A002
             partitionBy=None, **options) -> None:
        if format:
            self.format(format)
        if mode:
            self.mode(mode)
        if partitionBy is not None:
            self.partitionBy(partitionBy)
        self.options(**options)
        path = path or self._options.get("path")
        if path is None:
            raise ValueError("save() needs a path")
        fmt = str(self._options.get("__format", "parquet")).lower()
        if fmt not in ("parquet", "orc", "csv", "json", "text"):
            raise NotImplementedError(f"data source format {fmt!r} is not supported")
        getattr(self, fmt)(path)
