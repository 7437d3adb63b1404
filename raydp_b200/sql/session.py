"""``SparkSession`` / ``SparkContext`` of the built-in engine.

Parity: what the reference gets from pyspark after ``SparkSession.builder...master("ray://...")
.getOrCreate()`` (reference ``spark/ray_cluster.py:191-201``): ``range``, ``createDataFrame``,
``read`` (csv with header/inferSchema/sep/schema, parquet), ``conf.set/get``,
``sparkContext.{appName, applicationId, getConf, parallelize, setLogLevel, defaultParallelism}``,
``stop``.  ``enable_hive`` / ``spark.sql`` need a JVM and raise clearly.
"""
from __future__ import annotations

import os
import threading
import uuid
from typing import Any, Dict, Iterable, List, Optional, Sequence, Union

import numpy as np
import pyarrow as pa

from raydp_b200.sql import types as T
from raydp_b200.sql.backend import Backend, ClusterBackend, LocalBackend
from raydp_b200.sql.dataframe import DataFrame, _split_table
from raydp_b200.sql.plan import (CsvSource, FuncSource, LocalSource, ParquetSource, Plan,
                                 RangeSource, list_data_files)

_active_lock = threading.RLock()
_active_session: Optional["SparkSession"] = None
_live_sessions: List[Any] = []          # weak references, oldest first (the "default session" fallback)
_builder_options: Dict[str, str] = {}


class SparkConf:
    def __init__(self, values: Optional[Dict[str, str]] = None):
        self._values: Dict[str, str] = dict(values or {})

    def set(self, key: str, value) -> "SparkConf":
        self._values[key] = str(value)
        return self

    def setAll(self, pairs) -> "SparkConf":
        for k, v in pairs:
            self.set(k, v)
        return self

    def get(self, key: str, defaultValue: Optional[str] = None) -> Optional[str]:
        return self._values.get(key, defaultValue)

    def getAll(self):
        return list(self._values.items())

    def contains(self, key: str) -> bool:
        return key in self._values


class RuntimeConfig:
    def __init__(self, conf: SparkConf):
        self._conf = conf

    def set(self, key: str, value) -> None:
        self._conf.set(key, value)

    _UNSET = object()

    def get(self, key: str, default=_UNSET) -> Optional[str]:
        """Like ``spark.conf.get``: a missing key raises unless a default (``None`` included) is given."""
        if self._conf.contains(key):
            return self._conf.get(key)
        if default is RuntimeConfig._UNSET:
            raise KeyError(f"[SQL_CONF_NOT_FOUND] The SQL config \"{key}\" cannot be found.")
        return default

    def unset(self, key: str) -> None:
        self._conf._values.pop(key, None)


class SparkContext:
    def __init__(self, session: "SparkSession"):
        self._session = session
        self.appName = session._app_name
        self.applicationId = session._app_id
        self.master = session._master

    def getConf(self) -> SparkConf:
        return self._session._conf

    @property
    def defaultParallelism(self) -> int:
        return self._session._backend.default_parallelism

    def setLogLevel(self, level: str) -> None:
        self._session._conf.set("spark.log.level", level)

    def parallelize(self, data: Iterable, numSlices: Optional[int] = None):
        from raydp_b200.sql.rdd import RDD
        items = list(data)
        n = max(1, min(numSlices or self.defaultParallelism, max(1, len(items))))
        per, rem = divmod(len(items), n)
        parts, off = [], 0
        for i in range(n):
            size = per + (1 if i < rem else 0)
            parts.append(items[off:off + size])
            off += size
        return RDD(self._session, parts)

    @property
    def uiWebUrl(self):
        return None

    def stop(self) -> None:
        self._session.stop()

    @property
    def version(self) -> str:
        return self._session.version

    @property
    def defaultMinPartitions(self) -> int:
        return min(self.defaultParallelism, 2)

    @classmethod
    def getOrCreate(cls, conf: Optional[SparkConf] = None) -> "SparkContext":
        builder = SparkSession.builder
        if conf is not None:
            builder = builder.config(conf=conf)
        return builder.getOrCreate().sparkContext

    def range(self, start: int, end: Optional[int] = None, step: int = 1, numSlices: Optional[int] = None):  # noqa: A003
        if end is None:
            start, end = 0, start
        return self.parallelize(range(start, end, step), numSlices)

    def emptyRDD(self):
        from raydp_b200.sql.rdd import RDD
        return RDD(self._session, [[]])

    def union(self, rdds):
        from raydp_b200.sql.rdd import RDD
        return RDD(self._session, [p for r in rdds for p in r._collect_partitions()])

    def textFile(self, name: str, minPartitions: Optional[int] = None, use_unicode: bool = True):
        """One element per line; a directory or glob reads every file (one partition per file)."""
        from raydp_b200.sql.plan import list_data_files
        from raydp_b200.sql.rdd import RDD
        parts = []
        for path in list_data_files(name, ()):
            with open(path, "r", encoding="utf-8") as fh:
                parts.append(fh.read().splitlines())
        return RDD(self._session, parts or [[]])

    def wholeTextFiles(self, path: str, minPartitions: Optional[int] = None, use_unicode: bool = True):
        from raydp_b200.sql.plan import list_data_files
        from raydp_b200.sql.rdd import RDD
        parts = []
        for f in list_data_files(path, ()):
            with open(f, "r", encoding="utf-8") as fh:
                parts.append([("file:" + f, fh.read())])
        return RDD(self._session, parts or [[]])

    def pickleFile(self, name: str, minPartitions: Optional[int] = None):
        import pickle
        from raydp_b200.sql.plan import list_data_files
        from raydp_b200.sql.rdd import RDD
        parts = []
        for f in list_data_files(name, ()):
            with open(f, "rb") as fh:
                parts.append(pickle.load(fh))
        return RDD(self._session, parts or [[]])

    def broadcast(self, value) -> "Broadcast":
        return Broadcast(value)

    def accumulator(self, value, accum_param=None) -> "Accumulator":
        return Accumulator(value)

    def addPyFile(self, path: str) -> None:
        import sys as _sys
        target = os.path.dirname(os.path.abspath(path)) if path.endswith(".py") else os.path.abspath(path)
        if target not in _sys.path:
            _sys.path.insert(0, target)

    def addFile(self, path: str, recursive: bool = False) -> None:
        self._session._conf.set("spark.files", ",".join(filter(None, [self._session._conf.get("spark.files"), path])))

    def setJobGroup(self, groupId: str, description: str, interruptOnCancel: bool = False) -> None:
        self._session._conf.set("spark.jobGroup.id", groupId)

    def setJobDescription(self, value: str) -> None:
        self._session._conf.set("spark.job.description", value)

    def setLocalProperty(self, key: str, value: str) -> None:
        self._session._conf.set(key, value)

    def getLocalProperty(self, key: str) -> Optional[str]:
        return self._session._conf.get(key)

    def setCheckpointDir(self, dirName: str) -> None:
        self._session._conf.set("spark.checkpoint.dir", dirName)

    def cancelAllJobs(self) -> None:
        pass

    def cancelJobGroup(self, groupId: str) -> None:
        pass

    def statusTracker(self):
        return None

    # ---- the two py4j handles test code written against PySpark reaches for.  There is no JVM: they
    # answer the questions such code asks (force a GC; which recoverable RDDs are still pinned).
    @property
    def _jvm(self):
        return _JvmStandIn()

    @property
    def _jsc(self):
        return _JscStandIn(self._session)


class Broadcast:
    """``sc.broadcast(v)``: closures capture the wrapper and read ``.value`` on the executors (the value
    travels with the pickled closure; there is no separate torrent distribution to manage)."""

    def __init__(self, value):
        self._value = value

    @property
    def value(self):
        return self._value

    def unpersist(self, blocking: bool = False) -> None:
        pass

    def destroy(self, blocking: bool = False) -> None:
        self._value = None


class Accumulator:
    """Driver-side accumulator: ``+=`` / ``add`` from driver code (and from executors that run in this
    process); updates made in remote executor processes are not shipped back."""

    def __init__(self, value):
        self.value = value

    def add(self, term) -> None:
        self.value += term

    def __iadd__(self, term):
        self.add(term)
        return self

    def __repr__(self):
        return str(self.value)


class UDFRegistration:
    """``spark.udf.register(name, f, returnType)``: the function becomes callable from SQL text."""

    def __init__(self, session: "SparkSession"):
        self._session = session

    def register(self, name: str, f, returnType=None):
        from raydp_b200.sql import functions as F
        if isinstance(f, F.UserDefinedFunction):
            udf = f if returnType is None else F.UserDefinedFunction(f.func, returnType, f.vectorized)
        else:
            udf = F.UserDefinedFunction(f, returnType if returnType is not None else T.StringType())
        self._session._registered_udfs[name] = udf
        self._session._registered_udfs[name.lower()] = udf
        return udf

    def registerJavaFunction(self, name: str, javaClassName: str, returnType=None) -> None:
        raise NotImplementedError("there is no JVM in raydp_b200; register a Python function instead")


class _JvmStandIn:
    class System:
        @staticmethod
        def gc() -> None:
            import gc
            gc.collect()


class _PersistentRdds:
    def __init__(self, ids):
        self._ids = set(ids)

    def contains(self, rdd_id) -> bool:
        return int(rdd_id) in self._ids

    def __contains__(self, rdd_id) -> bool:
        return self.contains(rdd_id)

    def size(self) -> int:
        return len(self._ids)


class _JscStandIn:
    def __init__(self, session: "SparkSession"):
        self._session = session

    def sc(self) -> "_JscStandIn":
        return self

    def getPersistentRDDs(self) -> _PersistentRdds:
        """Ids of the recoverable RDDs whose lineage the master actor still pins."""
        backend = self._session._backend
        master = getattr(backend, "master", None)
        if master is None:
            return _PersistentRdds([])
        from raydp_b200 import runtime
        return _PersistentRdds(runtime.get(master.persistent_rdd_ids.remote()))


class DataFrameReader:
    def __init__(self, session: "SparkSession"):
        self._session = session
        self._format = "parquet"
        self._options: Dict[str, str] = {}
        self._schema: Optional[T.StructType] = None

    def format(self, source: str) -> "DataFrameReader":
        self._format = source.lower()
        return self

    def option(self, key: str, value) -> "DataFrameReader":
        self._options[key.lower()] = str(value)
        return self

    def options(self, **opts) -> "DataFrameReader":
        for k, v in opts.items():
            self.option(k, v)
        return self

    def schema(self, schema: Union[T.StructType, str]) -> "DataFrameReader":
        if isinstance(schema, str):
            schema = T.parse_ddl(schema)
        self._schema = schema
        return self

    def load(self, path: Optional[str] = None, format: Optional[str] = None, schema=None, **opts) -> DataFrame:  # noqa: A002
        if format:
            self.format(format)
        if schema is not None:
            self.schema(schema)
        self.options(**opts)
        path = path or self._options.get("path")
        if self._format in ("csv", "parquet", "orc", "json", "text"):
            paths = [path] if isinstance(path, str) else list(path)
            return getattr(self, self._format)(*paths) if self._format != "csv" else self.csv(paths)
        raise NotImplementedError(f"data source format {self._format!r} is not supported")

    def _files(self, fmt: str, paths, suffixes) -> DataFrame:
        import os
        from raydp_b200.sql.plan import FileSource, strip_scheme
        flat = [q for p in paths for q in (p if isinstance(p, (list, tuple)) else [p])]
        files = [f for p in flat for f in list_data_files(p, suffixes)]
        root = strip_scheme(flat[0]) if len(flat) == 1 and os.path.isdir(strip_scheme(flat[0])) else None
        base = self._options.get("basepath")
        return DataFrame(self._session, Plan(FileSource(files, fmt, strip_scheme(base) if base else root,
                                                        self._schema, self._options)))

    def json(self, *paths, schema=None, **opts) -> DataFrame:
        if schema is not None:
            self.schema(schema)
        self.options(**opts)
        return self._files("json", paths, ())

    def orc(self, *paths, **opts) -> DataFrame:
        self.options(**opts)
        return self._files("orc", paths, (".orc",))

    def text(self, *paths, wholetext: bool = False, **opts) -> DataFrame:
        self.option("wholetext", wholetext)
        return self._files("text", paths, ())

    def _flag(self, key: str, default: bool = False) -> bool:
        return str(self._options.get(key.lower(), default)).lower() == "true"

    def csv(self, path: Union[str, List[str]], header=None, inferSchema=None, sep=None,
            schema=None) -> DataFrame:
        if header is not None:
            self.option("header", header)
        if inferSchema is not None:
            self.option("inferSchema", inferSchema)
        if sep is not None:
            self.option("sep", sep)
        if schema is not None:
            self.schema(schema)
        paths = [path] if isinstance(path, str) else list(path)
        files = [f for p in paths for f in list_data_files(p, ())]
        delim = self._options.get("sep", self._options.get("delimiter", ","))
        src = CsvSource(files, self._flag("header"), self._flag("inferschema"), delim, self._schema)
        return DataFrame(self._session, Plan(src))

    def parquet(self, *paths: str, **opts) -> DataFrame:
        import os
        from raydp_b200.sql.plan import partition_values, strip_scheme
        self.options(**opts)
        flat = [q for p in paths for q in (p if isinstance(p, (list, tuple)) else [p])]
        files = [f for p in flat for f in list_data_files(p, (".parquet",))]
        root = strip_scheme(flat[0]) if len(flat) == 1 and os.path.isdir(strip_scheme(flat[0])) else None
        if self._schema is not None or any(partition_values(f, root) for f in files):
            return self._files("parquet", paths, (".parquet",))      # partition columns come from the paths
        return DataFrame(self._session, Plan(ParquetSource(files)))

    def table(self, name: str):
        return self._session.catalog.lookup(name)


class _Builder:
    def __init__(self):
        self._opts: Dict[str, str] = {}

    def master(self, url: str) -> "_Builder":
        self._opts["spark.master"] = url
        return self

    def appName(self, name: str) -> "_Builder":
        self._opts["spark.app.name"] = name
        return self

    def config(self, key=None, value=None, conf: Optional[SparkConf] = None, map=None) -> "_Builder":  # noqa: A002
        if conf is not None:
            self._opts.update(dict(conf.getAll()))
        if map is not None:
            self._opts.update({k: str(v) for k, v in map.items()})
        if key is not None:
            self._opts[key] = str(value)
        return self

    def enableHiveSupport(self) -> "_Builder":
        self._opts["spark.sql.catalogImplementation"] = "hive"
        return self

    def getOrCreate(self) -> "SparkSession":
        global _active_session
        with _active_lock:
            wanted = self._opts.get("spark.master")
            if _active_session is not None and not _active_session._stopped and (
                    wanted is None or wanted == _active_session._conf.get("spark.master", wanted)):
                for k, v in self._opts.items():
                    _active_session._conf.set(k, v)
                return _active_session
            # an explicit, different master (e.g. ``ray://<app master>`` from init_spark while a local[n]
            # session is still around) gets its own session instead of silently reusing the old backend
            opts = dict(_builder_options)
            submitted = os.environ.get("RAYDP_SUBMIT_CONF")
            if submitted:                       # running under raydp-submit
                import json
                opts.update(json.loads(submitted))
            opts.update(self._opts)
            if opts.get("spark.master", "") == "ray":
                # master URL "ray" = create the app master here (reference
                # RayCoarseGrainedSchedulerBackend.scala:84-96): delegate to init_spark
                from raydp_b200.context import init_spark
                extra = {k: v for k, v in opts.items() if k not in (
                    "spark.master", "spark.app.name", "spark.executor.instances", "spark.executor.cores",
                    "spark.executor.memory")}
                return init_spark(opts.get("spark.app.name", "raydp-submit"),
                                  int(opts.get("spark.executor.instances", "1")),
                                  int(opts.get("spark.executor.cores", "1")),
                                  opts.get("spark.executor.memory", "1g"), configs=extra)
            _builder_options.update(self._opts)
            session = SparkSession(opts)
            _active_session = session
            return session


class _BuilderDescriptor:
    def __get__(self, obj, owner) -> _Builder:
        return _Builder()


class SparkSession:
    builder = _BuilderDescriptor()

    def __init__(self, options: Dict[str, str]):
        self._conf = SparkConf(options)
        self._app_name = options.get("spark.app.name", "raydp-b200-app")
        self._master = options.get("spark.master", "local[*]")
        self._app_id = f"app-{uuid.uuid4().hex[:16]}"
        self._stopped = False
        self._tracked_refs: List[Any] = []
        self._conf.set("spark.app.id", self._app_id)
        self._conf.set("spark.app.name", self._app_name)
        self._conf.set("spark.master", self._master)
        self._conf.set("spark.sql.session.timeZone", options.get("spark.sql.session.timeZone", "UTC"))
        if self._master.startswith("local"):
            spec = self._master[len("local"):].strip("[]")
            threads = (os.cpu_count() or 1) if spec in ("", "*") else int(spec.split(",")[0])
            self._backend: Backend = LocalBackend(threads)
        elif self._master.startswith("ray"):
            self._backend = ClusterBackend(self._master, self._app_name, dict(options))
            self._app_id = self._backend.app_id
            self._conf.set("spark.app.id", self._app_id)
        else:
            raise ValueError(f"Could not parse Master URL: '{self._master}'")
        self.sparkContext = SparkContext(self)
        self.conf = RuntimeConfig(self._conf)
        self.version = "3.5.0-raydp-b200"
        self._registered_udfs: Dict[str, Any] = {}
        self.udf = UDFRegistration(self)
        import weakref
        _live_sessions[:] = [r for r in _live_sessions if r() is not None and not r()._stopped]
        _live_sessions.append(weakref.ref(self))

    # ------------------------------------------------------------------ factories
    @classmethod
    def getActiveSession(cls) -> Optional["SparkSession"]:
        if _active_session is not None and not _active_session._stopped:
            return _active_session
        for ref in reversed(_live_sessions):                 # another session that is still running
            other = ref()
            if other is not None and not other._stopped:
                return other
        return None

    @classmethod
    def active(cls) -> "SparkSession":
        session = cls.getActiveSession()
        if session is None:
            raise RuntimeError("No active or default Spark session found")
        return session

    @property
    def streams(self):
        raise NotImplementedError("structured streaming is not part of the built-in engine")

    @property
    def readStream(self):
        raise NotImplementedError("structured streaming is not part of the built-in engine")

    @property
    def _jvm(self):
        raise NotImplementedError("there is no JVM in raydp_b200; use the Python API")

    @property
    def read(self) -> DataFrameReader:
        return DataFrameReader(self)

    def range(self, start: int, end: Optional[int] = None, step: int = 1,
              numPartitions: Optional[int] = None) -> DataFrame:
        if end is None:
            start, end = 0, start
        n = numPartitions or self._backend.default_parallelism
        return DataFrame(self, Plan(RangeSource(start, end, step, n)))

    def _from_tables(self, tables: List[pa.Table]) -> DataFrame:
        return DataFrame(self, Plan(LocalSource(list(tables))))

    def _track_refs(self, refs: List[Any]) -> None:
        self._tracked_refs.extend(refs)

    def createDataFrame(self, data, schema=None, samplingRatio=None, verifySchema=True) -> DataFrame:
        from raydp_b200.sql.rdd import RDD
        names: Optional[List[str]] = None
        arrow_schema: Optional[pa.Schema] = None
        if isinstance(schema, T.StructType):
            arrow_schema = schema.to_arrow_schema()
            names = schema.names
        elif isinstance(schema, (list, tuple)):
            names = list(schema)
        elif isinstance(schema, str):
            st = T.parse_ddl(schema)
            arrow_schema, names = st.to_arrow_schema(), st.names
        if isinstance(data, RDD):
            parts = data._collect_partitions()
            tables = [_rows_to_table(p, names, arrow_schema) for p in parts if p]
            if not tables:
                tables = [_rows_to_table([], names, arrow_schema)]
            unified = pa.unify_schemas([t.schema for t in tables]) if arrow_schema is None else arrow_schema
            return self._from_tables([t.cast(unified) for t in tables])
        try:
            import pandas as pd
            if isinstance(data, pd.DataFrame):
                table = pa.Table.from_pandas(data, preserve_index=False)
                if names:
                    table = table.rename_columns(names)
                if arrow_schema is not None:
                    table = table.cast(arrow_schema, safe=False)
                return self._from_tables(_split_table(table, self._backend.default_parallelism))
        except ImportError:  # pragma: no cover
            pass
        if isinstance(data, pa.Table):
            return self._from_tables(_split_table(data, self._backend.default_parallelism))
        table = _rows_to_table(list(data), names, arrow_schema)
        return self._from_tables(_split_table(table, self._backend.default_parallelism))

    @property
    def catalog(self):
        """Temp views + parquet-backed warehouse tables (``raydp_b200.sql.sqlparser.Catalog``)."""
        if getattr(self, "_catalog", None) is None:
            from raydp_b200.sql.sqlparser import Catalog
            self._catalog = Catalog(self)
        return self._catalog

    def sql(self, query: str) -> DataFrame:
        """SELECT / SHOW / CREATE TABLE AS / INSERT / DROP over views and warehouse tables, parsed in
        Python (``raydp_b200.sql.sqlparser``); joins and sub-queries need the DataFrame API."""
        from raydp_b200.sql.sqlparser import run_sql
        return run_sql(self, query)

    def table(self, name: str) -> DataFrame:
        return self.catalog.lookup(name)

    def newSession(self) -> "SparkSession":
        return self

    def stop(self) -> None:
        global _active_session
        if self._stopped:
            return
        self._stopped = True
        try:
            if self._tracked_refs:
                from raydp_b200 import runtime
                if runtime.is_initialized():
                    runtime.free(self._tracked_refs)
            self._backend.stop()
        finally:
            with _active_lock:
                if _active_session is self:
                    _active_session = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.stop()


def reset_builder_singletons() -> None:
    """Forget builder options and the active session (reference ``context.py:33-51`` clears
    pyspark's builder ``_options`` so the next ``init_spark`` starts clean)."""
    global _active_session
    with _active_lock:
        _builder_options.clear()
        _active_session = None


def _rows_to_table(rows: List[Any], names: Optional[List[str]], schema: Optional[pa.Schema]) -> pa.Table:
    if rows and isinstance(rows[0], dict):
        names = names or list(rows[0].keys())
        rows = [tuple(r.get(n) for n in names) for r in rows]
    elif rows and isinstance(rows[0], T.Row) and rows[0].__fields__ and not names:
        names = list(rows[0].__fields__)
    elif rows and not isinstance(rows[0], (tuple, list)):
        rows = [(r,) for r in rows]
    width = len(rows[0]) if rows else (len(names) if names else 0)
    names = names or [f"_{i + 1}" for i in range(width)]
    cols = list(zip(*rows)) if rows else [[] for _ in names]
    arrays = []
    for i, name in enumerate(names):
        typ = schema.field(name).type if schema is not None else None
        values = list(cols[i]) if rows else []
        if typ is None and any(isinstance(v, (dict, T.Row)) or (isinstance(v, (list, tuple)) and v and
                                                                 isinstance(v[0], (dict, T.Row))) for v in values):
            typ = _infer_nested(values)          # PySpark's inference: dict -> map, Row -> struct
        if typ is not None and (pa.types.is_map(typ) or pa.types.is_struct(typ) or pa.types.is_list(typ)):
            values = [_arrow_ready(v, typ) for v in values]
        arrays.append(pa.array(values, type=typ))
    return pa.table(arrays, names=names)


def _infer_nested(values: List[Any]) -> pa.DataType:
    live = [v for v in values if v is not None]
    if not live:
        return pa.null()
    head = live[0]
    if isinstance(head, T.Row) and head.__fields__:
        return pa.struct([pa.field(f, _infer_nested([v[j] for v in live if isinstance(v, T.Row) and len(v) > j]))
                          for j, f in enumerate(head.__fields__)])
    if isinstance(head, dict):
        keys = [k for v in live if isinstance(v, dict) for k in v]
        items = [x for v in live if isinstance(v, dict) for x in v.values()]
        return pa.map_(_infer_nested(keys) if keys else pa.string(), _infer_nested(items) if items else pa.null())
    if isinstance(head, (list, tuple)):
        return pa.list_(_infer_nested([x for v in live if isinstance(v, (list, tuple)) for x in v]))
    return pa.array(live).type


def _arrow_ready(value, typ: pa.DataType):
    """Python value -> the shape ``pa.array`` wants for ``typ`` (Row -> dict, dict -> key/value pairs)."""
    if value is None:
        return None
    if pa.types.is_struct(typ):
        if isinstance(value, T.Row) and value.__fields__:
            value = value.asDict()
        elif isinstance(value, (tuple, list)):
            value = {f.name: v for f, v in zip(typ, value)}
        return {f.name: _arrow_ready(value.get(f.name), f.type) for f in typ}
    if pa.types.is_map(typ):
        pairs = value.items() if isinstance(value, dict) else value
        return [(_arrow_ready(k, typ.key_type), _arrow_ready(v, typ.item_type)) for k, v in pairs]
    if pa.types.is_list(typ) or pa.types.is_large_list(typ):
        return [_arrow_ready(v, typ.value_type) for v in value]
    return value
