"""Executor side of the DataFrame engine.

Parity: reference ``RayDPExecutor.scala:45-356`` -- an actor named ``raydp-executor-<id>`` that
registers with the app master (with restart detection ``:66-101``), runs partition tasks, caches
Arrow partitions in its block manager and serves them to fetch tasks through
``getRDDPartition`` (``:312-355``), asking for a recache when the block is missing
(``requestRecacheRDD`` ``:289-310``).  Here a "block" is an Arrow table held by the executor
process; a partition task is a pickled :class:`~raydp_b200.sql.plan.Plan` plus an action.
"""
from __future__ import annotations

import os
import time
import threading
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.exceptions import RayDPException
from raydp_b200.sql.plan import Plan

EXECUTOR_NAME_PREFIX = "raydp-executor-"


def _partial_agg(table: pa.Table, keys: List[str], aggs: List[Tuple[str, Optional[str]]]) -> pa.Table:
    """First stage of a two-stage aggregation; emits mergeable partial states.  Aggregates without a
    compact state (``sql/aggregates.py``) get their group's values gathered into a ``<col>_list`` column:
    the group-by collects ROW INDICES (a list of int64 works for every value type, structs included) and
    the values are taken by index afterwards."""
    specs = []
    gathered: List[str] = []
    for fn, col in aggs:
        if fn == "count":
            specs.append(([], "count_all") if col is None else (col, "count"))
        elif fn in ("sum", "min", "max"):
            specs.append((col, fn))
        elif fn == "mean":
            specs.extend([(col, "sum"), (col, "count")])
        elif fn in ("stddev", "variance"):
            sq = f"__sq_{col}"
            if sq not in table.column_names:
                c = pc.cast(table.column(col), pa.float64())
                table = table.append_column(sq, pc.multiply(c, c))
            specs.extend([(col, "sum"), (sq, "sum"), (col, "count")])
        elif fn == "count_distinct":
            specs.append((col, "distinct"))
        elif fn == "first":
            specs.append((col, "first"))
        elif col not in gathered:
            gathered.append(col)
    if gathered:
        table = table.append_column("__rowidx", pa.array(np.arange(table.num_rows, dtype=np.int64)))
        specs.append(("__rowidx", "list"))
    seen, uniq = set(), []
    for s in specs:
        k = (tuple(s[0]) if isinstance(s[0], list) else s[0], s[1])
        if k not in seen:
            seen.add(k)
            uniq.append(s)
    if keys:
        out = table.group_by(keys, use_threads=False).aggregate(uniq)
    elif table.num_rows == 0 and not gathered:
        out = pa.TableGroupBy(table, [], use_threads=False).aggregate(uniq)
    else:
        # a global aggregate is a group-by over one constant key (the scalar path lacks list / distinct)
        keyed = table.append_column("__all", pa.array(np.zeros(table.num_rows, dtype=np.int8)))
        out = keyed.group_by(["__all"], use_threads=False).aggregate(uniq).drop_columns(["__all"])
    if gathered:
        out = gather_by_index(out, "__rowidx_list", table, gathered)
    return out


def gather_by_index(grouped: pa.Table, index_col: str, source: pa.Table, cols: List[str]) -> pa.Table:
    """Replace ``index_col`` (list<int64> of row numbers of ``source``) by one ``<col>_list`` column per
    entry of ``cols`` holding the rows' values."""
    idx = grouped.column(index_col).combine_chunks()
    position = grouped.column_names.index(index_col)
    grouped = grouped.remove_column(position)
    flat = pc.list_flatten(idx)
    offsets = idx.offsets
    if len(idx) and offsets[0].as_py() != 0:
        offsets = pc.subtract(offsets, offsets[0])
    for col in cols:
        values = source.column(col).combine_chunks().take(flat)
        grouped = grouped.append_column(f"{col}_list", pa.ListArray.from_arrays(offsets, values))
    return grouped


def _write_one(table: pa.Table, directory: str, idx: int, args: Dict[str, Any]) -> None:
    fmt = args["format"]
    os.makedirs(directory, exist_ok=True)
    stem = os.path.join(directory, f"part-{idx:05d}-{args['job']}")
    if fmt == "parquet":
        import pyarrow.parquet as pq
        compression = args.get("compression") or "snappy"
        pq.write_table(table, stem + ".parquet",
                       compression=None if compression in ("none", "uncompressed") else compression)
    elif fmt == "orc":
        import pyarrow.orc as orc
        orc.write_table(table, stem + ".orc")
    elif fmt == "csv":
        import pyarrow.csv as pacsv
        pacsv.write_csv(table, stem + ".csv", pacsv.WriteOptions(include_header=bool(args.get("header", False)),
                                                                 delimiter=args.get("sep", ",")))
    elif fmt == "json":
        import json as _json
        from raydp_b200.sql.fnlib import _jsonable
        names = table.column_names
        cols = [c.to_pylist() for c in table.columns]
        with open(stem + ".json", "w", encoding="utf-8") as fh:
            for row in zip(*cols):
                fh.write(_json.dumps({k: _jsonable(v) for k, v in zip(names, row) if v is not None},
                                     separators=(",", ":")) + "\n")
    elif fmt == "text":
        if table.num_columns != 1 or not pa.types.is_string(table.schema.field(0).type):
            raise ValueError("Text data source supports only a single string column")
        with open(stem + ".txt", "w", encoding="utf-8") as fh:
            for line in table.column(0).to_pylist():
                fh.write(("" if line is None else line) + "\n")
    else:
        raise NotImplementedError(f"data source format {fmt!r} is not supported")


def write_partition(table: pa.Table, idx: int, args: Dict[str, Any]) -> int:
    """Write one partition under ``args["path"]``; with ``partition_by`` the rows go to
    ``key=value/...`` sub-directories and the partition columns are left out of the files."""
    keys: List[str] = list(args.get("partition_by") or [])
    if not keys:
        if table.num_rows or idx == 0:
            _write_one(table, args["path"], idx, args)
        return table.num_rows
    if table.num_rows == 0:
        return 0
    groups = table.group_by(keys, use_threads=False).aggregate([])
    key_cols = [table.column(k).combine_chunks() for k in keys]
    for g in range(groups.num_rows):
        mask = None
        parts = []
        for k, col in zip(keys, key_cols):
            v = groups.column(k)[g]
            hit = pc.is_null(col) if not v.is_valid else pc.fill_null(pc.equal(col, v), False)
            mask = hit if mask is None else pc.and_(mask, hit)
            parts.append(f"{k}={'__HIVE_DEFAULT_PARTITION__' if not v.is_valid else v.as_py()}")
        piece = table.filter(mask).drop_columns(keys)
        _write_one(piece, os.path.join(args["path"], *parts), idx, args)
    return table.num_rows


def _stable_hashes(col: pa.Array) -> np.ndarray:
    """Per-value 64-bit hashes that agree ACROSS PROCESSES (Python's ``hash(str)`` is salted per process, which
    would send equal keys to different shuffle partitions on different executors)."""
    import hashlib

    def digest(value) -> int:
        return int.from_bytes(hashlib.blake2b(repr(value).encode(), digest_size=8).digest(), "little")
    try:
        uniq = pc.unique(col)
        codes = pc.index_in(col, value_set=uniq).to_numpy(zero_copy_only=False)
        table = np.array([digest(v) for v in uniq.to_pylist()] + [0], dtype=np.uint64)
        return table[np.where(np.isnan(codes.astype(np.float64)), len(table) - 1, codes).astype(np.int64)]
    except (pa.ArrowNotImplementedError, pa.ArrowInvalid):        # nested types: hash row by row
        return np.array([digest(v) for v in col.to_pylist()], dtype=np.uint64)


def execute_action(plan: Plan, idx: int, action: str, args: Dict[str, Any], blocks: Dict) -> Any:
    """Run one partition task.  Shared by executor actors and the in-process local backend."""
    if action == "load_block":
        table = blocks.get((args["rdd_id"], idx))
        if table is None:
            table = plan.compute(idx)
            blocks[(args["rdd_id"], idx)] = table
        return table
    table = plan.compute(idx)
    if action == "count":
        return table.num_rows
    if action == "collect":
        return table
    if action == "take":
        return table.slice(0, args["n"])
    if action == "cache":
        blocks[(args["rdd_id"], idx)] = table.combine_chunks()
        return table.num_rows, int(table.nbytes)
    if action == "put":
        from raydp_b200 import runtime
        return runtime.put(table.combine_chunks()), table.num_rows
    if action == "write_parquet":
        import pyarrow.parquet as pq
        os.makedirs(args["path"], exist_ok=True)
        out = os.path.join(args["path"], f"part-{idx:05d}-{args['job']}.parquet")
        compression = args.get("compression") or "snappy"
        pq.write_table(table, out, compression=None if compression in ("none", "uncompressed")
                       else compression)
        return table.num_rows
    if action == "write_csv":
        import pyarrow.csv as pacsv
        os.makedirs(args["path"], exist_ok=True)
        out = os.path.join(args["path"], f"part-{idx:05d}-{args['job']}.csv")
        pacsv.write_csv(table, out, pacsv.WriteOptions(include_header=args.get("header", False)))
        return table.num_rows
    if action == "write_files":
        return write_partition(table, idx, args)
    if action == "shuffle_map":
        from raydp_b200 import runtime
        n_out = args["n_out"]
        keys = args.get("keys")
        if keys:
            h = np.zeros(table.num_rows, dtype=np.uint64)
            for k in keys:
                col = table.column(k).combine_chunks()
                if pa.types.is_integer(col.type) or pa.types.is_floating(col.type):
                    v = np.nan_to_num(col.to_numpy(zero_copy_only=False)).astype(np.int64).view(np.uint64)
                else:
                    v = _stable_hashes(col)
                h = (h ^ v) * np.uint64(0x9E3779B97F4A7C15)
            dest = (h % np.uint64(n_out)).astype(np.int64)
        else:
            # round-robin with a per-partition offset keeps output partitions balanced
            dest = (np.arange(table.num_rows, dtype=np.int64) + idx) % n_out
        out = []
        for d in range(n_out):
            piece = table.filter(pa.array(dest == d)) if table.num_rows else table
            out.append(runtime.put(piece.combine_chunks(), _owner=args.get("_shuffle_owner"))
                       if piece.num_rows else None)
        return out
    if action == "agg_partial":
        return _partial_agg(table, args["keys"], args["aggs"])
    if action == "py_partition":  # RDD.mapPartitions-style python callable on rows
        return args["fn"](table)
    raise NotImplementedError(f"action {action!r}")


class SparkExecutor:
    """The actor class; instances are created by the app master with
    ``max_restarts=-1, max_task_retries=-1`` and a concurrency of ``cores + 1`` so a data fetch
    can run while partition tasks occupy the cores (reference ``RayExecutorUtils.java:58-60``)."""

    def __init__(self, executor_id: int, app_id: str, master_name: str, cores: int,
                 conf: Optional[Dict[str, str]] = None):
        from raydp_b200 import runtime
        self.actor_executor_id = executor_id       # stable: part of the actor name
        self.executor_id = executor_id             # Spark-level id: changes after a restart
        self.app_id = app_id
        self.master_name = master_name
        self.cores = cores
        self.conf = conf or {}
        self.blocks: Dict[Tuple[int, int], pa.Table] = {}
        self._lock = threading.Lock()
        self.started = False
        ctx = runtime.get_runtime_context()
        self.restarted = ctx.was_current_actor_reconstructed
        self.node_ip = runtime.get_node_ip_address()
        self._register()

    def _register(self) -> None:
        """RegisterExecutor; a restarted actor first asks for a fresh executor id while keeping
        its actor name (reference ``RayDPExecutor.scala:82-94``)."""
        from raydp_b200 import runtime
        last_exc: Optional[BaseException] = None
        for _ in range(3):
            try:
                master = runtime.get_actor(self.master_name)
                if self.restarted:
                    self.executor_id = runtime.get(master.request_add_pending_restarted_executor.remote(
                        self.actor_executor_id), timeout=30)
                ok = runtime.get(master.register_executor.remote(
                    self.executor_id, self.node_ip, os.getpid(), self.actor_executor_id), timeout=30)
                if ok:
                    self.started = True
                    return
            except Exception as exc:  # noqa: BLE001
                last_exc = exc
        raise RayDPException(f"executor {self.executor_id} could not register with app master: {last_exc}")

    # ------------------------------------------------------------------ tasks
    def run_partition(self, plan: Plan, idx: int, action: str, args: Optional[dict] = None):
        args = dict(args or {})
        service = self._shuffle_service() if action == "shuffle_map" else None
        if service is not None:
            args["_shuffle_owner"] = service        # blocks outlive this executor
        out = execute_action(plan, idx, action, args, self.blocks)
        if service is not None:
            from raydp_b200 import runtime
            refs = [r for r in out if r is not None]
            if refs:
                runtime.get(service.register_blocks.remote(str(args.get("shuffle_id", "0")), refs), timeout=60)
        return out

    def _shuffle_service(self):
        """Handle of this node's external shuffle service when ``spark.shuffle.service.enabled``."""
        if str(self.conf.get("spark.shuffle.service.enabled", "false")).lower() != "true":
            return None
        if getattr(self, "_shuffle_handle", None) is None:
            from raydp_b200 import runtime
            from raydp_b200.spark.shuffle_service import service_name
            deadline = time.time() + 30
            while True:
                try:
                    self._shuffle_handle = runtime.get_actor(service_name(self.master_name, self.node_ip))
                    break
                except Exception:  # noqa: BLE001  (the master starts it right after registration)
                    if time.time() > deadline:
                        raise
                    time.sleep(0.05)
        return self._shuffle_handle

    def ping(self) -> Dict[str, Any]:
        return {"executor_id": self.executor_id, "actor_executor_id": self.actor_executor_id,
                "pid": os.getpid(), "node_ip": self.node_ip, "restarted": self.restarted,
                "cores": self.cores, "num_blocks": len(self.blocks)}

    # ------------------------------------------------------------------ data plane
    def get_block_locations(self, rdd_id: int, num_partitions: int) -> List[bool]:
        return [(rdd_id, p) in self.blocks for p in range(num_partitions)]

    def get_rdd_partition(self, rdd_id: int, partition_id: int) -> pa.Table:
        """Serve a cached Arrow partition; on a miss ask the driver agent (the master actor keeps
        the recoverable RDD's lineage) to recache it here, then retry once."""
        table = self.blocks.get((rdd_id, partition_id))
        if table is None:
            from raydp_b200 import runtime
            master = runtime.get_actor(self.master_name)
            plan = runtime.get(master.get_recoverable_plan.remote(rdd_id), timeout=60)
            if plan is None:
                raise RayDPException(
                    f"Still cannot get the block rdd_{rdd_id}_{partition_id} after recache: the "
                    "recoverable RDD was released")
            table = plan.compute(partition_id).combine_chunks()
            self.blocks[(rdd_id, partition_id)] = table
        return table.replace_schema_metadata()

    def drop_rdd(self, rdd_id: int) -> int:
        victims = [k for k in self.blocks if k[0] == rdd_id]
        for k in victims:
            self.blocks.pop(k, None)
        return len(victims)

    def block_ids(self) -> List[Tuple[int, int]]:
        return sorted(self.blocks)

    def stop(self) -> None:
        from raydp_b200 import runtime
        runtime.exit_actor()
