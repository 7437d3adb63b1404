"""Session lifecycle, placement, exchange and recovery.

Mirrors reference ``tests/test_spark_cluster.py`` (count ``:35-39``, fractional CPU ``:42-87``,
custom-resource affinity ``:90-110``, init inside an actor ``:113-134``, driver host ``:137-147``,
``to_spark`` round trips ``:150-216``, placement groups ``:219-259``, reconstruction ``:262-299``,
recoverable pin/release ``:302-366``), ``test_from_spark.py`` and ``test_spark_master_memory.py``.
"""
import time

import pyarrow as pa
import pytest

import raydp_b200 as raydp
from raydp_b200 import data as rdata
from raydp_b200 import runtime as rt
from raydp_b200.runtime.cluster_utils import Cluster
from raydp_b200.spark import (from_spark_recoverable, get_raydp_master_owner,
                              release_spark_recoverable, spark_dataframe_to_ray_dataset)
from raydp_b200.sql.functions import col, rand


def test_spark(spark_on_ray_small):
    spark = spark_on_ray_small
    assert spark.range(0, 10).count() == 10
    names = {a["name"] for a in rt.list_actors() if a["state"] == "ALIVE"}
    assert {"test_SPARK_MASTER", "raydp-executor-0"} <= names


def test_double_init_and_idempotent_stop(ray_cluster):
    raydp.init_spark("twice", 1, 1, "500M")
    with pytest.raises(Exception, match="The spark environment has inited."):
        raydp.init_spark("twice", 1, 1, "500M")
    raydp.stop_spark()
    raydp.stop_spark()
    spark = raydp.init_spark("again", 1, 1, "500 MB")
    assert spark.range(5).count() == 5
    assert spark.sparkContext.appName == "again"


def test_spark_driver_and_executor_hostname(spark_on_ray_small):
    conf = spark_on_ray_small.conf
    node_ip = rt.get_node_ip_address()
    assert conf.get("spark.driver.host") == node_ip
    assert conf.get("spark.driver.bindAddress") == node_ip
    assert conf.get("spark.executor.instances") == "1"
    assert conf.get("spark.executor.memory") == str(500 * 2 ** 20)


def test_legacy_spark_on_fractional_cpu():
    rt.shutdown()
    cluster = Cluster(initialize_head=True, head_node_args={"num_cpus": 2})
    try:
        spark = raydp.init_spark("frac", 1, 3, "500M",
                                 configs={"spark.ray.actor.resource.cpu": "0.1"})
        assert spark.range(0, 10).count() == 10
        ex = [a for a in rt.list_actors() if a["name"] == "raydp-executor-0"][0]
        assert ex["required_resources"]["CPU"] == pytest.approx(0.1)
        raydp.stop_spark()
        spark = raydp.init_spark("frac2", 1, 3, "500M", configs={
            "spark.ray.raydp_spark_executor.actor.resource.cpu": "0.25"})
        assert spark.range(0, 10).count() == 10
    finally:
        raydp.stop_spark()
        cluster.shutdown()


def test_spark_executor_node_affinity_and_master_resources():
    rt.shutdown()
    cluster = Cluster(initialize_head=True, head_node_args={"num_cpus": 1})
    try:
        cluster.add_node(num_cpus=3, resources={"spark_executor": 10, "master": 10},
                         memory=8 * 2 ** 30)
        spark = raydp.init_spark("aff", 2, 1, "500M", configs={
            "spark.ray.raydp_spark_executor.actor.resource.spark_executor": "1",
            "spark.ray.raydp_spark_master.actor.resource.CPU": "0.5",
            "spark.ray.raydp_spark_master.actor.resource.memory": str(100 * 2 ** 20),
            "spark.ray.raydp_spark_master.actor.resource.master": "1"})
        assert spark.range(0, 10).count() == 10
        actors = {a["name"]: a for a in rt.list_actors() if a["state"] == "ALIVE"}
        for name in ("raydp-executor-0", "raydp-executor-1", "aff_SPARK_MASTER"):
            assert actors[name]["node_ip"] == "127.0.0.2"
        need = actors["aff_SPARK_MASTER"]["required_resources"]
        assert need["CPU"] == 0.5 and need["memory"] == 100 * 2 ** 20 and need["master"] == 1
    finally:
        raydp.stop_spark()
        cluster.shutdown()


def test_spark_remote(ray_cluster):
    @rt.remote
    class SparkRemote:
        def __init__(self):
            import raydp_b200
            self.spark = raydp_b200.init_spark("remote", 1, 1, "500M")

        def run(self):
            return self.spark.range(0, 100).count()

        def stop(self):
            import raydp_b200
            raydp_b200.stop_spark()

    driver = SparkRemote.remote()
    assert rt.get(driver.run.remote(), timeout=120) == 100
    rt.get(driver.stop.remote())


@pytest.mark.parametrize("strategy", ["PACK", "SPREAD", "STRICT_PACK"])
def test_placement_group_strategy(ray_cluster, strategy):
    spark = raydp.init_spark("pg", 2, 1, "500M", placement_group_strategy=strategy)
    assert spark.range(0, 10).count() == 10
    table = rt.placement_group_table()
    assert sum(1 for row in table.values() if row["state"] == "CREATED") == 1
    raydp.stop_spark()
    assert all(row["state"] == "REMOVED" for row in rt.placement_group_table().values())


def test_existing_placement_group_and_bundle_indexes(ray_cluster):
    pg = rt.placement_group([{"CPU": 1, "memory": 600 * 2 ** 20}] * 3, strategy="PACK")
    assert pg.wait(10)
    spark = raydp.init_spark("pg2", 2, 1, "500M", placement_group=pg,
                             placement_group_bundle_indexes=[0, 2])
    assert spark.range(0, 10).count() == 10
    assert spark.conf.get("spark.ray.placement_group") == pg.id.hex()
    assert spark.conf.get("spark.ray.bundle_indexes") == "0,2"
    raydp.stop_spark()
    assert rt.placement_group_table(pg)["state"] == "CREATED"    # caller-owned group survives
    rt.remove_placement_group(pg)


def test_from_spark_to_spark_roundtrip(spark_on_ray_2_executors):
    spark = spark_on_ray_2_executors
    df = spark.createDataFrame([(i, float(i) * 2, str(i)) for i in range(300)], ["a", "b", "c"])
    ds = rdata.from_spark(df)
    assert ds.count() == 300
    assert ds.schema().base_schema.names == ["a", "b", "c"]
    back = ds.to_spark(spark)
    assert back.count() == 300
    assert sorted(r.a for r in back.collect()) == list(range(300))
    assert dict(back.dtypes) == dict(df.dtypes)
    # bytes blocks (Arrow IPC streams) take the other conversion branch
    table = pa.table({"x": list(range(50))})
    sink = pa.BufferOutputStream()
    with pa.ipc.new_stream(sink, table.schema) as w:
        w.write_table(table)
    from raydp_b200.spark import ray_dataset_to_spark_dataframe
    ref = rt.put(sink.getvalue().to_pybytes())
    assert ray_dataset_to_spark_dataframe(spark, table.schema, [ref]).count() == 50
    with pytest.raises(RuntimeError, match="required pyarrow.lib.Schema"):
        ray_dataset_to_spark_dataframe(spark, "not a schema", [ref])


def test_parallelism_and_api_compatibility(spark_on_ray_2_executors):
    spark = spark_on_ray_2_executors
    before = rt.available_resources()["CPU"]
    df = spark.range(0, 1000, numPartitions=3)
    ds = spark_dataframe_to_ray_dataset(df, parallelism=4, owner=get_raydp_master_owner(spark))
    assert ds.num_blocks() == 4 and ds.count() == 1000
    deadline = time.time() + 10
    while rt.available_resources()["CPU"] != before and time.time() < deadline:
        time.sleep(0.05)
    assert rt.available_resources()["CPU"] == before       # fetch tasks hold no CPUs afterwards


def test_release_spark_recoverable(spark_on_ray_2_executors):
    spark = spark_on_ray_2_executors
    master = rt.get_actor("test_SPARK_MASTER")
    df = spark.range(0, 1000, numPartitions=2)
    ds = from_spark_recoverable(df)
    assert ds.count() == 1000
    assert len(rt.get(master.persistent_rdd_ids.remote())) == 1
    ex = rt.get_actor("raydp-executor-0")
    assert len(rt.get(ex.block_ids.remote())) >= 1
    release_spark_recoverable(df)
    assert rt.get(master.persistent_rdd_ids.remote()) == []
    assert rt.get(ex.block_ids.remote()) == []
    release_spark_recoverable(df)      # second call is a no-op


def test_data_survives_stop_spark_without_cleanup(ray_cluster):
    spark = raydp.init_spark("keep", 1, 1, "500M")
    ds = rdata.from_spark(spark.range(0, 1000))
    raydp.stop_spark(cleanup_data=False)
    assert ds.count() == 1000 and sum(t.num_rows for t in ds.iter_tables()) == 1000
    assert any(a["name"] == "keep_SPARK_MASTER" and a["state"] == "ALIVE" for a in rt.list_actors())
    raydp.stop_spark()
    time.sleep(0.3)
    assert not any(a["name"] == "keep_SPARK_MASTER" and a["state"] == "ALIVE" for a in rt.list_actors())


def test_reconstruction():
    """Fault injection: lose the node holding an executor and its blocks, then read everything."""
    rt.shutdown()
    cluster = Cluster(initialize_head=True, head_node_args={"num_cpus": 1})
    try:
        cluster.add_node(num_cpus=2, resources={"spark_executor": 1}, memory=4 * 2 ** 30)
        victim = cluster.add_node(num_cpus=2, resources={"spark_executor": 1}, memory=4 * 2 ** 30)
        spark = raydp.init_spark("recon", 2, 1, "500M", configs={
            "spark.ray.raydp_spark_executor.actor.resource.spark_executor": "1"})
        df = spark.range(0, 100000, numPartitions=4).withColumn("x", rand(1))
        ds = spark_dataframe_to_ray_dataset(df)
        assert ds.count() == 100000
        on_victim = [a["name"] for a in rt.list_actors()
                     if a["node_id"] == victim.node_id and a["name"].startswith("raydp-executor")]
        assert len(on_victim) == 1
        cluster.remove_node(victim)
        cluster.add_node(num_cpus=2, resources={"spark_executor": 1}, memory=4 * 2 ** 30)
        total = 0
        for bundle in ds.iter_internal_ref_bundles():
            for ref, _meta in bundle.blocks:
                total += rt.get(ref, timeout=120).num_rows
        assert total == 100000
        master = rt.get_actor("recon_SPARK_MASTER")
        assert len(rt.get(master.get_restarted_executors.remote())) == 1
    finally:
        raydp.stop_spark()
        cluster.shutdown()


def test_dynamic_allocation_request_and_kill(ray_cluster):
    spark = raydp.init_spark("dyn", 1, 1, "300M", configs={
        "spark.dynamicAllocation.enabled": "true", "spark.dynamicAllocation.maxExecutors": "3",
        "spark.dynamicAllocation.initialExecutors": "1"})
    backend = spark._backend
    assert backend.executor_actor_ids() == [0]
    backend.request_total_executors(3)
    assert backend.executor_actor_ids() == [0, 1, 2]
    assert spark.range(0, 1000, numPartitions=6).count() == 1000
    backend.kill_executors([1])
    assert backend.executor_actor_ids() == [0, 2]
    assert spark.range(0, 1000, numPartitions=6).count() == 1000
    raydp.stop_spark()


def test_external_shuffle_service_keeps_blocks_after_executor_kill(ray_cluster):
    """``spark.shuffle.service.enabled``: one service actor per node owns the shuffle output, so it
    stays readable after dynamic allocation kills the executor that wrote it (reference
    ``RayAppMaster.scala:151-167``, ``RayExternalShuffleService.scala:26-57``)."""
    from raydp_b200.spark.shuffle_service import service_name
    spark = raydp.init_spark("shuf", 2, 1, "300M", configs={
        "spark.dynamicAllocation.enabled": "true", "spark.dynamicAllocation.maxExecutors": "2",
        "spark.dynamicAllocation.initialExecutors": "2", "spark.shuffle.service.enabled": "true"})
    master = rt.get_actor("shuf_SPARK_MASTER")
    nodes = rt.get(master.shuffle_service_nodes.remote())
    assert len(nodes) == 1
    service = rt.get_actor(service_name("shuf_SPARK_MASTER", nodes[0]))
    shuffled = spark.range(0, 4000, numPartitions=4).repartition(3, "id")
    assert rt.get(service.block_count.remote()) >= 3
    spark._backend.kill_executors([1])
    assert spark._backend.executor_actor_ids() == [0]
    assert shuffled.count() == 4000 and shuffled.rdd.getNumPartitions() == 3
    raydp.stop_spark()


def test_spark_defaults_conf_is_loaded_with_user_precedence(ray_cluster, tmp_path, monkeypatch):
    """``$SPARK_CONF_DIR/spark-defaults.conf`` feeds the session configuration; ``init_spark(configs=)``
    wins on conflicts (reference ``doc/spark_on_ray.md:92``, ``RayAppMaster.scala:61``)."""
    (tmp_path / "spark-defaults.conf").write_text(
        "# comment\nspark.sql.shuffle.partitions  7\nspark.custom.flag=from-file\nnot.spark.key 1\n")
    monkeypatch.setenv("SPARK_CONF_DIR", str(tmp_path))
    spark = raydp.init_spark("defaults", 1, 1, "300M", configs={"spark.custom.flag": "from-user"})
    try:
        conf = spark.sparkContext.getConf()
        assert conf.get("spark.sql.shuffle.partitions") == "7"
        assert conf.get("spark.custom.flag") == "from-user"
        assert conf.get("not.spark.key", "absent") == "absent"
    finally:
        raydp.stop_spark()


def test_wider_sql_surface_runs_on_executors(spark_on_ray_2_executors, tmp_path):
    """Gathered aggregates, higher-order functions, windows, partitioned writes and SQL text through the
    executor actors (expressions are pickled to the executors; the function library is imported there)."""
    from raydp_b200.sql import functions as F
    from raydp_b200.sql.window import Window
    spark = spark_on_ray_2_executors
    df = spark.createDataFrame([("a", 1, [1, 2]), ("a", 2, [3]), ("b", 5, []), (None, 9, None)],
                               ["k", "v", "xs"]).repartition(2)
    agg = df.groupBy("k").agg(F.collect_list("v").alias("l"), F.percentile_approx("v", 0.5).alias("p"),
                              F.max_by("v", "v").alias("m")).orderBy("k").collect()
    assert [(r.k, sorted(r.l), r.p, r.m) for r in agg] == [(None, [9], 9, 9), ("a", [1, 2], 1, 2), ("b", [5], 5, 5)]
    hof = df.select("k", "v", F.transform("xs", lambda x: x * F.col("v")).alias("t"),
                    F.concat_ws("-", "k", F.lit("z")).alias("c")).orderBy("v").collect()
    assert [(r.t, r.c) for r in hof] == [([1, 2], "a-z"), ([6], "a-z"), ([], "b-z"), (None, "z")]
    ranked = df.withColumn("r", F.rank().over(Window.partitionBy("k").orderBy(F.desc("v")))).orderBy("k", "r").collect()
    assert [(r.k, r.v, r.r) for r in ranked] == [(None, 9, 1), ("a", 2, 1), ("a", 1, 2), ("b", 5, 1)]
    df.select("k", "v").write.partitionBy("k").json(str(tmp_path / "o"))
    back = sorted((tuple(r) for r in spark.read.json(str(tmp_path / "o")).collect()), key=repr)
    assert back == [(1, "a"), (2, "a"), (5, "b"), (9, None)]
    df.createOrReplaceTempView("wide_t")
    rows = spark.sql("select k, sum(v) s from wide_t group by k having sum(v) > 2 order by s desc").collect()
    assert [(r.k, r.s) for r in rows] == [(None, 9), ("b", 5), ("a", 3)]


def test_shuffle_keys_hash_the_same_on_every_executor(spark_on_ray_2_executors):
    """String keys must land in the same shuffle partition whichever executor process hashes them
    (``hash(str)`` is salted per process): distinct / dropDuplicates / groupBy after repartition rely on it."""
    import subprocess
    import sys
    spark = spark_on_ray_2_executors
    words = [f"w{i % 7}" for i in range(200)]
    df = spark.createDataFrame([(w, i) for i, w in enumerate(words)], ["w", "i"]).repartition(4)
    assert df.select("w").distinct().count() == 7
    assert df.dropDuplicates(["w"]).count() == 7
    parts = df.repartition(3, "w").rdd.mapPartitions(lambda rows: [sorted({r.w for r in rows})]).collect()
    assert sorted(w for p in parts for w in p) == sorted(set(words))          # no key in two partitions
    code = ("import pyarrow as pa; from raydp_b200.sql.executor import _stable_hashes; "
            "print(_stable_hashes(pa.array(['a', 'b', None, 'a'])).tolist())")
    outs = {subprocess.run([sys.executable, "-c", code], capture_output=True, text=True,
                           env=dict(__import__('os').environ, PYTHONHASHSEED=seed)).stdout for seed in ("1", "2")}
    assert len(outs) == 1 and "[" in outs.pop()


def _backend_queries(spark):
    import numpy as np
    from raydp_b200.sql import functions as F
    from raydp_b200.sql.window import Window
    rng = np.random.default_rng(0)
    rows = [(f"k{int(rng.integers(0, 9))}", int(rng.integers(0, 50)), float(rng.normal()),
             None if i % 11 == 0 else f"2020-0{1 + i % 9}-1{i % 9}", [int(i % 3), int(i % 5)]) for i in range(500)]
    df = spark.createDataFrame(rows, ["k", "v", "x", "d", "a"]).repartition(5)
    other = spark.createDataFrame([(f"k{i}", i * 10) for i in range(0, 12, 2)], ["k", "w"])
    df.createOrReplaceTempView("t_diff")
    other.createOrReplaceTempView("o_diff")
    out = {
        "group": df.groupBy("k").agg(F.sum("v"), F.avg("x"), F.countDistinct("v"), F.collect_set("v"),
                                     F.percentile_approx("x", 0.5), F.stddev("x"), F.max_by("v", "x")),
        "distinct": df.select("k", "d").distinct(),
        "dropdup": df.dropDuplicates(["k"]).select("k"),
        "join_left_nested_payload": df.join(other, "k", "left").groupBy("w").count(),
        "join_full": df.select("k").distinct().join(other, "k", "full"),
        "window": df.withColumn("r", F.dense_rank().over(Window.partitionBy("k").orderBy(F.desc("v"))))
                    .filter("r <= 2").select("k", "v", "r").distinct(),
        "sql": spark.sql("select t.k, count(*) c, sum(w) sw from t_diff t join o_diff o on t.k = o.k "
                         "where v > 10 group by t.k having count(*) > 3"),
        "explode": df.select("k", F.explode("a").alias("e")).groupBy("k", "e").count(),
        "dates": df.groupBy(F.month(F.to_date("d")).alias("m")).agg(F.count("*").alias("n"), F.min("d")),
        "pivot": df.groupBy("k").pivot("v", [1, 2, 3]).agg(F.count("*")),
        "setops": df.select("k").subtract(other.select("k")),
        "rollup": df.rollup("k").agg(F.sum("v").alias("s")),
        "repartition_by_key": df.repartition(3, "k").groupBy(F.spark_partition_id().alias("p"))
                                .agg(F.countDistinct("k").alias("nk")).agg(F.sum("nk")),
        "top": df.orderBy("v", "x").limit(7).select("v", "x"),
    }

    def norm(frame):
        def key(v):
            if isinstance(v, float):
                return round(v, 9)
            if isinstance(v, list):
                return tuple(sorted(v, key=repr))
            return v
        return sorted((tuple(key(v) for v in r) for r in frame.collect()), key=repr)
    return {name: norm(frame) for name, frame in out.items()}


def test_local_and_cluster_backends_agree(spark_session, spark_on_ray_2_executors):
    """The same queries through the in-process backend and through two executor actors: aggregation,
    distinct, joins (nested payload columns included), windows, SQL, generators, pivots, set operations."""
    local = _backend_queries(spark_session)
    cluster = _backend_queries(spark_on_ray_2_executors)
    for name in local:
        assert local[name] == cluster[name], name
    assert local["repartition_by_key"] == [(9,)] and len(local["group"]) == 9 and len(local["sql"]) > 0
