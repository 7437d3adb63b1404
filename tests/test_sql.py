"""DataFrame engine on the local[2] backend (SURVEY.md Appendix C surface)."""
import os

import pyarrow as pa
import pytest

from raydp_b200.sql import SparkSession, Window, functions as F, types as T
from raydp_b200.sql.backend import get_initial_target_executor_number
from raydp_b200.sql.functions import col, lit, udf

REF_CSV = "/root/reference/examples/fake_nyctaxi.csv"


def test_range_count_partitions(spark_session):
    df = spark_session.range(0, 1000, numPartitions=4)
    assert df.count() == 1000
    assert df.rdd.getNumPartitions() == 4
    assert [f.name for f in df.schema] == ["id"]
    assert spark_session.range(10).count() == 10
    assert spark_session.range(0, 10, 3).collect()[-1].id == 9


def test_expressions_filter_select(spark_session):
    df = spark_session.range(0, 100).withColumn("x", col("id") * 2 + 1) \
        .withColumn("half", col("id") / 2).filter((col("x") > 10) & (col("id") < 50))
    rows = df.select("id", (col("x") - 1).alias("y"), F.abs(lit(-3)).alias("three")).collect()
    assert len(rows) == 45
    assert rows[0].y == rows[0].id * 2 and rows[0].three == 3
    assert dict(df.dtypes)["half"] == "double"
    assert df.drop("x").columns == ["id", "half"]
    assert df.withColumnRenamed("x", "xx").columns == ["id", "xx", "half"]


def test_rand_is_deterministic_and_split_is_a_partition(spark_session):
    df = spark_session.range(0, 2000, numPartitions=3).withColumn("r", F.rand(7))
    a = [r.r for r in df.collect()]
    b = [r.r for r in df.collect()]
    assert a == b and 0 <= min(a) and max(a) < 1
    s1, s2 = df.randomSplit([0.8, 0.2], seed=3)
    ids1, ids2 = {r.id for r in s1.collect()}, {r.id for r in s2.collect()}
    assert not (ids1 & ids2) and len(ids1 | ids2) == 2000
    assert 1400 < len(ids1) < 1800


def test_udf_and_time_functions(spark_session):
    @udf("int")
    def night(hour, weekday):
        return int(1) if (16 <= hour <= 20) and (weekday < 5) else int(0)

    df = spark_session.createDataFrame(
        [("2011-11-28 23:11:35 UTC",), ("2013-10-29 18:00:01 UTC",)], ["ts"])
    out = df.withColumn("h", F.hour(col("ts"))).withColumn("dow", F.dayofweek(col("ts")) - 2) \
        .withColumn("dom", F.dayofmonth(col("ts"))).withColumn("woy", F.weekofyear(col("ts"))) \
        .withColumn("q", F.quarter(col("ts"))).withColumn("y", F.year(col("ts"))) \
        .withColumn("m", F.month(col("ts"))).withColumn("night", night("h", "dow")).collect()
    assert (out[0].h, out[0].dow, out[0].dom, out[0].woy, out[0].q, out[0].y, out[0].m) == \
        (23, 0, 28, 48, 4, 2011, 11)
    assert (out[1].h, out[1].dow, out[1].night) == (18, 1, 1)


def test_groupby_agg_join_sort(spark_session):
    df = spark_session.range(0, 99).withColumn("k", col("id") % 3)
    g = df.groupBy("k").agg(F.count("*").alias("n"), F.avg("id"), F.max("id"), F.sum("id"))
    rows = {r.k: r for r in g.collect()}
    assert rows[0].n == 33 and rows[1]["avg(id)"] == 49.0 and rows[2]["max(id)"] == 98
    assert df.groupBy("k").count().count() == 3
    assert df.agg(F.count("*").alias("c")).collect()[0].c == 99
    j = df.join(g.select("k", "n"), "k", "left")
    assert j.count() == 99 and set(j.columns) == {"id", "k", "n"}
    top = df.orderBy(col("id").desc()).take(2)
    assert [r.id for r in top] == [98, 97]
    w = Window.partitionBy("k").orderBy(F.desc("id"))
    ranked = df.withColumn("rn", F.row_number().over(w)).filter(col("rn") == 1).collect()
    assert sorted(r.id for r in ranked) == [96, 97, 98]


def test_repartition_fillna_sample_distinct(spark_session):
    df = spark_session.range(0, 1000, numPartitions=2)
    rp = df.repartition(5)
    assert rp.rdd.getNumPartitions() == 5 and rp.count() == 1000
    assert sorted(r.id for r in rp.collect()) == list(range(1000))
    byk = df.withColumn("k", col("id") % 4).repartition(3, "k")
    assert byk.count() == 1000
    nulls = spark_session.createDataFrame([(1, None), (2, 3.0)], T.StructType([
        T.StructField("a", T.IntegerType()), T.StructField("b", T.DoubleType())]))
    assert [r.b for r in nulls.fillna(0, ["b"]).collect()] == [0.0, 3.0]
    assert nulls.dropna().count() == 1
    assert 50 < df.sample(fraction=0.1, seed=1).count() < 160
    assert df.withColumn("k", col("id") % 7).select("k").distinct().count() == 7
    assert df.sortWithinPartitions("id", ascending=False).take(1)[0].id in (499, 999)


def test_cache_persist_unpersist(spark_session):
    df = spark_session.range(0, 100).withColumn("r", F.rand())
    df.cache()
    assert df.is_cached
    first = [r.r for r in df.collect()]
    assert [r.r for r in df.collect()] == first
    df.unpersist()
    assert not df.is_cached and df.count() == 100


def test_csv_reader_infers_reference_schema(spark_session):
    if not os.path.exists(REF_CSV):
        pytest.skip("reference example csv not mounted")
    data = spark_session.read.format("csv").option("header", "true") \
        .option("inferSchema", "true").load("file://" + REF_CSV)
    types = dict(data.dtypes)
    assert types["fare_amount"] == "double" and types["passenger_count"] == "int"
    assert types["pickup_datetime"] == "timestamp" and types["key"] == "string"
    assert data.count() == 500


def test_parquet_roundtrip_and_modes(spark_session, tmp_path):
    path = str(tmp_path / "pq")
    df = spark_session.range(0, 100, numPartitions=3).withColumn("v", col("id") * 1.5)
    df.write.parquet(path, compression="snappy")
    with pytest.raises(FileExistsError):
        df.write.parquet(path)
    df.write.mode("overwrite").parquet(path)
    back = spark_session.read.parquet(path)
    assert back.count() == 100 and back.rdd.getNumPartitions() == 3
    assert os.path.exists(os.path.join(path, "_SUCCESS"))
    tsv = tmp_path / "x.tsv"
    tsv.write_text("1\t2.5\n3\t4.5\n")
    schema = T.StructType([T.StructField("a", T.IntegerType()), T.StructField("b", T.DoubleType())])
    got = spark_session.read.schema(schema).option("sep", "\t").csv(str(tsv)).collect()
    assert [(r.a, r.b) for r in got] == [(1, 2.5), (3, 4.5)]


def test_rdd_and_create_dataframe(spark_session):
    sc = spark_session.sparkContext
    rdd = sc.parallelize([1, 2, 3, 4, 5, 6], 3)
    assert rdd.getNumPartitions() == 3
    assert rdd.map(lambda x: x * 2).collect() == [2, 4, 6, 8, 10, 12]
    assert rdd.mapPartitions(lambda it: [sum(it)]).collect() == [3, 7, 11]
    df = spark_session.createDataFrame(sc.parallelize([(1, "a"), (2, "b")]), ["n", "s"])
    assert [tuple(r) for r in df.collect()] == [(1, "a"), (2, "b")]
    pdf = df.toPandas()
    assert list(pdf.columns) == ["n", "s"]
    assert spark_session.createDataFrame(pdf).count() == 2


def test_map_in_pandas_and_posexplode(spark_session):
    df = spark_session.range(0, 10)
    schema = T.StructType([T.StructField("id", T.LongType()), T.StructField("sq", T.LongType())])

    def fn(frames):
        for pdf in frames:
            pdf["sq"] = pdf["id"] ** 2
            yield pdf
    assert [r.sq for r in df.mapInPandas(fn, schema).collect()] == [i * i for i in range(10)]
    ex = df.filter(col("id") < 2).withColumn("b", col("id") + 10) \
        .select("id", F.posexplode(F.array("id", "b")))
    assert [(r.id, r.pos, r.col) for r in ex.collect()] == [(0, 0, 0), (0, 1, 10), (1, 0, 1), (1, 1, 11)]


def test_conf_and_context(spark_session):
    spark_session.conf.set("spark.sql.session.timeZone", "UTC")
    assert spark_session.conf.get("spark.sql.session.timeZone") == "UTC"
    sc = spark_session.sparkContext
    assert sc.appName == "RayDP test" and sc.applicationId.startswith("app-")
    assert sc.getConf().get("does.not.exist", "dflt") == "dflt"
    assert spark_session.sql("select 1 as one, 2 * 3 as six").collect()[0].six == 6


def test_initial_executor_number():
    # same four cases as reference TestRayCoarseGrainedSchedulerBackend.java:32-65
    assert get_initial_target_executor_number({}) == 2
    assert get_initial_target_executor_number({"spark.executor.instances": "-1"}) == 2
    assert get_initial_target_executor_number({"spark.executor.instances": "5"}) == 5
    dyn = {"spark.dynamicAllocation.enabled": "true", "spark.dynamicAllocation.minExecutors": "1",
           "spark.dynamicAllocation.initialExecutors": "3", "spark.dynamicAllocation.maxExecutors": "4"}
    assert get_initial_target_executor_number(dyn) == 3


def test_sql_front_end_views_tables_and_expressions(spark_session, tmp_path):
    """``spark.sql`` / ``filter(str)`` / ``selectExpr`` / ``F.expr`` / temp views / warehouse tables
    (reference usage: ``spark.sql("select * from db.xxx")`` with ``enable_hive``, doc/spark_on_ray.md:187-202)."""
    from raydp_b200.sql import functions as F
    spark = spark_session
    spark.conf.set("spark.sql.warehouse.dir", str(tmp_path / "wh"))
    df = spark.createDataFrame([(i, i % 3, float(i) * 0.5, "n%d" % (i % 2)) for i in range(30)], ["id", "k", "v", "s"])
    df.createOrReplaceTempView("t")
    got = spark.sql("SELECT id, v * 2 AS w FROM t WHERE id BETWEEN 5 AND 9 AND s = 'n1' ORDER BY id DESC LIMIT 2").collect()
    assert [(r.id, r.w) for r in got] == [(9, 9.0), (7, 7.0)]
    agg = spark.sql("select k, count(*) as n, sum(v) as total from t group by k having n > 0 order by k").collect()
    assert [(r.k, r.n) for r in agg] == [(0, 10), (1, 10), (2, 10)]
    assert abs(agg[1].total - sum(i * 0.5 for i in range(30) if i % 3 == 1)) < 1e-9
    assert spark.sql("select count(*) as n from t where not (k in (0, 1)) or id is null").collect()[0].n == 10
    assert spark.sql("select distinct s from t order by s").collect()[1].s == "n1"
    assert spark.sql("select case when k = 0 then 'zero' else 'other' end as c from t order by id limit 1").collect()[0].c == "zero"
    # string predicates / expressions on the DataFrame API
    assert df.filter("v >= 10 and k <> 2").count() == len([i for i in range(30) if i * 0.5 >= 10 and i % 3 != 2])
    row = df.selectExpr("id", "cast(v as int) + 1 as vi", "abs(id - 10) d").filter("id = 3").collect()[0]
    assert (row.id, row.vi, row.d) == (3, 2, 7)
    assert df.withColumn("e", F.expr("id % 4")).filter("e = 3").count() == 7
    # warehouse tables in a database, visible to a fresh lookup (what enable_hive sessions persist to)
    spark.sql("create database if not exists db")
    spark.sql("create table db.xxx as select id, v from t where id < 4")
    spark.sql("insert into db.xxx select id, v from t where id >= 28")
    assert sorted(r.id for r in spark.sql("select * from db.xxx").collect()) == [0, 1, 2, 3, 28, 29]
    assert spark.table("db.xxx").count() == 6 and spark.read.table("db.xxx").count() == 6
    df.limit(3).write.mode("overwrite").saveAsTable("plain")
    names = {(r.tableName, r.isTemporary) for r in spark.sql("show tables").collect()}
    assert ("plain", False) in names and ("t", True) in names
    assert ("xxx", "db", False) in spark.catalog.listTables("db") and "db" in spark.catalog.listDatabases()
    spark.sql("drop table if exists db.xxx")
    assert not spark.catalog.tableExists("db.xxx")
    # equi-joins in SQL text
    spark.createDataFrame([(0, "zero"), (1, "one"), (7, "seven")], ["kk", "name"]).createOrReplaceTempView("names")
    j = spark.sql("select id, name from t left join names n on t.k = n.kk where id < 3 order by id").collect()
    assert [(r.id, r.name) for r in j] == [(0, "zero"), (1, "one"), (2, None)]
    assert spark.sql("select count(*) as n from t join names on names.kk = t.k").collect()[0].n == 20
    # uncorrelated sub-queries are evaluated first
    assert spark.sql("select * from t where id in (select id from t where id < 4)").count() == 4
    import pytest as _pytest
    with _pytest.raises(ValueError):
        spark.sql("select * from t where id > interval 1 fortnight")


def test_map_in_pandas_accepts_ddl_schema(spark_session):
    df = spark_session.createDataFrame([(1, 2.0), (2, 3.0)], "a int, b double")

    def add(it):
        for pdf in it:
            yield pdf.assign(z=pdf["a"] * 10)

    out = df.mapInPandas(add, "a int, b double, z long").collect()
    assert [r.z for r in out] == [10, 20]


def test_drop_duplicates_subset_and_union_by_name(spark_session):
    df = spark_session.createDataFrame([(1, "a", 10), (1, "b", 11), (2, "c", 12), (2, "d", 13), (3, "e", 14)],
                                       ["k", "s", "v"])
    out = df.dropDuplicates(["k"]).collect()
    assert sorted(r.k for r in out) == [1, 2, 3] and {r.s for r in out} <= {"a", "b", "c", "d", "e"}
    assert df.dropDuplicates().count() == 5
    other = spark_session.createDataFrame([(99, "z")], ["v", "s"])
    merged = df.unionByName(other, allowMissingColumns=True)
    assert merged.count() == 6 and merged.columns == ["k", "s", "v"]
    last = [r for r in merged.collect() if r.v == 99][0]
    assert last.s == "z" and last.k is None
    import pytest as _pytest
    with _pytest.raises(ValueError):
        df.unionByName(other)


def test_join_variants(spark_session):
    a = spark_session.createDataFrame([(1, "x"), (2, "y"), (3, "z")], ["k", "a"]).repartition(2)
    b = spark_session.createDataFrame([(2, "B"), (3, "C"), (4, "D")], ["k", "b"])
    key = lambda r: tuple("" if v is None else str(v) for v in r)  # noqa: E731
    rows = lambda df: sorted((tuple(r) for r in df.collect()), key=key)  # noqa: E731
    assert rows(a.join(b, "k", "right")) == sorted([(2, "y", "B"), (3, "z", "C"), (4, None, "D")], key=key)
    assert len(rows(a.join(b, "k", "outer"))) == 4                      # unmatched rows exactly once each
    assert rows(a.join(b, a.k == b.k)) == [(2, "y", "B"), (3, "z", "C")]
    b2 = b.withColumnRenamed("k", "kk")
    assert a.join(b2, a.k == b2.kk, "left_semi").count() == 2
    assert a.crossJoin(b).count() == 9 and a.join(b, how="cross").count() == 9
    import pytest as _pytest
    with _pytest.raises(ValueError):                                  # same-named columns cannot be told apart
        a.join(b, a.k > b.k)
    assert rows(a.join(b2, a.k > b2.kk).select("k", "kk")) == [(3, 2)]   # a non-equality predicate


def test_frequency_dictionary_recipe(spark_session):
    """The DLRM notebook's categorical encoding (cells 6-7): posexplode(array(string columns)) alone in a
    select, groupBy/count, row_number window by descending count, join back, fillna."""
    df = spark_session.createDataFrame(
        [("a", "x"), ("a", "y"), ("b", "x"), ("a", None), ("c", "x")], ["c0", "c1"])
    pairs = (df.select(F.posexplode(F.array("c0", "c1")))
             .withColumnRenamed("pos", "column_id").withColumnRenamed("col", "value")
             .filter(col("value").isNotNull()).groupBy("column_id", "value").count())
    counts = {(r.column_id, r.value): r["count"] for r in pairs.collect()}
    assert counts == {(0, "a"): 3, (0, "b"): 1, (0, "c"): 1, (1, "x"): 3, (1, "y"): 1}
    ranked = pairs.withColumn("id", F.row_number().over(Window.partitionBy("column_id").orderBy(F.desc("count"))))
    top = {(r.column_id, r.value): r.id for r in ranked.collect()}
    assert top[(0, "a")] == 1 and top[(1, "x")] == 1 and top[(1, "y")] == 2
    d0 = ranked.filter(col("column_id") == 0).select("value", "id").withColumnRenamed("value", "c0")
    enc = df.join(d0.filter(col("id") <= 1), "c0", "left").drop("c0").withColumnRenamed("id", "c0").fillna(0, ["c0"])
    assert sorted(r.c0 for r in enc.collect()) == [0, 0, 1, 1, 1]


def test_null_conditions_follow_sql_three_valued_logic(spark_session):
    """A NULL condition in when()/CASE is "not matched"; IN / NOT IN over NULLs is NULL, never TRUE."""
    from raydp_b200.sql import functions as F
    spark = spark_session
    df = spark.createDataFrame([(1, 4.0), (2, None), (3, 2.0), (None, 7.0)], ["k", "v"])
    got = [r[0] for r in df.select(F.when(F.col("v") > 3, 1).otherwise(0).alias("w")).collect()]
    assert got == [1, 0, 0, 1]
    chained = df.select(F.when(F.col("v") > 3, "hi").when(F.col("v").isNull(), "n").alias("c")).collect()
    assert [r[0] for r in chained] == ["hi", "n", None, "hi"]
    df.createOrReplaceTempView("null_logic")
    rows = spark.sql("SELECT CASE WHEN v > 3 THEN 'hi' WHEN v IS NULL THEN 'n' END AS c FROM null_logic").collect()
    assert [r[0] for r in rows] == ["hi", "n", None, "hi"]
    assert df.filter(F.col("k").isin(1, None)).count() == 1          # NULL = NULL is not TRUE
    assert df.filter(~F.col("k").isin(1, None)).count() == 0         # 2 NOT IN (1, NULL) is NULL
    assert df.filter(~F.col("k").isin(1, 2)).count() == 1            # only k=3; the NULL key stays out
    assert spark.sql("SELECT * FROM null_logic WHERE k NOT IN (SELECT k FROM null_logic)").count() == 0
    assert spark.sql("SELECT * FROM null_logic WHERE k IN (SELECT k FROM null_logic)").count() == 3


def test_self_join_through_aliases(spark_session):
    from raydp_b200.sql import functions as F
    df = spark_session.createDataFrame([(1, 4.0), (2, 5.0), (3, 2.0)], ["k", "v"])
    a, b = df.alias("a"), df.alias("b")
    j = a.join(b, F.col("a.k") == F.col("b.k"))
    assert j.count() == 3 and j.columns == ["k", "v", "b.v"]
    assert sorted((r[0], r[1]) for r in j.select("a.v", "b.v").collect()) == [(2.0, 2.0), (4.0, 4.0), (5.0, 5.0)]
    shifted = a.join(b.withColumn("k", F.col("k") + 1).alias("b"), F.col("a.k") == F.col("b.k"), "left")
    rows = {r["k"]: r["b.v"] for r in shifted.collect()}
    assert rows == {1: None, 2: 4.0, 3: 5.0}
    other = spark_session.createDataFrame([(1, "x")], ["id", "tag"]).alias("o")
    assert a.join(other, F.col("a.k") == F.col("o.id")).select("k", "tag").collect()[0]["tag"] == "x"
    with pytest.raises(ValueError, match="does not relate"):
        a.join(b, F.col("a.k") == F.col("a.v"))


def test_div_first_and_distinct_details(spark_session):
    from raydp_b200.sql import functions as F
    spark = spark_session
    assert tuple(spark.sql("SELECT 7 div 2 AS a, -7 div 2 AS b").collect()[0]) == (3, -3)
    df = spark.createDataFrame([(1, None), (1, 2.0), (2, 3.0)], ["k", "v"]).coalesce(1)
    rows = df.groupBy("k").agg(F.first("v").alias("f"), F.first("v", True).alias("g")).orderBy("k").collect()
    assert [(r["f"], r["g"]) for r in rows] == [(None, 2.0), (3.0, 3.0)]
    assert df.agg(F.first("v")).columns == ["first(v)"]
    d = spark.createDataFrame([(float("nan"),), (float("nan"),), (0.0,), (-0.0,), (None,)], ["x"])
    assert d.distinct().count() == 3                                   # NaN, 0.0, NULL
