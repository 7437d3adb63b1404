"""Data sources (parquet / orc / json / csv / text, ``partitionBy`` + partition discovery), the RDD surface,
SparkContext helpers and ``spark.udf.register``."""
import os

import pytest

from raydp_b200.sql import functions as F
from raydp_b200.sql.session import SparkSession
from raydp_b200.sql.types import Row


@pytest.fixture(scope="module")
def spark(spark_session):
    return spark_session


@pytest.fixture(scope="module")
def df(spark):
    return spark.createDataFrame([("a", 1, 10.0), ("a", 2, 20.0), ("b", 5, None), (None, 9, 9.0)], ["k", "v", "x"])


def rows(frame, key="v"):
    return sorted((tuple(r) for r in frame.collect()), key=repr)


def test_partitioned_parquet_round_trip(spark, df, tmp_path):
    target = str(tmp_path / "p")
    df.write.partitionBy("k").parquet(target)
    assert sorted(os.listdir(target)) == ["_SUCCESS", "k=__HIVE_DEFAULT_PARTITION__", "k=a", "k=b"]
    back = spark.read.parquet(target)
    assert back.columns == ["v", "x", "k"]                       # partition columns come last, as in Spark
    assert rows(back) == rows(df.select("v", "x", "k"))
    assert rows(spark.read.parquet(os.path.join(target, "k=a"))) == [(1, 10.0), (2, 20.0)]
    assert back.filter("k = 'b'").count() == 1
    with pytest.raises(FileExistsError):
        df.write.parquet(target)
    df.write.mode("append").partitionBy("k").parquet(target)
    assert spark.read.parquet(target).count() == 8


def test_json_text_orc_csv_round_trips(spark, df, tmp_path):
    df.write.json(str(tmp_path / "j"))
    first = sorted(f for f in os.listdir(tmp_path / "j") if f.endswith(".json"))[0]
    assert open(tmp_path / "j" / first).read().splitlines()[0] == '{"k":"a","v":1,"x":10.0}'
    assert rows(spark.read.json(str(tmp_path / "j")).select("k", "v", "x")) == rows(df)
    typed = spark.read.schema("v int, k string").json(str(tmp_path / "j"))
    assert typed.dtypes == [("v", "int"), ("k", "string")] and typed.count() == 4
    df.select("k").fillna("?").write.text(str(tmp_path / "t"))
    assert sorted(r.value for r in spark.read.text(str(tmp_path / "t")).collect()) == ["?", "a", "a", "b"]
    df.write.format("orc").mode("overwrite").save(str(tmp_path / "o"))
    assert rows(spark.read.format("orc").load(str(tmp_path / "o"))) == rows(df)
    df.write.option("header", True).csv(str(tmp_path / "c"))
    back = spark.read.option("header", True).option("inferSchema", True).csv(str(tmp_path / "c"))
    assert back.columns == ["k", "v", "x"] and back.count() == 4
    df.write.partitionBy("k", "v").format("json").save(str(tmp_path / "pj"))
    nested = spark.read.json(str(tmp_path / "pj"))
    assert set(nested.columns) == {"x", "k", "v"} and rows(nested.select("k", "v", "x")) == rows(df)
    assert dict(nested.dtypes)["v"] == "int"                       # inferred from the directory names


def test_rdd_surface(spark):
    sc = spark.sparkContext
    r = sc.parallelize(range(10), 3)
    assert r.map(lambda x: x * 2).filter(lambda x: x % 3 == 0).collect() == [0, 6, 12, 18]
    assert (r.reduce(lambda a, b: a + b), r.sum(), r.max(), r.min(), r.mean(), r.count()) == (45, 45, 9, 0, 4.5, 10)
    assert r.take(3) == [0, 1, 2] and r.first() == 0 and r.top(2) == [9, 8] and r.takeOrdered(2, key=lambda x: -x) == [9, 8]
    assert r.glom().collect() == [[0, 1, 2, 3], [4, 5, 6], [7, 8, 9]]
    assert r.aggregate((0, 0), lambda a, x: (a[0] + x, a[1] + 1), lambda a, b: (a[0] + b[0], a[1] + b[1])) == (45, 10)
    assert r.fold(0, lambda a, b: a + b) == 45 and r.zipWithIndex().take(2) == [(0, 0), (1, 1)]
    assert r.union(r).distinct().count() == 10 and r.subtract(sc.parallelize([0, 1])).count() == 8
    assert r.sortBy(lambda x: -x).take(2) == [9, 8] and r.keyBy(lambda x: x % 2).countByKey() == {0: 5, 1: 5}
    kv = sc.parallelize([("a", 1), ("b", 2), ("a", 3)])
    assert sorted(kv.reduceByKey(lambda a, b: a + b).collect()) == [("a", 4), ("b", 2)]
    assert sorted(kv.groupByKey().mapValues(list).collect()) == [("a", [1, 3]), ("b", [2])]
    assert kv.join(sc.parallelize([("a", "x")])).collect() == [("a", (1, "x")), ("a", (3, "x"))]
    assert sorted(kv.leftOuterJoin(sc.parallelize([("a", "x")])).collect())[-1] == ("b", (2, None))
    assert kv.collectAsMap() == {"a": 3, "b": 2} and kv.keys().collect() == ["a", "b", "a"]
    assert sorted(kv.combineByKey(lambda v: [v], lambda c, v: c + [v], lambda a, b: a + b).collect()) == \
        [("a", [1, 3]), ("b", [2])]
    frame = spark.createDataFrame(kv.map(lambda t: Row(k=t[0], v=t[1])))
    assert frame.columns == ["k", "v"] and frame.count() == 3
    ids = spark.range(5).rdd
    assert ids.map(lambda row: row.id * 2).collect() == [0, 2, 4, 6, 8] and ids.first() == Row(id=0)
    assert ids.zipWithIndex().take(1) == [(Row(id=0), 0)]


def test_spark_context_helpers_and_udf_registration(spark, tmp_path):
    sc = spark.sparkContext
    sc.parallelize(["a b", "c"], 2).saveAsTextFile(str(tmp_path / "t"))
    assert sc.textFile(str(tmp_path / "t")).flatMap(str.split).collect() == ["a", "b", "c"]
    assert sc.range(3).collect() == [0, 1, 2] and sc.emptyRDD().isEmpty() and sc.version == spark.version
    lookup = sc.broadcast({0: "zero"})
    assert spark.range(2).rdd.map(lambda r: lookup.value.get(r.id)).collect() == ["zero", None]
    acc = sc.accumulator(0)
    acc += 5
    acc.add(2)
    assert acc.value == 7
    spark.udf.register("plus_one", lambda x: x + 1, "int")
    spark.range(3).createOrReplaceTempView("r_udf")
    assert [r.p for r in spark.sql("select plus_one(id) as p from r_udf").collect()] == [1, 2, 3]
    assert SparkSession.active() is spark


def test_rows_and_nested_schemas(spark):
    person = Row("name", "age")
    assert person("Al", 3) == Row(name="Al", age=3) and person("Al", 3).asDict() == {"name": "Al", "age": 3}
    assert Row(a=1, b=Row(c=2)).asDict(True) == {"a": 1, "b": {"c": 2}} and "a" in Row(a=1)
    nested = spark.createDataFrame([(1, [1, 2], {"k": 1.5})], "id int, xs array<int>, m map<string,double>")
    assert nested.dtypes == [("id", "int"), ("xs", "array<int>"), ("m", "map<string,double>")]
    assert nested.select(F.col("m")["k"].alias("v"), F.size("xs").alias("n")).collect()[0] == Row(v=1.5, n=2)


def test_pandas_on_spark_surface(spark):
    """``pyspark.pandas`` stand-in: lazy column expressions, reductions, accessors, groupby, merge."""
    from raydp_b200.sql import pandas_api as ps
    df = ps.DataFrame({"k": ["a", "a", "b", None], "v": [1, 2, 5, 9], "x": [10.0, 20.0, None, 9.0],
                       "t": ["2020-01-05 10:00:00"] * 4})
    df["y"] = df["v"] * 2 + 1
    assert df.shape == (4, 5) and df.dtypes.to_dict()["y"] == "bigint"
    assert df[df.v > 1][["k", "v"]].to_pandas()["v"].tolist() == [2, 5, 9]
    assert (df.v.sum(), df.v.mean(), df.x.max(), df.k.nunique()) == (17, 4.25, 20.0, 2)
    assert df.k.str.upper().to_pandas().tolist()[:3] == ["A", "A", "B"]
    assert df.t.astype("datetime64[ns]").dt.year.to_pandas().tolist() == [2020] * 4
    assert ps.to_datetime(df.t).dt.dayofweek.head(1).tolist() == [6]                       # a Sunday
    by_k = df.groupby("k").sum().to_pandas()
    assert by_k[by_k.k == "a"][["v", "x", "y"]].values.tolist() == [[3, 30.0, 8]]
    assert df.groupby("k")["v"].agg(["min", "max"]).to_pandas().values.tolist()[1:] == [["a", 1, 2], ["b", 5, 5]]
    assert df.groupby("k").agg(total=("v", "sum")).to_pandas()["total"].tolist() == [9, 3, 5]
    assert df.sort_values("v", ascending=False).head(1)["v"].tolist() == [9]
    assert df.fillna(0).x.sum() == 39.0 and df.dropna().shape == (2, 5) and "t" not in df.drop(columns=["t"]).columns
    other = ps.DataFrame({"k": ["a", "c"], "name": ["A", "C"], "v": [7, 8]})
    merged = df.merge(other, on="k", how="left")
    assert merged.columns == ["k", "v_x", "x", "t", "y", "name", "v_y"] and len(df.merge(other, on="k")) == 2
    assert df.v.map({1: "one"}).to_pandas().tolist()[0] == "one"
    assert df.x.fillna(-1).clip(0, 15).to_pandas().tolist() == [10.0, 15.0, 0.0, 9.0]
    assert df.assign(z=lambda d: d.v + d.y).z.sum() == 55 and df.describe().loc["mean", "v"] == 4.25
    assert ps.get_dummies(df[["k", "v"]]).columns == ["v", "k_a", "k_b"]
    assert ps.concat([df[["k"]], df[["k"]]]).shape == (8, 1) and df.loc[df.v > 2, ["k"]].shape == (2, 1)

    def row_total(r) -> float:
        return float(r["v"]) + float(r["y"])
    assert df.apply(row_total, axis=1).to_pandas().tolist() == [4.0, 7.0, 16.0, 28.0]
    assert ps.range(10).to_spark().count() == 10 and isinstance(df.to_spark(), type(spark.range(1)))


def test_catalog_api(spark, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)                                   # the default warehouse is ./spark-warehouse
    spark.range(3).createOrReplaceTempView("cat_v")
    cat = spark.catalog
    cols = cat.listColumns("cat_v")
    assert [(c.name, c.dataType) for c in cols] == [("id", "bigint")] and cat.getTable("cat_v").isTemporary
    assert cat.functionExists("concat_ws") and len(cat.listFunctions(pattern="array_*")) >= 15
    assert cat.databaseExists("default") and cat.getDatabase("default").name == "default"
    spark.range(5).write.parquet(str(tmp_path / "ext"))
    assert cat.createTable("cat_ext", path=str(tmp_path / "ext")).count() == 5
    assert spark.sql("select count(*) from cat_ext").collect()[0][0] == 5
    assert cat.createTable("cat_managed", schema="a int, b string").columns == ["a", "b"] and cat.tableExists("cat_managed")
    cat.cacheTable("cat_v")
    assert cat.isCached("cat_v")
    cat.clearCache()
    spark.sql("drop table cat_managed")
    assert cat.dropTempView("cat_v") and not cat.tableExists("cat_v")


def test_sql_ddl_and_dml_statements(spark, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)

    def q(text):
        return [tuple(r) for r in spark.sql(text).collect()]
    q("create table ddl_people (id int, name string, tags array<string>) using parquet")
    q("insert into ddl_people values (1, 'al', array('x')), (2, upper('bo'), array())")
    assert q("select * from ddl_people order by id") == [(1, "al", ["x"]), (2, "BO", [])]
    q("create or replace temp view ddl_young as select id, name from ddl_people where id < 2")
    assert q("select * from ddl_young") == [(1, "al")]
    assert q("describe ddl_people") == [("id", "int", None), ("name", "string", None), ("tags", "array<string>", None)]
    assert q("show columns in ddl_people") == [("id",), ("name",), ("tags",)]
    q("insert overwrite ddl_people select id + 10, name, tags from ddl_people")     # reads the table it replaces
    assert q("select id from ddl_people order by id") == [(11,), (12,)]
    assert dict(spark.table("ddl_people").dtypes)["id"] == "int"                     # typed by the table
    q("truncate table ddl_people")
    assert q("select count(*) from ddl_people") == [(0,)]
    q("create table if not exists ddl_people (id int)")                               # no-op
    assert q("set spark.sql.shuffle.partitions=7") == [("spark.sql.shuffle.partitions", "7")]
    assert q("set spark.sql.shuffle.partitions") == [("spark.sql.shuffle.partitions", "7")]
    assert q("values (1, 'a'), (2, 'b')") == [(1, "a"), (2, "b")]
    q("cache table ddl_young")
    q("drop table ddl_people")
    assert not spark.catalog.tableExists("ddl_people")


def test_nested_type_inference_from_python_objects(spark):
    import datetime as dt
    from raydp_b200.sql import functions as F
    df = spark.createDataFrame([Row(a=1, t=dt.datetime(2020, 1, 1), m={"k": 1.0}, l=[Row(z=1), Row(z=2)],
                                    s=Row(p="x", q=[1, 2]))])
    assert dict(df.dtypes) == {"a": "bigint", "t": "timestamp", "m": "map<string,double>",
                               "l": "array<struct<z:bigint>>", "s": "struct<p:string,q:array<bigint>>"}
    row = df.collect()[0]
    assert row.m == {"k": 1.0} and row.l == [Row(z=1), Row(z=2)] and row.s.q == [1, 2]
    d2 = spark.createDataFrame([{"a": 1, "b": {"c": [1, 2]}}, {"a": 2, "b": {"d": []}}])
    assert dict(d2.dtypes)["b"] == "map<string,array<bigint>>"
    assert [tuple(r) for r in d2.select(F.col("b")["c"].alias("c"), F.map_keys("b").alias("k")).collect()] == \
        [([1, 2], ["c"]), (None, ["d"])]
    d3 = spark.createDataFrame([(1, ("x", 2))], "id int, s struct<a:string,b:int>")
    assert d3.collect() == [Row(id=1, s=Row(a="x", b=2))] and tuple(d3.select("s.*").collect()[0]) == ("x", 2)
