"""The wider pyspark.sql surface: scalar functions, aggregates, windows, generators, DataFrame helpers.

Expected values are the ones Spark's own documentation / doctests give for the same inputs
(``pyspark.sql.functions`` docstrings), or are computed by hand in the test.
"""
import datetime as dt
import math

import pytest

from raydp_b200.sql import functions as F
from raydp_b200.sql.session import SparkSession
from raydp_b200.sql.window import Window


@pytest.fixture(scope="module")
def spark(spark_session):
    return spark_session


@pytest.fixture(scope="module")
def people(spark):
    return spark.createDataFrame(
        [("a", 1, 10.0), ("a", 2, 20.0), ("a", 2, None), ("b", 5, 1.0), ("b", 3, 2.0), (None, 9, 9.0)],
        ["k", "v", "x"])


def one(df, *cols):
    return df.select(*cols).collect()[0]


def test_string_functions(spark):
    df = spark.createDataFrame([("  Spark SQL  ", "2015-04-08 13:08:15", "a.b.c.d")], ["s", "t", "p"])
    row = one(df, F.trim("s").alias("trim"), F.ltrim("s").alias("l"), F.rtrim("s").alias("r"),
              F.substring("t", 1, 4).alias("sub"), F.substring_index("p", ".", 2).alias("si"),
              F.substring_index("p", ".", -3).alias("sin"), F.lpad(F.trim("s"), 12, "#").alias("lp"),
              F.rpad(F.lit("hi"), 5, "ab").alias("rp"), F.repeat(F.lit("ab"), 3).alias("rep"),
              F.reverse(F.lit("abc")).alias("rev"), F.initcap(F.lit("hello wORLD")).alias("ic"),
              F.instr("p", "b").alias("instr"), F.locate("c", "p", 1).alias("loc"),
              F.translate(F.lit("translate"), "rnlt", "123").alias("tr"),
              F.regexp_extract(F.lit("100-200"), r"(\d+)-(\d+)", 1).alias("re"),
              F.regexp_replace(F.lit("100-200"), r"(\d+)", "--").alias("rr"),
              F.split("p", r"\.").alias("sp"), F.concat_ws("-", F.lit("x"), F.lit("y")).alias("cw"),
              F.format_number(F.lit(1234567.891), 2).alias("fn"), F.format_string("%d %s", F.lit(5), F.lit("x")).alias("fs"),
              F.levenshtein(F.lit("kitten"), F.lit("sitting")).alias("lev"), F.soundex(F.lit("Peters")).alias("sx"),
              F.ascii(F.lit("A")).alias("asc"), F.base64(F.lit("Spark")).alias("b64"),
              F.md5(F.lit("Spark")).alias("md5"), F.sha2(F.lit("Spark"), 256).alias("sha2"),
              F.crc32(F.lit("ABC")).alias("crc"), F.hex(F.lit(255)).alias("hex"), F.conv(F.lit("100"), 2, 10).alias("conv"),
              F.bin(F.lit(5)).alias("bin"), F.overlay(F.lit("SPARK_SQL"), F.lit("CORE"), 7).alias("ov"),
              F.left(F.lit("Spark"), 3).alias("left"), F.right(F.lit("Spark"), 3).alias("right"))
    assert row["trim"] == "Spark SQL" and row["l"] == "Spark SQL  " and row["r"] == "  Spark SQL"
    assert row["sub"] == "2015" and row["si"] == "a.b" and row["sin"] == "b.c.d"
    assert row["lp"] == "###Spark SQL" and row["rp"] == "hiaba" and row["rep"] == "ababab" and row["rev"] == "cba"
    assert row["ic"] == "Hello World" and row["instr"] == 3 and row["loc"] == 5 and row["tr"] == "1a2s3ae"
    assert row["re"] == "100" and row["rr"] == "-----" and row["sp"] == ["a", "b", "c", "d"] and row["cw"] == "x-y"
    assert row["fn"] == "1,234,567.89" and row["fs"] == "5 x" and row["lev"] == 3 and row["sx"] == "P362"
    assert row["asc"] == 65 and row["b64"] == "U3Bhcms=" and row["md5"] == "8cde774d6f7333752ed72cacddb05126"
    assert row["sha2"] == "529bc3b07127ecb7e53a4dcf1991d9152c24537d919178022b2c42657f79a26b"
    assert row["crc"] == 2743272264 and row["hex"] == "FF" and row["conv"] == "4" and row["bin"] == "101"
    assert row["ov"] == "SPARK_CORE" and row["left"] == "Spa" and row["right"] == "ark"
    matches = df.select(F.col("p").like("a.%").alias("like"), F.col("p").rlike(r"^a\.b").alias("rl"),
                        F.col("p").startswith("a.").alias("sw"), F.col("p").endswith("x").alias("ew"),
                        F.col("p").contains("b.c").alias("ct"), F.col("p").substr(3, 3).alias("ss")).collect()[0]
    assert tuple(matches) == (True, True, True, False, True, "b.c")


def test_math_and_null_functions(spark):
    df = spark.createDataFrame([(2.5, -3, None)], "a double, b int, c double")
    row = one(df, F.bround("a").alias("br"), F.round("a").alias("rd"), F.rint("a").alias("ri"), F.signum("b").alias("sg"),
              F.pmod("b", F.lit(5)).alias("pm"), F.hypot(F.lit(3), F.lit(4)).alias("hy"), F.cbrt(F.lit(27.0)).alias("cb"),
              F.degrees(F.lit(math.pi)).alias("dg"), F.atan2(F.lit(1.0), F.lit(1.0)).alias("at"),
              F.factorial(F.lit(5)).alias("fa"), F.shiftleft(F.lit(21), 1).alias("sl"),
              F.nvl("c", "a").alias("nvl"), F.nvl2("c", F.lit(1), F.lit(0)).alias("nvl2"),
              F.nullif("b", F.lit(-3)).alias("nif"), F.nanvl(F.lit(float("nan")), F.lit(1.0)).alias("nan"),
              F.isnan(F.lit(float("nan"))).alias("isnan"), F.col("c").eqNullSafe(F.lit(None)).alias("ens"),
              F.greatest("a", "b").alias("gr"), F.expm1(F.lit(0.0)).alias("em"), F.col("b").bitwiseAND(6).alias("band"))
    assert row["br"] == 2.0 and row["rd"] == 3.0 and row["ri"] == 2.0 and row["sg"] == -1.0 and row["pm"] == 2
    assert row["hy"] == 5.0 and abs(row["cb"] - 3.0) < 1e-12 and abs(row["dg"] - 180.0) < 1e-9
    assert abs(row["at"] - math.pi / 4) < 1e-12 and row["fa"] == 120 and row["sl"] == 42
    assert row["nvl"] == 2.5 and row["nvl2"] == 0 and row["nif"] is None and row["nan"] == 1.0
    assert row["isnan"] is True and row["ens"] is True and row["gr"] == 2.5 and row["em"] == 0.0 and row["band"] == 4


def test_date_functions(spark):
    df = spark.createDataFrame([("2015-04-08 13:08:15", "2015-02-28", 1428476400)], ["ts", "d", "unix"])
    row = one(df, F.to_date("ts").alias("td"), F.date_add("d", 2).alias("da"), F.date_sub("d", 1).alias("ds"),
              F.datediff(F.to_date("ts"), "d").alias("dd"), F.add_months("d", 1).alias("am"),
              F.months_between(F.lit("1997-02-28 10:30:00"), F.lit("1996-10-30")).alias("mb"),
              F.last_day("d").alias("ld"), F.next_day(F.lit("2015-07-27"), "Sun").alias("nd"),
              F.date_format("ts", "MM/dd/yyyy").alias("df"), F.date_trunc("hour", "ts").alias("dt"),
              F.trunc("d", "year").alias("tr"), F.from_unixtime("unix").alias("fu"),
              F.to_timestamp(F.lit("08/04/2015 01:02"), "dd/MM/yyyy HH:mm").alias("tsf"),
              F.to_date(F.lit("04-08-2015"), "MM-dd-yyyy").alias("tdf"),
              F.from_utc_timestamp("ts", "Asia/Tokyo").alias("jst"), F.to_utc_timestamp("ts", "Asia/Tokyo").alias("utc"),
              F.col("ts").cast("timestamp").cast("long").alias("secs"), F.make_date(F.lit(2020), F.lit(2), F.lit(29)).alias("md"),
              F.year("d").alias("y"), F.dayofweek("d").alias("dow"))
    assert row["td"] == dt.date(2015, 4, 8) and row["da"] == dt.date(2015, 3, 2) and row["ds"] == dt.date(2015, 2, 27)
    assert row["dd"] == 39 and row["am"] == dt.date(2015, 3, 28) and row["mb"] == 3.94959677
    assert row["ld"] == dt.date(2015, 2, 28) and row["nd"] == dt.date(2015, 8, 2) and row["df"] == "04/08/2015"
    assert row["dt"] == dt.datetime(2015, 4, 8, 13, 0) and row["tr"] == dt.date(2015, 1, 1)
    assert row["fu"] == "2015-04-08 07:00:00" and row["tsf"] == dt.datetime(2015, 4, 8, 1, 2)
    assert row["tdf"] == dt.date(2015, 4, 8) and row["jst"] == dt.datetime(2015, 4, 8, 22, 8, 15)
    assert row["utc"] == dt.datetime(2015, 4, 8, 4, 8, 15) and row["secs"] == 1428498495
    assert row["md"] == dt.date(2020, 2, 29) and row["y"] == 2015 and row["dow"] == 7
    win = df.select(F.window("ts", "10 minutes").alias("w")).collect()[0]["w"]
    assert win["start"] == dt.datetime(2015, 4, 8, 13, 0) and win["end"] == dt.datetime(2015, 4, 8, 13, 10)


def test_array_map_struct_json_functions(spark):
    df = spark.createDataFrame([([3, 1, None, 2, 1], ["a", "b"], '{"a": {"b": [1, 2]}, "c": "x"}')], ["xs", "ys", "js"])
    row = one(df, F.size("xs").alias("size"), F.array_contains("xs", 2).alias("ac"), F.array_distinct("xs").alias("ad"),
              F.array_join("ys", ",").alias("aj"), F.array_max("xs").alias("amax"), F.array_min("xs").alias("amin"),
              F.sort_array("xs", asc=False).alias("sa"), F.array_sort("xs").alias("as_"), F.element_at("xs", -1).alias("ea"),
              F.col("xs")[0].alias("gi"), F.slice("xs", 2, 2).alias("sl"), F.array_position("xs", 1).alias("ap"),
              F.array_remove("xs", 1).alias("ar"), F.array_union("xs", F.array(F.lit(7))).alias("au"),
              F.array_intersect("xs", F.array(F.lit(1), F.lit(2))).alias("ai"), F.array_except("xs", F.array(F.lit(1))).alias("ae"),
              F.array_repeat(F.lit("z"), 2).alias("rp"), F.sequence(F.lit(1), F.lit(7), F.lit(3)).alias("sq"),
              F.flatten(F.array("ys", "ys")).alias("fl"), F.arrays_zip("ys", "ys").alias("az"),
              F.get_json_object("js", "$.a.b[1]").alias("gj"), F.to_json(F.struct("ys", F.lit(1).alias("n"))).alias("tj"),
              F.from_json("js", "c string, a struct<b: array<int>>").alias("fj"),
              F.create_map(F.lit("k"), F.lit(1)).alias("m"), F.array_compact("xs").alias("cp"), F.array_append("ys", "c").alias("app"))
    assert row["size"] == 5 and row["ac"] is True and row["ad"] == [3, 1, None, 2] and row["aj"] == "a,b"
    assert row["amax"] == 3 and row["amin"] == 1 and row["sa"] == [3, 2, 1, 1, None] and row["as_"] == [1, 1, 2, 3, None]
    assert row["ea"] == 1 and row["gi"] == 3 and row["sl"] == [1, None] and row["ap"] == 2 and row["ar"] == [3, None, 2]
    assert row["au"] == [3, 1, None, 2, 7] and row["ai"] == [1, 2] and row["ae"] == [3, None, 2] and row["rp"] == ["z", "z"]
    assert row["sq"] == [1, 4, 7] and row["fl"] == ["a", "b", "a", "b"] and row["az"][1].asDict() == {"0": "b", "1": "b"}
    assert row["gj"] == "2" and row["tj"] == '{"ys":["a","b"],"n":1}' and row["fj"].asDict(True) == {"c": "x", "a": {"b": [1, 2]}}
    assert dict(row["m"]) == {"k": 1} and row["cp"] == [3, 1, 2, 1] and row["app"] == ["a", "b", "c"]
    maps = df.select(F.create_map(F.lit("k"), F.size("xs"), F.lit("j"), F.lit(0)).alias("m")) \
        .select(F.map_keys("m").alias("k"), F.map_values("m").alias("v"), F.col("m")["k"].alias("at"),
                F.element_at("m", "j").alias("ej"), F.map_filter("m", lambda k, v: v > 0).alias("mf"),
                F.transform_values("m", lambda k, v: v + 1).alias("tv")).collect()[0]
    assert maps["k"] == ["k", "j"] and maps["v"] == [5, 0] and maps["at"] == 5 and maps["ej"] == 0
    assert dict(maps["mf"]) == {"k": 5} and dict(maps["tv"]) == {"k": 6, "j": 1}
    st = df.select(F.struct(F.size("xs").alias("n"), F.col("ys")).alias("s")) \
        .select(F.col("s").getField("n").alias("n"), F.col("s")["ys"].alias("ys"), F.col("s").withField("z", F.lit(1)).alias("w"),
                F.col("s").dropFields("ys").alias("d"), "s.*").collect()[0]
    assert st[0] == 5 and st[1] == ["a", "b"] and st["w"].asDict() == {"n": 5, "ys": ["a", "b"], "z": 1} and st["d"].asDict() == {"n": 5} and st["w"].z == 1
    assert st[4] == 5 and st[5] == ["a", "b"]


def test_higher_order_functions(spark):
    df = spark.createDataFrame([(1, [1, 2, 3, 4], [10, 20]), (2, [], [1]), (3, None, [1])], "k int, xs array<int>, ys array<int>")
    rows = df.select(F.transform("xs", lambda x: x * 2).alias("t"), F.transform("xs", lambda x, i: x + i).alias("ti"),
                     F.filter("xs", lambda x: x % 2 == 0).alias("f"), F.exists("xs", lambda x: x > 3).alias("e"),
                     F.forall("xs", lambda x: x > 0).alias("fa"), F.zip_with("xs", "ys", lambda a, b: F.coalesce(a, F.lit(0)) + F.coalesce(b, F.lit(0))).alias("z"),
                     F.aggregate("xs", F.lit(0), lambda acc, x: acc + x).alias("agg"),
                     F.aggregate("xs", F.lit(0), lambda acc, x: acc + x, lambda acc: acc * 10).alias("aggf"),
                     F.transform("xs", lambda x: x + F.col("k")).alias("outer")).collect()
    assert rows[0]["t"] == [2, 4, 6, 8] and rows[0]["ti"] == [1, 3, 5, 7] and rows[0]["f"] == [2, 4]
    assert rows[0]["e"] is True and rows[0]["fa"] is True and rows[0]["z"] == [11, 22, 3, 4]
    assert rows[0]["agg"] == 10 and rows[0]["aggf"] == 100 and rows[0]["outer"] == [2, 3, 4, 5]
    assert rows[1]["t"] == [] and rows[1]["e"] is False and rows[1]["fa"] is True and rows[1]["agg"] == 0
    assert rows[2]["t"] is None and rows[2]["e"] is None and rows[2]["agg"] is None


def test_generators(spark):
    df = spark.createDataFrame([(1, [10, 20], {"a": 1}), (2, [], None), (3, None, {"b": 2, "c": 3})],
                               "id int, xs array<int>, m map<string,int>")
    assert [tuple(r) for r in df.select("id", F.explode("xs").alias("x")).collect()] == [(1, 10), (1, 20)]
    assert [tuple(r) for r in df.select("id", F.explode_outer("xs")).collect()] == [(1, 10), (1, 20), (2, None), (3, None)]
    assert [tuple(r) for r in df.select(F.posexplode("xs").alias("p", "x"), (F.col("id") * 2).alias("d")).collect()] == \
        [(0, 10, 2), (1, 20, 2)]
    assert [tuple(r) for r in df.select("id", F.explode("m")).collect()] == [(1, "a", 1), (3, "b", 2), (3, "c", 3)]
    exploded = df.withColumn("x", F.explode("xs"))
    assert exploded.columns == ["id", "xs", "m", "x"] and [r["x"] for r in exploded.collect()] == [10, 20]


def test_gathered_aggregates(people):
    rows = {r["k"]: r for r in people.groupBy("k").agg(
        F.collect_list("v").alias("cl"), F.collect_set("v").alias("cs"), F.corr("v", "x").alias("corr"),
        F.percentile_approx("v", 0.5).alias("p50"), F.percentile_approx("v", [0.0, 1.0]).alias("pr"),
        F.max_by("x", "v").alias("maxby"), F.min_by("x", "v").alias("minby"), (F.sum("v") / F.count("v")).alias("ratio"),
        F.round(F.avg("x"), 1).alias("ravg"), F.countDistinct("v", "x").alias("cd2"), F.skewness("v").alias("skew"),
        F.kurtosis("v").alias("kurt"), F.last("x", True).alias("last"), F.median("v").alias("med"), F.mode("v").alias("mode"),
        F.stddev_pop("v").alias("sp"), F.var_pop("v").alias("vp"), F.sum_distinct("v").alias("sd"), F.product("v").alias("prod"),
        F.bool_and(F.col("v") > 1).alias("ba"), F.bool_or(F.col("v") > 4).alias("bo"), F.first("v").alias("first"),
        F.covar_samp("v", "x").alias("cov"), F.approx_count_distinct("v").alias("acd"), F.count_if(F.col("v") > 1).alias("ci"),
    ).collect()}
    a, b, n = rows["a"], rows["b"], rows[None]
    assert a["cl"] == [1, 2, 2] and sorted(a["cs"]) == [1, 2] and abs(a["corr"] - 1.0) < 1e-12 and a["p50"] == 2
    assert a["pr"] == [1, 2] and a["maxby"] == 20.0 and a["minby"] == 10.0 and abs(a["ratio"] - 5 / 3) < 1e-12
    assert a["ravg"] == 15.0 and a["cd2"] == 2 and abs(a["skew"] + 0.7071067811865) < 1e-9 and abs(a["kurt"] + 1.5) < 1e-9
    assert a["last"] == 20.0 and a["med"] == 2.0 and a["mode"] == 2 and abs(a["vp"] - 2 / 9) < 1e-12
    assert abs(a["sp"] - math.sqrt(2 / 9)) < 1e-12 and a["sd"] == 3 and a["prod"] == 4.0 and a["ba"] is False
    assert a["bo"] is False and a["first"] == 1 and abs(a["cov"] - 5.0) < 1e-12 and a["acd"] == 2 and a["ci"] == 2
    assert b["cl"] == [5, 3] and abs(b["corr"] + 1.0) < 1e-12 and b["bo"] is True and b["med"] == 4.0
    assert n["cl"] == [9] and n["corr"] is None and n["skew"] is None
    total = people.agg(F.collect_list("v").alias("cl"), F.countDistinct("k").alias("dk"), F.count("*").alias("n")).collect()[0]
    assert sorted(total["cl"]) == [1, 2, 2, 3, 5, 9] and total["dk"] == 2 and total["n"] == 6
    empty = people.filter("v > 100").agg(F.collect_list("v"), F.count("*"), F.sum("v")).collect()[0]
    assert tuple(empty) == ([], 0, None)


def test_pivot_rollup_cube(people):
    pv = {r["k"]: r for r in people.groupBy("k").pivot("v").agg(F.sum("x")).collect()}
    assert pv["a"]["1"] == 10.0 and pv["a"]["2"] == 20.0 and pv["b"]["5"] == 1.0 and pv[None]["9"] == 9.0 and pv["a"]["3"] is None
    two = people.groupBy("k").pivot("v", [1, 2]).agg(F.sum("x"), F.count("x"))
    assert two.columns == ["k", "1_sum(x)", "1_count(x)", "2_sum(x)", "2_count(x)"]
    assert people.groupBy().pivot("k").agg(F.sum("v")).collect()[0].asDict() == {"a": 5, "b": 8}
    roll = {(r["k"], r["v"]): r["s"] for r in people.filter("k is not null").rollup("k", "v").agg(F.sum("x").alias("s")).collect()}
    assert roll[("a", None)] == 30.0 and roll[("b", None)] == 3.0 and roll[(None, None)] == 33.0 and roll[("a", 1)] == 10.0
    assert people.cube("k", "v").count().count() == 14


def test_window_functions(people):
    w = Window.partitionBy("k").orderBy("v")
    rows = people.select(
        "k", "v", "x", F.row_number().over(w).alias("rn"), F.rank().over(w).alias("rk"), F.dense_rank().over(w).alias("dr"),
        F.lag("v").over(w).alias("lag"), F.lead("v", 1, -1).over(w).alias("lead"), F.sum("v").over(w).alias("run"),
        F.sum("x").over(w.rowsBetween(Window.unboundedPreceding, Window.currentRow)).alias("runx"),
        F.avg("x").over(Window.partitionBy("k")).alias("avgx"), F.collect_list("v").over(w).alias("cl"),
        (F.col("v") / F.sum("v").over(Window.partitionBy("k"))).alias("share"), F.ntile(2).over(w).alias("nt"),
        F.percent_rank().over(w).alias("pr"), F.cume_dist().over(w).alias("cd"), F.max("v").over(w.rowsBetween(-1, 1)).alias("mx"),
        F.first("v").over(w).alias("fv"), F.last("v").over(w.rowsBetween(Window.unboundedPreceding, Window.unboundedFollowing)).alias("lv"),
        F.sum("v").over(w.rangeBetween(-1, 0)).alias("rng"), F.count("*").over(w).alias("cnt"),
    ).collect()
    a = [r for r in rows if r["k"] == "a"]
    assert [r["rn"] for r in a] == [1, 2, 3] and [r["rk"] for r in a] == [1, 2, 2] and [r["dr"] for r in a] == [1, 2, 2]
    assert [r["lag"] for r in a] == [None, 1, 2] and [r["lead"] for r in a] == [2, 2, -1]
    assert [r["run"] for r in a] == [1, 5, 5]                   # RANGE frame: ties are peers
    assert [r["avgx"] for r in a] == [15.0] * 3 and [r["cl"] for r in a] == [[1], [1, 2, 2], [1, 2, 2]]
    assert [r["share"] for r in a] == [0.2, 0.4, 0.4] and [r["nt"] for r in a] == [1, 1, 2]
    assert [r["pr"] for r in a] == [0.0, 0.5, 0.5] and [round(r["cd"], 4) for r in a] == [0.3333, 1.0, 1.0]
    assert [r["mx"] for r in a] == [2, 2, 2] and [r["fv"] for r in a] == [1, 1, 1] and [r["lv"] for r in a] == [2, 2, 2]
    assert [r["rng"] for r in a] == [1, 5, 5] and [r["cnt"] for r in a] == [1, 3, 3]
    b = [r for r in rows if r["k"] == "b"]
    assert [r["v"] for r in b] == [3, 5] and [r["runx"] for r in b] == [2.0, 3.0] and [r["rng"] for r in b] == [3, 5]
    ranked = people.withColumn("r", F.row_number().over(Window.orderBy(F.desc("v")))).filter("r <= 2").collect()
    assert [r["v"] for r in ranked] == [9, 5]
    scaled = people.withColumn("z", F.col("v") - F.min("v").over(Window.partitionBy("k")))
    assert sorted(r["z"] for r in scaled.collect()) == [0, 0, 0, 1, 1, 2] and "__win_0" not in scaled.columns


def test_ordering_null_placement(people):
    assert [r["x"] for r in people.orderBy("x").collect()][:2] == [None, 1.0]                 # asc: nulls first
    assert [r["x"] for r in people.orderBy(F.desc("x")).collect()][-1] is None                # desc: nulls last
    assert [r["x"] for r in people.orderBy(F.col("x").asc_nulls_last()).collect()][-1] is None
    assert [r["x"] for r in people.orderBy(F.col("x").desc_nulls_first()).collect()][0] is None
    assert [r["v"] for r in people.orderBy(-F.col("v")).collect()] == [9, 5, 3, 2, 2, 1]      # expression keys


def test_dataframe_helpers(spark, people):
    other = spark.createDataFrame([("a", 1, 10.0), ("b", 3, 2.0), ("b", 3, 2.0)], ["k", "v", "x"])
    assert people.subtract(other).count() == 4 and people.exceptAll(other).count() == 4
    assert people.intersect(other).count() == 2 and other.intersectAll(other).count() == 3
    melted = people.melt(["k"], ["v", "x"], "var", "val")
    assert melted.count() == 12 and melted.columns == ["k", "var", "val"]
    summary = {r["summary"]: r for r in people.summary().collect()}
    assert summary["count"]["v"] == "6" and summary["50%"]["v"] == "2" and summary["max"]["x"] == "20.0"
    assert people.approxQuantile("v", [0.5, 1.0], 0) == [2.0, 9.0]
    assert abs(people.corr("v", "x") + 0.2692024941328392) < 1e-12 and abs(people.cov("v", "x") + 6.5) < 1e-12
    ct = {r["k_v"]: r for r in people.crosstab("k", "v").collect()}
    assert ct["a"]["2"] == 2 and ct["b"]["2"] == 0
    replaced = people.replace({"a": "z"}).replace(2, 22, "v").collect()
    assert replaced[1]["k"] == "z" and replaced[1]["v"] == 22
    assert people.toJSON().collect()[0] == '{"k":"a","v":1,"x":10.0}' and people.tail(1)[0]["v"] == 9
    odd = {r["odd"]: r["n"] for r in people.groupBy((F.col("v") % 2).alias("odd")).agg(F.count("*").alias("n")).collect()}
    assert odd == {1: 4, 0: 2}
    assert people.transform(lambda d: d.filter("v > 2")).count() == 3
    assert people.sampleBy("k", {"a": 1.0}, seed=1).count() == 3 and people.offset(4).count() == 2
    assert people.stat.freqItems(["k"], 0.3).collect()[0][0] == ["a", "b"]
    assert people.select(F.col("k").isNotNull().cast("int").alias("i")).agg(F.sum("i")).collect()[0][0] == 5


def test_sql_uses_the_same_functions(spark, people):
    people.createOrReplaceTempView("people_fn")
    row = spark.sql("SELECT k, collect_list(v) AS vs, percentile_approx(v, 0.5) AS p, max_by(x, v) AS mb, "
                    "concat_ws('-', k, 'z') AS c FROM people_fn WHERE k = 'a' GROUP BY k").collect()[0]
    assert row["vs"] == [1, 2, 2] and row["p"] == 2 and row["mb"] == 20.0 and row["c"] == "a-z"
    assert spark.sql("SELECT date_add('2020-02-27', 3) AS d, substring('Spark SQL', 5, 1) AS s, "
                     "array_contains(array(1, 2), 2) AS a").collect()[0].asDict() == \
        {"d": dt.date(2020, 3, 1), "s": "k", "a": True}


@pytest.fixture(scope="module")
def sales(spark):
    df = spark.createDataFrame(
        [("a", 1, 10.0, "2020-01-05"), ("a", 2, 20.0, "2020-02-01"), ("a", 2, None, "2021-01-01"),
         ("b", 5, 1.0, "2020-01-09"), ("b", 3, 2.0, "2020-03-01"), (None, 9, 9.0, None)], ["k", "v", "x", "d"])
    df.createOrReplaceTempView("sales")
    return df


def q(spark, text):
    return [tuple(r) for r in spark.sql(text).collect()]


def test_sql_grouping_having_order(spark, sales):
    assert q(spark, "SELECT k, count(*) AS n FROM sales GROUP BY k HAVING count(*) > 1 ORDER BY sum(v) DESC") == \
        [("b", 2), ("a", 3)]
    assert q(spark, "SELECT year(d) AS y, sum(v) FROM sales GROUP BY year(d) ORDER BY y") == \
        [(None, 9), (2020, 11), (2021, 2)]
    assert q(spark, "SELECT k, sum(v) / count(*) AS m, round(avg(x), 1) FROM sales GROUP BY 1 ORDER BY 1 NULLS LAST") == \
        [("a", 5 / 3, 15.0), ("b", 4.0, 1.5), (None, 9.0, 9.0)]
    assert q(spark, "SELECT k, first(x, true), last(v) FROM sales WHERE k IS NOT NULL GROUP BY k ORDER BY k") == \
        [("a", 10.0, 2), ("b", 1.0, 3)]
    assert q(spark, "SELECT k, v FROM sales ORDER BY 2 DESC NULLS LAST LIMIT 2") == [(None, 9), ("b", 5)]


def test_sql_predicates_and_operators(spark, sales):
    assert q(spark, "SELECT v FROM sales WHERE k IN ('a') AND v BETWEEN 1 AND 2 AND d LIKE '2020%' AND x IS NOT NULL") == \
        [(1,), (2,)]
    assert q(spark, "SELECT cast(v AS string) || '_' || k AS s FROM sales WHERE v NOT IN (2, 9) AND d NOT LIKE '%-03-%'") == \
        [("1_a",), ("5_b",)]
    assert q(spark, "SELECT split(d, '-')[0] AS y, DATE '2020-01-01' AS dd FROM sales LIMIT 1") == \
        [("2020", dt.date(2020, 1, 1))]
    assert q(spark, "SELECT count(*) FROM sales WHERE k <=> NULL OR d RLIKE '^2021'") == [(2,)]
    assert q(spark, "SELECT CASE WHEN v > 2 THEN 'hi' ELSE 'lo' END AS c, count(*) FROM sales GROUP BY 1 ORDER BY 1") == \
        [("hi", 3), ("lo", 3)]


def test_sql_subqueries_ctes_set_operations(spark, sales):
    assert q(spark, "SELECT k, v FROM (SELECT k, v FROM sales WHERE v > 1) sub WHERE k = 'a'") == [("a", 2), ("a", 2)]
    assert q(spark, "WITH big AS (SELECT * FROM sales WHERE v > 2), n AS (SELECT count(*) AS c FROM big) SELECT c FROM n") == [(3,)]
    assert q(spark, "SELECT k, v FROM sales WHERE v = (SELECT max(v) FROM sales)") == [(None, 9)]
    assert q(spark, "SELECT k, v FROM sales WHERE k IN (SELECT k FROM sales WHERE v > 4)") == [("b", 5), ("b", 3)]
    assert q(spark, "SELECT k FROM sales UNION SELECT k FROM sales ORDER BY k") == [(None,), ("a",), ("b",)]
    assert len(q(spark, "SELECT k FROM sales UNION ALL SELECT k FROM sales")) == 12
    assert q(spark, "SELECT k FROM sales EXCEPT SELECT k FROM sales WHERE v > 2") == [("a",)]
    assert q(spark, "SELECT k FROM sales INTERSECT SELECT k FROM sales WHERE v = 5") == [("b",)]
    assert q(spark, "SELECT EXISTS (SELECT 1 FROM sales WHERE v > 100) AS e") == [(False,)]


def test_sql_joins_with_aliases(spark, sales):
    assert sorted(q(spark, "SELECT a.k, b.v FROM sales a JOIN sales b ON a.k = b.k WHERE a.v < b.v")) == \
        [("a", 2), ("a", 2), ("b", 5)]
    assert q(spark, "SELECT s.k, u.n FROM sales s LEFT JOIN (SELECT k, count(*) AS n FROM sales GROUP BY k) u "
                    "ON s.k = u.k WHERE v = 1") == [("a", 3)]
    assert q(spark, "SELECT count(*) FROM sales a, sales b") == [(36,)]


def test_sql_window_clauses(spark, sales):
    rows = q(spark, "SELECT k, v, row_number() OVER (PARTITION BY k ORDER BY v DESC) AS rn, "
                    "sum(v) OVER (PARTITION BY k) AS tot, lag(v, 1) OVER (PARTITION BY k ORDER BY v DESC) AS p, "
                    "sum(v) OVER (PARTITION BY k ORDER BY v DESC ROWS BETWEEN UNBOUNDED PRECEDING AND CURRENT ROW) AS r "
                    "FROM sales WHERE k = 'a'")
    assert rows == [("a", 2, 1, 5, None, 2), ("a", 2, 2, 5, 2, 4), ("a", 1, 3, 5, 2, 5)]
    assert q(spark, "SELECT k, v FROM (SELECT k, v, rank() OVER (PARTITION BY k ORDER BY v DESC) AS r FROM sales) "
                    "WHERE r = 1 AND k = 'b'") == [("b", 5)]


def test_theta_joins_pandas_udfs_schema_json(spark, people):
    from raydp_b200.sql import types as T
    other = spark.createDataFrame([("a", "A"), ("c", "C")], ["kk", "name"])
    assert [tuple(r) for r in people.join(other, F.col("k") > F.col("kk")).collect()] == \
        [("b", 5, 1.0, "a", "A"), ("b", 3, 2.0, "a", "A")]
    left = people.join(other, (F.col("k") > F.col("kk")) & (F.col("v") > 3), "left")
    assert left.count() == 6 and left.filter("name is not null").count() == 1
    assert people.join(other, F.col("k") > F.col("kk"), "left_anti").count() == 4
    assert people.join(other, F.col("k") > F.col("kk"), "left_semi").count() == 2
    up = people.select(F.pandas_udf(lambda s: s.str.upper(), "string")("k").alias("u")).collect()
    assert [r.u for r in up] == ["A", "A", "A", "B", "B", None]

    @F.pandas_udf("double")
    def times(a, b):
        return a * b.fillna(0)
    assert [r.p for r in people.select(times("v", "x").alias("p")).collect()] == [10.0, 40.0, 0.0, 5.0, 6.0, 81.0]
    schema = people.schema
    assert T.StructType.fromJson(schema.jsonValue()) == schema and '"name":"k"' in schema.json()
    assert T.StructType().add("a", T.IntegerType()).add("b", "array<string>").simpleString() == \
        "struct<a:int,b:array<string>>"


def test_sql_interval_arithmetic(spark):
    spark.createDataFrame([("2020-01-31", "2020-01-31 10:00:00")], ["d", "t"]).createOrReplaceTempView("iv")
    row = spark.sql("select to_date(d) + interval 1 month as a, to_date(d) - interval '3' day as b, "
                    "to_timestamp(t) + interval 2 hours as c, to_timestamp(t) + interval '1 day 30 minutes' as e, "
                    "date '2020-03-01' - interval 1 year 1 month as f, to_date(d) + interval 1 hour as g from iv").collect()[0]
    assert row.a == dt.date(2020, 2, 29) and row.b == dt.date(2020, 1, 28) and row.c == dt.datetime(2020, 1, 31, 12, 0)
    assert row.e == dt.datetime(2020, 2, 1, 10, 30) and row.f == dt.date(2019, 2, 1)
    assert row.g == dt.datetime(2020, 1, 31, 1, 0)                       # a time part turns the date into a timestamp
    assert spark.sql("select count(*) from iv where to_date(d) >= date '2020-02-01' - interval 7 days").collect()[0][0] == 1


def test_time_windows_tumbling_and_sliding(spark):
    df = spark.createDataFrame([("2020-01-01 10:07:00", 1), ("2020-01-01 10:12:00", 2), (None, 3)], ["t", "v"])
    slid = df.groupBy(F.window("t", "10 minutes", "5 minutes")).agg(F.sum("v").alias("s")).orderBy("window").collect()
    assert [(r.window.start.minute, r.window.end.minute, r.s) for r in slid] == [(0, 10, 1), (5, 15, 3), (10, 20, 2)]
    tumbling = df.groupBy(F.window("t", "10 minutes")).agg(F.sum("v").alias("s")).orderBy("window").collect()
    assert [(r.window.start.minute, r.s) for r in tumbling] == [(0, 1), (10, 2)]          # the null timestamp drops out
    assert df.select("v", F.window("t", "10 minutes", "5 minutes")).count() == 4
    by_struct = df.select(F.struct("v", (F.col("v") % 2).alias("odd")).alias("s")).groupBy("s").count().collect()
    assert sorted((r.s.v, r.s.odd, r["count"]) for r in by_struct) == [(1, 1, 1), (2, 0, 1), (3, 1, 1)]


def test_second_batch_of_functions_and_generators(spark):
    df = spark.createDataFrame([("Ab1-x", 5, "a:1,b:2", "[1,2,3]", "2020-03-04 05:06:07")], ["s", "v", "m", "j", "t"])
    r = df.select(
        F.mask("s").alias("mk"), F.bit_count("v").alias("bc"), F.getbit("v", 2).alias("gb"), F.str_to_map("m").alias("sm"),
        F.json_array_length("j").alias("jl"), F.regexp_extract_all(F.lit("a1b22c333"), r"(\d+)", 1).alias("ra"),
        F.regexp_instr("s", r"\d").alias("ri"), F.replace("s", "b", "Z").alias("rp"), F.position(F.lit("1"), "s").alias("ps"),
        F.elt(F.lit(2), F.lit("x"), F.lit("y")).alias("el"), F.find_in_set(F.lit("b"), F.lit("a,b,c")).alias("fs"),
        F.width_bucket("v", 0, 10, 5).alias("wb"), F.weekday("t").alias("wd"), F.date_part("hour", "t").alias("dp"),
        F.unix_seconds(F.to_timestamp("t")).alias("us"), F.timestamp_millis(F.lit(1000)).alias("tm"),
        F.make_timestamp(F.lit(2020), F.lit(1), F.lit(2), F.lit(3), F.lit(4), F.lit(5.5)).alias("mt"),
        F.try_divide("v", F.lit(0)).alias("td"), F.try_add("v", F.lit(1)).alias("ta"), F.ceiling(F.lit(1.2)).alias("ce"),
        F.floor(F.lit(-1.2)).alias("fl"), F.ln(F.e()).alias("ln"), F.convert_timezone("UTC", "Asia/Tokyo", "t").alias("cz"),
        F.to_number(F.lit("$1,234.5")).alias("tn"), F.char(F.lit(65)).alias("ch")).first()
    assert (r.mk, r.bc, r.gb, r.sm, r.jl, r.ra) == ("Xxn-x", 2, 1, {"a": "1", "b": "2"}, 3, ["1", "22", "333"])
    assert (r.ri, r.rp, r.ps, r.el, r.fs, r.wb, r.wd, r.dp) == (3, "AZ1-x", 3, "y", 2, 3, 2, 5)
    assert r.us == 1583298367 and r.tm == dt.datetime(1970, 1, 1, 0, 0, 1) and r.mt == dt.datetime(2020, 1, 2, 3, 4, 5, 500000)
    assert r.td is None and r.ta == 6 and r.ce == 2 and isinstance(r.ce, int) and r.fl == -2 and r.ln == 1.0
    assert r.cz == dt.datetime(2020, 3, 4, 14, 6, 7) and r.tn == 1234.5 and r.ch == "A"
    stacked = df.select("v", F.stack(2, F.lit("a"), F.lit(1), F.lit("b"), F.lit(2))).collect()
    assert [tuple(x) for x in stacked] == [(5, "a", 1), (5, "b", 2)]
    assert [tuple(x) for x in df.select(F.inline(F.array(F.struct(F.lit(1).alias("x"), F.lit("p").alias("y"))))).collect()] == \
        [(1, "p")]
    assert tuple(spark.sql("select extract('year', to_timestamp('2020-03-04 05:06:07')) y, ceiling(1.5) c, power(2, 3) p, "
                           "ucase('a') u").collect()[0]) == (2020, 2, 8.0, "A")


def test_correlated_subquery_is_named_as_unsupported(spark):
    from raydp_b200.sql.sqlparser import SqlError
    spark.range(5).createOrReplaceTempView("corr_a")
    spark.range(3).withColumnRenamed("id", "k").createOrReplaceTempView("corr_b")
    with pytest.raises(SqlError, match="correlated sub-queries"):
        spark.sql("select * from corr_a where exists (select 1 from corr_b where corr_b.k = corr_a.id)")
    assert spark.sql("select corr_a.id from corr_a left semi join corr_b on corr_a.id = corr_b.k").count() == 3
    assert spark.sql("select * from corr_a where id not in (select k from corr_b)").count() == 2
