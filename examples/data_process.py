"""Feature engineering for the NYC-taxi examples, written against the PySpark-style DataFrame API.

Produces the same 20 features + label as the reference's ``examples/data_process.py`` (9 time
features ``:28-49``, 3 coordinate-difference features ``:63-68``, 8 landmark distances ``:69-84``;
raw columns dropped ``:87-95``), so every example model sees ``[B, 20]`` inputs.

Two of the time flags are written as Python UDFs on purpose (the reference does the same): the
executors run them row-wise, which exercises ``functions.udf``.  Everything else is a column
expression that the executors evaluate vectorised with ``pyarrow.compute``.
"""
import os

import _bootstrap  # noqa: F401  (source-checkout path setup)
from raydp.sql.functions import (abs as col_abs, col, dayofmonth, dayofweek, hour, lit, month, quarter,
                                 udf, weekofyear, year)

from random_nyctaxi import DEFAULT_PATH, ensure_csv

NYC_TRAIN_CSV = "file://" + DEFAULT_PATH
LABEL = "fare_amount"

# (longitude, latitude) of the places the distance features are measured against
LANDMARKS = {
    "jfk": (-73.7822222222, 40.6441666667),
    "ewr": (-74.175, 40.69),
    "lgr": (-73.87, 40.77),
    "downtown": (-74.0063889, 40.7141667),
}
RAW_COLUMNS = ["pickup_datetime", "pickup_longitude", "pickup_latitude", "dropoff_longitude",
               "dropoff_latitude", "passenger_count", "key"]
BOUNDS = {"pickup_longitude": (-76, -72), "dropoff_longitude": (-76, -72),
          "pickup_latitude": (38, 42), "dropoff_latitude": (38, 42), "passenger_count": (1, 6)}


def csv_path() -> str:
    """The example CSV, generated on first use (no dataset ships with the repository)."""
    return "file://" + ensure_csv()


def clean_up(df):
    for name, (lo, hi) in BOUNDS.items():
        df = df.filter((col(name) >= lo) & (col(name) <= hi))
    df = df.filter((col(LABEL) > 0) & (col(LABEL) < 250))
    return df.filter((col("dropoff_longitude") != col("pickup_longitude"))
                     & (col("dropoff_latitude") != col("pickup_latitude")))


@udf("int")
def _evening_rush(hour_of_day, weekday):
    return 1 if 16 <= hour_of_day <= 20 and weekday < 5 else 0


@udf("int")
def _late(hour_of_day):
    return 1 if hour_of_day <= 6 or hour_of_day >= 20 else 0


def add_time_features(df):
    ts = col("pickup_datetime")
    parts = {"day": dayofmonth(ts), "hour_of_day": hour(ts), "day_of_week": dayofweek(ts) - 2,
             "week_of_year": weekofyear(ts), "month_of_year": month(ts), "quarter_of_year": quarter(ts),
             "year": year(ts)}
    for name, expr in parts.items():
        df = df.withColumn(name, expr)
    df = df.withColumn("night", _evening_rush("hour_of_day", "day_of_week"))
    return df.withColumn("late_night", _late("hour_of_day"))


def add_distance_features(df):
    df = df.withColumn("abs_diff_longitude", col_abs(col("dropoff_longitude") - col("pickup_longitude")))
    df = df.withColumn("abs_diff_latitude", col_abs(col("dropoff_latitude") - col("pickup_latitude")))
    df = df.withColumn("manhattan", col("abs_diff_latitude") + col("abs_diff_longitude"))
    for place, (lon, lat) in LANDMARKS.items():
        for end in ("pickup", "dropoff"):
            # NB the reference passes (longitude, latitude) into a udf whose parameters are named
            # (lat, lon); the value is the same L1 distance either way.
            dist = col_abs(col(f"{end}_longitude") - lit(lon)) + col_abs(col(f"{end}_latitude") - lit(lat))
            df = df.withColumn(f"{end}_distance_{place}", dist.cast("float"))
    return df


def nyc_taxi_preprocess(df):
    df = add_distance_features(add_time_features(clean_up(df)))
    return df.drop(*RAW_COLUMNS)


def feature_names(df):
    return [f.name for f in df.schema if f.name != LABEL]


if __name__ == "__main__":
    import raydp
    from raydp import runtime as ray

    ray.init(address="local", num_cpus=6)
    spark = raydp.init_spark("NYCTAXI data processing", num_executors=1, executor_cores=1,
                             executor_memory="500M")
    data = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path())
    spark.conf.set("spark.sql.session.timeZone", "UTC")
    data = nyc_taxi_preprocess(data)
    data.show(5)
    ds = raydp.data.from_spark(data)
    print(ds.schema())
    print("rows:", ds.count(), "features:", len(feature_names(data)), "cwd:", os.getcwd())
    raydp.stop_spark()
    ray.shutdown()
