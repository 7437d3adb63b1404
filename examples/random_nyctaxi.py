"""Write a synthetic NYC-taxi-fare CSV with the Kaggle column layout.

Counterpart of the reference's ``examples/random_nyctaxi.py`` (same columns and value ranges:
``key, fare_amount, pickup_datetime, pickup_/dropoff_ longitude/latitude, passenger_count``;
timestamps formatted ``YYYY-mm-dd HH:MM:SS UTC``).  The fare is *not* pure noise here: it is a
noisy function of the trip's Manhattan distance, hour and passenger count, so the examples can
show a loss that actually goes down.

    python examples/random_nyctaxi.py --num-records 20000 --output examples/fake_nyctaxi.csv
"""
import argparse
import os

import numpy as np
import pyarrow as pa
import pyarrow.csv as pacsv

COLUMNS = ["key", "fare_amount", "pickup_datetime", "pickup_longitude", "pickup_latitude",
           "dropoff_longitude", "dropoff_latitude", "passenger_count"]
DEFAULT_PATH = os.path.join(os.path.dirname(os.path.realpath(__file__)), "fake_nyctaxi.csv")


def make_table(n: int, seed: int = 0) -> pa.Table:
    rng = np.random.default_rng(seed)
    lon0, lon1 = rng.uniform(-74.2, -73.8, size=(2, n))
    lat0, lat1 = rng.uniform(40.7, 40.8, size=(2, n))
    riders = rng.integers(1, 5, size=n)
    # five years of seconds after 2010-01-01, as in the reference generator
    when = np.datetime64("2010-01-01T00:00:00") + rng.integers(0, 157_680_000, size=n).astype("timedelta64[s]")
    hour = (when.astype("datetime64[h]") - when.astype("datetime64[D]")).astype(int)
    distance = np.abs(lon1 - lon0) + np.abs(lat1 - lat0)
    rush = ((hour >= 16) & (hour <= 20)).astype(float)
    fare = 3.0 + 55.0 * distance + 2.5 * rush + 0.4 * riders + rng.normal(0.0, 1.0, size=n)
    fare = np.clip(fare, 3.0, 60.0)
    stamps = [str(t).replace("T", " ") + " UTC" for t in when]
    return pa.table({"key": ["fake_key"] * n, "fare_amount": fare, "pickup_datetime": stamps,
                     "pickup_longitude": lon0, "pickup_latitude": lat0,
                     "dropoff_longitude": lon1, "dropoff_latitude": lat1,
                     "passenger_count": riders}).select(COLUMNS)


def ensure_csv(path: str = DEFAULT_PATH, n: int = 2000, seed: int = 0) -> str:
    """Create ``path`` if it does not exist yet and return it."""
    path = os.environ.get("NYC_TAXI_CSV", path)       # point the examples at a real dataset
    if not os.path.exists(path):
        pacsv.write_csv(make_table(n, seed), path)
    return path


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="synthetic NYC taxi fare generator")
    parser.add_argument("--num-records", type=int, default=2000, metavar="N")
    parser.add_argument("--seed", type=int, default=0)
    parser.add_argument("--output", default=DEFAULT_PATH)
    args = parser.parse_args()
    pacsv.write_csv(make_table(args.num_records, args.seed), args.output)
    print(f"wrote {args.num_records} rows to {args.output}")
