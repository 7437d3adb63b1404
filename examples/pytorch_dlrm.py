"""DLRM on a Criteo-shaped frame: Spark preprocessing -> ``TorchEstimator`` (``BASELINE.json`` config 3).

Script form of the reference's ``examples/pytorch_dlrm.ipynb``:

* cells 5-7  -- categorical columns are turned into dense ids on the executors.  The notebook
  builds frequency-ranked dictionaries with ``groupBy/count`` + ``row_number`` windows and joins
  them back; here the synthetic categories are hashed into a fixed vocabulary with
  ``functions.hash`` (the same thing Criteo pipelines do for the long tail), missing values are
  filled and the 13 integer features get the usual ``log(x + 1)`` transform;
* cell 9     -- the model is ``raydp.models.DLRM`` (bottom MLP 13-512-128-32, 26 tables of width 32,
  dot interaction, top MLP 1024-1024-512-256-1);
* cell 14    -- ``TorchEstimator`` with 13 float + 26 long feature columns, ``BCEWithLogitsLoss``, SGD.

    python examples/pytorch_dlrm.py --rows 20000                   # CPU workers, stock torch
    python examples/pytorch_dlrm.py --rows 2000000 --gpu --workers 2 --batch-size 8192

With ``--gpu`` the estimator runs the hybrid-parallel engine: dense layers data-parallel, embedding
tables sharded table-wise, lookups fused with the NVLink all-to-all.
"""
import argparse
import json
import time

import _bootstrap  # noqa: F401
import numpy as np
import pyarrow as pa
import torch
import torch.nn as nn

import raydp
from raydp import runtime as ray
from raydp.models import DLRM
from raydp.sql.functions import abs as col_abs, col, hash as col_hash, log
from raydp.torch import TorchEstimator
from raydp.utils import random_split

DENSE = [f"_c{i}" for i in range(1, 14)]        # Criteo naming: _c0 label, _c1.._c13 ints, _c14.._c39 cats
SPARSE = [f"_c{i}" for i in range(14, 40)]
LABEL = "_c0"


def synthetic_criteo(spark, rows: int, partitions: int, seed: int = 0):
    """Raw frame with the Criteo schema: int label, 13 nullable ints, 26 hex-string categories."""
    rng = np.random.default_rng(seed)
    cols = {}
    signal = np.zeros(rows)
    for i, name in enumerate(DENSE):
        values = rng.poisson(3.0 + i, rows).astype(np.float64)
        if i < 2:
            signal += (np.log1p(values) - np.log(4.0 + i)) * (1.0 if i == 0 else -0.7)
        mask = rng.random(rows) < 0.05           # Criteo has holes; fillna below closes them
        cols[name] = pa.array(values.astype(np.int64), mask=mask)
    for j, name in enumerate(SPARSE):
        ids = rng.integers(0, 1000 * (j + 1), rows)
        if j == 0:
            signal += (ids % 2) * 1.5 - 0.75
        cols[name] = pa.array([format(v * 2654435761 % (1 << 32), "08x") for v in ids])
    cols[LABEL] = pa.array((rng.random(rows) < 1 / (1 + np.exp(-signal))).astype(np.int32))
    table = pa.table(cols).select([LABEL] + DENSE + SPARSE)
    return spark.createDataFrame(table).repartition(partitions)


def preprocess(df, vocab: int):
    df = df.fillna(0, DENSE)
    for name in DENSE:
        df = df.withColumn(name, log(col(name) + 1.0).cast("float"))
    for name in SPARSE:
        df = df.withColumn(name, (col_abs(col_hash(col(name))) % vocab).cast("long"))
    return df.withColumn(LABEL, col(LABEL).cast("float"))


def preprocess_with_dictionaries(df, min_count: int = 2):
    """The notebook's Criteo recipe (cells 6-7) on this DataFrame API: count every (column, value)
    pair with ``posexplode(array(...))`` + ``groupBy``, rank values per column by frequency with a
    ``row_number`` window (rare values below ``min_count`` fall out and become id 0 after ``fillna``),
    join the dictionaries back column by column, log-transform the integer features and put the rows
    in a random order with a partition-local sort.  Returns ``(frame, table sizes)``."""
    from raydp.sql import Window
    from raydp.sql.functions import array, desc, posexplode, rand, row_number
    from raydp.sql.functions import count as count_rows

    pairs = (df.select(posexplode(array(*SPARSE)))
             .withColumnRenamed("pos", "column_id").withColumnRenamed("col", "value")
             .filter(col("value").isNotNull())
             .groupBy("column_id", "value").count()
             .filter(col("count") >= min_count))
    by_frequency = Window.partitionBy("column_id").orderBy(desc("count"))
    ranked = pairs.withColumn("id", row_number().over(by_frequency)).drop("count").cache()   # ids start at 1
    sizes = []
    for j, name in enumerate(SPARSE):
        dictionary = ranked.filter(col("column_id") == j).drop("column_id").withColumnRenamed("value", name)
        sizes.append(dictionary.agg(count_rows("*").alias("n")).collect()[0].n + 1)     # + the "rare" id 0
        df = df.join(dictionary, name, how="left").drop(name).withColumnRenamed("id", name)
    df = df.fillna(0, SPARSE + DENSE)
    for name in DENSE:
        df = df.withColumn(name, log(col(name) + 3.0).cast("float"))
    for name in SPARSE:
        df = df.withColumn(name, col(name).cast("long"))
    df = df.withColumn(LABEL, col(LABEL).cast("float"))
    ordinal = "__ordinal"
    df = df.withColumn(ordinal, (rand(7) * float(1 << 52)).cast("long"))
    df = df.repartition(4, ordinal).sortWithinPartitions(ordinal).drop(ordinal)
    return df.select(LABEL, *DENSE, *SPARSE), sizes


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=20000)
    ap.add_argument("--vocab", type=int, default=5000, help="rows per embedding table after hashing")
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--table-optimizer", default="sgd", choices=["sgd", "adagrad", "adam"],
                    help="sparse optimizer of the embedding tables on the GPU engine (row-wise Adagrad / lazy Adam)")
    ap.add_argument("--dictionary", action="store_true",
                    help="frequency-ranked dictionary encoding (the notebook's recipe) instead of hashing")
    args = ap.parse_args()

    ray.init(address="local", num_cpus=4 + args.workers)
    spark = raydp.init_spark("DLRM on Criteo-shaped data", num_executors=2, executor_cores=1,
                             executor_memory="1G")
    raw = synthetic_criteo(spark, args.rows, partitions=4)
    if args.dictionary:
        frame, table_sizes = preprocess_with_dictionaries(raw)
        print("table sizes:", table_sizes)
    else:
        frame, table_sizes = preprocess(raw, args.vocab), [args.vocab] * 26
    train_df, test_df = random_split(frame, [0.9, 0.1], 0)

    model = DLRM(num_dense=13, table_sizes=table_sizes, embedding_dim=32,
                 sparse_grad=False)               # dense grads: DDP over NCCL cannot reduce sparse ones
    estimator = TorchEstimator(num_workers=args.workers, model=model,
                               optimizer=torch.optim.SGD(model.parameters(), lr=args.lr),
                               loss=nn.BCEWithLogitsLoss(), feature_columns=DENSE + SPARSE,
                               feature_types=[torch.float] * 13 + [torch.long] * 26,
                               label_column=LABEL, label_type=torch.float, batch_size=args.batch_size,
                               num_epochs=args.epochs, drop_last=True, use_gpu=args.gpu,
                               resources_per_worker={"GPU": 1} if args.gpu else None,
                               metrics_name=["Accuracy"], embedding_optimizer=args.table_optimizer)
    t0 = time.perf_counter()
    estimator.fit_on_spark(train_df, test_df)
    wall = time.perf_counter() - t0
    history = estimator.results.metrics_history
    train = [h for h in history if "train_loss" in h]
    print(json.dumps({"example": "pytorch_dlrm", "engine": train[0].get("engine"), "workers": args.workers,
                      "first_train_loss": train[0]["train_loss"], "final_train_loss": train[-1]["train_loss"],
                      "accuracy": train[-1]["train_res"].get("Accuracy"), "fit_wall_s": wall}))
    trained = estimator.get_model()
    print(type(trained).__name__, "parameters:", sum(p.numel() for p in trained.parameters()))
    raydp.stop_spark()
    ray.shutdown()


if __name__ == "__main__":
    main()
