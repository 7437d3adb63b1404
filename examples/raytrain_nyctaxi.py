"""NYC-taxi regression with a hand-written training loop under ``TorchTrainer``.

Counterpart of the reference's ``examples/raytrain_nyctaxi.py``: Spark preprocessing, conversion
to a distributed dataset, then a user loop per worker that pulls its shard from the session,
wraps the model for data parallelism and reports metrics + a checkpoint every epoch.

    python examples/raytrain_nyctaxi.py --workers 2
"""
import argparse
import os
import tempfile

import _bootstrap  # noqa: F401
import torch
import torch.nn as nn

import raydp
from raydp import runtime as ray
from raydp import train
from raydp.train import Checkpoint, RunConfig, ScalingConfig, TorchTrainer
from raydp.utils import random_split

from data_process import LABEL, csv_path, feature_names, nyc_taxi_preprocess
from pytorch_nyctaxi import NYC_Model


def train_loop_per_worker(config):
    """Runs on every worker: one rank of a data-parallel job."""
    device = train.get_device()
    model = train.prepare_model(NYC_Model(len(config["features"])))
    optimizer = torch.optim.Adam(model.parameters(), lr=config["lr"])
    criterion = nn.SmoothL1Loss()
    shard = train.get_dataset_shard("train")
    test_shard = train.get_dataset_shard("test")
    rank = train.get_context().get_world_rank()

    def batches(ds):
        return ds.to_torch(feature_columns=config["features"], feature_column_dtypes=torch.float,
                           label_column=LABEL, label_column_dtype=torch.float,
                           batch_size=config["batch_size"], drop_last=False)

    for epoch in range(config["num_epochs"]):
        model.train()
        running, steps = 0.0, 0
        for x, y in batches(shard):
            x, y = x.to(device), y.to(device)
            loss = criterion(model(x), y)
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()
            running += loss.item()
            steps += 1
        model.eval()
        test_loss, test_steps = 0.0, 0
        with torch.no_grad():
            for x, y in batches(test_shard):
                test_loss += criterion(model(x.to(device)), y.to(device)).item()
                test_steps += 1
        metrics = {"epoch": epoch, "train_loss": running / max(1, steps),
                   "test_loss": test_loss / max(1, test_steps)}
        with tempfile.TemporaryDirectory() as tmp:
            checkpoint = None
            if rank == 0:
                module = model.module if hasattr(model, "module") else model
                torch.save(module.state_dict(), os.path.join(tmp, "model.pt"))
                checkpoint = Checkpoint.from_directory(tmp)
            train.report(metrics, checkpoint=checkpoint)
        if rank == 0:
            print(metrics, flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--gpu", action="store_true")
    args = ap.parse_args()

    ray.init(address="local", num_cpus=4 + args.workers)
    spark = raydp.init_spark("NYC Taxi Fare Prediction with RayDP", num_executors=1, executor_cores=1,
                             executor_memory="500M")
    frame = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path())
    spark.conf.set("spark.sql.session.timeZone", "UTC")
    frame = nyc_taxi_preprocess(frame)
    train_df, test_df = random_split(frame, [0.9, 0.1], 0)
    features = feature_names(train_df)
    train_ds = raydp.data.from_spark(train_df)
    test_ds = raydp.data.from_spark(test_df)
    # the datasets are owned by the master actor, so Spark's executors can go away before training
    raydp.stop_spark(cleanup_data=False)

    trainer = TorchTrainer(
        train_loop_per_worker,
        train_loop_config={"features": features, "num_epochs": args.epochs, "batch_size": 64, "lr": 1e-3},
        scaling_config=ScalingConfig(num_workers=args.workers, use_gpu=args.gpu),
        run_config=RunConfig(name="raytrain_nyctaxi"),
        datasets={"train": train_ds, "test": test_ds})
    result = trainer.fit()
    print("final metrics:", result.metrics)
    state = torch.load(os.path.join(result.checkpoint.path, "model.pt"), weights_only=True)
    restored = NYC_Model(len(features))
    restored.load_state_dict(state)
    print("restored", type(restored).__name__, "from", result.checkpoint.path)
    raydp.stop_spark()
    ray.shutdown()
