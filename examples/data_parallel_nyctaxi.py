"""User-managed data parallelism on the NYC-taxi model (what the reference does with Horovod).

The reference's ``examples/horovod_nyctaxi.py`` wraps the optimizer in
``hvd.DistributedOptimizer(op=hvd.Average)``, broadcasts the initial parameters and scales the
learning rate by the number of workers.  The same recipe here needs no extra dependency:

* on GPU workers the model is wrapped in ``raydp.parallel.FusedDataParallel``: gradients are views
  of NVLink-registered buckets, and one kernel per bucket all-reduces them across the ranks and
  applies the Adam/SGD update while the rest of backward is still running;
* on CPU workers gradients are averaged with ``torch.distributed.all_reduce`` over gloo.

    python examples/data_parallel_nyctaxi.py --workers 2
    python examples/data_parallel_nyctaxi.py --workers 2 --gpu
"""
import argparse

import _bootstrap  # noqa: F401
import torch
import torch.distributed as dist
import torch.nn as nn

import raydp
from raydp import runtime as ray
from raydp import train
from raydp.train import ScalingConfig, TorchTrainer

from data_process import LABEL, csv_path, feature_names, nyc_taxi_preprocess
from pytorch_nyctaxi import NYC_Model


def train_fn(config):
    ctx = train.get_context()
    world, rank, device = ctx.get_world_size(), ctx.get_world_rank(), train.get_device()
    torch.manual_seed(1234 + rank)                       # replicas start different on purpose ...
    model = NYC_Model(len(config["features"])).to(device)
    optimizer = torch.optim.Adam(model.parameters(), lr=config["base_lr"] * world)   # lr scaled by workers
    criterion = nn.SmoothL1Loss()
    fused = None
    if device.type == "cuda":
        from raydp.parallel import FusedDataParallel
        fused = FusedDataParallel(model, optimizer, device=device)     # ... and are broadcast from rank 0 here
    elif world > 1:
        for p in model.parameters():
            dist.broadcast(p.data, src=0)
    shard = train.get_dataset_shard("train")
    for epoch in range(config["num_epochs"]):
        total, steps = 0.0, 0
        for x, y in shard.to_torch(feature_columns=config["features"], feature_column_dtypes=torch.float,
                                   label_column=LABEL, label_column_dtype=torch.float,
                                   batch_size=config["batch_size"], drop_last=True):
            loss = criterion(model(x.to(device)), y.to(device))
            if fused is not None:
                fused.backward_and_step(loss)            # backward + all-reduce + optimizer, overlapped
            else:
                optimizer.zero_grad()
                loss.backward()
                if world > 1:
                    for p in model.parameters():
                        dist.all_reduce(p.grad)
                        p.grad /= world
                optimizer.step()
            total += loss.item()
            steps += 1
        train.report({"epoch": epoch, "train_loss": total / max(1, steps)})
    # replicas must have stayed identical
    flat = torch.cat([p.detach().reshape(-1).float().cpu() for p in model.parameters()])
    if world > 1:
        gathered = [torch.empty_like(flat) for _ in range(world)] if rank == 0 else None
        if dist.get_backend() == "nccl":
            flat_dev = flat.to(device)
            out = [torch.empty_like(flat_dev) for _ in range(world)]
            dist.all_gather(out, flat_dev)
            drift = max(float((o - out[0]).abs().max()) for o in out)
        else:
            dist.gather(flat, gathered, dst=0)
            drift = max(float((g - gathered[0]).abs().max()) for g in gathered) if rank == 0 else 0.0
        train.report({"replica_drift": drift})
    if fused is not None:
        fused.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--gpu", action="store_true")
    args = ap.parse_args()

    ray.init(address="local", num_cpus=4 + args.workers)
    spark = raydp.init_spark("NYC Taxi Fare Prediction with RayDP", num_executors=1, executor_cores=1,
                             executor_memory="500M")
    frame = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path())
    spark.conf.set("spark.sql.session.timeZone", "UTC")
    frame = nyc_taxi_preprocess(frame)
    features = feature_names(frame)
    dataset = raydp.data.from_spark(frame)

    trainer = TorchTrainer(train_fn,
                           train_loop_config={"features": features, "num_epochs": args.epochs,
                                              "batch_size": 64, "base_lr": 1e-3},
                           scaling_config=ScalingConfig(num_workers=args.workers, use_gpu=args.gpu),
                           datasets={"train": dataset})
    result = trainer.fit()
    for row in result.metrics_history:
        print(row)
    raydp.stop_spark()
    ray.shutdown()
