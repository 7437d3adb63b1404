"""NYC-taxi fare regression with ``TorchEstimator`` (``BASELINE.json`` config 1).

Same flow as the reference's ``examples/pytorch_nyctaxi.py``: start the runtime, ``init_spark``,
read the CSV, engineer 20 features on the executors, split 90/10, fit the
20-256-128-64-16-1 ReLU+BatchNorm MLP with SmoothL1 + Adam, fetch the trained model.

    python examples/pytorch_nyctaxi.py                     # 2 CPU workers over gloo
    python examples/pytorch_nyctaxi.py --gpu --workers 1   # tcgen05 engine on a B200

On GPU workers the estimator recognises the layer stack and runs it on the hand-written sm_100a
kernels; the checkpoint (``model.pt``) stays a plain ``state_dict`` of the module below.
"""
import argparse
import json
import time

import _bootstrap  # noqa: F401
import torch
import torch.nn as nn
import torch.nn.functional as F

import raydp
from raydp import runtime as ray
from raydp.torch import TorchEstimator
from raydp.utils import random_split

from data_process import LABEL, csv_path, feature_names, nyc_taxi_preprocess


class NYC_Model(nn.Module):
    """Linear -> ReLU -> BatchNorm1d x4, then a scalar head (reference ``pytorch_nyctaxi.py:69-92``)."""

    WIDTHS = (256, 128, 64, 16)

    def __init__(self, cols: int):
        super().__init__()
        dims = (cols,) + self.WIDTHS
        for i, (fan_in, fan_out) in enumerate(zip(dims[:-1], dims[1:]), start=1):
            setattr(self, f"fc{i}", nn.Linear(fan_in, fan_out))
            setattr(self, f"bn{i}", nn.BatchNorm1d(fan_out))
        self.fc5 = nn.Linear(dims[-1], 1)

    def forward(self, x):
        for i in range(1, 5):
            x = getattr(self, f"bn{i}")(F.relu(getattr(self, f"fc{i}")(x)))
        return self.fc5(x)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=30)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--gpu", action="store_true", help="one GPU per worker (fused sm_100a engine)")
    args = ap.parse_args()

    ray.init(address="local", num_cpus=4 + args.workers)
    spark = raydp.init_spark("NYC Taxi Fare Prediction with RayDP", num_executors=1, executor_cores=1,
                             executor_memory="500M")
    frame = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path())
    spark.conf.set("spark.sql.session.timeZone", "UTC")
    frame = nyc_taxi_preprocess(frame)
    train_df, test_df = random_split(frame, [0.9, 0.1], 0)
    features = feature_names(train_df)

    if args.gpu:
        from raydp.models import NYCTaxiMLP       # same layers + fused_spec() for the tcgen05 engine
        model = NYCTaxiMLP(len(features))
    else:
        model = NYC_Model(len(features))
    estimator = TorchEstimator(num_workers=args.workers, model=model,
                               optimizer=torch.optim.Adam(model.parameters(), lr=1e-3),
                               loss=nn.SmoothL1Loss(), feature_columns=features, feature_types=torch.float,
                               label_column=LABEL, label_type=torch.float, batch_size=args.batch_size,
                               num_epochs=args.epochs, use_gpu=args.gpu,
                               resources_per_worker={"GPU": 1} if args.gpu else None,
                               metrics_name=["MeanAbsoluteError", "MeanSquaredError"], use_ccl=False)
    t0 = time.perf_counter()
    estimator.fit_on_spark(train_df, test_df)
    wall = time.perf_counter() - t0

    history = estimator.results.metrics_history
    train = [h for h in history if "train_loss" in h]
    evals = [h for h in history if "test_loss" in h]
    samples = sum(h.get("samples_this_iter", 0) for h in train)
    busy = sum(h.get("time_this_iter_s", 0.0) for h in train)
    print(json.dumps({"example": "pytorch_nyctaxi", "workers": args.workers, "gpu": args.gpu,
                      "first_train_loss": train[0]["train_loss"], "final_train_loss": train[-1]["train_loss"],
                      "final_test_loss": evals[-1]["test_loss"] if evals else None,
                      "train_samples_per_s": samples * args.workers / busy if busy else None,
                      "fit_wall_s": wall}))
    trained = estimator.get_model()
    print(type(trained).__name__, "parameters:", sum(p.numel() for p in trained.parameters()))
    raydp.stop_spark()
    ray.shutdown()


if __name__ == "__main__":
    main()
