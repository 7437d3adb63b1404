"""TorchEstimator end to end on CPU/gloo workers.

Mirrors reference ``tests/test_torch.py:32-94`` (linear z = 3x + 4y + 5, ``num_workers=2``, with
and without ``fs_directory``) and ``test_torch_sequential.py:23-53`` (more workers than is
comfortable for 6 rows), and additionally asserts numerics: the loss goes down and the
checkpoint layout is ``model.pt`` = state_dict of the unwrapped module.
"""
import os

import pytest
import torch
import torch.nn as nn

import raydp_b200.sql.pandas_api as ps
from raydp_b200 import data as rdata
from raydp_b200.sql.functions import col, rand
from raydp_b200.torch import TorchEstimator
from raydp_b200.utils import random_split


class LinearModel(nn.Module):
    def __init__(self):
        super().__init__()
        self.linear = nn.Linear(2, 1)

    def forward(self, x):
        return self.linear(x)


def _frame(spark, n=12000):
    return spark.range(0, n).withColumn("x", rand(1) * 10).withColumn("y", rand(2) * 10) \
        .withColumn("z", col("x") * 3 + col("y") * 4 + 5).drop("id")


@pytest.mark.parametrize("use_fs_directory", [False, True])
def test_torch_estimator_two_workers(spark_on_ray_2_executors, tmp_path, use_fs_directory):
    spark = spark_on_ray_2_executors
    train_df, test_df = random_split(_frame(spark), [0.7, 0.3], 0)
    model = LinearModel()
    optimizer = torch.optim.Adam(model.parameters(), lr=0.2)
    est = TorchEstimator(num_workers=2, model=model, optimizer=optimizer, loss=nn.MSELoss(),
                         feature_columns=["x", "y"], feature_types=torch.float, label_column="z",
                         label_type=torch.float, batch_size=500, num_epochs=4, use_gpu=False,
                         metrics_name=["MeanAbsoluteError", "MeanSquaredError"], shuffle_seed=3)
    if use_fs_directory:
        est.fit_on_spark(train_df, test_df, fs_directory="file://" + str(tmp_path / "data"))
        assert any(f.endswith(".parquet") for _, _, fs in os.walk(str(tmp_path / "data")) for f in fs)
    else:
        est.fit_on_spark(train_df, test_df)
    history = est.results.metrics_history
    train = [h for h in history if "train_loss" in h]
    evals = [h for h in history if "test_loss" in h]
    assert len(train) == 4 and len(evals) == 4
    assert train[-1]["train_loss"] < train[0]["train_loss"] * 0.5
    assert set(train[0]["train_res"]) == {"MeanAbsoluteError", "MeanSquaredError"}
    trained = est.get_model()
    assert trained is model
    assert trained(torch.Tensor([[1, 1], [1, 2]])).shape == (2, 1)
    ckpt_dir = est.results.checkpoint.path
    state = torch.load(os.path.join(ckpt_dir, "model.pt"), weights_only=True)
    assert set(state) == {"linear.weight", "linear.bias"}          # unwrapped: no "module." prefix


def test_torch_estimator_creator_functions_and_koalas(spark_on_ray_small):
    pdf = ps.range(0, 2000)
    pdf["x"] = pdf["id"] * 0.001
    pdf["y"] = pdf["id"] * 0.002
    pdf["z"] = pdf["x"] * 3 + pdf["y"] * 4 + 5
    est = TorchEstimator(num_workers=1, model=lambda cfg: LinearModel(),
                         optimizer=lambda m, cfg: torch.optim.SGD(m.parameters(), lr=0.05),
                         loss=nn.SmoothL1Loss,                         # a class is instantiated
                         lr_scheduler_creator=lambda o, cfg: torch.optim.lr_scheduler.StepLR(o, 50, 0.9),
                         feature_columns=["x", "y"], feature_types=torch.float, label_column="z",
                         label_type=torch.float, batch_size=100, num_epochs=2, use_gpu=False)
    est.fit_on_spark(pdf)
    losses = [h["train_loss"] for h in est.results.metrics_history if "train_loss" in h]
    assert losses[1] < losses[0]
    assert isinstance(est.get_model(), LinearModel)


def test_torch_sequential_more_workers_than_rows(spark_on_ray_small):
    spark = spark_on_ray_small
    df = spark.createDataFrame([(0.0, 1.0, 2.0, 3.0)] * 6, ["a", "b", "c", "label"])
    model = nn.Sequential(nn.Linear(3, 2), nn.ReLU(), nn.Linear(2, 1))
    est = TorchEstimator(num_workers=3, model=model, optimizer=torch.optim.Adam(model.parameters()),
                         loss=nn.MSELoss(), feature_columns=["a", "b", "c"], feature_types=torch.float,
                         label_column="label", label_type=torch.float, batch_size=1, num_epochs=2,
                         use_gpu=False)
    est.fit_on_spark(df)
    assert est.get_model()(torch.zeros(1, 3)).shape == (1, 1)


def test_bad_arguments():
    with pytest.raises(TypeError):
        TorchEstimator(model=LinearModel(), batch_size=1, num_epochs=1, num_processes_for_data_loader=2)
    with pytest.raises(ValueError):
        TorchEstimator(model=LinearModel())
