"""TorchEstimator on GPU workers: fused engine end to end (ingest -> fused step -> checkpoint)."""
import os

import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


def _frame(spark, n, cols=20):
    from raydp_b200.sql.functions import col, rand
    df = spark.range(0, n)
    for i in range(cols):
        df = df.withColumn(f"f{i}", rand(i + 1) * 2 - 1)
    target = col("f0") * 3 + col("f1") * 2 - col("f2") + 5
    return df.withColumn("fare_amount", target).drop("id")


def test_fused_estimator_single_gpu(ray_cluster):
    import raydp_b200 as raydp
    from raydp_b200.models import NYCTaxiMLP
    from raydp_b200.torch import TorchEstimator
    spark = raydp.init_spark("gpu-est", 2, 1, "1GB")
    df = _frame(spark, 60000)
    train_df, test_df = df.randomSplit([0.8, 0.2], seed=1)
    features = [f"f{i}" for i in range(20)]
    model = NYCTaxiMLP(20)
    est = TorchEstimator(num_workers=1, resources_per_worker={"GPU": 1}, model=model,
                         optimizer=torch.optim.Adam(model.parameters(), lr=3e-3), loss=nn.SmoothL1Loss(),
                         feature_columns=features, feature_types=torch.float, label_column="fare_amount",
                         label_type=torch.float, batch_size=2048, num_epochs=6, shuffle=False,
                         metrics_name=["MeanAbsoluteError", "MeanSquaredError"],
                         lr_scheduler_creator=lambda o, c: torch.optim.lr_scheduler.StepLR(o, 1000, 0.9))
    est.fit_on_spark(train_df, test_df)
    hist = est.results.metrics_history
    train = [h for h in hist if "train_loss" in h]
    evals = [h for h in hist if "test_loss" in h]
    assert train[0]["engine"] == "fused"
    assert train[-1]["train_loss"] < train[0]["train_loss"] * 0.5
    assert evals[-1]["test_loss"] < evals[0]["test_loss"]
    assert train[-1]["h2d_bytes_this_iter"] > 0 and set(train[-1]["train_res"]) == {"MeanAbsoluteError", "MeanSquaredError"}
    trained = est.get_model()
    assert trained is model
    state = torch.load(os.path.join(est.results.checkpoint.path, "model.pt"), weights_only=True)
    assert "fc1.weight" in state and "bn4.running_var" in state and state["fc1.weight"].shape == (256, 20)
    # the trained plain module reproduces the evaluation loss (checkpoint round trip is faithful)
    pdf = test_df.toPandas()
    x = torch.tensor(pdf[features].values, dtype=torch.float32)
    y = torch.tensor(pdf["fare_amount"].values, dtype=torch.float32)
    trained.eval()
    with torch.no_grad():
        loss = nn.SmoothL1Loss()(trained(x)[:, 0], y).item()
    assert abs(loss - evals[-1]["test_loss"]) < 0.1 * max(1.0, evals[-1]["test_loss"])
    raydp.stop_spark()


@pytest.mark.multigpu
def test_fused_estimator_data_parallel(ray_cluster):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import raydp_b200 as raydp
    from raydp_b200 import runtime
    from raydp_b200.models import NYCTaxiMLP
    from raydp_b200.torch import TorchEstimator
    runtime.shutdown()
    runtime.init(num_cpus=6, num_gpus=2)
    spark = raydp.init_spark("gpu-dp", 1, 1, "1GB")
    df = _frame(spark, 80000)
    model = NYCTaxiMLP(20)
    est = TorchEstimator(num_workers=2, resources_per_worker={"GPU": 1}, model=model,
                         optimizer=torch.optim.Adam(model.parameters(), lr=3e-3), loss=nn.SmoothL1Loss(),
                         feature_columns=[f"f{i}" for i in range(20)], feature_types=torch.float,
                         label_column="fare_amount", label_type=torch.float, batch_size=2048, num_epochs=4,
                         shuffle=False, drop_last=True)
    est.fit_on_spark(df)
    train = [h for h in est.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused" and train[-1]["train_loss"] < train[0]["train_loss"] * 0.8
    raydp.stop_spark()


def test_dlrm_estimator_single_gpu(ray_cluster):
    """DLRM through TorchEstimator: dense + categorical columns, fused hybrid engine, full checkpoint."""
    import numpy as np
    import pyarrow as pa
    import raydp_b200 as raydp
    from raydp_b200 import data as rdata
    from raydp_b200.models import DLRM
    from raydp_b200.torch import TorchEstimator
    rng = np.random.default_rng(0)
    n, sizes = 16384, [50 + 7 * t for t in range(26)]
    cols = {f"d{i}": rng.standard_normal(n).astype(np.float32) for i in range(13)}
    for t, s in enumerate(sizes):
        cols[f"c{t}"] = rng.integers(0, s, n).astype(np.int64)
    logit = cols["d0"] * 1.5 - cols["d1"] + (cols["c0"] % 2) * 2.0 - 1.0
    cols["label"] = (rng.random(n) < 1 / (1 + np.exp(-logit))).astype(np.float32)
    ds = rdata.from_arrow([pa.table({k: v[i::2] for k, v in cols.items()}) for i in range(2)])
    torch.manual_seed(0)
    model = DLRM(13, sizes, 32, sparse_grad=False)
    features = [f"d{i}" for i in range(13)] + [f"c{t}" for t in range(26)]
    est = TorchEstimator(num_workers=1, resources_per_worker={"GPU": 1}, model=model,
                         optimizer=torch.optim.SGD(model.parameters(), lr=0.1), loss=nn.BCEWithLogitsLoss(),
                         feature_columns=features, feature_types=[torch.float] * 13 + [torch.long] * 26,
                         label_column="label", label_type=torch.float, batch_size=1024, num_epochs=12,
                         shuffle=False, drop_last=True, metrics_name=["Accuracy"])
    est.fit(ds, ds)
    train = [h for h in est.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused-dlrm"
    # Plain SGD at lr 0.1 is the stable regime on this data (the PyTorch model goes 0.694 -> 0.686 in 8
    # epochs).  lr 0.5, or 0.05 with momentum 0.9, sit at the edge of instability: with the same
    # initialisation the final loss then lands anywhere between 0.56 and 0.71 depending on the order
    # of the atomic accumulations.  Step-by-step numerics against PyTorch are in test_gpu_dlrm.py;
    # here the loss must go down and, below, the checkpoint must reproduce the engine's evaluation.
    assert all(h["train_loss"] == h["train_loss"] for h in train)
    assert train[-1]["train_loss"] < train[0]["train_loss"] - 0.002
    trained = est.get_model()
    state = torch.load(os.path.join(est.results.checkpoint.path, "model.pt"), weights_only=True)
    assert state["embeddings.25.weight"].shape == (sizes[25], 32) and "top.4.bias" in state
    trained.eval()
    # checkpoint fidelity: the plain torch module restored from model.pt must score like the fused
    # engine did on the same rows in its last evaluation epoch
    x = torch.tensor(np.stack([cols[f"d{i}"] for i in range(13)], 1))
    s = torch.tensor(np.stack([cols[f"c{t}"] for t in range(26)], 1))
    with torch.no_grad():
        acc = ((trained(x, s)[:, 0] > 0).float() == torch.tensor(cols["label"])).float().mean().item()
    evals = [h for h in est.results.metrics_history if "eval_res" in h]
    assert abs(acc - evals[-1]["eval_res"]["Accuracy"]) < 0.02, (acc, evals[-1]["eval_res"])
    assert acc > 0.5


def test_dlrm_estimator_with_momentum_is_bit_reproducible(ray_cluster):
    """Two fits from the same initialisation (SGD momentum 0.9 on the dense layers, row-wise Adagrad on the
    tables, a multi-hot table, a ragged last batch) write the same BITS into model.pt: weight gradients are
    ordered split sums, bias gradients ordered column sums, duplicate table rows are combined in fixed point."""
    import numpy as np
    import pyarrow as pa
    import raydp_b200 as raydp  # noqa: F401
    from raydp_b200 import data as rdata
    from raydp_b200.models import DLRM
    from raydp_b200.torch import TorchEstimator
    rng = np.random.default_rng(1)
    n, sizes, bags = 5000, [11, 300, 57, 1000, 8], [1, 3, 1, 1, 2]
    cols = {f"d{i}": rng.standard_normal(n).astype(np.float32) for i in range(13)}
    names = []
    for t, (s, L) in enumerate(zip(sizes, bags)):
        for l in range(L):
            v = rng.integers(0, s, n).astype(np.int64)
            if l > 0:
                v[rng.random(n) < 0.3] = -1                      # ragged bags
            cols[f"c{t}_{l}"] = v
            names.append(f"c{t}_{l}")
    logit = cols["d0"] * 1.5 - cols["d1"] + (cols["c0_0"] % 2) * 2.0 - 1.0
    cols["label"] = (rng.random(n) < 1 / (1 + np.exp(-logit))).astype(np.float32)
    ds = rdata.from_arrow([pa.table({k: v[i::2] for k, v in cols.items()}) for i in range(2)])
    features = [f"d{i}" for i in range(13)] + names
    states = []
    for run in range(2):
        torch.manual_seed(0)
        model = DLRM(13, sizes, 32, bottom=(64, 32), top=(128, 64, 1), sparse_grad=False, bag_sizes=bags,
                     pooling="mean")
        est = TorchEstimator(num_workers=1, resources_per_worker={"GPU": 1}, model=model,
                             optimizer=torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9),
                             loss=nn.BCEWithLogitsLoss(), feature_columns=features,
                             feature_types=[torch.float] * 13 + [torch.long] * len(names), label_column="label",
                             label_type=torch.float, batch_size=512, num_epochs=4, shuffle=False, drop_last=False,
                             metrics_name=["Accuracy"], embedding_optimizer="adagrad")
        est.fit(ds)
        train = [h for h in est.results.metrics_history if "train_loss" in h]
        assert train[0]["engine"] == "fused-dlrm" and train[-1]["train_loss"] < train[0]["train_loss"]
        states.append(torch.load(os.path.join(est.results.checkpoint.path, "model.pt"), weights_only=True))
    assert states[0].keys() == states[1].keys()
    for k in states[0]:
        assert torch.equal(states[0][k], states[1][k]), k
    assert not torch.equal(states[0]["embeddings.1.weight"], model.embeddings[1].weight.detach().cpu())   # it trained


@pytest.mark.multigpu
def test_rank_killed_mid_training_is_detected_and_the_group_restarts(ray_cluster, tmp_path):
    """Failure path of the peer-memory collectives: rank 1 dies in the middle of an epoch.  Rank 0's next
    all-reduce must not hang on the dead peer (bounded waits), the trainer must see the failure, tear the group
    down and -- ``max_retries=1`` -- run the fit again to completion (reference recovery semantics:
    ``FailureConfig(max_failures=max_retries)``, torch/estimator.py:334)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    import time
    import raydp_b200 as raydp
    from raydp_b200 import runtime
    from raydp_b200.torch import TorchEstimator
    from user_models import NYC_Model
    os.environ["RAYDP_B200_COMM_TIMEOUT_S"] = "5"
    runtime.shutdown()
    runtime.init(num_cpus=6, num_gpus=2)
    flag = str(tmp_path / "rank1_died")
    try:
        spark = raydp.init_spark("gpu-kill", 1, 1, "1GB")
        df = _frame(spark, 64 * 2048)

        def scheduler_that_kills_rank_1(optimizer, config):
            from raydp_b200.train import get_context

            class Sched(torch.optim.lr_scheduler.LambdaLR):
                def step(self, *a, **kw):
                    super().step(*a, **kw)
                    if self.last_epoch == 9 and get_context().get_world_rank() == 1 and not os.path.exists(flag):
                        open(flag, "w").close()
                        os._exit(7)
            return Sched(optimizer, lambda step: 1.0)
        model = NYC_Model(20)
        est = TorchEstimator(num_workers=2, resources_per_worker={"GPU": 1}, model=model,
                             optimizer=torch.optim.Adam(model.parameters(), lr=3e-3), loss=nn.SmoothL1Loss(),
                             feature_columns=[f"f{i}" for i in range(20)], feature_types=torch.float,
                             label_column="fare_amount", label_type=torch.float, batch_size=2048, num_epochs=3,
                             shuffle=False, drop_last=True, lr_scheduler_creator=scheduler_that_kills_rank_1)
        t0 = time.time()
        est.fit_on_spark(df, max_retries=1)
        assert os.path.exists(flag), "the failure was never injected"
        assert time.time() - t0 < 240
        train = [h for h in est.results.metrics_history if "train_loss" in h]
        assert len(train) == 3 and train[0]["engine"] == "fused"
        assert train[-1]["train_loss"] < train[0]["train_loss"]
        raydp.stop_spark()
    finally:
        os.environ.pop("RAYDP_B200_COMM_TIMEOUT_S", None)
