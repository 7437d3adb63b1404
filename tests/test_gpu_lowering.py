"""Unchanged user modules on GPU workers: the reference's ``NYC_Model`` and a notebook-style ``Dlrm`` must
select the fused engines with no ``fused_spec()``; arbitrary modules must run on the generic GPU engine
(tcgen05 Linear layers + fused data parallel + GPU ingest), and after N steps the weights must agree with
plain fp32 PyTorch run on the same batches in the same order."""
import copy
import os

import numpy as np
import pyarrow as pa
import pytest
import torch
import torch.nn as nn

from user_models import Dlrm, GeluNet, NYC_Model

pytestmark = pytest.mark.gpu

F = 20
NAMES = [f"f{i}" for i in range(F)]


def _regression_columns(n, seed=0):
    rng = np.random.default_rng(seed)
    cols = {name: rng.standard_normal(n).astype(np.float32) for name in NAMES}
    y = 3 * cols["f0"] + 2 * cols["f1"] - cols["f2"] + 0.5 * cols["f3"] * cols["f4"] + 5
    cols["fare_amount"] = y.astype(np.float32)
    return cols


def _dataset(cols, n_blocks=2):
    from raydp_b200 import data as rdata
    n = len(next(iter(cols.values())))
    per = n // n_blocks
    return rdata.from_arrow([pa.table({k: v[b * per:(b + 1) * per] for k, v in cols.items()})
                             for b in range(n_blocks)])


def _torch_reference(model, make_opt, loss_fn, cols, batch, epochs):
    """The reference loop (torch/estimator.py:272-293) in fp32 on the CPU, same batch order."""
    x = torch.tensor(np.stack([cols[n] for n in NAMES], 1))
    y = torch.tensor(cols["fare_amount"]).unsqueeze(1)
    opt = make_opt(model)
    model.train()
    losses = []
    for _ in range(epochs):
        total, steps = 0.0, 0
        for lo in range(0, len(x) - batch + 1, batch):
            out = model(x[lo:lo + batch])
            loss = loss_fn(out, y[lo:lo + batch])
            total += loss.item()
            opt.zero_grad()
            loss.backward()
            opt.step()
            steps += 1
        losses.append(total / steps)
    return losses


def _update_agreement(init_state, ours, ref, keys):
    """cosine similarity and relative error of the total parameter update, over the given tensors."""
    d_ours = torch.cat([(ours[k].float() - init_state[k].float()).reshape(-1) for k in keys])
    d_ref = torch.cat([(ref[k].float() - init_state[k].float()).reshape(-1) for k in keys])
    cos = torch.nn.functional.cosine_similarity(d_ours, d_ref, dim=0).item()
    rel = ((d_ours - d_ref).norm() / d_ref.norm()).item()
    return cos, rel


def _fit(model, make_opt, loss, cols, batch, epochs, **kw):
    from raydp_b200.torch import TorchEstimator
    est = TorchEstimator(num_workers=1, resources_per_worker={"GPU": 1}, model=model, optimizer=make_opt(model),
                         loss=loss, feature_columns=NAMES, feature_types=torch.float, label_column="fare_amount",
                         label_type=torch.float, batch_size=batch, num_epochs=epochs, shuffle=False, drop_last=True,
                         metrics_name=["MeanAbsoluteError", "MeanSquaredError"], **kw)
    est.fit(_dataset(cols))
    return est


def test_reference_nyc_model_runs_fused_and_matches_fp32_torch(ray_cluster):
    cols = _regression_columns(8 * 4096)
    torch.manual_seed(7)
    model = NYC_Model(F)
    assert not hasattr(model, "fused_spec")
    init = copy.deepcopy(model.state_dict())
    make_opt = lambda m: torch.optim.SGD(m.parameters(), lr=0.05)          # noqa: E731
    ref_model = copy.deepcopy(model)
    ref_losses = _torch_reference(ref_model, make_opt, nn.SmoothL1Loss(), cols, 4096, 2)
    est = _fit(model, make_opt, nn.SmoothL1Loss(), cols, 4096, 2)
    train = [h for h in est.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused", train[0]
    for got, want in zip(train, ref_losses):
        assert abs(got["train_loss"] - want) < 0.03 * want + 0.02, (got["train_loss"], want)
    ours = torch.load(os.path.join(est.results.checkpoint.path, "model.pt"), weights_only=True)
    ref = ref_model.state_dict()
    assert set(ours) == set(ref)
    weights = [k for k in ref if k.endswith("weight") or k.endswith("bias")]
    cos, rel = _update_agreement(init, ours, ref, weights)
    assert cos > 0.98 and rel < 0.2, (cos, rel)
    for k in ref:
        if "running_" in k:
            assert torch.allclose(ours[k], ref[k], rtol=0.05, atol=0.02), k
    assert int(ours["bn1.num_batches_tracked"]) == int(ref["bn1.num_batches_tracked"]) == 16
    # and Adam, the example's optimizer: same engine, loss follows the fp32 loop
    torch.manual_seed(7)
    model2 = NYC_Model(F)
    ref2 = copy.deepcopy(model2)
    adam = lambda m: torch.optim.Adam(m.parameters(), lr=1e-3)             # noqa: E731
    ref_losses = _torch_reference(ref2, adam, nn.SmoothL1Loss(), cols, 4096, 2)
    est2 = _fit(model2, adam, nn.SmoothL1Loss(), cols, 4096, 2)
    train = [h for h in est2.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused"
    assert abs(train[-1]["train_loss"] - ref_losses[-1]) < 0.03 * ref_losses[-1] + 0.02


def test_arbitrary_module_runs_on_the_generic_gpu_engine_and_matches_fp32_torch(ray_cluster):
    cols = _regression_columns(8 * 2048, seed=3)
    torch.manual_seed(11)
    model = GeluNet(F, width=64, p=0.0)
    init = copy.deepcopy(model.state_dict())
    make_opt = lambda m: torch.optim.SGD(m.parameters(), lr=0.005, momentum=0.9)  # noqa: E731
    ref_model = copy.deepcopy(model)
    ref_losses = _torch_reference(ref_model, make_opt, nn.MSELoss(), cols, 2048, 2)
    est = _fit(model, make_opt, nn.MSELoss(), cols, 2048, 2)
    train = [h for h in est.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused-generic", train[0]
    assert train[0]["fused_linears"] == 4 and train[0]["kernel_launches"] > 0
    assert train[0]["h2d_bytes_this_iter"] > 0
    for got, want in zip(train, ref_losses):
        assert abs(got["train_loss"] - want) < 0.05 * want + 0.05, (got["train_loss"], want)
    ours = torch.load(os.path.join(est.results.checkpoint.path, "model.pt"), weights_only=True)
    ref = ref_model.state_dict()
    assert set(ours) == set(ref)
    cos, rel = _update_agreement(init, ours, ref, list(ref))
    assert cos > 0.97 and rel < 0.25, (cos, rel)
    trained = est.get_model()
    assert trained is model


def test_generic_engine_with_dropout_and_an_optimizer_it_cannot_fuse(ray_cluster):
    cols = _regression_columns(8 * 2048, seed=5)
    torch.manual_seed(13)
    model = GeluNet(F, width=48, p=0.1)                     # width 48: not a multiple of 16 * anything special
    est = _fit(model, lambda m: torch.optim.Adagrad(m.parameters(), lr=0.05), nn.SmoothL1Loss(), cols, 2048, 4,
               lr_scheduler_creator=lambda o, c: torch.optim.lr_scheduler.StepLR(o, 100, 0.9))
    train = [h for h in est.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused-generic"
    assert train[-1]["train_loss"] < 0.7 * train[0]["train_loss"], [h["train_loss"] for h in train]


def _criteo_like(n, sizes, seed=0):
    rng = np.random.default_rng(seed)
    cols = {f"_c{i}": rng.standard_normal(n).astype(np.float32) for i in range(1, 14)}
    for t, s in enumerate(sizes):
        cols[f"_c{14 + t}"] = rng.integers(0, s, n).astype(np.int64)
    logit = cols["_c1"] * 1.5 - cols["_c2"] + (cols["_c14"] % 2) * 2.0 - 1.0
    cols["_c0"] = (rng.random(n) < 1 / (1 + np.exp(-logit))).astype(np.int32)     # IntegerType label, as in the notebook
    return cols


def test_notebook_style_dlrm_runs_on_the_fused_dlrm_engine(ray_cluster):
    from raydp_b200 import data as rdata
    from raydp_b200.torch import TorchEstimator
    sizes = [50 + 7 * t for t in range(26)]
    cols = _criteo_like(16384, sizes)
    ds = rdata.from_arrow([pa.table({k: v[i::2] for k, v in cols.items()}) for i in range(2)])
    torch.manual_seed(0)
    model = Dlrm(13, sizes, [512, 128, 32], [1024, 1024, 512, 256, 1])
    features = [f"_c{i}" for i in range(1, 40)]
    est = TorchEstimator(num_workers=1, resources_per_worker={"GPU": 1}, model=model,
                         optimizer=torch.optim.SGD(model.parameters(), lr=0.1), loss=nn.BCEWithLogitsLoss(reduction="mean"),
                         feature_columns=features, feature_types=[torch.float32] * 13 + [torch.long] * 26,
                         label_column="_c0", label_type=torch.float, batch_size=1024, num_epochs=10,
                         shuffle=False, drop_last=True, metrics_name=["Accuracy"])
    est.fit(ds, ds)
    train = [h for h in est.results.metrics_history if "train_loss" in h]
    assert train[0]["engine"] == "fused-dlrm", train[0]
    assert train[-1]["train_loss"] < train[0]["train_loss"] - 0.002
    state = torch.load(os.path.join(est.results.checkpoint.path, "model.pt"), weights_only=True)
    assert set(state) == set(model.state_dict())            # the USER's parameter names
    trained = est.get_model()
    trained.eval()
    inputs = [torch.tensor(cols[f"_c{i}"]).unsqueeze(1) for i in range(1, 40)]
    with torch.no_grad():
        acc = ((trained(*inputs)[:, 0] > 0).float() == torch.tensor(cols["_c0"]).float()).float().mean().item()
    evals = [h for h in est.results.metrics_history if "eval_res" in h]
    assert abs(acc - evals[-1]["eval_res"]["Accuracy"]) < 0.02, (acc, evals[-1]["eval_res"])


def test_per_column_dtypes_feed_a_generic_module(ray_cluster):
    """A list of feature dtypes -> model(*columns) (reference docstring torch/estimator.py:115-121) with
    float columns from the packed matrix and integer columns from the index matrix, all on the GPU loader."""
    from raydp_b200 import data as rdata
    from raydp_b200.torch import TorchEstimator

    class Wide(nn.Module):
        def __init__(self):
            super().__init__()
            self.emb_a, self.emb_b = nn.Embedding(11, 8), nn.Embedding(7, 8)
            self.mlp = nn.Sequential(nn.Linear(3 + 16, 64), nn.ReLU(), nn.Linear(64, 1))

        def forward(self, x0, a, x1, b, x2):
            h = torch.cat([x0, x1, x2, self.emb_a(a[:, 0]), self.emb_b(b[:, 0])], dim=1)
            return self.mlp(h)
    rng = np.random.default_rng(1)
    n = 8192
    cols = {"x0": rng.standard_normal(n).astype(np.float32), "a": rng.integers(0, 11, n).astype(np.int64),
            "x1": rng.standard_normal(n), "b": rng.integers(0, 7, n).astype(np.int32),
            "x2": rng.standard_normal(n).astype(np.float32)}
    cols["y"] = (cols["x0"] - 2 * cols["x1"] + (cols["a"] % 3) + 0.5 * cols["b"]).astype(np.float32)
    ds = rdata.from_arrow([pa.table(cols)])
    torch.manual_seed(2)
    model = Wide()
    est = TorchEstimator(num_workers=1, resources_per_worker={"GPU": 1}, model=model,
                         optimizer=torch.optim.Adam(model.parameters(), lr=1e-2), loss=nn.MSELoss(),
                         feature_columns=["x0", "a", "x1", "b", "x2"],
                         feature_types=[torch.float, torch.long, torch.float, torch.long, torch.float],
                         label_column="y", label_type=torch.float, batch_size=512, num_epochs=6, shuffle=False)
    est.fit(ds, ds)
    hist = est.results.metrics_history
    train = [h for h in hist if "train_loss" in h]
    evals = [h for h in hist if "test_loss" in h]
    assert train[0]["engine"] == "fused-generic"
    assert train[-1]["train_loss"] < 0.3 * train[0]["train_loss"]
    trained = est.get_model().eval()
    with torch.no_grad():
        out = trained(torch.tensor(cols["x0"]).unsqueeze(1), torch.tensor(cols["a"]).unsqueeze(1),
                      torch.tensor(cols["x1"], dtype=torch.float32).unsqueeze(1),
                      torch.tensor(cols["b"]).long().unsqueeze(1), torch.tensor(cols["x2"]).unsqueeze(1))
        loss = nn.MSELoss()(out[:, 0], torch.tensor(cols["y"])).item()
    assert abs(loss - evals[-1]["test_loss"]) < 0.1 * evals[-1]["test_loss"] + 0.02, (loss, evals[-1]["test_loss"])
