"""Lowering of plain user modules onto the fused engines (raydp_b200/torch/lowering.py): what is
recognised, what is refused, and that the recognised restatement equals the module (CPU only)."""
import ast
import json
import os

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from raydp_b200.torch.lowering import dlrm_restatement, lower_module

from user_models import Dlrm, GeluNet, NYC_Model

REF = "/root/reference/examples"


def test_nyc_model_is_lowered_with_the_users_names():
    plan, why = lower_module(NYC_Model(20))
    assert plan is not None, why
    assert plan.kind == "mlp" and plan.how == "fx"
    spec = plan.spec
    assert spec.in_features == 20 and list(spec.hidden) == [256, 128, 64, 16] and spec.batch_norm
    assert list(spec.linear_names) == ["fc1", "fc2", "fc3", "fc4", "fc5"]
    assert list(spec.bn_names) == ["bn1", "bn2", "bn3", "bn4"]


def test_sequential_and_method_relu_variants():
    seq = nn.Sequential(nn.Linear(12, 64), nn.ReLU(), nn.BatchNorm1d(64), nn.Dropout(0.0),
                        nn.Linear(64, 32), nn.ReLU(), nn.BatchNorm1d(32), nn.Linear(32, 1))
    plan, why = lower_module(seq)
    assert plan is not None, why
    assert list(plan.spec.linear_names) == ["0", "4", "7"] and list(plan.spec.bn_names) == ["2", "6"]

    class M(nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.b = nn.Linear(8, 32), nn.Linear(32, 1)

        def forward(self, x):
            return self.b(self.a(x).relu().float())
    plan, why = lower_module(M())
    assert plan is not None and not plan.spec.batch_norm and list(plan.spec.hidden) == [32], why


@pytest.mark.parametrize("model, fragment", [
    (GeluNet(20), "gelu"),
    (nn.Sequential(nn.Linear(8, 32), nn.BatchNorm1d(32), nn.ReLU(), nn.Linear(32, 1)), "followed by ReLU"),
    (nn.Sequential(nn.Linear(8, 32), nn.ReLU(), nn.Dropout(0.2), nn.Linear(32, 1)), "Dropout"),
    (nn.Sequential(nn.Linear(8, 32), nn.ReLU(), nn.Linear(32, 3)), "one output"),
    (nn.Sequential(nn.Linear(8, 40), nn.ReLU(), nn.Linear(40, 1)), "last hidden width"),
    (nn.Sequential(nn.Linear(8, 32, bias=False), nn.ReLU(), nn.Linear(32, 1)), "bias=False"),
    (nn.Sequential(nn.Linear(8, 32), nn.ReLU(), nn.BatchNorm1d(32, affine=False), nn.Linear(32, 1)), "affine"),
])
def test_unsupported_shapes_are_refused_with_a_reason(model, fragment):
    plan, why = lower_module(model)
    assert plan is None and fragment.lower() in why.lower(), why


def test_data_dependent_control_flow_is_refused():
    class Branchy(nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.b = nn.Linear(8, 16), nn.Linear(16, 1)

        def forward(self, x):
            h = F.relu(self.a(x))
            if h.sum() > 0:
                h = h * 2
            return self.b(h)
    plan, why = lower_module(Branchy())
    assert plan is None and "trace" in why


def test_shared_layer_is_refused():
    class Twice(nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.b = nn.Linear(16, 16), nn.Linear(16, 1)

        def forward(self, x):
            return self.b(F.relu(self.a(F.relu(self.a(x)))))
    plan, why = lower_module(Twice())
    assert plan is None and "twice" in why


def _dlrm(sizes=None):
    torch.manual_seed(0)
    sizes = sizes or [40 + 3 * t for t in range(26)]
    return Dlrm(13, sizes, [512, 128, 32], [1024, 1024, 512, 256, 1]), sizes


def test_notebook_style_dlrm_is_lowered_and_proved():
    model, sizes = _dlrm()
    plan, why = lower_module(model)
    assert plan is not None, why
    assert plan.kind == "dlrm" and plan.per_column_inputs
    assert plan.spec.num_dense == 13 and tuple(plan.spec.table_sizes) == tuple(sizes)
    assert plan.spec.bottom == (512, 128, 32) and plan.spec.top == (1024, 1024, 512, 256, 1)
    assert plan.bottom_names[0] == "bottom_model.mlp.layers.0" and plan.top_names[-1] == "top_model.out"
    assert plan.embedding_names[25] == "bottom_model.embeddings.embeddings.25"
    # the restatement is the module's function, on inputs the probe never saw
    bottom, top, embs = plan.modules(model)
    dense = torch.randn(7, 13)
    sparse = torch.stack([torch.randint(0, n, (7,)) for n in sizes], 1)
    cols = [dense[:, i:i + 1] for i in range(13)] + [sparse[:, t:t + 1] for t in range(26)]
    with torch.no_grad():
        assert torch.allclose(model(*cols), dlrm_restatement(bottom, top, embs, dense, sparse), atol=1e-5)
    # engine-side names map back onto the user's parameter names
    state = {f"bottom.{i}.{k}": torch.zeros(1) for i in range(3) for k in ("weight", "bias")}
    state.update({f"top.{i}.{k}": torch.zeros(1) for i in range(5) for k in ("weight", "bias")})
    state.update({f"embeddings.{t}.weight": torch.zeros(1) for t in range(26)})
    assert set(plan.rename_state(state)) == set(model.state_dict())


def test_dlrm_lookalike_with_other_math_fails_the_probe():
    model, _ = _dlrm()

    class Scaled(Dlrm):
        def forward(self, *inputs):
            return super().forward(*inputs) * 0.5
    other = Scaled(13, [40 + 3 * t for t in range(26)], [512, 128, 32], [1024, 1024, 512, 256, 1])
    plan, why = lower_module(other)
    assert plan is None and "probe" in why


def test_self_describing_models_keep_their_spec():
    from raydp_b200.models import DLRM, NYCTaxiMLP
    plan, _ = lower_module(NYCTaxiMLP(20))
    assert plan.kind == "mlp" and plan.how == "fused_spec"
    plan, _ = lower_module(DLRM(13, [10] * 26))
    assert plan.kind == "dlrm" and plan.embedding_names[3] == "embeddings.3"


@pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not present")
def test_the_references_own_classes_are_lowered():
    tree = ast.parse(open(os.path.join(REF, "pytorch_nyctaxi.py")).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "NYC_Model")
    ns = {"nn": nn, "F": F, "torch": torch}
    exec(compile(ast.Module([cls], []), "reference", "exec"), ns)
    plan, why = lower_module(ns["NYC_Model"](20))
    assert plan is not None and plan.kind == "mlp", why
    nb = json.load(open(os.path.join(REF, "pytorch_dlrm.ipynb")))
    cell = next("".join(c["source"]) for c in nb["cells"]
                if c["cell_type"] == "code" and "class Dlrm(" in "".join(c["source"]))
    ns2 = {}
    exec(cell, ns2)
    model = ns2["Dlrm"](13, [30 + t for t in range(26)], [512, 128, 32], [1024, 1024, 512, 256, 1])
    plan, why = lower_module(model)
    assert plan is not None and plan.kind == "dlrm" and plan.per_column_inputs, why


def test_select_fused_checks_loss_optimizer_and_columns():
    from raydp_b200.torch import fused_engine as fe
    model = NYC_Model(20)
    cfg = {"feature_columns": [f"f{i}" for i in range(20)], "feature_types": torch.float}
    adam = torch.optim.Adam(model.parameters())
    orig = fe.ops.kernels_available
    fe.ops.kernels_available = lambda: True
    try:
        assert fe.select_fused(cfg, model, adam, nn.SmoothL1Loss())[0] is not None
        assert "loss" in fe.select_fused(cfg, model, adam, nn.L1Loss())[1]
        assert "optimizer" in fe.select_fused(cfg, model, torch.optim.RMSprop(model.parameters()), nn.MSELoss())[1]
        assert "per-column" in fe.select_fused(dict(cfg, feature_types=[torch.float] * 20), model, adam, nn.MSELoss())[1]
    finally:
        fe.ops.kernels_available = orig
