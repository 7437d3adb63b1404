"""The tutorial script runs end to end on CPU (capability of the reference's tutorials/*.ipynb): CSV with string
categories and "N/A" values -> describe / rollup / UDF clean-up -> random_split / unionByName -> estimator with
per-column dtypes feeding ``model(*columns)``, and the same data through a hand-written TorchTrainer loop."""
import importlib.util
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load():
    spec = importlib.util.spec_from_file_location("stroke_prediction", os.path.join(ROOT, "tutorials", "stroke_prediction.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_tutorial_with_the_estimator(capsys):
    out = _load().main(["--rows", "1200", "--epochs", "3"])
    assert out["losses"][-1] < out["losses"][0] and out["train_rows"] > out["test_rows"] > 100
    model = out["model"]
    assert isinstance(model, torch.nn.Module) and len(model.tables) == 8
    text = capsys.readouterr().out
    assert "rows with a missing bmi:" in text and "|summary|" in text and "| Other|    1|" in text


def test_tutorial_with_a_trainer_loop():
    out = _load().main(["--rows", "1000", "--epochs", "2", "--trainer", "--workers", "2"])
    assert len(out["losses"]) == 2 and out["losses"][-1] < out["losses"][0]


def test_prepare_data_loader_shards_over_the_group(ray_cluster):
    from raydp_b200 import train
    from raydp_b200.train import ScalingConfig, TorchTrainer

    def loop():
        from torch.utils.data import DataLoader, TensorDataset
        ds = TensorDataset(torch.arange(40).float().view(-1, 1))
        loader = train.torch.prepare_data_loader(DataLoader(ds, batch_size=5))
        seen = torch.cat([b[0].view(-1) for b in loader]).tolist()
        train.report({"n": len(seen), "first": seen[0], "device": str(train.torch.get_device())})

    res = TorchTrainer(loop, scaling_config=ScalingConfig(num_workers=2))
    result = res.fit()
    assert result.metrics["n"] == 20 and result.metrics["device"] == "cpu"     # each rank sees half of the rows
