"""Hyper-parameter search (the ``ray.tune`` slice) and the first-generation ``Trainer`` API."""
import pytest

from raydp_b200 import tune
from raydp_b200.train import Trainer, TrainingCallback
from raydp_b200.train import session


def _trainable(config):
    for i in range(3):
        tune.report(loss=(config["lr"] - 0.01) ** 2 + 1.0 / (i + 1), it=i)


def test_search_space_expansion():
    space = {"a": tune.grid_search([1, 2]), "b": {"c": tune.grid_search(["x", "y"]), "d": tune.uniform(0, 1)}, "e": 5}
    configs = tune.generate_configs(space, num_samples=2, seed=0)
    assert len(configs) == 8 and {(c["a"], c["b"]["c"]) for c in configs} == {(1, "x"), (1, "y"), (2, "x"), (2, "y")}
    assert all(0 <= c["b"]["d"] <= 1 and c["e"] == 5 for c in configs)
    assert len({c["b"]["d"] for c in configs}) == 8                      # every trial draws its own sample
    assert tune.generate_configs(space, seed=0) == tune.generate_configs(space, seed=0)
    draws = tune.generate_configs({"n": tune.randint(0, 3), "c": tune.choice(["p", "q"]), "l": tune.loguniform(1e-4, 1e-1)},
                                  num_samples=20, seed=1)
    assert all(d["n"] in (0, 1, 2) and d["c"] in ("p", "q") and 1e-4 <= d["l"] <= 1e-1 for d in draws)


def test_run_in_process_and_best_config():
    analysis = tune.run(_trainable, config={"lr": tune.grid_search([0.005, 0.01, 0.05])}, metric="loss", mode="min")
    assert analysis.get_best_config() == {"lr": 0.01} and abs(analysis.best_result["loss"] - 1 / 3) < 1e-12
    assert analysis.get_best_config("loss", "max") == {"lr": 0.05}
    frame = analysis.dataframe()
    assert frame.shape[0] == 3 and "config/lr" in frame.columns and list(frame["it"]) == [2, 2, 2]
    with pytest.raises(tune.TuneError):
        tune.run(lambda config: 1 / 0, config={})


def test_tuner_runs_trials_as_runtime_tasks(ray_cluster):
    grid = tune.Tuner(_trainable, param_space={"lr": tune.grid_search([0.01, 0.1])},
                      tune_config=tune.TuneConfig(metric="loss", mode="min")).fit()
    best = grid.get_best_result()
    assert len(grid) == 2 and best.config == {"lr": 0.01} and abs(best.metrics["loss"] - 1 / 3) < 1e-12
    assert grid.num_errors == 0 and len(best.metrics_history) == 3
    failed = tune.Tuner(lambda config: 1 / 0, param_space={}).fit()
    assert failed.num_errors == 1 and "ZeroDivisionError" in str(failed.errors[0])


def test_legacy_trainer_run_callbacks_and_tune_trainable(ray_cluster):
    seen = []

    class Collect(TrainingCallback):
        def handle_result(self, results, **info):
            seen.append(results)

    def train_func(config):
        rank = session.get_context().get_world_rank()
        session.report(loss=config["x"] * (rank + 1))
        return rank

    trainer = Trainer(backend="torch", num_workers=2)
    with pytest.raises(RuntimeError):
        trainer.run(train_func, config={"x": 1.0})
    trainer.start()
    assert trainer.run(train_func, config={"x": 2.0}, callbacks=[Collect()]) == [0, 1]
    assert seen == [[{"loss": 2.0}, {"loss": 4.0}]]
    trainable = trainer.to_tune_trainable(train_func)
    trainer.shutdown()
    analysis = tune.run(trainable, config={"x": tune.grid_search([3.0, 1.0])}, metric="loss", mode="min",
                        max_concurrent_trials=1)
    assert analysis.get_best_config() == {"x": 1.0}


def test_trainer_resumes_from_last_reported_checkpoint(ray_cluster, tmp_path):
    """A worker failure mid-run restarts the group (``FailureConfig``); ``train.get_checkpoint()`` then
    returns the newest checkpoint reported before the failure, so the loop continues instead of restarting.
    ``resume_from_checkpoint`` seeds the first attempt the same way."""
    import os
    from raydp_b200.train import (Checkpoint, FailureConfig, RunConfig, ScalingConfig, TorchTrainer, get_checkpoint,
                                  report)
    flag = str(tmp_path / "crashed_once")

    def loop(config):
        start = 0
        ckpt = get_checkpoint()
        if ckpt is not None:
            start = ckpt.to_dict()["epoch"] + 1
        for epoch in range(start, 4):
            if epoch == 2 and not os.path.exists(config["flag"]):
                open(config["flag"], "w").close()
                raise RuntimeError("injected failure")
            report({"epoch": epoch, "started_at": start}, checkpoint=Checkpoint.from_dict({"epoch": epoch}))

    trainer = TorchTrainer(loop, train_loop_config={"flag": flag}, scaling_config=ScalingConfig(num_workers=1),
                           run_config=RunConfig(storage_path=str(tmp_path), name="resume",
                                                failure_config=FailureConfig(max_failures=1)))
    result = trainer.fit()
    assert result.metrics["epoch"] == 3 and result.metrics["started_at"] == 2          # resumed after epoch 1
    assert [m["epoch"] for m in result.metrics_history] == [2, 3]
    again = TorchTrainer(loop, train_loop_config={"flag": flag}, scaling_config=ScalingConfig(num_workers=1),
                         run_config=RunConfig(storage_path=str(tmp_path), name="seeded"),
                         resume_from_checkpoint=Checkpoint.from_dict({"epoch": 2})).fit()
    assert [m["epoch"] for m in again.metrics_history] == [3] and again.checkpoint.to_dict() == {"epoch": 3}


def test_failure_of_a_non_zero_rank_surfaces_at_once_and_the_group_restarts(ray_cluster, tmp_path):
    """Rank 1 raises while rank 0 is inside a collective that can no longer complete.  The driver must
    see the failure without waiting on rank 0 (which would block until the backend's timeout), tear the
    group down and let ``FailureConfig`` start the next attempt."""
    import os
    import time
    from raydp_b200.train import FailureConfig, RunConfig, ScalingConfig, TorchTrainer, get_context, report
    flag = str(tmp_path / "rank1_failed_once")

    def loop(config):
        import torch
        import torch.distributed as dist
        rank = get_context().get_world_rank()
        if rank == 1 and not os.path.exists(config["flag"]):
            open(config["flag"], "w").close()
            raise RuntimeError("boom on rank 1")
        t = torch.ones(1)
        dist.all_reduce(t)              # first attempt: rank 0 waits here for a peer that is gone
        report({"sum": float(t.item()), "rank": rank})

    t0 = time.time()
    result = TorchTrainer(loop, train_loop_config={"flag": flag}, scaling_config=ScalingConfig(num_workers=2),
                          run_config=RunConfig(storage_path=str(tmp_path), name="rank1",
                                               failure_config=FailureConfig(max_failures=1))).fit()
    assert result.metrics["sum"] == 2.0 and os.path.exists(flag)
    assert time.time() - t0 < 120, "the driver waited for the stuck rank instead of the failed one"


def test_a_rank_that_dies_without_raising_reports_the_tail_of_its_log(ray_cluster, tmp_path):
    """A worker process that disappears (here: ``os._exit``) leaves no traceback; the error carries the last
    lines the worker printed (per-worker log files, ``runtime.worker_logs``)."""
    import os
    import pytest
    from raydp_b200.train import RunConfig, ScalingConfig, TorchTrainer, get_context
    from raydp_b200.train.trainer import TrainingFailedError

    def loop(config):
        import torch
        import torch.distributed as dist
        if get_context().get_world_rank() == 1:
            print("about to crash: marker-7f3a", flush=True)
            os._exit(3)
        t = torch.ones(1)
        dist.all_reduce(t)

    with pytest.raises(TrainingFailedError) as err:
        TorchTrainer(loop, train_loop_config={}, scaling_config=ScalingConfig(num_workers=2),
                     run_config=RunConfig(storage_path=str(tmp_path), name="dies")).fit()
    assert "rank 1 failed" in str(err.value) and "marker-7f3a" in str(err.value), str(err.value)[-600:]


def test_horovod_surface_averages_gradients_over_the_group(ray_cluster):
    """``hvd.DistributedOptimizer``: replicas that start from rank 0's parameters and see different data end the
    step with identical parameters (gradients averaged in ``step()``)."""
    from raydp_b200 import train
    from raydp_b200.train import ScalingConfig, TorchTrainer

    def loop():
        import torch
        from raydp_b200.train import horovod_api as hvd
        hvd.init()
        torch.manual_seed(hvd.rank())                      # different initial weights per rank
        model = torch.nn.Linear(3, 1)
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        hvd.broadcast_optimizer_state(opt, root_rank=0)
        opt = hvd.DistributedOptimizer(opt, named_parameters=model.named_parameters(), op=hvd.Average)
        for step in range(3):
            x = torch.full((4, 3), float(hvd.rank() + 1 + step))
            opt.zero_grad()
            (model(x) ** 2).mean().backward()
            opt.step()
        w = model.weight.detach().reshape(-1)
        spread = hvd.allreduce(w, op=hvd.Average) - w
        train.report({"spread": float(spread.abs().max()), "size": hvd.size(), "w0": float(w[0])})

    res = TorchTrainer(loop, scaling_config=ScalingConfig(num_workers=2)).fit()
    assert res.metrics["size"] == 2 and res.metrics["spread"] < 1e-7
