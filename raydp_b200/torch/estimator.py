"""``TorchEstimator``: scikit-learn-style distributed PyTorch training from Spark data.

Parity: reference ``python/raydp/torch/estimator.py:41-396`` -- same constructor arguments
(``:76-97``), ``fit(train_ds, evaluate_ds=None, max_retries=3)`` (``:312-356``),
``fit_on_spark(train_df, evaluate_df=None, max_retries=3, fs_directory=None, compression=None,
stop_spark_after_conversion=False)`` (``:358-390``), ``get_model()`` (``:392-396``) and the
training-function semantics of ``:177-310``: model / optimizer / loss given as instances or
creator functions, one global ``random_shuffle`` before training, per-epoch reports with keys
``train_res``/``train_loss`` and ``eval_res``/``test_loss``, loss = mean of per-batch means, a
single end-of-training checkpoint ``model.pt`` written by world rank 0.

Differences, on purpose: a loss given as a *class* is instantiated;
workers are placed on GPUs when ``resources_per_worker`` has ``GPU`` (or ``use_gpu=True`` is
passed -- the reference silently drops that kwarg); on GPU workers the step is driven by the
fused engine (``raydp_b200.torch.engine``) whose kernels replace the per-step host sync,
DDP/NCCL all-reduce and optimizer of the stock loop.
"""
from __future__ import annotations

import inspect
import os
import tempfile
from typing import Any, Callable, Dict, List, Optional, Union

import torch

from raydp_b200.estimator import EstimatorInterface
from raydp_b200.spark.interfaces import DF, OPTIONAL_DF, SparkEstimatorInterface
from raydp_b200.torch.config import TorchConfig
from raydp_b200.torch.torch_metrics import TorchMetric
from raydp_b200.train import (Checkpoint, FailureConfig, RunConfig, ScalingConfig, TorchCheckpoint,
                              TorchTrainer)
from raydp_b200.train import session as _session


class TorchEstimator(EstimatorInterface, SparkEstimatorInterface):
    """Train a torch model data-parallel on ``num_workers`` processes.

    ``model`` is an ``nn.Module`` or ``fn(config) -> nn.Module``; ``optimizer`` an optimizer
    instance or ``fn(model, config) -> optimizer``; ``loss`` a loss instance, class or
    ``fn(config) -> loss``; ``lr_scheduler_creator`` is ``fn(optimizer, config)``.  With one
    ``feature_types`` dtype the model receives a single ``[B, F]`` tensor; with a list of
    dtypes it receives one ``[B, 1]`` tensor per feature column (``model(*inputs)``).
    """

    def __init__(self,
                 num_workers: int = 1,
                 resources_per_worker: Optional[Dict[str, float]] = None,
                 model: Union[torch.nn.Module, Callable] = None,
                 optimizer: Union[torch.optim.Optimizer, Callable] = None,
                 loss: Union[torch.nn.modules.loss._Loss, Callable] = None,
                 lr_scheduler_creator: Optional[Callable] = None,
                 scheduler_step_freq="batch",
                 feature_columns: List[str] = None,
                 feature_types: Optional[Union[torch.dtype, List[torch.dtype]]] = None,
                 label_column: str = None,
                 label_type: Optional[torch.dtype] = None,
                 batch_size: int = None,
                 drop_last: bool = False,
                 num_epochs: int = None,
                 shuffle: bool = True,
                 num_processes_for_data_loader: int = 0,
                 placement_strategy: str = "PACK",
                 metrics_name: Optional[List[Union[str, Callable]]] = None,
                 metrics_config: Optional[Dict[str, Dict[str, Any]]] = None,
                 use_ccl: bool = False,
                 **extra_config):
        self._num_workers = num_workers
        self._resources_per_worker = resources_per_worker
        self._model = model
        self._optimizer = optimizer
        self._loss = loss
        self._lr_scheduler_creator = lr_scheduler_creator
        self._feature_columns = feature_columns
        self._feature_types = feature_types
        self._label_column = label_column
        self._label_type = label_type
        self._batch_size = batch_size
        self._drop_last = drop_last
        self._num_epochs = num_epochs
        self._shuffle = shuffle
        self._placement_strategy = placement_strategy
        self._metrics = TorchMetric(metrics_name, metrics_config)
        self._use_ccl = use_ccl
        self._extra_config = extra_config
        # extras understood by this implementation (the reference stores and ignores them)
        self._use_gpu = extra_config.get("use_gpu")
        self._engine = extra_config.get("engine", "auto")          # auto | fused | stock
        self._compute_dtype = extra_config.get("compute_dtype")     # e.g. torch.bfloat16
        if self._num_epochs is None or self._batch_size is None:
            raise ValueError("batch_size and num_epochs are required")
        if num_processes_for_data_loader > 0:
            raise TypeError("multiple processes for data loader has not supported")
        self._trainer: Optional[TorchTrainer] = None
        self._trained_results = None

    # ------------------------------------------------------------------ train function
    @staticmethod
    def _build_model(spec, config):
        if isinstance(spec, torch.nn.Module):
            return spec
        if callable(spec):
            return spec(config)
        raise Exception("Unsupported parameter, we only support torch.nn.Model instance "
                        "or a function(dict -> model)")

    @staticmethod
    def _build_optimizer(spec, model, config):
        if isinstance(spec, torch.optim.Optimizer):
            # Rebuild the same optimizer class around *this* worker's parameters and carry the
            # hyper-parameters / state over (the pickled instance points at stale tensors).
            saved = spec.state_dict()
            fresh = spec.__class__(model.parameters(), lr=0.1)
            fresh.load_state_dict(saved)
            return fresh
        if callable(spec):
            return spec(model, config)
        raise Exception("Unsupported parameter, we only support torch.optim.Optimizer subclass "
                        "instance or a function((models, dict) -> optimizer)")

    @staticmethod
    def _build_loss(spec, config):
        base = torch.nn.modules.loss._Loss
        if inspect.isclass(spec) and issubclass(spec, base):
            return spec()
        if isinstance(spec, base):
            return spec
        if callable(spec):
            return spec(config)
        raise Exception("Unsupported parameter, we only support torch.nn.modules.loss._Loss "
                        "subclass, subclass instance or a function(dict -> loss)")

    @staticmethod
    def train_func(config: Dict[str, Any]) -> None:
        from raydp_b200.torch.engine import make_engine
        sess = _session.get_session()
        model = TorchEstimator._build_model(config["model"], config)
        optimizer = TorchEstimator._build_optimizer(config["optimizer"], model, config)
        criterion = TorchEstimator._build_loss(config["loss"], config)
        scheduler = (config["lr_scheduler_creator"](optimizer, config)
                     if config["lr_scheduler_creator"] else None)
        engine = make_engine(config, sess, model, optimizer, criterion, scheduler, config["metrics"])
        train_shard = _session.get_dataset_shard("train")
        eval_shard = _session.get_dataset_shard("evaluate") if config["evaluate"] else None
        for epoch in range(config["num_epochs"]):
            train_res, train_loss, timing = engine.train_epoch(train_shard)
            _session.report(dict(epoch=epoch, train_res=train_res, train_loss=train_loss, **timing))
            if eval_shard is not None:
                eval_res, test_loss = engine.evaluate_epoch(eval_shard)
                _session.report(dict(epoch=epoch, eval_res=eval_res, test_loss=test_loss))
        states = engine.state_dict()
        with tempfile.TemporaryDirectory() as tmp:
            checkpoint = None
            if sess.world_rank == 0:      # replicas are identical under data parallelism
                torch.save(states, os.path.join(tmp, "model.pt"))
                checkpoint = Checkpoint.from_directory(tmp)
            _session.report({}, checkpoint=checkpoint)
        engine.close()

    # ------------------------------------------------------------------ fit
    def fit(self, train_ds, evaluate_ds=None, max_retries: int = 3):
        config = {
            "model": self._model, "optimizer": self._optimizer, "loss": self._loss,
            "lr_scheduler_creator": self._lr_scheduler_creator,
            "feature_columns": self._feature_columns, "feature_types": self._feature_types,
            "label_column": self._label_column, "label_type": self._label_type,
            "batch_size": self._batch_size, "num_epochs": self._num_epochs,
            "drop_last": self._drop_last, "evaluate": evaluate_ds is not None,
            "metrics": self._metrics, "engine": self._engine, "compute_dtype": self._compute_dtype,
            "extra": {k: v for k, v in self._extra_config.items()
                      if k not in ("use_gpu", "engine", "compute_dtype")},
        }
        scaling = ScalingConfig(num_workers=self._num_workers, use_gpu=self._use_gpu,
                                resources_per_worker=self._resources_per_worker,
                                placement_strategy=self._placement_strategy)
        run_config = RunConfig(failure_config=FailureConfig(max_failures=max_retries))
        if self._shuffle:
            seed = self._extra_config.get("shuffle_seed")
            train_ds = train_ds.random_shuffle(seed)
            if evaluate_ds is not None:
                evaluate_ds = evaluate_ds.random_shuffle(seed)
        datasets = {"train": train_ds}
        if evaluate_ds is not None:
            datasets["evaluate"] = evaluate_ds
        torch_config = TorchConfig(backend="ccl") if self._use_ccl else TorchConfig()
        self._trainer = TorchTrainer(TorchEstimator.train_func, train_loop_config=config,
                                     scaling_config=scaling, run_config=run_config,
                                     torch_config=torch_config, datasets=datasets)
        if self._extra_config.get("spmd_local_shards"):
            self._trainer._local_shards = True
        self._trained_results = self._trainer.fit()
        return self._trained_results

    def fit_on_spark(self, train_df: DF, evaluate_df: OPTIONAL_DF = None, max_retries: int = 3,
                     fs_directory: Optional[str] = None, compression: Optional[str] = None,
                     stop_spark_after_conversion: bool = False):
        from raydp_b200.context import stop_spark
        from raydp_b200.spark.dataset import (get_raydp_master_owner, read_spark_parquet,
                                              spark_dataframe_to_ray_dataset)
        super().fit_on_spark(train_df, evaluate_df)
        train_df = self._check_and_convert(train_df)
        evaluate_ds = None
        if evaluate_df is not None:
            evaluate_df = self._check_and_convert(evaluate_df)
        if fs_directory is not None:
            app_id = train_df.sql_ctx.sparkSession.sparkContext.applicationId
            path = fs_directory.rstrip("/") + f"/{app_id}"
            train_df.write.parquet(path + "/train", compression=compression)
            train_ds = read_spark_parquet(path + "/train")
            if evaluate_df is not None:
                evaluate_df.write.parquet(path + "/test", compression=compression)
                evaluate_ds = read_spark_parquet(path + "/test")
        else:
            owner = None
            if stop_spark_after_conversion:
                owner = get_raydp_master_owner(train_df.sql_ctx.sparkSession)
            train_ds = spark_dataframe_to_ray_dataset(train_df, owner=None)
            if evaluate_df is not None:
                evaluate_ds = spark_dataframe_to_ray_dataset(evaluate_df, owner=None)
            del owner
        if stop_spark_after_conversion:
            stop_spark(cleanup_data=False)
        return self.fit(train_ds, evaluate_ds, max_retries)

    def get_model(self):
        assert self._trainer is not None, "Must call fit first"
        if self._trained_results.checkpoint is None:
            raise RuntimeError("this rank holds no checkpoint (only world rank 0 writes model.pt)")
        model = self._model if isinstance(self._model, torch.nn.Module) else None
        if model is None:
            model = self._model({"extra": self._extra_config})
        return TorchCheckpoint(self._trained_results.checkpoint.to_directory()).get_model(model)

    @property
    def results(self):
        return self._trained_results
