"""``TFEstimator``: Keras training from Spark data (import-guarded).

Parity: reference ``python/raydp/tf/estimator.py:41-310`` -- ctor ``(num_workers, resources_per_worker,
model, optimizer, loss, metrics, feature_columns, label_columns, merge_feature_columns, batch_size,
drop_last, num_epochs, shuffle, callbacks, **extra_config)`` serialises the Keras model / optimizer /
loss (``:93-149``); ``train_func`` builds them under ``MultiWorkerMirroredStrategy`` and feeds
``dataset_shard.to_tf`` (``:171-210``); rank 0 reports a SavedModel checkpoint; ``fit``,
``fit_on_spark(train_df, evaluate_df, fs_directory, compression, max_retries,
stop_spark_after_conversion)`` and ``get_model`` (``:212-310``).

TensorFlow is not part of this image (the framework targets PyTorch + hand-written CUDA on B200),
so the estimator is complete but untested here: construction raises a clear ImportError when
``tensorflow`` cannot be imported.  With TensorFlow installed, each worker process (one per
``num_workers``) gets a ``TF_CONFIG`` cluster spec and trains under the multi-worker strategy.
"""
from __future__ import annotations

import json
import os
import tempfile
from typing import Any, Dict, List, Optional, Union

import numpy as np

from raydp_b200 import runtime
from raydp_b200.estimator import EstimatorInterface
from raydp_b200.spark.interfaces import SparkEstimatorInterface
from raydp_b200.train.checkpoint import Checkpoint
from raydp_b200.utils import find_free_port


def _require_tf():
    try:
        import tensorflow as tf  # type: ignore
        return tf
    except ImportError as exc:
        raise ImportError("raydp_b200.tf.TFEstimator needs the 'tensorflow' package, which is not "
                          "installed in this environment; use raydp_b200.torch.TorchEstimator") from exc


def _tf_worker(rank: int, world: int, ports: List[int], config: Dict[str, Any], shard, eval_shard, out_dir: str):
    tf = _require_tf()
    os.environ["TF_CONFIG"] = json.dumps({"cluster": {"worker": [f"127.0.0.1:{p}" for p in ports]},
                                          "task": {"type": "worker", "index": rank}})
    strategy = tf.distribute.MultiWorkerMirroredStrategy() if world > 1 else tf.distribute.get_strategy()
    with strategy.scope():
        model = tf.keras.models.model_from_json(config["model"])
        optimizer = tf.keras.optimizers.get(config["optimizer"])
        loss = tf.keras.losses.get(config["loss"])
        model.compile(optimizer=optimizer, loss=loss, metrics=config["metrics"])

    def to_tf(ds):
        feats, labels = config["feature_columns"], config["label_columns"]
        table = ds.to_arrow()
        x = np.column_stack([table.column(c).to_numpy(zero_copy_only=False) for c in feats]).astype(np.float32)
        y = np.column_stack([table.column(c).to_numpy(zero_copy_only=False) for c in labels]).astype(np.float32)
        if not config["merge_feature_columns"]:
            x = tuple(x[:, i:i + 1] for i in range(x.shape[1]))
        out = tf.data.Dataset.from_tensor_slices((x, y))
        return out.batch(config["batch_size"], drop_remainder=config["drop_last"])

    history = []
    train = to_tf(shard)
    evaluate = to_tf(eval_shard) if eval_shard is not None else None
    for epoch in range(config["num_epochs"]):
        hist = model.fit(train, epochs=1, verbose=0, callbacks=config.get("callbacks"))
        entry = {"epoch": epoch, "train": {k: float(v[-1]) for k, v in hist.history.items()}}
        if evaluate is not None:
            entry["test"] = [float(v) for v in np.atleast_1d(model.evaluate(evaluate, verbose=0))]
        history.append(entry)
    model.save(os.path.join(out_dir, f"rank{rank}") if rank else out_dir)   # every rank must call save
    return history if rank == 0 else None


class TFEstimator(EstimatorInterface, SparkEstimatorInterface):
    def __init__(self, num_workers: int = 1, resources_per_worker: Optional[Dict[str, float]] = None,
                 model=None, optimizer=None, loss=None, metrics: Optional[List] = None,
                 feature_columns: Union[str, List[str]] = None, label_columns: Union[str, List[str]] = None,
                 merge_feature_columns: bool = True, batch_size: int = 128, drop_last: bool = False,
                 num_epochs: int = 1, shuffle: bool = True, callbacks: Optional[List] = None, **extra_config):
        tf = _require_tf()
        self._num_workers = num_workers
        self._resources_per_worker = resources_per_worker
        self._serialized_model = model.to_json() if hasattr(model, "to_json") else model
        self._serialized_optimizer = (tf.keras.optimizers.serialize(optimizer)
                                      if isinstance(optimizer, tf.keras.optimizers.Optimizer) else optimizer)
        self._serialized_loss = (tf.keras.losses.serialize(loss)
                                 if isinstance(loss, tf.keras.losses.Loss) else loss)
        self._metrics = [m if isinstance(m, str) else tf.keras.metrics.serialize(m) for m in (metrics or [])]
        self._feature_columns = [feature_columns] if isinstance(feature_columns, str) else list(feature_columns)
        self._label_columns = [label_columns] if isinstance(label_columns, str) else list(label_columns)
        self._merge_feature_columns = merge_feature_columns
        self._batch_size = batch_size
        self._drop_last = drop_last
        self._num_epochs = num_epochs
        self._shuffle = shuffle
        self._callbacks = callbacks
        self._extra_config = extra_config
        self._checkpoint: Optional[Checkpoint] = None
        self._results = None

    def fit(self, train_ds, evaluate_ds=None, max_retries: int = 0):
        if self._shuffle:
            train_ds = train_ds.random_shuffle()
            if evaluate_ds is not None:
                evaluate_ds = evaluate_ds.random_shuffle()
        config = {"model": self._serialized_model, "optimizer": self._serialized_optimizer,
                  "loss": self._serialized_loss, "metrics": self._metrics,
                  "feature_columns": self._feature_columns, "label_columns": self._label_columns,
                  "merge_feature_columns": self._merge_feature_columns, "batch_size": self._batch_size,
                  "drop_last": self._drop_last, "num_epochs": self._num_epochs, "callbacks": self._callbacks}
        n = self._num_workers
        shards = train_ds.split(n, equal=True) if n > 1 else [train_ds]
        eval_shards = (evaluate_ds.split(n, equal=True) if n > 1 else [evaluate_ds]) if evaluate_ds is not None \
            else [None] * n
        ports = [find_free_port() for _ in range(n)]
        out_dir = tempfile.mkdtemp(prefix="raydp_b200_tf_")
        res = dict(self._resources_per_worker or {})
        task = runtime.remote(_tf_worker).options(num_cpus=res.get("CPU", 1), num_gpus=res.get("GPU"),
                                                  max_retries=max_retries)
        refs = [task.remote(r, n, ports, config, shards[r], eval_shards[r], out_dir) for r in range(n)]
        self._results = runtime.get(refs)[0]
        self._checkpoint = Checkpoint.from_directory(out_dir)
        return self._results

    def fit_on_spark(self, train_df, evaluate_df=None, fs_directory: Optional[str] = None,
                     compression: Optional[str] = None, max_retries: int = 3,
                     stop_spark_after_conversion: bool = False):
        from raydp_b200.context import stop_spark
        from raydp_b200.spark.dataset import read_spark_parquet, spark_dataframe_to_ray_dataset
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
            train_ds = spark_dataframe_to_ray_dataset(train_df)
            if evaluate_df is not None:
                evaluate_ds = spark_dataframe_to_ray_dataset(evaluate_df)
        if stop_spark_after_conversion:
            stop_spark(cleanup_data=False)
        return self.fit(train_ds, evaluate_ds, max_retries)

    def get_model(self):
        assert self._checkpoint is not None, "Must call fit first"
        tf = _require_tf()
        return tf.keras.models.load_model(self._checkpoint.path)
