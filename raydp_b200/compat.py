"""Import-name compatibility for scripts written against the reference's stack.

Reference programs start with ``import ray`` / ``from pyspark.sql.functions import col`` (every file
under ``examples/`` does).  Neither package exists here -- the runtime, the DataFrame engine, the
dataset layer and the trainer are this framework's own -- so :func:`install` registers thin module
objects named ``ray`` and ``pyspark`` whose attributes point at the equivalents:

    ray.init / get / put / remote / get_actor / ...      -> raydp_b200.runtime
    ray.util.placement_group, ray.util.state.list_actors -> raydp_b200.runtime
    ray.data.from_spark / from_arrow_refs / Dataset      -> raydp_b200.data
    ray.train.torch.TorchTrainer, ray.train.ScalingConfig, ray.air.session, ...  -> raydp_b200.train
    ray.cluster_utils.Cluster                            -> raydp_b200.runtime.cluster_utils
    pyspark.sql.{SparkSession,DataFrame,functions,types,Window}, pyspark.pandas   -> raydp_b200.sql

A real installation always wins: a name is only registered when ``importlib`` cannot find the
genuine package (``force=True`` overrides).  ``import raydp`` calls ``install()``; ``import raydp_b200``
does not touch ``sys.modules``.
"""
from __future__ import annotations

import importlib
import importlib.machinery
import importlib.util
import sys
import types
from typing import Dict, List

_INSTALLED: List[str] = []


class _ExistingModuleLoader:
    """Hands the import system the module :func:`install` has just registered under this name."""

    def create_module(self, spec):
        return sys.modules[spec.name]

    def exec_module(self, module) -> None:
        pass


def _module(name: str, doc: str, **attrs) -> types.ModuleType:
    mod = types.ModuleType(name, doc)
    mod.__dict__.update(attrs)
    mod.__raydp_b200_alias__ = True
    # a spec keeps ``importlib.util.find_spec(name)`` (used by other libraries to probe for optional
    # dependencies) from raising on a module that sits in sys.modules without one
    # the loader matters when the module is registered WHILE an import of its name is in flight
    # (install_lazy): importlib then loads through sys.modules[name].__spec__
    mod.__spec__ = importlib.machinery.ModuleSpec(name, loader=_ExistingModuleLoader())
    mod.__path__ = []           # lets ``import <alias>.<child>`` resolve children registered in sys.modules
    return mod


def _register(tree: Dict[str, types.ModuleType]) -> None:
    """``tree`` maps dotted names to modules; children become attributes of their parents."""
    for name in sorted(tree):
        sys.modules[name] = tree[name]
        _INSTALLED.append(name)
        parent, _, leaf = name.rpartition(".")
        if parent and parent in tree:
            # ``ray.util.placement_group`` is a function AND the name of a sub-module: keep the function
            existing = getattr(tree[parent], leaf, None)
            if existing is None or isinstance(existing, types.ModuleType):
                setattr(tree[parent], leaf, tree[name])


def _absent(name: str) -> bool:
    if name in sys.modules:
        return bool(getattr(sys.modules[name], "__raydp_b200_alias__", False))
    try:
        return importlib.util.find_spec(name) is None
    except (ImportError, ValueError):
        return True


class JobConfig:
    """Stand-in for ``ray.job_config.JobConfig``: settings are recorded, ``code_search_path`` entries are
    put on ``sys.path`` (workers inherit the driver's path)."""

    def __init__(self, **settings):
        self.settings = settings
        for path in settings.get("code_search_path") or []:
            if path not in sys.path:
                sys.path.append(path)


def _passthrough_decorator(*args, **kwargs):
    """``@PublicAPI`` / ``@PublicAPI(stability="beta")``: annotations only."""
    if len(args) == 1 and callable(args[0]) and not kwargs:
        return args[0]
    return lambda obj: obj


def _global_gc() -> None:
    import gc
    gc.collect()


class _NodeView:
    """``ray.worker.global_worker.node``: what the reference's submit recipe reads to write ``ray.conf``
    (reference doc/spark_on_ray.md, "spark-submit" section)."""

    @staticmethod
    def _info() -> dict:
        from raydp_b200 import runtime
        if not runtime.is_initialized():
            raise RuntimeError("the runtime is not initialised: call ray.init() first")
        return runtime._info

    @property
    def node_ip_address(self) -> str:
        return self._info()["node_ip"]

    @property
    def address(self) -> str:
        return self._info()["address"]

    @property
    def unique_id(self) -> str:
        return self._info()["node_id"]

    def get_session_dir_path(self) -> str:
        return self._info()["dir"]

    def get_logs_dir_path(self) -> str:
        return self._info()["log_dir"]


class _GlobalWorker:
    node = _NodeView()

    @property
    def connected(self) -> bool:
        from raydp_b200 import runtime
        return runtime.is_initialized()

    @property
    def namespace(self):
        return _NodeView._info().get("namespace")


def _ray_tree() -> Dict[str, types.ModuleType]:
    from raydp_b200 import data, runtime, train, tune
    from raydp_b200.train import legacy
    from raydp_b200.runtime import util as rt_util
    from raydp_b200.spark.parallel_iterator_worker import ParallelIteratorWorkerWithLen
    from raydp_b200.torch.torch_ml_dataset import TorchMLDataset
    from raydp_b200.runtime import cluster_utils, protocol
    from raydp_b200.train import checkpoint, config, session, torch_utils, trainer

    top = {k: getattr(runtime, k) for k in (
        "init", "shutdown", "is_initialized", "get", "put", "wait", "remote", "get_actor", "kill", "nodes",
        "cluster_resources", "available_resources", "get_runtime_context", "ObjectRef")}
    def ray_init(*args, **kwargs):
        """``ray.init``: like :func:`raydp_b200.runtime.init`, with Ray's default of echoing worker output on the
        driver (scripts written against Ray print their training progress from the workers)."""
        import os
        if "log_to_driver" not in kwargs and "RAYDP_B200_LOG_TO_DRIVER" not in os.environ:
            kwargs["log_to_driver"] = True
        return runtime.init(*args, **kwargs)
    top["init"] = ray_init
    ray = _module("ray", "alias of raydp_b200.runtime (see raydp_b200.compat)", __version__="2.50.0", **top)
    ray.__path__ = []           # lets ``import ray.data`` style imports resolve through sys.modules

    class _Client:              # ``ray.util.client.ray.is_connected()`` (reference context.py:219)
        @staticmethod
        def is_connected() -> bool:
            return False

    tree = {
        "ray": ray,
        "ray.actor": _module("ray.actor", "", exit_actor=runtime.exit_actor, ActorHandle=runtime.ActorHandle),
        "ray.exceptions": _module("ray.exceptions", "", RayError=protocol.RuntimeErrorBase,
                                  RayTaskError=protocol.TaskError, RayActorError=protocol.ActorDiedError,
                                  GetTimeoutError=protocol.GetTimeoutError, ObjectLostError=protocol.ObjectLostError),
        "ray.util": _module("ray.util", "", placement_group=runtime.placement_group,
                            remove_placement_group=runtime.remove_placement_group,
                            placement_group_table=runtime.placement_group_table,
                            get_placement_group=runtime.get_placement_group,
                            get_node_ip_address=runtime.get_node_ip_address),
        "ray.util.placement_group": _module("ray.util.placement_group", "",
                                            placement_group=runtime.placement_group,
                                            remove_placement_group=runtime.remove_placement_group,
                                            placement_group_table=runtime.placement_group_table,
                                            get_placement_group=runtime.get_placement_group,
                                            PlacementGroup=runtime.PlacementGroup),
        "ray.util.scheduling_strategies": _module(
            "ray.util.scheduling_strategies", "",
            PlacementGroupSchedulingStrategy=runtime.PlacementGroupSchedulingStrategy),
        "ray.util.state": _module("ray.util.state", "", list_actors=runtime.list_actors),
        "ray.util.client": _module("ray.util.client", "", ray=_Client),
        "ray.cluster_utils": _module("ray.cluster_utils", "", Cluster=cluster_utils.Cluster),
        # a few private hooks that test code written against Ray pokes at
        "ray.job_config": _module("ray.job_config", "", JobConfig=JobConfig),
        "ray._private": _module("ray._private", ""),
        "ray.worker": _module("ray.worker", "", global_worker=_GlobalWorker()),
        "ray._private.worker": _module("ray._private.worker", "", global_worker=_GlobalWorker()),
        "ray._private.client_mode_hook": _module("ray._private.client_mode_hook", "",
                                                 client_mode_wrap=lambda fn: fn),
        "ray._private.internal_api": _module("ray._private.internal_api", "", global_gc=_global_gc),
        "ray.data": data,
        "ray.data.aggregate": _module("ray.data.aggregate", "", **{k: getattr(data, k) for k in (
            "AggregateFn", "Count", "Sum", "Min", "Max", "Mean", "Std", "AbsMax", "Unique", "Quantile")}),
        "ray.data.preprocessors": __import__("raydp_b200.data.preprocessors", fromlist=["x"]),
        "ray.data.preprocessor": _module("ray.data.preprocessor", "", Preprocessor=__import__(
            "raydp_b200.data.preprocessors", fromlist=["x"]).Preprocessor),
        "ray.data.context": _module("ray.data.context", "", DataContext=data.DataContext),
        "ray.train": _module("ray.train", "", ScalingConfig=config.ScalingConfig, RunConfig=config.RunConfig,
                             FailureConfig=config.FailureConfig, CheckpointConfig=config.CheckpointConfig,
                             Checkpoint=checkpoint.Checkpoint, get_dataset_shard=session.get_dataset_shard,
                             report=session.report, get_context=session.get_context,
                             get_checkpoint=session.get_checkpoint, Result=trainer.Result,
                             Trainer=legacy.Trainer, TrainingCallback=legacy.TrainingCallback,
                             world_rank=lambda: session.get_context().get_world_rank(),
                             local_rank=lambda: session.get_context().get_local_rank(),
                             world_size=lambda: session.get_context().get_world_size()),
        "ray.train.torch": _module("ray.train.torch", "", TorchTrainer=trainer.TorchTrainer,
                                   TorchConfig=config.TorchConfig, TorchCheckpoint=checkpoint.TorchCheckpoint,
                                   prepare_model=torch_utils.prepare_model, get_device=torch_utils.get_device,
                                   prepare_data_loader=torch_utils.prepare_data_loader),
        "ray.air": _module("ray.air", "", session=session, Checkpoint=checkpoint.Checkpoint,
                           ScalingConfig=config.ScalingConfig, RunConfig=config.RunConfig,
                           FailureConfig=config.FailureConfig, CheckpointConfig=config.CheckpointConfig),
        "ray.air.config": _module("ray.air.config", "", ScalingConfig=config.ScalingConfig,
                                  RunConfig=config.RunConfig, FailureConfig=config.FailureConfig,
                                  CheckpointConfig=config.CheckpointConfig),
        "ray.air.session": session,
        "ray.util.actor_pool": _module("ray.util.actor_pool", "", ActorPool=rt_util.ActorPool),
        "ray.util.queue": _module("ray.util.queue", "", Queue=rt_util.Queue, Empty=rt_util.Empty, Full=rt_util.Full),
        "ray.util.iter": _module("ray.util.iter", "", ParallelIteratorWorker=ParallelIteratorWorkerWithLen),
        "ray.util.data": _module("ray.util.data", "", MLDataset=TorchMLDataset),
        "ray.util.annotations": _module("ray.util.annotations", "", PublicAPI=_passthrough_decorator,
                                        DeveloperAPI=_passthrough_decorator, Deprecated=_passthrough_decorator),
        "ray.tune": tune,
        "ray.tune.search": _module("ray.tune.search", ""),
        "ray.tune.search.sample": _module("ray.tune.search.sample", "", Domain=tune.Domain),
        "ray.data.dataset": _module("ray.data.dataset", "", Dataset=data.Dataset),
        "ray.types": _module("ray.types", "", ObjectRef=runtime.ObjectRef),
        "ray.cloudpickle": __import__("cloudpickle"),
        "ray.runtime_context": _module("ray.runtime_context", "", get_runtime_context=runtime.get_runtime_context),
    }
    tree["ray.train"].torch = tree["ray.train.torch"]
    ray.tune, ray.cloudpickle = tune, tree["ray.cloudpickle"]
    # ray.data / ray.air.session are real modules of this package: register under the alias names too
    ray.data, ray.train = data, tree["ray.train"]
    del train
    return tree


def _pyspark_tree() -> Dict[str, types.ModuleType]:
    from raydp_b200 import sql
    from raydp_b200.sql import column, dataframe, functions, pandas_api, session, types as sql_types, window

    pyspark = _module("pyspark", "alias of raydp_b200.sql (see raydp_b200.compat)", __version__="3.5.0",
                      SparkConf=session.SparkConf, SparkContext=session.SparkContext,
                      StorageLevel=dataframe.StorageLevel, Row=sql_types.Row)
    pyspark.__path__ = []
    return {
        "pyspark": pyspark,
        "pyspark.sql": sql,
        "pyspark.sql.functions": functions,
        "pyspark.sql.types": sql_types,
        "pyspark.sql.window": window,
        "pyspark.sql.session": session,
        "pyspark.sql.dataframe": dataframe,
        "pyspark.sql.column": column,
        "pyspark.pandas": pandas_api,
        "pyspark.storagelevel": _module("pyspark.storagelevel", "", StorageLevel=dataframe.StorageLevel),
        **_pyspark_ml_tree(),
    }


def _pyspark_ml_tree() -> Dict[str, types.ModuleType]:
    from raydp_b200.sql import ml
    from raydp_b200.sql.ml import base, feature, linalg, models

    def pick(mod, *names):
        return {n: getattr(mod, n) for n in names}
    return {
        "pyspark.ml": _module("pyspark.ml", "alias of raydp_b200.sql.ml", **pick(
            base, "Pipeline", "PipelineModel", "Transformer", "Estimator", "Model")),
        "pyspark.ml.feature": feature,
        "pyspark.ml.linalg": linalg,
        "pyspark.ml.pipeline": _module("pyspark.ml.pipeline", "", **pick(base, "Pipeline", "PipelineModel")),
        "pyspark.ml.regression": _module("pyspark.ml.regression", "", **pick(
            models, "LinearRegression", "LinearRegressionModel")),
        "pyspark.ml.classification": _module("pyspark.ml.classification", "", **pick(
            models, "LogisticRegression", "LogisticRegressionModel")),
        "pyspark.ml.clustering": _module("pyspark.ml.clustering", "", **pick(models, "KMeans", "KMeansModel")),
        "pyspark.ml.evaluation": _module("pyspark.ml.evaluation", "", **pick(
            models, "Evaluator", "RegressionEvaluator", "BinaryClassificationEvaluator",
            "MulticlassClassificationEvaluator")),
        "pyspark.ml.tuning": _module("pyspark.ml.tuning", "", **pick(
            models, "ParamGridBuilder", "CrossValidator", "CrossValidatorModel", "TrainValidationSplit",
            "TrainValidationSplitModel")),
        # vectors are array<double> columns already
        "pyspark.ml.functions": _module("pyspark.ml.functions", "", vector_to_array=lambda col, dtype="float64": col,
                                        array_to_vector=lambda col: col),
    }


def install(ray: bool = True, pyspark: bool = True, force: bool = False) -> List[str]:
    """Register the ``ray`` / ``pyspark`` aliases that are missing; returns the top-level names added."""
    added = []
    if ray and (force or _absent("ray")) and not getattr(sys.modules.get("ray"), "__raydp_b200_alias__", False):
        _register(_ray_tree())
        added.append("ray")
    if pyspark and (force or _absent("pyspark")) and \
            not getattr(sys.modules.get("pyspark"), "__raydp_b200_alias__", False):
        _register(_pyspark_tree())
        added.append("pyspark")
    if ray and _absent("horovod") and "horovod" not in sys.modules:
        # examples/horovod_nyctaxi.py: `import horovod.torch as hvd`, `from ray.train.horovod import HorovodTrainer`
        from raydp_b200.train import horovod_api
        hvd_root = _module("horovod", "alias package (see raydp_b200.train.horovod_api)")
        hvd_root.__path__ = []
        _register({"horovod": hvd_root, "horovod.torch": horovod_api,
                   "ray.train.horovod": _module("ray.train.horovod", "",
                                                HorovodTrainer=horovod_api.horovod_trainer_class())})
        if "ray.train" in sys.modules:
            sys.modules["ray.train"].horovod = sys.modules["ray.train.horovod"]
        added.append("horovod")
    if ray and _absent("xgboost_ray") and "xgboost_ray" not in sys.modules:
        # the reference's XGBoost example talks to xgboost_ray directly (examples/xgboost_ray_nyctaxi.py:5)
        from raydp_b200.xgboost import ray_api
        _register({"xgboost_ray": _module("xgboost_ray", "alias of raydp_b200.xgboost.ray_api", RayDMatrix=ray_api.RayDMatrix,
                                          RayParams=ray_api.RayParams, train=ray_api.train, predict=ray_api.predict)})
        added.append("xgboost_ray")
    if _INSTALLED:
        import os
        os.environ["RAYDP_B200_COMPAT_ALIASES"] = "1"      # worker processes install the same aliases
    return added


class _LazyAliasFinder:
    """Installs the aliases the first time something imports ``ray`` / ``pyspark`` (or a sub-module).  Worker
    processes use this: most of them never touch the alias names, and building the alias trees imports
    torch, pandas and the whole DataFrame engine, which would slow every worker start."""

    _busy = False

    def find_spec(self, name, path=None, target=None):
        root = name.partition(".")[0]
        if root not in ("ray", "pyspark", "xgboost_ray", "horovod") or _LazyAliasFinder._busy:
            return None
        _LazyAliasFinder._busy = True
        try:
            if self in sys.meta_path:
                sys.meta_path.remove(self)
            install()
        finally:
            _LazyAliasFinder._busy = False
        if name not in sys.modules:
            return None
        return importlib.machinery.ModuleSpec(name, _ExistingModuleLoader())


def install_lazy() -> None:
    """Arrange for :func:`install` to run on the first import of an alias name."""
    if any(isinstance(f, _LazyAliasFinder) for f in sys.meta_path):
        return
    if getattr(sys.modules.get("ray"), "__raydp_b200_alias__", False):
        return
    sys.meta_path.insert(0, _LazyAliasFinder())


def uninstall() -> None:
    """Remove every alias :func:`install` registered."""
    for name in _INSTALLED:
        mod = sys.modules.get(name)
        if mod is not None and (getattr(mod, "__raydp_b200_alias__", False) or
                                getattr(mod, "__name__", "").startswith("raydp_b200")):
            del sys.modules[name]
    _INSTALLED.clear()


def installed() -> List[str]:
    return sorted(set(n.split(".")[0] for n in _INSTALLED))


def _main(argv: List[str]) -> int:
    """``python -m raydp_b200.compat script.py [args...]``: run a script written against the reference
    (``import ray`` / ``import pyspark`` / ``import raydp`` in any order) with the aliases in place."""
    import os
    import runpy
    if not argv:
        print("usage: python -m raydp_b200.compat script.py [args...]", file=sys.stderr)
        return 2
    install()
    import raydp  # noqa: F401  (registers the raydp.* alias finder)
    script = os.path.abspath(argv[0])
    sys.argv = [script] + list(argv[1:])
    sys.path.insert(0, os.path.dirname(script))
    runpy.run_path(script, run_name="__main__")
    return 0


if __name__ == "__main__":
    sys.exit(_main(sys.argv[1:]))
