"""Configuration keys honoured (or knowingly ignored) by the Spark-on-runtime layer.

Parity: reference ``SparkOnRayConfigs.java:4-128`` and SURVEY.md section 5.6.  RayDP keys keep
the ``spark.`` prefix so one flat ``Dict[str, str]`` carries everything.
"""
# master actor resources
SPARK_MASTER_ACTOR_RESOURCE_PREFIX = "spark.ray.raydp_spark_master.actor.resource"
SPARK_MASTER_RESOURCE_PREFIX_DEPRECATED = "spark.ray.raydp_spark_master.resource"
# executor actor resources
SPARK_EXECUTOR_ACTOR_RESOURCE_PREFIX = "spark.ray.raydp_spark_executor.actor.resource"
SPARK_EXECUTOR_ACTOR_CPU_RESOURCE = SPARK_EXECUTOR_ACTOR_RESOURCE_PREFIX + ".cpu"
RAY_ACTOR_CPU_RESOURCE_DEPRECATED = "spark.ray.actor.resource.cpu"
# fetch-task resources of the recoverable conversion
RECOVERABLE_FETCH_TASK_RESOURCE_PREFIX = "spark.ray.raydp_recoverable_fetch.task.resource"
# placement
RAY_PLACEMENT_GROUP = "spark.ray.placement_group"
RAY_BUNDLE_INDEXES = "spark.ray.bundle_indexes"
# dynamic allocation / shuffle service
DYN_ALLOC_ENABLED = "spark.dynamicAllocation.enabled"
DYN_ALLOC_MAX_EXECUTORS = "spark.dynamicAllocation.maxExecutors"
SHUFFLE_SERVICE_ENABLED = "spark.shuffle.service.enabled"
# JVM-only plumbing: accepted, logged at debug level, ignored (there is no JVM)
JVM_ONLY_KEYS = (
    "spark.javaagent", "spark.ray.log4j.factory.class", "spark.preferClassPath",
    "spark.ray.preferClassPath", "spark.log4j.config.file.name",
    "spark.ray.log4j.config.file.name", "spark.ray.raydp_app_master.extraJavaOptions",
    "raydp.executor.extraClassPath", "spark.executor.extraJavaOptions",
    "spark.driver.extraJavaOptions", "spark.driver.extraClassPath", "spark.jars",
)

RAYDP_SPARK_MASTER_SUFFIX = "_SPARK_MASTER"
EXECUTOR_NAME_PREFIX = "raydp-executor-"


def resources_from_prefix(conf, prefix: str):
    """``{name: float}`` for every ``<prefix>.<name>`` key."""
    out = {}
    plen = len(prefix) + 1
    for key, value in conf.items():
        if key.startswith(prefix + "."):
            out[key[plen:]] = float(value)
    return out
