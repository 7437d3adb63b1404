"""``raydp-submit``: run a PySpark-style script against a fresh Spark-on-runtime session.

Parity: reference ``bin/raydp-submit:25-188`` wraps ``spark-submit --master ray``: it resolves the
package homes, forwards ``--conf`` pairs, optionally reads a JSON runtime config (``--ray-conf``,
format in reference ``doc/spark_on_ray.md:116-143``) and makes ``--py-files`` importable.  Master URL
``ray`` (no actor name) means "create the app master here" (reference
``RayCoarseGrainedSchedulerBackend.scala:84-96``): the submitted script just calls
``SparkSession.builder.getOrCreate()``.
"""
from __future__ import annotations

import argparse
import json
import os
import runpy
import sys
from typing import Dict, List

ENV_CONF = "RAYDP_SUBMIT_CONF"


def parse(argv: List[str]):
    ap = argparse.ArgumentParser(prog="raydp-submit", description=__doc__.split("\n\n")[0])
    ap.add_argument("--ray-conf", default=None, help="JSON file with runtime settings (num_cpus, num_gpus, ...)")
    ap.add_argument("--conf", action="append", default=[], metavar="KEY=VALUE")
    ap.add_argument("--py-files", default="", help="comma separated .py/.zip files or directories to import")
    ap.add_argument("--master", default="ray")
    ap.add_argument("--name", default=None)
    ap.add_argument("--num-executors", type=int, default=None)
    ap.add_argument("--executor-cores", type=int, default=None)
    ap.add_argument("--executor-memory", default=None)
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    return ap.parse_args(argv)


def main(argv=None) -> int:
    args = parse(sys.argv[1:] if argv is None else argv)
    conf: Dict[str, str] = {}
    for pair in args.conf:
        if "=" not in pair:
            raise SystemExit(f"--conf expects KEY=VALUE, got {pair!r}")
        k, v = pair.split("=", 1)
        conf[k.strip()] = v.strip()
    if args.name:
        conf["spark.app.name"] = args.name
    if args.num_executors is not None:
        conf["spark.executor.instances"] = str(args.num_executors)
    if args.executor_cores is not None:
        conf["spark.executor.cores"] = str(args.executor_cores)
    if args.executor_memory is not None:
        conf["spark.executor.memory"] = args.executor_memory
    conf["spark.master"] = args.master
    os.environ[ENV_CONF] = json.dumps(conf)
    for item in [p for p in args.py_files.split(",") if p]:
        path = os.path.abspath(item)
        sys.path.insert(0, path if os.path.isdir(path) or path.endswith(".zip") else os.path.dirname(path))
    from raydp_b200 import runtime
    ray_conf = {}
    if args.ray_conf:
        with open(args.ray_conf) as fh:
            ray_conf = json.load(fh)
    runtime.init(num_cpus=ray_conf.get("num_cpus"), num_gpus=ray_conf.get("num_gpus"),
                 resources=ray_conf.get("resources"), namespace=ray_conf.get("namespace"))
    sys.argv = [args.script] + list(args.script_args)
    try:
        runpy.run_path(args.script, run_name="__main__")
    finally:
        from raydp_b200.context import stop_spark
        stop_spark()
        runtime.shutdown()
    return 0


if __name__ == "__main__":
    sys.exit(main())
