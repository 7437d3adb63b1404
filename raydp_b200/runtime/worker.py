"""Worker process of the actor runtime: hosts one actor, or runs tasks one after another.

Started by the head as ``python -m raydp_b200.runtime.worker`` with the connection details in
the environment.  It keeps two connections: a *control* channel on which the head pushes
work (``exec``) and the worker answers (``done``), and the ordinary API client used when the
hosted code itself calls ``runtime.get/put/remote`` (e.g. ``init_spark`` inside an actor,
reference ``test_spark_cluster.py:113-134``).
"""
from __future__ import annotations

import os
import sys
import threading
import traceback
from concurrent.futures import ThreadPoolExecutor
from multiprocessing.connection import Client

from raydp_b200.runtime import protocol as P


def _resolve_args(args, kwargs):
    """Top-level ObjectRef arguments are replaced by their values (Ray semantics)."""
    from raydp_b200 import runtime
    def conv(v):
        return runtime.get(v) if isinstance(v, runtime.ObjectRef) else v
    return [conv(a) for a in args], {k: conv(v) for k, v in kwargs.items()}


def main() -> None:
    if os.environ.get("RAYDP_B200_COMPAT_ALIASES") == "1":
        # the driver runs with the ``ray`` / ``pyspark`` import aliases: closures it ships refer to
        # modules by those names
        try:
            from raydp_b200 import compat
            compat.install_lazy()               # on first use: building the alias trees imports torch & co
        except Exception:  # noqa: BLE001
            traceback.print_exc()
    head = os.environ[P.ENV_HEAD]
    auth = bytes.fromhex(os.environ[P.ENV_AUTH])
    worker_id = os.environ[P.ENV_WORKER]
    node_id = os.environ[P.ENV_NODE]
    session = os.environ[P.ENV_SESSION]

    from raydp_b200 import runtime
    runtime._connect_from_env()

    conn = Client(head, family="AF_UNIX", authkey=auth)
    conn.send(("worker", worker_id))
    send_lock = threading.Lock()
    state = {"actor": None, "pool": ThreadPoolExecutor(max_workers=1), "exiting": False}

    def reply(call_id, desc):
        try:
            with send_lock:
                conn.send(("done", call_id, desc, node_id))
        except (OSError, ValueError):
            os._exit(0)

    def run(call_id, kind, a, b, c):
        try:
            if kind == "create":
                cls = P.loads(a)
                args, kwargs = P.loads(b)
                args, kwargs = _resolve_args(args, kwargs)
                runtime._set_actor_context(c)
                state["actor"] = cls(*args, **kwargs)
                desc = ("inline", P.dumps(True))
            elif kind == "call":
                args, kwargs = P.loads(b)
                args, kwargs = _resolve_args(args, kwargs)
                if a == "__rdp_ready__":
                    value = True
                else:
                    value = getattr(state["actor"], a)(*args, **kwargs)
                desc = P.make_descriptor(value, session, node_id, c)
            else:  # task
                func = P.loads(a)
                args, kwargs = P.loads(b)
                args, kwargs = _resolve_args(args, kwargs)
                value = func(*args, **kwargs)
                desc = P.make_descriptor(value, session, node_id, c)
        except runtime._ActorExit:
            state["exiting"] = True
            try:
                with send_lock:
                    conn.send(("exit_actor",))
                    conn.send(("done", call_id, ("inline", P.dumps(None)), node_id))
            finally:
                os._exit(0)
        except BaseException as exc:  # noqa: BLE001
            desc = P.error_descriptor(exc, traceback.format_exc())
        reply(call_id, desc)

    try:
        while True:
            msg = conn.recv()
            if msg[0] == "exit":
                break
            _, call_id, kind, a, b, c = msg
            if kind == "create":
                conc = int(c.get("max_concurrency", 1) or 1)
                state["pool"] = ThreadPoolExecutor(max_workers=max(1, conc),
                                                   thread_name_prefix="rdp-actor")
                # construct synchronously so method calls can never overtake __init__
                run(call_id, kind, a, b, c)
            else:
                state["pool"].submit(run, call_id, kind, a, b, c)
    except (EOFError, OSError):
        pass
    finally:
        os._exit(0)


if __name__ == "__main__":
    try:
        main()
    except Exception:  # noqa: BLE001
        traceback.print_exc()
        sys.exit(1)
