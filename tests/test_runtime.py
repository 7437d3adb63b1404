"""Actor runtime: tasks, actors, restart policy, placement groups, node loss."""
import os
import signal
import time

import pyarrow as pa
import pytest

from raydp_b200 import runtime as rt
from raydp_b200.runtime.cluster_utils import Cluster


@rt.remote
def _add(a, b):
    return a + b


@rt.remote
class _Counter:
    def __init__(self, start=0):
        self.v = start

    def inc(self, n=1):
        self.v += n
        return self.v

    def pid(self):
        return os.getpid()

    def table(self, n):
        return pa.table({"x": list(range(n))})

    def restarted(self):
        return rt.get_runtime_context().was_current_actor_reconstructed

    def boom(self):
        raise ValueError("boom")

    def leave(self):
        rt.exit_actor()


def test_tasks_and_object_refs(ray_cluster):
    assert rt.get(_add.remote(1, 2)) == 3
    ref = rt.put(40)
    assert rt.get(_add.remote(ref, 2)) == 42          # top-level refs are resolved
    refs = [_add.remote(i, i) for i in range(4)]
    ready, rest = rt.wait(refs, num_returns=4, timeout=30)
    assert len(ready) == 4 and not rest
    assert rt.get(refs) == [0, 2, 4, 6]


def test_actor_basics_and_errors(ray_cluster):
    c = _Counter.options(name="cnt").remote(10)
    assert rt.get(c.inc.remote()) == 11
    assert rt.get(rt.get_actor("cnt").inc.remote(4)) == 15
    t = rt.get(c.table.remote(50000))                 # large result travels through shm
    assert isinstance(t, pa.Table) and t.num_rows == 50000
    with pytest.raises(rt.TaskError, match="boom"):
        rt.get(c.boom.remote())
    with pytest.raises(ValueError):
        _Counter.options(name="cnt").remote()
    rt.get(c.leave.remote())
    time.sleep(0.3)
    with pytest.raises(rt.ActorDiedError):
        rt.get(c.inc.remote())
    with pytest.raises(ValueError):
        rt.get_actor("cnt")


def test_actor_restart_keeps_name(ray_cluster):
    c = _Counter.options(name="phoenix", max_restarts=-1, max_task_retries=-1).remote(5)
    pid = rt.get(c.pid.remote())
    os.kill(pid, signal.SIGKILL)
    assert rt.get(c.inc.remote(), timeout=60) == 6     # state rebuilt from ctor args
    assert rt.get(c.restarted.remote()) is True
    assert rt.get(c.pid.remote()) != pid
    assert rt.list_actors([("name", "=", "phoenix")])[0]["num_restarts"] == 1


def test_resources_and_placement_groups(ray_cluster):
    base = rt.available_resources()["CPU"]
    c = _Counter.options(num_cpus=0.5, resources={}).remote()
    rt.get(c.inc.remote())
    assert rt.available_resources()["CPU"] == pytest.approx(base - 0.5)
    rt.kill(c)
    for strategy in ("PACK", "SPREAD", "STRICT_PACK"):
        pg = rt.placement_group([{"CPU": 1}, {"CPU": 1}], strategy=strategy)
        assert pg.wait(10)
        assert rt.placement_group_table(pg)["state"] == "CREATED"
        assert pg.bundle_count == 2
        rt.remove_placement_group(pg)
        assert rt.placement_group_table(pg)["state"] == "REMOVED"
    deadline = time.time() + 5
    while rt.available_resources().get("CPU") != base and time.time() < deadline:
        time.sleep(0.05)
    assert rt.available_resources()["CPU"] == base
    # STRICT_SPREAD needs as many nodes as bundles: stays pending on a single node
    pg = rt.placement_group([{"CPU": 1}, {"CPU": 1}], strategy="STRICT_SPREAD")
    assert not pg.wait(0.5)
    rt.remove_placement_group(pg)


def test_custom_resource_node_affinity_and_node_loss():
    rt.shutdown()
    cluster = Cluster(initialize_head=True, head_node_args={"num_cpus": 2})
    try:
        node = cluster.add_node(num_cpus=2, resources={"special": 2})
        c = _Counter.options(resources={"special": 1}, max_restarts=-1).remote()
        rt.get(c.inc.remote())
        rec = [a for a in rt.list_actors() if a["actor_id"] == c._actor_id][0]
        assert rec["node_id"] == node.node_id and rec["node_ip"] == "127.0.0.2"

        @rt.remote(resources={"special": 0.001}, max_retries=-1)
        def make():
            return pa.table({"v": list(range(100000))})

        ref = make.remote()
        assert rt.get(ref).num_rows == 100000
        cluster.remove_node(node)                       # block and actor are lost
        cluster.add_node(num_cpus=2, resources={"special": 2})
        assert rt.get(ref, timeout=60).num_rows == 100000   # rebuilt from lineage
        assert rt.get(c.inc.remote(), timeout=60) == 1      # actor restarted elsewhere
    finally:
        cluster.shutdown()


def test_actor_pool_and_queue(ray_cluster):
    """``ActorPool`` keeps submission order in ``map`` and hands queued work to freed actors; ``Queue`` is
    shared state in an actor, usable from tasks."""
    from raydp_b200 import runtime
    from raydp_b200.runtime.util import ActorPool, Empty, Full, Queue

    @runtime.remote
    class Doubler:
        def f(self, v):
            import time
            time.sleep(0.01 * (5 - v))
            return v * 2

    pool = ActorPool([Doubler.remote(), Doubler.remote()])
    assert list(pool.map(lambda a, v: a.f.remote(v), range(5))) == [0, 2, 4, 6, 8]
    assert sorted(pool.map_unordered(lambda a, v: a.f.remote(v), range(5))) == [0, 2, 4, 6, 8]
    assert not pool.has_next() and pool.has_free()

    q = Queue(maxsize=2)
    q.put(1)
    q.put(2)
    with pytest.raises(Full):
        q.put(3, block=False)
    assert (q.get(), q.size(), q.get(), q.empty()) == (1, 1, 2, True)
    with pytest.raises(Empty):
        q.get(timeout=0.05)

    @runtime.remote
    def producer(queue):
        for i in range(3):
            queue.put(i)

    shared = Queue()
    runtime.get(producer.remote(shared))
    assert [shared.get() for _ in range(3)] == [0, 1, 2]
    shared.shutdown()


def _shm_segments(prefix="rdp_"):
    import os
    return [n for n in os.listdir("/dev/shm") if n.startswith(prefix)]


def test_shared_memory_is_clean_after_shutdown_and_after_a_killed_driver(tmp_path):
    """Objects are POSIX shared-memory segments: ``shutdown`` must remove every one of the session, and a
    session whose driver was killed (no shutdown ran) must be reaped by the next ``init``."""
    import subprocess
    import sys
    import numpy as np
    from raydp_b200 import runtime
    runtime.shutdown()
    runtime.init(num_cpus=2)
    session = runtime._head.session
    refs = [runtime.put(np.arange(200000 + i)) for i in range(5)]

    @runtime.remote
    def produce(i):
        return np.full(150000, i)
    refs += [produce.remote(i) for i in range(4)]
    assert int(runtime.get(refs)[-1][0]) == 3
    assert len(_shm_segments(f"rdp_{session}_")) >= 9
    runtime.shutdown()
    assert _shm_segments(f"rdp_{session}_") == [] and _shm_segments(f"rdphead_{session}_") == []
    # a driver that dies without shutting down
    code = "\n".join(["import os, numpy as np", "from raydp_b200 import runtime", "runtime.init(num_cpus=1)",
                      "refs = [runtime.put(np.arange(200000)) for _ in range(7)]",
                      "print(runtime._head.session, flush=True)", "os._exit(3)"])
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=root, timeout=120)
    assert out.returncode == 3, out.stderr[-1000:]
    dead = out.stdout.strip().splitlines()[-1]
    assert len(_shm_segments(f"rdp_{dead}_")) == 7           # leaked, as expected
    runtime.init(num_cpus=1)                                  # the next head reaps them
    try:
        assert _shm_segments(f"rdp_{dead}_") == [] and _shm_segments(f"rdphead_{dead}_") == []
    finally:
        runtime.shutdown()


def test_worker_logs_have_a_job_header_and_can_be_followed_from_the_driver(tmp_path):
    """Capability of the reference's Java agent (core/agent/.../Agent.java:49-97): worker stdout / stderr land in
    per-process log files that start with a ``:job_id:`` header; ``log_to_driver`` echoes them on the driver."""
    import subprocess
    import sys
    code = "\n".join([
        "import sys, time",
        "from raydp_b200 import runtime",
        "runtime.init(num_cpus=2, log_to_driver=True)",
        "@runtime.remote",
        "class Talker:",
        "    def say(self, what):",
        "        print('worker says', what, flush=True)",
        "        return what",
        "t = Talker.options(name='talker').remote()",
        "assert runtime.get(t.say.remote('hello-42')) == 'hello-42'",
        "logs = runtime.worker_logs('talker')",
        "assert len(logs) == 1, logs",
        "rec = next(iter(logs.values()))",
        "assert rec['text'].startswith(':job_id:' + rec['job_id']), rec['text'][:80]",
        "assert ':actor_name:talker' in rec['text'] and 'worker says hello-42' in rec['text']",
        "time.sleep(0.8)",
        "runtime.shutdown()",
        "print('OK')",
    ])
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120,
                         cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert out.returncode == 0 and "OK" in out.stdout, out.stdout[-800:] + out.stderr[-1500:]
    assert "(actor-talker pid=" in out.stderr and "worker says hello-42" in out.stderr, out.stderr[-800:]
