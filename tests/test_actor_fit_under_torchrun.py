"""An actor-mode fit started from a process that itself runs under torchrun (what bench.py's actor arm does at
N > 1): the workers inherit the launcher's TORCHELASTIC_* environment and must still form their own group."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_actor_workers_form_their_own_group_under_torchrun(tmp_path):
    script = tmp_path / "driver.py"
    script.write_text(textwrap.dedent(f"""
        import datetime, os, sys
        sys.path.insert(0, {ROOT!r})
        import torch, torch.distributed as dist
        dist.init_process_group("gloo")
        store = dist.distributed_c10d._get_default_store()
        if dist.get_rank() == 0:
            os.environ["RAYDP_B200_SPMD"] = "0"
            from raydp_b200 import runtime, train
            from raydp_b200.train import ScalingConfig, TorchTrainer
            runtime.init(num_cpus=4)

            def loop():
                import torch.distributed as d
                t = torch.ones(1) * (d.get_rank() + 1)
                d.all_reduce(t)
                train.report({{"sum": float(t.item()), "world": d.get_world_size()}})
            res = TorchTrainer(loop, scaling_config=ScalingConfig(num_workers=2, use_gpu=False)).fit()
            print("RESULT", res.metrics, flush=True)
            runtime.shutdown()
            store.set("done", "1")
        else:
            store.wait(["done"], datetime.timedelta(minutes=3))
        dist.barrier()
        dist.destroy_process_group()
    """))
    env = dict(os.environ, RAYDP_B200_PG_TIMEOUT_S="60")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29713", str(script)],
                         capture_output=True, text=True, timeout=240, env=env)
    assert out.returncode == 0, out.stdout[-1500:] + out.stderr[-1500:]
    assert "RESULT {'sum': 3.0, 'world': 2}" in out.stdout
