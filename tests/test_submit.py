"""raydp-submit runs a script whose session is created from the submit configuration
(reference CI runs ``examples/raydp-submit.py`` the same way, ``raydp.yml:112-122``)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_raydp_submit_example(tmp_path):
    helper = tmp_path / "helper_mod.py"
    helper.write_text("VALUE = 41\n")
    script = tmp_path / "job.py"
    script.write_text(
        "import helper_mod\n"
        "from raydp_b200.sql import SparkSession\n"
        "spark = SparkSession.builder.appName('job').getOrCreate()\n"
        "print('RESULT', spark.range(0, 100).count() + helper_mod.VALUE, spark.conf.get('spark.executor.cores'))\n")
    out = subprocess.run([os.path.join(ROOT, "bin", "raydp-submit"), "--conf", "spark.executor.instances=1",
                          "--conf", "spark.executor.cores=1", "--conf", "spark.executor.memory=300M",
                          "--py-files", str(helper), str(script)], capture_output=True, text=True, timeout=180)
    assert out.returncode == 0, out.stdout[-1500:] + out.stderr[-1500:]
    assert "RESULT 141 1" in out.stdout
