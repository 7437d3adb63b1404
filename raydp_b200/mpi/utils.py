"""Helpers of the MPI launcher.

Parity: reference ``python/raydp/mpi/utils.py:30-119`` -- ``StoppableThread``, ``run_cmd`` (process
in its own session + watchdog thread + line-by-line output forwarder), ``get_node_ip_address``.
"""
from __future__ import annotations

import os
import select
import signal
import subprocess
import sys
import threading
from typing import Callable, Dict, List, Optional

import psutil


class StoppableThread(threading.Thread):
    def __init__(self, group=None, target=None, name=None, args=(), kwargs=None):
        super().__init__(group=group, target=target, name=name, args=args, kwargs=kwargs or {}, daemon=True)
        self._stop_event = threading.Event()

    def stop(self) -> None:
        self._stop_event.set()

    def stopped(self) -> bool:
        return self._stop_event.is_set()


def run_cmd(cmd, env: Optional[Dict[str, str]] = None, failed_callback: Optional[Callable] = None,
            shell: bool = True, prefix: str = ""):
    """Start ``cmd`` in a new session; forward its output; call ``failed_callback`` on non-zero exit."""
    proc = subprocess.Popen(cmd, shell=shell, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                            env=env, start_new_session=True)

    def check_failed():
        while not threading.current_thread().stopped():
            rc = proc.poll()
            if rc is not None:
                if rc != 0 and failed_callback is not None:
                    failed_callback()
                return
            threading.current_thread()._stop_event.wait(0.1)

    def forward():
        for line in iter(proc.stdout.readline, b""):
            sys.stdout.write(prefix + line.decode(errors="replace"))
            sys.stdout.flush()

    watchdog = StoppableThread(target=check_failed, name="raydp-mpi-watchdog")
    watchdog.start()
    threading.Thread(target=forward, name="raydp-mpi-output", daemon=True).start()
    return proc, watchdog


def terminate_process_group(proc: subprocess.Popen) -> None:
    if proc is None or proc.poll() is not None:
        return
    try:
        os.killpg(proc.pid, signal.SIGTERM)
    except (ProcessLookupError, PermissionError):
        pass
    try:
        proc.wait(timeout=3)
    except subprocess.TimeoutExpired:
        try:
            os.killpg(proc.pid, signal.SIGKILL)
        except (ProcessLookupError, PermissionError):
            pass


def get_environ_value(key: str) -> str:
    assert key in os.environ, f"{key} should be set in the environ"
    return os.environ[key]


def get_node_ip_address(node_addresses: List[str]) -> str:
    """Pick the local NIC address that appears in the job's node list (reference ``:113-119``)."""
    local = {a.address for addrs in psutil.net_if_addrs().values() for a in addrs}
    local.update({"127.0.0.1"})
    for ip in node_addresses:
        if ip in local or ip.startswith("127."):
            return ip
    raise RuntimeError(f"none of the job's node addresses {node_addresses} is local to this host")
