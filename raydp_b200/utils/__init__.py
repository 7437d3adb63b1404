"""Pure helpers shared by every layer.

Parity: reference ``python/raydp/utils.py:27-222`` (``parse_memory_size`` :125-146,
``random_split`` :67-83, ``convert_to_spark``/``df_type_check`` :86-122,
``divide_blocks`` :149-222, ``get_node_address`` :34-58,
``register_exit_handler`` :61-64).  Behaviour (accepted spellings, error text,
the equal-sample invariant) is matched; the implementations are new.
"""
from __future__ import annotations

import atexit
import math
import re
import signal
import socket
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

MEMORY_SIZE_UNITS = {"K": 2 ** 10, "M": 2 ** 20, "G": 2 ** 30, "T": 2 ** 40}
# Kept for API compatibility with the reference constant of the same name.
BLOCK_SIZE_BIT = 10

_SIZE_RE = re.compile(r"^\s*([0-9]*\.?[0-9]+)\s*([KMGT]?)I?B?\s*$")


def parse_memory_size(memory_size) -> int:
    """``"10KB"``, ``"10 g"``, ``"500M"``, ``"12"`` or an int -> bytes (powers of 1024)."""
    if isinstance(memory_size, (int, np.integer)):
        return int(memory_size)
    if isinstance(memory_size, float):
        return int(memory_size)
    text = str(memory_size).strip().upper()
    match = _SIZE_RE.match(text)
    if match is None:
        raise ValueError(f"can not parse memory size: {memory_size!r}")
    number, unit = match.groups()
    scale = MEMORY_SIZE_UNITS[unit] if unit else 1
    if not unit and "." not in number:
        return int(number)
    return int(float(number) * scale)


def format_memory_size(num_bytes: int) -> str:
    """Inverse of :func:`parse_memory_size` for logs (``1536`` -> ``"1.5K"``)."""
    value = float(num_bytes)
    for unit in ("", "K", "M", "G", "T"):
        if abs(value) < 1024.0 or unit == "T":
            return f"{value:.4g}{unit}"
        value /= 1024.0
    return f"{value:.4g}T"


def _unsupported(df):
    return Exception(
        f"The type: {type(df)} is not supported, only support "
        "pyspark.sql.DataFrame and pyspark.pandas.DataFrame")


def _classify_frame(df) -> str:
    """Return ``"spark"`` or ``"pandas_on_spark"``; raise the reference's exact error otherwise."""
    # The built-in engine first; real pyspark objects are also accepted when importable.
    from raydp_b200.sql.dataframe import DataFrame as _DF
    from raydp_b200.sql.pandas_api import DataFrame as _PSDF
    if isinstance(df, _DF):
        return "spark"
    if isinstance(df, _PSDF):
        return "pandas_on_spark"
    try:  # pragma: no cover - pyspark is not part of this image
        import pyspark  # type: ignore
        if isinstance(df, pyspark.sql.DataFrame):
            return "spark"
        import pyspark.pandas as ps  # type: ignore
        if isinstance(df, ps.DataFrame):
            return "pandas_on_spark"
    except Exception:  # noqa: BLE001
        pass
    raise _unsupported(df)


def df_type_check(df) -> bool:
    """True for a Spark or pandas-on-Spark frame, raises otherwise (reference utils.py:106-112)."""
    _classify_frame(df)
    return True


def convert_to_spark(df):
    """``(spark_df, was_already_spark)`` (reference utils.py:115-122)."""
    kind = _classify_frame(df)
    if kind == "spark":
        return df, True
    return df.to_spark(), False


def random_split(df, weights: Sequence[float], seed: Optional[int] = None):
    """Random split of a Spark / pandas-on-Spark frame (reference utils.py:67-83)."""
    sdf, is_spark = convert_to_spark(df)
    parts = sdf.randomSplit(list(weights), seed)
    if is_spark:
        return parts
    from raydp_b200.sql import pandas_api as ps
    return [ps.DataFrame(p) for p in parts]


def divide_blocks(blocks: List[int], world_size: int, shuffle: bool = False,
                  shuffle_seed: Optional[int] = None) -> Dict[int, List[Tuple[int, int]]]:
    """Give every rank the same number of samples out of ``blocks`` (sizes).

    Returns ``{rank: [(block_index, samples_taken_from_that_block), ...]}``.  Blocks are
    dealt round-robin (after an optional seeded shuffle); a rank that falls short wraps
    around over its own blocks, so ``sum(samples) == ceil(total / world_size)`` for every
    rank -- the invariant tested by reference ``test_spark_utils.py:116-160``.
    """
    if len(blocks) < world_size:
        raise Exception("do not have enough blocks to divide")
    per_rank_blocks = int(math.ceil(len(blocks) / world_size))
    per_rank_samples = int(math.ceil(sum(blocks) / world_size))
    order = list(range(len(blocks)))
    padded = per_rank_blocks * world_size
    order += order[: padded - len(order)]
    rng = np.random.RandomState(shuffle_seed if shuffle_seed else 0)
    if shuffle:
        rng.shuffle(order)
    out: Dict[int, List[Tuple[int, int]]] = {}
    for rank in range(world_size):
        mine = order[rank::world_size]
        taken: List[Tuple[int, int]] = []
        have = 0
        cursor = 0
        # First pass over the assigned blocks, then wrap around until full.
        while have < per_rank_samples:
            idx = mine[cursor % len(mine)]
            size = blocks[idx]
            if size > 0:
                take = min(size, per_rank_samples - have)
                taken.append((idx, take))
                have += take
            cursor += 1
            if cursor > 4 * len(mine) + per_rank_samples:  # all-empty guard
                raise Exception("blocks assigned to a rank hold no samples")
        out[rank] = taken
    return out


def split_rows_evenly(block_rows: List[int], world_size: int, equal: bool = True
                      ) -> List[List[Tuple[int, int, int]]]:
    """Contiguous equal split used by the trainer: ``[(block, start, stop), ...]`` per rank.

    With ``equal=True`` every rank gets exactly ``total // world_size`` rows (the remainder
    is dropped) so data-parallel ranks take the same number of steps.
    """
    total = sum(block_rows)
    per = total // world_size if equal else int(math.ceil(total / world_size))
    out: List[List[Tuple[int, int, int]]] = [[] for _ in range(world_size)]
    blk, off = 0, 0
    for rank in range(world_size):
        need = per if equal else min(per, total - rank * per)
        while need > 0 and blk < len(block_rows):
            avail = block_rows[blk] - off
            if avail <= 0:
                blk, off = blk + 1, 0
                continue
            take = min(avail, need)
            out[rank].append((blk, off, off + take))
            off += take
            need -= take
    return out


def get_node_address() -> str:
    """Best-effort primary IP of this host (reference utils.py:34-58 scans raylet cmdlines)."""
    from raydp_b200 import runtime
    if runtime.is_initialized():
        return runtime.get_node_ip_address()
    sock = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    try:
        sock.connect(("10.255.255.255", 1))
        return sock.getsockname()[0]
    except OSError:
        return "127.0.0.1"
    finally:
        sock.close()


def register_exit_handler(func) -> None:
    """Run ``func`` at interpreter exit and on SIGTERM/SIGINT (reference utils.py:61-64)."""
    atexit.register(func)
    try:
        signal.signal(signal.SIGTERM, lambda *_: func())
        signal.signal(signal.SIGINT, lambda *_: func())
    except ValueError:  # not the main thread
        pass


def find_free_port(host: str = "127.0.0.1") -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as sock:
        sock.bind((host, 0))
        return sock.getsockname()[1]
