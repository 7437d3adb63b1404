"""Pure helpers; same behaviour tables as reference ``tests/test_spark_utils.py:30-160``."""
import math

import pytest

import raydp_b200.sql.pandas_api as ps
from raydp_b200 import utils
from raydp_b200.sql import DataFrame


def test_df_type_check(spark_session):
    assert utils.df_type_check(spark_session.range(0, 10))
    assert utils.df_type_check(ps.range(0, 10))
    other = "df"
    msg = (f"The type: {type(other)} is not supported, only support "
           "pyspark.sql.DataFrame and pyspark.pandas.DataFrame")
    with pytest.raises(Exception) as info:
        utils.df_type_check(other)
    assert str(info.value) == msg


def test_convert_to_spark(spark_session):
    sdf = spark_session.range(0, 10)
    converted, is_spark = utils.convert_to_spark(sdf)
    assert is_spark and converted is sdf
    converted, is_spark = utils.convert_to_spark(ps.range(0, 10))
    assert not is_spark
    assert isinstance(converted, DataFrame)
    assert converted.count() == 10


def test_random_split(spark_session):
    splits = utils.random_split(spark_session.range(0, 100), [0.7, 0.3], 1)
    assert len(splits) == 2
    assert splits[0].count() + splits[1].count() == 100
    assert 50 < splits[0].count() < 90
    splits = utils.random_split(ps.range(0, 10), [0.7, 0.3])
    assert all(isinstance(s, ps.DataFrame) for s in splits) and len(splits) == 2


def test_memory_size_parser():
    units = ["", "K", "M", "G", "T"]
    expected = [10 * math.pow(2, 10 * p) for p in range(len(units))]
    for fmt in ("10{}", "10 {}", "10  {}"):
        assert [utils.parse_memory_size(fmt.format(u)) for u in units] == expected
        assert [utils.parse_memory_size(fmt.format(u.lower())) for u in units] == expected
    units_b = ["B", "KB", "MB", "GB", "TB"]
    for fmt in ("10{}", "10 {}", "10  {}"):
        assert [utils.parse_memory_size(fmt.format(u)) for u in units_b] == expected
        assert [utils.parse_memory_size(fmt.format(u.lower())) for u in units_b] == expected
    assert utils.parse_memory_size(4096) == 4096
    assert utils.parse_memory_size("1.5g") == int(1.5 * 2 ** 30)
    with pytest.raises(ValueError):
        utils.parse_memory_size("ten gigs")


@pytest.mark.parametrize("blocks", [[5, 1, 2, 3, 5, 6, 2, 1, 2], [5, 1, 2, 3, 5, 6, 2, 2, 2]])
@pytest.mark.parametrize("shuffle", [False, True])
def test_divide_blocks(blocks, shuffle):
    divided = utils.divide_blocks(blocks, 3, shuffle, 7 if shuffle else None)
    assert len(divided) == 3
    totals = []
    for rank in range(3):
        n = 0
        for index, num in divided[rank]:
            assert num <= blocks[index]
            n += num
        totals.append(n)
    assert totals[0] == totals[1] == totals[2] == math.ceil(sum(blocks) / 3)
    with pytest.raises(Exception):
        utils.divide_blocks([1, 2], 3)


def test_split_rows_evenly():
    plan = utils.split_rows_evenly([10, 3, 7, 5], 3)
    assert [sum(stop - start for _, start, stop in p) for p in plan] == [8, 8, 8]
    seen = sorted((b, s) for p in plan for b, s, _ in p)
    assert seen[0] == (0, 0)


def test_reference_import_names_resolve_to_this_framework():
    """``import raydp`` provides ``ray`` / ``pyspark`` aliases when the real packages are absent, so
    scripts written against the reference (``import ray``; ``from pyspark.sql.functions import col``)
    import unchanged; a genuine installation is never shadowed."""
    import importlib
    import sys
    import raydp  # noqa: F401
    from raydp_b200 import compat, data as rdata, runtime
    from raydp_b200.sql import functions as F
    if "ray" in compat.installed():
        import ray
        from ray.train.torch import TorchTrainer
        from ray.util.placement_group import placement_group
        # ray.init is runtime.init with Ray's default of echoing worker output on the driver
        assert ray.init.__name__ == "ray_init" and ray.shutdown is runtime.shutdown
        assert ray.data is rdata and placement_group is runtime.placement_group
        assert TorchTrainer.__module__.startswith("raydp_b200.train")
        assert ray.util.client.ray.is_connected() is False
    if "pyspark" in compat.installed():
        from pyspark.sql.functions import col
        from pyspark.sql import SparkSession
        import pyspark.pandas as ps
        assert col is F.col and SparkSession.__module__.startswith("raydp_b200.sql") and hasattr(ps, "range")
    assert compat.install() == []                      # idempotent
    real = [n for n in ("ray", "pyspark") if n not in compat.installed()]
    for n in real:                                      # never shadow a genuine package
        assert not getattr(importlib.import_module(n), "__raydp_b200_alias__", False)
    assert all(getattr(sys.modules[n], "__raydp_b200_alias__", False) for n in compat.installed())
