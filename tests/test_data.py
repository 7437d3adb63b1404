"""Dataset surface (ray.data stand-in)."""
import numpy as np
import pyarrow as pa
import pytest
import torch

from raydp_b200 import data as rdata


def test_constructors_and_metadata(ray_cluster, tmp_path):
    ds = rdata.from_items([{"a": i, "b": float(i)} for i in range(100)], parallelism=4)
    assert ds.count() == 100 and ds.num_blocks() == 4
    assert ds.schema().names == ["a", "b"]
    assert ds.schema().base_schema.field("b").type == pa.float64()
    assert ds.take(2) == [{"a": 0, "b": 0.0}, {"a": 1, "b": 1.0}]
    assert rdata.range(10, parallelism=2).count() == 10
    ds.write_parquet(str(tmp_path / "pq"))
    back = rdata.read_parquet(str(tmp_path / "pq"))
    assert back.count() == 100 and sorted(r["a"] for r in back.take_all()) == list(range(100))


def test_split_shuffle_and_batches(ray_cluster):
    ds = rdata.from_arrow([pa.table({"x": np.arange(i * 100, (i + 1) * 100)}) for i in range(5)])
    parts = ds.split(3, equal=True)
    assert [p.count() for p in parts] == [166, 166, 166]
    seen = np.concatenate([p.to_arrow().column("x").to_numpy() for p in parts])
    assert len(set(seen.tolist())) == 498                  # views, no duplicated rows
    a, b = ds.split_at_indices([120])
    assert a.count() == 120 and b.count() == 380 and b.take(1)[0]["x"] == 120
    sh = ds.random_shuffle(seed=5)
    vals = sh.to_arrow().column("x").to_numpy()
    assert sorted(vals.tolist()) == list(range(500)) and vals.tolist() != list(range(500))
    assert sh.to_arrow().column("x").to_numpy().tolist() != ds.random_shuffle(seed=6).to_arrow().column("x").to_numpy().tolist()
    sizes = [b.num_rows for b in ds.iter_batches(batch_size=128)]
    assert sizes == [128, 128, 128, 116]
    assert [b.num_rows for b in ds.iter_batches(batch_size=128, drop_last=True)] == [128, 128, 128]
    mapped = ds.map_batches(lambda t: t.append_column("y", pa.array(t.column("x").to_numpy() * 2)))
    assert mapped.take(1)[0] == {"x": 0, "y": 0}


def test_to_torch_single_and_per_column_dtypes(ray_cluster):
    ds = rdata.from_arrow(pa.table({"f1": np.arange(10, dtype=np.float64),
                                    "f2": np.arange(10, dtype=np.int64),
                                    "y": np.arange(10, dtype=np.float64) * 3}))
    it = ds.to_torch(label_column="y", feature_columns=["f1", "f2"], feature_column_dtypes=torch.float,
                     label_column_dtype=torch.float, batch_size=4)
    batches = list(it)
    assert len(batches) == 3 == len(it)
    x, y = batches[0]
    assert x.shape == (4, 2) and x.dtype == torch.float32 and y.shape == (4, 1)
    assert torch.equal(y[:, 0], torch.tensor([0., 3., 6., 9.]))
    multi = ds.to_torch(label_column="y", feature_columns=["f1", "f2"],
                        feature_column_dtypes=[torch.float, torch.long],
                        label_column_dtype=torch.float, batch_size=5, drop_last=True)
    feats, y = next(iter(multi))
    assert isinstance(feats, list) and feats[0].dtype == torch.float32 and feats[1].dtype == torch.int64
    assert feats[1].shape == (5, 1)


def test_legacy_torch_ml_dataset_and_prefetcher(ray_cluster):
    """Reference names ``TorchMLDataset`` / ``PrefetchedDataLoader`` (``torch_ml_dataset.py:30-111``)."""
    import pyarrow as pa
    import torch
    from torch.utils.data import DataLoader
    from raydp_b200 import data as rdata
    from raydp_b200.torch.torch_ml_dataset import PrefetchedDataLoader, TorchMLDataset
    ds = rdata.from_arrow(pa.table({"a": [float(i) for i in range(100)], "b": [2.0 * i for i in range(100)],
                                    "y": [3.0 * i for i in range(100)]}))
    parts = [TorchMLDataset(ds, world_size=2, world_rank=r, feature_columns=["a", "b"], label_column="y")
             for r in range(2)]
    assert len(parts[0]) + len(parts[1]) == 100
    seen = 0
    for x, y in PrefetchedDataLoader(DataLoader(parts[0], batch_size=16), max_size=2):
        assert x.shape[1] == 2 and torch.allclose(x[:, 0] * 3, y.reshape(-1))
        seen += x.shape[0]
    assert seen == len(parts[0])


def test_legacy_parallel_iterator_worker():
    from raydp_b200.spark.parallel_iterator_worker import ParallelIteratorWorkerWithLen
    w = ParallelIteratorWorkerWithLen(lambda: range(3), repeat=True, num_records=3)
    w.par_iter_init([lambda it: (v * 10 for v in it)])
    assert [w.par_iter_next() for _ in range(5)] == [0, 10, 20, 0, 10] and len(w) == 3


def test_row_wise_map_filter_add_column_sort(ray_cluster):
    import pyarrow as pa
    from raydp_b200 import data as rdata
    ds = rdata.from_arrow([pa.table({"a": list(range(0, 5))}), pa.table({"a": list(range(5, 10))})])
    out = ds.map(lambda r: {"a": r["a"], "sq": r["a"] ** 2}).filter(lambda r: r["sq"] % 2 == 1)
    assert [r["sq"] for r in out.take_all()] == [1, 9, 25, 49, 81]
    plus = ds.add_column("b", lambda pdf: pdf["a"] * 10).sort("b", descending=True)
    assert [r["b"] for r in plus.take(3)] == [90, 80, 70]
    import pytest
    try:
        import tensorflow  # noqa: F401
    except ImportError:
        with pytest.raises(ImportError):
            ds.to_tf("a", "a", batch_size=2)


def test_aggregations_groupby_zip_join_and_io(ray_cluster, tmp_path):
    """The wider ray.data surface: global / grouped aggregates, custom AggregateFn, map_groups, zip, join,
    sampling, extra readers and writers."""
    from raydp_b200 import data
    ds = data.from_items([{"k": "a" if i % 2 else "b", "v": i, "x": float(i) * 0.5} for i in range(10)], parallelism=3)
    assert (ds.sum("v"), ds.min("v"), ds.max("x"), ds.mean("v")) == (45, 0, 4.5, 4.5)
    assert abs(ds.std("v") - 3.0276503540974917) < 1e-12 and sorted(ds.unique("k")) == ["a", "b"]
    assert ds.groupby("k").count().take_all() == [{"k": "a", "count()": 5}, {"k": "b", "count()": 5}]
    assert ds.groupby("k").sum("v").take_all() == [{"k": "a", "sum(v)": 25}, {"k": "b", "sum(v)": 20}]
    assert ds.groupby("k").mean(["v", "x"]).take_all()[0] == {"k": "a", "mean(v)": 5.0, "mean(x)": 2.5}
    assert ds.aggregate(data.Count(), data.Sum("v"), data.Quantile("v", q=0.5), data.AbsMax("x")) == \
        {"count()": 10, "sum(v)": 45, "quantile(v)": 4.5, "absmax(x)": 4.5}
    total = data.AggregateFn(init=lambda k: 0, accumulate_row=lambda a, r: a + r["v"], merge=lambda a, b: a + b,
                             name="tot")
    assert ds.groupby("k").aggregate(data.Max("v"), total).take_all() == \
        [{"k": "a", "max(v)": 9, "tot": 25}, {"k": "b", "max(v)": 8, "tot": 20}]
    shifted = ds.groupby("k").map_groups(lambda p: p.assign(v=p.v - p.v.min())).take_all()
    assert [r["v"] for r in shifted if r["k"] == "a"] == [0, 2, 4, 6, 8]
    assert ds.flat_map(lambda r: [r, r]).count() == 20 and ds.rename_columns({"v": "w"}).columns() == ["k", "w", "x"]
    assert list(ds.take_batch(2)["v"]) == [0, 1]
    assert ds.zip(ds.select_columns(["v"])).columns() == ["k", "v", "x", "v_1"]
    assert ds.join(data.from_items([{"k": "a", "n": 1}]), "inner", on="k").count() == 5
    assert 0 < ds.random_sample(0.5, seed=1).count() < 10 and ds.repeat(2).count() == 20
    shards = ds.streaming_split(2)
    assert sum(s.count() for s in shards) == 10 and sum(len(b["v"]) for b in shards[0].iter_batches(batch_format="numpy")) > 0
    ds.write_json(str(tmp_path / "j"))
    ds.write_csv(str(tmp_path / "c"))
    assert data.read_json(str(tmp_path / "j")).sum("v") == 45 and data.read_csv(str(tmp_path / "c")).count() == 10
    assert data.read_text(str(tmp_path / "c")).count() == 13          # 3 blocks: 3 headers + 10 rows
    assert data.range_tensor(4, shape=(2, 2)).take(1) == [{"data": [[0, 0], [0, 0]]}]
    assert data.DataContext.get_current() is data.DataContext.get_current()


def test_preprocessors(ray_cluster):
    import pandas as pd
    from raydp_b200 import data
    from raydp_b200.data.preprocessors import (Chain, Concatenator, LabelEncoder, MinMaxScaler, OneHotEncoder,
                                               OrdinalEncoder, PreprocessorNotFittedException, SimpleImputer,
                                               StandardScaler)
    ds = data.from_items([{"k": "a" if i % 2 else "b", "v": float(i), "x": None if i == 3 else float(i) * 2,
                           "y": "p" if i < 5 else "n"} for i in range(10)], parallelism=3)
    scaled = StandardScaler(["v"]).fit_transform(ds).to_pandas()["v"]
    assert abs(scaled.mean()) < 1e-12 and abs(scaled.std(ddof=0) - 1.0) < 1e-12
    assert MinMaxScaler(["v"]).fit_transform(ds).max("v") == 1.0
    assert abs(SimpleImputer(["x"]).fit_transform(ds).take(4)[3]["x"] - 28.0 / 3) < 1e-12
    assert SimpleImputer(["x"], "constant", fill_value=-1.0).transform(ds).take(4)[3]["x"] == -1.0
    chain = Chain(SimpleImputer(["x"]), OneHotEncoder(["k"]), LabelEncoder("y"), MinMaxScaler(["v", "x"]),
                  Concatenator(output_column_name="features", exclude=["y"]))
    out = chain.fit_transform(ds)
    assert out.columns() == ["y", "features"] and out.take(1) == [{"y": 1, "features": [0.0, 0.0, 0.0, 1.0]}]
    served = chain.transform_batch(pd.DataFrame({"k": ["a"], "v": [9.0], "x": [None], "y": ["n"]}))
    assert served["y"].tolist() == [0] and served["features"][0].tolist()[0] == 1.0
    assert [r["k"] for r in OrdinalEncoder(["k"]).fit_transform(ds).take(2)] == [1, 0]
    enc = LabelEncoder("y").fit(ds)
    assert enc.inverse_transform(enc.transform(ds)).take(1)[0]["y"] == "p"
    with pytest.raises(PreprocessorNotFittedException):
        StandardScaler(["v"]).transform(ds)


def test_batch_iteration_options(ray_cluster):
    import numpy as np
    import torch
    from raydp_b200 import data
    ds = data.from_items([{"i": i, "f": [float(i), float(-i)]} for i in range(100)], parallelism=4)
    batches = list(ds.iter_batches(batch_size=16, batch_format="numpy", local_shuffle_buffer_size=32,
                                   local_shuffle_seed=0))
    seen = np.concatenate([b["i"] for b in batches])
    assert sorted(seen.tolist()) == list(range(100)) and seen[:16].tolist() != list(range(16))     # all rows, shuffled
    assert batches[0]["f"].shape == (16, 2)                                       # vector column -> [n, width]
    first = next(iter(ds.iter_torch_batches(batch_size=8, dtypes={"i": torch.float32, "f": torch.float32})))
    assert first["f"].shape == (8, 2) and first["i"].dtype == torch.float32
    assert next(iter(ds.iter_torch_batches(batch_size=4, collate_fn=lambda b: int(b["i"].sum())))) == 6


def test_to_torch_batches_straddle_blocks_in_row_order(ray_cluster):
    """Batches are slices of per-block tensors; a batch that spans two (or three) blocks is stitched in row order,
    for the single-matrix and the per-column (list of dtypes) conventions alike."""
    import numpy as np
    import pyarrow as pa
    import torch
    from raydp_b200 import data as rdata
    sizes = [7, 0, 3, 25, 1, 10]                               # uneven blocks, one empty, one smaller than a batch
    start, tables = 0, []
    for n in sizes:
        ids = np.arange(start, start + n, dtype=np.float32)
        tables.append(pa.table({"a": ids, "b": ids * 2, "y": ids + 0.5}))
        start += n
    ds = rdata.from_arrow(tables)
    for drop_last in (False, True):
        for dtypes in (torch.float, [torch.float, torch.long]):
            it = ds.to_torch(label_column="y", feature_columns=["a", "b"], batch_size=8, drop_last=drop_last,
                             feature_column_dtypes=dtypes, label_column_dtype=torch.float)
            seen, shapes = [], []
            for feats, label in it:
                a = feats[0].view(-1) if isinstance(feats, list) else feats[:, 0]
                b = feats[1].view(-1) if isinstance(feats, list) else feats[:, 1]
                assert torch.equal(b.float(), a.float() * 2) and torch.equal(label.view(-1), a.float() + 0.5)
                seen += a.tolist()
                shapes.append(len(a))
            total = sum(sizes)
            want = total // 8 * 8 if drop_last else total
            assert seen == [float(i) for i in range(want)]
            assert shapes[:-1] == [8] * (len(shapes) - 1) and len(it) == len(shapes)


def test_iter_torch_batches_slices_blocks_and_keeps_row_order(ray_cluster):
    import numpy as np
    import pyarrow as pa
    import torch
    from raydp_b200 import data as rdata
    ds = rdata.from_arrow([pa.table({"a": np.arange(7.0), "i": np.arange(7)}),
                           pa.table({"a": np.arange(7.0, 10.0), "i": np.arange(7, 10)}),
                           pa.table({"a": np.arange(10.0, 30.0), "i": np.arange(10, 30)})])
    got = list(ds.iter_torch_batches(batch_size=8, dtypes={"a": torch.float32, "i": torch.int64}))
    assert [len(b["a"]) for b in got] == [8, 8, 8, 6]
    assert torch.equal(torch.cat([b["i"] for b in got]), torch.arange(30)) and got[0]["a"].dtype == torch.float32
    assert [len(b["a"]) for b in ds.iter_torch_batches(batch_size=8, drop_last=True)] == [8, 8, 8]
    assert list(ds.iter_torch_batches(batch_size=16, collate_fn=lambda b: int(b["i"].sum()))) == [120, 315]
