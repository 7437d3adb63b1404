"""K5 ingestion.  CPU: the host-side copy plan of ``ArrowGpuLoader`` (addresses, byte counts, validity
bitmap windows, vector columns, chunking across record batches) is replayed with plain memory reads.
GPU: the loader's batches equal what the CPU ``to_torch`` path yields for the same dataset (nulls as NaN or
as a fill value, vector columns, mixed dtypes, batches straddling chunks and blocks)."""
import ctypes

import numpy as np
import pyarrow as pa
import pytest
import torch

_NP = {0: np.float32, 1: np.float64, 2: np.int32, 3: np.int64, 4: np.int16, 5: np.int8}


def _table(n=1000, seed=0):
    rng = np.random.default_rng(seed)
    mask = rng.random(n) < 0.1
    lmask = rng.random(n) < 0.05
    cols = {"a": pa.array(rng.standard_normal(n).astype(np.float32)),
            "b": pa.array(rng.standard_normal(n), mask=mask),
            "v": pa.FixedSizeListArray.from_arrays(pa.array(rng.standard_normal(n * 3).astype(np.float32)), 3),
            "w": pa.array(list(rng.standard_normal((n, 2)))),                  # list<double>, uniform length
            "s": pa.array(rng.integers(-5, 5, n).astype(np.int16)),
            "i": pa.array(rng.integers(0, 50, n).astype(np.int64), mask=lmask),
            "y": pa.array(rng.standard_normal(n).astype(np.float32))}
    return pa.table(cols), mask, lmask


def test_copy_plan_replayed_on_the_host(ray_cluster):
    from raydp_b200 import data as rdata
    from raydp_b200.ops.ingest import ArrowGpuLoader
    t, mask, lmask = _table()
    n = t.num_rows
    # slices give non-zero Arrow offsets (bit offsets into the validity bitmaps included)
    ds = rdata.from_arrow([t.slice(0, 613), t.slice(613)])
    ld = ArrowGpuLoader(ds, ["a", "b", "v", "w", "s"], "y", 256, False, torch.device("cpu"),
                        sparse_columns=["i"], table_rows=[0], chunk_rows=300)
    ld._build_plan()
    assert ld.n_features == 1 + 1 + 3 + 2 + 1 and ld.ld == 8 and len(ld._chunks) == 5
    ref = {name: (np.asarray(t.column(name).combine_chunks().flatten()).reshape(n, -1) if name in ("v", "w")
                  else np.asarray(t.column(name).combine_chunks().fill_null(0)).reshape(n, 1)) for name in ld._cols}
    row = 0
    for ch in ld._chunks:
        stage = np.zeros(ch.total, dtype=np.uint8)
        for src, nb, off in zip(ch.src, ch.nbytes, ch.dst_off):
            stage[off:off + nb] = np.frombuffer((ctypes.c_uint8 * nb).from_address(src), dtype=np.uint8)
        for name, (voff, valoff, vbit0, w, code) in zip(ld._cols, ch.cols):
            dt = np.dtype(_NP[code])
            vals = np.frombuffer(stage[voff:voff + ch.rows * w * dt.itemsize].tobytes(), dtype=dt).reshape(ch.rows, w)
            if name in ("b", "i"):
                assert valoff >= 0
                bits = np.unpackbits(stage[valoff:valoff + (vbit0 + ch.rows + 7) // 8], bitorder="little")
                valid = bits[vbit0:vbit0 + ch.rows].astype(bool)
                want = ~(mask if name == "b" else lmask)[row:row + ch.rows]
                assert (valid == want).all(), name
                vals = vals * valid[:, None]
            else:
                assert valoff < 0
            assert np.allclose(vals, ref[name][row:row + ch.rows]), name
        row += ch.rows
    assert row == n


def test_plan_rejects_ragged_vectors_and_strings(ray_cluster):
    from raydp_b200 import data as rdata
    from raydp_b200.ops.ingest import ArrowGpuLoader
    ragged = pa.table({"v": pa.array([[1.0, 2.0], [3.0]]), "y": pa.array([0.0, 1.0])})
    with pytest.raises(TypeError, match="ONE length"):
        ArrowGpuLoader(rdata.from_arrow([ragged]), ["v"], "y", 2, False, torch.device("cpu"))._build_plan()
    strings = pa.table({"c": pa.array(["a", "b"]), "y": pa.array([0.0, 1.0])})
    with pytest.raises(TypeError, match="cannot be ingested"):
        ArrowGpuLoader(rdata.from_arrow([strings]), ["c"], "y", 2, False, torch.device("cpu"))._build_plan()


def test_cpu_path_null_semantics(ray_cluster):
    """NULL -> NaN in floating tensors by default (the reference's pandas path), a fill value on request, 0 in
    integer tensors."""
    from raydp_b200 import data as rdata
    t = pa.table({"x": pa.array([1.0, None, 3.0]), "k": pa.array([1, None, 3]), "y": pa.array([0.0, 1.0, None])})
    ds = rdata.from_arrow([t])
    f, l = next(iter(ds.to_torch(label_column="y", feature_columns=["x", "k"], batch_size=3)))
    assert torch.isnan(f[1, 0]) and torch.isnan(f[1, 1]) and torch.isnan(l[2, 0]) and f[0, 1] == 1.0
    f, l = next(iter(ds.to_torch(label_column="y", feature_columns=["x", "k"], batch_size=3, null_fill=-1.0)))
    assert f[1, 0] == -1.0 and l[2, 0] == -1.0
    (fx, fk), _ = next(iter(ds.to_torch(label_column="y", feature_columns=["x", "k"],
                                         feature_column_dtypes=[torch.float, torch.long], batch_size=3)))
    assert torch.isnan(fx[1, 0]) and fk[1, 0] == 0 and fk.dtype == torch.long


@pytest.mark.gpu
@pytest.mark.parametrize("null_fill", [float("nan"), 0.0, -2.5])
def test_gpu_loader_matches_cpu_to_torch(ray_cluster, null_fill):
    from raydp_b200 import data as rdata
    from raydp_b200.ops.ingest import ArrowGpuLoader
    t, _, _ = _table(n=5000, seed=3)
    ds = rdata.from_arrow([t.slice(0, 1777), t.slice(1777, 2100), t.slice(3877)])
    dev = torch.device("cuda", 0)
    feats = ["a", "b", "v", "w", "s"]
    ld = ArrowGpuLoader(ds, feats, "y", 512, False, dev, out_dtype=torch.float32, sparse_columns=["i"],
                        table_rows=[7], chunk_rows=700, null_fill=null_fill)
    cpu = list(ds.to_torch(label_column="y", feature_columns=feats, batch_size=512, null_fill=null_fill))
    idx = np.asarray(t.column("i").combine_chunks().fill_null(0)) % 7
    for epoch in range(2):                       # the second epoch replays the prebuilt plan
        got = [(x.cpu().clone(), s.cpu().clone(), y.cpu().clone()) for x, s, y in ld]
        assert len(got) == len(cpu) == 10
        row = 0
        for (x, s, y), (cf, cl) in zip(got, cpu):
            assert x.shape == (cf.shape[0], 8) and cf.shape[1] == 8        # a, b, v[3], w[2], s
            assert torch.allclose(x, cf.float(), rtol=1e-6, atol=1e-6, equal_nan=True)
            assert torch.allclose(y, cl[:, 0].float(), equal_nan=True)
            assert np.array_equal(s[:, 0].numpy(), idx[row:row + x.shape[0]])
            row += x.shape[0]
    assert ld.h2d_bytes > 2 * 5000 * (4 + 8 + 12 + 16 + 2 + 8 + 4)
    # bf16 output with a wider pitch, as the fused MLP consumes it
    ld16 = ArrowGpuLoader(ds, feats, "y", 512, True, dev, ld=16, null_fill=0.0)
    x, y = next(iter(ld16))
    want = next(iter(ds.to_torch(label_column="y", feature_columns=feats, batch_size=512, null_fill=0.0)))[0]
    assert x.dtype == torch.bfloat16 and x.shape == (512, 16) and float(x[:, 8:].float().abs().max()) == 0.0
    assert torch.allclose(x[:, :8].float().cpu(), want.float(), rtol=1e-2, atol=1e-2)
