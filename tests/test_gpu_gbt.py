"""Histogram gradient boosting on the GPU (`tree_method="gpu_hist"`): kernels against numpy, booster against the
CPU booster of the same module."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(n, F, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, F)).astype(np.float32)
    x[rng.random((n, F)) < 0.03] = np.nan                      # missing values
    y = (2 * np.nan_to_num(x[:, 0]) - np.nan_to_num(x[:, 1]) ** 2 + 0.5 * np.nan_to_num(x[:, 2] * x[:, 3])
         + 0.1 * rng.standard_normal(n))
    return x, y


@pytest.mark.parametrize("n_slots,n_bins", [(1, 257), (8, 65), (64, 257), (200, 257)])   # last one: no shared memory
def test_hist_kernel_matches_bincount(n_slots, n_bins):
    from raydp_b200 import ops
    n, F = 50_003, 7
    rng = np.random.default_rng(n_slots)
    binned = rng.integers(0, n_bins, (n, F)).astype(np.uint16)
    g = rng.standard_normal(n).astype(np.float32)
    h = rng.random(n).astype(np.float32)
    slot = rng.integers(-1, n_slots, n).astype(np.int32)
    dev = torch.device("cuda")
    hist = torch.zeros(n_slots, 2, F, n_bins, dtype=torch.float64, device=dev)
    ops.gbt_hist(torch.from_numpy(np.ascontiguousarray(binned.T).view(np.int16)).to(dev), torch.from_numpy(g).to(dev),
                 torch.from_numpy(h).to(dev), torch.from_numpy(slot).to(dev), hist)
    ref = np.zeros((n_slots, 2, F, n_bins))
    keep = slot >= 0
    for f in range(F):
        flat = slot[keep].astype(np.int64) * n_bins + binned[keep, f]
        ref[:, 0, f, :] = np.bincount(flat, weights=g[keep].astype(np.float64), minlength=n_slots * n_bins).reshape(n_slots, n_bins)
        ref[:, 1, f, :] = np.bincount(flat, weights=h[keep].astype(np.float64), minlength=n_slots * n_bins).reshape(n_slots, n_bins)
    got = hist.cpu().numpy()
    assert np.allclose(got, ref, rtol=1e-4, atol=1e-3)
    assert abs(got[:, 1].sum() - F * h[keep].astype(np.float64).sum()) < 1e-2 * F


def test_partition_kernel_moves_rows_like_numpy():
    from raydp_b200 import ops
    n, F, n_bins = 20_001, 5, 33
    rng = np.random.default_rng(3)
    binned = rng.integers(0, n_bins, (n, F)).astype(np.uint16)
    node_of = rng.integers(0, 3, n).astype(np.int32)                # nodes 0..2; 0 and 2 are split, 1 is a leaf
    in_sample = (rng.random(n) < 0.7).astype(np.uint8)
    feat = np.array([2, -1, 4, -1, -1, -1, -1], dtype=np.int32)
    thr = np.array([10, 0, 20, 0, 0, 0, 0], dtype=np.int32)
    left = np.array([3, 0, 5, 0, 0, 0, 0], dtype=np.int32)
    right = np.array([4, 0, 6, 0, 0, 0, 0], dtype=np.int32)
    dl = np.array([1, 0, 0, 0, 0, 0, 0], dtype=np.uint8)
    nxt = np.array([-1, -1, -1, 0, 1, 2, 3], dtype=np.int32)
    miss = n_bins - 1
    want = node_of.copy()
    for nid in (0, 2):
        here = node_of == nid
        b = binned[here, feat[nid]]
        go_left = np.where(b == miss, bool(dl[nid]), b <= thr[nid])
        want[here] = np.where(go_left, left[nid], right[nid])
    want_slot = np.where(in_sample == 1, nxt[want], -1)
    dev = torch.device("cuda")
    up = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    node_d, slot_d = up(node_of.copy()), torch.zeros(n, dtype=torch.int32, device=dev)
    ops.gbt_partition(up(np.ascontiguousarray(binned.T).view(np.int16)), node_d, slot_d, up(in_sample), up(feat), up(thr),
                      up(left), up(right), up(dl), up(nxt), miss)
    assert np.array_equal(node_d.cpu().numpy(), want) and np.array_equal(slot_d.cpu().numpy(), want_slot)


@pytest.mark.parametrize("objective", ["reg:squarederror", "binary:logistic"])
def test_gpu_booster_tracks_the_cpu_booster(objective):
    from raydp_b200.xgboost import hist_gbt
    x, y = _data(20_000, 10)
    if objective == "binary:logistic":
        y = (y > np.median(y)).astype(np.float64)
    params = {"objective": objective, "max_depth": 5, "eta": 0.3, "subsample": 0.8, "seed": 7}
    cpu, hist_cpu = hist_gbt.train(dict(params, device="cpu"), x, y, 8)
    gpu, hist_gpu = hist_gbt.train(dict(params, tree_method="gpu_hist"), x, y, 8)
    assert cpu.attributes["tree_method"] == "hist" and gpu.attributes["tree_method"] == "gpu_hist"
    metric = list(hist_cpu["train"])[0]
    a, b = hist_cpu["train"][metric], hist_gpu["train"][metric]
    assert b[-1] < b[0] and abs(a[-1] - b[-1]) < 0.02 * max(abs(a[-1]), 1e-3) + 1e-3, (a, b)
    # fp32 gradients on the device can flip a near-tie split; the ensembles must still agree closely
    pc, pg = cpu.predict(x), gpu.predict(x)
    assert np.mean(np.abs(pc - pg)) < 0.02 * (np.std(pc) + 1e-6)


def test_xgboost_estimator_on_a_gpu_worker(ray_cluster):
    """A boosting worker that was given a GPU grows its trees with the histogram kernels."""
    import pyarrow as pa
    from raydp_b200 import data as rdata
    from raydp_b200.xgboost import XGBoostEstimator
    x, y = _data(6000, 6, seed=5)
    cols = {f"f{i}": x[:, i] for i in range(6)}
    cols["label"] = y.astype(np.float32)
    ds = rdata.from_arrow([pa.table({k: v[i::2] for k, v in cols.items()}) for i in range(2)])
    est = XGBoostEstimator({"objective": "reg:squarederror", "max_depth": 4, "eta": 0.3}, label_column="label",
                           num_workers=1, resources_per_worker={"GPU": 1, "CPU": 1}, num_boost_round=6)
    res = est.fit(ds, ds)
    assert res["tree_method"] == "gpu_hist" and res["rounds"] == 6
    rmse = res["history"]["train"]["rmse"]
    assert rmse[-1] < 0.6 * rmse[0]
    model = est.get_model()
    assert model.attributes["tree_method"] == "gpu_hist" and model.num_boosted_rounds() == 6
    pred = model.predict(np.stack([cols[f"f{i}"] for i in range(6)], 1))
    assert np.sqrt(np.mean((pred - y) ** 2)) < 1.2 * rmse[-1] + 1e-6
