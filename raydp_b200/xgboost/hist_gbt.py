"""Distributed histogram gradient boosting: the training algorithm behind ``XGBoostEstimator`` when the
boosting runs on this framework's own worker group.

The reference hands the work to ``ray.train.xgboost.XGBoostTrainer`` (``python/raydp/xgboost/estimator.py:58-81``):
``num_workers`` processes each hold a row shard, build per-node gradient histograms and all-reduce them
(rabit) so every worker grows the same tree; a checkpoint is kept every iteration
(``CheckpointConfig(checkpoint_frequency=1, num_to_keep=1)``).  This module restates that algorithm
(XGBoost's ``tree_method="hist"``) on numpy + ``torch.distributed``:

1. **sketch**: every worker proposes per-feature quantile cuts of its shard, the proposals are
   all-gathered and merged into global bin edges (``max_bin``, default 256) -- all workers bin alike;
2. **per boosting round**: gradients / hessians of the objective at the current margin; the tree is grown
   level by level: local histograms of (sum g, sum h) per (node, feature, bin) with ``np.bincount``, ONE
   all-reduce per level for all nodes of the level, split choice by the XGBoost gain
   ``1/2 [G_L^2/(H_L+lambda) + G_R^2/(H_R+lambda) - G^2/(H+lambda)] - gamma`` under ``min_child_weight``
   and ``max_depth``, missing values (NaN) sent to the side with the larger gain, leaf weight
   ``-eta * G/(H+lambda)``;
3. metrics (``rmse`` / ``logloss`` / ``error``) are all-reduced sums, so every rank reports the same
   numbers; rank 0 checkpoints the model after every round, and a restarted group resumes after the
   last checkpointed round.

Supported parameters (XGBoost names): ``objective`` (``reg:squarederror``, ``reg:logistic``,
``binary:logistic``, ``reg:pseudohubererror``), ``eta``/``learning_rate``, ``max_depth``, ``lambda``/
``reg_lambda``, ``gamma``/``min_split_loss``, ``min_child_weight``, ``max_bin``, ``base_score``,
``subsample`` (row sampling, seeded identically per round on every worker's own shard), ``seed``,
``eval_metric``.  Anything else is rejected loudly rather than ignored.
"""
from __future__ import annotations

import json
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

_KNOWN = {"objective", "eta", "learning_rate", "max_depth", "lambda", "reg_lambda", "gamma", "min_split_loss",
          "min_child_weight", "max_bin", "base_score", "subsample", "seed", "random_state", "eval_metric",
          "tree_method", "nthread", "n_jobs", "verbosity", "num_boost_round", "n_estimators", "device"}


class _Comm:
    """Sum / gather over the worker group; the identity when there is one worker."""

    def __init__(self):
        try:
            import torch.distributed as dist
            self.dist = dist if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1 else None
        except ImportError:
            self.dist = None
        self.world = self.dist.get_world_size() if self.dist else 1
        self.rank = self.dist.get_rank() if self.dist else 0

    def allreduce(self, arr: np.ndarray) -> np.ndarray:
        if self.dist is None:
            return arr
        import torch
        t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64))
        if t.device.type == "cpu" and self.dist.get_backend() == "nccl":
            t = t.cuda()
        self.dist.all_reduce(t)
        return t.cpu().numpy()

    def allgather(self, obj) -> list:
        if self.dist is None:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out


# ------------------------------------------------------------------------------------------- objective
def _objective(name: str):
    if name in ("reg:squarederror", "reg:linear"):
        return (lambda m, y: (m - y, np.ones_like(m))), (lambda m: m), "rmse"
    if name in ("binary:logistic", "reg:logistic"):
        def gh(m, y):
            p = 1.0 / (1.0 + np.exp(-m))
            return p - y, np.maximum(p * (1.0 - p), 1e-16)
        return gh, (lambda m: 1.0 / (1.0 + np.exp(-m))), "logloss"
    if name == "reg:pseudohubererror":
        def gh(m, y):
            r = m - y
            s = np.sqrt(1.0 + r * r)
            return r / s, 1.0 / (s * s * s)
        return gh, (lambda m: m), "rmse"
    raise ValueError(f"objective {name!r} is not supported by the built-in distributed booster "
                     "(reg:squarederror, reg:logistic, binary:logistic, reg:pseudohubererror)")


def _metric_sums(metric: str, pred: np.ndarray, y: np.ndarray) -> np.ndarray:
    n = float(len(y))
    if metric == "rmse":
        return np.array([np.sum((pred - y) ** 2), n])
    if metric == "mae":
        return np.array([np.sum(np.abs(pred - y)), n])
    if metric == "logloss":
        p = np.clip(pred, 1e-12, 1 - 1e-12)
        return np.array([-np.sum(y * np.log(p) + (1 - y) * np.log(1 - p)), n])
    if metric == "error":
        return np.array([np.sum((pred > 0.5) != (y > 0.5)), n])
    raise ValueError(f"eval_metric {metric!r} is not supported (rmse, mae, logloss, error)")


def _metric_value(metric: str, sums: np.ndarray) -> float:
    v = sums[0] / max(1.0, sums[1])
    return float(np.sqrt(v)) if metric == "rmse" else float(v)


# ----------------------------------------------------------------------------------------------- model
class HistGBTModel:
    """The trained ensemble.  Trees are flat arrays: ``feature[i] < 0`` marks a leaf with ``value[i]``;
    otherwise rows with ``x[feature] < threshold`` go to ``left[i]``, NaNs follow ``default_left[i]``."""

    def __init__(self, objective: str, base_margin: float, feature_names: Optional[Sequence[str]] = None):
        self.objective = objective
        self.base_margin = float(base_margin)
        self.feature_names = list(feature_names) if feature_names is not None else None
        self.trees: List[Dict[str, list]] = []
        self.best_iteration = 0
        self.attributes: Dict[str, Any] = {}          # e.g. {"tree_method": "gpu_hist"}: how the model was grown

    # xgboost.Booster-flavoured surface ------------------------------------------------------------
    def num_boosted_rounds(self) -> int:
        return len(self.trees)

    def margin(self, x: np.ndarray, trees: Optional[Sequence[Dict[str, list]]] = None) -> np.ndarray:
        x = np.asarray(x, dtype=np.float32)
        out = np.full(x.shape[0], self.base_margin, dtype=np.float64)
        for tree in (self.trees if trees is None else trees):
            out += _tree_predict(tree, x)
        return out

    def predict(self, x, output_margin: bool = False) -> np.ndarray:
        x = x.get_data() if hasattr(x, "get_data") else x
        m = self.margin(np.asarray(x))
        return m if output_margin else _objective(self.objective)[1](m)

    inplace_predict = predict

    def save_model(self, path: str) -> None:
        with open(path, "w", encoding="utf-8") as fh:
            json.dump({"format": "raydp_b200.hist_gbt/1", "objective": self.objective, "base_margin": self.base_margin,
                       "feature_names": self.feature_names, "trees": self.trees, "attributes": self.attributes}, fh)

    @classmethod
    def load_model(cls, path: str) -> "HistGBTModel":
        with open(path, "r", encoding="utf-8") as fh:
            blob = json.load(fh)
        if blob.get("format") != "raydp_b200.hist_gbt/1":
            raise ValueError(f"{path} is not a raydp_b200 hist_gbt model")
        model = cls(blob["objective"], blob["base_margin"], blob["feature_names"])
        model.trees = blob["trees"]
        model.attributes = dict(blob.get("attributes") or {})
        return model

    def get_dump(self) -> List[str]:
        names = self.feature_names
        out = []
        for tree in self.trees:
            lines = []
            for i, f in enumerate(tree["feature"]):
                if f < 0:
                    lines.append(f"{i}:leaf={tree['value'][i]:.6g}")
                else:
                    fname = names[f] if names else f"f{f}"
                    miss = tree["left"][i] if tree["default_left"][i] else tree["right"][i]
                    lines.append(f"{i}:[{fname}<{tree['threshold'][i]:.6g}] yes={tree['left'][i]},"
                                 f"no={tree['right'][i]},missing={miss}")
            out.append("\n".join(lines))
        return out


def _tree_predict(tree: Dict[str, list], x: np.ndarray) -> np.ndarray:
    feature = np.asarray(tree["feature"])
    thr = np.asarray(tree["threshold"], dtype=np.float32)
    left, right = np.asarray(tree["left"]), np.asarray(tree["right"])
    dleft = np.asarray(tree["default_left"], dtype=bool)
    value = np.asarray(tree["value"], dtype=np.float64)
    node = np.zeros(x.shape[0], dtype=np.int64)
    active = feature[node] >= 0
    while active.any():
        idx = np.nonzero(active)[0]
        f = feature[node[idx]]
        v = x[idx, f]
        go_left = np.where(np.isnan(v), dleft[node[idx]], v < thr[node[idx]])
        node[idx] = np.where(go_left, left[node[idx]], right[node[idx]])
        active = feature[node] >= 0
    return value[node]


# --------------------------------------------------------------------------------------------- training
def _local_quantiles_gpu(xd, qs: np.ndarray) -> List[Tuple[np.ndarray, int]]:
    """``np.quantile(col[~isnan], qs)`` (linear interpolation) for every column of the device matrix ``xd``: one
    device sort per feature instead of a host sort (the host sorts of a 1 M x 28 table take longer than 50
    boosting rounds on the GPU)."""
    import torch
    out = []
    q = torch.from_numpy(qs).to(xd.device)
    for f in range(xd.shape[1]):
        col = xd[:, f]
        col = torch.sort(col[~torch.isnan(col)]).values.double()
        m = col.numel()
        if m == 0:
            out.append((np.zeros(0), 0))
            continue
        pos = q * (m - 1)
        lo = pos.floor().long()
        hi = torch.clamp(lo + 1, max=m - 1)
        vals = col[lo] + (col[hi] - col[lo]) * (pos - lo.double())
        out.append((vals.cpu().numpy(), m))
    return out


def _global_cuts(x: np.ndarray, max_bin: int, comm: _Comm, xd=None) -> List[np.ndarray]:
    """Per-feature bin edges shared by all workers: merge of every worker's local quantile proposals
    (computed on the device when ``xd``, the feature matrix on the GPU, is given)."""
    F = x.shape[1]
    qs = np.linspace(0.0, 1.0, 2 * max_bin + 1)[1:-1]
    local = []
    if xd is not None:
        local = _local_quantiles_gpu(xd, qs)
    for f in range(F if xd is None else 0):
        col = x[:, f]
        col = col[~np.isnan(col)]
        local.append((np.quantile(col, qs) if len(col) else np.zeros(0), len(col)))
    gathered = comm.allgather(local)
    cuts = []
    for f in range(F):
        cand = np.concatenate([np.asarray(g[f][0], dtype=np.float64) for g in gathered if g[f][1] > 0] or [np.zeros(0)])
        w = np.concatenate([np.full(len(g[f][0]), g[f][1] / max(1, len(g[f][0]))) for g in gathered if g[f][1] > 0]
                           or [np.zeros(0)])
        if len(cand) == 0:
            cuts.append(np.zeros(0, dtype=np.float32))
            continue
        order = np.argsort(cand, kind="stable")
        cand, w = cand[order], w[order]
        cdf = np.cumsum(w) / np.sum(w)
        picks = np.searchsorted(cdf, np.linspace(0, 1, max_bin + 1)[1:-1], side="left")
        edges = np.unique(cand[np.clip(picks, 0, len(cand) - 1)]).astype(np.float32)
        cuts.append(edges)
    return cuts


def _bin(x: np.ndarray, cuts: List[np.ndarray]) -> Tuple[np.ndarray, int]:
    """uint16 bin ids; bin ``b`` holds values in ``[cut[b-1], cut[b])``; the LAST bin id is 'missing'."""
    nb = max((len(c) for c in cuts), default=0) + 1          # value bins per feature (at most)
    out = np.empty(x.shape, dtype=np.uint16)
    for f, c in enumerate(cuts):
        col = x[:, f]
        b = np.searchsorted(c, col, side="right").astype(np.uint16)
        b[np.isnan(col)] = nb                                 # missing
        out[:, f] = b
    return out, nb + 1


def _split_level(hist: np.ndarray, level: List[int], tree: Dict[str, list], depth: int, cuts: List[np.ndarray],
                 n_bins: int, p: Dict[str, float]) -> List[Tuple[int, int, int, bool, int, int]]:
    """Leaf values and best splits of one level from its (already all-reduced) histogram
    ``[len(level), 2, F, n_bins]``.  Appends the children to ``tree``; returns ``(node, feature, bin, default_left,
    left child, right child)`` for every node that was split."""
    lam, gamma, mcw, eta = p["lambda"], p["gamma"], p["min_child_weight"], p["eta"]
    miss = n_bins - 1
    F = hist.shape[2]
    splits = []
    for i, nid in enumerate(level):
        hg, hh = hist[i, 0], hist[i, 1]
        G, H = hg[0].sum(), hh[0].sum()                                      # any feature's bins sum to the node total
        tree["value"][nid] = float(-eta * G / (H + lam))
        if depth == int(p["max_depth"]) or H < 2 * mcw:
            continue
        parent = G * G / (H + lam)
        gm, hm = hg[:, miss], hh[:, miss]
        cg, chh = np.cumsum(hg[:, :miss], axis=1), np.cumsum(hh[:, :miss], axis=1)    # left sums for "x < cut[b]" after bin b
        best = (0.0, -1, -1, True)
        for default_left in (True, False):
            GL = cg[:, :-1] + (gm[:, None] if default_left else 0.0)
            HL = chh[:, :-1] + (hm[:, None] if default_left else 0.0)
            GR, HR = G - GL, H - HL
            gain = 0.5 * (GL * GL / (HL + lam) + GR * GR / (HR + lam) - parent) - gamma
            ok = (HL >= mcw) & (HR >= mcw)
            for f in range(F):                                                  # bins beyond this feature's cuts are empty
                ok[f, len(cuts[f]):] = False
            gain = np.where(ok, gain, -np.inf)
            k = int(np.argmax(gain))
            f, b = divmod(k, gain.shape[1])
            if gain[f, b] > best[0] + 1e-12:
                best = (float(gain[f, b]), f, b, default_left)
        if best[1] < 0:
            continue
        _, f, b, default_left = best
        li, ri = len(tree["feature"]), len(tree["feature"]) + 1
        for key, val in (("feature", -1), ("threshold", 0.0), ("left", -1), ("right", -1), ("default_left", True),
                         ("value", 0.0)):
            tree[key] += [val, val]
        tree["feature"][nid], tree["threshold"][nid] = int(f), float(cuts[f][b])
        tree["left"][nid], tree["right"][nid], tree["default_left"][nid] = li, ri, bool(default_left)
        splits.append((nid, int(f), int(b), bool(default_left), li, ri))
    return splits


def _new_tree() -> Dict[str, list]:
    return {"feature": [-1], "threshold": [0.0], "left": [-1], "right": [-1], "default_left": [True], "value": [0.0]}


def _grow_tree(binned: np.ndarray, n_bins: int, cuts: List[np.ndarray], g: np.ndarray, h: np.ndarray,
               rows: np.ndarray, p: Dict[str, float], comm: _Comm) -> Tuple[Dict[str, list], np.ndarray]:
    """One tree, level-wise.  Returns the tree and the leaf index of every LOCAL row (all rows, so the
    margin update needs no second traversal)."""
    n, F = binned.shape
    tree = _new_tree()
    node_of = np.zeros(n, dtype=np.int64)
    in_sample = np.zeros(n, dtype=bool)
    in_sample[rows] = True
    level = [0]
    offsets = (np.arange(F, dtype=np.int64) * n_bins)[None, :]
    miss = n_bins - 1
    for depth in range(int(p["max_depth"]) + 1):
        if not level:
            break
        slot = {nid: i for i, nid in enumerate(level)}
        hist = np.zeros((len(level), 2, F * n_bins))
        sel = np.nonzero(in_sample & np.isin(node_of, level))[0]
        if len(sel):
            s = np.vectorize(slot.get, otypes=[np.int64])(node_of[sel]) if len(level) > 1 else np.zeros(len(sel), np.int64)
            flat = (binned[sel].astype(np.int64) + offsets) + (s * (F * n_bins))[:, None]
            size = len(level) * F * n_bins
            hg = np.bincount(flat.ravel(), weights=np.repeat(g[sel], F), minlength=size)
            hh = np.bincount(flat.ravel(), weights=np.repeat(h[sel], F), minlength=size)
            hist[:, 0, :] = hg.reshape(len(level), F * n_bins)
            hist[:, 1, :] = hh.reshape(len(level), F * n_bins)
        hist = comm.allreduce(hist).reshape(len(level), 2, F, n_bins)         # ONE collective per level
        nxt = []
        for nid, f, b, default_left, li, ri in _split_level(hist, level, tree, depth, cuts, n_bins, p):
            here = np.nonzero(node_of == nid)[0]
            bins = binned[here, f]
            go_left = np.where(bins == miss, default_left, bins <= b)
            node_of[here] = np.where(go_left, li, ri)
            nxt += [li, ri]
        level = nxt
    return tree, node_of


class _GpuRows:
    """The row-sized state of the booster on the GPU (``tree_method="gpu_hist"``): feature-major bin ids, sampled
    mask, node and histogram-slot of every row.  The two passes over the rows of every tree level -- histogram
    build and row partition -- run as CUDA kernels (``csrc/kernels/gbt.cu``); the small per-level histogram comes
    back to the host for the fp64 split search and the cross-worker all-reduce."""

    def __init__(self, binned: Optional[np.ndarray], n_bins: int, device, x: Optional[np.ndarray] = None,
                 cuts: Optional[List[np.ndarray]] = None):
        """From host bin ids (``binned`` [n, F] uint16), or -- ``binned=None`` -- straight from the raw features:
        ``x`` goes to the device once and is binned there (``torch.bucketize`` per feature, NaN -> the missing bin),
        feature-major, so the host never builds or transposes the bin matrix."""
        import torch
        self.torch = torch
        self.device = device
        self.n_bins = n_bins
        if binned is not None:
            self.n, self.F = binned.shape
            self.binned_t = torch.from_numpy(np.ascontiguousarray(binned.T).view(np.int16)).to(device)
        else:
            self.n, self.F = x.shape
            xd = x if torch.is_tensor(x) else torch.from_numpy(x).to(device)
            self.binned_t = torch.empty(self.F, self.n, dtype=torch.int16, device=device)
            for f, c in enumerate(cuts):
                col = xd[:, f].contiguous()
                b = torch.bucketize(col, torch.from_numpy(np.asarray(c, dtype=np.float32)).to(device), right=True)
                b = torch.where(torch.isnan(col), torch.full_like(b, n_bins - 1), b)
                self.binned_t[f] = b.to(torch.int16)
            del xd
        self.node_of = torch.zeros(self.n, dtype=torch.int32, device=device)
        self.slot = torch.zeros(self.n, dtype=torch.int32, device=device)
        self.in_sample = torch.ones(self.n, dtype=torch.uint8, device=device)

    def grow(self, cuts: List[np.ndarray], g: np.ndarray, h: np.ndarray, rows: np.ndarray, p: Dict[str, float],
             comm: _Comm) -> Tuple[Dict[str, list], np.ndarray]:
        from raydp_b200 import ops
        torch = self.torch
        dev = self.device
        tree = _new_tree()
        gd = torch.from_numpy(np.ascontiguousarray(g, dtype=np.float32)).to(dev)
        hd = torch.from_numpy(np.ascontiguousarray(h, dtype=np.float32)).to(dev)
        if len(rows) == self.n:
            self.in_sample.fill_(1)
        else:
            mask = np.zeros(self.n, dtype=np.uint8)
            mask[rows] = 1
            self.in_sample.copy_(torch.from_numpy(mask))
        self.node_of.zero_()
        self.slot.copy_(self.in_sample.to(torch.int32) - 1)         # level 0: slot 0 for sampled rows, -1 otherwise
        level = [0]
        miss = self.n_bins - 1
        i32 = dict(dtype=torch.int32, device=dev)
        for depth in range(int(p["max_depth"]) + 1):
            if not level:
                break
            hist_d = torch.zeros(len(level), 2, self.F, self.n_bins, dtype=torch.float64, device=dev)
            if self.n:
                ops.gbt_hist(self.binned_t, gd, hd, self.slot, hist_d)
            hist = comm.allreduce(hist_d.cpu().numpy().reshape(len(level), 2, self.F * self.n_bins))
            hist = hist.reshape(len(level), 2, self.F, self.n_bins)
            splits = _split_level(hist, level, tree, depth, cuts, self.n_bins, p)
            nodes = len(tree["feature"])
            feat = np.full(nodes, -1, dtype=np.int32)
            thr = np.zeros(nodes, dtype=np.int32)
            left = np.zeros(nodes, dtype=np.int32)
            right = np.zeros(nodes, dtype=np.int32)
            dl = np.zeros(nodes, dtype=np.uint8)
            nxt_slot = np.full(nodes, -1, dtype=np.int32)
            nxt = []
            for nid, f, b, default_left, li, ri in splits:
                feat[nid], thr[nid], left[nid], right[nid], dl[nid] = f, b, li, ri, int(default_left)
                nxt_slot[li], nxt_slot[ri] = len(nxt), len(nxt) + 1
                nxt += [li, ri]
            if self.n:
                up = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
                ops.gbt_partition(self.binned_t, self.node_of, self.slot, self.in_sample, up(feat), up(thr), up(left),
                                  up(right), up(dl), up(nxt_slot), miss)
            level = nxt
        return tree, self.node_of.cpu().numpy().astype(np.int64)


def _gpu_device(p: Dict[str, Any]):
    """The CUDA device to boost on, or None.  A GPU is demanded by ``tree_method="gpu_hist"`` or ``device="cuda"``
    (the xgboost 1.x and 2.x spellings) and refused by ``device="cpu"``; otherwise the histogram kernels are used
    whenever this worker process has a CUDA device and the kernels library."""
    method, where = str(p.get("tree_method") or "auto"), str(p.get("device") or "").lower()
    wanted = method == "gpu_hist" or where.startswith(("cuda", "gpu"))
    if where == "cpu" and not wanted:
        return None
    try:
        import torch
        from raydp_b200 import ops
        if torch.cuda.is_available() and ops.kernels_available():
            return torch.device("cuda", torch.cuda.current_device())
    except Exception:  # noqa: BLE001 - no torch / no driver: CPU path
        pass
    if wanted:
        raise RuntimeError('tree_method="gpu_hist" / device="cuda" need a CUDA device and the built kernels library')
    return None


def normalise_params(params: Dict[str, Any]) -> Dict[str, Any]:
    unknown = sorted(set(params) - _KNOWN)
    if unknown:
        raise ValueError(f"xgboost parameters {unknown} are not supported by the built-in distributed booster")
    tm = params.get("tree_method", "auto")
    if tm not in ("hist", "auto", "approx", "gpu_hist"):
        raise ValueError(f"tree_method={tm!r}: the built-in distributed booster implements 'hist' / 'gpu_hist'")
    return {"tree_method": str(tm), "device": str(params.get("device") or ""),
            "objective": str(params.get("objective", "reg:squarederror")),
            "eta": float(params.get("eta", params.get("learning_rate", 0.3))),
            "max_depth": int(params.get("max_depth", 6)),
            "lambda": float(params.get("lambda", params.get("reg_lambda", 1.0))),
            "gamma": float(params.get("gamma", params.get("min_split_loss", 0.0))),
            "min_child_weight": float(params.get("min_child_weight", 1.0)),
            "max_bin": int(params.get("max_bin", 256)),
            "base_score": params.get("base_score"),
            "subsample": float(params.get("subsample", 1.0)),
            "seed": int(params.get("seed", params.get("random_state", 0))),
            "eval_metric": params.get("eval_metric")}


def train(params: Dict[str, Any], x: np.ndarray, y: np.ndarray, num_boost_round: int,
          evals: Sequence[Tuple[np.ndarray, np.ndarray, str]] = (), feature_names: Optional[Sequence[str]] = None,
          resume: Optional[HistGBTModel] = None, callback=None) -> Tuple[HistGBTModel, Dict[str, Dict[str, List[float]]]]:
    """Boost on this worker's shard ``(x, y)``; call it on every worker of the group (or alone).
    ``callback(round, model, metrics)`` runs after every round on every rank."""
    p = normalise_params(params)
    comm = _Comm()
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.asarray(y, dtype=np.float64)
    grad_hess, _link, default_metric = _objective(p["objective"])
    metrics = p["eval_metric"] or default_metric
    metrics = [metrics] if isinstance(metrics, str) else list(metrics)
    if resume is not None:
        model = resume
    else:
        if p["base_score"] is not None:
            base = float(p["base_score"])
        else:
            tot = comm.allreduce(np.array([y.sum(), float(len(y))]))
            base = tot[0] / max(1.0, tot[1])
        if p["objective"] in ("binary:logistic", "reg:logistic"):
            base = min(max(base, 1e-6), 1 - 1e-6)
            base = float(np.log(base / (1 - base)))
        model = HistGBTModel(p["objective"], base, feature_names)
    device = _gpu_device(p)
    if device is not None:
        import torch
        xd = torch.from_numpy(x).to(device)                                     # the features go to the GPU once
        cuts = _global_cuts(x, p["max_bin"], comm, xd=xd)
        binned, n_bins = None, max((len(c) for c in cuts), default=0) + 2       # same bin count as _bin()
        gpu_rows = _GpuRows(None, n_bins, device, x=xd, cuts=cuts)
        del xd
    else:
        cuts = _global_cuts(x, p["max_bin"], comm)
        binned, n_bins = _bin(x, cuts)
        gpu_rows = None
    model.attributes["tree_method"] = "gpu_hist" if gpu_rows is not None else "hist"
    margin = model.margin(x)
    eval_margins = [model.margin(np.asarray(ex, dtype=np.float32)) for ex, _, _ in evals]
    history: Dict[str, Dict[str, List[float]]] = {}
    link = _objective(p["objective"])[1]
    for rnd in range(len(model.trees), num_boost_round):
        g, h = grad_hess(margin, y)
        if p["subsample"] < 1.0:
            rng = np.random.default_rng([p["seed"], rnd, comm.rank])
            rows = np.nonzero(rng.random(len(y)) < p["subsample"])[0]
        else:
            rows = np.arange(len(y))
        if gpu_rows is not None:
            tree, leaf = gpu_rows.grow(cuts, g, h, rows, p, comm)
        else:
            tree, leaf = _grow_tree(binned, n_bins, cuts, g, h, rows, p, comm)
        model.trees.append(tree)
        margin += np.asarray(tree["value"])[leaf]
        round_metrics: Dict[str, float] = {}
        sets = [("train", margin, y)] + [(name, None, ey) for _, ey, name in evals]
        for k, (ex, _, _) in enumerate(evals):
            eval_margins[k] += _tree_predict(tree, np.asarray(ex, dtype=np.float32))
        for k, (name, m, ty) in enumerate(sets):
            m = margin if k == 0 else eval_margins[k - 1]
            for metric in metrics:
                sums = comm.allreduce(_metric_sums(metric, link(m), np.asarray(ty, dtype=np.float64)))
                value = _metric_value(metric, sums)
                history.setdefault(name, {}).setdefault(metric, []).append(value)
                round_metrics[f"{name}-{metric}"] = value
        if callback is not None:
            callback(rnd, model, round_metrics)
    return model, history
