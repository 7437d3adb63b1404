"""Lowering of plain ``nn.Module``s onto the fused sm_100a engines.

The reference hands whatever module the user wrote to DDP (``python/raydp/torch/estimator.py:180-187,
243``): its flagship example defines ``NYC_Model`` inline (``examples/pytorch_nyctaxi.py:69-92``) and the
DLRM notebook defines ``Dlrm`` / ``MultiTableEmbeddings`` / ``DotInteraction`` inline
(``examples/pytorch_dlrm.ipynb`` cell 9).  Neither knows anything about this framework, so the engines
must recognise the computation from the module itself:

* **MLP chains** are recognised exactly, from the ``torch.fx`` graph of the module: a single-input chain
  ``(Linear -> ReLU [-> BatchNorm1d])* -> Linear(-> 1)`` (ReLU as module, function or method; identity /
  ``Dropout(p=0)`` / ``.float()`` nodes are skipped).  The result is an :class:`MLPSpec` whose
  ``linear_names`` / ``bn_names`` are the user's own sub-module paths, so the ``model.pt`` written after
  training loads straight back into the user's class.
* **DLRM-shaped models** take ``*inputs`` and cannot be traced; they are recognised structurally
  (``nn.Embedding`` tables of one width, a bottom ``Linear`` chain ending at that width, a top chain
  starting at ``S(S-1)/2 + D + pad`` and ending at one logit, nothing else that holds parameters) and the
  guess is then *proved* on probe batches: the restatement that the fused kernels implement (bottom MLP,
  lookups, pairwise dots in lower-triangular row-major order, concat, top MLP) must reproduce the
  module's own output.  A model that merely looks similar fails the probe and runs on the generic engine.

``lower_module`` returns ``(plan, reason)``; ``plan is None`` means "not one of the fully fused shapes",
and the caller falls back to :class:`raydp_b200.torch.generic_engine.GenericGpuEngine`, which still runs
every ``nn.Linear`` on the tcgen05 kernels.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


class NotLowerable(Exception):
    pass


@dataclass
class MLPPlan:
    spec: Any                      # raydp_b200.ops.fused_mlp.MLPSpec
    kind: str = "mlp"
    how: str = "fx"


@dataclass
class DLRMPlan:
    spec: Any                      # raydp_b200.models.dlrm.DLRMSpec
    embedding_names: List[str] = field(default_factory=list)
    bottom_names: List[str] = field(default_factory=list)
    top_names: List[str] = field(default_factory=list)
    per_column_inputs: bool = True     # model(*columns) (the notebook) vs model(dense, sparse)
    kind: str = "dlrm"
    how: str = "structure+probe"

    def modules(self, model: nn.Module):
        get = model.get_submodule
        return ([get(n) for n in self.bottom_names], [get(n) for n in self.top_names],
                [get(n) for n in self.embedding_names])

    def rename_state(self, state: dict) -> dict:
        """Engine-side names (``bottom.i`` / ``top.i`` / ``embeddings.t``) -> the user's parameter names."""
        out = {}
        for i, n in enumerate(self.bottom_names):
            out[f"{n}.weight"], out[f"{n}.bias"] = state[f"bottom.{i}.weight"], state[f"bottom.{i}.bias"]
        for i, n in enumerate(self.top_names):
            out[f"{n}.weight"], out[f"{n}.bias"] = state[f"top.{i}.weight"], state[f"top.{i}.bias"]
        for t, n in enumerate(self.embedding_names):
            if f"embeddings.{t}.weight" in state:
                out[f"{n}.weight"] = state[f"embeddings.{t}.weight"]
        return out


# ---------------------------------------------------------------------------------------------- MLP
_RELU_FUNCS = {F.relu, torch.relu, torch.relu_, torch.nn.functional.relu_}
_SKIP_METHODS = {"float", "contiguous", "detach_"}
_RELU_METHODS = {"relu", "relu_"}


def _classify(node, modules) -> Tuple[str, Optional[str]]:
    if node.op == "call_module":
        m = modules[node.target]
        if isinstance(m, nn.Linear):
            return "linear", node.target
        if isinstance(m, nn.ReLU):
            return "relu", None
        if isinstance(m, nn.BatchNorm1d):
            return "bn", node.target
        if isinstance(m, nn.Identity) or (isinstance(m, nn.Dropout) and m.p == 0.0):
            return "skip", None
        raise NotLowerable(f"module {node.target} ({type(m).__name__}) is not part of a fusable MLP chain")
    if node.op == "call_function":
        if node.target in _RELU_FUNCS:
            return "relu", None
        if node.target is F.dropout and (node.kwargs.get("p", node.args[1] if len(node.args) > 1 else 0.5) == 0.0):
            return "skip", None
        raise NotLowerable(f"function {getattr(node.target, '__name__', node.target)} is not part of a fusable MLP chain")
    if node.op == "call_method":
        if node.target in _RELU_METHODS:
            return "relu", None
        if node.target in _SKIP_METHODS:
            return "skip", None
        raise NotLowerable(f"tensor method .{node.target}() is not part of a fusable MLP chain")
    raise NotLowerable(f"graph node {node.op} is not part of a fusable MLP chain")


def _mlp_plan_fx(model: nn.Module) -> MLPPlan:
    import torch.fx as fx
    from raydp_b200.ops.fused_mlp import MLPSpec
    try:
        gm = fx.symbolic_trace(model)
    except Exception as exc:  # noqa: BLE001 - any tracing failure means "not a static chain"
        raise NotLowerable(f"torch.fx cannot trace the module ({type(exc).__name__}: {exc})") from None
    modules = dict(gm.named_modules())
    cur, steps, done = None, [], False
    for node in gm.graph.nodes:
        if node.op == "placeholder":
            if cur is not None:
                raise NotLowerable("the module takes more than one input")
            cur = node
            continue
        if cur is None:
            raise NotLowerable("the module has no tensor input")
        if node.op == "output":
            if node.args[0] is not cur:
                raise NotLowerable("the output is not the end of a single chain")
            done = True
            break
        tensor_in = [a for a in list(node.args) + list(node.kwargs.values()) if isinstance(a, fx.Node)]
        if tensor_in != [cur]:
            raise NotLowerable("the graph is not a single chain of layers")
        kind, name = _classify(node, modules)
        if kind != "skip":
            steps.append((kind, name))
        cur = node
    if not done:
        raise NotLowerable("no output node")
    # grammar: (linear relu [bn])* linear
    lin_names, bn_names, has_bn, i = [], [], [], 0
    while True:
        if i >= len(steps) or steps[i][0] != "linear":
            raise NotLowerable("expected a Linear layer in the chain")
        lin_names.append(steps[i][1])
        i += 1
        if i == len(steps):
            break                                   # that was the head
        if steps[i][0] != "relu":
            raise NotLowerable("a hidden Linear must be followed by ReLU (Linear -> ReLU -> BatchNorm1d order)")
        i += 1
        if i < len(steps) and steps[i][0] == "bn":
            bn_names.append(steps[i][1])
            has_bn.append(True)
            i += 1
        else:
            has_bn.append(False)
    if len(lin_names) < 2:
        raise NotLowerable("a single Linear is not an MLP")
    if any(has_bn) and not all(has_bn):
        raise NotLowerable("BatchNorm1d after some hidden layers but not all")
    lins = [modules[n] for n in lin_names]
    if len(set(lin_names)) != len(lin_names) or len(set(bn_names)) != len(bn_names):
        raise NotLowerable("a layer is applied twice (shared weights)")
    for a, b in zip(lins, lins[1:]):
        if a.out_features != b.in_features:
            raise NotLowerable("layer widths do not chain")
    if any(l.bias is None for l in lins):
        raise NotLowerable("Linear(bias=False) is not on the fused path")
    bns = [modules[n] for n in bn_names]
    for l, bn in zip(lins, bns):
        if bn.num_features != l.out_features or not bn.affine or not bn.track_running_stats or bn.momentum is None:
            raise NotLowerable("BatchNorm1d must be affine, track running stats and use a fixed momentum")
    if bns and (len({bn.eps for bn in bns}) > 1 or len({bn.momentum for bn in bns}) > 1):
        raise NotLowerable("BatchNorm layers differ in eps / momentum")
    claimed = {id(p) for m in lins + bns for p in m.parameters()}
    if any(id(p) not in claimed for p in model.parameters()):
        raise NotLowerable("the module owns parameters outside the recognised chain")
    try:
        spec = MLPSpec(lins[0].in_features, [l.out_features for l in lins[:-1]], lins[-1].out_features,
                       batch_norm=bool(bns), bn_eps=bns[0].eps if bns else 1e-5,
                       bn_momentum=bns[0].momentum if bns else 0.1, linear_names=lin_names, bn_names=bn_names)
    except ValueError as exc:
        raise NotLowerable(str(exc)) from None
    return MLPPlan(spec)


# --------------------------------------------------------------------------------------------- DLRM
_PARAM_FREE_LEAVES = (nn.ReLU, nn.Identity)


def _r8(v: int) -> int:
    return (v + 7) // 8 * 8


def dlrm_restatement(bottom: Sequence[nn.Linear], top: Sequence[nn.Linear], embs: Sequence[nn.Embedding],
                     dense: torch.Tensor, sparse: torch.Tensor) -> torch.Tensor:
    """The computation the fused DLRM kernels implement, in plain torch, on the user's own layers."""
    x = dense
    for lin in bottom:
        x = F.relu(lin(x))
    feats = torch.stack([x] + [e(sparse[:, t]) for t, e in enumerate(embs)], dim=1)
    S = feats.shape[1]
    inter = torch.bmm(feats, feats.transpose(1, 2))
    li, lj = torch.tril_indices(S, S, offset=-1)
    flat = inter[:, li, lj]
    width = top[0].in_features
    pad = torch.zeros(x.shape[0], width - flat.shape[1] - x.shape[1], dtype=x.dtype, device=x.device)
    z = torch.cat([x, flat, pad], dim=1)
    for i, lin in enumerate(top):
        z = lin(z)
        if i < len(top) - 1:
            z = F.relu(z)
    return z


def _dlrm_plan(model: nn.Module) -> DLRMPlan:
    from raydp_b200.models.dlrm import DLRMSpec
    embs = [(n, m) for n, m in model.named_modules() if isinstance(m, nn.Embedding)]
    lins = [(n, m) for n, m in model.named_modules() if isinstance(m, nn.Linear)]
    if not embs or len(lins) < 2:
        raise NotLowerable("no embedding tables")
    for _, m in model.named_modules():
        if m is not model and not list(m.children()) \
                and not isinstance(m, (nn.Embedding, nn.Linear) + _PARAM_FREE_LEAVES):
            raise NotLowerable(f"layer type {type(m).__name__} is not part of the DLRM shape")
    claimed = {id(p) for _, m in embs + lins for p in m.parameters()}
    if any(id(p) not in claimed for p in model.parameters()):
        raise NotLowerable("the module owns parameters outside Embedding / Linear layers")
    D = embs[0][1].embedding_dim
    for _, e in embs:
        if e.embedding_dim != D or e.padding_idx is not None or e.max_norm is not None or e.scale_grad_by_freq:
            raise NotLowerable("embedding tables must share one width and use plain lookups")
    if any(l.bias is None for _, l in lins):
        raise NotLowerable("Linear(bias=False) is not on the fused path")
    split = next((i for i in range(1, len(lins)) if lins[i][1].in_features != lins[i - 1][1].out_features), None)
    if split is None:
        raise NotLowerable("no bottom / top MLP boundary found")
    bottom, top = lins[:split], lins[split:]
    for a, b in zip(top, top[1:]):
        if a[1].out_features != b[1].in_features:
            raise NotLowerable("top MLP widths do not chain")
    S = len(embs) + 1
    inter_w = _r8(S * (S - 1) // 2 + D + 1)
    if bottom[-1][1].out_features != D:
        raise NotLowerable("the bottom MLP must end at the embedding width")
    if top[0][1].in_features != inter_w:
        raise NotLowerable(f"the top MLP must start at the padded interaction width {inter_w}")
    if top[-1][1].out_features != 1:
        raise NotLowerable("the top MLP must end at one logit")
    if D % 8 != 0 or D > 256:
        raise NotLowerable("embedding width must be a multiple of 8 (<= 256)")
    plan = DLRMPlan(DLRMSpec(bottom[0][1].in_features, tuple(e.num_embeddings for _, e in embs), D,
                             tuple(l.out_features for _, l in bottom), tuple(l.out_features for _, l in top)),
                    [n for n, _ in embs], [n for n, _ in bottom], [n for n, _ in top])
    _probe_dlrm(model, plan)
    return plan


def _probe_dlrm(model: nn.Module, plan: DLRMPlan) -> None:
    """Run the module and the restatement on probe batches; accept the plan only if they agree."""
    bottom, top, embs = plan.modules(model)
    p = next(model.parameters())
    was_training = model.training
    g = torch.Generator().manual_seed(20240607)
    ok_convention = None
    try:
        model.eval()
        with torch.no_grad():
            for trial in range(2):
                B = 5 + trial
                dense = torch.randn(B, plan.spec.num_dense, generator=g).to(p.device, p.dtype)
                sparse = torch.stack([torch.randint(0, n, (B,), generator=g) for n in plan.spec.table_sizes], 1).to(p.device)
                want = dlrm_restatement(bottom, top, embs, dense, sparse)
                got = None
                conventions = [True, False] if ok_convention is None else [ok_convention]
                for per_column in conventions:
                    try:
                        if per_column:
                            cols = [dense[:, i:i + 1] for i in range(dense.shape[1])] + \
                                   [sparse[:, t:t + 1] for t in range(sparse.shape[1])]
                            out = model(*cols)
                        else:
                            out = model(dense, sparse)
                    except Exception:  # noqa: BLE001 - wrong calling convention for this module
                        continue
                    if isinstance(out, torch.Tensor) and out.shape == want.shape and \
                            torch.allclose(out, want, rtol=1e-4, atol=1e-5 * (1.0 + float(want.abs().max()))):
                        got, ok_convention = out, per_column
                        break
                if got is None:
                    raise NotLowerable("the module's output differs from the DLRM restatement on a probe batch")
    finally:
        model.train(was_training)
    plan.per_column_inputs = bool(ok_convention)


# ------------------------------------------------------------------------------------------- public
def lower_module(model: nn.Module) -> Tuple[Optional[Any], str]:
    """``(plan, "")`` for a module one of the fully fused engines can run, else ``(None, why)``."""
    if hasattr(model, "fused_spec"):                     # the module describes itself
        from raydp_b200.models.dlrm import DLRMSpec
        try:
            spec = model.fused_spec()
        except ValueError as exc:
            return None, str(exc)
        if isinstance(spec, DLRMSpec):
            T = len(spec.table_sizes)
            return DLRMPlan(spec, [f"embeddings.{t}" for t in range(T)],
                            [f"bottom.{i}" for i in range(len(spec.bottom))],
                            [f"top.{i}" for i in range(len(spec.top))], True, how="fused_spec"), ""
        return MLPPlan(spec, how="fused_spec"), ""
    has_embedding = any(isinstance(m, (nn.Embedding, nn.EmbeddingBag)) for m in model.modules())
    try:
        return (_dlrm_plan(model) if has_embedding else _mlp_plan_fx(model)), ""
    except NotLowerable as exc:
        return None, str(exc)
