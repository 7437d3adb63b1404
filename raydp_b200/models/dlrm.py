"""DLRM, Criteo-shaped (workload B of SURVEY.md section 2.6).

Parity: reference ``examples/pytorch_dlrm.ipynb`` cell 9 (``MultiTableEmbeddings``, ``TorchMlp``,
``DotInteraction``, ``Dlrm``) and cell 14 (13 dense + 26 categorical features, ``embedding_dim=32``,
bottom MLP 13 -> 512 -> 128 -> 32, top MLP 384 -> 1024 -> 1024 -> 512 -> 256 -> 1 where
384 = 27*26/2 + 32 + 1 pad, ``BCEWithLogitsLoss``, SGD lr 0.01).  Plain torch; ``fused_spec()`` lets the
GPU engine run it on the hybrid-parallel fused path (dense layers data-parallel, embedding tables
sharded table-wise with the all-to-all fused into the lookup).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class DLRMSpec:
    num_dense: int
    table_sizes: Sequence[int]
    embedding_dim: int
    bottom: Sequence[int]
    top: Sequence[int]
    # multi-hot tables: table t reads bag_sizes[t] consecutive index columns per sample (negative = padding)
    # and pools them ("sum" or "mean"); None = one index per table, as in the reference notebook
    bag_sizes: Optional[Sequence[int]] = None
    pooling: str = "sum"

    @property
    def index_columns(self) -> int:
        return sum(self.bag_sizes) if self.bag_sizes else len(self.table_sizes)


class DLRM(nn.Module):
    def __init__(self, num_dense: int = 13, table_sizes: Sequence[int] = (1000,) * 26, embedding_dim: int = 32,
                 bottom: Sequence[int] = (512, 128, 32), top: Sequence[int] = (1024, 1024, 512, 256, 1),
                 sparse_grad: bool = True, bag_sizes: Optional[Sequence[int]] = None, pooling: str = "sum"):
        super().__init__()
        assert pooling in ("sum", "mean")
        assert bag_sizes is None or len(bag_sizes) == len(table_sizes)
        self.bag_sizes = tuple(int(b) for b in bag_sizes) if bag_sizes else None
        self.pooling = pooling
        assert bottom[-1] == embedding_dim, "the bottom MLP must end at the embedding width"
        self.num_dense, self.table_sizes, self.embedding_dim = num_dense, tuple(table_sizes), embedding_dim
        self.bottom_dims, self.top_dims = tuple(bottom), tuple(top)
        self.embeddings = nn.ModuleList(nn.Embedding(n, embedding_dim, sparse=sparse_grad) for n in table_sizes)
        dims = [num_dense] + list(bottom)
        self.bottom = nn.ModuleList(nn.Linear(dims[i], dims[i + 1]) for i in range(len(bottom)))
        S = 1 + len(table_sizes)
        self.num_interactions = S * (S - 1) // 2 + embedding_dim
        self.interaction_width = (self.num_interactions + 1 + 7) // 8 * 8       # + zero pad (reference pads 1)
        tdims = [self.interaction_width] + list(top)
        self.top = nn.ModuleList(nn.Linear(tdims[i], tdims[i + 1]) for i in range(len(top)))
        li, lj = torch.tril_indices(S, S, offset=-1)
        self.register_buffer("_li", li, persistent=False)
        self.register_buffer("_lj", lj, persistent=False)

    def forward(self, dense, sparse=None, *more):
        if sparse is not None and more:            # per-column tensors: 13 x [B,1] float, 26 x [B,1] long
            cols = [dense, sparse, *more]
            dense = torch.cat(cols[:self.num_dense], dim=1)
            sparse = torch.cat(cols[self.num_dense:], dim=1)
        x = dense
        for lin in self.bottom:
            x = F.relu(lin(x))
        if self.bag_sizes is None:
            emb = [e(sparse[:, t].long()) for t, e in enumerate(self.embeddings)]
        else:
            emb, col = [], 0
            for e, n in zip(self.embeddings, self.bag_sizes):
                bag = sparse[:, col:col + n].long()
                col += n
                keep = (bag >= 0).to(x.dtype).unsqueeze(-1)
                pooled = (e(bag.clamp_min(0)) * keep).sum(1)
                if self.pooling == "mean":
                    pooled = pooled / keep.sum(1).clamp_min(1.0)
                emb.append(pooled)
        feats = torch.stack([x] + emb, dim=1)
        inter = torch.bmm(feats, feats.transpose(1, 2))
        flat = inter[:, self._li, self._lj]
        pad = torch.zeros(x.shape[0], self.interaction_width - self.num_interactions, device=x.device, dtype=x.dtype)
        z = torch.cat([x, flat, pad], dim=1)
        for i, lin in enumerate(self.top):
            z = lin(z)
            if i < len(self.top) - 1:
                z = F.relu(z)
        return z

    def fused_spec(self) -> DLRMSpec:
        return DLRMSpec(self.num_dense, self.table_sizes, self.embedding_dim, self.bottom_dims, self.top_dims,
                        self.bag_sizes, self.pooling)
