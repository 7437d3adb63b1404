"""``TorchMLDataset`` / ``PrefetchedDataLoader`` under their reference names.

Parity: reference ``python/raydp/torch/torch_ml_dataset.py:30-111``.  The reference classes sit on
``ray.util.data.MLDataset``, which Ray removed; nothing live imports them.
Here the same two names work against the current data layer: ``TorchMLDataset`` is a
``torch.utils.data.IterableDataset`` over one shard of a :class:`raydp_b200.data.Dataset`
(yielding per-row ``(features, label)`` tensors, so ``DataLoader(batch_size=...)`` collates them),
and ``PrefetchedDataLoader`` is the background-thread prefetcher with a bounded queue.
"""
from __future__ import annotations

from typing import Any, List, Optional

import torch
from torch.utils.data import IterableDataset

from raydp_b200.data.torch_iter import PrefetchedLoader, TorchBatchIterable


class TorchMLDataset(IterableDataset):
    def __init__(self, ds, world_size: int = 1, world_rank: int = 0, shuffle: bool = False,
                 shuffle_seed: Optional[int] = None, feature_columns: Optional[List[str]] = None,
                 feature_types: Any = torch.float, label_column: Optional[str] = None,
                 label_type: Any = torch.float, block_rows: int = 4096):
        super().__init__()
        if not 0 <= world_rank < world_size:
            raise ValueError(f"world_rank {world_rank} out of range for world_size {world_size}")
        if shuffle:
            ds = ds.random_shuffle(shuffle_seed)
        self._shard = ds.split(world_size, equal=True)[world_rank] if world_size > 1 else ds
        self._kw = dict(label_column=label_column, feature_columns=feature_columns,
                        label_column_dtype=label_type, feature_column_dtypes=feature_types,
                        batch_size=block_rows, drop_last=False)

    def __len__(self) -> int:
        return self._shard.count()

    def __iter__(self):
        for feats, label in TorchBatchIterable(self._shard, **self._kw):
            if isinstance(feats, (list, tuple)):
                rows = zip(*[f.unbind(0) for f in feats])
                feats_iter = ([*r] for r in rows)
            else:
                feats_iter = iter(feats.unbind(0))
            if label is None:
                yield from feats_iter
            else:
                yield from zip(feats_iter, label.unbind(0))


class PrefetchedDataLoader(PrefetchedLoader):
    """``PrefetchedDataLoader(base_data_loader, max_size=5)`` -- reference ``:69-111``."""

    def __init__(self, base_data_loader, max_size: int = 5):
        super().__init__(base_data_loader, max_size=max_size)
