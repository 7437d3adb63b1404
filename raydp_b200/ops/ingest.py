"""GPU ingestion pipeline: pinned Arrow blocks in shared memory -> training tensors on device.

Replaces ``dataset.to_torch(...)`` of the reference (``torch/estimator.py:226-232``: Arrow ->
pandas -> numpy -> ``torch.as_tensor`` -> concat/cast on the CPU, then a blocking copy in any
GPU-patched loop).  Here the host only *addresses* data:

1. every block of the shard is mapped from the object store and page-locked once
   (``cudaHostRegister`` on the shared-memory mapping, so Arrow column buffers are DMA-able
   in place);
2. per chunk of rows, one ``cudaMemcpyAsync`` per column moves the raw column bytes to a
   double-buffered device staging area on a side stream;
3. the ingest kernel (``rdp_ingest_pack`` / ``rdp_ingest_indices``) decodes the mixed-dtype
   columns, normalises, zero-pads and writes the row-major bf16 feature matrix that the
   tcgen05 GEMM reads through TMA, plus the fp32 label vector, on the compute stream.

Batches may straddle chunk and block boundaries; they are assembled directly in the output
buffer by packing at a row offset, so there is no concat.
"""
from __future__ import annotations

from typing import Iterator, List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import torch

from raydp_b200 import ops

_ARROW_CODES = {pa.float32(): (0, 4), pa.float64(): (1, 8), pa.int32(): (2, 4), pa.int64(): (3, 8),
                pa.int16(): (4, 2), pa.int8(): (5, 1), pa.uint8(): (5, 1)}


def _plain_array(col: pa.ChunkedArray) -> List[pa.Array]:
    """Fixed-width, null-free chunks whose data buffer can be copied as raw bytes."""
    out = []
    for chunk in col.chunks:
        t = chunk.type
        if pa.types.is_boolean(t):
            chunk = chunk.cast(pa.int8())
        elif pa.types.is_timestamp(t) or pa.types.is_date64(t):
            chunk = chunk.cast(pa.int64())
        elif pa.types.is_date32(t) or pa.types.is_uint16(t):
            chunk = chunk.cast(pa.int32())
        elif pa.types.is_uint32(t) or pa.types.is_uint64(t):
            chunk = chunk.cast(pa.int64())
        elif pa.types.is_float16(t):
            chunk = chunk.cast(pa.float32())
        if chunk.null_count:
            chunk = chunk.fill_null(0)
        if chunk.type not in _ARROW_CODES:
            raise TypeError(f"column type {chunk.type} cannot be ingested; encode it on the executors "
                            "(strings/dictionaries must become numeric ids before training)")
        out.append(chunk)
    return out


class _ColumnSpan:
    """A run of rows of one column that is contiguous in (pinned) host memory."""
    __slots__ = ("address", "rows", "code", "width", "keep")

    def __init__(self, arr: pa.Array):
        if arr.type not in _ARROW_CODES:
            raise TypeError(f"the GPU ingest path reads fixed-width numeric columns, not {arr.type}: expand vector "
                            "columns into scalar columns (F.col('features')[i]) or use a model without fused_spec()")
        code, width = _ARROW_CODES[arr.type]
        buf = arr.buffers()[1]
        self.address = buf.address + arr.offset * width
        self.rows = len(arr)
        self.code, self.width = code, width
        self.keep = arr            # keeps the Arrow buffer (and the shm mapping) alive


class ArrowGpuLoader:
    """Iterable of ``(features, label)`` device tensors for one dataset shard.

    ``features`` is bf16 ``[B, ld]`` (zero padded to ``ld``), or int32 ``[B, n_sparse]`` index rows
    when ``sparse_columns`` is given the iterator yields ``(dense, sparse, label)``.
    Output tensors are reused between iterations (consume them before the next ``next()``).
    """

    def __init__(self, dataset, feature_columns: Sequence[str], label_column: Optional[str], batch_size: int,
                 drop_last: bool, device: torch.device, ld: Optional[int] = None,
                 out_dtype: torch.dtype = torch.bfloat16, chunk_rows: Optional[int] = None,
                 sparse_columns: Optional[Sequence[str]] = None, table_rows: Optional[Sequence[int]] = None,
                 shift: Optional[torch.Tensor] = None, scale: Optional[torch.Tensor] = None):
        self.ds = dataset
        self.features = list(feature_columns)
        self.sparse = list(sparse_columns or [])
        self.table_rows = list(table_rows or [0] * len(self.sparse))
        self.label = label_column
        self.batch = int(batch_size)
        self.drop_last = drop_last
        self.device = device
        self.ld = ld or (max(8, len(self.features)) + 7) // 8 * 8
        self.out_dtype = out_dtype
        self.chunk_rows = int(chunk_rows or max(self.batch, 1 << 18))   # staging granularity of the DMA pipeline
        self.shift, self.scale = shift, scale
        self.copy_stream = torch.cuda.Stream(device=device)
        self._cols = self.features + self.sparse + ([self.label] if self.label else [])
        self._blocks: Optional[List[List[List[_ColumnSpan]]]] = None   # [block][col][span]
        self._stage = [None, None]
        self._stage_free = [None, None]        # events: compute done reading slot
        self._x = torch.zeros(self.batch, self.ld, dtype=out_dtype, device=device)
        self._y = torch.zeros(self.batch, dtype=torch.float32, device=device)
        self._s = torch.zeros(self.batch, max(1, len(self.sparse)), dtype=torch.int32, device=device) \
            if self.sparse else None
        self.h2d_bytes = 0
        self.rows_total = dataset.count()

    def __len__(self) -> int:
        n = self.rows_total
        return n // self.batch if self.drop_last else -(-n // self.batch)

    # ------------------------------------------------------------------ host side
    def _prepare(self) -> None:
        """Map + pin every block once; later epochs reuse the mappings."""
        if self._blocks is not None:
            return
        blocks = []
        for blk in self.ds._blocks:
            table = blk.load(pin=True).select(self._cols)
            cols = []
            for name in self._cols:
                cols.append([_ColumnSpan(a) for a in _plain_array(table.column(name))])
            blocks.append(cols)
        self._blocks = blocks

    def _chunks(self) -> Iterator[Tuple[List[Tuple[int, int, int, int]], int]]:
        """Yield (per-column copy lists, rows) for successive row ranges of <= chunk_rows."""
        for cols in self._blocks:
            total = sum(s.rows for s in cols[0])
            start = 0
            while start < total:
                rows = min(self.chunk_rows, total - start)
                plan = []
                for spans in cols:
                    pieces, off, need, pos = [], start, rows, 0
                    for s in spans:
                        if off >= s.rows:
                            off -= s.rows
                            continue
                        take = min(s.rows - off, need)
                        pieces.append((s.address + off * s.width, take * s.width, pos * s.width, s.code, s.width))
                        pos += take
                        need -= take
                        off = 0
                        if need == 0:
                            break
                    plan.append(pieces)
                yield plan, rows
                start += rows

    def _staging(self, slot: int, nbytes: int) -> torch.Tensor:
        buf = self._stage[slot]
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=self.device)
            self._stage[slot] = buf
        return buf

    # ------------------------------------------------------------------ iteration
    def __iter__(self):
        self._prepare()
        compute = torch.cuda.current_stream(self.device)
        nf, ns = len(self.features), len(self.sparse)
        filled = 0
        slot = 0
        for plan, rows in self._chunks():
            # ---- stage the chunk (side stream) ----
            offsets, total = [], 0
            for pieces in plan:
                width = pieces[0][4]
                offsets.append(total)
                total += (rows * width + 255) // 256 * 256
            stage = self._staging(slot, total)
            if self._stage_free[slot] is not None:
                self.copy_stream.wait_event(self._stage_free[slot])
            base = stage.data_ptr()
            for pieces, off in zip(plan, offsets):
                for src, nbytes, dst_off, _code, _w in pieces:
                    ops.memcpy_h2d_async(base + off + dst_off, src, nbytes, self.copy_stream)
                    self.h2d_bytes += nbytes
            ready = torch.cuda.Event()
            ready.record(self.copy_stream)
            compute.wait_event(ready)
            col_ptrs = [base + off for off in offsets]
            codes = [pieces[0][3] for pieces in plan]
            # ---- carve batches out of the staged chunk (compute stream) ----
            done = 0
            while done < rows:
                take = min(self.batch - filled, rows - done)
                label_ptr = col_ptrs[nf + ns] if self.label else 0
                label_code = codes[nf + ns] if self.label else 0
                ops.ingest_pack(col_ptrs[:nf], codes[:nf], take, done, self._x[filled:], self.shift, self.scale,
                                label_ptr, label_code, self._y[filled:] if self.label else None)
                if ns:
                    ops.ingest_indices(col_ptrs[nf:nf + ns], codes[nf:nf + ns], take, done, self.table_rows,
                                       self._s[filled:])
                filled += take
                done += take
                if filled == self.batch:
                    yield self._emit(filled)
                    filled = 0
            free = torch.cuda.Event()
            free.record(compute)
            self._stage_free[slot] = free
            slot ^= 1
        if filled and not self.drop_last:
            yield self._emit(filled)

    def _emit(self, rows: int):
        x = self._x[:rows]
        y = self._y[:rows] if self.label else None
        if self.sparse:
            return x, self._s[:rows], y
        return x, y
