"""GPU ingestion pipeline: pinned Arrow blocks in shared memory -> training tensors on device.

Replaces ``dataset.to_torch(...)`` of the reference (``torch/estimator.py:226-232``: Arrow ->
pandas -> numpy -> ``torch.as_tensor`` -> concat/cast on the CPU, then a blocking copy in any
GPU-patched loop).  Here the host only *addresses* data:

1. every block of the shard is mapped from the object store and page-locked once
   (``cudaHostRegister`` on the shared-memory mapping, so Arrow buffers are DMA-able in place) and
   the copy plan of every chunk -- (device offset, host address, bytes) of each value buffer and
   validity bitmap -- is computed ONCE; later epochs replay it;
2. per chunk of rows ONE native call (``rdp_ingest_stage``) issues all ``cudaMemcpyAsync`` of the chunk
   on a copy stream into a double-buffered device staging area, one chunk ahead of the packing;
3. the ingest kernel (``rdp_ingest_pack_v2`` / ``rdp_ingest_indices_v2``), on a second side stream,
   decodes the mixed-dtype columns -- validity bitmaps included: a null feature becomes ``null_fill``
   (default NaN, which is what the reference's pandas/numpy path produces for a null float; pass
   ``null_fill=0.0`` to train through missing values), vector columns (``fixed_size_list`` /
   uniform ``list``, e.g. a ``VectorAssembler`` output) contribute their width -- normalises, zero-pads
   and writes the row-major bf16 feature matrix that the tcgen05 GEMM reads through TMA, plus the
   fp32 label vector, into one of TWO output slots;
4. the compute stream waits for the slot's "ready" event, the consumer trains on it, and the side
   stream only refills a slot after the "free" event recorded behind the consumer's work: batch i+1
   is copied and packed while step i runs.

Batches may straddle chunk and block boundaries; they are assembled directly in the output
slot by packing at a row offset, so there is no concat.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import torch

from raydp_b200 import ops

_ARROW_CODES = {pa.float32(): (0, 4), pa.float64(): (1, 8), pa.int32(): (2, 4), pa.int64(): (3, 8),
                pa.int16(): (4, 2), pa.int8(): (5, 1), pa.uint8(): (5, 1)}


def _raw_copyable(chunk: pa.Array) -> pa.Array:
    """Bring a scalar chunk to a fixed-width type whose value buffer can be copied as raw bytes
    (casts of views such as timestamp -> int64 are zero-copy; bool / unsigned / half are re-encoded)."""
    t = chunk.type
    if pa.types.is_boolean(t):
        chunk = chunk.cast(pa.int8())
    elif pa.types.is_timestamp(t) or pa.types.is_date64(t) or pa.types.is_duration(t) or pa.types.is_time64(t):
        chunk = chunk.cast(pa.int64())
    elif pa.types.is_date32(t) or pa.types.is_uint16(t) or pa.types.is_time32(t):
        chunk = chunk.cast(pa.int32())
    elif pa.types.is_uint32(t) or pa.types.is_uint64(t):
        chunk = chunk.cast(pa.int64())
    elif pa.types.is_float16(t):
        chunk = chunk.cast(pa.float32())
    elif pa.types.is_decimal(t):
        chunk = chunk.cast(pa.float64())
    if chunk.type not in _ARROW_CODES:
        raise TypeError(f"column type {chunk.type} cannot be ingested; encode it on the executors "
                        "(strings/dictionaries must become numeric ids before training)")
    return chunk


class _Span:
    """Rows of one column that are contiguous in (pinned) host memory: value buffer + validity bitmap."""
    __slots__ = ("address", "rows", "code", "elem", "width", "valid_address", "valid_bit0", "keep")

    def __init__(self, arr: pa.Array):
        t = arr.type
        width, first = 1, arr.offset
        top = arr
        if pa.types.is_fixed_size_list(t):
            width = t.list_size
            values = arr.values
            first = arr.offset * width
        elif pa.types.is_list(t) or pa.types.is_large_list(t):
            offs = np.asarray(arr.offsets)                 # zero-copy view of the offsets (already sliced)
            width = int(offs[1] - offs[0]) if len(arr) else 1
            if len(arr) and (width <= 0 or not np.array_equal(offs, offs[0] + width * np.arange(len(offs), dtype=offs.dtype))):
                raise TypeError("a vector feature column must hold vectors of ONE length (no empty or ragged rows)")
            values = arr.values
            first = int(offs[0]) if len(arr) else 0
        else:
            values = None
        if values is not None:
            if values.null_count:
                raise TypeError("a vector feature column must not have nulls inside its vectors")
            values = _raw_copyable(values)
            code, elem = _ARROW_CODES[values.type]
            buf = values.buffers()[1]
            self.address = buf.address + (values.offset + first) * elem
            self.keep = (arr, values)
        else:
            top = _raw_copyable(arr)
            code, elem = _ARROW_CODES[top.type]
            buf = top.buffers()[1]
            self.address = buf.address + top.offset * elem
            self.keep = (arr, top)
        self.rows, self.code, self.elem, self.width = len(arr), code, elem, max(1, width)
        vbuf = top.buffers()[0] if top.null_count else None
        self.valid_address = vbuf.address if vbuf is not None else 0
        self.valid_bit0 = top.offset if vbuf is not None else 0


class _Chunk:
    """One staged row range: the prebuilt copy plan (per staging slot) and the kernel descriptors."""
    __slots__ = ("rows", "src", "nbytes", "dst_off", "n_copies", "total", "cols", "bytes_total")


class ArrowGpuLoader:
    """Iterable of ``(features, label)`` device tensors for one dataset shard.

    ``features`` is bf16 (or fp32) ``[B, ld]``, zero padded to ``ld``; when ``sparse_columns`` is given
    the iterator yields ``(dense, sparse, label)`` with int32 ``[B, n_sparse]`` index rows.
    Output tensors alternate between two slots: consume a batch before asking for the one after next.
    """

    def __init__(self, dataset, feature_columns: Sequence[str], label_column: Optional[str], batch_size: int,
                 drop_last: bool, device: torch.device, ld: Optional[int] = None,
                 out_dtype: torch.dtype = torch.bfloat16, chunk_rows: Optional[int] = None,
                 sparse_columns: Optional[Sequence[str]] = None, table_rows: Optional[Sequence[int]] = None,
                 shift: Optional[torch.Tensor] = None, scale: Optional[torch.Tensor] = None,
                 null_fill: float = float("nan")):
        self.ds = dataset
        self.features = list(feature_columns)
        self.sparse = list(sparse_columns or [])
        self.table_rows = list(table_rows or [0] * len(self.sparse))
        self.label = label_column
        self.batch = int(batch_size)
        self.drop_last = drop_last
        self.device = device
        self._ld_arg = ld
        self.out_dtype = out_dtype
        # staging granularity of the DMA pipeline: >= 1 MB per fp32 column copy (0.5 MB copies reach 36 GB/s on
        # this box, 1 MB copies 44.6 GB/s; profiles/ingest_timeline.md)
        self.chunk_rows = int(chunk_rows or max(self.batch, 1 << 18))
        self.shift, self.scale = shift, scale
        self.null_fill = float(null_fill)
        import os
        # the side stream outranks the training stream: its short pack kernels must not wait for a gap
        # between the persistent GEMM kernels of the step that is running
        prio = -1 if os.environ.get("RAYDP_B200_INGEST_PRIO", "1") != "0" else 0
        self.copy_stream = torch.cuda.Stream(device=device, priority=prio) if device.type == "cuda" else None
        self.pack_stream = torch.cuda.Stream(device=device, priority=prio) if device.type == "cuda" else None
        self._stage_ready = [torch.cuda.Event(), torch.cuda.Event()] if device.type == "cuda" else [None, None]
        self._stage_free: List[Optional[torch.cuda.Event]] = [None, None]
        self._cols = self.features + self.sparse + ([self.label] if self.label else [])
        self._chunks: Optional[List[_Chunk]] = None
        self._stage: List[Optional[torch.Tensor]] = [None, None]
        self._slots = None
        self._ready = [torch.cuda.Event(), torch.cuda.Event()] if device.type == "cuda" else [None, None]
        self._free: List[Optional[torch.cuda.Event]] = [None, None]
        self.h2d_bytes = 0
        self.rows_total = dataset.count()
        self.n_features: Optional[int] = None
        self.ld: Optional[int] = ld

    def __len__(self) -> int:
        n = self.rows_total
        return n // self.batch if self.drop_last else -(-n // self.batch)

    # ------------------------------------------------------------------ host side, once
    def _prepare(self) -> None:
        """Map + pin every block once, build the copy plan of every chunk, allocate the device buffers;
        epochs replay the plan."""
        if self._slots is not None:
            return
        stage_bytes = self._build_plan()
        for slot in range(2):
            self._stage[slot] = torch.empty(max(stage_bytes, 1 << 20), dtype=torch.uint8, device=self.device)
        B, ns = self.batch, len(self.sparse)
        self._slots = [(torch.zeros(B, self.ld, dtype=self.out_dtype, device=self.device),
                        torch.zeros(B, dtype=torch.float32, device=self.device) if self.label else None,
                        torch.zeros(B, max(1, ns), dtype=torch.int32, device=self.device) if ns else None)
                       for _ in range(2)]

    def _build_plan(self) -> int:
        """Host-only part of :meth:`_prepare`: returns the staging bytes one chunk needs at most."""
        chunks: List[_Chunk] = []
        self._keep = []                 # spans own the Arrow arrays, which own the pinned mappings
        stage_bytes = 0
        widths_seen = None
        for blk in self.ds._blocks:
            table = blk.load(pin=self.device.type == "cuda").select(self._cols)
            if table.num_rows == 0:
                continue
            cols = [[_Span(a) for a in table.column(name).chunks if len(a)] for name in self._cols]
            self._keep.append(cols)
            widths = [spans[0].width for spans in cols]
            for spans in cols:
                if any(s.width != spans[0].width or s.code != spans[0].code for s in spans):
                    raise TypeError("a column changes its vector width or type between record batches")
            if widths_seen is None:
                widths_seen = widths
            elif widths != widths_seen:
                raise TypeError("a vector column changes its width between blocks")
            # chunk boundaries: every span boundary of every column, then at most chunk_rows apart,
            # so each chunk lies inside exactly one span of each column (one copy per buffer)
            cuts = {0, table.num_rows}
            for spans in cols:
                pos = 0
                for s in spans:
                    pos += s.rows
                    cuts.add(pos)
            edges = sorted(cuts)
            bounds = []
            for lo, hi in zip(edges, edges[1:]):
                start = lo
                while start < hi:
                    bounds.append((start, min(hi, start + self.chunk_rows)))
                    start += self.chunk_rows
            for lo, hi in bounds:
                ch = _Chunk()
                ch.rows = hi - lo
                src, nbytes, dst_off, descr, total = [], [], [], [], 0
                for spans in cols:
                    pos = 0
                    for s in spans:
                        if lo < pos + s.rows:
                            break
                        pos += s.rows
                    off = lo - pos                                   # first row inside span s
                    vb = ch.rows * s.width * s.elem
                    src.append(s.address + off * s.width * s.elem)
                    nbytes.append(vb)
                    dst_off.append(total)
                    val_off, vbit0 = -1, 0
                    total += (vb + 255) // 256 * 256
                    if s.valid_address:
                        bit = s.valid_bit0 + off
                        nb = (bit % 8 + ch.rows + 7) // 8
                        src.append(s.valid_address + bit // 8)
                        nbytes.append(nb)
                        dst_off.append(total)
                        val_off, vbit0 = total, bit % 8
                        total += (nb + 255) // 256 * 256
                    descr.append((dst_off[-1] if val_off < 0 else dst_off[-2], val_off, vbit0, s.width, s.code))
                ch.src, ch.nbytes, ch.dst_off, ch.cols, ch.total = src, nbytes, dst_off, descr, total
                ch.n_copies = len(src)
                ch.bytes_total = sum(nbytes)
                stage_bytes = max(stage_bytes, total)
                chunks.append(ch)
        self._chunks = chunks
        nf = len(self.features)
        self.n_features = sum((widths_seen or [1] * len(self._cols))[:nf])
        if self.sparse and widths_seen and any(w != 1 for w in widths_seen[nf:nf + len(self.sparse)]):
            raise TypeError("categorical (index) columns must be scalar integer columns")
        self.ld = self._ld_arg or (max(8, self.n_features) + 7) // 8 * 8
        if self.ld < self.n_features:
            raise ValueError(f"the model takes {self.ld} input features but the feature columns provide "
                             f"{self.n_features}")
        return stage_bytes

    # ------------------------------------------------------------------ iteration
    def _stage_chunk(self, ci: int) -> None:
        """All host->device copies of chunk ``ci`` (copy stream), one chunk ahead of the packing."""
        ch = self._chunks[ci]
        slot = ci & 1
        if self._stage_free[slot] is not None:            # the packs that read this staging slot (chunk ci-2)
            self.copy_stream.wait_event(self._stage_free[slot])
        base = self._stage[slot].data_ptr()
        ops.ingest_stage([base + o for o in ch.dst_off], ch.src, ch.nbytes, self.copy_stream)
        self._stage_ready[slot].record(self.copy_stream)
        self.h2d_bytes += ch.bytes_total

    def __iter__(self):
        """Three streams.  The COPY stream only ever waits for a staging slot to be drained, so the DMA of
        chunk c+1 runs while chunk c is packed and trained on; the PACK stream waits for staged data and for a
        free output slot; the consumer's stream waits for a packed batch.  (With copies and packs on one
        stream, the next chunk's DMA sat behind a pack that was waiting for an output slot, i.e. behind a whole
        training step -- profiles/ingest_timeline.md.)"""
        self._prepare()
        compute = torch.cuda.current_stream(self.device)
        pack = self.pack_stream
        nf, ns = len(self.features), len(self.sparse)
        out, filled = 0, 0
        if self._free[out] is not None:
            pack.wait_event(self._free[out])
        n_chunks = len(self._chunks)
        if n_chunks:
            self._stage_chunk(0)
        for ci, ch in enumerate(self._chunks):
            if ci + 1 < n_chunks:
                self._stage_chunk(ci + 1)
            slot = ci & 1
            base = self._stage[slot].data_ptr()
            pack.wait_event(self._stage_ready[slot])
            cols = ch.cols
            f_ptr = [base + c[0] for c in cols[:nf]]
            f_val = [base + c[1] if c[1] >= 0 else 0 for c in cols[:nf]]
            f_vb = [c[2] for c in cols[:nf]]
            f_w = [c[3] for c in cols[:nf]]
            f_code = [c[4] for c in cols[:nf]]
            lab = cols[nf + ns] if self.label else None
            done = 0
            while done < ch.rows:
                take = min(self.batch - filled, ch.rows - done)
                x, y, s = self._slots[out]
                ops.ingest_pack_v2(f_ptr, f_val, f_vb, f_w, f_code, take, done, x[filled:], self.shift, self.scale,
                                   self.null_fill, base + lab[0] if lab else 0,
                                   base + lab[1] if (lab and lab[1] >= 0) else 0, lab[2] if lab else 0,
                                   lab[4] if lab else 0, y[filled:] if lab else None, stream=pack)
                if ns:
                    sc = cols[nf:nf + ns]
                    ops.ingest_indices_v2([base + c[0] for c in sc], [base + c[1] if c[1] >= 0 else 0 for c in sc],
                                          [c[2] for c in sc], [c[4] for c in sc], take, done, self.table_rows,
                                          s[filled:], stream=pack)
                filled += take
                done += take
                if done == ch.rows:                       # last reader of this staging slot is enqueued
                    if self._stage_free[slot] is None:
                        self._stage_free[slot] = torch.cuda.Event()
                    self._stage_free[slot].record(pack)
                if filled == self.batch:
                    yield self._emit(out, filled, compute, pack)
                    out, filled = self._release(out, compute, pack)
        if filled and not self.drop_last:
            yield self._emit(out, filled, compute, pack)
            self._release(out, compute, pack)

    def _emit(self, slot: int, rows: int, compute, pack):
        self._ready[slot].record(pack)
        compute.wait_event(self._ready[slot])
        x, y, s = self._slots[slot]
        xv = x[:rows]
        yv = y[:rows] if y is not None else None
        if self.sparse:
            return xv, s[:rows], yv
        return xv, yv

    def _release(self, slot: int, compute, pack) -> Tuple[int, int]:
        """The consumer's work on ``slot`` has been enqueued: mark it, move to the other slot."""
        if self._free[slot] is None:
            self._free[slot] = torch.cuda.Event()
        self._free[slot].record(compute)
        other = slot ^ 1
        if self._free[other] is not None:
            pack.wait_event(self._free[other])
        return other, 0
