"""Python bindings of the sm_100a kernels (``librdp_kernels.so``, C ABI in ``include/raydp_b200.h``).

Every wrapper takes torch CUDA tensors, passes raw device pointers plus the *current* torch
stream, and raises with the library's message on failure -- there is no eager/PyTorch
fallback: on a GPU box a missing or broken extension is an error, not a slow path.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from raydp_b200._native import kernels_available, kernels_lib  # noqa: F401

_initialised_devices = set()
_launches = 0


def launch_count() -> int:
    """Kernels launched through this module so far (graph replays are added by their owner)."""
    return _launches


def add_launches(n: int) -> None:
    global _launches
    _launches += int(n)


class KernelError(RuntimeError):
    pass


def lib():
    handle = kernels_lib()
    dev = torch.cuda.current_device()
    if dev not in _initialised_devices:
        rc = handle.rdp_cuda_init(dev)
        if rc < 0:
            raise KernelError(handle.rdp_last_error().decode())
        _initialised_devices.add(dev)
    return handle


def _check(rc: int) -> None:
    global _launches
    _launches += 1
    if rc != 0:
        raise KernelError(f"raydp_b200 kernel call failed ({rc}): {kernels_lib().rdp_last_error().decode()}")


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def num_sms() -> int:
    return lib().rdp_num_sms()


def _pitch(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "row-major 2-D tensor required"
    return t.stride(0)


def gemm_tn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, *,
            bias: Optional[torch.Tensor] = None, relu: bool = False, aux: Optional[torch.Tensor] = None,
            aux_relu_mask: bool = False, stats_mode: int = 0, stats: Optional[torch.Tensor] = None,
            out_dtype: torch.dtype = torch.bfloat16, n: Optional[int] = None, k: Optional[int] = None) -> torch.Tensor:
    """``out[M,N] = a[M,K] @ b[N,K].T`` on tcgen05 tensor cores with a fused epilogue."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    M = a.shape[0]
    K = k if k is not None else a.shape[1]
    N = n if n is not None else b.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    _check(lib().rdp_gemm_tn(_ptr(a), _pitch(a), _ptr(b), _pitch(b), _ptr(out), _pitch(out),
                             1 if out.dtype == torch.float32 else 0, M, N, K, _ptr(bias),
                             1 if relu else 0, _ptr(aux), _pitch(aux) if aux is not None else 0,
                             1 if aux_relu_mask else 0, stats_mode, _ptr(stats), _stream()))
    return out


def gemm_wgrad(g: torch.Tensor, x: torch.Tensor, out: torch.Tensor, *, mw: Optional[int] = None,
               nw: Optional[int] = None) -> torch.Tensor:
    """``out[Mw,Nw] += g[:, :Mw].T @ x[:, :Nw]`` (fp32 accumulate over the batch rows)."""
    assert g.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and out.dtype == torch.float32
    rows = g.shape[0]
    Mw = mw if mw is not None else g.shape[1]
    Nw = nw if nw is not None else x.shape[1]
    _check(lib().rdp_gemm_wgrad(_ptr(g), _pitch(g), _ptr(x), _pitch(x), _ptr(out), _pitch(out),
                                rows, Mw, Nw, _stream()))
    return out


def gemm_wgrad_det(g: torch.Tensor, x: torch.Tensor, out: torch.Tensor, ws: Optional[torch.Tensor], *,
                   mw: Optional[int] = None, nw: Optional[int] = None) -> torch.Tensor:
    """``out[Mw,Nw] = g[:, :Mw].T @ x[:, :Nw]``, bit-reproducible: batch splits are stored into ``ws`` (fp32
    scratch, any size >= 0; it bounds the number of splits) and added in split order.  Overwrites ``out``."""
    global _launches
    assert g.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and out.dtype == torch.float32
    rows = g.shape[0]
    Mw = mw if mw is not None else g.shape[1]
    Nw = nw if nw is not None else x.shape[1]
    _check(lib().rdp_gemm_wgrad_det(_ptr(g), _pitch(g), _ptr(x), _pitch(x), _ptr(out), _pitch(out), rows, Mw, Nw,
                                    _ptr(ws), ws.numel() if ws is not None else 0, _stream()))
    _launches += 1
    return out


def gemm_wgrad_partial(g: torch.Tensor, x: torch.Tensor, out: torch.Tensor, ws: torch.Tensor, *,
                       mw: Optional[int] = None, nw: Optional[int] = None) -> int:
    """First phase of the deterministic wgrad: split tiles of ``g[:, :Mw].T @ x[:, :Nw]`` stored into ``ws``;
    returns the number of splits (1: the product is already in ``out``).  See :func:`wgrad_reduce_many`."""
    rows = g.shape[0]
    Mw = mw if mw is not None else g.shape[1]
    Nw = nw if nw is not None else x.shape[1]
    splits = ctypes.c_int(0)
    _check(lib().rdp_gemm_wgrad_partial(_ptr(g), _pitch(g), _ptr(x), _pitch(x), _ptr(out), _pitch(out), rows, Mw, Nw,
                                        _ptr(ws), ws.numel(), ctypes.byref(splits), _stream()))
    return splits.value


def wgrad_reduce_many(wss: Sequence[torch.Tensor], splits: Sequence[int], outs: Sequence[torch.Tensor]) -> None:
    """Second phase: ``outs[l] = sum over splits of wss[l]`` in split order, all layers in one launch."""
    for i in range(0, len(outs), 16):
        w, s_, o = wss[i:i + 16], splits[i:i + 16], outs[i:i + 16]
        if all(v <= 1 for v in s_):
            continue
        elems = (ctypes.c_longlong * len(o))(*[t.numel() for t in o])
        _check(lib().rdp_wgrad_reduce_many(_vp_arr([t.data_ptr() for t in w]), elems, _i_arr(s_),
                                           _vp_arr([t.data_ptr() for t in o]), len(o), _stream()))


def colsum_ordered(mats: Sequence[torch.Tensor], widths: Sequence[int], outs: Sequence[torch.Tensor],
                   ws: Optional[torch.Tensor] = None, counters: Optional[torch.Tensor] = None) -> None:
    """``outs[i][:widths[i]] = mats[i][:, :widths[i]].sum(0)`` with a fixed summation order (bias gradients of a
    bit-reproducible step); bf16 matrices with the same number of rows, up to 16 per launch.  ``ws`` (fp32) and
    ``counters`` (int32, zero) let the rows be split over CTAs (two-level ordered tree)."""
    rows = mats[0].shape[0]
    for i in range(0, len(mats), 16):
        m, w, o = mats[i:i + 16], widths[i:i + 16], outs[i:i + 16]
        assert all(t.dtype == torch.bfloat16 and t.shape[0] == rows for t in m)
        _check(lib().rdp_colsum_ordered(_vp_arr([t.data_ptr() for t in m]), _i_arr([_pitch(t) for t in m]),
                                        _i_arr(w), _vp_arr([t.data_ptr() for t in o]), len(m), rows, _ptr(ws),
                                        ws.numel() if ws is not None else 0, _ptr(counters),
                                        counters.numel() if counters is not None else 0, _stream()))


def set_dynamic_tiles(on: bool) -> None:
    """Large-GEMM tile schedule: hardware work stealing (``True``) for GEMMs that overlap with an
    all-reduce kernel, static persistent loop (``False``, default) otherwise.  Process-wide."""
    lib().rdp_gemm_set_dynamic_tiles(1 if on else 0)


def gemm2_tn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, *,
             bias: Optional[torch.Tensor] = None, relu: bool = False, mh: int = 0) -> torch.Tensor:
    """``gemm_tn`` forced onto the CTA-pair (cta_group::2) kernel; ``b.shape[0] % 256 == 0``.
    ``mh``: 0 = choose, 1 = 256-row cluster tiles, 2 = 512-row cluster tiles."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    M, K = a.shape
    N = b.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    _check(lib().rdp_gemm2_tn(_ptr(a), _pitch(a), _ptr(b), _pitch(b), _ptr(out), _pitch(out), M, N, K,
                              _ptr(bias), 1 if relu else 0, mh, _stream()))
    return out


def gemm2_wgrad(g: torch.Tensor, x: torch.Tensor, out: torch.Tensor, mh: int = 0) -> torch.Tensor:
    """``gemm_wgrad`` forced onto the CTA-pair kernel."""
    assert g.dtype == torch.bfloat16 and x.dtype == torch.bfloat16 and out.dtype == torch.float32
    _check(lib().rdp_gemm2_wgrad(_ptr(g), _pitch(g), _ptr(x), _pitch(x), _ptr(out), _pitch(out),
                                 g.shape[0], g.shape[1], x.shape[1], mh, _stream()))
    return out


def bn_fold(stats, count, gamma, beta, running_mean, running_var, momentum, eps, use_running,
            a, c, mean, rstd, w, b, w_folded, w_folded_t, b_folded) -> None:
    N, K = w.shape
    _check(lib().rdp_bn_fold(_ptr(stats), float(count), _ptr(gamma), _ptr(beta), _ptr(running_mean),
                             _ptr(running_var), float(momentum), float(eps), 1 if use_running else 0,
                             _ptr(a), _ptr(c), _ptr(mean), _ptr(rstd), _ptr(w), _ptr(b), N, K,
                             _ptr(w_folded), _pitch(w_folded), _ptr(w_folded_t),
                             _pitch(w_folded_t) if w_folded_t is not None else 0, _ptr(b_folded), _stream()))


def cast_weight(w: torch.Tensor, out: torch.Tensor, out_t: Optional[torch.Tensor] = None) -> None:
    N, K = w.shape
    _check(lib().rdp_cast_weight(_ptr(w), N, K, _ptr(out), _pitch(out), _ptr(out_t),
                                 _pitch(out_t) if out_t is not None else 0, _stream()))


def cast_weight_many(ws, outs, outs_t, biases, bias_outs) -> None:
    """fp32 weights ``ws[l]`` [N, K] -> bf16 operands ``outs[l]`` (and transposes ``outs_t[l]``), biases copied into
    their padded operands, for up to 16 layers per launch."""
    for i in range(0, len(ws), 16):
        w, o, t, b, bo = ws[i:i + 16], outs[i:i + 16], outs_t[i:i + 16], biases[i:i + 16], bias_outs[i:i + 16]
        _check(lib().rdp_cast_weight_many(_vp_arr([x.data_ptr() for x in w]), _i_arr([x.shape[0] for x in w]),
                                          _i_arr([x.shape[1] for x in w]), _vp_arr([x.data_ptr() for x in o]),
                                          _i_arr([_pitch(x) for x in o]), _vp_arr([x.data_ptr() for x in t]),
                                          _i_arr([_pitch(x) for x in t]), _vp_arr([x.data_ptr() for x in b]),
                                          _vp_arr([x.data_ptr() for x in bo]), len(w), _stream()))


def bn_relu_bwd(dy, z, dpre, stats2, mean, rstd, a, dgamma, dbeta, dbias, has_bn: bool = True) -> None:
    B, N = dy.shape
    _check(lib().rdp_bn_relu_bwd(_ptr(dy), _ptr(z), _ptr(dpre), B, N, _pitch(dy), _ptr(stats2), _ptr(mean),
                                 _ptr(rstd), _ptr(a), _ptr(dgamma), _ptr(dbeta), _ptr(dbias),
                                 1 if has_bn else 0, _stream()))


def gemm_tn_bnbwd(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, aux: torch.Tensor, alpha: torch.Tensor,
                  beta: torch.Tensor, delta: torch.Tensor, colsum: torch.Tensor) -> torch.Tensor:
    """dgrad fused with the BN + ReLU backward of the layer it feeds: ``out = (aux > 0) ? alpha * (a @ b.T) +
    beta * aux + delta : 0`` and ``colsum += out.sum(0)``."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and aux.dtype == torch.bfloat16
    M, K = a.shape
    N = b.shape[0]
    _check(lib().rdp_gemm_tn_bnbwd(_ptr(a), _pitch(a), _ptr(b), _pitch(b), _ptr(out), _pitch(out), M, N, K,
                                   _ptr(aux), _pitch(aux), _ptr(alpha), _ptr(beta), _ptr(delta), _ptr(colsum),
                                   _stream()))
    return out


def bn_bwd_coef(dw, db, w_t, a, c, mean, rstd, count, dgamma, dbeta, alpha, beta, delta) -> None:
    """BN backward statistics of layer l-1 from Linear l's raw weight / bias gradients (no batch pass); ``dw`` is
    rewritten into the gradient of the unfolded weight."""
    N, K = dw.shape
    _check(lib().rdp_bn_bwd_coef(_ptr(dw), _pitch(dw), _ptr(db), _ptr(w_t), _pitch(w_t), N, K, _ptr(a), _ptr(c),
                                 _ptr(mean), _ptr(rstd), float(count), _ptr(dgamma), _ptr(dbeta), _ptr(alpha),
                                 _ptr(beta), _ptr(delta), _stream()))


def grad_fixup(dw, db, a, c) -> None:
    N, K = dw.shape
    _check(lib().rdp_grad_fixup(_ptr(dw), _ptr(db), _ptr(a), _ptr(c), N, K, _stream()))


LOSS_KINDS = {"mse": 0, "smooth_l1": 1, "bce_logits": 2}


def mlp_head(z, a, c, w, bias, target, loss_kind: int, grad_scale: float, pred, accum, dy, stats2, dw, db) -> None:
    B, K = z.shape
    _check(lib().rdp_mlp_head(_ptr(z), B, K, _pitch(z), _ptr(a), _ptr(c), _ptr(w), _ptr(bias), _ptr(target),
                              loss_kind, float(grad_scale), _ptr(pred), _ptr(accum), _ptr(dy), _ptr(stats2),
                              _ptr(dw), _ptr(db), _stream()))


def adam_step(p, g, m, v, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0, adamw=False,
              lr_dev=None, step_dev=None, shadow=None) -> None:
    """Fused Adam/AdamW over flat fp32 buffers; ``shadow`` (bf16, same numel) receives the updated
    parameters in the same pass."""
    _check(lib().rdp_adam_step(_ptr(p), _ptr(g), _ptr(m), _ptr(v), p.numel(), lr, beta1, beta2, eps,
                               weight_decay, step, grad_scale, 1 if adamw else 0, _ptr(lr_dev),
                               _ptr(step_dev), _ptr(shadow), _stream()))


def sgd_step(p, g, buf, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False,
             first_step=False, grad_scale=1.0, lr_dev=None, step_dev=None, shadow=None) -> None:
    _check(lib().rdp_sgd_step(_ptr(p), _ptr(g), _ptr(buf), p.numel(), lr, momentum, dampening, weight_decay,
                              1 if nesterov else 0, 1 if first_step else 0, grad_scale, _ptr(lr_dev),
                              _ptr(step_dev), _ptr(shadow), _stream()))


def relu_bwd_bias(g: torch.Tensor, act: Optional[torch.Tensor], dbias: Optional[torch.Tensor],
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out = where(act > 0, g, 0)`` (skipped when ``act`` is None) and ``dbias += out.sum(0)`` in one pass."""
    assert g.dtype == torch.bfloat16 and g.dim() == 2
    if act is not None and out is None:
        out = torch.empty_like(g)
    M, N = g.shape
    _check(lib().rdp_relu_bwd_bias(_ptr(g), _ptr(act), _ptr(out), _ptr(dbias), M, N, _pitch(g), _stream()))
    return out if act is not None else g


def gemm2_nn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, mh: int = 0) -> torch.Tensor:
    """``out[M,N] = a[M,K] @ b[K,N]`` (``b`` row-major: MN-major UMMA operand) on the CTA-pair kernel."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    M, K = a.shape
    N = b.shape[1]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=a.device)
    _check(lib().rdp_gemm2_nn(_ptr(a), _pitch(a), _ptr(b), _pitch(b), _ptr(out), _pitch(out), M, N, K, mh,
                              _stream()))
    return out


def add_scalar(x: torch.Tensor, v: float) -> None:
    _check(lib().rdp_add_scalar(_ptr(x), float(v), _stream()))


def publish_host(src: torch.Tensor, host_pinned: torch.Tensor) -> None:
    """``host_pinned[:] = src`` by a device-side store into pinned host memory (no memcpy in the copy queue)."""
    assert host_pinned.is_pinned() and src.dtype == torch.float32 and host_pinned.dtype == torch.float32
    _check(lib().rdp_publish_host(_ptr(src), ctypes.c_void_p(host_pinned.data_ptr()), src.numel(), _stream()))


def memcpy_h2d_async(dst_ptr: int, src_host_ptr: int, nbytes: int, stream: Optional[torch.cuda.Stream] = None) -> None:
    s = ctypes.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
    _check(lib().rdp_memcpy_h2d_async(ctypes.c_void_p(dst_ptr), ctypes.c_void_p(src_host_ptr), nbytes, s))


_DTYPE_CODE = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.int16: 4,
               torch.int8: 5, torch.uint8: 5, torch.bool: 5}


def dtype_code(dtype) -> int:
    return _DTYPE_CODE[dtype]


def ingest_pack(col_ptrs: Sequence[int], dtype_codes: Sequence[int], rows: int, src_offset: int,
                out: torch.Tensor, shift=None, scale=None, label_ptr: int = 0, label_code: int = 0,
                label_out: Optional[torch.Tensor] = None) -> None:
    n = len(col_ptrs)
    ptrs = (ctypes.c_void_p * max(1, n))(*col_ptrs)
    codes = (ctypes.c_int * max(1, n))(*dtype_codes)
    _check(lib().rdp_ingest_pack(ptrs, codes, n, rows, src_offset, _ptr(out), _pitch(out),
                                 1 if out.dtype == torch.float32 else 0, _ptr(shift), _ptr(scale),
                                 ctypes.c_void_p(label_ptr) if label_ptr else None, label_code,
                                 _ptr(label_out), _stream()))


def ingest_indices(col_ptrs, dtype_codes, rows, src_offset, table_rows, out: torch.Tensor,
                   crosses: Sequence = ()) -> None:
    n = len(col_ptrs)
    ptrs = (ctypes.c_void_p * n)(*col_ptrs)
    codes = (ctypes.c_int * n)(*dtype_codes)
    trows = (ctypes.c_int * n)(*[int(t) for t in table_rows])
    nc = len(crosses)
    ca = (ctypes.c_int * max(1, nc))(*[c[0] for c in crosses])
    cb = (ctypes.c_int * max(1, nc))(*[c[1] for c in crosses])
    cr = (ctypes.c_int * max(1, nc))(*[c[2] for c in crosses])
    _check(lib().rdp_ingest_indices(ptrs, codes, n, rows, src_offset, trows, nc, ca, cb, cr, _ptr(out), _stream()))


def ingest_pack_v2(col_ptrs: Sequence[int], valid_ptrs: Sequence[int], vbit0: Sequence[int], widths: Sequence[int],
                   dtype_codes: Sequence[int], rows: int, src_offset: int, out: torch.Tensor, shift=None, scale=None,
                   null_value: float = float("nan"), label_ptr: int = 0, label_valid_ptr: int = 0,
                   label_vbit0: int = 0, label_code: int = 0, label_out: Optional[torch.Tensor] = None,
                   stream: Optional[torch.cuda.Stream] = None) -> None:
    """Pack staged Arrow columns (scalar or fixed-width vector, with validity bitmaps) into ``out[rows, ld]``."""
    n = len(col_ptrs)
    m = max(1, n)
    ptrs = (ctypes.c_void_p * m)(*col_ptrs)
    valid = (ctypes.c_void_p * m)(*[v or None for v in valid_ptrs])
    vb = (ctypes.c_int * m)(*vbit0)
    ws = (ctypes.c_int * m)(*widths)
    codes = (ctypes.c_int * m)(*dtype_codes)
    s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _stream()
    _check(lib().rdp_ingest_pack_v2(ptrs, valid, vb, ws, codes, n, rows, src_offset, _ptr(out), _pitch(out),
                                    1 if out.dtype == torch.float32 else 0, _ptr(shift), _ptr(scale),
                                    float(null_value), ctypes.c_void_p(label_ptr) if label_ptr else None,
                                    ctypes.c_void_p(label_valid_ptr) if label_valid_ptr else None, label_vbit0,
                                    label_code, _ptr(label_out), s))


def ingest_indices_v2(col_ptrs, valid_ptrs, vbit0, dtype_codes, rows, src_offset, table_rows, out: torch.Tensor,
                      crosses: Sequence = (), stream: Optional[torch.cuda.Stream] = None) -> None:
    n = len(col_ptrs)
    ptrs = (ctypes.c_void_p * n)(*col_ptrs)
    valid = (ctypes.c_void_p * n)(*[v or None for v in valid_ptrs])
    vb = (ctypes.c_int * n)(*vbit0)
    codes = (ctypes.c_int * n)(*dtype_codes)
    trows = (ctypes.c_int * n)(*[int(t) for t in table_rows])
    nc = len(crosses)
    ca = (ctypes.c_int * max(1, nc))(*[c[0] for c in crosses])
    cb = (ctypes.c_int * max(1, nc))(*[c[1] for c in crosses])
    cr = (ctypes.c_int * max(1, nc))(*[c[2] for c in crosses])
    s = ctypes.c_void_p(stream.cuda_stream) if stream is not None else _stream()
    _check(lib().rdp_ingest_indices_v2(ptrs, valid, vb, codes, n, rows, src_offset, trows, nc, ca, cb, cr,
                                       _ptr(out), s))


def ingest_stage(dst: Sequence[int], src: Sequence[int], nbytes: Sequence[int], stream: torch.cuda.Stream) -> None:
    """Issue every host->device copy of a chunk from one native call (``cudaMemcpyAsync`` per span)."""
    global _launches
    n = len(dst)
    if n == 0:
        return
    arr = ctypes.c_uint64 * n
    rc = lib().rdp_ingest_stage(arr(*dst), arr(*src), arr(*nbytes), n, ctypes.c_void_p(stream.cuda_stream))
    if rc != 0:
        raise KernelError(f"rdp_ingest_stage failed ({rc}): {kernels_lib().rdp_last_error().decode()}")


# ---- DLRM: embedding exchange, dot interaction, BCE loss ------------------------------------------
def _vp_arr(ptrs):
    return (ctypes.c_void_p * max(1, len(ptrs)))(*[ctypes.c_void_p(int(x)) for x in ptrs])


def _i_arr(vals):
    return (ctypes.c_int * max(1, len(vals)))(*[int(v) for v in vals])


def embedding_fwd(table_ptrs, slots, cols, idx_ptrs, feats_ptrs, batches, idx_ld: int, S: int, D: int) -> None:
    _check(lib().rdp_embedding_fwd(_vp_arr(table_ptrs), _i_arr(slots), _i_arr(cols), len(table_ptrs),
                                   _vp_arr(idx_ptrs), _vp_arr(feats_ptrs), _i_arr(batches), len(idx_ptrs),
                                   idx_ld, S, D, _stream()))


def embedding_bwd_sgd(table_ptrs, slots, cols, idx_ptrs, dfeats_ptrs, batches, idx_ld: int, S: int, D: int,
                      lr: float, grad_scale: float, lr_dev: Optional[torch.Tensor] = None) -> None:
    _check(lib().rdp_embedding_bwd_sgd(_vp_arr(table_ptrs), _i_arr(slots), _i_arr(cols), len(table_ptrs),
                                       _vp_arr(idx_ptrs), _vp_arr(dfeats_ptrs), _i_arr(batches), len(idx_ptrs),
                                       idx_ld, S, D, float(lr), float(grad_scale), _ptr(lr_dev), _stream()))


def embedding_bag_fwd(table_ptrs, slots, cols, lens, idx_t_ptrs, feats_ptrs, batches, batch_ptrs, ldb: int, S: int,
                      D: int, mean: bool) -> None:
    """K4 v2 forward: pooled lookup + dispatch (``rdp_embedding_bag_fwd``).  ``batch_ptrs`` may be None."""
    _check(lib().rdp_embedding_bag_fwd(_vp_arr(table_ptrs), _i_arr(slots), _i_arr(cols), _i_arr(lens), len(table_ptrs),
                                       _vp_arr(idx_t_ptrs), _vp_arr(feats_ptrs), _i_arr(batches),
                                       _vp_arr(batch_ptrs) if batch_ptrs else None, len(idx_t_ptrs), int(ldb), S, D,
                                       1 if mean else 0, _stream()))


EMB_OPT_KINDS = {"sgd": 0, "adagrad": 1, "rowwise_adagrad": 1, "adam": 2}


def embedding_bag_bwd(table_ptrs, slots, cols, lens, idx_t_ptrs, dfeats_ptrs, batches, batch_ptrs, ldb: int, S: int,
                      D: int, mean: bool, grad_scale: float, acc: torch.Tensor, flag_ptrs, touched: torch.Tensor,
                      count: torch.Tensor, state_ptrs, state_strides, kind: int, lr: float, eps: float = 1e-8,
                      beta1: float = 0.9,
                      beta2: float = 0.999, weight_decay: float = 0.0, lr_dev: Optional[torch.Tensor] = None,
                      step_dev: Optional[torch.Tensor] = None) -> None:
    """K4 v2 backward: deterministic duplicate combine + one update per touched row (three kernels).
    ``acc``: int64 ``[slots, D]`` (zero), ``touched``: int32 ``[slots, 2]``, ``count``: int32 ``[1]``;
    ``slots`` >= number of indices a step can carry for the local tables."""
    global _launches
    n = len(table_ptrs)
    strides = (ctypes.c_int64 * max(1, n))(*[int(v) for v in (state_strides or [0] * n)])
    _check(lib().rdp_embedding_bag_bwd(_vp_arr(table_ptrs), _i_arr(slots), _i_arr(cols), _i_arr(lens), n,
                                       _vp_arr(idx_t_ptrs), _vp_arr(dfeats_ptrs), _i_arr(batches),
                                       _vp_arr(batch_ptrs) if batch_ptrs else None, len(idx_t_ptrs), int(ldb), S, D,
                                       1 if mean else 0, float(grad_scale), _ptr(acc), acc.shape[0],
                                       _vp_arr(flag_ptrs), _ptr(touched), _ptr(count),
                                       _vp_arr(state_ptrs) if state_ptrs else None,
                                       strides, int(kind), float(lr), float(eps), float(beta1), float(beta2),
                                       float(weight_decay), _ptr(lr_dev), _ptr(step_dev), _stream()))
    _launches += 2          # claim + combine + update (the counter reset is a memset node)


def gbt_hist(binned_t: torch.Tensor, g: torch.Tensor, h: torch.Tensor, slot: torch.Tensor, hist: torch.Tensor) -> None:
    """``hist[slot, 0|1, f, bin] += g|h`` over all rows with ``slot >= 0`` (``binned_t``: uint16-as-int16 ``[F, n]``;
    ``hist``: fp64 ``[n_slots, 2, F, n_bins]``, zeroed by the caller)."""
    F, n = binned_t.shape
    n_slots, _, F2, n_bins = hist.shape
    assert F2 == F and binned_t.element_size() == 2 and hist.dtype == torch.float64 and slot.dtype == torch.int32
    assert g.dtype == torch.float32 and h.dtype == torch.float32 and hist.is_contiguous() and binned_t.is_contiguous()
    _check(lib().rdp_gbt_hist(_ptr(binned_t), _ptr(g), _ptr(h), _ptr(slot), _ptr(hist), n, F, n_bins, n_slots,
                              _stream()))


def gbt_partition(binned_t: torch.Tensor, node_of: torch.Tensor, slot: torch.Tensor, in_sample: torch.Tensor,
                  feat: torch.Tensor, thr_bin: torch.Tensor, left: torch.Tensor, right: torch.Tensor,
                  default_left: torch.Tensor, next_slot: torch.Tensor, miss_bin: int) -> None:
    """Rows move to the child node chosen by their level's split; ``slot`` becomes the next level's histogram slot."""
    F, n = binned_t.shape
    assert node_of.dtype == torch.int32 and slot.dtype == torch.int32 and in_sample.element_size() == 1
    assert all(t.dtype == torch.int32 for t in (feat, thr_bin, left, right, next_slot)) and default_left.element_size() == 1
    _check(lib().rdp_gbt_partition(_ptr(binned_t), _ptr(node_of), _ptr(slot), _ptr(in_sample), _ptr(feat),
                                   _ptr(thr_bin), _ptr(left), _ptr(right), _ptr(default_left), _ptr(next_slot), n,
                                   int(miss_bin), _stream()))


def dot_interaction_fwd(feats: torch.Tensor, out: torch.Tensor) -> None:
    B, S, D = feats.shape
    _check(lib().rdp_dot_interaction_fwd(_ptr(feats), _ptr(out), B, S, D, _pitch(out), _stream()))


def dot_interaction_bwd(feats: torch.Tensor, dout: torch.Tensor, dfeats: torch.Tensor) -> None:
    B, S, D = feats.shape
    _check(lib().rdp_dot_interaction_bwd(_ptr(feats), _ptr(dout), _ptr(dfeats), B, S, D, _pitch(dout), _stream()))


def bce_logits(logits: torch.Tensor, target: torch.Tensor, dlogits: Optional[torch.Tensor], accum: torch.Tensor,
               grad_scale: float, dbias: Optional[torch.Tensor] = None) -> None:
    B = logits.shape[0]
    _check(lib().rdp_bce_logits(_ptr(logits), _pitch(logits), _ptr(target), _ptr(dlogits),
                                _pitch(dlogits) if dlogits is not None else 0,
                                dlogits.shape[1] if dlogits is not None else 0, _ptr(accum), B, float(grad_scale),
                                _ptr(dbias), _stream()))


def comm_barrier(bases, flag_offset: int, rank: int, world: int, epoch: torch.Tensor) -> None:
    _check(lib().rdp_comm_barrier(_vp_arr(bases), int(flag_offset), rank, world, _ptr(epoch), _stream()))
