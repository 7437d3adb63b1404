"""The allocation-free GEMM + batch-statistics entry point, driven through the C ABI.

Who needs it: code that may not call cudaMalloc -- a plugin hosted by an application that owns all
device memory through its own pool, or a latency-critical loop that cannot absorb cudaMalloc's
implicit device synchronisation.  ``rdp_gemm_tn_bnstats_noalloc`` keeps its [2, N] accumulator in a
``__device__`` array inside the library instead of taking a caller-allocated scratch buffer.

The test calls it with plain ctypes (no wrapper) twice in a row with different inputs -- the shared
array keeps the previous call's sums, so the second call only matches if the entry point clears
it -- and compares the output matrix and the per-column mean / variance with the existing entry
point ``rdp_gemm_tn`` fed a caller-allocated, caller-zeroed statistics buffer.  It also checks the
size limit.  The file name sorts last so a ``-x`` run has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

VP, I, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_float


def _lib():
    from raydp_b200._native import KERNELS_LIB
    lib = ctypes.CDLL(KERNELS_LIB)
    lib.rdp_last_error.restype = ctypes.c_char_p
    lib.rdp_cuda_init.argtypes = [I]
    lib.rdp_cuda_init.restype = I
    lib.rdp_gemm_tn.argtypes = [VP, I, VP, I, VP, I, I, I, I, I, VP, I, VP, I, I, I, VP, VP]
    lib.rdp_gemm_tn.restype = I
    lib.rdp_gemm_tn_bnstats_noalloc.argtypes = [VP, I, VP, I, VP, I, I, I, I, I, VP, I, VP, VP, VP]
    lib.rdp_gemm_tn_bnstats_noalloc.restype = I
    assert lib.rdp_cuda_init(torch.cuda.current_device()) > 0, lib.rdp_last_error()
    return lib


def _p(t):
    return VP(t.data_ptr())


def _run_pair(lib, M, N, K, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = (torch.randn(M, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
    b = (torch.randn(N, K, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda", generator=g)
    stream = VP(torch.cuda.current_stream().cuda_stream)
    # existing entry point: caller-allocated, caller-zeroed scratch
    out_ref = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    stats = torch.zeros(2, N, device="cuda")
    rc = lib.rdp_gemm_tn(_p(a), K, _p(b), K, _p(out_ref), N, 0, M, N, K, _p(bias), 1, None, 0, 0, 1,
                         _p(stats), stream)
    assert rc == 0, lib.rdp_last_error()
    # new entry point: no scratch argument at all
    out_new = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    mean = torch.full((N,), float("nan"), device="cuda")
    var = torch.full((N,), float("nan"), device="cuda")
    rc = lib.rdp_gemm_tn_bnstats_noalloc(_p(a), K, _p(b), K, _p(out_new), N, 0, M, N, K, _p(bias), 1,
                                         _p(mean), _p(var), stream)
    assert rc == 0, lib.rdp_last_error()
    torch.cuda.synchronize()
    mean_ref = stats[0] / M
    var_ref = (stats[1] / M - mean_ref * mean_ref).clamp_min(0)
    assert torch.equal(out_new, out_ref)
    assert torch.allclose(mean, mean_ref, rtol=2e-3, atol=1e-3)
    assert torch.allclose(var, var_ref, rtol=2e-3, atol=1e-3)
    # and both agree with plain PyTorch fp32 on the same operands
    full = torch.relu(a.float() @ b.float().t() + bias)
    assert (out_new.float() - full).abs().max().item() < 0.05
    assert torch.allclose(mean, full.mean(0), rtol=2e-2, atol=2e-2)
    assert torch.allclose(var, full.var(0, unbiased=False), rtol=3e-2, atol=2e-2)


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    lib = _lib()
    _run_pair(lib, 3000, 256, 64, seed=1)
    _run_pair(lib, 1500, 128, 32, seed=2)      # smaller N: stale sums of call 1 must not leak in
    _run_pair(lib, 3000, 256, 64, seed=3)


def test_noalloc_rejects_sizes_beyond_the_builtin_scratch():
    lib = _lib()
    M, N, K = 128, 2048, 64
    a = torch.zeros(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(N, K, device="cuda", dtype=torch.bfloat16)
    out = torch.zeros(M, N, device="cuda", dtype=torch.bfloat16)
    mean, var = torch.zeros(N, device="cuda"), torch.zeros(N, device="cuda")
    rc = lib.rdp_gemm_tn_bnstats_noalloc(_p(a), K, _p(b), K, _p(out), N, 0, M, N, K, None, 0, _p(mean),
                                         _p(var), VP(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    assert b"exceeds the built-in scratch" in lib.rdp_last_error()
