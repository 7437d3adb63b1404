"""tcgen05 GEMM kernels vs a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from raydp_b200 import ops
    return ops


def _rand(*shape):
    return (torch.randn(*shape, device="cuda") * 0.5).to(torch.bfloat16)


@pytest.mark.parametrize("M,N,K", [(128, 16, 64), (256, 256, 32), (1000, 128, 256), (4096, 64, 128),
                                   (333, 256, 24), (65536, 256, 32), (2048, 1024, 384), (512, 512, 1024)])
def test_gemm_tn_plain(M, N, K):
    ops = _ops()
    a, b = _rand(M, K), _rand(N, K)
    out = ops.gemm_tn(a, b)
    ref = a.float() @ b.float().t()
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 0.05, (M, N, K, err)


def test_gemm_tn_bias_relu_stats():
    ops = _ops()
    M, N, K = 5000, 256, 32
    a, b = _rand(M, K), _rand(N, K)
    bias = torch.randn(N, device="cuda")
    stats = torch.zeros(2, N, device="cuda")
    out = ops.gemm_tn(a, b, bias=bias, relu=True, stats_mode=1, stats=stats)
    ref = torch.relu(a.float() @ b.float().t() + bias)
    assert (out.float() - ref).abs().max().item() < 0.05
    z = out.float()
    assert torch.allclose(stats[0], z.sum(0), rtol=2e-3, atol=0.5)
    assert torch.allclose(stats[1], (z * z).sum(0), rtol=2e-3, atol=0.5)


def test_gemm_tn_aux_stats_and_mask_fp32_out():
    ops = _ops()
    M, N, K = 3000, 128, 256
    a, b, aux = _rand(M, K), _rand(N, K), _rand(M, N)
    stats = torch.zeros(2, N, device="cuda")
    out = ops.gemm_tn(a, b, aux=aux, stats_mode=2, stats=stats, out_dtype=torch.float32)
    ref = a.float() @ b.float().t()
    assert torch.allclose(out, ref, rtol=1e-2, atol=5e-2)
    assert torch.allclose(stats[0], ref.sum(0), rtol=5e-3, atol=1.0)
    assert torch.allclose(stats[1], (ref * aux.float()).sum(0), rtol=5e-3, atol=1.0)
    masked = ops.gemm_tn(a, b, aux=aux, aux_relu_mask=True)
    assert torch.allclose(masked.float(), ref * (aux.float() > 0), rtol=2e-2, atol=6e-2)


@pytest.mark.parametrize("rows,Mw,Nw", [(256, 128, 64), (4096, 256, 32), (10000, 128, 256), (777, 16, 64),
                                       (65536, 64, 128), (3000, 512, 384)])
def test_gemm_wgrad(rows, Mw, Nw):
    ops = _ops()
    g, x = _rand(rows, Mw), _rand(rows, Nw)
    out = torch.zeros(Mw, Nw, device="cuda")
    ops.gemm_wgrad(g, x, out)
    ref = g.float().t() @ x.float()
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err <= 5e-3 * ref.abs().max().item() + 0.05, (rows, Mw, Nw, err)


# ---- CTA-pair (cta_group::2) kernels ------------------------------------------------------------
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 512, 320), (300, 256, 128), (4096, 2560, 256),
                                   (1000, 768, 1032)])
def test_gemm2_tn_pair_kernel(M, N, K):
    ops = _ops()
    a, b = _rand(M, K), _rand(N, K)
    bias = torch.randn(N, device="cuda")
    out = ops.gemm2_tn(a, b, bias=bias, relu=True)
    ref = torch.relu(a.float() @ b.float().t() + bias)
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 0.05, (M, N, K, err)
    plain = ops.gemm2_tn(a, b)
    ref2 = a.float() @ b.float().t()
    assert (plain.float() - ref2).abs().max().item() <= 0.02 * ref2.abs().max().item() + 0.05


@pytest.mark.parametrize("rows,Mw,Nw", [(64, 256, 256), (1000, 512, 768), (4096, 384, 264), (2048, 2304, 2048)])
def test_gemm2_wgrad_pair_kernel(rows, Mw, Nw):
    ops = _ops()
    g, x = _rand(rows, Mw), _rand(rows, Nw)
    out = torch.full((Mw, Nw), 0.25, device="cuda")
    ops.gemm2_wgrad(g, x, out)
    ref = g.float().t() @ x.float() + 0.25
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-2, (rows, Mw, Nw, err)


def test_large_gemm_dispatches_to_pair_kernel_and_matches():
    """rdp_gemm_tn / rdp_gemm_wgrad choose the pair kernels for big plain problems."""
    ops = _ops()
    a, b = _rand(2048, 1024), _rand(4096, 1024)
    out = ops.gemm_tn(a, b)
    ref = a.float() @ b.float().t()
    assert (out.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 0.05
    g, x = _rand(1024, 4096), _rand(1024, 2048)
    w = torch.zeros(4096, 2048, device="cuda")
    ops.gemm_wgrad(g, x, w)
    refw = g.float().t() @ x.float()
    assert (w - refw).abs().max().item() <= 2e-3 * refw.abs().max().item() + 1e-2
