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
