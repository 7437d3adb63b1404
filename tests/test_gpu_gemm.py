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
@pytest.mark.parametrize("mh", [1, 2])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 512, 320), (300, 256, 128), (4096, 2560, 256),
                                   (1000, 768, 1032), (9000, 3072, 128)])
def test_gemm2_tn_pair_kernel(M, N, K, mh):
    ops = _ops()
    a, b = _rand(M, K), _rand(N, K)
    bias = torch.randn(N, device="cuda")
    out = ops.gemm2_tn(a, b, bias=bias, relu=True, mh=mh)
    ref = torch.relu(a.float() @ b.float().t() + bias)
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 0.05, (M, N, K, err)
    plain = ops.gemm2_tn(a, b, mh=mh)
    ref2 = a.float() @ b.float().t()
    assert (plain.float() - ref2).abs().max().item() <= 0.02 * ref2.abs().max().item() + 0.05


@pytest.mark.parametrize("mh", [1, 2])
@pytest.mark.parametrize("rows,Mw,Nw", [(64, 256, 256), (1000, 512, 768), (4096, 384, 264), (2048, 2304, 2048),
                                       (512, 6144, 3328)])
def test_gemm2_wgrad_pair_kernel(rows, Mw, Nw, mh):
    ops = _ops()
    g, x = _rand(rows, Mw), _rand(rows, Nw)
    out = torch.full((Mw, Nw), 0.25, device="cuda")
    ops.gemm2_wgrad(g, x, out, mh=mh)
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


@pytest.mark.parametrize("mh", [1, 2])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (700, 512, 200), (4096, 1024, 768), (1536, 3328, 512)])
def test_gemm2_nn_pair_kernel(M, N, K, mh):
    """dgrad shape: out = a @ b with b row-major [K, N] (MN-major B operand, no transposed weight copy)."""
    ops = _ops()
    a, b = _rand(M, K), _rand(K, N)
    out = ops.gemm2_nn(a, b, mh=mh)
    ref = a.float() @ b.float()
    torch.cuda.synchronize()
    err = (out.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 0.05, (M, N, K, err)


def test_fused_linear_matches_torch_on_pair_kernel_shapes():
    """FusedLinear fwd/bwd (pair kernels: TN fwd, NN dgrad, MN-major wgrad, fused ReLU-mask + bias grad)."""
    from raydp_b200.ops.linear import FusedLinear
    torch.manual_seed(0)
    lin = FusedLinear(1024, 2560, relu=True).cuda()
    with torch.no_grad():      # bf16-representable values: the ReLU masks of both sides then agree
        lin.weight.copy_(lin.weight.to(torch.bfloat16).float())
    ref = torch.nn.Linear(1024, 2560).cuda()
    ref.load_state_dict(lin.state_dict())
    x = torch.randn(4096, 1024, device="cuda").to(torch.bfloat16).float()
    xr = x.clone().requires_grad_(True)
    xf = x.clone().requires_grad_(True)
    y = lin(xf.to(torch.bfloat16))
    yr = torch.relu(ref(xr))
    go = torch.randn_like(yr).to(torch.bfloat16).float()
    y.backward(go.to(torch.bfloat16))
    yr.backward(go)
    rel = lambda a, b: ((a.float() - b.float()).norm() / b.float().norm()).item()  # noqa: E731
    assert rel(y, yr) < 2e-2
    assert rel(xf.grad, xr.grad) < 3e-2
    assert rel(lin.weight.grad, ref.weight.grad) < 3e-2
    assert rel(lin.bias.grad, ref.bias.grad) < 3e-2


def test_pair_kernels_with_dynamic_tile_schedule():
    """Cluster-launch-control work stealing: same results as the static persistent loop."""
    ops = _ops()
    a, b = _rand(4096, 512), _rand(5120, 512)
    bias = torch.randn(5120, device="cuda")
    g, x = _rand(1024, 2304), _rand(1024, 2048)
    static_out = ops.gemm2_tn(a, b, bias=bias, relu=True, mh=1)
    ops.set_dynamic_tiles(True)
    try:
        for mh in (1, 2):
            out = ops.gemm2_tn(a, b, bias=bias, relu=True, mh=mh)
            assert torch.equal(out, static_out), mh
            w = torch.zeros(2304, 2048, device="cuda")
            ops.gemm2_wgrad(g, x, w, mh=mh)
            ref = g.float().t() @ x.float()
            assert (w - ref).abs().max().item() <= 2e-3 * ref.abs().max().item() + 1e-2
            bt = _rand(512, 5120)
            nn = ops.gemm2_nn(a, bt, mh=mh)
            refn = a.float() @ bt.float()
            assert (nn.float() - refn).abs().max().item() <= 0.02 * refn.abs().max().item() + 0.05
        torch.cuda.synchronize()
    finally:
        ops.set_dynamic_tiles(False)
