"""Fused MLP step vs the same model in plain PyTorch fp32 (autograd, BatchNorm1d, SmoothL1)."""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


def _setup(B=4096, cols=20, loss="smooth_l1", seed=0):
    from raydp_b200.models import NYCTaxiMLP
    from raydp_b200.ops.fused_mlp import FusedMLP
    torch.manual_seed(seed)
    model = NYCTaxiMLP(cols).cuda()
    fused = FusedMLP(model.fused_spec(), B, torch.device("cuda"), loss=loss)
    fused.load_state_dict(model.state_dict())
    x = torch.randn(B, cols, device="cuda")
    y = (x[:, :3].sum(1) * 2 + 1 + 0.1 * torch.randn(B, device="cuda"))
    return model, fused, x, y


def _rel(a, b):
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("loss_name,crit", [("smooth_l1", nn.SmoothL1Loss()), ("mse", nn.MSELoss())])
def test_forward_backward_matches_autograd(loss_name, crit):
    # B = 16384: bias / beta gradients in front of a BatchNorm are near-cancelling sums whose relative error
    # against an fp32 model falls with 1/sqrt(B) (bf16 activations); the flagship batch is 8x larger still
    model, fused, x, y = _setup(B=16384, loss=loss_name)
    xb = fused.pack_input(x)
    fused.train_step(xb, y)
    # the reference sees the same bf16-rounded inputs
    out = model(xb[:, :20].float())
    loss = crit(out[:, 0], y)
    loss.backward()
    torch.cuda.synchronize()
    acc = fused.accum.cpu()
    assert acc[3].item() == x.shape[0]
    assert abs(acc[0].item() / acc[3].item() - loss.item()) < 2e-2 * max(1.0, abs(loss.item()))
    assert _rel(fused.pred[:x.shape[0]], out[:, 0].detach()) < 2e-2
    sd = dict(model.named_parameters())
    worst = {}
    for l in range(4):
        worst[f"W{l}"] = _rel(fused.view(f"W{l}", True), sd[f"fc{l + 1}.weight"].grad)
        worst[f"b{l}"] = _rel(fused.view(f"b{l}", True), sd[f"fc{l + 1}.bias"].grad)
        worst[f"g{l}"] = _rel(fused.view(f"g{l}", True), sd[f"bn{l + 1}.weight"].grad)
        worst[f"be{l}"] = _rel(fused.view(f"be{l}", True), sd[f"bn{l + 1}.bias"].grad)
    worst["Wh"] = _rel(fused.view("Wh", True).reshape(-1), sd["fc5.weight"].grad.reshape(-1))
    worst["bh"] = _rel(fused.view("bh", True), sd["fc5.bias"].grad)
    bad = {k: v for k, v in worst.items() if not v < 5e-2}
    assert not bad, worst
    for l in range(4):
        bn = getattr(model, f"bn{l + 1}")
        assert _rel(fused.running_mean[l], bn.running_mean) < 2e-2
        assert _rel(fused.running_var[l], bn.running_var) < 2e-2


def test_training_converges_like_torch_adam():
    from raydp_b200 import ops
    model, fused, x, y = _setup(B=8192, seed=1)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    crit = nn.SmoothL1Loss()
    m = torch.zeros_like(fused.params)
    v = torch.zeros_like(fused.params)
    xb = fused.pack_input(x)
    ref_losses, fused_losses = [], []
    for step in range(1, 41):
        opt.zero_grad()
        loss = crit(model(xb[:, :20].float())[:, 0], y)
        loss.backward()
        opt.step()
        ref_losses.append(loss.item())
        fused.accum.zero_()
        fused.train_step(xb, y)
        ops.adam_step(fused.params, fused.grads, m, v, 1e-3, 0.9, 0.999, 1e-8, 0.0, step)
        fused_losses.append((fused.accum[0] / fused.accum[3]).item())
    assert fused_losses[-1] < fused_losses[0] * 0.9
    assert abs(fused_losses[-1] - ref_losses[-1]) < 0.05 * ref_losses[0]
    # parameters stay close to the fp32 reference trajectory
    assert _rel(fused.view("W1"), model.fc2.weight.detach()) < 5e-2
    # checkpoint layout is loadable by the plain module
    from raydp_b200.models import NYCTaxiMLP
    fresh = NYCTaxiMLP(20)
    fresh.load_state_dict(fused.state_dict())
    fresh = fresh.cuda().eval()
    pred_ref = fresh(xb[:, :20].float())[:, 0]
    pred_fused = fused.predict(xb)
    assert _rel(pred_fused, pred_ref.detach()) < 3e-2


def test_bce_head_and_eval_mode():
    model, fused, x, y = _setup(loss="bce_logits", seed=2)
    t = (y > y.median()).float()
    xb = fused.pack_input(x)
    fused.train_step(xb, t)
    out = model(xb[:, :20].float())[:, 0]
    loss = nn.BCEWithLogitsLoss()(out, t)
    torch.cuda.synchronize()
    assert abs((fused.accum[0] / fused.accum[3]).item() - loss.item()) < 2e-2
    model.eval()
    fused.load_state_dict(model.state_dict())
    fused.accum.zero_()
    pred = fused.eval_step(xb, t)
    assert _rel(pred, model(xb[:, :20].float())[:, 0].detach()) < 3e-2


def test_optimizer_kernels_match_torch():
    from raydp_b200 import ops
    torch.manual_seed(0)
    n = 100003
    p0 = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda")
    for wd, adamw in ((0.0, False), (0.01, False), (0.01, True)):
        p_ref = p0.clone().requires_grad_(True)
        opt = (torch.optim.AdamW if adamw else torch.optim.Adam)([p_ref], lr=1e-2, weight_decay=wd)
        p, m, v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
        for step in range(1, 4):
            p_ref.grad = g.clone()
            opt.step()
            ops.adam_step(p, g, m, v, 1e-2, 0.9, 0.999, 1e-8, wd, step, 1.0, adamw)
        assert torch.allclose(p, p_ref.detach(), rtol=1e-5, atol=1e-6)
    p_ref = p0.clone().requires_grad_(True)
    opt = torch.optim.SGD([p_ref], lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-3)
    p, buf = p0.clone(), torch.zeros(n, device="cuda")
    for step in range(3):
        p_ref.grad = g.clone()
        opt.step()
        ops.sgd_step(p, g, buf, 0.1, 0.9, 0.0, 1e-3, True, step == 0)
    assert torch.allclose(p, p_ref.detach(), rtol=1e-5, atol=1e-6)


def test_optimizer_kernels_write_bf16_shadow():
    """The fused optimizer writes the bf16 GEMM-operand copy of the parameters in the same pass
    (vectorised body + scalar tail: n is not a multiple of 4)."""
    from raydp_b200 import ops
    torch.manual_seed(1)
    n = 40007
    p, g = torch.randn(n, device="cuda"), torch.randn(n, device="cuda")
    m, v = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    shadow = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    ops.adam_step(p, g, m, v, 1e-2, 0.9, 0.999, 1e-8, 0.0, 1, 1.0, False, shadow=shadow)
    assert torch.equal(shadow, p.to(torch.bfloat16))
    buf = torch.zeros(n, device="cuda")
    ops.sgd_step(p, g, buf, 0.1, 0.9, 0.0, 0.0, False, True, shadow=shadow)
    assert torch.equal(shadow, p.to(torch.bfloat16))


def test_relu_bwd_bias_kernel():
    from raydp_b200 import ops
    torch.manual_seed(2)
    M, N = 3001, 520
    g = torch.randn(M, N, device="cuda").to(torch.bfloat16)
    act = torch.relu(torch.randn(M, N, device="cuda")).to(torch.bfloat16)
    db = torch.full((N,), 0.5, device="cuda")
    out = ops.relu_bwd_bias(g, act, db)
    ref = torch.where(act > 0, g, torch.zeros_like(g))
    assert torch.equal(out, ref)
    assert torch.allclose(db, ref.float().sum(0) + 0.5, rtol=1e-3, atol=5e-2)
    db2 = torch.zeros(N, device="cuda")
    same = ops.relu_bwd_bias(g, None, db2)
    assert same is g and torch.allclose(db2, g.float().sum(0), rtol=1e-3, atol=5e-2)


def test_ingest_pack_matches_torch():
    from raydp_b200 import ops
    rows = 10007
    cols = [torch.randn(rows + 5, device="cuda", dtype=torch.float64),
            torch.randn(rows + 5, device="cuda", dtype=torch.float32),
            torch.randint(-50, 50, (rows + 5,), device="cuda", dtype=torch.int32),
            torch.randint(-50, 50, (rows + 5,), device="cuda", dtype=torch.int64)]
    label = torch.randn(rows + 5, device="cuda", dtype=torch.float64)
    shift = torch.tensor([0.5, 0.0, 1.0, -2.0], device="cuda")
    scale = torch.tensor([2.0, 1.0, 0.1, 0.5], device="cuda")
    out = torch.full((rows, 8), 7.0, device="cuda", dtype=torch.bfloat16)
    lab = torch.empty(rows, device="cuda")
    ops.ingest_pack([c.data_ptr() for c in cols], [ops.dtype_code(c.dtype) for c in cols], rows, 5, out,
                    shift, scale, label.data_ptr(), ops.dtype_code(label.dtype), lab)
    ref = torch.stack([(c[5:].float() - shift[i]) * scale[i] for i, c in enumerate(cols)], 1)
    assert torch.allclose(out[:, :4].float(), ref.to(torch.bfloat16).float())
    assert (out[:, 4:] == 0).all()
    assert torch.allclose(lab, label[5:].float())
    idx_out = torch.empty(rows, 3, device="cuda", dtype=torch.int32)
    ops.ingest_indices([cols[2].data_ptr(), cols[3].data_ptr()], [2, 3], rows, 5, [7, 0], idx_out,
                       crosses=[(0, 1, 1000)])
    assert torch.equal(idx_out[:, 0].long(), cols[2][5:].long() % 7)
    assert torch.equal(idx_out[:, 1].long(), cols[3][5:])
    assert int(idx_out[:, 2].min()) >= 0 and int(idx_out[:, 2].max()) < 1000


def test_cast_weight_plain_and_transposed():
    from raydp_b200 import ops
    torch.manual_seed(4)
    for N, K, ldo, ldt in ((100, 52, 56, 104), (1024, 384, 384, 1024), (33, 7, 8, 40)):
        w = torch.randn(N, K, device="cuda")
        out = torch.zeros(N, ldo, dtype=torch.bfloat16, device="cuda")[:, :K] if ldo != K else \
            torch.zeros(N, K, dtype=torch.bfloat16, device="cuda")
        out_t = torch.zeros(K, ldt, dtype=torch.bfloat16, device="cuda")[:, :N] if ldt != N else \
            torch.zeros(K, N, dtype=torch.bfloat16, device="cuda")
        ops.cast_weight(w, out, out_t)
        assert torch.equal(out, w.to(torch.bfloat16)) and torch.equal(out_t, w.t().to(torch.bfloat16))
        only = torch.zeros(N, K, dtype=torch.bfloat16, device="cuda")
        ops.cast_weight(w, only)
        assert torch.equal(only, w.to(torch.bfloat16))


def test_bn_backward_without_a_batch_pass_matches_fp32():
    """rdp_bn_bwd_coef + rdp_gemm_tn_bnbwd against the textbook BN+ReLU backward in fp32 PyTorch: the batch sums
    come from the weight gradient (sum_b dy = db . W, sum_b dy*r = sum_i W * G), never from a pass over dy."""
    from raydp_b200 import ops
    torch.manual_seed(3)
    B, K, N = 3000, 128, 64            # layer l: K inputs (width of layer l-1), N outputs; B not a tile multiple
    dev = "cuda"
    r = torch.relu(torch.randn(B, K, device=dev)).to(torch.bfloat16)                 # ReLU output of layer l-1
    dpre = (0.1 * torch.randn(B, N, device=dev)).to(torch.bfloat16)                  # dL/dpre of layer l
    W = (torch.randn(N, K, device=dev) / K ** 0.5)
    w_t = W.t().contiguous().to(torch.bfloat16)                                      # [K, N] dgrad operand
    gamma = 1 + 0.1 * torch.randn(K, device=dev)
    bn_bias = 0.1 * torch.randn(K, device=dev)
    rf = r.float()
    mean, var = rf.mean(0), rf.var(0, unbiased=False)
    rstd = (var + 1e-5).rsqrt()
    a, c = gamma * rstd, bn_bias - gamma * rstd * mean
    # reference: dy = dpre @ W(bf16), textbook BN backward, ReLU mask
    dy = dpre.float() @ w_t.float().t()
    zhat = (rf - mean) * rstd
    s1, s2 = dy.sum(0), (dy * zhat).sum(0)
    dr = gamma * rstd * (dy - s1 / B - zhat * s2 / B)
    want = torch.where(rf > 0, dr, torch.zeros_like(dr))
    G = dpre.float().t() @ rf                                                        # raw wgrad of layer l
    db = dpre.float().sum(0)
    dw = G.clone()
    dgamma, dbeta = torch.zeros(K, device=dev), torch.zeros(K, device=dev)
    al, be, de = torch.zeros(K, device=dev), torch.zeros(K, device=dev), torch.zeros(K, device=dev)
    ops.bn_bwd_coef(dw, db, w_t, a, c, mean, rstd, B, dgamma, dbeta, al, be, de)
    assert _rel(dbeta, s1) < 1e-4 and _rel(dgamma, s2) < 1e-3
    assert _rel(dw, a[None, :] * G + c[None, :] * db[:, None]) < 1e-5
    out = torch.empty(B, K, dtype=torch.bfloat16, device=dev)
    colsum = torch.zeros(K, device=dev)
    ops.gemm_tn_bnbwd(dpre, w_t, out, r, al, be, de, colsum)
    torch.cuda.synchronize()
    assert _rel(out.float(), want) < 1e-2
    assert torch.equal(out.float() == 0, ~(rf > 0) | (want.to(torch.bfloat16).float() == 0))
    assert _rel(colsum, out.float().sum(0)) < 1e-4          # tensor-core column sums of the stored tile
    assert _rel(colsum, want.sum(0)) < 5e-2
