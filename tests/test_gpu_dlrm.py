"""DLRM kernels and the hybrid-parallel step vs the plain PyTorch model."""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu

SIZES = [97, 1000, 13, 512, 64, 2048, 5, 300, 77, 1024, 31, 256, 999, 128, 17, 4096, 50, 60, 70, 80, 90, 100, 110,
         120, 130, 140]


def _rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def _data(B, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dense = torch.randn(B, 13, device="cuda", generator=g)
    idx = torch.stack([torch.randint(0, n, (B,), device="cuda", generator=g) for n in SIZES], 1).int()
    y = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    return dense, idx, y


def _pack_dense(dense):
    out = torch.zeros(dense.shape[0], 16, device="cuda", dtype=torch.bfloat16)
    out[:, :13] = dense.to(torch.bfloat16)
    return out


def test_dot_interaction_matches_bmm_tril():
    from raydp_b200 import ops
    B, S, D = 777, 27, 32
    feats = (torch.randn(B, S, D, device="cuda") * 0.5).to(torch.bfloat16)
    out = torch.empty(B, 384, device="cuda", dtype=torch.bfloat16)
    ops.dot_interaction_fwd(feats, out)
    f = feats.float().requires_grad_(True)
    inter = torch.bmm(f, f.transpose(1, 2))
    li, lj = torch.tril_indices(S, S, offset=-1)
    ref = torch.cat([f[:, 0, :], inter[:, li, lj], torch.zeros(B, 1, device="cuda")], 1)
    assert _rel(out, ref.detach()) < 1e-2 and (out[:, 383] == 0).all()
    dout = (torch.randn(B, 384, device="cuda") * 0.1).to(torch.bfloat16)
    dfeats = torch.empty_like(feats)
    ops.dot_interaction_bwd(feats, dout, dfeats)
    ref.backward(dout.float())
    assert _rel(dfeats, f.grad) < 1.5e-2


def test_embedding_fwd_bwd_sharded_like_a_single_table_set():
    from raydp_b200 import ops
    B, T, D = 1000, 5, 32
    sizes = [50, 60, 70, 80, 90]
    tables = [torch.randn(n, D, device="cuda") for n in sizes]
    idx = torch.stack([torch.randint(0, n, (B,), device="cuda") for n in sizes], 1).int()
    feats = torch.zeros(B, T + 1, D, device="cuda", dtype=torch.bfloat16)
    # two "ranks": tables 0,2,4 and 1,3 push into the same slab
    for owned in ([0, 2, 4], [1, 3]):
        ops.embedding_fwd([tables[t].data_ptr() for t in owned], [t + 1 for t in owned], owned,
                          [idx.data_ptr()], [feats.data_ptr()], [B], T, T + 1, D)
    ref = torch.stack([tables[t][idx[:, t].long()] for t in range(T)], 1)
    assert torch.allclose(feats[:, 1:].float(), ref.to(torch.bfloat16).float())
    g = (torch.randn(B, T + 1, D, device="cuda") * 0.1).to(torch.bfloat16)
    expect = [t.clone() for t in tables]
    for t in range(T):
        expect[t].index_add_(0, idx[:, t].long(), -0.5 * 0.25 * g[:, t + 1].float())
    for owned in ([0, 2, 4], [1, 3]):
        ops.embedding_bwd_sgd([tables[t].data_ptr() for t in owned], [t + 1 for t in owned], owned,
                              [idx.data_ptr()], [g.data_ptr()], [B], T, T + 1, D, 0.5, 0.25)
    for t in range(T):
        assert torch.allclose(tables[t], expect[t], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_hybrid_step_matches_torch_sgd(world):
    from raydp_b200.models import DLRM
    from raydp_b200.parallel.hybrid import LocalWorld
    torch.manual_seed(0)
    B = 512 * world
    model = DLRM(13, SIZES, 32, sparse_grad=False).cuda()
    emu = LocalWorld(model.fused_spec(), B // world, world, torch.device("cuda"))
    emu.load_from_module(model)
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    crit = nn.BCEWithLogitsLoss()
    losses_ref, losses = [], []
    for step in range(6):
        dense, idx, y = _data(B, 10 + step)
        dq = dense.to(torch.bfloat16).float()
        opt.zero_grad()
        loss = crit(model(dq, idx.long())[:, 0], y)
        loss.backward()
        opt.step()
        losses_ref.append(loss.item())
        for r in emu.ranks:
            r.accum.zero_()
        per = B // world
        emu.train_step([_pack_dense(dense[i * per:(i + 1) * per]) for i in range(world)],
                       [idx[i * per:(i + 1) * per] for i in range(world)],
                       [y[i * per:(i + 1) * per] for i in range(world)], lr=0.05)
        acc = torch.stack([r.accum for r in emu.ranks]).sum(0)
        losses.append((acc[0] / acc[3]).item())
    assert abs(losses[0] - losses_ref[0]) < 2e-2
    assert abs(losses[-1] - losses_ref[-1]) < 3e-2
    r0 = emu.ranks[0]
    assert _rel(r0.weight(0), model.bottom[0].weight.detach()) < 2e-2
    assert _rel(r0.weight(len(r0.layers) - 1), model.top[-1].weight.detach()) < 2e-2
    t = emu.ranks[0].owned[0]
    assert _rel(emu.ranks[0].tables[t], model.embeddings[t].weight.detach()) < 2e-2
    for r in emu.ranks[1:]:
        assert torch.equal(r.params, r0.params)                      # dense replicas stay identical
    dense, idx, _ = _data(B, 99)
    per = B // world
    pred = torch.cat(emu.predict([_pack_dense(dense[i * per:(i + 1) * per]) for i in range(world)],
                                 [idx[i * per:(i + 1) * per] for i in range(world)]))
    model.eval()
    ref = model(dense.to(torch.bfloat16).float(), idx.long())[:, 0]
    assert (pred - ref.detach()).abs().max().item() < 0.1


def test_graph_replayed_step_tracks_the_eager_step():
    """DistributedDLRM replays one CUDA graph per slab parity; parameters and tables must follow the
    eager step (differences only from atomic accumulation order), including an lr change mid-way
    (lr is a device scalar inside the graph)."""
    from raydp_b200.models import DLRM
    from raydp_b200.parallel.hybrid import DistributedDLRM
    torch.manual_seed(3)
    B = 512
    model = DLRM(13, SIZES, 32, sparse_grad=False).cuda()
    engines = []
    for use_graph in (False, True):
        eng = DistributedDLRM(model.fused_spec(), B, torch.device("cuda"), momentum=0.9, use_graph=use_graph)
        eng.core.load_from_module(model)
        engines.append(eng)
    for step in range(8):
        dense, idx, y = _data(B, 50 + step)
        lr = 0.05 if step < 5 else 0.01
        for eng in engines:
            eng.train_step(_pack_dense(dense), idx, y, lr)
    torch.cuda.synchronize()
    eager, graphed = engines
    assert graphed._graphs[0] is not None and graphed._graphs[1] is not None and eager._graphs[0] is None
    assert _rel(graphed.core.params, eager.core.params) < 2e-3
    for t in (0, 5, 15):
        assert _rel(graphed.core.tables[t], eager.core.tables[t]) < 2e-3
    for eng in engines:
        eng.close()


@pytest.mark.multigpu
def test_multigpu_hybrid_dlrm_graph_vs_eager():
    """Multi-process hybrid DLRM: graph-replayed steps track eager steps, dense replicas stay bit-identical."""
    import os
    import subprocess
    import sys
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("needs >= 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={min(n_gpus, 4)}",
           "--master-addr", "127.0.0.1", "--master-port", "29538", os.path.join(root, "bench", "dlrm_bench.py"),
           "--check"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "CHECK OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
