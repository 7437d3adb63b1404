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


@pytest.mark.parametrize("S,D", [(27, 32), (7, 32), (32, 32), (27, 16)])
def test_dot_interaction_matches_bmm_tril(S, D):
    """D = 32 runs on the tensor cores (mma.sync + ldmatrix, one warp per sample); other widths on the scalar
    kernels.  Both against torch.bmm + tril gather in fp32."""
    from raydp_b200 import ops
    B = 777
    pairs = S * (S - 1) // 2
    ld = (D + pairs + 1 + 7) // 8 * 8
    feats = (torch.randn(B, S, D, device="cuda") * 0.5).to(torch.bfloat16)
    out = torch.full((B, ld), 7.0, device="cuda", dtype=torch.bfloat16)
    ops.dot_interaction_fwd(feats, out)
    f = feats.float().requires_grad_(True)
    inter = torch.bmm(f, f.transpose(1, 2))
    li, lj = torch.tril_indices(S, S, offset=-1)
    ref = torch.cat([f[:, 0, :], inter[:, li, lj], torch.zeros(B, ld - D - pairs, device="cuda")], 1)
    assert _rel(out, ref.detach()) < 1e-2 and (out[:, D + pairs:] == 0).all()
    dout = (torch.randn(B, ld, device="cuda") * 0.1).to(torch.bfloat16)
    dfeats = torch.full_like(feats, 3.0)
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
    eager step, including an lr change mid-way
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


# ---- K4 v2: bags, uneven batches, deterministic combine, row-wise optimizers -----------------------------
def _bag_setup(seed=0):
    torch.manual_seed(seed)
    D = 32
    sizes, lens = [50, 600, 7, 90], [1, 3, 4, 2]
    tables = [torch.randn(n, D, device="cuda") for n in sizes]
    col0 = [sum(lens[:t]) for t in range(len(lens))]
    return D, sizes, lens, col0, tables


def _bag_indices(B, sizes, lens, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    cols = []
    for n, L in zip(sizes, lens):
        bag = torch.randint(0, n, (B, L), device="cuda", generator=g)
        if L > 1:                                  # ragged bags: pad some positions (never position 0)
            pad = torch.rand(B, L, device="cuda", generator=g) < 0.3
            pad[:, 0] = False
            bag[pad] = -1
        cols.append(bag)
    return torch.cat(cols, 1).int()


def _pool_ref(table, bag, mean):
    keep = (bag >= 0).float().unsqueeze(-1)
    out = (table[bag.clamp_min(0).long()] * keep).sum(1)
    return out / keep.sum(1).clamp_min(1.0) if mean else out


@pytest.mark.parametrize("mean", [False, True])
def test_embedding_bag_forward_pools_like_torch_with_uneven_batches(mean):
    """Two 'requesting ranks' with different batches (capacity 256; 250 and 37 valid rows, read from device
    memory), two owners; transposed index matrices read with 16-byte loads; negative index = padding."""
    from raydp_b200 import ops
    D, sizes, lens, col0, tables = _bag_setup()
    T, cap = len(sizes), 256
    batches = [250, 37]
    idx = [_bag_indices(b, sizes, lens, 5 + i) for i, b in enumerate(batches)]
    idx_t = [torch.full((sum(lens), cap), 12345, dtype=torch.int32, device="cuda") for _ in batches]   # stale tail
    for it, i in zip(idx_t, idx):
        it[:, :i.shape[0]] = i.t()
    feats = [torch.zeros(cap, T + 1, D, device="cuda", dtype=torch.bfloat16) for _ in batches]
    counts = [torch.tensor([b], dtype=torch.int32, device="cuda") for b in batches]
    for owned in ([0, 2], [1, 3]):
        ops.embedding_bag_fwd([tables[t].data_ptr() for t in owned], [t + 1 for t in owned],
                              [col0[t] for t in owned], [lens[t] for t in owned],
                              [t.data_ptr() for t in idx_t], [f.data_ptr() for f in feats], [cap, cap],
                              [c.data_ptr() for c in counts], cap, T + 1, D, mean)
    for f, i, b in zip(feats, idx, batches):
        for t in range(T):
            ref = _pool_ref(tables[t], i[:, col0[t]:col0[t] + lens[t]], mean)
            assert _rel(f[:b, t + 1], ref) < 4e-3
        assert (f[b:] == 0).all()                   # rows past the valid batch are never written


@pytest.mark.parametrize("kind", ["sgd", "adagrad", "adam"])
def test_embedding_bag_backward_combines_duplicates_then_updates_each_row_once(kind):
    from raydp_b200 import ops
    D, sizes, lens, col0, tables = _bag_setup(1)
    T, cap, B, mean = len(sizes), 512, 509, True
    idx = _bag_indices(B, sizes, lens, 9)           # table 2 has 7 rows: heavy duplication
    idx_t = torch.zeros(sum(lens), cap, dtype=torch.int32, device="cuda")
    idx_t[:, :B] = idx.t()
    g = (torch.randn(cap, T + 1, D, device="cuda") * 0.1).to(torch.bfloat16)
    count = torch.tensor([B], dtype=torch.int32, device="cuda")
    lr, scale, eps, b1, b2 = 0.05, 0.5, 1e-8, 0.9, 0.999
    start = [t.clone() for t in tables]
    acc = torch.zeros(cap * sum(lens), D, dtype=torch.int64, device="cuda")
    flag = [torch.zeros(n, dtype=torch.int32, device="cuda") for n in sizes]
    lst = torch.zeros(cap * sum(lens), 2, dtype=torch.int32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    state = {"sgd": None, "adagrad": [torch.zeros(n, device="cuda") for n in sizes],
             "adam": [torch.zeros(2, n, D, device="cuda") for n in sizes]}[kind]
    step = torch.zeros(1, device="cuda")

    def run():
        step.add_(1.0)
        own = list(range(T))
        ops.embedding_bag_bwd([tables[t].data_ptr() for t in own], [t + 1 for t in own], col0, lens,
                              [idx_t.data_ptr()], [g.data_ptr()], [cap], [count.data_ptr()], cap, T + 1, D, mean,
                              scale, acc, [f.data_ptr() for f in flag], lst, cnt,
                              [s.data_ptr() for s in state] if state else None,
                              [n * D for n in sizes] if kind == "adam" else None, ops.EMB_OPT_KINDS[kind], lr, eps,
                              b1, b2, 0.0, None, step)

    # reference: dense gradient per table by index_add, then the optimizer on touched rows only
    ref = [t.clone() for t in start]
    ref_state = [torch.zeros(n, device="cuda") for n in sizes] if kind == "adagrad" else \
        [torch.zeros(2, n, D, device="cuda") for n in sizes]
    for it in range(1, 3):
        run()
        n_touched = 0
        for t in range(T):
            bag = idx[:, col0[t]:col0[t] + lens[t]]
            keep = bag >= 0
            gt = g[:B, t + 1].float() * scale / keep.sum(1, keepdim=True).clamp_min(1).float()
            G = torch.zeros(sizes[t], D, device="cuda")
            G.index_add_(0, bag.clamp_min(0).long().reshape(-1),
                         (gt.unsqueeze(1) * keep.unsqueeze(-1).float()).reshape(-1, D))
            touched = torch.zeros(sizes[t], dtype=torch.bool, device="cuda")
            touched[bag[keep].long()] = True
            if kind == "sgd":
                ref[t] -= lr * G
            elif kind == "adagrad":
                ref_state[t][touched] += (G[touched] ** 2).mean(1)
                ref[t][touched] -= lr * G[touched] / (ref_state[t][touched].sqrt().unsqueeze(1) + 1e-8)
            else:
                m, v = ref_state[t][0], ref_state[t][1]
                m[touched] = b1 * m[touched] + (1 - b1) * G[touched]
                v[touched] = b2 * v[touched] + (1 - b2) * G[touched] ** 2
                mh, vh = m[touched] / (1 - b1 ** it), v[touched] / (1 - b2 ** it)
                ref[t][touched] -= lr * mh / (vh.sqrt() + eps)
            n_touched += int(touched.sum())
        assert int(cnt[0]) == n_touched
    assert (acc == 0).all()                                            # scratch is clean for the next step
    for t in range(T):
        assert _rel(tables[t] - start[t], ref[t] - start[t]) < (5e-3 if kind == "adam" else 2e-3), (kind, t)
        assert (flag[t] == 0).all()
    # order independence: the same update again from the same start gives the same BITS
    first = [t.clone() for t in tables]
    for t in range(T):
        tables[t].copy_(start[t])
    if state:
        for s_ in state:
            s_.zero_()
    step.zero_()
    run(); run()
    for t in range(T):
        assert torch.equal(tables[t], first[t])


def _bag_model(pooling="mean"):
    from raydp_b200.models import DLRM
    sizes = [97, 1000, 13, 512, 64, 300]
    return DLRM(13, sizes, 32, bottom=(64, 32), top=(128, 64, 1), sparse_grad=False, bag_sizes=[1, 3, 2, 1, 4, 1],
                pooling=pooling).cuda(), sizes


def _bag_batch(model, sizes, B, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dense = torch.randn(B, 13, device="cuda", generator=g)
    idx = _bag_indices(B, sizes, list(model.bag_sizes), seed + 1000)
    y = (torch.rand(B, device="cuda", generator=g) < 0.3).float()
    return dense, idx, y


def test_hybrid_step_with_bags_and_uneven_rank_batches_matches_torch():
    """Three model-parallel ranks with 256, 131 and 0 rows: DDP semantics (mean of the per-rank mean losses over
    the ranks; an empty rank contributes zero gradients but still serves its tables)."""
    from raydp_b200.parallel.hybrid import LocalWorld
    torch.manual_seed(0)
    model, sizes = _bag_model()
    world, cap = 3, 256
    emu = LocalWorld(model.fused_spec(), cap, world, torch.device("cuda"))
    emu.load_from_module(model)
    opt = torch.optim.SGD(model.parameters(), lr=0.05)
    crit = nn.BCEWithLogitsLoss()
    batches = [256, 131, 0]
    for step in range(5):
        parts = [_bag_batch(model, sizes, b, 100 * step + r) for r, b in enumerate(batches)]
        opt.zero_grad()
        loss = sum(crit(model(d.to(torch.bfloat16).float(), i.long())[:, 0], y) for d, i, y in parts if d.shape[0])
        (loss / world).backward()
        opt.step()
        emu.train_step([_pack_dense(d) for d, _, _ in parts], [i for _, i, _ in parts], [y for _, _, y in parts],
                       lr=0.05)
    r0 = emu.ranks[0]
    assert _rel(r0.weight(0), model.bottom[0].weight.detach()) < 2e-2
    assert _rel(r0.weight(len(r0.layers) - 1), model.top[-1].weight.detach()) < 2e-2
    for r in emu.ranks:
        for t in r.owned:
            assert _rel(r.tables[t], model.embeddings[t].weight.detach()) < 2e-2
    for r in emu.ranks[1:]:
        assert torch.equal(r.params, r0.params)


@pytest.mark.parametrize("table_optimizer", ["sgd", "adagrad"])
def test_dlrm_steps_with_momentum_are_bit_reproducible(table_optimizer):
    """Two runs of the same graph-replayed steps (momentum SGD on the dense layers, heavy index duplication)
    end in the same BITS: ordered split sums for weight gradients, ordered column sums for bias gradients,
    fixed-point combine for table rows."""
    from raydp_b200.parallel.hybrid import DistributedDLRM
    finals = []
    for run in range(2):
        torch.manual_seed(3)
        model, sizes = _bag_model("sum")
        eng = DistributedDLRM(model.fused_spec(), 512, torch.device("cuda"), momentum=0.9, use_graph=True,
                              table_optimizer=table_optimizer, deterministic=True)
        eng.core.load_from_module(model)
        for step in range(8):
            dense, idx, y = _bag_batch(model, sizes, 512, 50 + step)
            eng.train_step(_pack_dense(dense), idx, y, 0.05)
        torch.cuda.synchronize()
        assert eng._graphs[0] is not None
        finals.append((eng.core.params.clone(), {t: w.clone() for t, w in eng.core.tables.items()}))
        eng.close()
    assert torch.equal(finals[0][0], finals[1][0])
    for t in finals[0][1]:
        assert torch.equal(finals[0][1][t], finals[1][1][t])
