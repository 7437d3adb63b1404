"""Every hand-written kernel of the hot paths, once, inside a cudaProfilerStart/Stop window -- the target of the
`ncu --set full` capture whose summary is profiles/ncu_all_kernels.md.

    python bench/profile_targets.py                                   # must exit 0 first
    ncu --set full --clock-control none --import-source on --profile-from-start off \
        -o gpurun_out/prof_all -f python bench/profile_targets.py
    ncu -i gpurun_out/prof_all.ncu-rep --page raw --csv > /tmp/raw.csv && python bench/ncu_summarize.py /tmp/raw.csv

Sections (each warmed up outside the window, then ONE instance inside it):
  step    one eager training step of the flagship NYC-taxi MLP (B = 131072): every kernel of the step
  dlrm    one eager step of the hybrid DLRM (26 tables, B = 8192): embedding_bag_*, dot_interaction_*, ...
  comm    the all-reduce kernels on 4 emulated ranks (cooperative launch; LocalEmulation): one-shot, LL,
          two-shot, bf16-wire -- same device code as the multi-GPU path, peers are local
  ingest  ingest_pack_kernel / ingest_indices_kernel on one 131072-row chunk
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("RAYDP_B200_DLRM_GRAPH", "0")

import torch  # noqa: E402

from raydp_b200 import ops  # noqa: E402


def window(fn):
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()


def section_step(dev):
    from raydp_b200.models import NYCTaxiMLP
    from raydp_b200.ops.fused_mlp import FusedMLP
    B = 131072
    torch.manual_seed(0)
    model = NYCTaxiMLP(20)
    fused = FusedMLP(model.fused_spec(), B, dev)
    fused.load_state_dict({k: v.to(dev) for k, v in model.state_dict().items()})
    x = fused.pack_input(torch.randn(B, 20, device=dev))
    y = torch.randn(B, device=dev)
    m, v = torch.zeros_like(fused.params), torch.zeros_like(fused.params)
    step = [0]

    def one():
        step[0] += 1
        fused.train_step(x, y)
        ops.adam_step(fused.params, fused.grads, m, v, 1e-3, 0.9, 0.999, 1e-8, 0.0, step[0])
    for _ in range(3):
        one()
    window(one)


def section_dlrm(dev):
    from raydp_b200.models import DLRM
    from raydp_b200.parallel.hybrid import DistributedDLRM
    B, rows = 8192, 100000
    torch.manual_seed(0)
    model = DLRM(13, [rows] * 26, 32, sparse_grad=False)
    eng = DistributedDLRM(model.fused_spec(), B, dev, use_graph=False, table_optimizer="adagrad")
    eng.core.load_from_module(model)
    d = torch.zeros(B, eng.core.kpad, device=dev, dtype=torch.bfloat16)
    d[:, :13] = torch.randn(B, 13, device=dev).to(torch.bfloat16)
    idx = torch.randint(0, rows, (B, 26), device=dev, dtype=torch.int32)
    y = (torch.rand(B, device=dev) < 0.25).float()
    for _ in range(3):
        eng.train_step(d, idx, y, 0.01)
    window(lambda: eng.train_step(d, idx, y, 0.01))
    eng.close()


def section_comm(dev):
    from raydp_b200.comm import LocalEmulation
    world = 4
    for algo, n in (("one_shot", 48516), ("ll", 48516), ("two_shot", 4 << 20), ("bf16", 4 << 20)):
        emu = LocalEmulation(world, n, dev)
        ins = [torch.randn(n, device=dev) for _ in range(world)]
        outs = [torch.empty(n, device=dev) for _ in range(world)]
        for _ in range(2):
            emu.allreduce(ins, outs, scale=1.0 / world, algo=algo)
        # (ncu saves and restores device memory around every replay pass, so the epoch words and flags start
        #  each pass from the same state; all "ranks" live in this one cooperative grid)
        window(lambda: emu.allreduce(ins, outs, scale=1.0 / world, algo=algo))


def section_ingest(dev):
    rows, ncols = 131072, 14
    cols = [torch.randn(rows, device=dev, dtype=torch.float64 if i % 3 == 0 else torch.float32)
            for i in range(ncols)]
    codes = [1 if c.dtype == torch.float64 else 0 for c in cols]
    valid = [torch.full(((rows + 7) // 8,), 0xFF, device=dev, dtype=torch.uint8) if i % 2 == 0 else None
             for i in range(ncols)]
    label = torch.randn(rows, device=dev)
    out = torch.empty(rows, 32, device=dev, dtype=torch.bfloat16)
    lab = torch.empty(rows, device=dev)
    keys = [torch.randint(0, 1 << 30, (rows,), device=dev, dtype=torch.int64) for _ in range(26)]
    idx = torch.empty(rows, 26, device=dev, dtype=torch.int32)

    def run():
        ops.ingest_pack_v2([c.data_ptr() for c in cols], [v.data_ptr() if v is not None else 0 for v in valid],
                           [0] * ncols, [1] * ncols, codes, rows, 0, out, null_value=0.0,
                           label_ptr=label.data_ptr(), label_code=0, label_out=lab)
        ops.ingest_indices_v2([k.data_ptr() for k in keys], [0] * 26, [0] * 26, [3] * 26, rows, 0, [100000] * 26, idx)
    for _ in range(2):
        run()
    window(run)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sections", default="step,dlrm,comm,ingest")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    for name in args.sections.split(","):
        globals()[f"section_{name}"](dev)
        print(f"[profile_targets] {name} done", flush=True)


if __name__ == "__main__":
    main()
