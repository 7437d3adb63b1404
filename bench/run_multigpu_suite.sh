#!/usr/bin/env bash
# Multi-GPU measurement suite on ONE node: bash bench/run_multigpu_suite.sh N [outdir]
# Writes JSON results under outdir (default gpurun_out/suite_N).  Every command has its own timeout.
set -u
N=${1:-2}
OUT=${2:-gpurun_out/suite_${N}}
mkdir -p "$OUT"
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29600
step() {   # name timeout cmd...
  local name=$1 limit=$2; shift 2
  port=$((port + 1))
  echo "== $name"
  timeout "$limit" "$@" > "$OUT/$name.log" 2>&1 || echo "   ($name exited with $?)"
  tail -n 2 "$OUT/$name.log"
}
step allreduce 240 $RUN --master-port $((port + 1)) bench/allreduce_bench.py --json "$OUT/allreduce.json"
step alltoall 200 $RUN --master-port $((port + 2)) bench/alltoall_bench.py --json "$OUT/alltoall.json"
step bigmlp_ours 200 $RUN --master-port $((port + 3)) bench/bigmlp_bench.py --steps 6 --warmup 3
step bigmlp_standin 200 $RUN --master-port $((port + 4)) bench/bigmlp_bench.py --steps 6 --warmup 3 --impl standin
step dlrm_ours 200 $RUN --master-port $((port + 5)) bench/dlrm_bench.py
step dlrm_check 200 $RUN --master-port $((port + 7)) bench/dlrm_bench.py --check
step dlrm_variants 200 $RUN --master-port $((port + 8)) bench/dlrm_variants.py
step collectives_stress 200 $RUN --master-port $((port + 9)) bench/allreduce_bench.py --stress 300
step ingest 200 $RUN --master-port $((port + 10)) bench/ingest_bench.py --gpus "$N" --json "$OUT/ingest.json"
step bench_ours 400 $RUN --master-port $((port + 6)) bench.py --gpus "$N" --steps 30 --warmup 5
grep -h '^{' "$OUT"/bigmlp_ours.log "$OUT"/bigmlp_standin.log "$OUT"/dlrm_ours.log "$OUT"/dlrm_variants.log "$OUT"/bench_ours.log > "$OUT/lines.json" 2>/dev/null
echo "suite done -> $OUT"
