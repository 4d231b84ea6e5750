#!/bin/bash
# Developer tool (N-GPU box): the round's multi-GPU measurements into gpurun_out/. Usage: tools/final_multi.sh N [c2|c4|c5 ...]
N=$1; shift
mkdir -p gpurun_out
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port=29600
run() {  # tag args...
  tag=$1; shift
  port=$((port + 1))
  $T --master-port $port bench.py --gpus $N "$@" > gpurun_out/r02_${tag}_n$N.json 2> gpurun_out/r02_${tag}_n$N.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_${tag}_n$N.json").read().strip().splitlines()[-1])
    print("${tag} N=$N", d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"), d.get("allreduce"))
except Exception as ex:
    print("${tag} N=$N FAILED", ex)
PY
}
for what in "$@"; do
  case $what in
    c2) run bench_ours --steps 20 --warmup 5; run bench_ref --impl reference --steps 20 --warmup 5 ;;
    c2d0) run bench_ours_d0 --steps 20 --warmup 5 --sh-degree 0 --no-e2e ;;
    c4) run c4_ours --config C4 --steps 10 --warmup 3 --no-e2e; run c4_ref --config C4 --impl reference --steps 4 --warmup 3 ;;
    c5) run c5_ours --config C5 --steps 30 --warmup 5; run c5_ref --config C5 --impl reference --steps 6 --warmup 3 ;;
    comm) timeout 300 python -m pytest tests/test_gpu_comm.py -m gpu -q 2>&1 | tail -2 ;;
  esac
done
