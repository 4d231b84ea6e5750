#!/bin/bash
# Developer tool (GPU box with N GPUs): bench.py at the driver's command line for N = 1,2,4,8 plus variants of the
# gradient-bucket slicing. Lines go to gpurun_out/sweep_*.json.
mkdir -p gpurun_out
MAXN=${1:-8}
run() {  # n tag extra-args...
  n=$1; tag=$2; shift 2
  if [ "$n" = 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 --no-e2e --no-cpu-baseline "$@" > gpurun_out/sweep_${tag}_n$n.json 2> gpurun_out/sweep_${tag}_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $n --steps 20 --warmup 5 --no-e2e "$@" > gpurun_out/sweep_${tag}_n$n.json 2> gpurun_out/sweep_${tag}_n$n.err
  fi
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/sweep_${tag}_n$n.json").read().strip().splitlines()[-1])
    print("${tag}", "N=$n", d["value"], d["ms_per_step"], d.get("allreduce", {}).get("alone_ms"))
except Exception as ex:
    print("${tag} N=$n FAILED", ex)
PY
}
run 1 base
for n in 2 4 8; do
  [ $n -le $MAXN ] || continue
  run $n s4
  run $n s1 --slices 1
  run $n s2 --slices 2
  run $n s8 --slices 8
  run $n s4np --nccl-normal-priority
  run $n ref --impl reference
done
run 1 ref --impl reference
