#!/bin/bash
# Developer tool (1-GPU box): the round's single-GPU measurements into gpurun_out/.
mkdir -p gpurun_out
B="timeout 600 python bench.py"
$B > gpurun_out/r02_bench_ours.json 2> gpurun_out/r02_bench_ours.err
$B --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_ref.json 2>/dev/null
$B --sh-degree 0 --steps 100 > gpurun_out/r02_bench_ours_d0.json 2>/dev/null
$B --sh-degree 0 --impl reference --steps 20 --warmup 5 > gpurun_out/r02_bench_ref_d0.json 2>/dev/null
$B --sh-degree 1 --steps 100 --no-cpu-baseline > gpurun_out/r02_bench_ours_d1.json 2>/dev/null
$B --config C5 --steps 50 --no-cpu-baseline > gpurun_out/r02_c5_ours_n1.json 2>/dev/null
$B --config C5 --impl reference --steps 10 --warmup 3 > gpurun_out/r02_c5_ref_n1.json 2>/dev/null
$B --config C4 --steps 20 --no-cpu-baseline > gpurun_out/r02_c4_ours_n1.json 2> gpurun_out/r02_c4_ours_n1.err
$B --config C4 --impl reference --steps 5 --warmup 3 > gpurun_out/r02_c4_ref_n1.json 2>/dev/null
$B --config C3 --steps 30 --no-cpu-baseline > gpurun_out/r02_c3_ours_n1.json 2>/dev/null
$B --config C3 --impl reference --steps 5 --warmup 3 > gpurun_out/r02_c3_ref_n1.json 2>/dev/null
$B --config C1 --steps 100 --views-per-step 4 --no-cpu-baseline > gpurun_out/r02_c1_ours_n1.json 2>/dev/null
$B --config C1 --impl reference --steps 20 --views-per-step 4 > gpurun_out/r02_c1_ref_n1.json 2>/dev/null
for f in gpurun_out/r02_*.json; do python - <<PY
import json
try:
    d = json.loads(open("$f").read().strip().splitlines()[-1])
    print("$f", d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value"), (d.get("roofline") or {}).get("view", {}).get("frac"))
except Exception as ex:
    print("$f FAILED", ex)
PY
done
