#!/bin/bash
# Developer tool (GPU box): bench.py stage times for every library variant in var/ (built by tools/variants.sh).
mkdir -p gpurun_out
cp das3r_b200/libdas3r_raster.so /tmp/base_lib.so
for so in var/*.so; do
  name=$(basename $so .so)
  cp $so das3r_b200/libdas3r_raster.so
  timeout 200 python bench.py --steps 20 --warmup 5 --no-e2e --no-cpu-baseline "$@" > gpurun_out/var_$name.json 2> gpurun_out/var_$name.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/var_$name.json").read().strip().splitlines()[-1])
    st = d["roofline"]["stage_ms"]
    print("$name", d["value"], d["ms_per_step"], {k: round(v, 3) for k, v in st.items()})
except Exception as ex:
    print("$name FAILED", ex)
PY
done
cp /tmp/base_lib.so das3r_b200/libdas3r_raster.so
