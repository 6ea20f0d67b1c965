#!/bin/bash
# parity tests + bench A/B over an environment switch.  usage: bash tools/gpu_ab.sh <tag> <ENVVAR> <valA> <valB>
tag=$1; var=$2; a=$3; b=$4
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit=$?"
tail -n 12 gpurun_out/${tag}_pytest.log
for v in $a $b; do
  env $var=$v timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench_${var}_$v.json 2> gpurun_out/${tag}_bench_$v.err
  echo "bench $var=$v exit=$?"; tail -n 3 gpurun_out/${tag}_bench_$v.err
  python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench_${var}_$v.json"))
print("$var=$v", round(d["value"],1), "frames/s  e2e", round(d["e2e"]["value"],1), "conv TF/s", round(d["roofline"]["achieved"],1), d["clocks"], d["kernel_time_share"])
PY
done
