#!/bin/bash
for cfg in "4 128" "4 96" "3 128" "5 128" "6 96" "8 64" "4 160" "2 256"; do
  set -- $cfg
  python bench.py --steps 20 --warmup 5 --no-cpu --no-workloads --hamming-keyframes 0 --sustained-seconds 0 --streams $1 --chunk $2 > gpurun_out/sw_$1_$2.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/sw_$1_$2.json"))
print("streams $1 chunk $2 value %.0f e2e %.0f" % (d["value"], d["e2e"]["value"]))
PY
done
