#!/bin/bash
# usage: tools/variant_run.sh <tag> "<nvcc -D flags>"   -- rebuilds liborbgpu.so with the flags on the GPU box, checks parity, runs a short bench
tag=$1; flags=$2
make -s -B -f my-slam_b200/csrc/Makefile EXTRA="$flags" > gpurun_out/${tag}_build.log 2>&1 || { echo "$tag BUILD FAILED"; tail -5 gpurun_out/${tag}_build.log; exit 1; }
python -m pytest tests/test_gpu_extract.py -x -q -m gpu 2>&1 | tail -n 1
python bench.py --steps 10 --warmup 3 --no-cpu --no-workloads --hamming-keyframes 0 --sustained-seconds 0 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { echo "$tag BENCH FAILED"; tail -3 gpurun_out/${tag}_bench.err; }
python tools/latency_probe.py C1 > gpurun_out/${tag}_lat.json 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench.json")); l=json.load(open("gpurun_out/${tag}_lat.json"))
print("$tag value %.0f e2e %.0f pyramid_ser %.3f lat %.4f" % (d["value"], d["e2e"]["value"], d["roofline"]["stage_ms_per_step_serialized"]["pyramid"], l["latency_ms"]["median"]))
PY
