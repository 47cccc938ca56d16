#!/bin/bash
# usage (on the GPU box): tools/sweep.sh "<EXTRA flags>" [bench args...]  -> rebuilds liborbgpu.so with the flags, runs bench, prints one line
flags="$1"; shift
touch my-slam_b200/csrc/orb_kernels.cu
make -s -f my-slam_b200/csrc/Makefile EXTRA="$flags" > /dev/null 2>gpurun_out/sweep_build.err || { echo "BUILD FAILED $flags"; exit 1; }
python bench.py --no-cpu "$@" > gpurun_out/sw.json 2>gpurun_out/sw.err || { echo "RUN FAILED $flags"; tail -3 gpurun_out/sw.err; exit 1; }
python - "$flags" "$*" <<'PY'
import json,sys
d=json.load(open("gpurun_out/sw.json")); st=d["roofline"]["stage_ms_per_step_serialized"]
print("%-40s %-22s dev %7d e2e %7d | %s" % (sys.argv[1], sys.argv[2], d["value"], d["e2e"]["value"], " ".join("%s %.3f" % (k[:4], v) for k, v in st.items() if k not in ("upload","download"))))
PY
