#!/bin/bash
# usage: tools/bench_n.sh <N> <tag> [bench.py flags]   -- one bench line under torchrun on N GPUs, JSON to gpurun_out/<tag>.json
N=$1; tag=$2; shift 2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N "$@" \
    > gpurun_out/$tag.json 2> gpurun_out/$tag.err || { echo "$tag FAILED"; tail -5 gpurun_out/$tag.err; exit 1; }
python - <<PY
import json
d=json.load(open("gpurun_out/$tag.json"))
e=d["e2e"]; cc=e.get("copy_ceiling",{})
print("$tag N=%d value %.0f e2e %.0f sync %.0f ceiling %.0f (h2d %.1f GB/s) sustained %.0f" % (d["n_gpus"], d["value"], e["value"], e["synchronous_call_value"], cc.get("frames_per_s",0), cc.get("h2d_gbs",0), d.get("sustained",{}).get("frames_per_s",0)))
h=d.get("hamming")
if h: print("   hamming %.1f G dist/s kf %d kernel %.2f ms gather %.3f ms imbalance %.3f" % (h["distance_evals_per_s"]/1e9, h["keyframes"], h["kernel_ms"], h["gather_ms"], h["breakdown"]["imbalance_ms"]))
w=d.get("workloads",{}).get("C3")
if w: print("   C3 seq value %.0f e2e %.0f parity %s" % (w["value"], w["e2e"]["value"], w["parity_vs_one_gpu"]))
PY
