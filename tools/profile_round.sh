#!/bin/bash
# usage (on the GPU box, one GPU): tools/profile_round.sh <tag>
# 1. plain run of the profiled command (must exit 0), 2. ncu launch list of the same command, 3. one ncu --set full
# capture of one device-path chunk (64 frames: 7 resize + FAST + octree + blur + describe).  Outputs in gpurun_out/.
tag="${1:-rX}"
CMD="python bench.py --frames 192 --steps 2 --warmup 3 --no-cpu --hamming-keyframes 0 --no-workloads --sustained-seconds 0"
$CMD > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "plain run failed"; tail -3 gpurun_out/${tag}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${tag}_launches.csv $CMD > gpurun_out/${tag}_ncu1.log 2>&1
# device-path warm-up: 3 steps x 3 chunks x 11 launches = 99 launches; capture the first timed chunk
ncu --set full --import-source on --clock-control none --launch-skip 99 -c 11 \
    -k regex:"k_resize_level|k_fast_cells|k_blur|k_describe|k_octree" -f -o gpurun_out/${tag}_full $CMD > gpurun_out/${tag}_ncu2.log 2>&1
ls -la gpurun_out/${tag}_*
