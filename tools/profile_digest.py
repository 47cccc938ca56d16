#!/usr/bin/env python
"""Turns the outputs of tools/profile_round.sh into the files kept under profiles/:
   <tag>_launches.csv (copied), <tag>_launch_shares.json, <tag>_ncu_full_summary.txt, ncu_traffic.json.
usage: python tools/profile_digest.py <tag>   (run where ncu is installed; reads gpurun_out/<tag>_*)"""
import collections
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CMD = "python bench.py --frames 192 --steps 2 --warmup 3 --no-cpu --hamming-keyframes 0 --no-workloads --sustained-seconds 0"
STAGE = {"k_resize_level": "pyramid", "k_fast_cells": "fast_cells", "k_octree": "octree", "k_blur": "blur", "k_describe": "describe"}


def kernel_base(name):
    """'void k_fast_cells<48>(OrbPlanDev)' -> 'k_fast_cells'"""
    n = name.split("(")[0].strip()
    if n.startswith("void "):
        n = n[5:]
    return n.split("<")[0]


def main(tag):
    go, pr = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
    # ---- launch list -> shares
    src = os.path.join(go, tag + "_launches.csv")
    lines = [l for l in open(src) if not l.startswith("==")]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    tot = collections.defaultdict(float); cnt = collections.Counter()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = kernel_base(r["Kernel Name"])
        v = float(r["Metric Value"].replace(",", ""))
        v = v / 1000.0 if r["Metric Unit"] in ("ns", "nsecond") else v
        tot[k] += v; cnt[k] += 1
    path = {k: v for k, v in tot.items() if k in STAGE}
    s = sum(path.values())
    shares = {"command": CMD + " (default batching: device path 4 streams x 64-frame chunks)",
              "note": "ncu --metrics gpu__time_duration.sum --clock-control none: per-launch times are cold-cache and serialised; compare SHARES",
              "kernels": {k: {"launches": cnt[k], "total_us": v, "avg_us": v / cnt[k], "share": v / s} for k, v in sorted(path.items(), key=lambda kv: -kv[1])},
              "other_kernels": {k: {"launches": cnt[k], "total_us": v} for k, v in tot.items() if k not in STAGE}}
    json.dump(shares, open(os.path.join(pr, tag + "_launch_shares.json"), "w"), indent=1)
    shutil.copy(src, os.path.join(pr, tag + "_launches.csv"))
    # ---- full capture -> summary + dram traffic per frame
    rep = os.path.join(go, tag + "_full.ncu-rep")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    open("/tmp/_raw.csv", "w").write(raw)
    summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), "/tmp/_raw.csv"], capture_output=True, text=True).stdout
    open(os.path.join(pr, tag + "_ncu_full_summary.txt"), "w").write(summ)
    rr = list(csv.reader(io.StringIO(raw)))
    hdr, units = rr[0], rr[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        v = float(r[ix[name]].replace(",", "")); u = units[ix[name]]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    frames = None
    dram = collections.defaultdict(float); inst = collections.defaultdict(float)
    for r in rr[2:]:
        k = kernel_base(r[ix["Kernel Name"]])
        g = r[ix["Grid Size"]].strip("()").split(",")
        frames = int(g[1])
        dram[STAGE[k]] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
        inst[STAGE[k]] += val(r, "smsp__inst_executed.sum")
    traffic = {"source": "ncu --set full --clock-control none, one chunk (%d frames, C1) of `%s`; profiles/%s_ncu_full_summary.txt" % (frames, CMD, tag),
               "frames_per_launch": frames, "dram_bytes_per_frame": {k: v / frames for k, v in dram.items()},
               "warp_instructions_per_frame": {k: v / frames for k, v in inst.items()}}
    json.dump(traffic, open(os.path.join(pr, "ncu_traffic.json"), "w"), indent=1)
    print(json.dumps({"shares": {k: round(v["share"], 3) for k, v in shares["kernels"].items()}, "traffic": traffic}, indent=1))


if __name__ == "__main__":
    main(sys.argv[1])
