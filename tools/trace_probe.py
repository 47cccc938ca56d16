#!/usr/bin/env python
"""Where the microseconds of a BoW-gated search go: runs SearchByBoW and SearchForTriangulation with host arrays and with the
keyframes resident in HBM (orbgpu_kfstore) under ORBGPU_TRACE_SEARCH=1, which makes the library print the host-clock phases
of every call (merge walk, staging, enqueue, wait, finish) and the kernel time between two CUDA events to stderr.
    python tools/trace_probe.py 2>&1 | grep orbgpu
The scene builders live under tests/ (they use the oracle for nothing but the level tables here)."""
import os, sys
os.environ["ORBGPU_TRACE_SEARCH"] = "1"
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import __graft_entry__ as g
import oracle_lib, test_search as ts, test_triangulation as tt
pkg = g.load_package(); po = oracle_lib.OraclePort()
F1, F2, bounds = ts.two_frames(pkg, ts.device_extract(pkg))
(k1, d1), (k2, d2) = F1, F2
s1, s2, keep = ts.bow_sides(pkg, F1, F2, 3, False, 100)
store = pkg.KeyFrameStore()
sc8 = po.tables()[0][0]; sg8 = (sc8 * sc8).astype(np.float32)
a1, a2 = keep
i1 = store.add(k1, d1, None, a1["node_ids"], a1["node_start"], a1["indices"], sc8, sg8)
i2 = store.add(k2, d2, None, a2["node_ids"], a2["node_start"], a2["indices"], sc8, sg8)
print("--- plain bow", file=sys.stderr)
for _ in range(6): pkg.search_by_bow(s1, s2)
print("--- resident bow", file=sys.stderr)
for _ in range(6): store.search_by_bow(i1, a1["valid"], i2, a2["valid"])
G1, G2, _ = ts.two_frames(pkg, ts.device_extract(pkg), 640, 480, 0x7A1A + 1000, 1000)
args, keep, T1 = tt.tri_scene(pkg, po, G1, G2, 4, 100, False, True, 0.5)
sc, sg = keep[2]
j1 = store.add(keep[0][0], keep[0][1], keep[0][3], keep[0][4], keep[0][5], keep[0][6], sc, sg, tt.CAM)
j2 = store.add(keep[1][0], keep[1][1], keep[1][3], keep[1][4], keep[1][5], keep[1][6], sc, sg, tt.CAM)
job, kj = store.triangulation_job(j1, keep[0][2], j2, keep[1][2], np.ctypeslib.as_array(args.F12), np.ctypeslib.as_array(args.Cw1), np.ctypeslib.as_array(args.T2w), False, True)
print("--- plain tri", file=sys.stderr)
for _ in range(6): pkg.search_for_triangulation(args)
print("--- resident tri", file=sys.stderr)
for _ in range(6): store.search_for_triangulation(job)
print("--- resident bow, 400 calls back to back (last three)", file=sys.stderr)
for _ in range(400): store.search_by_bow(i1, a1["valid"], i2, a2["valid"])
import subprocess
print(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,pstate", "--format=csv,noheader"], capture_output=True, text=True).stdout, file=sys.stderr)
