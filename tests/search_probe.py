#!/usr/bin/env python
"""Small driver for profiling the frame-side search rows under ncu (uses the oracle scene builders, hence under tests/):
   ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/x.csv python tests/search_probe.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g          # noqa: E402
import oracle_lib                     # noqa: E402
import test_search as ts              # noqa: E402
import test_local_points as tl        # noqa: E402

pkg = g.load_package()
po = oracle_lib.OraclePort()
F1, F2, bounds = ts.two_frames(pkg, ts.device_extract(pkg))
(k1, d1), (k2, d2) = F1, F2
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
for rep, th in [(1, 1.0), (4, 3.0)]:
    G1 = (np.concatenate([k1] * rep), np.concatenate([d1] * rep))
    args, keep = tl.local_scene(pkg, po, G1, F2, bounds, th, 1)
    for _ in range(reps):
        pkg.search_local_points(args)
args, keep = ts.projection_scene(pkg, oracle_lib, F1, F2, bounds, 0.0, True)
for _ in range(reps):
    pkg.search_by_projection(args)
pm = np.stack([k1["x"], k1["y"]], 1)
for _ in range(reps):
    pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm)
