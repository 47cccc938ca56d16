#!/usr/bin/env python
"""Phase trace (ORBGPU_TRACE_SEARCH=1) of the two matcher calls of every tracked frame: SearchByProjection(CurrentFrame,
LastFrame) and Tracking::SearchLocalPoints in one call.  python tools/trace_probe2.py 2>&1 | grep orbgpu"""
import os, sys
os.environ["ORBGPU_TRACE_SEARCH"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as g
import oracle_lib, test_search as ts, test_local_map as tm
pkg = g.load_package(); po = oracle_lib.OraclePort()
F1, F2, bounds = ts.two_frames(pkg, ts.device_extract(pkg))
(k1, d1), (k2, d2) = F1, F2
args, keep = ts.projection_scene(pkg, oracle_lib, F1, F2, bounds, 0.0, True)
print("--- search_by_projection 1000x1000", file=sys.stderr)
for _ in range(6): pkg.search_by_projection(args)
G1 = (np.concatenate([k1] * 4), np.concatenate([d1] * 4))
args2, keep2 = tm.local_map_scene(pkg, po, G1, F2, bounds, 1.0, 1)
print("--- search_local_map 4000x1000", file=sys.stderr)
for _ in range(6): pkg.search_local_map(args2)
