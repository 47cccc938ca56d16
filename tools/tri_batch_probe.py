#!/usr/bin/env python
"""SearchForTriangulation for 20 neighbouring keyframes in one call (LocalMapping::CreateNewMapPoints), per pair: host arrays,
keyframes resident in HBM, and one resident pair per call.  python tools/tri_batch_probe.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as g
import oracle_lib, test_search as ts, test_triangulation as tt
pkg = g.load_package(); po = oracle_lib.OraclePort()
G1, G2, _ = ts.two_frames(pkg, ts.device_extract(pkg), 640, 480, 0x7A1A + 1000, 1000)
many = [tt.tri_scene(pkg, po, G1, G2, 4 + i, 100, False, True, 0.5) for i in range(20)]
bl = [m[0] for m in many]


def med(f, reps=30):
    for _ in range(3): f()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); f(); t.append(time.perf_counter() - t0)
    return float(np.median(t)) * 1e3


want = [po.search_for_triangulation(b) for b in bl]
got = pkg.search_for_triangulation_batch(bl)
assert all(x[0] == y[0] and np.array_equal(x[1], y[1]) for x, y in zip(got, want))
print("batch, arrays, per pair    %.4f ms" % (med(lambda: pkg.search_for_triangulation_batch(bl)) / 20))
store = pkg.KeyFrameStore(); jl = []; kl = []
for m in many:
    kp = m[1]; sc, sg = kp[2]
    x1 = store.add(kp[0][0], kp[0][1], kp[0][3], kp[0][4], kp[0][5], kp[0][6], sc, sg, tt.CAM)
    x2 = store.add(kp[1][0], kp[1][1], kp[1][3], kp[1][4], kp[1][5], kp[1][6], sc, sg, tt.CAM)
    jb = store.triangulation_job(x1, kp[0][2], x2, kp[1][2], np.ctypeslib.as_array(m[0].F12), np.ctypeslib.as_array(m[0].Cw1),
                                 np.ctypeslib.as_array(m[0].T2w), False, True)
    jl.append(jb[0]); kl.append(jb[1])
got = store.search_for_triangulation_batch(jl)
assert all(x[0] == y[0] and np.array_equal(x[1], y[1]) for x, y in zip(got, want))
print("batch, resident, per pair  %.4f ms" % (med(lambda: store.search_for_triangulation_batch(jl)) / 20))
print("one resident pair per call %.4f ms" % med(lambda: store.search_for_triangulation(jl[0])))
print("CPU restatement, per pair  %.4f ms" % (med(lambda: [po.search_for_triangulation(b) for b in bl], 10) / 20))
