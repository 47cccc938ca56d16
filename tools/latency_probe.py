#!/usr/bin/env python
"""Single-frame latency of orbgpu_extract (the live-SLAM case: one frame in flight) and of the submit/collect queue.
usage: python tools/latency_probe.py [C1|C2|C3|C4]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g   # noqa: E402

SHAPES = {"C1": (640, 480, 1000), "C2": (1241, 376, 2000), "C3": (752, 480, 1000), "C4": (1920, 1080, 4000)}


def main():
    import torch
    pkg = g.load_package()
    name = sys.argv[1] if len(sys.argv) > 1 else "C1"
    w, h, nf = SHAPES[name]
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7)
    cap = ex.max_keypoints(w, h)
    frames = [torch.from_numpy(pkg.synth.synth(w, h, 0x5EED0000 + i)).pin_memory() for i in range(8)]
    kps = torch.empty((cap, 28), dtype=torch.uint8).pin_memory(); desc = torch.empty((cap, 32), dtype=torch.uint8).pin_memory()
    nout = torch.zeros(1, dtype=torch.int32).pin_memory()
    import ctypes as C
    L = pkg.lib()

    def one(i):
        f = frames[i % len(frames)]
        rc = L.orbgpu_extract(ex._h, f.data_ptr(), w, h, w, kps.data_ptr(), desc.data_ptr(), cap, C.cast(nout.data_ptr(), C.POINTER(C.c_int)))
        assert rc == 0, rc
    for i in range(20):
        one(i)
    t = []
    for i in range(200):
        t0 = time.perf_counter(); one(i); t.append(time.perf_counter() - t0)
    t = np.array(t) * 1e3
    ex.set_profiling(True)
    for i in range(50):
        one(i)
    st = ex.stage_ms(reset=True)
    ex.set_profiling(False)
    out = {"workload": name, "api": "orbgpu_extract, pinned host buffers, one frame in flight",
           "latency_ms": {"median": float(np.median(t)), "p10": float(np.percentile(t, 10)), "p90": float(np.percentile(t, 90))},
           "stage_ms_per_frame": {k: v / 50 for k, v in st.items()}, "keypoints": int(nout[0])}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
