#!/usr/bin/env python
"""A short pass over every kernel family for compute-sanitizer (memcheck / racecheck): the smoke extraction (C1) and match,
one 1920x1080 frame with 4000 features (C4), a batch through the graph-replayed host path, the octree sizes that spill from
shared memory to the global scratch area, one grid-gated search, one BoW search, the all-pairs count and the no-allocation
match.  Every result is checked against the oracle, so a sanitizer-clean run is also a correct one.
usage (one tool per gpurun call): compute-sanitizer --tool memcheck python tools/sanitize_cases.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g   # noqa: E402
import oracle_lib             # noqa: E402


def main():
    pkg = g.load_package()
    synth = pkg.synth
    po = oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)
    # C1 single frame + brute-force match (what smoke() runs)
    img = synth.synth(640, 480, 0x5EED0000)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    k, d = ex(img)
    ok, od = po.extract(img)
    assert k.tobytes() == ok.tobytes() and np.array_equal(d, od)
    m = pkg.ORBmatcher(0.6, True)
    got = m.match_bruteforce(d[:256], d[256:768]); want = po.match_bruteforce(d[:256], d[256:768])
    assert all(np.array_equal(a, b) for a, b in zip(got, want))
    print("C1 frame + match ok", flush=True)
    # a small batch through the host path (graph replay, several chunk sizes)
    frames = np.stack([synth.synth(640, 480, 0x5EED0000 + i) for i in range(5)])
    kb, db, nb = ex.extract_batch(frames)
    kb2, db2, nb2 = ex.extract_batch(frames[:3])
    assert nb[0] == len(ok) and kb[0, :nb[0]].tobytes() == ok.tobytes() and np.array_equal(nb[:3], nb2)
    print("batch ok", flush=True)
    # C4: one 1920 x 1080 frame, 4000 features
    big = synth.synth(1920, 1080, 0x5EED3000)
    ex4 = pkg.ORBextractor(4000, 1.2, 8, 20, 7)
    k4, d4 = ex4(big)
    o4k, o4d = oracle_lib.OraclePort(4000, 1.2, 8, 20, 7).extract(big)
    assert k4.tobytes() == o4k.tobytes() and np.array_equal(d4, o4d)
    print("C4 frame ok (%d keypoints)" % len(k4), flush=True)
    # octree on candidate lists that spill keys / node tables to global scratch
    for n, N, W, H in [(5000, 217, 608, 448), (20000, 869, 1888, 1048), (1500, 4000, 608, 448)]:
        rng = np.random.default_rng(n + N)
        flat = rng.choice(W * H, size=min(n, W * H), replace=False)
        xs = (flat % W).astype(np.int32); ys = (flat // W).astype(np.int32); resp = rng.integers(0, 256, len(flat)).astype(np.int32)
        assert np.array_equal(ex.octree(xs, ys, resp, W, H, N), oracle_lib.OraclePort().octree(xs, ys, resp, W, H, N))
    print("octree spill cases ok", flush=True)
    # one grid-gated search and one BoW-gated search (section 8(f) rows)
    from test_search import bow_sides, two_frames
    F1, F2, bounds = two_frames(pkg, lambda im, nf: ex(im))
    (k1, d1), (k2, d2) = F1, F2
    pm0 = np.stack([k1["x"], k1["y"]], 1)
    gi = pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm0, 100, 0.9, True)
    wi = po.search_for_initialization(k1, d1, k2, d2, bounds, pm0, 100, 0.9, True)
    assert gi[0] == wi[0] and np.array_equal(gi[1], wi[1])
    s1, s2, keep = bow_sides(pkg, F1, F2, 3, True, 24)
    gb = pkg.search_by_bow(s1, s2, 0.7, True, True); wb = po.search_by_bow(s1, s2, 0.7, True, True)
    assert gb[0] == wb[0] and np.array_equal(gb[1], wb[1])
    print("search rows ok", flush=True)
    # all-pairs counts and the allocation-free match
    desc = synth.keyframe_descriptors(6, 300)
    m.set_keyframes(desc)
    pairs = pkg.shard.all_pairs(6)
    cnt = m.count_pairs(pairs[:, 0], pairs[:, 1])
    assert cnt.tolist() == [po.match_count(desc[a], desc[b], 50, 0.6) for a, b in pairs]
    L = pkg.lib()
    A = np.ascontiguousarray(d[:200]); B = np.ascontiguousarray(d[200:700])
    bi = np.zeros(200, np.int32); b1 = np.zeros(200, np.uint16); b2 = np.zeros(200, np.uint16)
    assert L.orbgpu_match_bruteforce_noalloc(0, A.ctypes.data, 200, B.ctypes.data, 500, bi.ctypes.data, b1.ctypes.data, b2.ctypes.data) == 0
    w = po.match_bruteforce(A, B)
    assert np.array_equal(bi, w[0]) and np.array_equal(b1, w[1]) and np.array_equal(b2, w[2])
    print("all-pairs + noalloc ok", flush=True)
    print("SANITIZE CASES PASSED", flush=True)


if __name__ == "__main__":
    main()
