#!/usr/bin/env python
"""Timing of the SURVEY section-8(f) rows next to their CPU restatements (one thread, the reference's execution model
for these functions): python tests/rows_bench.py > profiles/<round>_rows.json (lives under tests/ because it uses the oracle as checker and CPU baseline).  Each GPU figure is the wall time of the
public C-ABI call (host buffers in, host buffers out; synchronous), median of `reps` calls after warm-up; the CPU
figure is oracle/frame_port.cc / orb_port.cc on the same inputs.  Results of both are compared before timing."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))   # this file's own directory
import __graft_entry__ as g          # noqa: E402
import oracle_lib                     # noqa: E402  (checker + CPU baseline only)
import test_search as ts              # noqa: E402  (scene builders)


def med(f, reps=30, warm=3):
    for _ in range(warm):
        f()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); f(); t.append(time.perf_counter() - t0)
    return float(np.median(t)) * 1e3


def run():
    """-> {row: {"gpu_ms": ..., "cpu_ms": ...}}; every row's results are checked against the oracle before it is timed."""
    pkg = g.load_package()
    po = oracle_lib.OraclePort()
    out = {"unit": "ms per call (median)", "cpu": "oracle restatement, 1 thread"}
    # --- stereo: Frame::ComputeStereoMatches
    left, right = pkg.synth.stereo_pair(640, 480, 0x57E2E001)
    exL = pkg.ORBextractor(1000, 1.2, 8, 20, 7); exR = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    kL, dL = exL(left); kR, dR = exR(right)
    pl = oracle_lib.OraclePort(); pr = oracle_lib.OraclePort()
    pl.extract(left); pr.extract(right)
    lv = [pl.level_pixels(l) for l in range(8)], [pr.level_pixels(l) for l in range(8)]
    bf, b = 386.1448, 0.5372
    u, d = pkg.stereo_match(exL, exR, bf, b)
    ou, od = po.stereo_matches(kL, dL, kR, dR, lv[0], lv[1], bf, b)
    assert u.tobytes() == ou.tobytes() and d.tobytes() == od.tobytes()
    out["stereo_match_640x480_1000kp"] = {"gpu_ms": med(lambda: pkg.stereo_match(exL, exR, bf, b)),
                                          "cpu_ms": med(lambda: po.stereo_matches(kL, dL, kR, dR, lv[0], lv[1], bf, b), 10, 1),
                                          "matched": int((u >= 0).sum())}
    # --- undistort
    K, D = (517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
    import ctypes as C
    L = C.CDLL(oracle_lib.PORT_SO)
    L.orbo_undistort_points.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    Kf, Df = np.array(K, np.float32), np.array(D, np.float32)

    def cpu_und(p):
        o = np.zeros_like(p); L.orbo_undistort_points(p.ctypes.data, len(p), Kf.ctypes.data, Df.ctypes.data, len(Df), o.ctypes.data); return o
    pts = np.ascontiguousarray(np.stack([kL["x"], kL["y"]], 1))
    assert np.array_equal(np.stack([exL.undistort_keypoints(K, D)[f] for f in ("x", "y")], 1), cpu_und(pts))
    out["undistort_keypoints_1000"] = {"gpu_ms": med(lambda: exL.undistort_keypoints(K, D)), "cpu_ms": med(lambda: cpu_und(pts))}
    # the Frame constructor's ExtractORB + UndistortKeyPoints: two calls against the fused one (the extraction is in both)
    fk, fu, fd = exL.extract_undistorted(left, K, D)
    assert fk.tobytes() == kL.tobytes() and fu.tobytes() == exL.undistort_keypoints(K, D).tobytes()
    out["extract_then_undistort_640x480_1000kp"] = {
        "gpu_ms_two_calls": med(lambda: (exL(left), exL.undistort_keypoints(K, D))),
        "gpu_ms": med(lambda: exL.extract_undistorted(left, K, D)),
        "gpu_ms_extract_alone": med(lambda: exL(left)),
        "cpu_ms": med(lambda: (pl.extract(left), cpu_und(pts)), 5, 1)}
    big = (np.random.default_rng(1).random((1 << 22, 2)) * np.array([640.0, 480.0])).astype(np.float32)
    assert pkg.undistort_points(big, K, D).tobytes() == cpu_und(big).tobytes()
    out["undistort_points_4M"] = {"gpu_ms": med(lambda: pkg.undistort_points(big, K, D), 5, 1), "cpu_ms": med(lambda: cpu_und(big), 3, 1)}
    # --- searches
    F1, F2, bounds = ts.two_frames(pkg, ts.device_extract(pkg))
    (k1, d1), (k2, d2) = F1, F2
    pm = np.stack([k1["x"], k1["y"]], 1)
    a = pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm); o = po.search_for_initialization(k1, d1, k2, d2, bounds, pm)
    assert a[0] == o[0] and np.array_equal(a[1], o[1])
    out["search_for_initialization_1000x1000"] = {"gpu_ms": med(lambda: pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm)),
                                                  "cpu_ms": med(lambda: po.search_for_initialization(k1, d1, k2, d2, bounds, pm)),
                                                  "nmatches": a[0]}
    args, keep = ts.projection_scene(pkg, oracle_lib, F1, F2, bounds, 0.0, True)
    a = pkg.search_by_projection(args); o = po.search_by_projection(args)
    assert a[0] == o[0] and np.array_equal(a[1], o[1])
    out["search_by_projection_1000x1000"] = {"gpu_ms": med(lambda: pkg.search_by_projection(args)),
                                             "cpu_ms": med(lambda: po.search_by_projection(args)), "nmatches": a[0]}
    s1, s2, keep = ts.bow_sides(pkg, F1, F2, 3, False, 100)
    a = pkg.search_by_bow(s1, s2); o = po.search_by_bow(s1, s2)
    assert a[0] == o[0] and np.array_equal(a[1], o[1])
    out["search_by_bow_1000x1000_100nodes"] = {"gpu_ms": med(lambda: pkg.search_by_bow(s1, s2)), "cpu_ms": med(lambda: po.search_by_bow(s1, s2)),
                                              "nmatches": a[0]}
    # the same searches with the keyframe(s) resident in HBM (orbgpu_kfstore): keyframe/frame (tracking) and keyframe/keyframe
    store = pkg.KeyFrameStore()
    sc8 = po.tables()[0][0]; sg8 = (sc8 * sc8).astype(np.float32)
    a1, a2 = keep
    i1 = store.add(k1, d1, None, a1["node_ids"], a1["node_start"], a1["indices"], sc8, sg8)
    i2 = store.add(k2, d2, None, a2["node_ids"], a2["node_start"], a2["indices"], sc8, sg8)
    r = store.search_by_bow(i1, a1["valid"], -1, None, s2)
    assert r[0] == o[0] and np.array_equal(r[1], o[1])
    out["search_by_bow_1000x1000_100nodes"]["gpu_ms_keyframe_resident"] = med(lambda: store.search_by_bow(i1, a1["valid"], -1, None, s2))
    r = store.search_by_bow(i1, a1["valid"], i2, a2["valid"])
    assert r[0] == o[0] and np.array_equal(r[1], o[1])
    out["search_by_bow_1000x1000_100nodes"]["gpu_ms_both_resident"] = med(lambda: store.search_by_bow(i1, a1["valid"], i2, a2["valid"]))
    # --- local-map search: SearchByProjection(Frame, vector<MapPoint*>, th), 1000 and 4000 map points against 1000 features
    import test_local_points as tl
    for rep, th in [(1, 1.0), (4, 3.0)]:
        G1 = (np.concatenate([k1] * rep), np.concatenate([d1] * rep))
        args, keep = tl.local_scene(pkg, po, G1, F2, bounds, th, 1)
        a = pkg.search_local_points(args); o = po.search_local_points(args)
        assert a[0] == o[0] and np.array_equal(a[1], o[1])
        out["search_local_points_%dx%d_th%d" % (len(G1[0]), len(k2), int(th))] = {
            "gpu_ms": med(lambda: pkg.search_local_points(args)), "cpu_ms": med(lambda: po.search_local_points(args)), "nmatches": a[0]}
    # --- SearchForTriangulation: 1000 x 1000 and 2000 x 2000 features in 100 vocabulary nodes
    import test_triangulation as tt
    for nf in (1000, 2000):
        G1, G2, _ = ts.two_frames(pkg, ts.device_extract(pkg), 640, 480, 0x7A1A + nf, nf)
        args, keep, T1 = tt.tri_scene(pkg, po, G1, G2, 4, 100, False, True, 0.5)
        a = pkg.search_for_triangulation(args); o = po.search_for_triangulation(args)
        assert a[0] == o[0] and np.array_equal(a[1], o[1])
        out["search_for_triangulation_%dx%d_100nodes" % (len(G1[0]), len(G2[0]))] = {
            "gpu_ms": med(lambda: pkg.search_for_triangulation(args)), "cpu_ms": med(lambda: po.search_for_triangulation(args)), "nmatches": a[0]}
        sc, sg = keep[2]
        j1 = store.add(keep[0][0], keep[0][1], keep[0][3], keep[0][4], keep[0][5], keep[0][6], sc, sg, tt.CAM)
        j2 = store.add(keep[1][0], keep[1][1], keep[1][3], keep[1][4], keep[1][5], keep[1][6], sc, sg, tt.CAM)
        job, kj = store.triangulation_job(j1, keep[0][2], j2, keep[1][2], np.ctypeslib.as_array(args.F12), np.ctypeslib.as_array(args.Cw1),
                                          np.ctypeslib.as_array(args.T2w), False, True)
        r = store.search_for_triangulation(job)
        assert r[0] == o[0] and np.array_equal(r[1], o[1])
        out["search_for_triangulation_%dx%d_100nodes" % (len(G1[0]), len(G2[0]))]["gpu_ms_resident"] = med(lambda: store.search_for_triangulation(job))
        if nf == 1000:          # one keyframe against 20 neighbours in one call (CreateNewMapPoints), per pair
            many = [tt.tri_scene(pkg, po, G1, G2, 4 + i, 100, False, True, 0.5) for i in range(20)]
            bl = [m[0] for m in many]
            got = pkg.search_for_triangulation_batch(bl)
            assert all(g[0] == po.search_for_triangulation(b)[0] for g, b in zip(got, bl))
            out["search_for_triangulation_batch_20_pairs_%dx%d_per_pair" % (len(G1[0]), len(G2[0]))] = {
                "gpu_ms": med(lambda: pkg.search_for_triangulation_batch(bl)) / 20,
                "cpu_ms": med(lambda: [po.search_for_triangulation(b) for b in bl], 10, 1) / 20}
            jl, kl = [], []
            for m in many:
                kp = m[1]
                x1 = store.add(kp[0][0], kp[0][1], kp[0][3], kp[0][4], kp[0][5], kp[0][6], sc, sg, tt.CAM)
                x2 = store.add(kp[1][0], kp[1][1], kp[1][3], kp[1][4], kp[1][5], kp[1][6], sc, sg, tt.CAM)
                jb = store.triangulation_job(x1, kp[0][2], x2, kp[1][2], np.ctypeslib.as_array(m[0].F12), np.ctypeslib.as_array(m[0].Cw1),
                                             np.ctypeslib.as_array(m[0].T2w), False, True)
                jl.append(jb[0]); kl.append(jb[1])
            rg = store.search_for_triangulation_batch(jl)
            assert all(x[0] == y[0] and np.array_equal(x[1], y[1]) for x, y in zip(rg, got))
            out["search_for_triangulation_batch_20_pairs_%dx%d_per_pair" % (len(G1[0]), len(G2[0]))]["gpu_ms_resident"] = \
                med(lambda: store.search_for_triangulation_batch(jl)) / 20
    # --- Tracking::SearchLocalPoints in one call: isInFrustum for 4000 local map points + the search against 1000 features
    import test_local_map as tm
    G1 = (np.concatenate([k1] * 4), np.concatenate([d1] * 4))
    args, keep = tm.local_map_scene(pkg, po, G1, F2, bounds, 1.0, 1)
    a = pkg.search_local_map(args); o = po.search_local_map(args, pkg.track_fields)
    assert tm.same(a, o)
    out["search_local_map_%dx%d_th1" % (len(G1[0]), len(k2))] = {
        "gpu_ms": med(lambda: pkg.search_local_map(args)), "cpu_ms": med(lambda: po.search_local_map(args, pkg.track_fields)),
        "in_view": a[2], "nmatches": a[0]}
    # --- the search of ORBmatcher::Fuse: 4000 map points into a 1000-feature keyframe
    import test_fuse as tf
    fa, fkeep = tf.fuse_from_local_map(pkg, po, keep, 3.0, bounds)
    a = pkg.fuse_search(fa); o = po.fuse_search(fa)
    assert np.array_equal(a, o)
    out["fuse_search_%dx%d_th3" % (fa.n_points, fa.n_kf)] = {
        "gpu_ms": med(lambda: pkg.fuse_search(fa)), "cpu_ms": med(lambda: po.fuse_search(fa)), "fused": int((a >= 0).sum())}
    fl = [fa] * 20               # 20 target keyframes in one call (SearchInNeighbors), per keyframe
    assert all(np.array_equal(g, o) for g in pkg.fuse_search_batch(fl))
    out["fuse_search_batch_20_keyframes_%dx%d_per_keyframe" % (fa.n_points, fa.n_kf)] = {
        "gpu_ms": med(lambda: pkg.fuse_search_batch(fl)) / 20, "cpu_ms": med(lambda: [po.fuse_search(x) for x in fl], 10, 1) / 20}
    # --- MapPoint::ComputeDistinctiveDescriptors for the 2000 map points a new keyframe touches (up to 12 observations each)
    import test_distinctive as td
    dd, gs = td.groups(1, 2000, 12)
    a = pkg.distinctive_descriptors(dd, gs); o = po.distinctive_descriptors(dd, gs)
    assert np.array_equal(a, o)
    out["distinctive_descriptors_2000_points_%d_observations" % len(dd)] = {
        "gpu_ms": med(lambda: pkg.distinctive_descriptors(dd, gs)), "cpu_ms": med(lambda: po.distinctive_descriptors(dd, gs), 10, 1)}
    # --- the Sim3 routines of loop closing on the same 4000-point / 1000-feature snapshot, and SearchBySim3 between the two frames
    import test_sim3 as t3
    S = t3.sim3_of(keep["Tcw"], 1.1)
    a = pkg.fuse_search_sim3(fa, S); o = po.fuse_search_sim3(fa, S)
    assert np.array_equal(a, o)
    out["fuse_search_sim3_%dx%d" % (fa.n_points, fa.n_kf)] = {
        "gpu_ms": med(lambda: pkg.fuse_search_sim3(fa, S)), "cpu_ms": med(lambda: po.fuse_search_sim3(fa, S)), "fused": int((a >= 0).sum())}
    mi = np.zeros(fa.n_kf, np.uint8)
    a = pkg.search_by_projection_sim3(fa, S, mi); o = po.search_by_projection_sim3(fa, S, mi)
    assert a[0] == o[0] and np.array_equal(a[1], o[1])
    out["search_by_projection_sim3_%dx%d" % (fa.n_points, fa.n_kf)] = {
        "gpu_ms": med(lambda: pkg.search_by_projection_sim3(fa, S, mi)), "cpu_ms": med(lambda: po.search_by_projection_sim3(fa, S, mi)),
        "nmatches": a[0]}
    a12, a21, s12, R12, t12, keep3 = t3.sim3_pair(pkg, po, F1, F2, bounds, 7.5, 1)
    a = pkg.search_by_sim3(a12, a21, s12, R12, t12); o = po.search_by_sim3(a12, a21, s12, R12, t12)
    assert a[0] == o[0] and np.array_equal(a[1], o[1])
    out["search_by_sim3_%dx%d" % (a12.n_points, a21.n_points)] = {
        "gpu_ms": med(lambda: pkg.search_by_sim3(a12, a21, s12, R12, t12)), "cpu_ms": med(lambda: po.search_by_sim3(a12, a21, s12, R12, t12)),
        "nfound": a[0]}
    # --- MapPoint::UpdateNormalAndDepth for 2000 map points (about 10 observations each)
    import test_normals as tn
    sn = tn.scene(3, 2000, 20)
    nargs = (sn["pos"], sn["start"], sn["centers"], sn["ref_center"], sn["ref_scale"], sn["last_scale"], sn["normal"], sn["maxd"], sn["mind"])
    assert tn.same(pkg.update_normals_and_depths(*nargs), oracle_lib.port_update_normal_and_depth(*nargs))
    out["update_normal_and_depth_2000_points_%d_observations" % int(sn["start"][-1])] = {
        "gpu_ms": med(lambda: pkg.update_normals_and_depths(*nargs)), "cpu_ms": med(lambda: oracle_lib.port_update_normal_and_depth(*nargs), 10, 1)}
    # --- Frame::ComputeBoW on a vocabulary of ORBvoc.txt's shape (k = 10, L = 6: 1 111 111 nodes, 10^6 words), and the L1 score
    #     of one BowVector against 2000 keyframes (KeyFrameDatabase::DetectLoopCandidates / DetectRelocalizationCandidates)
    import test_vocab as tv
    rng = np.random.default_rng(77)
    nn = (10 ** 7 - 1) // 9
    parent = (np.arange(nn, dtype=np.int64) - 1) // 10
    parent[0] = 0
    leaf = (np.arange(nn) >= (10 ** 6 - 1) // 9).astype(np.uint8)
    vdesc = rng.integers(0, 256, (nn, 32), dtype=np.uint8)
    wts = np.where(leaf > 0, rng.uniform(0.1, 9.0, nn), 0.0)
    voc = pkg.ORBVocabulary(10, 6, parent.astype(np.int32), leaf, vdesc, wts)
    pvoc = oracle_lib.PortVocabulary(10, 6, parent.astype(np.int32), leaf, vdesc, wts)
    for dsc in (dL, np.concatenate([dL, dR, dL[::-1] ^ np.uint8(0x5A), dR[::-1] ^ np.uint8(0xA5)])):
        assert tv.same_bow(voc.transform(dsc, 4), pvoc.transform(dsc, 4))
        out["compute_bow_%d_descriptors_1111111_nodes" % len(dsc)] = {
            "gpu_ms": med(lambda: voc.transform(dsc, 4)), "cpu_ms": med(lambda: pvoc.transform(dsc, 4), 10, 2),
            "words": len(voc.transform(dsc, 4)[0])}
    batch = [np.roll(dL, i, axis=0) ^ np.uint8(i & 0xFF) for i in range(256)]
    bb = voc.transform_batch(batch, 4)
    assert tv.same_bow(bb[17], pvoc.transform(batch[17], 4))
    out["compute_bow_batch_256_frames_x_%d_descriptors" % len(dL)] = {
        "gpu_ms_per_frame": med(lambda: voc.transform_batch(batch, 4), 10, 2) / 256,
        "cpu_ms_per_frame": med(lambda: pvoc.transform(batch[17], 4), 10, 2)}
    q = tv.random_bow_vectors(rng, 10 ** 6, 1, 500, 900)[0]
    kw, kv, ks = tv.csr(tv.random_bow_vectors(rng, 10 ** 6, 2000, 500, 900, share=q[0]))
    a = voc.score_csr(q[0], q[1], kw, kv, ks); o = pvoc.score_csr(q[0], q[1], kw, kv, ks)
    assert np.array_equal(a.view(np.uint64), o.view(np.uint64))
    out["bow_score_1_vs_2000_keyframes"] = {"gpu_ms": med(lambda: voc.score_csr(q[0], q[1], kw, kv, ks)),
                                            "cpu_ms": med(lambda: pvoc.score_csr(q[0], q[1], kw, kv, ks), 10, 2)}
    # --- KeyFrameDatabase: 2000 keyframes resident on the device; DetectRelocalizationCandidates for one frame (device query +
    #     host thresholds / covisibility accumulation) against the restatement (inverted file already built, as in the reference)
    kfs, erased, qq, connected, neighbours, prev, min_score = tv.db_scene(5, 10 ** 6, n_kf=2000, reloc=True)
    dargs, dkeep = oracle_lib.db_args(kfs, erased, qq, connected, neighbours, prev, min_score, True)
    db = pkg.KeyFrameDatabase()
    for kf in kfs:
        db.add(*kf)
    for i in np.nonzero(erased)[0]:
        db.erase(int(i))
    db.reloc_score[:] = list(prev)
    want = oracle_lib.port_db_detect(dargs)
    got = db.detect_relocalization_candidates(qq, lambda k: neighbours[k])
    assert got == want[3].tolist()
    out["keyframe_database_query_2000_keyframes"] = {
        "gpu_ms_device_query": med(lambda: db.query(*qq)),
        "gpu_ms_detect_with_python_host_part": med(lambda: db.detect_relocalization_candidates(qq, lambda k: neighbours[k])),
        "cpu_ms": med(lambda: oracle_lib.port_db_detect(dargs), 10, 2), "candidates": len(got),
        "sharing_words": int(want[0].sum())}
    return out


def main():
    print(json.dumps(run(), indent=1))


if __name__ == "__main__":
    main()
