"""Resident keyframe store (orbgpu_kfstore_*, orbgpu_kf_*): keyframes uploaded once, SearchByBoW (src/ORBmatcher.cc:159-288,
:522-655) and SearchForTriangulation (:657-820) by id.  Every result is checked bit-for-bit against the oracle restatement
(pinned against the reference's own code in tests/test_ref_frame.py) on the same inputs, and the store's bookkeeping
(id reuse, erase, errors, many keyframes, threads) on its own."""
import threading

import numpy as np
import pytest

import oracle_lib
from test_search import bow_sides, device_extract, feature_vector, random_frame, two_frames
from test_triangulation import CAM, camera_center, compute_f12, poses, random_tri_args, tri_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


def level_tables(port):
    s = port.tables()[0][0]
    return s, (s * s).astype(np.float32)


def add_tri_side(store, arrs, sc, sg, cam=CAM):
    """arrs = the keep-alive list of pkg.triangulation_args: keys, desc, has_mp, u_right, node_ids, node_start, indices"""
    return store.add(arrs[0], arrs[1], arrs[3], arrs[4], arrs[5], arrs[6], sc, sg, cam)


def test_triangulation_by_id_equals_oracle(pkg, port):
    store = pkg.KeyFrameStore()
    total = 0
    for seed in range(31, 43):
        args, keep, _ = random_tri_args(pkg, port, seed)
        sc, sg = keep[2]
        i1 = add_tri_side(store, keep[0], sc, sg); i2 = add_tri_side(store, keep[1], sc, sg)
        job, kj = store.triangulation_job(i1, keep[0][2], i2, keep[1][2], np.ctypeslib.as_array(args.F12), np.ctypeslib.as_array(args.Cw1),
                                          np.ctypeslib.as_array(args.T2w), args.only_stereo, args.check_orientation)
        want = port.search_for_triangulation(args)
        got = store.search_for_triangulation(job)
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
        plain = pkg.search_for_triangulation(args)
        assert plain[0] == want[0] and np.array_equal(plain[1], want[1])
        total += want[0]
    assert total > 200 and len(store) == 24


def test_triangulation_batch_by_id_one_keyframe_against_its_neighbours(pkg, port):
    """CreateNewMapPoints (src/LocalMapping.cc:239-270): the current keyframe against each neighbour, one call."""
    store = pkg.KeyFrameStore()
    sc, sg = level_tables(port)
    F = [two_frames(pkg, device_extract(pkg), 640, 480, 0x7A1A + s, 1000) for s in (1000, 1001)]
    frames = [F[0][0], F[0][1], F[1][0], F[1][1]]
    rng = np.random.default_rng(5)
    sides, ids = [], []
    for k, d in frames:
        ur = np.where(rng.random(len(k)) < 0.4, k["x"] - 3.0, -1.0).astype(np.float32)
        fv = feature_vector(d, 24)
        sides.append((k, d, ur) + fv)
        ids.append(store.add(k, d, ur, fv[0], fv[1], fv[2], sc, sg, CAM))
    jobs, wants, keep_all = [], [], []
    for n, mz in [(1, 0.0), (2, 0.5), (3, -0.4)]:
        T1, T2 = poses(mz)
        has1 = rng.random(len(frames[0][0])) < 0.3; has2 = rng.random(len(frames[n][0])) < 0.3
        F12 = compute_f12(T1, T2, CAM); Cw1 = camera_center(T1)
        s1 = (sides[0][0], sides[0][1], has1, sides[0][2]) + sides[0][3:]
        s2 = (sides[n][0], sides[n][1], has2, sides[n][2]) + sides[n][3:]
        args, keep = pkg.triangulation_args(s1, s2, F12, Cw1, T2, CAM, sc, sg, False, True)
        wants.append(port.search_for_triangulation(args))
        j, kj = store.triangulation_job(ids[0], has1, ids[n], has2, F12, Cw1, T2, False, True)
        jobs.append(j); keep_all.append((keep, kj))
    got = store.search_for_triangulation_batch(jobs)
    assert len(got) == 3
    for g, w in zip(got, wants):
        assert g[0] == w[0] and np.array_equal(g[1], w[1])
    assert sum(w[0] for w in wants) > 40
    assert store.search_for_triangulation_batch([]) == []


@pytest.mark.parametrize("kf_kf", [False, True], ids=["keyframe-frame", "keyframe-keyframe"])
def test_search_by_bow_by_id_equals_oracle(pkg, oracle, port, kf_kf):
    po = oracle.OraclePort()
    store = pkg.KeyFrameStore()
    sc, sg = level_tables(port)
    checked = 0
    for nf, seed in [(1000, 3), (2000, 4)]:
        F1, F2, _ = two_frames(pkg, device_extract(pkg), 640, 480, 0xB0B0 + nf, nf)
        s1, s2, keep = bow_sides(pkg, F1, F2, seed, kf_kf, 24 if nf == 1000 else 100)
        (k1, d1), (k2, d2) = F1, F2
        a1, a2 = keep
        i1 = store.add(k1, d1, None, a1["node_ids"], a1["node_start"], a1["indices"], sc, sg)
        i2 = store.add(k2, d2, None, a2["node_ids"], a2["node_start"], a2["indices"], sc, sg) if kf_kf else -1
        for ratio, ori in [(0.7, True), (0.9, False), (0.6, True)]:
            want = po.search_by_bow(s1, s2, ratio, ori, kf_kf)
            if kf_kf:
                got = store.search_by_bow(i1, a1["valid"], i2, a2["valid"], None, ratio, ori, kf_kf)
            else:
                got = store.search_by_bow(i1, a1["valid"], -1, None, s2, ratio, ori, kf_kf)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
            plain = pkg.search_by_bow(s1, s2, ratio, ori, kf_kf)
            assert plain[0] == want[0] and np.array_equal(plain[1], want[1]) and np.array_equal(plain[2], want[2])
            checked += want[0]
    assert checked > 400


def test_store_bookkeeping_and_errors(pkg, port):
    store = pkg.KeyFrameStore()
    sc, sg = level_tables(port)
    rng = np.random.default_rng(11)
    pool = rng.integers(0, 256, (9, 32), dtype=np.uint8)

    def kf(n, words=12):
        k, d = random_frame(rng, n, 640, 480, pool)
        return (k, d) + feature_vector(d, words)
    A, B, Cc = kf(300), kf(500), kf(40)
    ia = store.add(A[0], A[1], None, A[2], A[3], A[4], sc, sg, CAM)
    ib = store.add(B[0], B[1], None, B[2], B[3], B[4], sc, sg, CAM)
    assert (ia, ib) == (0, 1) and len(store) == 2
    store.erase(ia)
    assert len(store) == 1
    with pytest.raises(pkg.OrbGpuError):
        store.erase(ia)                                            # already gone
    with pytest.raises(pkg.OrbGpuError):
        store.search_by_bow(ia, np.ones(300, np.uint8), ib, np.ones(500, np.uint8))
    ic = store.add(Cc[0], Cc[1], None, Cc[2], Cc[3], Cc[4], sc, sg, CAM)
    assert ic == ia and len(store) == 2                            # the id is reused, with the new keyframe's contents
    po = oracle_lib.OraclePort()
    s1, k1 = pkg.bow_side(Cc[1], np.ones(40, np.uint8), Cc[0]["angle"], Cc[2], Cc[3], Cc[4])
    s2, k2 = pkg.bow_side(B[1], np.ones(500, np.uint8), B[0]["angle"], B[2], B[3], B[4])
    want = po.search_by_bow(s1, s2, 0.9, True, True)
    got = store.search_by_bow(ic, np.ones(40, np.uint8), ib, np.ones(500, np.uint8), None, 0.9, True, True)
    assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    # both forms of side 2 at once, or neither, is an argument error; so is an octave outside the level tables
    with pytest.raises(pkg.OrbGpuError):
        store.search_by_bow(ic, np.ones(40, np.uint8), ib, np.ones(500, np.uint8), s2)
    with pytest.raises(pkg.OrbGpuError):
        store.search_by_bow(ic, np.ones(40, np.uint8))
    bad = B[0].copy(); bad["octave"][7] = 8
    with pytest.raises(pkg.OrbGpuError):
        store.add(bad, B[1], None, B[2], B[3], B[4], sc, sg, CAM)
    with pytest.raises(pkg.OrbGpuError):
        store.add(B[0], B[1], None, B[2], B[3], (B[4] + 500).astype(np.uint32), sc, sg, CAM)     # index past N
    # an empty keyframe and a keyframe without a FeatureVector are legal and match nothing
    e = (np.zeros(0, np.uint32), np.zeros(1, np.int32), np.zeros(0, np.uint32))
    ie = store.add(B[0][:0], B[1][:0], None, e[0], e[1], e[2], sc, sg, CAM)
    inofv = store.add(B[0], B[1], None, e[0], e[1], e[2], sc, sg, CAM)
    assert store.search_by_bow(ie, np.zeros(0, np.uint8), ib, np.ones(500, np.uint8))[0] == 0
    assert store.search_by_bow(inofv, np.ones(500, np.uint8), ib, np.ones(500, np.uint8))[0] == 0
    store.close()


def test_search_by_bow_kernel_paths(pkg, port):
    """The three device paths of SearchByBoW against the oracle: fixed-point rounds (ordinary vocabularies), one warp per
    node with the features read from memory (a node of side 2 larger than the shared-memory slice, few queries), and the
    two-kernel path (one node holding everything); heavy contention: many queries for few features, pooled descriptors."""
    po = oracle_lib.OraclePort()
    store = pkg.KeyFrameStore()
    sc, sg = level_tables(port)
    rng = np.random.default_rng(77)
    pool = rng.integers(0, 256, (6, 32), dtype=np.uint8)

    def near(n):                                                   # descriptors a few bits away from a small pool: many ties
        d = pool[rng.integers(0, len(pool), n)].copy()
        flips = rng.integers(0, 256, (n, 3))
        for i in range(n):
            for b in flips[i]:
                d[i, b >> 3] ^= np.uint8(1 << (b & 7))
        return d

    def one_node(n, node=5):
        return np.array([node], np.uint32), np.array([0, n], np.int32), rng.permutation(n).astype(np.uint32)
    total = 0
    for n1, n2 in [(3, 600), (200, 30), (64, 128), (256, 32), (7, 257), (900, 700)]:
        k1, _ = random_frame(rng, n1, 640, 480, pool); k2, _ = random_frame(rng, n2, 640, 480, pool)
        d1, d2 = near(n1), near(n2)
        f1, f2 = one_node(n1), one_node(n2)
        v1 = (rng.random(n1) < 0.9).astype(np.uint8); v2 = (rng.random(n2) < 0.9).astype(np.uint8)
        s1, keep1 = pkg.bow_side(d1, v1, k1["angle"], *f1); s2, keep2 = pkg.bow_side(d2, v2, k2["angle"], *f2)
        i1 = store.add(k1, d1, None, f1[0], f1[1], f1[2], sc, sg); i2 = store.add(k2, d2, None, f2[0], f2[1], f2[2], sc, sg)
        for ratio, strict in [(0.9, True), (0.7, False), (1.5, True)]:
            want = po.search_by_bow(s1, s2, ratio, True, strict)
            got = store.search_by_bow(i1, v1, i2, v2, None, ratio, True, strict)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2]), (n1, n2, ratio)
            plain = pkg.search_by_bow(s1, s2, ratio, True, strict)
            assert plain[0] == want[0] and np.array_equal(plain[1], want[1])
            total += want[0]
    assert total > 100


def test_many_keyframes_stay_intact_and_threads_can_share_the_store(pkg, port):
    """600 keyframes (two 32 MB slabs), a third erased and re-added in another order; searches from four threads."""
    store = pkg.KeyFrameStore()
    sc, sg = level_tables(port)
    rng = np.random.default_rng(21)
    pool = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    kfs, ids = [], []
    for i in range(600):
        n = int(rng.integers(700, 1400))
        k, d = random_frame(rng, n, 640, 480, pool)
        fv = feature_vector(d, 30)
        kfs.append((k, d) + fv)
        ids.append(store.add(k, d, None, fv[0], fv[1], fv[2], sc, sg, CAM))
    for i in range(0, 600, 3):
        store.erase(ids[i])
    for i in reversed(range(0, 600, 3)):
        k, d, a, b, c = kfs[i]
        ids[i] = store.add(k, d, None, a, b, c, sc, sg, CAM)
    assert len(store) == 600 and len(set(ids)) == 600
    po = oracle_lib.OraclePort()
    pairs = [(int(a), int(b)) for a, b in rng.integers(0, 600, (24, 2))]
    wants = []
    for a, b in pairs:
        s1, k1 = pkg.bow_side(kfs[a][1], np.ones(len(kfs[a][0]), np.uint8), kfs[a][0]["angle"], *kfs[a][2:])
        s2, k2 = pkg.bow_side(kfs[b][1], np.ones(len(kfs[b][0]), np.uint8), kfs[b][0]["angle"], *kfs[b][2:])
        wants.append(po.search_by_bow(s1, s2, 0.75, True, True))
    assert sum(w[0] for w in wants) > 200
    errors = []

    def worker(t):
        try:
            for rep in range(3):
                for (a, b), w in list(zip(pairs, wants))[t::4]:
                    g = store.search_by_bow(ids[a], np.ones(len(kfs[a][0]), np.uint8), ids[b], np.ones(len(kfs[b][0]), np.uint8),
                                            None, 0.75, True, True)
                    if not (g[0] == w[0] and np.array_equal(g[1], w[1]) and np.array_equal(g[2], w[2])):
                        errors.append((a, b))
        except Exception as ex:                                    # noqa: BLE001
            errors.append(repr(ex))
    th = [threading.Thread(target=worker, args=(t,)) for t in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert errors == []
