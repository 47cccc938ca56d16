"""Grid-gated searches between two frames: Frame::GetFeaturesInArea (src/Frame.cc:327-380),
ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520), ORBmatcher::SearchByProjection(Frame, Frame)
(src/ORBmatcher.cc:1328-1470).  The oracle restatement (oracle/frame_port.cc) is pinned bit-for-bit against the reference's
own code in tests/test_ref_frame.py (oracle/_ref/libframeref.so); the CPU tests here check it against the geometry of a
synthetic two-view scene and a brute-force window search, the GPU tests check the device path bit-for-bit against it."""
import numpy as np
import pytest

SHIFT = (12, -7)              # frame 2 = frame 1's scene moved by this many pixels


def two_frames(pkg, extract, w=640, h=480, seed=0x2F2A3E, nf=1000):
    """Two views of one synthetic scene: the second is the first shifted by SHIFT, with fresh pixel noise."""
    pad = 32
    scene = pkg.synth.synth(w + 2 * pad, h + 2 * pad, seed)
    f1 = np.ascontiguousarray(scene[pad:pad + h, pad:pad + w])
    f2 = scene[pad - SHIFT[1]:pad - SHIFT[1] + h, pad - SHIFT[0]:pad - SHIFT[0] + w].astype(np.int32)
    nz = (pkg.synth.splitmix64(seed ^ 0xABCD, 0, w * h) % np.uint64(7)).astype(np.int32).reshape(h, w) - 3
    f2 = np.clip(f2 + nz, 0, 255).astype(np.uint8)
    k1, d1 = extract(f1, nf)
    k2, d2 = extract(f2, nf)
    return (k1, d1), (k2, d2), (0.0, float(w), 0.0, float(h))


def oracle_extract(oracle):
    def run(img, nf):
        return oracle.OraclePort(nf, 1.2, 8, 20, 7).extract(img)
    return run


def device_extract(pkg):
    def run(img, nf):
        return pkg.ORBextractor(nf, 1.2, 8, 20, 7)(img)
    return run


def projection_scene(pkg, oracle, F1, F2, bounds, motion, mono, seed=7):
    """Last frame = F1 with map points at depth ~Z; current frame = F2 seen after a camera translation that moves the
    projections by about SHIFT (plus `motion` along z for the forward / backward branches)."""
    (k1, d1), (k2, d2) = F1, F2
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = 517.3, 516.5, 318.6, 255.3
    bf, b = 40.0, 40.0 / fx
    Z = 25.0
    n1, n2 = len(k1), len(k2)
    z = (Z + rng.uniform(-0.4, 0.4, n1)).astype(np.float32)
    Tl = np.eye(4, dtype=np.float32)
    Tl[:3, 3] = [0.3, -0.2, 0.1]
    ang = 0.002
    Tl[:3, :3] = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], np.float32)
    # points in the last camera, then in the world
    pc = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], 1).astype(np.float64)
    Rl = Tl[:3, :3].astype(np.float64); tl = Tl[:3, 3].astype(np.float64)
    pw = ((pc - tl) @ Rl).astype(np.float32)                  # Rl^T (pc - tl)
    Tc = Tl.copy()
    Tc[:3, 3] += np.array([SHIFT[0] * Z / fx, SHIFT[1] * Z / fy, motion], np.float32)
    s4 = oracle.OraclePort(1000, 1.2, 8, 20, 7).tables()[0]
    valid = (rng.random(n1) < 0.85).astype(np.uint8)
    observed = (rng.random(n1) < 0.7).astype(np.uint8)
    occupied = (rng.random(n2) < 0.05).astype(np.uint8)
    ur = np.where(rng.random(n2) < 0.6, k2["x"] - bf / Z + rng.normal(0, 3.0, n2), -1.0).astype(np.float32)
    if mono:
        ur[:] = -1.0
    args, keep = pkg.projection_args(pw, d1, valid, observed, k1["octave"], k1["angle"], Tl, k2, d2, ur, occupied, Tc,
                                     (fx, fy, cx, cy), bf, b, bounds, s4[0], 15.0 if mono else 7.0, mono)
    return args, keep


# ---------------------------------------------------------------------------------------------------------- CPU: oracle
def test_oracle_features_in_area_equals_window_scan_as_a_set(pkg, oracle):
    (k1, d1), _, bounds = two_frames(pkg, oracle_extract(oracle))
    po = oracle.OraclePort()
    rng = np.random.default_rng(1)
    for _ in range(60):
        x, y = float(rng.uniform(-20, 660)), float(rng.uniform(-20, 500))
        r = float(rng.choice([7.0, 15.0, 43.2, 100.0]))
        lo, hi = [(-1, -1), (0, 0), (1, 3), (2, -1), (0, 4)][int(rng.integers(5))]
        got = po.features_in_area(k1, bounds, x, y, r, lo, hi)
        inside = (np.abs(k1["x"] - np.float32(x)) < r) & (np.abs(k1["y"] - np.float32(y)) < r)
        if lo > 0 or hi >= 0:
            inside &= k1["octave"] >= lo
            if hi >= 0:
                inside &= k1["octave"] <= hi
        # the grid only bounds the scan; every keypoint of the open window is returned (a window may not reach
        # past the cells it touches, so the result is exactly the window content)
        assert sorted(got.tolist()) == np.nonzero(inside)[0].tolist()
        assert len(set(got.tolist())) == len(got)


def test_oracle_initialization_matches_follow_the_shift(pkg, oracle):
    (k1, d1), (k2, d2), bounds = two_frames(pkg, oracle_extract(oracle))
    po = oracle.OraclePort()
    pm0 = np.stack([k1["x"], k1["y"]], 1)
    nm, m12, pm = po.search_for_initialization(k1, d1, k2, d2, bounds, pm0)
    ok = m12 >= 0
    assert nm == ok.sum() and nm > 60
    assert (k1["octave"][ok] == 0).all() and (k2["octave"][m12[ok]] == 0).all()
    assert len(set(m12[ok].tolist())) == ok.sum()             # vnMatches21 keeps matches one-to-one
    dx = k2["x"][m12[ok]] - k1["x"][ok]; dy = k2["y"][m12[ok]] - k1["y"][ok]
    assert (np.abs(dx - SHIFT[0]) < 2).mean() > 0.9 and (np.abs(dy - SHIFT[1]) < 2).mean() > 0.9
    assert np.array_equal(pm[ok], np.stack([k2["x"][m12[ok]], k2["y"][m12[ok]]], 1)) and np.array_equal(pm[~ok], pm0[~ok])


def test_oracle_projection_matches_follow_the_motion(pkg, oracle):
    F1, F2, bounds = two_frames(pkg, oracle_extract(oracle))
    po = oracle.OraclePort()
    for motion, mono in [(0.0, True), (0.0, False), (0.5, False), (-0.5, False)]:
        args, keep = projection_scene(pkg, oracle, F1, F2, bounds, motion, mono)
        nm, cm = po.search_by_projection(args)
        got = cm >= 0
        assert got.sum() > 150 and nm >= got.sum() - 5
        assert not (got & (keep["cur_occupied"] > 0)).any()   # occupied features are never taken
        assert keep["mp_valid"][cm[got]].all()
        k1, k2 = F1[0], F2[0]
        dx = k2["x"][got] - k1["x"][cm[got]]; dy = k2["y"][got] - k1["y"][cm[got]]
        assert (np.abs(dx - SHIFT[0]) < 3).mean() > 0.85 and (np.abs(dy - SHIFT[1]) < 3).mean() > 0.85


# ------------------------------------------------------------------------------------------------- GPU: device == oracle
@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(640, 480, 1000, 100, 0.9, True), (640, 480, 2000, 100, 0.9, True), (752, 480, 1000, 30, 0.6, False),
                                 (333, 217, 700, 100, 0.9, True)])
def test_device_search_for_initialization_equals_oracle(pkg, oracle, cfg):
    w, h, nf, win, ratio, ori = cfg
    (k1, d1), (k2, d2), bounds = two_frames(pkg, device_extract(pkg), w, h, 0x2F2A3E + nf, nf)
    po = oracle.OraclePort()
    pm = np.stack([k1["x"], k1["y"]], 1)
    for _ in range(2):                                        # second round starts from the updated vbPrevMatched
        onm, om12, opm = po.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, ori)
        nm, m12, npm = pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, ori)
        assert nm == onm and np.array_equal(m12, om12) and npm.tobytes() == opm.tobytes()
        pm = npm
    assert nm > 30


@pytest.mark.gpu
def test_device_search_for_initialization_with_duplicates_steals_like_the_reference(pkg, oracle):
    """Many identical descriptors in F1 competing for the same F2 feature: exercises vMatchedDistance / vnMatches21."""
    (k1, d1), (k2, d2), bounds = two_frames(pkg, device_extract(pkg))
    rng = np.random.default_rng(9)
    lvl0 = np.nonzero(k1["octave"] == 0)[0]
    d1 = d1.copy(); k1 = k1.copy()
    for i in lvl0[::3]:                                        # every third level-0 keypoint becomes a near copy of a neighbour
        j = lvl0[int(rng.integers(len(lvl0)))]
        d1[i] = d1[j]; d1[i, int(rng.integers(32))] ^= np.uint8(1 << int(rng.integers(8)))
        k1["x"][i] = k1["x"][j] + np.float32(rng.integers(-3, 4)); k1["y"][i] = k1["y"][j]
    pm = np.stack([k1["x"], k1["y"]], 1)
    po = oracle.OraclePort()
    for ori in (True, False):
        onm, om12, opm = po.search_for_initialization(k1, d1, k2, d2, bounds, pm, 100, 0.9, ori)
        nm, m12, npm = pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm, 100, 0.9, ori)
        assert nm == onm and np.array_equal(m12, om12) and npm.tobytes() == opm.tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("case", [(0.0, True), (0.0, False), (0.5, False), (-0.5, False)], ids=["mono", "stereo", "forward", "backward"])
def test_device_search_by_projection_equals_oracle(pkg, oracle, case):
    motion, mono = case
    F1, F2, bounds = two_frames(pkg, device_extract(pkg))
    po = oracle.OraclePort()
    for seed in (7, 8):
        args, keep = projection_scene(pkg, oracle, F1, F2, bounds, motion, mono, seed)
        onm, ocm = po.search_by_projection(args)
        nm, cm = pkg.search_by_projection(args)
        assert nm == onm and np.array_equal(cm, ocm)
        args.check_orientation = 0
        onm, ocm = po.search_by_projection(args)
        nm, cm = pkg.search_by_projection(args)
        assert nm == onm and np.array_equal(cm, ocm)
    assert (cm >= 0).sum() > 150


@pytest.mark.gpu
def test_device_search_edge_cases(pkg, oracle):
    (k1, d1), (k2, d2), bounds = two_frames(pkg, device_extract(pkg), 333, 217, 5, 500)
    pm = np.stack([k1["x"], k1["y"]], 1)
    nm, m12, _ = pkg.search_for_initialization(k1, d1, k2[:0], d2[:0], bounds, pm)          # empty F2
    assert nm == 0 and (m12 == -1).all()
    nm, m12, _ = pkg.search_for_initialization(k1[:0], d1[:0], k2, d2, bounds, pm[:0])      # empty F1
    assert nm == 0 and len(m12) == 0
    far = pm + np.float32(5000)                                                            # windows outside the grid
    nm, m12, _ = pkg.search_for_initialization(k1, d1, k2, d2, bounds, far)
    onm, om12, _ = oracle.OraclePort().search_for_initialization(k1, d1, k2, d2, bounds, far)
    assert nm == onm == 0 and np.array_equal(m12, om12)
    with pytest.raises(RuntimeError):
        pkg.search_for_initialization(k1, d1, k2, d2, (0, 0, 0, 217), pm)                  # degenerate bounds


# ------------------------------------------------------------------------------------------------ SearchByBoW (f3)
def feature_vector(desc, n_words):
    """A stand-in for DBoW2's FeatureVector (the vocabulary tree stays on the reference path): node id = a few leading
    descriptor bits + a coarse hash, features listed per node in index order, nodes ascending -- the layout of
    std::map<NodeId, std::vector<unsigned>>."""
    level_bits = 3 if n_words <= 32 else 5
    node = ((desc[:, 0].astype(np.uint32) >> (8 - level_bits)) * 7 + (desc[:, 5].astype(np.uint32) >> 6)) % np.uint32(n_words)
    ids = np.unique(node)
    start, idx = [0], []
    for i in ids:
        members = np.nonzero(node == i)[0]
        idx.extend(members.tolist()); start.append(len(idx))
    return ids.astype(np.uint32), np.array(start, np.int32), np.array(idx, np.uint32)


def bow_sides(pkg, F1, F2, seed, kf_kf, n_words=24):
    (k1, d1), (k2, d2) = F1, F2
    rng = np.random.default_rng(seed)
    v1 = (rng.random(len(k1)) < 0.8).astype(np.uint8)
    v2 = (rng.random(len(k2)) < 0.8).astype(np.uint8) if kf_kf else np.ones(len(k2), np.uint8)
    s1, keep1 = pkg.bow_side(d1, v1, k1["angle"], *feature_vector(d1, n_words))
    s2, keep2 = pkg.bow_side(d2, v2, k2["angle"], *feature_vector(d2, n_words))
    return s1, s2, (keep1, keep2)


def test_oracle_bow_matches_are_one_to_one_and_follow_the_shift(pkg, oracle):
    F1, F2, _ = two_frames(pkg, oracle_extract(oracle))
    po = oracle.OraclePort()
    for kf_kf in (False, True):
        s1, s2, keep = bow_sides(pkg, F1, F2, 3, kf_kf)
        nm, m12, m21 = po.search_by_bow(s1, s2, 0.7, True, kf_kf)
        ok = m12 >= 0
        assert nm == ok.sum() == (m21 >= 0).sum() and nm > 100
        assert np.array_equal(m21[m12[ok]], np.nonzero(ok)[0])             # the two views of the pair list agree
        assert keep[0]["valid"][ok].all() and keep[1]["valid"][m12[ok]].all()
        k1, k2 = F1[0], F2[0]
        dx = k2["x"][m12[ok]] / 1.0 - k1["x"][ok]; dy = k2["y"][m12[ok]] - k1["y"][ok]
        assert (np.abs(dx - SHIFT[0]) < 4).mean() > 0.8 and (np.abs(dy - SHIFT[1]) < 4).mean() > 0.8


@pytest.mark.gpu
@pytest.mark.parametrize("kf_kf", [False, True], ids=["keyframe-frame", "keyframe-keyframe"])
def test_device_search_by_bow_equals_oracle(pkg, oracle, kf_kf):
    po = oracle.OraclePort()
    for nf, seed in [(1000, 3), (2000, 4)]:
        F1, F2, _ = two_frames(pkg, device_extract(pkg), 640, 480, 0xB0B0 + nf, nf)
        s1, s2, keep = bow_sides(pkg, F1, F2, seed, kf_kf, 24 if nf == 1000 else 100)
        for ratio, ori in [(0.7, True), (0.9, False), (0.6, True)]:
            onm, om12, om21 = po.search_by_bow(s1, s2, ratio, ori, kf_kf)
            nm, m12, m21 = pkg.search_by_bow(s1, s2, ratio, ori, kf_kf)
            assert nm == onm and np.array_equal(m12, om12) and np.array_equal(m21, om21)
    assert nm > 80
    # a single shared node (every feature against every feature) and disjoint vocabularies
    (k1, d1), (k2, d2) = F1, F2
    one = (np.array([5], np.uint32), np.array([0, len(k1)], np.int32), np.arange(len(k1), dtype=np.uint32))
    two = (np.array([5], np.uint32), np.array([0, len(k2)], np.int32), np.arange(len(k2), dtype=np.uint32))
    a, ka = pkg.bow_side(d1, np.ones(len(k1), np.uint8), k1["angle"], *one)
    b, kb = pkg.bow_side(d2, np.ones(len(k2), np.uint8), k2["angle"], *two)
    got, want = pkg.search_by_bow(a, b), po.search_by_bow(a, b)
    assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    two = (np.array([9], np.uint32), two[1], two[2])
    b, kb = pkg.bow_side(d2, np.ones(len(k2), np.uint8), k2["angle"], *two)
    nm, m12, m21 = pkg.search_by_bow(a, b)
    assert nm == 0 and (m12 == -1).all() and (m21 == -1).all()


# --------------------------------------------------------------------------------------------- randomised stress
def knife_edge_projection_args(pkg, port):
    """A pair of poses whose tlc.z differs in the last bit between the float evaluation of "-Rcw.t()*tcw" (what
    cv::MatExpr does, src/ORBmatcher.cc:1341) and a double-accumulating one; the baseline b is set to the smaller value so
    that bForward = tlc.z > b holds for one evaluation only.  -> (args, keepalive, (Tcw_cur, Tcw_last, tlc.z in float))"""
    f32 = np.float32
    rng = np.random.default_rng(2024)
    for _ in range(20000):
        q, _ = np.linalg.qr(rng.normal(0, 1, (3, 3)))
        Tc = np.eye(4, dtype=f32); Tc[:3, :3] = q.astype(f32); Tc[:3, 3] = rng.normal(0, 0.5, 3).astype(f32)
        q2, _ = np.linalg.qr(rng.normal(0, 1, (3, 3)))
        Tl = np.eye(4, dtype=f32); Tl[:3, :3] = q2.astype(f32); Tl[:3, 3] = rng.normal(0, 0.5, 3).astype(f32)
        R, t = Tc[:3, :3], Tc[:3, 3]
        twc_f = np.array([-f32(f32(f32(R[0, i] * t[0]) + f32(R[1, i] * t[1])) + f32(R[2, i] * t[2])) for i in range(3)], f32)
        twc_d = (-(R.astype(np.float64).T @ t.astype(np.float64))).astype(f32)

        def tlc_z(twc):
            r = Tl[2, :3]
            s = f32(f32(f32(r[0] * twc[0]) + f32(r[1] * twc[1])) + f32(r[2] * twc[2]))
            return f32(np.float64(s) + np.float64(Tl[2, 3]))
        zf, zd = tlc_z(twc_f), tlc_z(twc_d)
        if zf != zd and min(zf, zd) > 0.05:
            break
    else:
        raise AssertionError("no knife-edge pose found")
    b = float(min(zf, zd))
    rng = np.random.default_rng(3)
    pool = rng.integers(0, 256, (30, 32), dtype=np.uint8)
    n1, n2 = 1200, 1200
    k1, d1 = random_frame(rng, n1, 640, 480, pool); k2, d2 = random_frame(rng, n2, 640, 480, pool)
    fx, fy, cx, cy = 500.0, 505.0, 320.0, 240.0
    z = rng.uniform(2, 12, n1)                    # world points that project into the current image
    pc = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], 1)
    pw = ((pc - Tc[:3, 3].astype(np.float64)) @ Tc[:3, :3].astype(np.float64)).astype(f32)
    s4 = port.tables()[0]
    args, keep = pkg.projection_args(pw, d1, np.ones(n1, bool), np.ones(n1, bool), k1["octave"], k1["angle"], Tl, k2, d2,
                                     -np.ones(n2, f32), np.zeros(n2, bool), Tc, (fx, fy, cx, cy), 40.0, b,
                                     (0.0, 640.0, 0.0, 480.0), s4[0], 15.0, False)
    return args, keep, (Tc, Tl, zf)


@pytest.mark.gpu
def test_device_search_by_projection_forward_flag_on_a_knife_edge(pkg, oracle):
    """The host side of orbgpu_search_by_projection forms tlc with cv::MatExpr's arithmetic (float "-R.t()*t")."""
    po = oracle.OraclePort()
    args, keep, _ = knife_edge_projection_args(pkg, po)
    want = po.search_by_projection(args); got = pkg.search_by_projection(args)
    assert want[0] > 20 and got[0] == want[0] and np.array_equal(got[1], want[1])


def random_frame(rng, n, w, h, pool):
    """Keypoints anywhere in and slightly outside the image bounds (undistortion can push them out, src/Frame.cc:387),
    all octaves, descriptors drawn from a small pool with a few flipped bits (many ties and competing queries)."""
    KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
    k = np.zeros(n, KP)
    k["x"] = rng.uniform(-12, w + 12, n).astype(np.float32); k["y"] = rng.uniform(-12, h + 12, n).astype(np.float32)
    k["octave"] = rng.integers(0, 8, n); k["angle"] = rng.uniform(0, 360, n).astype(np.float32)
    k["size"] = 31; k["response"] = rng.integers(7, 200, n); k["class_id"] = -1
    d = pool[rng.integers(0, len(pool), n)].copy()
    flips = rng.integers(0, 4, n)
    for i in range(n):
        for _ in range(int(flips[i])):
            d[i, int(rng.integers(32))] ^= np.uint8(1 << int(rng.integers(8)))
    return k, d


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [11, 12, 13])
def test_device_searches_equal_oracle_on_random_frames(pkg, oracle, seed):
    rng = np.random.default_rng(seed)
    w, h = 640, 480
    pool = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    n1, n2 = int(rng.integers(300, 1500)), int(rng.integers(300, 1500))
    k1, d1 = random_frame(rng, n1, w, h, pool); k2, d2 = random_frame(rng, n2, w, h, pool)
    k1["octave"][rng.random(n1) < 0.6] = 0                    # SearchForInitialization only uses level 0
    bounds = (-8.5, w + 7.25, -6.0, h + 9.5)                  # undistorted image bounds need not be the pixel grid
    po = oracle.OraclePort()
    pm = np.stack([k1["x"], k1["y"]], 1) + rng.normal(0, 20, (n1, 2)).astype(np.float32)
    for win, ratio, ori in [(100, 0.9, True), (37, 0.75, False)]:
        want = po.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, ori)
        got = pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, ori)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()
    # projection: random world points in front of / behind the camera, random validity, occupancy and stereo coordinates
    s4 = po.tables()[0]
    for mono, motion in [(True, 0.0), (False, 0.3), (False, -0.3), (False, 0.0)]:
        Tl = np.eye(4, dtype=np.float32); Tc = np.eye(4, dtype=np.float32)
        Tc[:3, 3] = [0.05, -0.02, motion]
        z = rng.uniform(-2, 12, n1).astype(np.float32); z[np.abs(z) < 0.2] = 0.5
        fx, fy, cx, cy = 500.0, 505.0, 320.0, 240.0
        pw = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], 1).astype(np.float32)
        ur = np.where(rng.random(n2) < 0.5, k2["x"] - rng.uniform(0, 30, n2), -1.0).astype(np.float32)
        args, keep = pkg.projection_args(pw, d1, rng.random(n1) < 0.8, rng.random(n1) < 0.6, k1["octave"], k1["angle"], Tl, k2, d2, ur,
                                         rng.random(n2) < 0.1, Tc, (fx, fy, cx, cy), 40.0, 0.08, bounds, s4[0], 15.0 if mono else 7.0, mono)
        want = po.search_by_projection(args); got = pkg.search_by_projection(args)
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
    # BoW: random vocabulary words, validity on both sides
    def fv(n, words):
        node = rng.integers(0, words, n)
        ids = np.unique(node); start = [0]; idx = []
        for i in ids:
            idx.extend(np.nonzero(node == i)[0].tolist()); start.append(len(idx))
        return ids.astype(np.uint32), np.array(start, np.int32), np.array(idx, np.uint32)
    for words, strict in [(7, False), (90, True)]:
        s1, keep1 = pkg.bow_side(d1, rng.random(n1) < 0.7, k1["angle"], *fv(n1, words))
        s2, keep2 = pkg.bow_side(d2, rng.random(n2) < 0.7, k2["angle"], *fv(n2, words))
        want = po.search_by_bow(s1, s2, 0.8, True, strict); got = pkg.search_by_bow(s1, s2, 0.8, True, strict)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
