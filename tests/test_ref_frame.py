"""Pins of the SURVEY.md section-8(f) rows against the UNMODIFIED reference code.

oracle/_ref/libframeref.so holds the reference's own src/Frame.cc, src/ORBmatcher.cc, src/KeyFrame.cc, src/MapPoint.cc
(compiled unchanged against oracle/cvshim2, see oracle/Makefile and oracle/ref_frame_wrap.cc).  These CPU tests run the
same inputs through it and through the restatements the GPU tests compare the device path with
(oracle/frame_port.cc, oracle/orb_port.cc:orbport_stereo_matches): restatement == reference, bit for bit.  The library
is built in the development container (where /root/reference exists) and travels prebuilt; without it the tests skip."""
import numpy as np
import pytest

import oracle_lib
from test_search import SHIFT, bow_sides, oracle_extract, projection_scene, random_frame, two_frames

pytestmark = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")

GEMM_1_T = 1


@pytest.fixture(scope="module")
def ref():
    s4 = oracle_lib.OraclePort(1000, 1.2, 8, 20, 7).tables()[0]
    return oracle_lib.OracleFrameRef((s4[0], s4[1]))


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


# ------------------------------------------------------------------------------------------- cv::gemm / cv::MatExpr pins
def test_shim_gemm_equals_cv2_golden(ref):
    """The shim's cv::gemm (3x3 * 3x1 float: the small-matrix case and the generic double-accumulating case) against
    cv2 4.13.0 outputs (tests/golden/gemm_golden.npz, written by tests/golden/make_golden.py)."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "gemm_golden.npz"))
    A, x, c = g["A"], g["x"], g["c"]
    g1, g2, g3, g4 = g["ax_plus_c"], g["neg_ax"], g["neg_atx"], g["atx_plus_c"]
    for i in range(len(A)):
        assert ref.gemm3(A[i], x[i], c[i], 1.0, 1.0, 0).tobytes() == g1[i].tobytes()
        assert ref.gemm3(A[i], x[i], None, -1.0, 0.0, 0).tobytes() == g2[i].tobytes()
        assert ref.gemm3(A[i], x[i], None, -1.0, 0.0, GEMM_1_T).tobytes() == g3[i].tobytes()
        assert ref.gemm3(A[i], x[i], c[i], 1.0, 1.0, GEMM_1_T).tobytes() == g4[i].tobytes()


def test_neg_rt_t_is_the_small_matrix_gemm_of_the_transposed_copy(ref):
    """-Rcw.t()*tcw (src/ORBmatcher.cc:1341): cv::operator-(MatExpr) materialises the transpose (MatOp::subtract), so the
    product is gemm(R^T copy, t, alpha=-1, flags=0) -- float accumulation -- and not a GEMM_1_T call."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "gemm_golden.npz"))
    rng = np.random.default_rng(5)
    differs = 0
    for i in range(400):
        T = np.eye(4, dtype=np.float32); T[:3, :3] = g["A"][i]; T[:3, 3] = g["x"][i]
        Tl = np.eye(4, dtype=np.float32); Tl[:3, :3] = g["A"][(i + 1) % len(g["A"])]; Tl[:3, 3] = g["c"][i]
        twc, tlc = ref.neg_rt_t(T, Tl)
        # cv2 itself, on the transposed copy: the golden "neg_ax" of A^T
        want = ref.gemm3(np.ascontiguousarray(g["A"][i].T), g["x"][i], None, -1.0, 0.0, 0)
        assert twc.tobytes() == want.tobytes()
        differs += twc.tobytes() != g["neg_atx"][i].tobytes()
        assert tlc.tobytes() == ref.gemm3(Tl[:3, :3], twc, Tl[:3, 3], 1.0, 1.0, 0).tobytes()
    assert differs > 0          # the two evaluations are really different arithmetic


def test_shim_mat_div_equals_cv2_golden(ref):
    """"M / s" (src/ORBmatcher.cc:301-302) = convertTo with scale 1/s: v * (float)(1.0 / s) in float, as cv2 4.13.0 computes it
    (tests/golden/scale_golden.npz); the double product float(double(v) / s) differs in the last bit for most vectors."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "scale_golden.npz"))
    src, sc, dst = g["src"], g["s"], g["dst"]
    other = 0
    for i in range(len(sc)):
        got = ref.mat_div(src[i], sc[i])
        assert got.tobytes() == dst[i].tobytes()
        other += (src[i].astype(np.float64) * (1.0 / np.float64(sc[i]))).astype(np.float32).tobytes() != dst[i].tobytes()
    assert other > 1000


def test_shim_mat_add_div_equals_cv2_golden(ref):
    """"acc + v / s" (src/MapPoint.cc:355) = cv::scaleAdd(v, 1 / s, acc): fmaf(v, (float)(1.0 / s), acc), as cv2 4.13.0
    computes it (tests/golden/scale_add_golden.npz); an unfused multiply-then-add differs for about half of the vectors."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "scale_add_golden.npz"))
    v, acc, s, out = g["v"], g["acc"], g["s"], g["out"]
    other = 0
    for i in range(len(s)):
        got = ref.mat_add_div(acc[i], v[i], s[i])
        assert got.tobytes() == out[i].tobytes()
        other += (v[i] * np.float32(1.0 / s[i]) + acc[i]).astype(np.float32).tobytes() != out[i].tobytes()
    assert other > 1000


# ------------------------------------------------------------------------------------------------ M1 / M3: host pieces
def test_descriptor_distance_and_three_maxima_equal_reference(ref, port):
    """ORBmatcher::DescriptorDistance (the reference's SWAR popcount) and ComputeThreeMaxima executed by the reference,
    against the oracle restatement and the python restatement the host C++ test uses."""
    import random
    from test_host_cpp import three_maxima_oracle
    rng = np.random.default_rng(17)
    d = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    d[1] = d[0]; d[2] = ~d[0]; d[3] = 0; d[4] = 255
    for i in range(len(d) - 1):
        want = ref.descriptor_distance(d[i], d[i + 1])
        assert port.hamming(d[i], d[i + 1]) == want == int(np.unpackbits(d[i] ^ d[i + 1]).sum())
    assert ref.descriptor_distance(d[0], d[1]) == 0 and ref.descriptor_distance(d[0], d[2]) == 256
    rnd = random.Random(7)
    cases = [[0] * 30, [5] * 30, [1, 0, 0, 50] + [0] * 26, [3, 5, 2, 9, 0, 1], [10, 1, 1], [100, 10, 9], [100, 9, 9]]
    for _ in range(200):
        cases.append([rnd.choice([0, 0, 1, 2, 3, 5, 8, 13, 40, 90]) for _ in range(30)])
    for sizes in cases:
        assert ref.three_maxima(sizes) == three_maxima_oracle(sizes), sizes


# --------------------------------------------------------------------------------------------- f2: grid-gated searches
def test_features_in_area_restatement_equals_reference(ref, port, scene):
    (k1, d1), _, bounds = scene
    rng = np.random.default_rng(1)
    frames = [(k1, bounds)]
    pool = rng.integers(0, 256, (8, 32), dtype=np.uint8)
    frames.append((random_frame(rng, 900, 640, 480, pool)[0], (-8.5, 647.25, -6.0, 489.5)))
    for k, b in frames:
        for _ in range(150):
            x, y = float(rng.uniform(-40, 690)), float(rng.uniform(-40, 520))
            r = float(rng.choice([0.5, 7.0, 15.0, 43.2, 100.0, 700.0]))
            lo, hi = [(-1, -1), (0, 0), (1, 3), (2, -1), (0, 4), (7, 7), (-1, 2)][int(rng.integers(7))]
            want = ref.features_in_area(k, b, x, y, r, lo, hi)
            got = port.features_in_area(k, b, x, y, r, lo, hi)
            assert np.array_equal(got, want)                      # same members in the same visiting order


def test_search_for_initialization_restatement_equals_reference(ref, port, scene):
    (k1, d1), (k2, d2), bounds = scene
    pm0 = np.stack([k1["x"], k1["y"]], 1)
    total = 0
    for win, ratio, ori in [(100, 0.9, True), (30, 0.6, False), (100, 0.75, True), (7, 0.9, True)]:
        want = ref.search_for_initialization(k1, d1, k2, d2, bounds, pm0, win, ratio, ori)
        got = port.search_for_initialization(k1, d1, k2, d2, bounds, pm0, win, ratio, ori)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()
        total += want[0]
    assert total > 150
    # duplicated features on the searched side: later queries steal earlier matches (vnMatches21, src/ORBmatcher.cc:463-471)
    k2d = np.concatenate([k2, k2[:300]]); d2d = np.concatenate([d2, d2[:300]])
    want = ref.search_for_initialization(k1, d1, k2d, d2d, bounds, pm0, 100, 0.9, True)
    got = port.search_for_initialization(k1, d1, k2d, d2d, bounds, pm0, 100, 0.9, True)
    assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()


@pytest.mark.parametrize("case", [(0.0, True), (0.0, False), (0.5, False), (-0.5, False)], ids=["mono", "stereo", "forward", "backward"])
def test_search_by_projection_restatement_equals_reference(pkg, oracle, ref, port, scene, case):
    motion, mono = case
    F1, F2, bounds = scene
    for seed, ori in [(7, True), (8, False)]:
        args, keep = projection_scene(pkg, oracle, F1, F2, bounds, motion, mono, seed)
        args.check_orientation = int(ori)
        wn, wm = ref.search_by_projection(args)
        gn, gm = port.search_by_projection(args)
        assert gn == wn and wn > 100
        assert np.array_equal(np.where(gm == -2, -1, gm), wm)


@pytest.mark.parametrize("seed", [11, 12, 13, 14, 15, 16])
def test_searches_on_random_frames_restatement_equals_reference(pkg, ref, port, seed):
    """Keypoints in and slightly outside the image bounds, all octaves, pooled descriptors with many ties, random
    validity / occupancy / stereo coordinates: the order-dependent rules of the reference decide most results."""
    rng = np.random.default_rng(seed)
    w, h = 640, 480
    pool = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    n1, n2 = int(rng.integers(300, 1500)), int(rng.integers(300, 1500))
    k1, d1 = random_frame(rng, n1, w, h, pool); k2, d2 = random_frame(rng, n2, w, h, pool)
    k1["octave"][rng.random(n1) < 0.6] = 0
    bounds = (-8.5, w + 7.25, -6.0, h + 9.5)
    pm = np.stack([k1["x"], k1["y"]], 1) + rng.normal(0, 20, (n1, 2)).astype(np.float32)
    for win, ratio, ori in [(100, 0.9, True), (37, 0.75, False)]:
        want = ref.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, ori)
        got = port.search_for_initialization(k1, d1, k2, d2, bounds, pm, win, ratio, ori)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()
    s4 = port.tables()[0]
    for mono, motion in [(True, 0.0), (False, 0.3), (False, -0.3), (False, 0.0)]:
        Tl = np.eye(4, dtype=np.float32); Tc = np.eye(4, dtype=np.float32)
        ang = float(rng.uniform(-0.05, 0.05))
        Tc[:3, :3] = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], np.float32)
        Tc[:3, 3] = [0.05, -0.02, motion]
        z = rng.uniform(-2, 12, n1).astype(np.float32); z[np.abs(z) < 0.2] = 0.5
        fx, fy, cx, cy = 500.0, 505.0, 320.0, 240.0
        pw = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], 1).astype(np.float32)
        ur = np.where(rng.random(n2) < 0.5, k2["x"] - rng.uniform(0, 30, n2), -1.0).astype(np.float32)
        args, keep = pkg.projection_args(pw, d1, rng.random(n1) < 0.8, rng.random(n1) < 0.6, k1["octave"], k1["angle"], Tl, k2, d2, ur,
                                         rng.random(n2) < 0.1, Tc, (fx, fy, cx, cy), 40.0, 0.08, bounds, s4[0], 15.0 if mono else 7.0, mono)
        wn, wm = ref.search_by_projection(args)
        gn, gm = port.search_by_projection(args)
        assert gn == wn and np.array_equal(np.where(gm == -2, -1, gm), wm)

    def fv(n, words):
        node = rng.integers(0, words, n)
        ids = np.unique(node); start = [0]; idx = []
        for i in ids:
            idx.extend(np.nonzero(node == i)[0].tolist()); start.append(len(idx))
        return ids.astype(np.uint32), np.array(start, np.int32), np.array(idx, np.uint32)
    for words, strict in [(7, False), (90, True), (30, False), (5, True)]:
        v2 = rng.random(n2) < 0.7 if strict else np.ones(n2, bool)        # a Frame has no map-point validity
        s1, keep1 = pkg.bow_side(d1, rng.random(n1) < 0.7, k1["angle"], *fv(n1, words))
        s2, keep2 = pkg.bow_side(d2, v2, k2["angle"], *fv(n2, words))
        for ratio, ori in [(0.8, True), (0.6, False)]:
            want = ref.search_by_bow(s1, s2, ratio, ori, strict)
            got = port.search_by_bow(s1, s2, ratio, ori, strict)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])


# ---------------------------------------------------------------------------------------------- f3: BoW-gated search
@pytest.mark.parametrize("kf_kf", [False, True], ids=["keyframe-frame", "keyframe-keyframe"])
def test_search_by_bow_restatement_equals_reference(pkg, ref, port, scene, kf_kf):
    F1, F2, _ = scene
    total = 0
    for seed, words in [(3, 24), (4, 100), (5, 1)]:
        s1, s2, keep = bow_sides(pkg, F1, F2, seed, kf_kf, words)
        for ratio, ori in [(0.7, True), (0.9, False), (0.6, True)]:
            want = ref.search_by_bow(s1, s2, ratio, ori, kf_kf)
            got = port.search_by_bow(s1, s2, ratio, ori, kf_kf)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
            total += want[0]
    assert total > 500


# ------------------------------------------------------------------------------------------------- f1: stereo matcher
@pytest.mark.parametrize("shape", [(640, 480, 1000, (6, 14, 27)), (1241, 376, 2000, (3, 40)), (333, 217, 700, (0, 9))])
def test_stereo_matches_restatement_equals_reference(pkg, oracle, ref, shape):
    from test_stereo import _oracle_pair
    w, h, nf, disp = shape
    left, right, po, args = _oracle_pair(pkg, oracle, w, h, 0x57E2E002 + w, nf, disp)
    for bf, b in [(386.1448, 0.5372), (10.0, 0.8), (40.0, 0.08)]:
        wu, wd = ref.stereo_matches(*args, bf, b)
        gu, gd = po.stereo_matches(*args, bf, b)
        assert (wu >= 0).sum() > 20
        assert gu.tobytes() == wu.tobytes() and gd.tobytes() == wd.tobytes()


# ------------------------------------------------------------------------------------------------- f4: undistortion
def test_undistort_keypoints_through_the_reference_frame_code(oracle, ref, port):
    """Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:404-463) executed by the reference itself on top of the
    cv2-pinned undistortPoints primitive: the mat.reshape / copy-back plumbing around it."""
    import ctypes as C
    rng = np.random.default_rng(9)
    n = 500
    k = np.zeros(n, oracle_lib.KP_DTYPE)
    k["x"] = rng.uniform(0, 640, n).astype(np.float32); k["y"] = rng.uniform(0, 480, n).astype(np.float32)
    k["octave"] = rng.integers(0, 8, n); k["angle"] = rng.uniform(0, 360, n).astype(np.float32); k["size"] = 31
    K4 = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)          # TUM1.yaml
    dist = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
    ku, b4 = ref.undistort_keypoints(k, K4, dist, 640, 480)
    L = port.L
    L.orbo_undistort_points.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    pts = np.ascontiguousarray(np.stack([k["x"], k["y"]], 1), np.float32); out = np.zeros_like(pts)
    L.orbo_undistort_points(pts.ctypes.data, n, K4.ctypes.data, dist.ctypes.data, 5, out.ctypes.data)
    assert ku["x"].tobytes() == out[:, 0].tobytes() and ku["y"].tobytes() == out[:, 1].tobytes()
    for f in ("size", "angle", "octave", "response", "class_id"):
        assert np.array_equal(ku[f], k[f])
    corners = np.array([[0, 0], [640, 0], [0, 480], [640, 480]], np.float32); co = np.zeros_like(corners)
    L.orbo_undistort_points(corners.ctypes.data, 4, K4.ctypes.data, dist.ctypes.data, 5, co.ctypes.data)
    want = np.array([min(co[0, 0], co[2, 0]), max(co[1, 0], co[3, 0]), min(co[0, 1], co[1, 1]), max(co[2, 1], co[3, 1])], np.float32)
    assert b4.tobytes() == want.tobytes()
    # zero distortion: keypoints pass through, bounds are the image rectangle
    ku0, b0 = ref.undistort_keypoints(k, K4, np.zeros(4, np.float32), 640, 480)
    assert ku0.tobytes() == k.tobytes() and b0.tolist() == [0.0, 640.0, 0.0, 480.0]


def test_search_by_projection_forward_flag_on_a_knife_edge(pkg, ref, port):
    """bForward = tlc.z > mb with tlc from "-Rcw.t()*tcw": the restatement must form it in float like cv::MatExpr does."""
    from test_search import knife_edge_projection_args
    args, keep, (Tc, Tl, tlc_z) = knife_edge_projection_args(pkg, port)
    twc, tlc = ref.neg_rt_t(Tc, Tl)
    assert tlc[2] == tlc_z                                       # the numpy float32 evaluation is the reference's
    wn, wm = ref.search_by_projection(args)
    gn, gm = port.search_by_projection(args)
    assert wn > 20
    assert gn == wn and np.array_equal(np.where(gm == -2, -1, gm), wm)
