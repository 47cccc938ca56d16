"""GPU: the reference's own Frame / KeyFrame / MapPoint objects driven through ORB_SLAM2::ORBmatcher twice --
  ref  oracle/_ref/libframeref.so   the UNMODIFIED reference src/ORBmatcher.cc
  gpu  oracle/_ref/libframegpu.so   the same tree with src/ORBmatcher.cc replaced by my-slam_b200/host/ORBmatcher_gpu.cc, the
                                    drop-in whose ten Search* / Fuse members keep the reference's signatures
                                    (include/ORBmatcher.h:41-81) and call liborbgpu.so
-- on the same scenarios, through the same harness (oracle/ref_frame_wrap.cc builds the objects with the reference's
constructors and reads the mutated objects back).  One test per routine: reference-lib -> device directly, no restatement in
between.  The scenario builders are the ones of the CPU tests (real two-view scenes + randomised frames)."""
import numpy as np
import pytest

import oracle_lib
from test_fuse import fuse_from_local_map
from test_local_map import local_map_scene, random_local_map_args, same
from test_local_points import local_scene, random_local_args
from test_reloc_projection import reloc_from_local_map
from test_search import bow_sides, oracle_extract, projection_scene, random_frame, two_frames
from test_sim3 import fuse_args_from, sim3_of, sim3_pair
from test_triangulation import random_tri_args, tri_scene

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oracle_lib.have_frame_gpu(), reason="oracle/_ref/libframegpu.so not built")]


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def ref(port):
    s4 = port.tables()[0]
    return oracle_lib.OracleFrameRef((s4[0], s4[1]))


@pytest.fixture(scope="module")
def gpu(port):
    s4 = port.tables()[0]
    return oracle_lib.OracleFrameRef((s4[0], s4[1]), so=oracle_lib.FRAME_GPU_SO)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


def test_patched_tree_descriptor_distance(ref, gpu):
    rng = np.random.default_rng(5)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert gpu.descriptor_distance(a, b) == ref.descriptor_distance(a, b)
    for sizes in ([0] * 30, [5] * 30, list(range(30)), [3, 3, 40, 3, 39, 41] + [0] * 24, [100, 9, 9] + [0] * 27):
        assert gpu.three_maxima(sizes) == ref.three_maxima(sizes)


def test_patched_search_for_initialization(ref, gpu, scene):
    (k1, d1), (k2, d2), bounds = scene
    pm0 = np.stack([k1["x"], k1["y"]], 1)
    total = 0
    for win, ratio, ori in [(100, 0.9, True), (30, 0.6, False), (7, 0.9, True)]:
        want = ref.search_for_initialization(k1, d1, k2, d2, bounds, pm0, win, ratio, ori)
        got = gpu.search_for_initialization(k1, d1, k2, d2, bounds, pm0, win, ratio, ori)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()
        total += want[0]
    assert total > 100
    k2d = np.concatenate([k2, k2[:300]]); d2d = np.concatenate([d2, d2[:300]])          # later queries steal earlier matches
    want = ref.search_for_initialization(k1, d1, k2d, d2d, bounds, pm0, 100, 0.9, True)
    got = gpu.search_for_initialization(k1, d1, k2d, d2d, bounds, pm0, 100, 0.9, True)
    assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()


@pytest.mark.parametrize("case", [(0.0, True), (0.0, False), (0.5, False), (-0.5, False)], ids=["mono", "stereo", "forward", "backward"])
def test_patched_search_by_projection_last_frame(pkg, oracle, ref, gpu, scene, case):
    motion, mono = case
    F1, F2, bounds = scene
    for seed, ori in [(7, True), (8, False)]:
        args, keep = projection_scene(pkg, oracle, F1, F2, bounds, motion, mono, seed)
        args.check_orientation = int(ori)
        wn, wm = ref.search_by_projection(args)
        gn, gm = gpu.search_by_projection(args)
        assert gn == wn and wn > 100 and np.array_equal(gm, wm)


@pytest.mark.parametrize("kf_kf", [False, True], ids=["keyframe-frame", "keyframe-keyframe"])
def test_patched_search_by_bow(pkg, ref, gpu, scene, kf_kf):
    F1, F2, _ = scene
    total = 0
    for seed, words in [(3, 24), (4, 100), (5, 1)]:
        s1, s2, keep = bow_sides(pkg, F1, F2, seed, kf_kf, words)
        for ratio, ori in [(0.7, True), (0.9, False)]:
            want = ref.search_by_bow(s1, s2, ratio, ori, kf_kf)
            got = gpu.search_by_bow(s1, s2, ratio, ori, kf_kf)
            assert got[0] == want[0] and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
            total += want[0]
    assert total > 300


def test_patched_search_by_bow_on_the_same_keyframes_again(pkg, ref, gpu, scene):
    """The patched matcher keeps keyframes resident in HBM: the same two KeyFrame objects searched before pKF2 has a
    FeatureVector, after ComputeBoW filled it in (the resident copy must be refreshed) and after map points were erased from
    pKF1 (only the flags change) -- each time what the unmodified reference returns."""
    F1, F2, _ = scene
    for seed, words in [(3, 24), (4, 100)]:
        s1, s2, keep = bow_sides(pkg, F1, F2, seed, True, words)
        want = ref.search_by_bow_thrice(s1, s2, 0.75, True)
        got = gpu.search_by_bow_thrice(s1, s2, 0.75, True)
        assert want[0][0] == 0 and want[1][0] > 100 and 0 < want[2][0] < want[1][0]
        for g, w in zip(got, want):
            assert g[0] == w[0] and np.array_equal(g[1], w[1])


@pytest.mark.parametrize("th", [1.0, 3.0])
def test_patched_search_local_points_and_local_map(pkg, port, ref, gpu, scene, th):
    F1, F2, bounds = scene
    for seed, ratio, stereo in [(1, 0.8, True), (2, 0.6, False)]:
        args, keep = local_scene(pkg, port, F1, F2, bounds, th, seed, ratio, stereo)
        wn, wm = ref.search_local_points(args)
        gn, gm = gpu.search_local_points(args)
        assert wn > 100 and gn == wn and np.array_equal(gm, wm)
        args, keep = local_map_scene(pkg, port, F1, F2, bounds, th, seed, stereo)       # Frame::isInFrustum (reference) + the search
        want = ref.search_local_map(args, pkg.track_fields)
        got = gpu.search_local_map(args, pkg.track_fields)
        assert want[2] > 300 and same(got, want)
    args, keep = random_local_args(pkg, port, 21 + int(th))
    wn, wm = ref.search_local_points(args)
    gn, gm = gpu.search_local_points(args)
    assert gn == wn and np.array_equal(gm, wm)


@pytest.mark.parametrize("cfg", [(3, 24, False, True, 0.0), (4, 100, False, False, 0.5), (5, 1, True, True, -0.4)], ids=["sideways", "forward", "stereo-backward"])
def test_patched_search_for_triangulation(pkg, port, ref, gpu, scene, cfg):
    seed, words, only_stereo, ori, mz = cfg
    F1, F2, _ = scene
    args, keep, T1 = tri_scene(pkg, port, F1, F2, seed, words, only_stereo, ori, mz)
    wn, wm, cw = ref.search_for_triangulation(args, T1)
    gn, gm, cg = gpu.search_for_triangulation(args, T1)
    assert wn > 15 and gn == wn and np.array_equal(gm, wm) and np.array_equal(cg, cw)
    args, keep, T1 = random_tri_args(pkg, port, 31 + seed)
    wn, wm, cw = ref.search_for_triangulation(args, T1)
    gn, gm, cg = gpu.search_for_triangulation(args, T1)
    assert gn == wn and np.array_equal(gm, wm)


@pytest.mark.parametrize("cfg", [(3.0, 1.0, (0.0, 640.0, 0.0, 480.0)), (3.0, 50.0, (-8.5, 647.25, -6.75, 489.5))], ids=["th3", "wide-gate-bounds-out"])
def test_patched_fuse(pkg, port, ref, gpu, scene, cfg):
    th, sig, bounds = cfg
    F1, F2, _ = scene
    for seed, stereo in [(1, True), (2, False)]:
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed, stereo)
        args, fkeep = fuse_from_local_map(pkg, port, keep, th, bounds, sig)
        want, ow = ref.fuse_search(args)
        got, og = gpu.fuse_search(args)
        assert (want >= 0).sum() > 200 and np.array_equal(got, want) and np.array_equal(og, ow)
    _, keep = random_local_map_args(pkg, port, 51)
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, (-8.5, 647.25, -6.0, 489.5), 30.0, seed=51)
    want, ow = ref.fuse_search(args)
    got, og = gpu.fuse_search(args)
    assert (want >= 0).sum() > 20 and np.array_equal(got, want)


@pytest.mark.parametrize("cfg", [(10.0, 100, True), (3.0, 64, False)], ids=["th10-ori", "th3-64"])
def test_patched_search_by_projection_keyframe(pkg, port, ref, gpu, scene, cfg):
    th, orb_dist, ori = cfg
    F1, F2, bounds = scene
    for seed in (1, 2):
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, rkeep = reloc_from_local_map(pkg, port, keep, F1[0]["angle"], th, orb_dist, bounds, ori, seed)
        wn, wm = ref.search_by_projection_keyframe(args)
        gn, gm = gpu.search_by_projection_keyframe(args)
        assert wn > 150 and gn == wn and np.array_equal(gm, wm)


@pytest.mark.parametrize("cfg", [(1.0, 4.0, (0.0, 640.0, 0.0, 480.0)), (0.83, 4.0, (-8.5, 647.25, -6.75, 489.5)), (1.37, 10.0, (3.9, 636.2, 2.6, 477.9))],
                         ids=["s1", "s0.83", "s1.37"])
def test_patched_sim3_routines(pkg, port, ref, gpu, scene, cfg):
    sc, th, bounds = cfg
    F1, F2, _ = scene
    for seed in (1, 2):
        rng = np.random.default_rng(seed)
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, fkeep = fuse_args_from(pkg, port, keep, th, bounds, seed)
        S = sim3_of(keep["Tcw"], sc)
        want = ref.fuse_search_sim3(args, S)
        got = gpu.fuse_search_sim3(args, S)
        assert (want >= 0).sum() > 100 and np.array_equal(got, want)
        args, fkeep = fuse_args_from(pkg, port, keep, int(th), bounds, seed)
        mi = rng.random(args.n_kf) < 0.15
        wn, wm = ref.search_by_projection_sim3(args, S, mi)
        gn, gm = gpu.search_by_projection_sim3(args, S, mi)
        assert wn > 100 and gn == wn and np.array_equal(gm, wm)
    total = 0
    for seed, s12, jit in [(1, 1.0, 0.0), (2, 1.0, 0.02), (1, 1.05, 0.0)]:
        a12, a21, s, R12, t12, keep = sim3_pair(pkg, port, F1, F2, bounds, 7.5, seed, s12, jit)
        wn, wm = ref.search_by_sim3(a12, a21, s, R12, t12)
        gn, gm = gpu.search_by_sim3(a12, a21, s, R12, t12)
        assert gn == wn and np.array_equal(gm, wm)
        total += wn
    assert total > 100


# ---- the Frame-constructor rows, reference library -> device directly --------------------------------------------------------
@pytest.mark.parametrize("shape", [(640, 480, 1000, (6, 14, 27)), (1241, 376, 2000, (3, 40))])
def test_device_stereo_matches_equal_reference_frame_code(pkg, oracle, ref, shape):
    """f1: Frame::ComputeStereoMatches (src/Frame.cc:466-640) executed by the reference's own Frame.cc vs orbgpu_stereo_match."""
    from test_stereo import _oracle_pair
    w, h, nf, disp = shape
    left, right, po, args = _oracle_pair(pkg, oracle, w, h, 0x57E2E002 + w, nf, disp)
    exL = pkg.ORBextractor(nf, 1.2, 8, 20, 7); exR = pkg.ORBextractor(nf, 1.2, 8, 20, 7)
    exL(left); exR(right)
    for bf, b in [(386.1448, 0.5372), (10.0, 0.8)]:
        wu, wd = ref.stereo_matches(*args, bf, b)
        u, d = pkg.stereo_match(exL, exR, bf, b)
        assert (wu >= 0).sum() > 20 and u.tobytes() == wu.tobytes() and d.tobytes() == wd.tobytes()


def test_device_undistort_keypoints_equal_reference_frame_code(pkg, ref):
    """f4: Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:404-463) executed by the reference's own Frame.cc on the
    keypoints the device extracted, vs orbgpu_undistort_keypoints (device-resident keypoints) and orbgpu_undistort_points."""
    img = pkg.synth.synth(640, 480, 0x5EED0000 + 77)
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    k, _ = ex(img)
    K4 = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)          # TUM1.yaml
    dist = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
    want, b4 = ref.undistort_keypoints(k, K4, dist, 640, 480)
    got = ex.undistort_keypoints(K4, dist)
    assert len(got) == len(k) > 900 and got.tobytes() == want.tobytes()
    corners = np.array([[0, 0], [640, 0], [0, 480], [640, 480]], np.float32)
    co = pkg.undistort_points(corners, K4, dist)
    mine = np.array([min(co[0, 0], co[2, 0]), max(co[1, 0], co[3, 0]), min(co[0, 1], co[1, 1]), max(co[2, 1], co[3, 1])], np.float32)
    assert mine.tobytes() == b4.tobytes()


def test_device_extractor_equals_unmodified_reference_extractor(pkg):
    """E7: ORBextractor::operator() executed by the UNMODIFIED reference src/ORBextractor.cc (oracle/_ref/liborbref.so) vs the
    device, on frames that are not among the golden cases."""
    if not oracle_lib.have_ref():
        pytest.skip("oracle/_ref/liborbref.so not built")
    for (w, h, nf, seed) in ((640, 480, 1000, 0xABC001), (752, 480, 1000, 0xABC002), (1241, 376, 2000, 0xABC003)):
        img = pkg.synth.synth(w, h, seed)
        rk, rd = oracle_lib.OracleRef(nf, 1.2, 8, 20, 7).extract(img)
        k, d = pkg.ORBextractor(nf, 1.2, 8, 20, 7)(img)
        assert len(k) == len(rk) > 0.9 * nf and k.tobytes() == rk.tobytes() and np.array_equal(d, rd)
