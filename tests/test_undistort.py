"""Frame::UndistortKeyPoints / ComputeImageBounds (reference src/Frame.cc:404-463): cv::undistortPoints(pts, K, D, noArray(), K).
Parity is pinned: tests/golden/undistort_golden.npz holds cv2 4.13.0 outputs (tests/golden/make_golden.py undistort)."""
import ctypes as C
import os

import numpy as np
import pytest

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "undistort_golden.npz"))
CAMS = list(range(int(GOLD["count"])))


def oracle_undistort(oracle, pts, K, D):
    L = C.CDLL(oracle.PORT_SO)
    L.orbo_undistort_points.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    pts = np.ascontiguousarray(pts, np.float32); K = np.ascontiguousarray(K, np.float32); D = np.ascontiguousarray(D, np.float32)
    out = np.zeros_like(pts)
    L.orbo_undistort_points(pts.ctypes.data, len(pts), K.ctypes.data, D.ctypes.data, len(D), out.ctypes.data)
    return out


@pytest.mark.parametrize("c", CAMS)
def test_oracle_equals_cv2_golden(oracle, c):
    und = oracle_undistort(oracle, GOLD["pts_%d" % c], GOLD["K_%d" % c], GOLD["D_%d" % c])
    assert und.tobytes() == GOLD["und_%d" % c].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("c", CAMS)
def test_device_points_equal_cv2_golden(pkg, c):
    und = pkg.undistort_points(GOLD["pts_%d" % c], GOLD["K_%d" % c], GOLD["D_%d" % c])
    assert und.tobytes() == GOLD["und_%d" % c].tobytes()


@pytest.mark.gpu
def test_device_keypoints_undistorted_in_place_of_the_host_pass(pkg, oracle):
    img = pkg.synth.synth(752, 480, 0x0D157)
    ex = pkg.ORBextractor(1200, 1.2, 8, 20, 7)
    kps, _ = ex(img)
    K, D = GOLD["K_1"], GOLD["D_1"]                          # EuRoC: 4 coefficients
    un = ex.undistort_keypoints(K, D)
    assert len(un) == len(kps)
    want = oracle_undistort(oracle, np.stack([kps["x"], kps["y"]], 1), K, D)
    assert np.stack([un["x"], un["y"]], 1).tobytes() == want.tobytes()
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(un[f], kps[f])
    assert np.abs(un["x"] - kps["x"]).max() > 1.0            # the distortion does move points
    # mDistCoef[0] == 0: mvKeysUn = mvKeys (src/Frame.cc:406-410)
    same = ex.undistort_keypoints(K, np.zeros(4, np.float32))
    assert same.tobytes() == kps.tobytes()
    # a million-point list through the generic entry point: bit-equal to the oracle at size
    rng = np.random.default_rng(5)
    pts = (rng.random((1 << 20, 2)) * np.array([752.0, 480.0])).astype(np.float32)
    assert pkg.undistort_points(pts, GOLD["K_0"], GOLD["D_0"]).tobytes() == oracle_undistort(oracle, pts, GOLD["K_0"], GOLD["D_0"]).tobytes()


@pytest.mark.gpu
def test_extract_and_undistort_in_one_call_equals_the_two_calls(pkg, oracle):
    """orbgpu_extract_undistorted: ExtractORB + UndistortKeyPoints of the Frame constructor behind one synchronisation"""
    ex = pkg.ORBextractor(1200, 1.2, 8, 20, 7); two = pkg.ORBextractor(1200, 1.2, 8, 20, 7)
    for seed, (w, h), c in [(0x0D157, (752, 480), 1), (0x0D158, (640, 480), 0), (0x0D159, (752, 480), 1), (0x0D15A, (203, 149), 0)]:
        img = pkg.synth.synth(w, h, seed)
        K, D = GOLD["K_%d" % c], GOLD["D_%d" % c]
        kps, un, desc = ex.extract_undistorted(img, K, D)
        k2, d2 = two(img)
        assert kps.tobytes() == k2.tobytes() and desc.tobytes() == d2.tobytes() and len(kps) > 50
        assert un.tobytes() == two.undistort_keypoints(K, D).tobytes()
        want = oracle_undistort(oracle, np.stack([kps["x"], kps["y"]], 1), K, D)
        assert np.stack([un["x"], un["y"]], 1).tobytes() == want.tobytes()
    # the frame stays the handle's most recent frame; a zero first coefficient copies the keypoints; an empty image is a no-op
    assert ex.pyramid_level(1).shape == two.pyramid_level(1).shape
    kps, un, _ = ex.extract_undistorted(img, K, np.zeros(4, np.float32))
    assert un.tobytes() == kps.tobytes()
    assert ex.extract_undistorted(np.zeros((0, 0), np.uint8), K, D) == (None, None, None)
    with pytest.raises(pkg.OrbGpuError):
        ex.extract_undistorted(img, K, np.zeros(3, np.float32))


@pytest.mark.gpu
def test_undistort_argument_errors(pkg):
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7)
    with pytest.raises(RuntimeError):
        ex.undistort_keypoints(GOLD["K_0"], GOLD["D_0"])    # nothing extracted yet
    with pytest.raises(RuntimeError):
        pkg.undistort_points(np.zeros((4, 2), np.float32), GOLD["K_0"], np.zeros(3, np.float32))   # ndist < 4
