"""The local-map search ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:44-125), the
matcher call Tracking::SearchLocalPoints makes on every tracked frame (src/Tracking.cc:1150-1200).  The map points arrive
as the plain arrays Frame::isInFrustum left in them (mTrackProjX/Y/XR, mnTrackScaleLevel, mTrackViewCos).
CPU: the restatement (oracle/frame_port.cc:frameport_search_local_points) against the reference's own code
(oracle/_ref/libframeref.so) and against the geometry of a two-view scene; GPU: orbgpu_search_local_points bit-for-bit
against the restatement."""
import numpy as np
import pytest

import oracle_lib
from test_search import SHIFT, device_extract, oracle_extract, random_frame, two_frames

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")


def local_scene(pkg, port, F1, F2, bounds, th, seed, nnratio=0.8, stereo=True, observed_p=0.9):
    """Map points = the features of frame 1 (their descriptors), predicted in frame 2 at the shifted position with a little
    noise, at the octave they were seen in (sometimes one off); frame 2 is the searched frame."""
    (k1, d1), (k2, d2) = F1, F2
    rng = np.random.default_rng(seed)
    n1, n2 = len(k1), len(k2)
    px = (k1["x"] + SHIFT[0] + rng.normal(0, 0.8, n1)).astype(np.float32)
    py = (k1["y"] + SHIFT[1] + rng.normal(0, 0.8, n1)).astype(np.float32)
    lvl = np.clip(k1["octave"] + rng.choice([0, 0, 0, 1, -1], n1), 0, 7).astype(np.int32)
    cos = rng.choice(np.array([0.5, 0.9, 0.998, 0.99800003, 0.9985, 1.0], np.float32), n1)
    disp = 3.5
    pxr = (px - disp + rng.normal(0, 2.0, n1)).astype(np.float32)
    ur = np.where(rng.random(n2) < 0.6, k2["x"] - disp + rng.normal(0, 2.0, n2), -1.0).astype(np.float32)
    if not stereo:
        ur[:] = -1.0
    s4 = port.tables()[0]
    return pkg.local_points_args(rng.random(n1) < 0.85, px, py, pxr, lvl, cos, d1, rng.random(n1) < observed_p, k2, d2, ur,
                                 rng.random(n2) < 0.05, bounds, s4[0], th, nnratio)


def random_local_args(pkg, port, seed):
    """Pooled descriptors (many equal distances), keypoints in and outside the grid, every octave, random flags."""
    rng = np.random.default_rng(seed)
    w, h = 640, 480
    pool = rng.integers(0, 256, (30, 32), dtype=np.uint8)
    n1, n2 = int(rng.integers(200, 2500)), int(rng.integers(300, 1500))
    k1, d1 = random_frame(rng, n1, w, h, pool); k2, d2 = random_frame(rng, n2, w, h, pool)
    # a third of the points aim exactly at a feature of the searched frame
    hit = rng.integers(0, n2, n1)
    aim = rng.random(n1) < 0.35
    px = np.where(aim, k2["x"][hit] + rng.normal(0, 2, n1), k1["x"]).astype(np.float32)
    py = np.where(aim, k2["y"][hit] + rng.normal(0, 2, n1), k1["y"]).astype(np.float32)
    lvl = np.where(aim, np.clip(k2["octave"][hit] + rng.integers(0, 2, n1), 0, 7), k1["octave"]).astype(np.int32)
    d1 = np.where(aim[:, None], d2[hit], d1)                      # ... and carry its descriptor, up to a few bits
    flip = rng.random(n1) < 0.5
    d1[flip, rng.integers(0, 32, flip.sum())] ^= np.uint8(1) << rng.integers(0, 8, flip.sum()).astype(np.uint8)
    cos = rng.uniform(0.99, 1.0, n1).astype(np.float32)
    pxr = (px - rng.uniform(0, 20, n1)).astype(np.float32)
    ur = np.where(rng.random(n2) < 0.5, k2["x"] - rng.uniform(0, 20, n2), -1.0).astype(np.float32)
    bounds = (-8.5, w + 7.25, -6.0, h + 9.5)
    th = float(rng.choice([1.0, 3.0, 5.0, 2.5]))
    ratio = float(rng.choice([0.8, 0.6, 0.95]))
    return pkg.local_points_args(rng.random(n1) < 0.8, px, py, pxr, lvl, cos, d1, rng.random(n1) < 0.7, k2, d2, ur,
                                 rng.random(n2) < 0.1, bounds, port.tables()[0][0], th, ratio)


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


# ------------------------------------------------------------------------------------------------------------- CPU side
def test_oracle_local_matches_follow_the_shift(pkg, port, scene):
    """Geometry: an accepted feature lies inside the search window of its map point, at an allowed octave, nobody is
    assigned twice, and most in-view points of the shifted scene find their feature."""
    F1, F2, bounds = scene
    args, keep = local_scene(pkg, port, F1, F2, bounds, 1.0, 3, stereo=False, observed_p=2.0)
    nm, cm = port.search_local_points(args)
    k2 = F2[0]
    s = port.tables()[0][0]
    got = np.nonzero(cm >= 0)[0]
    assert nm == len(got) > 300
    assert len(set(cm[got].tolist())) == len(got)
    for idx in got:
        i = cm[idx]
        assert keep["in_view"][i]
        r = (2.5 if np.float64(keep["view_cos"][i]) > 0.998 else 4.0) * s[keep["scale_level"][i]]
        assert abs(k2["x"][idx] - keep["proj_x"][i]) < r and abs(k2["y"][idx] - keep["proj_y"][i]) < r
        assert keep["scale_level"][i] - 1 <= k2["octave"][idx] <= keep["scale_level"][i]
        assert not keep["cur_occupied"][idx]
    # a wider window can only be asked for more candidates: th = 5 finds at least 80 % as many (ratio test may drop a few)
    args5, keep5 = local_scene(pkg, port, F1, F2, bounds, 5.0, 3, stereo=False, observed_p=2.0)
    assert port.search_local_points(args5)[0] > 0.8 * nm


@needs_ref
@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
def test_local_points_restatement_equals_reference(pkg, port, scene, th):
    F1, F2, bounds = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    total = 0
    for seed, ratio, stereo in [(1, 0.8, True), (2, 0.6, False), (3, 0.95, True)]:
        args, keep = local_scene(pkg, port, F1, F2, bounds, th, seed, ratio, stereo)
        wn, wm = ref.search_local_points(args)
        gn, gm = port.search_local_points(args)
        assert gn == wn and np.array_equal(gm, wm)
        total += wn
    assert total > 600


@needs_ref
@pytest.mark.parametrize("seed", [21, 22, 23, 24, 25, 26])
def test_local_points_on_random_frames_restatement_equals_reference(pkg, port, seed):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    args, keep = random_local_args(pkg, port, seed)
    wn, wm = ref.search_local_points(args)
    gn, gm = port.search_local_points(args)
    assert wn > 20
    assert gn == wn and np.array_equal(gm, wm)


# ------------------------------------------------------------------------------------------------------------- GPU side
@pytest.mark.gpu
@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
def test_device_local_points_equal_oracle(pkg, oracle, port, th):
    F1, F2, bounds = two_frames(pkg, device_extract(pkg))
    for seed, ratio, stereo in [(1, 0.8, True), (2, 0.6, False), (3, 0.95, True)]:
        args, keep = local_scene(pkg, port, F1, F2, bounds, th, seed, ratio, stereo)
        want = port.search_local_points(args)
        got = pkg.search_local_points(args)
        assert want[0] > 200
        assert got[0] == want[0] and np.array_equal(got[1], want[1])


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [21, 22, 23, 24, 25, 26, 27, 28])
def test_device_local_points_equal_oracle_on_random_frames(pkg, port, seed):
    args, keep = random_local_args(pkg, port, seed)
    want = port.search_local_points(args)
    got = pkg.search_local_points(args)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])


@pytest.mark.gpu
def test_device_local_points_edge_cases(pkg, port):
    rng = np.random.default_rng(5)
    pool = rng.integers(0, 256, (4, 32), dtype=np.uint8)
    k2, d2 = random_frame(rng, 400, 640, 480, pool)
    k1, d1 = random_frame(rng, 300, 640, 480, pool)
    s = port.tables()[0][0]
    bounds = (0.0, 640.0, 0.0, 480.0)

    def build(in_view, lvl, n1=300, n2=400, occupied=None, observed=None):
        return pkg.local_points_args(in_view[:n1], k1["x"][:n1], k1["y"][:n1], k1["x"][:n1] - 5, lvl[:n1], np.full(n1, 0.9, np.float32),
                                     d1[:n1], np.ones(n1, np.uint8) if observed is None else observed, k2[:n2], d2[:n2],
                                     np.full(n2, -1.0, np.float32), np.zeros(n2, np.uint8) if occupied is None else occupied,
                                     bounds, s, 5.0, 0.8)
    ones = np.ones(300, np.uint8); lvl = k1["octave"].astype(np.int32)
    # no points / no features / nobody in view
    for a, keep in [build(ones, lvl, n1=0), build(ones, lvl, n2=0), build(np.zeros(300, np.uint8), lvl)]:
        nm, cm = pkg.search_local_points(a)
        assert nm == 0 and (cm == -1).all()
        assert port.search_local_points(a)[0] == 0
    # every feature already holds an observed map point: nothing can be assigned
    a, keep = build(ones, lvl, occupied=np.ones(400, np.uint8))
    assert pkg.search_local_points(a)[0] == 0 == port.search_local_points(a)[0]
    # unobserved map points do not block the feature they were given: a later point may take it again
    a, keep = build(ones, lvl, observed=np.zeros(300, np.uint8))
    want = port.search_local_points(a); got = pkg.search_local_points(a)
    assert want[0] > (want[1] >= 0).sum() > 0                 # more acceptances than features holding a match
    assert got[0] == want[0] and np.array_equal(got[1], want[1])
    # a predicted level outside the scale table is an argument error, not a read past the table
    bad = lvl.copy(); bad[7] = 8
    a, keep = build(ones, bad)
    with pytest.raises(pkg.OrbGpuError):
        pkg.search_local_points(a)
    bad[7] = 99; inv = ones.copy(); inv[7] = 0                # ... unless that point is not in view
    a, keep = build(inv, bad)
    want = port.search_local_points(a); got = pkg.search_local_points(a)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])


@pytest.mark.gpu
def test_device_crowded_grid_cells_keep_index_order(pkg, port):
    """Hundreds of features inside a few grid cells (k_grid_build's long-range path): candidates must still be visited in
    index order, which decides ties between equal descriptors."""
    rng = np.random.default_rng(77)
    pool = rng.integers(0, 256, (3, 32), dtype=np.uint8)
    k2, d2 = random_frame(rng, 900, 640, 480, pool)
    crowd = rng.permutation(900)[:500]
    k2["x"][crowd] = (300 + rng.uniform(0, 25, 500)).astype(np.float32); k2["y"][crowd] = (200 + rng.uniform(0, 12, 500)).astype(np.float32)
    k1, d1 = random_frame(rng, 600, 640, 480, pool)
    k1["x"][:400] = (305 + rng.uniform(0, 15, 400)).astype(np.float32); k1["y"][:400] = (203 + rng.uniform(0, 6, 400)).astype(np.float32)
    s = port.tables()[0][0]
    bounds = (0.0, 640.0, 0.0, 480.0)
    for th, ratio in [(1.0, 0.8), (5.0, 0.95)]:
        a, keep = pkg.local_points_args(np.ones(600, np.uint8), k1["x"], k1["y"], k1["x"] - 4, k1["octave"], np.full(600, 0.9, np.float32),
                                        d1, rng.random(600) < 0.8, k2, d2, np.full(900, -1.0, np.float32), np.zeros(900, np.uint8),
                                        bounds, s, th, ratio)
        want = port.search_local_points(a); got = pkg.search_local_points(a)
        assert want[0] > 50
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
    k1["octave"][:] = 0
    pm = np.stack([k1["x"], k1["y"]], 1)
    want = port.search_for_initialization(k1, d1, k2, d2, bounds, pm, 20, 0.9, True)
    got = pkg.search_for_initialization(k1, d1, k2, d2, bounds, pm, 20, 0.9, True)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])
