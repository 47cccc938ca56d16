"""Tracking::SearchLocalPoints (src/Tracking.cc:1150-1200) in one call: Frame::isInFrustum (src/Frame.cc:269-325) with
MapPoint::PredictScale (src/MapPoint.cc:402-417) for every candidate local map point, then the local-map
SearchByProjection (src/ORBmatcher.cc:44-125).
CPU: the restatement (oracle/frame_port.cc:frameport_search_local_map) against the reference's own Frame::isInFrustum /
MapPoint / ORBmatcher code (oracle/_ref/libframeref.so) and against the geometry of a two-view scene.
GPU: orbgpu_search_local_map bit-for-bit against the restatement (all six mTrack* fields and the matches), and the
device logf against glibc's over a sweep of float bit patterns."""
import numpy as np
import pytest

import oracle_lib
from test_search import SHIFT, device_extract, oracle_extract, random_frame, two_frames

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")
CAM = (517.3, 516.5, 318.6, 255.3)
LOG12 = float(np.log(np.float32(1.2)).astype(np.float32))        # Frame: mfLogScaleFactor = log(mfScaleFactor)


def local_map_scene(pkg, port, F1, F2, bounds, th, seed, stereo=True, Z=25.0):
    """Local map = the features of frame 1 lifted to depth ~Z; the current frame sees them shifted by SHIFT.  Normals point
    from the point towards a reference camera near the current one; the scale-invariance distances put the predicted
    level at the octave the feature was seen in (sometimes beside it, sometimes out of range)."""
    (k1, d1), (k2, d2) = F1, F2
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = CAM
    n1, n2 = len(k1), len(k2)
    z = (Z + rng.uniform(-0.5, 0.5, n1)).astype(np.float32)
    Tl = np.eye(4, dtype=np.float32)
    ang = 0.003
    Tl[:3, :3] = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], np.float32)
    Tl[:3, 3] = [0.3, -0.2, 0.1]
    pc = np.stack([(k1["x"] - cx) * z / fx, (k1["y"] - cy) * z / fy, z], 1).astype(np.float64)
    Rl, tl = Tl[:3, :3].astype(np.float64), Tl[:3, 3].astype(np.float64)
    pw = ((pc - tl) @ Rl).astype(np.float32)
    Tc = Tl.copy()
    Tc[:3, 3] += np.array([SHIFT[0] * Z / fx, SHIFT[1] * Z / fy, 0.05], np.float32)
    Ow = -(Rl.T @ tl)
    po = pw.astype(np.float64) - Ow
    dist = np.linalg.norm(po, axis=1)
    nrm = po / dist[:, None]
    tilt = rng.random(n1)
    nrm[tilt < 0.08] *= -1.0                                   # seen from behind: viewCos < 0.5
    side = np.cross(nrm, [0.0, 1.0, 0.0]); side /= np.linalg.norm(side, axis=1)[:, None]
    mid = (tilt > 0.08) & (tilt < 0.5)                         # around the 0.998 radius switch and the 0.5 limit
    a = rng.choice([0.03, 0.0632, 0.0633, 0.2, 1.0, 1.047, 1.048], n1)
    nrm[mid] = (np.cos(a)[:, None] * nrm + np.sin(a)[:, None] * side)[mid]
    octv = k1["octave"].astype(np.float64) + rng.choice([-0.5, -0.5, -0.5, 0.0, -1.0, 0.4], n1)
    maxd = (dist * 1.2 ** octv).astype(np.float32)
    mind = (maxd / np.float32(1.2 ** 7)).astype(np.float32)
    far = rng.random(n1) < 0.06
    maxd[far] *= np.float32(0.2)                               # dist > 1.2 * mfMaxDistance
    near = rng.random(n1) < 0.04
    mind[near] = maxd[near]                                    # dist < 0.8 * mfMinDistance for the upper octaves
    ur = np.where(rng.random(n2) < 0.6, k2["x"] - 40.0 / Z + rng.normal(0, 1.0, n2), -1.0).astype(np.float32)
    if not stereo:
        ur[:] = -1.0
    s4 = port.tables()[0]
    return pkg.local_map_args(rng.random(n1) < 0.9, pw, nrm, maxd, mind, d1, rng.random(n1) < 0.9, Tc, CAM, 40.0, k2, d2, ur,
                              rng.random(n2) < 0.05, bounds, s4[0], LOG12, th, 0.8)


def random_local_map_args(pkg, port, seed, n_cur=None):
    """Points all around the camera (behind it, outside the image, at any distance), arbitrary normals and distance bands,
    pooled descriptors."""
    rng = np.random.default_rng(seed)
    w, h = 640, 480
    pool = rng.integers(0, 256, (30, 32), dtype=np.uint8)
    n1 = int(rng.integers(500, 4000)); n2 = int(rng.integers(300, 1500)) if n_cur is None else n_cur
    k2, d2 = random_frame(rng, max(n2, 1), w, h, pool)
    k2, d2 = k2[:n2], d2[:n2]
    fx, fy, cx, cy = CAM
    ang = float(rng.uniform(-0.3, 0.3))
    Tc = np.eye(4, dtype=np.float32)
    Tc[:3, :3] = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], np.float32)
    Tc[:3, 3] = rng.uniform(-1, 1, 3).astype(np.float32)
    # points: most aimed at a feature of the frame (so that there is something to match), the rest anywhere
    z = rng.uniform(-3, 30, n1).astype(np.float32); z[np.abs(z) < 0.05] = 0.0        # some exactly on the camera plane
    hit = rng.integers(0, max(n2, 1), n1)
    aim = (rng.random(n1) < 0.6) & (n2 > 0)
    px = np.where(aim, k2["x"][hit] if n2 else 0 + rng.normal(0, 1.5, n1), rng.uniform(-200, w + 200, n1))
    py = np.where(aim, k2["y"][hit] if n2 else 0 + rng.normal(0, 1.5, n1), rng.uniform(-200, h + 200, n1))
    pc = np.stack([(px - cx) * z / fx, (py - cy) * z / fy, z], 1).astype(np.float64)
    R, t = Tc[:3, :3].astype(np.float64), Tc[:3, 3].astype(np.float64)
    pw = ((pc - t) @ R).astype(np.float32)
    po = pw.astype(np.float64) + (R.T @ t)
    dist = np.linalg.norm(po, axis=1) + 1e-9
    nrm = po / dist[:, None] + rng.normal(0, 0.6, (n1, 3))
    nrm /= np.linalg.norm(nrm, axis=1)[:, None] + 1e-12
    lvl = np.where(aim & (n2 > 0), k2["octave"][hit] if n2 else 0, rng.integers(0, 8, n1)).astype(np.float64)
    maxd = (dist * 1.2 ** (lvl - rng.uniform(0, 1, n1)) * rng.choice([1.0, 1.0, 1.0, 0.1, 30.0], n1)).astype(np.float32)
    mind = (maxd * rng.choice([0.2, 0.279, 1.0], n1).astype(np.float32)).astype(np.float32)
    d1 = pool[rng.integers(0, len(pool), n1)].copy()
    if n2:
        d1 = np.where(aim[:, None], d2[hit], d1)
    ur = np.where(rng.random(n2) < 0.5, k2["x"] - rng.uniform(0, 10, n2), -1.0).astype(np.float32)
    bounds = (-8.5, w + 7.25, -6.0, h + 9.5)
    th = float(rng.choice([1.0, 3.0, 5.0]))
    return pkg.local_map_args(rng.random(n1) < 0.85, pw, nrm, maxd, mind, d1, rng.random(n1) < 0.8, Tc, CAM, float(rng.choice([40.0, 386.0])),
                              k2, d2, ur, rng.random(n2) < 0.1, bounds, port.tables()[0][0], LOG12, th, float(rng.choice([0.8, 0.6])))


def same(a, b):
    """(nmatches, cur_match, n_in_view, fields): everything bit for bit; returns True or raises with the first difference"""
    assert a[0] == b[0] and a[2] == b[2], ("counts", a[0], b[0], a[2], b[2])
    assert np.array_equal(a[1], b[1]), ("cur_match", np.nonzero(a[1] != b[1])[0][:5])
    for k in a[3]:
        if a[3][k].tobytes() != b[3][k].tobytes():
            bad = np.nonzero(a[3][k].view(np.uint8).reshape(len(a[3][k]), -1) != b[3][k].view(np.uint8).reshape(len(b[3][k]), -1))[0]
            i = int(bad[0])
            raise AssertionError((k, i, a[3][k][i], b[3][k][i], {f: (a[3][f][i], b[3][f][i]) for f in a[3]}))
    return True


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


# ------------------------------------------------------------------------------------------------------------- CPU side
def test_oracle_local_map_follows_the_geometry(pkg, port, scene):
    F1, F2, bounds = scene
    args, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, 3, stereo=False)
    nm, cm, nv, f = port.search_local_map(args, pkg.track_fields)
    iv = f["in_view"].astype(bool)
    assert nv == iv.sum() > 500 and nm > 250
    assert not iv[~keep["candidate"].astype(bool)].any()
    k1 = F1[0]
    assert np.abs(f["proj_x"][iv] - (k1["x"][iv] + SHIFT[0])).max() < 1.5      # depth jitter + the small forward motion
    assert np.abs(f["proj_y"][iv] - (k1["y"][iv] + SHIFT[1])).max() < 1.5
    assert (np.abs(f["scale_level"][iv] - k1["octave"][iv]) <= 1).all()
    assert (f["view_cos"][iv] >= 0.5).all() and ((f["view_cos"][iv] > 0.998).sum() > 50) and ((f["view_cos"][iv] <= 0.998).sum() > 50)
    assert (f["proj_x"][~iv] == 0).all() and (f["scale_level"][~iv] == 0).all()
    got = np.nonzero(cm >= 0)[0]
    assert iv[cm[got]].all()


def test_logf_restatement_equals_glibc(port):
    """The algorithm k_frustum runs (glibc 2.39 logf, separate multiplies and adds) against the host's logf: every 13th
    positive float pattern from the subnormals to inf (165 M values; the full sweep of all 2.1 G was run once, 0
    mismatches with and without FMA contraction), and the negative / NaN patterns sampled."""
    assert port.logf_port_mismatches(0, 0x7f800001 // 13 + 1, 13) == 0
    assert port.logf_port_mismatches(0x7f800000, 1 << 16, 127) == 0          # inf, NaNs, -0, negatives
    assert port.logf_port_mismatches(0x3f000000, 1 << 24, 1) == 0            # every float in [0.5, 2)


@needs_ref
@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
def test_local_map_restatement_equals_reference(pkg, port, scene, th):
    F1, F2, bounds = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    for seed, stereo in [(1, True), (2, False)]:
        args, keep = local_map_scene(pkg, port, F1, F2, bounds, th, seed, stereo)
        want = ref.search_local_map(args, pkg.track_fields)
        got = port.search_local_map(args, pkg.track_fields)
        assert want[0] > 200 and want[2] > 500
        assert same(got, want)


@needs_ref
@pytest.mark.parametrize("seed", [41, 42, 43, 44, 45, 46, 47, 48])
def test_local_map_on_random_points_restatement_equals_reference(pkg, port, seed):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    args, keep = random_local_map_args(pkg, port, seed)
    want = ref.search_local_map(args, pkg.track_fields)
    got = port.search_local_map(args, pkg.track_fields)
    assert want[2] > 30
    assert same(got, want)


@needs_ref
def test_local_map_degenerate_points_restatement_equals_reference(pkg, port):
    """Points at the camera centre (dist = 0 -> logf(inf), the int conversion of inf) and zero distance bands."""
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    args, keep = random_local_map_args(pkg, port, 63)
    R, t = keep["Tcw"][:3, :3], keep["Tcw"][:3, 3]
    keep["world_pos"][:40] = -(R.T @ t)
    keep["min_distance"][:80] = 0.0
    keep["max_distance"][40:60] = 0.0
    want = ref.search_local_map(args, pkg.track_fields)
    got = port.search_local_map(args, pkg.track_fields)
    assert same(got, want)


# ------------------------------------------------------------------------------------------------------------- GPU side
@pytest.mark.gpu
def test_device_logf_equals_glibc(pkg, port):
    """MapPoint::PredictScale calls glibc's logf; the device restatement must agree on every float it can meet."""
    bits = np.arange(0, 0x7f800000, 509, dtype=np.uint32)                      # 4.2 M positive floats, every exponent
    special = np.array([0, 1, 0x007fffff, 0x00800000, 0x3f7fffff, 0x3f800000, 0x3f800001, 0x3f330000, 0x3f32ffff, 0x7f7fffff,
                        0x7f800000, 0x7fc00000, 0x80000000, 0xbf800000, 0xff800000], np.uint32)
    rng = np.random.default_rng(3)
    near1 = (np.float32(1) + rng.normal(0, 0.3, 500000).astype(np.float32)).view(np.uint32)
    x = np.concatenate([bits, special, near1, np.arange(1, 1 << 23, 4099, dtype=np.uint32)]).view(np.float32)
    want = port.logf(x)
    got = pkg.debug_logf(x)
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert got[~nan].tobytes() == want[~nan].tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("th", [1.0, 3.0, 5.0])
def test_device_local_map_equals_oracle(pkg, port, th):
    F1, F2, bounds = two_frames(pkg, device_extract(pkg))
    for seed, stereo in [(1, True), (2, False)]:
        args, keep = local_map_scene(pkg, port, F1, F2, bounds, th, seed, stereo)
        want = port.search_local_map(args, pkg.track_fields)
        got = pkg.search_local_map(args)
        assert want[0] > 200 and want[2] > 500
        assert same(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [41, 42, 43, 44, 45, 46, 47, 48, 49, 50])
def test_device_local_map_equals_oracle_on_random_points(pkg, port, seed):
    args, keep = random_local_map_args(pkg, port, seed)
    want = port.search_local_map(args, pkg.track_fields)
    got = pkg.search_local_map(args)
    assert same(got, want)


@pytest.mark.gpu
def test_device_local_map_edge_cases(pkg, port):
    # no features: a batched isInFrustum
    args, keep = random_local_map_args(pkg, port, 61, n_cur=0)
    want = port.search_local_map(args, pkg.track_fields); got = pkg.search_local_map(args)
    assert want[2] > 30 and want[0] == 0 and same(got, want)
    # no candidate points
    args, keep = random_local_map_args(pkg, port, 62)
    keep["candidate"][:] = 0
    want = port.search_local_map(args, pkg.track_fields); got = pkg.search_local_map(args)
    assert want[2] == 0 and want[0] == 0 and same(got, want)
    # points exactly at the camera centre (dist = 0: logf(inf), the cvttss2si of inf) and with zero distance bands
    args, keep = random_local_map_args(pkg, port, 63)
    R, t = keep["Tcw"][:3, :3], keep["Tcw"][:3, 3]
    keep["world_pos"][:40] = -(R.T @ t)
    keep["min_distance"][:80] = 0.0
    keep["max_distance"][40:60] = 0.0
    want = port.search_local_map(args, pkg.track_fields); got = pkg.search_local_map(args)
    assert same(got, want)
