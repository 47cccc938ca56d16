"""The three Sim3 routines of loop closing: ORBmatcher::Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint)
(src/ORBmatcher.cc:969-1097), SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (:290-403) and SearchBySim3
(:1102-1326).  CPU: the restatements (oracle/frame_port.cc) against the reference's own code on KeyFrames / MapPoints built by
its constructors (oracle/_ref/libframeref.so); GPU: the device entry points bit-for-bit against the restatements."""
import numpy as np
import pytest

import oracle_lib
from test_local_map import CAM, LOG12, local_map_scene, random_local_map_args
from test_search import SHIFT, device_extract, oracle_extract, two_frames

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")


def sim3_of(T, s):
    S = T.astype(np.float32).copy()
    S[:3, :] = (S[:3, :] * np.float32(s)).astype(np.float32)
    return S


def fuse_args_from(pkg, port, keep, th, bounds, seed=0, present_p=0.85):
    rng = np.random.default_rng(seed)
    s = port.tables()[0][0]
    inv_sigma2 = (np.float32(1.0) / (s * s)).astype(np.float32)
    n1 = len(keep["candidate"])
    return pkg.fuse_args(rng.random(n1) < present_p, keep["world_pos"], keep["normal"], keep["max_distance"], keep["min_distance"],
                         keep["mp_desc"], keep["Tcw"], np.zeros(3, np.float32), CAM, 40.0, keep["cur_keys_un"], keep["cur_desc"],
                         keep["cur_u_right"], bounds, s, inv_sigma2, LOG12, th)


def lift(k, T, rng, Z=25.0):
    """Map points of a keyframe: its keypoints at depth ~Z in its own camera, as world points; plus the distance band that
    predicts the keypoint's octave (sometimes a neighbouring one) from about that distance."""
    fx, fy, cx, cy = CAM
    n = len(k)
    z = (Z + rng.uniform(-0.5, 0.5, n)).astype(np.float32)
    pc = np.stack([(k["x"] - cx) * z / fx, (k["y"] - cy) * z / fy, z], 1).astype(np.float64)
    R, t = T[:3, :3].astype(np.float64), T[:3, 3].astype(np.float64)
    pw = ((pc - t) @ R).astype(np.float32)
    dist = np.linalg.norm(pc, axis=1)
    octv = k["octave"].astype(np.float64) + rng.choice([-0.5, -0.5, -0.5, 0.0, -1.0, 0.4], n)
    maxd = (dist * 1.2 ** octv).astype(np.float32)
    mind = (maxd / np.float32(1.2 ** 7)).astype(np.float32)
    far = rng.random(n) < 0.05
    maxd[far] *= np.float32(0.2)
    return pw, maxd, mind


def sim3_pair(pkg, port, F1, F2, bounds, th, seed, s12=1.0, jitter=0.0):
    (k1, d1), (k2, d2) = F1, F2
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy = CAM
    Z = 25.0
    T1 = np.eye(4, dtype=np.float32)
    ang = 0.003
    T1[:3, :3] = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], np.float32)
    T1[:3, 3] = [0.3, -0.2, 0.1]
    T2 = T1.copy()
    T2[:3, 3] += np.array([SHIFT[0] * Z / fx, SHIFT[1] * Z / fy, 0.05], np.float32)
    pw1, maxd1, mind1 = lift(k1, T1, rng)
    pw2, maxd2, mind2 = lift(k2, T2, rng)
    s = port.tables()[0][0]
    zeros3 = np.zeros(3, np.float32)
    a12, keep12 = pkg.fuse_args(rng.random(len(k1)) < 0.8, pw1, np.zeros_like(pw1), maxd1, mind1, d1, T1, zeros3, CAM, 40.0, k2, d2,
                                np.full(len(k2), -1.0, np.float32), bounds, s, s, LOG12, th)
    a21, keep21 = pkg.fuse_args(rng.random(len(k2)) < 0.8, pw2, np.zeros_like(pw2), maxd2, mind2, d2, T2, zeros3, CAM, 40.0, k1, d1,
                                np.full(len(k1), -1.0, np.float32), bounds, s, s, LOG12, th)
    R1, t1, R2, t2 = T1[:3, :3].astype(np.float64), T1[:3, 3].astype(np.float64), T2[:3, :3].astype(np.float64), T2[:3, 3].astype(np.float64)
    R12 = R1 @ R2.T
    t12 = t1 - R12 @ t2 * s12 + rng.normal(0, jitter, 3)
    return a12, a21, np.float32(s12), R12.astype(np.float32), t12.astype(np.float32), (keep12, keep21)


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


CASES = [(1.0, 3.0, (0.0, 640.0, 0.0, 480.0)), (0.83, 4.0, (-8.5, 647.25, -6.75, 489.5)), (1.37, 10.0, (3.9, 636.2, 2.6, 477.9))]


# ------------------------------------------------------------------------------------------------------------- CPU side
@needs_ref
@pytest.mark.parametrize("cfg", CASES, ids=["s1", "s0.83", "s1.37"])
def test_fuse_sim3_restatement_equals_reference(pkg, port, scene, cfg):
    sc, th, bounds = cfg
    F1, F2, _ = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    for seed in (1, 2):
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, fkeep = fuse_args_from(pkg, port, keep, th, bounds, seed)
        S = sim3_of(keep["Tcw"], sc)
        want = ref.fuse_search_sim3(args, S)
        got = port.fuse_search_sim3(args, S)
        assert (want >= 0).sum() > 200
        assert np.array_equal(got, want)


@needs_ref
@pytest.mark.parametrize("cfg", CASES, ids=["s1", "s0.83", "s1.37"])
def test_projection_sim3_restatement_equals_reference(pkg, port, scene, cfg):
    sc, th, bounds = cfg
    F1, F2, _ = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    for seed in (1, 2):
        rng = np.random.default_rng(seed)
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, fkeep = fuse_args_from(pkg, port, keep, int(th), bounds, seed)
        S = sim3_of(keep["Tcw"], sc)
        mi = rng.random(args.n_kf) < 0.15
        wn, wm = ref.search_by_projection_sim3(args, S, mi)
        gn, gm = port.search_by_projection_sim3(args, S, mi)
        assert wn > 150
        assert gn == wn and np.array_equal(gm, wm)


@needs_ref
@pytest.mark.parametrize("cfg", [(1.0, 7.5, 0.0), (1.0, 7.5, 0.02), (1.05, 10.0, 0.0)], ids=["exact", "jitter", "scaled"])
def test_search_by_sim3_restatement_equals_reference(pkg, port, scene, cfg):
    s12, th, jit = cfg
    F1, F2, bounds = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    total = 0
    for seed in (1, 2):
        a12, a21, s, R12, t12, keep = sim3_pair(pkg, port, F1, F2, bounds, th, seed, s12, jit)
        wn, wm = ref.search_by_sim3(a12, a21, s, R12, t12)
        gn, gm = port.search_by_sim3(a12, a21, s, R12, t12)
        assert gn == wn and np.array_equal(gm, wm)
        total += wn
    assert total > (300 if s12 == 1.0 else 20)


@needs_ref
@pytest.mark.parametrize("seed", [81, 82, 83, 84])
def test_sim3_routines_on_random_points_restatement_equals_reference(pkg, port, seed):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    rng = np.random.default_rng(seed)
    _, keep = random_local_map_args(pkg, port, seed)
    bounds = (-8.5, 647.25, -6.0, 489.5)
    args, fkeep = fuse_args_from(pkg, port, keep, float(rng.choice([3.0, 4.0, 10.0])), bounds, seed)
    S = sim3_of(keep["Tcw"], float(rng.uniform(0.5, 2.0)))
    assert np.array_equal(port.fuse_search_sim3(args, S), ref.fuse_search_sim3(args, S))
    mi = rng.random(args.n_kf) < 0.2
    w = ref.search_by_projection_sim3(args, S, mi); g = port.search_by_projection_sim3(args, S, mi)
    assert g[0] == w[0] and np.array_equal(g[1], w[1])


# ------------------------------------------------------------------------------------------------------------- GPU side
@pytest.mark.gpu
@pytest.mark.parametrize("cfg", CASES, ids=["s1", "s0.83", "s1.37"])
def test_device_fuse_and_projection_sim3_equal_oracle(pkg, port, cfg):
    sc, th, bounds = cfg
    F1, F2, _ = two_frames(pkg, device_extract(pkg))
    for seed in (1, 2):
        rng = np.random.default_rng(seed)
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, fkeep = fuse_args_from(pkg, port, keep, int(th), bounds, seed)
        S = sim3_of(keep["Tcw"], sc)
        want = port.fuse_search_sim3(args, S)
        assert (want >= 0).sum() > 200
        assert np.array_equal(pkg.fuse_search_sim3(args, S), want)
        mi = rng.random(args.n_kf) < 0.15
        w = port.search_by_projection_sim3(args, S, mi); g = pkg.search_by_projection_sim3(args, S, mi)
        assert w[0] > 150 and g[0] == w[0] and np.array_equal(g[1], w[1])


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(1.0, 7.5, 0.0), (1.0, 7.5, 0.02), (1.05, 10.0, 0.0)], ids=["exact", "jitter", "scaled"])
def test_device_search_by_sim3_equals_oracle(pkg, port, cfg):
    s12, th, jit = cfg
    F1, F2, bounds = two_frames(pkg, device_extract(pkg))
    for seed in (1, 2):
        a12, a21, s, R12, t12, keep = sim3_pair(pkg, port, F1, F2, bounds, th, seed, s12, jit)
        w = port.search_by_sim3(a12, a21, s, R12, t12); g = pkg.search_by_sim3(a12, a21, s, R12, t12)
        assert g[0] == w[0] and np.array_equal(g[1], w[1])
    assert w[0] > (150 if s12 == 1.0 else 5)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [81, 82, 83, 84, 85, 86])
def test_device_sim3_routines_equal_oracle_on_random_points(pkg, port, seed):
    rng = np.random.default_rng(seed)
    _, keep = random_local_map_args(pkg, port, seed)
    bounds = (-8.5, 647.25, -6.0, 489.5)
    args, fkeep = fuse_args_from(pkg, port, keep, float(rng.choice([3.0, 4.0, 10.0])), bounds, seed)
    S = sim3_of(keep["Tcw"], float(rng.uniform(0.5, 2.0)))
    assert np.array_equal(pkg.fuse_search_sim3(args, S), port.fuse_search_sim3(args, S))
    mi = rng.random(args.n_kf) < 0.2
    w = port.search_by_projection_sim3(args, S, mi); g = pkg.search_by_projection_sim3(args, S, mi)
    assert g[0] == w[0] and np.array_equal(g[1], w[1])
