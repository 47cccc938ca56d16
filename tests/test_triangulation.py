"""ORBmatcher::SearchForTriangulation + CheckDistEpipolarLine (src/ORBmatcher.cc:657-820, :140-157): the BoW-gated,
epipolar-checked search LocalMapping::CreateNewMapPoints runs against every neighbouring keyframe.
CPU: the restatement (oracle/frame_port.cc:frameport_search_for_triangulation) against the reference's own code on
KeyFrames built by the reference's constructor (oracle/_ref/libframeref.so) and against the epipolar geometry of a two-view
scene; GPU: orbgpu_search_for_triangulation bit-for-bit against the restatement."""
import numpy as np
import pytest

import oracle_lib
from test_search import SHIFT, device_extract, feature_vector, oracle_extract, random_frame, two_frames

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")
CAM = (517.3, 516.5, 318.6, 255.3)


def skew(t):
    return np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float64)


def compute_f12(T1w, T2w, cam):
    """LocalMapping::ComputeF12 (stays host in the reference): F12 = K1^-T [t12]x R12 K2^-1, rounded to float."""
    fx, fy, cx, cy = cam
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float64)
    R1, t1 = T1w[:3, :3].astype(np.float64), T1w[:3, 3].astype(np.float64)
    R2, t2 = T2w[:3, :3].astype(np.float64), T2w[:3, 3].astype(np.float64)
    R12 = R1 @ R2.T
    t12 = -R12 @ t2 + t1
    return (np.linalg.inv(K).T @ skew(t12) @ R12 @ np.linalg.inv(K)).astype(np.float32)


def poses(motion_z=0.0, Z=25.0):
    """Camera 2 = camera 1 moved so that points at depth Z shift by SHIFT pixels (+ motion_z along the optical axis)."""
    fx, fy, cx, cy = CAM
    T1 = np.eye(4, dtype=np.float32)
    ang = 0.002
    T1[:3, :3] = np.array([[np.cos(ang), -np.sin(ang), 0], [np.sin(ang), np.cos(ang), 0], [0, 0, 1]], np.float32)
    T1[:3, 3] = [0.3, -0.2, 0.1]
    T2 = T1.copy()
    T2[:3, 3] += np.array([SHIFT[0] * Z / fx, SHIFT[1] * Z / fy, motion_z], np.float32)
    return T1, T2


def camera_center(T):
    R, t = T[:3, :3], T[:3, 3]
    return (-(R.T.astype(np.float32) @ t.astype(np.float32))).astype(np.float32)


def tri_scene(pkg, port, F1, F2, seed, n_words=24, only_stereo=False, ori=True, motion_z=0.0, Cw1=None, F12=None):
    (k1, d1), (k2, d2) = F1, F2
    rng = np.random.default_rng(seed)
    n1, n2 = len(k1), len(k2)
    T1, T2 = poses(motion_z)
    s4 = port.tables()[0]
    sigma2 = (s4[0] * s4[0]).astype(np.float32)
    ur1 = np.where(rng.random(n1) < 0.4, k1["x"] - 3.0, -1.0).astype(np.float32)
    ur2 = np.where(rng.random(n2) < 0.4, k2["x"] - 3.0, -1.0).astype(np.float32)
    side1 = (k1, d1, rng.random(n1) < 0.3, ur1) + feature_vector(d1, n_words)
    side2 = (k2, d2, rng.random(n2) < 0.3, ur2) + feature_vector(d2, n_words)
    args, keep = pkg.triangulation_args(side1, side2, compute_f12(T1, T2, CAM) if F12 is None else F12,
                                        camera_center(T1) if Cw1 is None else Cw1, T2, CAM, s4[0], sigma2, only_stereo, ori)
    return args, keep, T1


def random_tri_args(pkg, port, seed):
    """Pooled descriptors (distance ties: the LAST passing candidate wins), the epipole inside the image, random flags."""
    rng = np.random.default_rng(seed)
    pool = rng.integers(0, 256, (25, 32), dtype=np.uint8)
    n1, n2 = int(rng.integers(300, 1800)), int(rng.integers(300, 1800))
    k1, d1 = random_frame(rng, n1, 640, 480, pool); k2, d2 = random_frame(rng, n2, 640, 480, pool)
    T1 = np.eye(4, dtype=np.float32); T2 = np.eye(4, dtype=np.float32)
    T2[:3, 3] = [float(rng.uniform(-0.3, 0.3)), float(rng.uniform(-0.2, 0.2)), float(rng.choice([1.0, -0.8, 0.5]))]
    s4 = port.tables()[0]
    sigma2 = (s4[0] * s4[0]).astype(np.float32) * np.float32(rng.choice([1.0, 40.0, 400.0]))     # wide bands: many candidates pass
    words = int(rng.choice([5, 24, 80]))

    def fv(n):
        node = rng.integers(0, words, n)
        ids = np.unique(node); start = [0]; idx = []
        for i in ids:
            idx.extend(np.nonzero(node == i)[0].tolist()); start.append(len(idx))
        return ids.astype(np.uint32), np.array(start, np.int32), np.array(idx, np.uint32)
    ur1 = np.where(rng.random(n1) < 0.5, k1["x"] - 3.0, -1.0).astype(np.float32)
    ur2 = np.where(rng.random(n2) < 0.5, k2["x"] - 3.0, -1.0).astype(np.float32)
    ur1[rng.random(n1) < 0.05] = 0.0                               # mvuRight == 0 counts as stereo (">= 0")
    side1 = (k1, d1, rng.random(n1) < 0.3, ur1) + fv(n1)
    side2 = (k2, d2, rng.random(n2) < 0.3, ur2) + fv(n2)
    F12 = compute_f12(T1, T2, CAM) if seed % 3 else rng.normal(0, 1, (3, 3)).astype(np.float32) * np.float32(1e-3)
    args, keep = pkg.triangulation_args(side1, side2, F12, camera_center(T1), T2, CAM, s4[0], sigma2, bool(seed % 2), bool(seed % 4 < 2))
    return args, keep, T1


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


# ------------------------------------------------------------------------------------------------------------- CPU side
def test_oracle_triangulation_matches_obey_the_epipolar_geometry(pkg, port, scene):
    F1, F2, _ = scene
    args, keep, T1 = tri_scene(pkg, port, F1, F2, 3)
    nm, m12 = port.search_for_triangulation(args)
    ok = np.nonzero(m12 >= 0)[0]
    assert nm == len(ok) > 60
    (k1, _), (k2, _) = F1, F2
    has1, has2 = keep[0][2], keep[1][2]
    assert not has1[ok].any() and not has2[m12[ok]].any()         # only features without a map point, on both sides
    F12 = compute_f12(*poses(), CAM).astype(np.float64)
    sig = (port.tables()[0][0].astype(np.float64)) ** 2
    for i in ok:
        j = m12[i]
        l = np.array([k1["x"][i], k1["y"][i], 1.0]) @ F12
        d2 = (l[0] * k2["x"][j] + l[1] * k2["y"][j] + l[2]) ** 2 / (l[0] ** 2 + l[1] ** 2)
        assert d2 < 3.84 * sig[k2["octave"][j]] * 1.001
    dx = k2["x"][m12[ok]] - k1["x"][ok]; dy = k2["y"][m12[ok]] - k1["y"][ok]
    assert (np.abs(dx - SHIFT[0]) < 4).mean() > 0.7 and (np.abs(dy - SHIFT[1]) < 4).mean() > 0.7
    # bOnlyStereo keeps only pairs whose two features have a right coordinate
    args, keep, _ = tri_scene(pkg, port, F1, F2, 3, only_stereo=True)
    nm2, m2 = port.search_for_triangulation(args)
    ok2 = np.nonzero(m2 >= 0)[0]
    assert 0 < nm2 < nm and (keep[0][3][ok2] >= 0).all() and (keep[1][3][m2[ok2]] >= 0).all()


@needs_ref
@pytest.mark.parametrize("cfg", [(3, 24, False, True, 0.0), (4, 100, False, False, 0.5), (5, 1, True, True, -0.4), (6, 24, False, True, 2.0)],
                         ids=["sideways", "forward", "stereo-backward", "fast-forward"])
def test_triangulation_restatement_equals_reference(pkg, port, scene, cfg):
    seed, words, only_stereo, ori, mz = cfg
    F1, F2, _ = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    args, keep, T1 = tri_scene(pkg, port, F1, F2, seed, words, only_stereo, ori, mz)
    wn, wm, cw = ref.search_for_triangulation(args, T1)
    args, keep, T1 = tri_scene(pkg, port, F1, F2, seed, words, only_stereo, ori, mz, Cw1=cw)      # the centre the reference derived
    gn, gm = port.search_for_triangulation(args)
    assert wn > 15
    assert gn == wn and np.array_equal(gm, wm)


@needs_ref
@pytest.mark.parametrize("seed", [31, 32, 33, 34, 35, 36, 37, 38])
def test_triangulation_on_random_frames_restatement_equals_reference(pkg, port, seed):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    args, keep, T1 = random_tri_args(pkg, port, seed)
    wn, wm, cw = ref.search_for_triangulation(args, T1)
    assert np.array_equal(cw, camera_center(T1))                  # identity rotation: the centre is exact either way (up to -0)
    gn, gm = port.search_for_triangulation(args)
    assert gn == wn and np.array_equal(gm, wm)


def test_random_triangulation_cases_are_not_trivial(pkg, port):
    total = 0
    for seed in range(31, 39):
        args, keep, T1 = random_tri_args(pkg, port, seed)
        total += port.search_for_triangulation(args)[0]
    assert total > 400


# ------------------------------------------------------------------------------------------------------------- GPU side
@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(3, 24, False, True, 0.0), (4, 100, False, False, 0.5), (5, 1, True, True, -0.4), (6, 24, False, True, 2.0)],
                         ids=["sideways", "forward", "stereo-backward", "fast-forward"])
def test_device_triangulation_equals_oracle(pkg, port, cfg):
    seed, words, only_stereo, ori, mz = cfg
    for nf in (1000, 2000):
        F1, F2, _ = two_frames(pkg, device_extract(pkg), 640, 480, 0x7A1A + nf, nf)
        args, keep, T1 = tri_scene(pkg, port, F1, F2, seed, words, only_stereo, ori, mz)
        want = port.search_for_triangulation(args)
        got = pkg.search_for_triangulation(args)
        assert want[0] > 15
        assert got[0] == want[0] and np.array_equal(got[1], want[1])


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [31, 32, 33, 34, 35, 36, 37, 38, 39, 40])
def test_device_triangulation_equals_oracle_on_random_frames(pkg, port, seed):
    args, keep, T1 = random_tri_args(pkg, port, seed)
    want = port.search_for_triangulation(args)
    got = pkg.search_for_triangulation(args)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])


@pytest.mark.gpu
def test_device_triangulation_batch_equals_oracle(pkg, port):
    """One keyframe against its neighbours in one call, as CreateNewMapPoints visits them (src/LocalMapping.cc:239-270)"""
    jobs = [random_tri_args(pkg, port, seed) for seed in range(50, 62)]
    F1, F2, _ = two_frames(pkg, device_extract(pkg), 640, 480, 0x7A1A + 1000, 1000)
    jobs += [tri_scene(pkg, port, F1, F2, seed, 24, False, True, mz) for seed, mz in [(3, 0.0), (4, 0.5), (5, -0.4)]]
    rng = np.random.default_rng(3)
    k1, d1 = random_frame(rng, 50, 640, 480, rng.integers(0, 256, (3, 32), dtype=np.uint8))
    empty = (np.zeros(0, np.uint32), np.zeros(1, np.int32), np.zeros(0, np.uint32))
    s = port.tables()[0][0]
    T1, T2 = poses(0.3)
    side = (k1, d1, np.zeros(50, bool), np.full(50, -1.0, np.float32))
    jobs.insert(4, pkg.triangulation_args(side + empty, side + empty, compute_f12(T1, T2, CAM), camera_center(T1), T2, CAM, s,
                                          (s * s).astype(np.float32), False, True) + (T1,))     # nothing shared: no queries
    got = pkg.search_for_triangulation_batch([j[0] for j in jobs])
    assert len(got) == len(jobs)
    total = 0
    for j, g in zip(jobs, got):
        want = port.search_for_triangulation(j[0])
        assert g[0] == want[0] and np.array_equal(g[1], want[1])
        total += want[0]
    assert total > 300 and pkg.search_for_triangulation_batch([]) == []


@pytest.mark.gpu
def test_device_triangulation_edge_cases(pkg, port):
    rng = np.random.default_rng(9)
    pool = rng.integers(0, 256, (3, 32), dtype=np.uint8)
    k1, d1 = random_frame(rng, 200, 640, 480, pool); k2, d2 = random_frame(rng, 300, 640, 480, pool)
    s = port.tables()[0][0]; sig = (s * s).astype(np.float32) * np.float32(1e6)
    T1, T2 = poses(0.5)
    one1 = (np.array([5], np.uint32), np.array([0, 200], np.int32), np.arange(200, dtype=np.uint32))
    one2 = (np.array([5], np.uint32), np.array([0, 300], np.int32), np.arange(300, dtype=np.uint32))
    none = np.full(300, -1.0, np.float32)

    def run(side1, side2, F12=None, **kw):
        a, keep = pkg.triangulation_args(side1, side2, compute_f12(T1, T2, CAM) if F12 is None else F12, camera_center(T1), T2, CAM, s, sig, **kw)
        want = port.search_for_triangulation(a); got = pkg.search_for_triangulation(a)
        assert got[0] == want[0] and np.array_equal(got[1], want[1])
        return got
    z1, z2 = np.zeros(200, np.uint8), np.zeros(300, np.uint8)
    # one shared node: every feature against every feature, equal descriptors everywhere -> the last passing candidate wins
    nm, m = run((k1, d1, z1, none[:200]) + one1, (k2, d2, z2, none) + one2)
    assert nm > 50
    # disjoint vocabularies, every feature already holding a map point, a zero F12 (den == 0 rejects everything)
    other = (np.array([9], np.uint32),) + one2[1:]
    assert run((k1, d1, z1, none[:200]) + one1, (k2, d2, z2, none) + other)[0] == 0
    assert run((k1, d1, z1 + 1, none[:200]) + one1, (k2, d2, z2, none) + one2)[0] == 0
    assert run((k1, d1, z1, none[:200]) + one1, (k2, d2, z2 + 1, none) + one2)[0] == 0
    assert run((k1, d1, z1, none[:200]) + one1, (k2, d2, z2, none) + one2, F12=np.zeros((3, 3), np.float32))[0] == 0
    # bOnlyStereo with no stereo feature at all; empty keyframes
    assert run((k1, d1, z1, none[:200]) + one1, (k2, d2, z2, none) + one2, only_stereo=True)[0] == 0
    e = (np.zeros(0, np.uint32), np.zeros(1, np.int32), np.zeros(0, np.uint32))
    assert run((k1[:0], d1[:0], z1[:0], none[:0]) + e, (k2, d2, z2, none) + one2)[0] == 0
    # an octave outside the scale table is an argument error
    bad = k2.copy(); bad["octave"][3] = 8
    a, keep = pkg.triangulation_args((k1, d1, z1, none[:200]) + one1, (bad, d2, z2, none) + one2, compute_f12(T1, T2, CAM), camera_center(T1), T2, CAM, s, sig)
    with pytest.raises(pkg.OrbGpuError):
        pkg.search_for_triangulation(a)
