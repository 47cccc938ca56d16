"""The search inside ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:825-957, called from
LocalMapping::SearchInNeighbors): projection, KeyFrame::IsInImage / GetFeaturesInArea (src/KeyFrame.cc:569-613, int image
bounds), distance and viewing-angle tests, PredictScale, the chi-square gated best-descriptor scan.
CPU: the restatement (oracle/frame_port.cc:frameport_fuse_search) against the reference's own Fuse run one map point per
call on a KeyFrame built by the reference's constructor (oracle/_ref/libframeref.so); GPU: orbgpu_fuse_search bit-for-bit
against the restatement."""
import numpy as np
import pytest

import oracle_lib
from test_local_map import CAM, LOG12, local_map_scene, random_local_map_args
from test_search import SHIFT, device_extract, oracle_extract, two_frames

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")


def camera_center(T):
    R, t = T[:3, :3], T[:3, 3]
    return (-(R.T.astype(np.float32) @ t.astype(np.float32))).astype(np.float32)


def fuse_from_local_map(pkg, port, keep, th, bounds, sigma_scale=1.0, Ow=None, seed=0):
    """Re-uses a local-map scene (points, normals, distance bands, descriptors; the frame becomes the keyframe)."""
    rng = np.random.default_rng(seed)
    s = port.tables()[0][0]
    inv_sigma2 = (np.float32(1.0) / (s * s)).astype(np.float32) / np.float32(sigma_scale)
    n1 = len(keep["candidate"])
    return pkg.fuse_args(rng.random(n1) < 0.9, keep["world_pos"], keep["normal"], keep["max_distance"], keep["min_distance"], keep["mp_desc"],
                         keep["Tcw"], camera_center(keep["Tcw"]) if Ow is None else Ow, CAM, 40.0, keep["cur_keys_un"], keep["cur_desc"],
                         keep["cur_u_right"], bounds, s, inv_sigma2, LOG12, th)


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


# ------------------------------------------------------------------------------------------------------------- CPU side
def test_oracle_fuse_follows_the_geometry(pkg, port, scene):
    F1, F2, bounds = scene
    _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, 3)
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds)
    best = port.fuse_search(args)
    ok = np.nonzero(best >= 0)[0]
    assert len(ok) > 250
    k1, k2 = F1[0], F2[0]
    assert fkeep["present"][ok].all()
    dx = k2["x"][best[ok]] - k1["x"][ok]; dy = k2["y"][best[ok]] - k1["y"][ok]
    assert np.abs(dx - SHIFT[0]).max() < 8 and np.abs(dy - SHIFT[1]).max() < 8           # inside the chi-square gate
    assert (np.abs(dx - SHIFT[0]) < 2).mean() > 0.8


@needs_ref
@pytest.mark.parametrize("cfg", [(3.0, 1.0, (0.0, 640.0, 0.0, 480.0)), (3.0, 50.0, (-8.5, 647.25, -6.75, 489.5)), (2.5, 4.0, (3.9, 636.2, 2.6, 477.9))],
                         ids=["th3", "wide-gate-bounds-out", "bounds-in"])
def test_fuse_restatement_equals_reference(pkg, port, scene, cfg):
    th, sig, bounds = cfg
    F1, F2, _ = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    for seed, stereo in [(1, True), (2, False)]:
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed, stereo)
        args, fkeep = fuse_from_local_map(pkg, port, keep, th, bounds, sig)
        want, ow = ref.fuse_search(args)
        args, fkeep = fuse_from_local_map(pkg, port, keep, th, bounds, sig, Ow=ow)           # the centre the reference derived
        got = port.fuse_search(args)
        assert (want >= 0).sum() > 200
        assert np.array_equal(got, want)


@needs_ref
@pytest.mark.parametrize("seed", [51, 52, 53, 54, 55, 56])
def test_fuse_on_random_points_restatement_equals_reference(pkg, port, seed):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    _, keep = random_local_map_args(pkg, port, seed)
    bounds = (-8.5, 647.25, -6.0, 489.5)
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0, seed=seed)
    want, ow = ref.fuse_search(args)
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0, Ow=ow, seed=seed)
    got = port.fuse_search(args)
    assert (want >= 0).sum() > 20
    assert np.array_equal(got, want)


# ------------------------------------------------------------------------------------------------------------- GPU side
@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(3.0, 1.0, (0.0, 640.0, 0.0, 480.0)), (3.0, 50.0, (-8.5, 647.25, -6.75, 489.5)), (2.5, 4.0, (3.9, 636.2, 2.6, 477.9))],
                         ids=["th3", "wide-gate-bounds-out", "bounds-in"])
def test_device_fuse_equals_oracle(pkg, port, cfg):
    th, sig, bounds = cfg
    F1, F2, _ = two_frames(pkg, device_extract(pkg))
    for seed, stereo in [(1, True), (2, False)]:
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed, stereo)
        args, fkeep = fuse_from_local_map(pkg, port, keep, th, bounds, sig)
        want = port.fuse_search(args)
        got = pkg.fuse_search(args)
        assert (want >= 0).sum() > 200
        assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [51, 52, 53, 54, 55, 56, 57, 58])
def test_device_fuse_equals_oracle_on_random_points(pkg, port, seed):
    _, keep = random_local_map_args(pkg, port, seed)
    bounds = (-8.5, 647.25, -6.0, 489.5)
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0, seed=seed)
    want = port.fuse_search(args)
    got = pkg.fuse_search(args)
    assert np.array_equal(got, want)


@pytest.mark.gpu
def test_device_fuse_batch_equals_oracle(pkg, port):
    """Several target keyframes in one call, as SearchInNeighbors visits them (src/LocalMapping.cc:485-492)"""
    bounds = (-8.5, 647.25, -6.0, 489.5)
    jobs = []
    for seed in range(70, 80):
        _, keep = random_local_map_args(pkg, port, seed, n_cur=0 if seed == 73 else None)     # one keyframe without features
        jobs.append(fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0, seed=seed))
    jobs[5][1]["present"][:] = 0                                                               # one job without points to fuse
    got = pkg.fuse_search_batch([j[0] for j in jobs])
    assert len(got) == len(jobs) and pkg.fuse_search_batch([]) == []
    fused = 0
    for j, g in zip(jobs, got):
        want = port.fuse_search(j[0])
        assert np.array_equal(g, want)
        fused += int((want >= 0).sum())
    assert fused > 200


@pytest.mark.gpu
def test_device_fuse_edge_cases(pkg, port):
    _, keep = random_local_map_args(pkg, port, 59)
    bounds = (0.0, 640.0, 0.0, 480.0)
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0)
    fkeep["present"][:] = 0
    assert (pkg.fuse_search(args) == -1).all() and (port.fuse_search(args) == -1).all()
    _, keep = random_local_map_args(pkg, port, 60, n_cur=0)                                  # a keyframe without features
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0)
    assert (pkg.fuse_search(args) == -1).all()
    _, keep = random_local_map_args(pkg, port, 61)
    keep["cur_keys_un"]["octave"][5] = 8                                                     # outside the scale table
    args, fkeep = fuse_from_local_map(pkg, port, keep, 3.0, bounds, 30.0)
    with pytest.raises(pkg.OrbGpuError):
        pkg.fuse_search(args)
