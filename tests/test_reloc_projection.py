"""ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const set<MapPoint*>& sAlreadyFound, th, ORBdist)
(src/ORBmatcher.cc:1472-1597), the guided search of Tracking::Relocalization.
CPU: the restatement (oracle/frame_port.cc:frameport_search_by_projection_keyframe) against the reference's own code on a
KeyFrame / Frame pair built by the reference's constructors (oracle/_ref/libframeref.so); GPU:
orbgpu_search_by_projection_keyframe bit-for-bit against the restatement."""
import numpy as np
import pytest

import oracle_lib
from test_local_map import CAM, LOG12, local_map_scene, random_local_map_args
from test_search import SHIFT, device_extract, oracle_extract, two_frames

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")


def reloc_from_local_map(pkg, port, keep, angles, th, orb_dist, bounds, ori=True, seed=0):
    """Re-uses a local-map scene: its points become the keyframe's map points (one per keyframe feature)."""
    rng = np.random.default_rng(seed)
    n1 = len(keep["candidate"]); n2 = len(keep["cur_keys_un"])
    ang = angles if angles is not None else rng.uniform(0, 360, n1).astype(np.float32)
    return pkg.reloc_projection_args(rng.random(n1) < 0.8, keep["world_pos"], keep["max_distance"], keep["min_distance"], keep["mp_desc"], ang,
                                     keep["cur_keys_un"], keep["cur_desc"], rng.random(n2) < 0.1, keep["Tcw"], CAM, bounds,
                                     port.tables()[0][0], LOG12, th, orb_dist, ori)


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


@pytest.fixture(scope="module")
def scene(pkg, oracle):
    return two_frames(pkg, oracle_extract(oracle))


def test_oracle_reloc_matches_follow_the_shift(pkg, port, scene):
    F1, F2, bounds = scene
    _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, 3)
    args, rkeep = reloc_from_local_map(pkg, port, keep, F1[0]["angle"], 10.0, 100, bounds)
    nm, cm = port.search_by_projection_keyframe(args)
    got = np.nonzero(cm >= 0)[0]
    assert nm == len(got) > 300 and len(set(cm[got].tolist())) == len(got)
    k1, k2 = F1[0], F2[0]
    dx = k2["x"][got] - k1["x"][cm[got]]; dy = k2["y"][got] - k1["y"][cm[got]]
    assert (np.abs(dx - SHIFT[0]) < 3).mean() > 0.8 and (np.abs(dy - SHIFT[1]) < 3).mean() > 0.8
    assert rkeep["mp_valid"][cm[got]].all() and not rkeep["cur_occupied"][got].any()


@needs_ref
@pytest.mark.parametrize("cfg", [(10.0, 100, True), (3.0, 64, False), (10.0, 100, False)], ids=["th10-ori", "th3-64", "th10"])
def test_reloc_projection_restatement_equals_reference(pkg, port, scene, cfg):
    th, orb_dist, ori = cfg
    F1, F2, bounds = scene
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    for seed in (1, 2):
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, rkeep = reloc_from_local_map(pkg, port, keep, F1[0]["angle"], th, orb_dist, bounds, ori, seed)
        wn, wm = ref.search_by_projection_keyframe(args)
        gn, gm = port.search_by_projection_keyframe(args)
        assert wn > 150
        assert gn == wn and np.array_equal(np.where(gm == -2, -1, gm), wm)


@needs_ref
@pytest.mark.parametrize("seed", [71, 72, 73, 74, 75, 76])
def test_reloc_projection_on_random_points_restatement_equals_reference(pkg, port, seed):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    _, keep = random_local_map_args(pkg, port, seed)
    bounds = (-8.5, 647.25, -6.0, 489.5)
    args, rkeep = reloc_from_local_map(pkg, port, keep, None, 10.0 if seed % 2 else 3.0, 100 if seed % 3 else 64, bounds, bool(seed % 2), seed)
    wn, wm = ref.search_by_projection_keyframe(args)
    gn, gm = port.search_by_projection_keyframe(args)
    assert gn == wn and np.array_equal(np.where(gm == -2, -1, gm), wm)


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(10.0, 100, True), (3.0, 64, False), (10.0, 100, False)], ids=["th10-ori", "th3-64", "th10"])
def test_device_reloc_projection_equals_oracle(pkg, port, cfg):
    th, orb_dist, ori = cfg
    F1, F2, bounds = two_frames(pkg, device_extract(pkg))
    for seed in (1, 2):
        _, keep = local_map_scene(pkg, port, F1, F2, bounds, 1.0, seed)
        args, rkeep = reloc_from_local_map(pkg, port, keep, F1[0]["angle"], th, orb_dist, bounds, ori, seed)
        want = port.search_by_projection_keyframe(args)
        got = pkg.search_by_projection_keyframe(args)
        assert want[0] > 150
        assert got[0] == want[0] and np.array_equal(got[1], want[1])


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [71, 72, 73, 74, 75, 76, 77, 78])
def test_device_reloc_projection_equals_oracle_on_random_points(pkg, port, seed):
    _, keep = random_local_map_args(pkg, port, seed)
    bounds = (-8.5, 647.25, -6.0, 489.5)
    args, rkeep = reloc_from_local_map(pkg, port, keep, None, 10.0 if seed % 2 else 3.0, 100 if seed % 3 else 64, bounds, bool(seed % 2), seed)
    want = port.search_by_projection_keyframe(args)
    got = pkg.search_by_projection_keyframe(args)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])
