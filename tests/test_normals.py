"""MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371), batched over map points: mean viewing direction, and the
scale-invariance distance band from the reference keyframe.
CPU: the restatement (oracle/frame_port.cc:frameport_update_normal_and_depth) against the reference's own MapPoint.cc on
KeyFrames built by its constructor (oracle/_ref/libframeref.so), whose cv::Mat expressions run through the shim pinned to
cv2 4.13.0 (tests/golden/scale_add_golden.npz, scale_golden.npz); GPU: orbgpu_update_normals_and_depths bit-for-bit against
the restatement."""
import numpy as np
import pytest

import oracle_lib

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")
SCALES = (np.float32(1.2) ** np.arange(8)).astype(np.float32)


def scene(seed, n_points, max_obs=40):
    """Points in front of a camera trajectory; 0..max_obs observing keyframes each (some points without any); a few
    observations from the point's own position (norm 0 -> inf / NaN, which must come out the same)."""
    rng = np.random.default_rng(seed)
    pos = rng.normal(0, 4, (n_points, 3)).astype(np.float32) + np.float32([0, 0, 12])
    nobs = rng.integers(0, max_obs + 1, n_points)
    nobs[rng.random(n_points) < 0.05] = 0
    start = np.zeros(n_points + 1, np.int32); start[1:] = np.cumsum(nobs)
    centers = (rng.normal(0, 2.5, (int(start[-1]), 3)) + rng.choice([0.0, 30.0], (int(start[-1]), 1))).astype(np.float32)
    for g in rng.choice(n_points, min(3, n_points), replace=False):
        if nobs[g] > 1:
            centers[start[g] + 1] = pos[g]
    ref_slot = np.array([int(rng.integers(0, max(n, 1))) for n in nobs], np.int32)
    ref_level = rng.integers(0, 8, n_points).astype(np.int32)
    ref_center = np.zeros((n_points, 3), np.float32)
    has = nobs > 0
    ref_center[has] = centers[start[:-1][has] + ref_slot[has]]
    normal0 = rng.normal(0, 1, (n_points, 3)).astype(np.float32)
    max0 = rng.uniform(1, 50, n_points).astype(np.float32); min0 = rng.uniform(0.1, 1, n_points).astype(np.float32)
    return dict(pos=pos, start=start, centers=centers, ref_slot=ref_slot, ref_level=ref_level, ref_center=ref_center,
                ref_scale=SCALES[ref_level], last_scale=np.full(n_points, SCALES[-1], np.float32), normal=normal0, maxd=max0, mind=min0)


def run_port(s):
    return oracle_lib.port_update_normal_and_depth(s["pos"], s["start"], s["centers"], s["ref_center"], s["ref_scale"], s["last_scale"],
                                                   s["normal"], s["maxd"], s["mind"])


def same(a, b):
    for x, y, nm in zip(a, b, ("normal", "max", "min")):
        xb = np.ascontiguousarray(x, np.float32).reshape(-1).view(np.uint32)
        yb = np.ascontiguousarray(y, np.float32).reshape(-1).view(np.uint32)
        nan = np.isnan(x.reshape(-1)) & np.isnan(y.reshape(-1))          # x86 and CUDA differ in the NaN payload only
        bad = np.nonzero((xb != yb) & ~nan)[0]
        assert len(bad) == 0, (nm, bad[:5], x.reshape(-1)[bad[:5]], y.reshape(-1)[bad[:5]])
    return True


def test_port_normals_follow_the_geometry():
    s = scene(1, 500)
    nv, mx, mn = run_port(s)
    nobs = np.diff(s["start"])
    keep = nobs == 0
    assert keep.sum() > 5 and np.array_equal(nv[keep], s["normal"][keep]) and np.array_equal(mx[keep], s["maxd"][keep])
    ok = (nobs > 0) & np.isfinite(nv).all(1)
    g = int(np.nonzero(ok & (nobs > 3))[0][0])
    d = s["pos"][g].astype(np.float64) - s["centers"][s["start"][g]:s["start"][g + 1]].astype(np.float64)
    want = (d / np.linalg.norm(d, axis=1)[:, None]).mean(0)
    assert np.allclose(nv[g], want, atol=1e-5)
    dist = np.linalg.norm(s["pos"][g].astype(np.float64) - s["ref_center"][g])
    assert abs(mx[g] - dist * s["ref_scale"][g]) < 1e-3 * dist and abs(mn[g] - mx[g] / SCALES[-1]) < 1e-5 * mx[g] + 1e-6
    assert (np.linalg.norm(nv[ok], axis=1) <= 1.0 + 1e-5).all()


@needs_ref
@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_normals_restatement_equals_reference(seed):
    s = scene(seed, 400)
    ref = oracle_lib.OracleFrameRef()
    want = ref.update_normal_and_depth(s["pos"], s["start"], s["centers"], s["ref_slot"], s["ref_level"], SCALES, s["normal"], s["maxd"], s["mind"])
    got = run_port(s)
    assert same(got, want)
    for a, b in zip(got, want):                                       # on one machine even the NaNs are the same bits
        assert a.tobytes() == b.tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(1, 400, 40), (2, 5000, 12), (3, 1, 3), (4, 20000, 6), (5, 300, 400)])
def test_device_normals_equal_oracle(pkg, cfg):
    seed, n, max_obs = cfg
    s = scene(seed, n, max_obs)
    want = run_port(s)
    got = pkg.update_normals_and_depths(s["pos"], s["start"], s["centers"], s["ref_center"], s["ref_scale"], s["last_scale"],
                                        s["normal"], s["maxd"], s["mind"])
    assert same(got, want)


@pytest.mark.gpu
def test_device_normals_edge_cases(pkg):
    z3 = np.zeros((0, 3), np.float32); z = np.zeros(0, np.float32)
    out = pkg.update_normals_and_depths(z3, np.zeros(1, np.int32), z3, z3, z, z, z3, z, z)
    assert len(out[0]) == 0
    s = scene(9, 50)
    s["start"][:] = 0                                                 # nobody observed: everything stays
    got = pkg.update_normals_and_depths(s["pos"], s["start"], np.zeros((0, 3), np.float32), s["ref_center"], s["ref_scale"],
                                        s["last_scale"], s["normal"], s["maxd"], s["mind"])
    assert np.array_equal(got[0], s["normal"]) and np.array_equal(got[1], s["maxd"]) and np.array_equal(got[2], s["mind"])
    bad = np.array([0, 3, 2], np.int32)
    with pytest.raises(pkg.OrbGpuError):
        pkg.update_normals_and_depths(s["pos"][:2], bad, np.zeros((3, 3), np.float32), s["ref_center"][:2], s["ref_scale"][:2],
                                      s["last_scale"][:2], s["normal"][:2], s["maxd"][:2], s["mind"][:2])
