"""Frame::ComputeStereoMatches (reference src/Frame.cc:466-640), SURVEY.md section 8(f) first "next" row.
The oracle restatement is pinned bit-for-bit against the reference's own Frame::ComputeStereoMatches in
tests/test_ref_frame.py (oracle/_ref/libframeref.so); the CPU test here checks it against the geometry of a synthetic
rectified pair, the GPU test checks the device kernel bit-for-bit against the restatement."""
import numpy as np
import pytest


def _oracle_pair(pkg, oracle, w, h, seed, nf, disp):
    left, right = pkg.synth.stereo_pair(w, h, seed, disp)
    pl = oracle.OraclePort(nf, 1.2, 8, 20, 7); pr = oracle.OraclePort(nf, 1.2, 8, 20, 7)
    kL, dL = pl.extract(left); kR, dR = pr.extract(right)
    lvlL = [pl.level_pixels(l) for l in range(8)]; lvlR = [pr.level_pixels(l) for l in range(8)]
    return left, right, pl, (kL, dL, kR, dR, lvlL, lvlR)


def test_oracle_recovers_band_disparities(pkg, oracle):
    w, h, disp = 640, 480, (6, 14, 27)
    bf, b = 386.1448, 0.5372                               # KITTI 00-02 (Examples/Stereo/KITTI00-02.yaml: Camera.bf; b = bf / fx)
    left, right, po, args = _oracle_pair(pkg, oracle, w, h, 0x57E2E001, 1000, disp)
    u, d = po.stereo_matches(*args, bf, b)
    kL = args[0]
    ok = u >= 0
    assert ok.sum() > 0.4 * len(kL)                        # most keypoints of a clean synthetic pair find their match
    assert ((u < 0) == (d < 0)).all()
    band = np.minimum((kL["y"] * len(disp) / h).astype(int), len(disp) - 1)
    true_d = np.array(disp, np.float32)[band]
    meas = kL["x"] - u
    inside = ok & (np.abs((kL["y"] * len(disp) / h) % 1.0 - 0.5) < 0.4)     # away from the band seams
    err = np.abs(meas - true_d)[inside]
    assert np.median(err) < 0.35 and (err < 1.5).mean() > 0.9
    assert np.allclose(d[ok], np.float32(bf) / meas[ok], rtol=1e-6)
    assert (meas[ok] < np.float32(bf) / np.float32(b)).all() and (meas[ok] > 0).all()


def test_oracle_empty_inputs(pkg, oracle):
    po = oracle.OraclePort(1000, 1.2, 8, 20, 7)
    img = pkg.synth.flat(320, 240)
    k, d = po.extract(img)
    lv = [po.level_pixels(l) for l in range(8)]
    u, dep = po.stereo_matches(k, d, k, d, lv, lv, 100.0, 0.5)
    assert len(u) == 0 and len(dep) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(640, 480, 1000, (6, 14, 27)), (1241, 376, 2000, (3, 40)), (333, 217, 700, (0, 9))])
def test_device_stereo_matches_equal_oracle(pkg, oracle, shape):
    w, h, nf, disp = shape
    bf, b = 386.1448, 0.5372
    left, right, po, args = _oracle_pair(pkg, oracle, w, h, 0x57E2E002 + w, nf, disp)
    exL = pkg.ORBextractor(nf, 1.2, 8, 20, 7); exR = pkg.ORBextractor(nf, 1.2, 8, 20, 7)
    kL, dL = exL(left); kR, dR = exR(right)
    assert kL.tobytes() == args[0].tobytes() and kR.tobytes() == args[2].tobytes()
    u, d = pkg.stereo_match(exL, exR, bf, b)
    ou, od = po.stereo_matches(*args, bf, b)
    assert len(u) == len(kL)
    assert (u >= 0).sum() > 0.25 * len(kL)
    assert u.tobytes() == ou.tobytes() and d.tobytes() == od.tobytes()
    # a narrow baseline (small bf/b) exercises the uR window and the disparity < maxD rejection
    u2, d2 = pkg.stereo_match(exL, exR, 10.0, 0.8)
    ou2, od2 = po.stereo_matches(*args, 10.0, 0.8)
    assert u2.tobytes() == ou2.tobytes() and d2.tobytes() == od2.tobytes()


@pytest.mark.gpu
def test_device_stereo_without_frames_is_a_state_error(pkg):
    exL = pkg.ORBextractor(500, 1.2, 8, 20, 7); exR = pkg.ORBextractor(500, 1.2, 8, 20, 7)
    with pytest.raises(RuntimeError):
        pkg.stereo_match(exL, exR, 100.0, 0.5)
