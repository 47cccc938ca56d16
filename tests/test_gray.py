"""Colour -> gray conversion ahead of the extractor (reference src/Tracking.cc:244-257, cv::cvtColor)."""
import os

import numpy as np
import pytest

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gray_golden.npz"))


def gray_oracle(img, rgb_order):
    """OpenCV 4.13 8-bit RGB2Gray fixed point: (c0*k0 + c1*19235 + c2*k2 + 2^14) >> 15, (k0,k2) = (B,R) weights."""
    k0, k2 = (9798, 3735) if rgb_order else (3735, 9798)
    c = img.astype(np.int64)
    return ((c[..., 0] * k0 + c[..., 1] * 19235 + c[..., 2] * k2 + (1 << 14)) >> 15).astype(np.uint8)


@pytest.mark.parametrize("src,key,rgb", [("c3", "c3_bgr", False), ("c3", "c3_rgb", True), ("c4", "c4_bgra", False),
                                         ("c4", "c4_rgba", True), ("ramp", "ramp_bgr", False), ("ramp", "ramp_rgb", True)])
def test_gray_oracle_matches_cv2_golden(src, key, rgb):
    assert np.array_equal(gray_oracle(GOLD[src], rgb), GOLD[key])


@pytest.mark.gpu
@pytest.mark.parametrize("channels,rgb", [(3, False), (3, True), (4, False), (4, True)])
def test_colour_input_matches_gray_pipeline(pkg, oracle, channels, rgb):
    synth = pkg.synth
    w, h = 500, 300                                              # width not a multiple of 16: exercises the pitched layout too
    planes = [synth.synth(w, h, 4242 + i) for i in range(3)] + [synth.noise(w, h, 5)]
    img = np.ascontiguousarray(np.stack(planes[:channels], 2))
    gray = gray_oracle(img, rgb)
    ex = pkg.ORBextractor(800, 1.2, 8, 20, 7)
    ex.set_input_format(channels, rgb)
    kps, desc = ex(img)
    assert np.array_equal(ex.pyramid_level(0), gray)             # mImGray
    okps, odesc = oracle.OraclePort(800, 1.2, 8, 20, 7).extract(gray)
    assert len(kps) > 300 and kps.tobytes() == okps.tobytes() and np.array_equal(desc, odesc)
    # strided rows (padded colour rows) and back to gray on the same handle
    pad = np.zeros((h, w + 9, channels), np.uint8); pad[:, :w] = img
    k2, d2 = ex(pad[:, :w])
    assert k2.tobytes() == okps.tobytes() and np.array_equal(d2, odesc)
    ex.set_input_format(1)
    k3, d3 = ex(gray)
    assert k3.tobytes() == okps.tobytes() and np.array_equal(d3, odesc)


@pytest.mark.gpu
def test_cv2_golden_through_the_device_conversion(pkg):
    ex = pkg.ORBextractor(100, 1.2, 2, 20, 7)
    for src, key, ch, rgb in [("c3", "c3_bgr", 3, False), ("c3", "c3_rgb", 3, True), ("c4", "c4_bgra", 4, False),
                              ("c4", "c4_rgba", 4, True), ("ramp", "ramp_bgr", 3, False), ("ramp", "ramp_rgb", 3, True)]:
        ex.set_input_format(ch, rgb)
        ex(np.ascontiguousarray(GOLD[src]))
        assert np.array_equal(ex.pyramid_level(0), GOLD[key]), key
