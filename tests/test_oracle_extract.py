"""CPU (not gpu): the oracle port against (a) golden results of the UNMODIFIED reference extractor
(tests/golden/extract_golden.json, octree_golden.npz) and (b) the reference itself when
oracle/_ref/liborbref.so is present (it is built in the dev container and travels prebuilt)."""
import hashlib
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(__file__)
GOLD = json.load(open(os.path.join(HERE, "golden", "extract_golden.json")))
OCT = np.load(os.path.join(HERE, "golden", "octree_golden.npz"))


def frame_for(synth, case):
    fam, w, h, seed = case["family"], case["w"], case["h"], case["seed"]
    if fam == "synth": return synth.synth(w, h, seed)
    if fam == "flat": return synth.flat(w, h)
    if fam == "lowtex": return synth.lowtex(w, h, seed)
    return synth.noise(w, h, seed)


CPU_CASES = [c for c in GOLD["cases"] if c["w"] * c["h"] <= 752 * 480]


@pytest.mark.parametrize("case", CPU_CASES, ids=lambda c: "%s-%x" % (c["name"], c["seed"]))
def test_port_matches_reference_golden(pkg, oracle, case):
    img = frame_for(pkg.synth, case)
    assert hashlib.sha256(img.tobytes()).hexdigest() == case["image_sha256"]     # generator is deterministic
    kps, desc = oracle.OraclePort(case["nfeatures"], 1.2, 8, 20, 7).extract(img)
    assert len(kps) == case["n"]
    assert (np.bincount(kps["octave"], minlength=8).tolist() if len(kps) else [0] * 8) == case["per_level"]
    assert hashlib.sha256(kps.tobytes()).hexdigest() == case["kps_sha256"]
    assert hashlib.sha256(desc.tobytes()).hexdigest() == case["desc_sha256"]


def test_constructor_tables_match_reference_golden(oracle):
    s4, feats, umax = oracle.OraclePort(1000, 1.2, 8, 20, 7).tables()
    t = GOLD["tables_nf1000"]
    assert [float(x).hex() for x in s4[0]] == t["scale_hex"]
    assert [float(x).hex() for x in s4[1]] == t["inv_scale_hex"]
    assert [float(x).hex() for x in s4[2]] == t["sigma2_hex"]
    assert [float(x).hex() for x in s4[3]] == t["inv_sigma2_hex"]
    assert feats.tolist() == t["features_per_level"] == [217, 181, 151, 126, 105, 87, 73, 60]
    assert umax.tolist() == t["umax"] == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def test_port_octree_matches_reference_golden(oracle):
    port = oracle.OraclePort()
    n_cases = int(OCT["count"])
    for k in range(n_cases):
        xyr = OCT["oct_%d_in" % k].astype(np.int32); W, H, N = OCT["oct_%d_WHN" % k].tolist()
        ref = OCT["oct_%d_out" % k].astype(np.int32).reshape(-1, 3)
        idx = port.octree(xyr[:, 0], xyr[:, 1], xyr[:, 2], W, H, N)
        assert np.array_equal(xyr[idx].reshape(-1, 3), ref), (k, W, H, N, len(xyr))
    assert n_cases > 100


def test_port_matches_reference_live(pkg, oracle):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref/liborbref.so not present")
    synth = pkg.synth
    for (w, h, nf, seeds) in [(640, 480, 1000, range(0x5EED0100, 0x5EED0104)), (333, 217, 500, range(40, 43)), (752, 480, 1500, [77])]:
        port = oracle.OraclePort(nf, 1.2, 8, 20, 7); ref = oracle.OracleRef(nf, 1.2, 8, 20, 7)
        for seed in seeds:
            img = synth.synth(w, h, seed)
            kp, dp = port.extract(img); kr, dr = ref.extract(img, arena_mode=1)
            assert kp.tobytes() == kr.tobytes() and dp.tobytes() == dr.tobytes(), (w, h, seed)
            for l, (a, b) in enumerate(zip(port.pyramid(img), ref.pyramid(img))):
                assert np.array_equal(a, b), l


def test_port_other_parameters_live(pkg, oracle):
    if not oracle.have_ref():
        pytest.skip("oracle/_ref/liborbref.so not present")
    img = pkg.synth.synth(480, 360, 11)
    for (nf, sc, nl, ini, mn) in [(500, 1.2, 4, 20, 7), (800, 1.5, 5, 30, 10), (300, 1.1, 8, 12, 5), (50, 1.2, 8, 20, 7)]:
        kp, dp = oracle.OraclePort(nf, sc, nl, ini, mn).extract(img)
        kr, dr = oracle.OracleRef(nf, sc, nl, ini, mn).extract(img, arena_mode=1)
        assert kp.tobytes() == kr.tobytes() and dp.tobytes() == dr.tobytes(), (nf, sc, nl, ini, mn)


def test_hamming_port_equals_bit_count(oracle):
    rng = np.random.default_rng(0)
    port = oracle.OraclePort()
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert port.hamming(a, b) == int(np.unpackbits(a ^ b).sum())
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    assert port.hamming(z, z) == 0 and port.hamming(z, o) == 256
