"""GPU: Hamming kernels through the C ABI against the oracle's restatement of
ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663) and of the best/second-best scan (:201-226)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def correlated_descriptors(rng, K, n, pool=4096):
    """BASELINE config 5 generator shape: descriptors = pool word XOR sparse noise (p = 1/16)."""
    base = rng.integers(0, 256, (pool, 32), dtype=np.uint8)
    idx = (977 * np.arange(K)[:, None] + 31 * np.arange(n)[None, :]) % pool
    noise = rng.integers(0, 256, (4, K, n, 32), dtype=np.uint8)
    return base[idx] ^ (noise[0] & noise[1] & noise[2] & noise[3])


def test_bruteforce_matches_oracle(pkg, oracle):
    rng = np.random.default_rng(1)
    m = pkg.ORBmatcher(0.6, True); po = oracle.OraclePort()
    for (nA, nB) in [(1, 1), (7, 300), (1000, 1000), (129, 127), (500, 0)]:
        A = rng.integers(0, 256, (nA, 32), dtype=np.uint8); B = rng.integers(0, 256, (max(nB, 1), 32), dtype=np.uint8)[:nB]
        if nB == 0:
            bi, b1, b2 = m.match_bruteforce(A, np.zeros((0, 32), np.uint8))
            assert (bi == -1).all() and (b1 == 256).all() and (b2 == 256).all()
            continue
        got = m.match_bruteforce(A, B); want = po.match_bruteforce(A, B)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)


def test_ties_keep_lowest_index(pkg, oracle):
    rng = np.random.default_rng(2)
    B = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    B = np.concatenate([B, B, B])                      # every candidate appears three times
    A = B[:64].copy(); A[:, 0] ^= 1                    # distance 1 to three candidates each
    bi, b1, b2 = pkg.ORBmatcher().match_bruteforce(A, B)
    assert np.array_equal(bi, np.arange(64)) and (b1 == 1).all() and (b2 == 1).all()
    want = oracle.OraclePort().match_bruteforce(A, B)
    assert np.array_equal(bi, want[0]) and np.array_equal(b2, want[2])


def test_real_descriptors_between_consecutive_frames(pkg, oracle):
    synth = pkg.synth
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    a = synth.synth(640, 480, 4242)
    _, d1 = ex(a)
    _, d2 = ex(np.roll(a, 3, axis=1))
    got = pkg.ORBmatcher().match_bruteforce(d1, d2); want = oracle.OraclePort().match_bruteforce(d1, d2)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    assert (got[1] <= 50).sum() > 200                    # the shifted frame really does match


def test_allpairs_counts(pkg, oracle):
    rng = np.random.default_rng(3)
    K, n = 12, 250
    desc = correlated_descriptors(rng, K, n)
    m = pkg.ORBmatcher(0.6, True)
    m.set_keyframes(desc)
    a, b = np.triu_indices(K, 1)
    got = m.count_pairs(a, b)
    po = oracle.OraclePort()
    want = np.array([po.match_count(desc[i], desc[j], 50, 0.6) for i, j in zip(a, b)], np.int32)
    assert np.array_equal(got, want)
    assert want.max() > 0
    # asymmetric: (b, a) is a different scan
    got_ba = m.count_pairs(b, a)
    want_ba = np.array([po.match_count(desc[j], desc[i], 50, 0.6) for i, j in zip(a, b)], np.int32)
    assert np.array_equal(got_ba, want_ba)
    assert m.last_kernel_ms() > 0


def test_allpairs_properties_at_full_keyframe_size(pkg):
    """1000 descriptors per keyframe (BASELINE config 5 size): a keyframe against itself has best = 0 for
    every query; against its bit-complement nothing is accepted."""
    rng = np.random.default_rng(4)
    kf = rng.integers(0, 256, (1000, 32), dtype=np.uint8)
    desc = np.stack([kf, kf ^ 0xFF, kf[::-1].copy()])
    m = pkg.ORBmatcher(0.6, True)
    m.set_keyframes(desc)
    got = m.count_pairs([0, 0, 0, 2], [0, 1, 2, 0])
    assert got[0] == 1000 and got[1] == 0 and got[2] == 1000 and got[3] == 1000
    with pytest.raises(pkg.OrbGpuError):
        m.count_pairs([0], [3])                         # keyframe index out of range
