"""MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-303), batched over map points: the N x N table of
ORBmatcher::DescriptorDistance (:281) of a map point's observed descriptors, the median of every row, the first row with
the smallest median.  CPU: the restatement against the reference's own MapPoint code (oracle/_ref/libframeref.so) and a
numpy evaluation; GPU: orbgpu_distinctive_descriptors against the restatement."""
import numpy as np
import pytest

import oracle_lib

needs_ref = pytest.mark.skipif(not oracle_lib.have_frame_ref(), reason="oracle/_ref/libframeref.so not built")


def groups(seed, n_groups, nmax, distinct=False):
    """Observation sets of map points: a base descriptor seen with a few flipped bits (real observations of one point),
    some exact repeats (equal rows -> equal medians -> the first wins), some unrelated outliers, empty groups."""
    rng = np.random.default_rng(seed)
    sizes = rng.integers(0, nmax + 1, n_groups)
    sizes[rng.random(n_groups) < 0.1] = 1
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    desc = np.zeros((int(start[-1]), 32), np.uint8)
    for g in range(n_groups):
        base = rng.integers(0, 256, 32, dtype=np.uint8)
        for j in range(start[g], start[g + 1]):
            d = base.copy()
            kind = rng.random()
            if kind < 0.15:
                d = rng.integers(0, 256, 32, dtype=np.uint8)                 # an outlier observation
            elif kind < 0.9 or distinct:
                for _ in range(int(rng.integers(1 if distinct else 0, 40))):
                    d[int(rng.integers(32))] ^= np.uint8(1 << int(rng.integers(8)))
            desc[j] = d
        if distinct:                                                          # make every row of the group unique
            for j in range(start[g], start[g + 1]):
                desc[j, 31] = (j - start[g]) & 0xFF; desc[j, 30] = (j - start[g]) >> 8
    return desc, start


def numpy_best(desc, start):
    out = []
    for g in range(len(start) - 1):
        d = desc[start[g]:start[g + 1]]
        if len(d) == 0:
            out.append(-1); continue
        D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(2)
        med = np.sort(D, axis=1)[:, int(0.5 * (len(d) - 1))]
        out.append(int(np.argmin(med)))                                       # argmin = the first minimum
    return np.array(out, np.int32)


@pytest.fixture(scope="module")
def port():
    return oracle_lib.OraclePort(1000, 1.2, 8, 20, 7)


def test_restatement_equals_numpy(port):
    for seed, ng, nmax in [(1, 300, 12), (2, 60, 70), (3, 5, 300)]:
        desc, start = groups(seed, ng, nmax)
        assert np.array_equal(port.distinctive_descriptors(desc, start), numpy_best(desc, start))


@needs_ref
def test_restatement_equals_reference(port):
    ref = oracle_lib.OracleFrameRef((port.tables()[0][0], port.tables()[0][1]))
    for seed, ng, nmax, distinct in [(4, 200, 12, True), (5, 40, 70, True), (6, 200, 15, False)]:
        desc, start = groups(seed, ng, nmax, distinct)
        want = ref.distinctive_descriptors(desc, start)
        got = port.distinctive_descriptors(desc, start)
        # a row equal to an earlier row has the same median, so the first of equal rows is also the reference's BestIdx
        assert np.array_equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(1, 2000, 12), (2, 300, 70), (3, 9, 600), (7, 1, 1), (8, 5000, 6)])
def test_device_distinctive_descriptors_equal_oracle(pkg, port, cfg):
    seed, ng, nmax = cfg
    desc, start = groups(seed, ng, nmax)
    assert np.array_equal(pkg.distinctive_descriptors(desc, start), port.distinctive_descriptors(desc, start))


@pytest.mark.gpu
def test_device_distinctive_descriptors_edge_cases(pkg, port):
    assert len(pkg.distinctive_descriptors(np.zeros((0, 32), np.uint8), [0])) == 0
    assert pkg.distinctive_descriptors(np.zeros((0, 32), np.uint8), [0, 0, 0]).tolist() == [-1, -1]
    d = np.zeros((4, 32), np.uint8); d[1] = 255; d[3] = 255                     # two pairs of equal rows: the first row wins
    assert pkg.distinctive_descriptors(d, [0, 4]).tolist() == port.distinctive_descriptors(d, [0, 4]).tolist() == [0]
    with pytest.raises(pkg.OrbGpuError):
        pkg.distinctive_descriptors(d, [0, 3, 2])                               # decreasing offsets
