"""GPU: the octree kernel on arbitrary candidate lists (C ABI orbgpu_octree) against the golden
results of the reference's DistributeOctTree and against the oracle port on larger random sets."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
OCT = np.load(os.path.join(os.path.dirname(__file__), "golden", "octree_golden.npz"))


@pytest.fixture(scope="module")
def ex(pkg):
    return pkg.ORBextractor(1000, 1.2, 8, 20, 7)


def test_octree_matches_reference_golden(ex):
    for k in range(int(OCT["count"])):
        xyr = OCT["oct_%d_in" % k].astype(np.int32); W, H, N = OCT["oct_%d_WHN" % k].tolist()
        ref = OCT["oct_%d_out" % k].astype(np.int32).reshape(-1, 3)
        idx = ex.octree(xyr[:, 0], xyr[:, 1], xyr[:, 2], W, H, N)
        assert np.array_equal(xyr[idx].reshape(-1, 3), ref), (k, W, H, N, len(xyr))


@pytest.mark.parametrize("n,N,W,H", [(5000, 217, 608, 448), (20000, 869, 1888, 1048), (60000, 300, 1209, 344),
                                       (3000, 2500, 608, 448), (1500, 4000, 608, 448), (9, 4, 4000, 40), (12000, 1200, 700, 700)])
def test_octree_matches_port_large(ex, oracle, n, N, W, H):
    """Sizes that spill keys and/or node tables from shared memory to the global scratch area."""
    rng = np.random.default_rng(n + N)
    flat = rng.choice(W * H, size=min(n, W * H), replace=False)
    xs = (flat % W).astype(np.int32); ys = (flat // W).astype(np.int32)
    resp = rng.integers(0, 256, len(flat)).astype(np.int32)
    got = ex.octree(xs, ys, resp, W, H, N)
    want = oracle.OraclePort().octree(xs, ys, resp, W, H, N)
    assert np.array_equal(got, want)


def test_octree_equal_responses_pick_first_in_input_order(ex, oracle):
    rng = np.random.default_rng(5)
    W, H = 300, 200
    flat = rng.choice(W * H, size=4000, replace=False)
    xs = (flat % W).astype(np.int32); ys = (flat // W).astype(np.int32)
    resp = np.full(len(flat), 40, np.int32)
    for N in (10, 100, 1000):
        assert np.array_equal(ex.octree(xs, ys, resp, W, H, N), oracle.OraclePort().octree(xs, ys, resp, W, H, N))
