"""GPU: orbgpu_match_bruteforce_noalloc (the allocation-free variant with a fixed __device__ scratch array; no
counterpart in the reference) against orbgpu_match_bruteforce through the
C ABI.  Integer results: the bar is bit-exact, as in test_gpu_match.py.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX = 4096          # ORBGPU_NOALLOC_MAX_DESCRIPTORS in include/orbgpu.h


def _call(L, name, A, B):
    A = np.ascontiguousarray(A, np.uint8); B = np.ascontiguousarray(B, np.uint8)
    bi = np.full(len(A), -7, np.int32); b1 = np.full(len(A), 999, np.uint16); b2 = np.full(len(A), 999, np.uint16)
    rc = getattr(L, name)(0, A.ctypes.data, len(A), B.ctypes.data if len(B) else A.ctypes.data, len(B),
                          bi.ctypes.data, b1.ctypes.data, b2.ctypes.data)
    return rc, bi, b1, b2


def test_noalloc_equals_existing_entry_on_consecutive_calls(pkg):
    L = pkg.lib()
    rng = np.random.default_rng(77)
    # consecutive calls share the scratch array: big call first, then smaller ones that must not see its leftovers,
    # then the maximum size, an empty candidate set and the first shape again
    shapes = [(1000, 1000), (37, 129), (1, 1), (MAX, MAX), (500, 0), (1000, 1000)]
    pool = rng.integers(0, 256, (512, 32), dtype=np.uint8)
    for k, (nA, nB) in enumerate(shapes):
        A = pool[rng.integers(0, 512, nA)] ^ (rng.integers(0, 256, (nA, 32), dtype=np.uint8) & rng.integers(0, 256, (nA, 32), dtype=np.uint8) & rng.integers(0, 256, (nA, 32), dtype=np.uint8))
        B = pool[rng.integers(0, 512, nB)] if nB else np.zeros((0, 32), np.uint8)     # many exact ties among candidates
        rc0, *want = _call(L, "orbgpu_match_bruteforce", A, B)
        rc1, *got = _call(L, "orbgpu_match_bruteforce_noalloc", A, B)
        assert rc0 == 0 and rc1 == 0, (k, rc0, rc1, L.orbgpu_last_cuda_error())
        for g, w in zip(got, want):
            assert np.array_equal(g, w), "call %d (%d x %d) differs from orbgpu_match_bruteforce" % (k, nA, nB)
        rc2, *again = _call(L, "orbgpu_match_bruteforce_noalloc", A, B)               # same inputs, second call in a row
        assert rc2 == 0
        for g, w in zip(again, want):
            assert np.array_equal(g, w)
        if nB == 0:
            assert (got[0] == -1).all() and (got[1] == 256).all() and (got[2] == 256).all()


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(pkg):
    L = pkg.lib()
    A = np.zeros((MAX + 1, 32), np.uint8); B = np.zeros((8, 32), np.uint8)
    rc, bi, b1, b2 = _call(L, "orbgpu_match_bruteforce_noalloc", A, B)
    assert rc == -2 and (bi == -7).all() and (b1 == 999).all()                       # ORBGPU_ERR_CAPACITY, outputs untouched
    assert b"4096" in L.orbgpu_last_cuda_error()
    rc, bi, _, _ = _call(L, "orbgpu_match_bruteforce_noalloc", B, A)
    assert rc == -2 and (bi == -7).all()
    assert L.orbgpu_match_bruteforce_noalloc(0, None, 1, B.ctypes.data, 8, bi.ctypes.data, b1.ctypes.data, b2.ctypes.data) == -1
    # a valid call right after the rejected ones still works
    rc, bi, b1, b2 = _call(L, "orbgpu_match_bruteforce_noalloc", B, B)
    assert rc == 0 and (bi == 0).all() and (b1 == 0).all() and (b2 == 0).all()
