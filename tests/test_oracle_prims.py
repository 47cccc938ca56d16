"""CPU (not gpu): the oracle's primitives against golden vectors produced by cv2 4.13.0 / glibc
(tests/golden/make_golden.py).  This is what pins the oracle's pixel arithmetic."""
import ctypes as C
import os

import numpy as np
import pytest

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "prims_golden.npz"))


@pytest.fixture(scope="module")
def L(oracle):
    lib = oracle.OraclePort().L
    lib.orbo_resize_linear_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_int, C.c_int, C.c_size_t]
    lib.orbo_gauss7_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_size_t]
    lib.orbo_fast9_16.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    lib.orbo_fast_atan2.restype = C.c_float
    lib.orbo_fast_atan2.argtypes = [C.c_float, C.c_float]
    lib.orbo_sincosf.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.orbo_sincos_checksum.restype = C.c_uint64
    lib.orbo_sincos_checksum.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
    lib.orbo_border_reflect101_u8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int]
    return lib


def test_resize_matches_cv2(L):
    n = 0
    for key in GOLD.files:
        if key.startswith("resize_") and key.endswith("_src"):
            src = np.ascontiguousarray(GOLD[key]); ref = GOLD[key.replace("_src", "_dst")]
            out = np.zeros_like(ref)
            L.orbo_resize_linear_u8(src.ctypes.data, src.shape[1], src.shape[0], src.shape[1], out.ctypes.data, ref.shape[1], ref.shape[0], ref.shape[1])
            assert np.array_equal(out, ref), key
            n += 1
    assert n == 10


def test_blur_matches_cv2(L):
    n = 0
    for key in GOLD.files:
        if key.startswith("blur_") and key.endswith("_src"):
            src = np.ascontiguousarray(GOLD[key]); ref = GOLD[key.replace("_src", "_dst")]
            out = np.zeros_like(ref)
            L.orbo_gauss7_u8(src.ctypes.data, src.shape[1], src.shape[0], src.shape[1], out.ctypes.data, src.shape[1])
            assert np.array_equal(out, ref), key
            # in place, as the reference calls it (src/ORBextractor.cc:1088)
            buf = src.copy()
            L.orbo_gauss7_u8(buf.ctypes.data, src.shape[1], src.shape[0], src.shape[1], buf.ctypes.data, src.shape[1])
            assert np.array_equal(buf, ref), key
            n += 1
    assert n == 10


def test_fast_matches_cv2_including_order(L):
    cnt = int(GOLD["fast_count"])
    total = 0
    for k in range(cnt):
        src = np.ascontiguousarray(GOLD["fast_%d_src" % k]); th = int(GOLD["fast_%d_th" % k]); ref = GOLD["fast_%d_kps" % k]
        cap = 4096
        xs = np.zeros(cap, np.int32); ys = np.zeros(cap, np.int32); sc = np.zeros(cap, np.int32)
        n = L.orbo_fast9_16(src.ctypes.data, src.shape[1], src.shape[0], src.shape[1], th, 1, xs.ctypes.data, ys.ctypes.data, sc.ctypes.data, cap)
        got = np.stack([xs[:n], ys[:n], sc[:n]], 1)
        assert got.shape == ref.shape and np.array_equal(got, ref), k
        total += n
    assert total > 1000


def test_fast_atan2_matches_cv2_bitwise(L):
    y, x, ref = GOLD["atan2_y"], GOLD["atan2_x"], GOLD["atan2_deg"]
    got = np.array([L.orbo_fast_atan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert got.tobytes() == ref.tobytes()


def test_sincosf_restatement_matches_glibc_golden(L):
    x, rs, rc = GOLD["sincos_x"], GOLD["sincos_sin"], GOLD["sincos_cos"]
    s = C.c_float(); c = C.c_float()
    for i in range(len(x)):
        L.orbo_sincosf(float(x[i]), C.byref(s), C.byref(c))
        assert np.float32(s.value).tobytes() == rs[i].tobytes() and np.float32(c.value).tobytes() == rc[i].tobytes(), float(x[i])


def test_sincosf_restatement_matches_this_hosts_glibc(L):
    # every 61st float in [0, 2*pi]: the restated algorithm == the libm the reference would call here
    two_pi_bits = int(np.float32(6.2831855).view(np.uint32))
    count = two_pi_bits // 61
    assert L.orbo_sincos_checksum(0, count, 61, 0) == L.orbo_sincos_checksum(0, count, 61, 1)


def test_border_reflect101(L):
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, (11, 23), dtype=np.uint8)
    out = np.zeros((11 + 38, 23 + 38), np.uint8)
    L.orbo_border_reflect101_u8(src.ctypes.data, 23, 11, 23, out.ctypes.data, out.shape[1], 19)
    # explicit REFLECT_101 index map (the 19-px pad exceeds the 11-px height, so it reflects twice)
    def r101(i, n):
        while i < 0 or i >= n:
            i = -i if i < 0 else 2 * (n - 1) - i
        return i
    ys = [r101(y - 19, 11) for y in range(out.shape[0])]; xs = [r101(x - 19, 23) for x in range(out.shape[1])]
    assert np.array_equal(out, src[np.ix_(ys, xs)])
