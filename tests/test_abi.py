"""CPU (not gpu): the C-ABI library loads, exports exactly what include/orbgpu.h declares, and its
host-only entry points behave; compute entry points fail loudly without a device (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "orbgpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(orbgpu_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()
    names = header_functions()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), "liborbgpu.so does not export %s" % n
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.lib_path()], capture_output=True, text=True, check=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert exported == names, "exports and header differ: %s" % (set(exported) ^ set(names))


def test_python_binding_covers_header(pkg):
    assert sorted(pkg.lib()._orbgpu_symbols) == header_functions()


def test_keypoint_record_is_28_bytes(pkg):
    assert pkg.KP_DTYPE.itemsize == 28
    assert [pkg.KP_DTYPE.fields[n][1] for n in ("x", "y", "size", "angle", "response", "octave", "class_id")] == [0, 4, 8, 12, 16, 20, 24]


def test_strerror_and_version(pkg):
    L = pkg.lib()
    assert L.orbgpu_strerror(0) == b"ok"
    for code in range(-6, 0):
        assert len(L.orbgpu_strerror(code)) > 3
    assert b"sm_100a" in L.orbgpu_version()


def test_bad_arguments_are_rejected_before_touching_cuda(pkg):
    L = pkg.lib()
    h = C.c_void_p()
    assert L.orbgpu_create(None, 0, 1000, 1.2, 8, 20, 7) == -1
    assert L.orbgpu_create(C.byref(h), 0, 0, 1.2, 8, 20, 7) == -1          # nfeatures < 1
    assert L.orbgpu_create(C.byref(h), 0, 1000, 1.0, 8, 20, 7) == -1        # scale must be > 1
    assert L.orbgpu_create(C.byref(h), 0, 1000, 1.2, 99, 20, 7) == -1       # too many levels
    n = C.c_int(5)
    assert L.orbgpu_extract(None, None, 0, 0, 0, None, None, 0, C.byref(n)) == -1
    assert L.orbgpu_hamming256(np.zeros(32, np.uint8).ctypes.data, np.full(32, 255, np.uint8).ctypes.data) == 256


def test_no_cpu_fallback_without_device(pkg):
    if pkg.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.OrbGpuError):
        pkg.ORBextractor(1000, 1.2, 8, 20, 7)
    m = pkg.ORBmatcher()
    with pytest.raises(pkg.OrbGpuError):
        m.match_bruteforce(np.zeros((4, 32), np.uint8), np.zeros((4, 32), np.uint8))


@settings(max_examples=200, deadline=None)
@given(st.binary(min_size=32, max_size=32), st.binary(min_size=32, max_size=32))
def test_descriptor_distance_host(pkg, oracle, a, b):
    """ORBmatcher::DescriptorDistance drop-in (host popcount) == the reference's SWAR formula."""
    a = np.frombuffer(a, np.uint8); b = np.frombuffer(b, np.uint8)
    d = pkg.ORBmatcher.DescriptorDistance(a, b)
    assert d == int(np.unpackbits(a ^ b).sum())
    assert d == oracle.OraclePort().hamming(a, b)


def test_pattern_table_hash():
    src = open(os.path.join(ROOT, "my-slam_b200", "csrc", "orb_pattern.h")).read()
    body = src[src.index("{") + 1:src.index("};")]
    vals = [int(v) for v in re.findall(r"-?\d+", body)]
    assert len(vals) == 1024 and max(map(abs, vals)) == 13
    import hashlib
    assert hashlib.sha256(bytes((v + 256) % 256 for v in vals)).hexdigest() == \
        "2164181aea6ff9ac426ca512d5130d15e1f6e3cd47b1cbdd568bbe1e55d49023"
