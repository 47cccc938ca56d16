"""GPU parity of the C++ drop-in classes (ORB_SLAM2::ORBextractor / ORBmatcher / FrameQueue) vs the oracle."""
import os
import subprocess

import numpy as np
import pytest

from test_host_cpp import cli  # noqa: F401  (fixture)

pytestmark = pytest.mark.gpu
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def read_records(buf, off):
    n = int(np.frombuffer(buf, np.int32, 1, off)[0]); off += 4
    kps = np.frombuffer(buf, KP, n, off); off += 28 * n
    desc = np.frombuffer(buf, np.uint8, 32 * n, off).reshape(n, 32); off += 32 * n
    return kps, desc, off


def run_extract(cli, tmp_path, img, nf, mode):
    raw = tmp_path / "frame.gray"
    raw.write_bytes(np.ascontiguousarray(img).tobytes())
    out = tmp_path / ("out_%s.bin" % mode)
    h, w = img.shape
    r = subprocess.run([cli, "extract", str(raw), str(w), str(h), str(nf), "1.2", "8", "20", "7", str(out), mode],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out.read_bytes(), r.stdout


@pytest.mark.parametrize("mode,copies", [("call", 1), ("queue", 3), ("batch", 4)])
def test_cpp_extractor_matches_oracle(cli, pkg, oracle, tmp_path, mode, copies):
    img = pkg.synth.synth(640, 480, 0x5EED0000 + 3)
    okps, odesc = oracle.OraclePort(1000, 1.2, 8, 20, 7).extract(img)
    buf, stdout = run_extract(cli, tmp_path, img, 1000, mode)
    off = 0
    for _ in range(copies):
        kps, desc, off = read_records(buf, off)
        assert kps.tobytes() == okps.tobytes()
        assert np.array_equal(desc, odesc)
    assert off == len(buf)
    # getters copied by the Frame constructors (src/Frame.cc:182-188)
    s4, _, _ = oracle.OraclePort(1000, 1.2, 8, 20, 7).tables()
    vals = stdout.split()
    assert vals[1] == "8"
    got = np.array(list(map(float, vals[4:])), np.float32).reshape(8, 4).T
    assert np.array_equal(got, s4.astype(np.float32))


def test_cpp_flat_image_releases_descriptors(cli, pkg, tmp_path):
    buf, _ = run_extract(cli, tmp_path, pkg.synth.flat(320, 240), 500, "call")
    kps, desc, off = read_records(buf, 0)
    assert len(kps) == 0 and off == len(buf)


def test_cpp_pyramid_download_matches_reference_borders(cli, pkg, oracle, tmp_path):
    """mvImagePyramid levels incl. the 19-px REFLECT_101 frame (src/ORBextractor.cc:1109-1137)."""
    img = pkg.synth.synth(752, 480, 0x5EED2000)
    buf, _ = run_extract(cli, tmp_path, img, 1000, "pyramid")
    _, _, off = read_records(buf, 0)
    levels = oracle.OraclePort(1000, 1.2, 8, 20, 7).pyramid(img)
    for l in range(8):
        bw, bh = np.frombuffer(buf, np.int32, 2, off); off += 8
        got = np.frombuffer(buf, np.uint8, int(bw) * int(bh), off).reshape(bh, bw); off += int(bw) * int(bh)
        assert np.array_equal(got, np.pad(levels[l], 19, mode="reflect")), l
    assert off == len(buf)


def test_cpp_matcher_matches_oracle(cli, pkg, oracle, tmp_path):
    rng = np.random.default_rng(5)
    A = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    B = rng.integers(0, 256, (517, 32), dtype=np.uint8)
    B[40:140] = A[100:200] ^ (rng.integers(0, 256, (100, 32), dtype=np.uint8) & rng.integers(0, 256, (100, 32), dtype=np.uint8)
                              & rng.integers(0, 256, (100, 32), dtype=np.uint8) & rng.integers(0, 256, (100, 32), dtype=np.uint8))
    (tmp_path / "A.bin").write_bytes(A.tobytes()); (tmp_path / "B.bin").write_bytes(B.tobytes())
    out = tmp_path / "m.bin"
    r = subprocess.run([cli, "match", str(tmp_path / "A.bin"), "300", str(tmp_path / "B.bin"), "517", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    v = np.frombuffer(out.read_bytes(), np.int32)
    idx, d1, d2, m12, acc, d00 = v[:300], v[300:600], v[600:900], v[900:1200], v[1200], v[1201]
    port = oracle.OraclePort()
    oi, o1, o2 = port.match_bruteforce(A, B)
    assert np.array_equal(idx, oi) and np.array_equal(d1, o1) and np.array_equal(d2, o2)
    ok = (o1 <= 50) & (o1.astype(np.float32) < np.float32(0.6) * o2.astype(np.float32))
    assert acc == int(ok.sum()) == port.match_count(A, B, 50, 0.6) and acc >= 90
    assert np.array_equal(m12, np.where(ok, oi, -1))
    assert d00 == port.hamming(A[0], B[0])


def test_cpp_stereo_matches_equal_oracle(cli, pkg, oracle, tmp_path):
    """Two ORBextractor instances on two threads (src/Frame.cc:78-81), then the device ComputeStereoMatches."""
    left, right = pkg.synth.stereo_pair(640, 480, 0x57E2E0AA)
    (tmp_path / "l.gray").write_bytes(left.tobytes()); (tmp_path / "r.gray").write_bytes(right.tobytes())
    out = tmp_path / "stereo.bin"
    bf, b = 386.1448, 0.5372
    r = subprocess.run([cli, "stereo", str(tmp_path / "l.gray"), str(tmp_path / "r.gray"), "640", "480", "1000", repr(bf), repr(b), str(out)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = out.read_bytes()
    n = int(np.frombuffer(buf, np.int32, 1)[0])
    u = np.frombuffer(buf, np.float32, n, 4); d = np.frombuffer(buf, np.float32, n, 4 + 4 * n)
    pl = oracle.OraclePort(1000, 1.2, 8, 20, 7); pr = oracle.OraclePort(1000, 1.2, 8, 20, 7)
    kL, dL = pl.extract(left); kR, dR = pr.extract(right)
    ou, od = pl.stereo_matches(kL, dL, kR, dR, [pl.level_pixels(l) for l in range(8)], [pr.level_pixels(l) for l in range(8)],
                               np.float32(bf), np.float32(b))
    assert n == len(kL) and u.tobytes() == ou.tobytes() and d.tobytes() == od.tobytes()
    assert (u >= 0).sum() > 300
    # two live extractors = a stereo pair: mvImagePyramid of both is filled by default (src/Frame.cc:473, :563-580 read it)
    vals = list(map(int, r.stdout.split()))
    assert vals[:2] == [len(kL), len(kR)] and vals[2:4] == [640, 480] and vals[4:6] == [179, 134]


@pytest.mark.parametrize("name,levelsup", [("k10-depth4", 2), ("ragged", 2), ("ties", 4)])
def test_cpp_vocabulary_and_bow_database_match_oracle(cli, pkg, tmp_path, name, levelsup):
    """ORB_SLAM2::ORBVocabulary (loadFromTextFile, both transform overloads, score) and BowDatabase through the CLI, against the
    restatement of DBoW2 (oracle/vocab_port.cc)"""
    import oracle_lib
    import test_vocab as tv
    t = tv.make_tree(**tv.TREES[name])
    path = tmp_path / "voc.txt"
    pkg.write_vocabulary_text(path, t["k"], t["L"], t["parent"], t["is_leaf"], t["desc"], t["weight"], trailing_newline=True)
    d = tv.features_for(t, 21, 900)
    raw = tmp_path / "desc.bin"; raw.write_bytes(d.tobytes())
    out = tmp_path / "bow.bin"
    r = subprocess.run([cli, "bow", str(path), str(raw), str(len(d)), str(levelsup), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert r.stdout.startswith("%d words in the vocabulary" % int(t["is_leaf"].sum()))
    buf = out.read_bytes()
    nw, nn, ni = np.frombuffer(buf, np.int32, 3, 0); off = 12
    words = np.frombuffer(buf, np.uint32, nw, off); off += 4 * nw
    values = np.frombuffer(buf, np.float64, nw, off); off += 8 * nw
    nodes = np.frombuffer(buf, np.uint32, nn, off); off += 4 * nn
    start, idx = [0], []
    for _ in range(nn):
        c = int(np.frombuffer(buf, np.int32, 1, off)[0]); off += 4
        idx.extend(np.frombuffer(buf, np.uint32, c, off).tolist()); off += 4 * c
        start.append(len(idx))
    self_score = np.frombuffer(buf, np.float64, 1, off)[0]; off += 8
    common = int(np.frombuffer(buf, np.int32, 1, off)[0]); off += 4
    db_score = np.frombuffer(buf, np.float64, 1, off)[0]
    port = oracle_lib.PortVocabulary(*tv.vocab_args(t))
    want = port.transform(d, levelsup)
    assert tv.same_bow((words, values, nodes, np.array(start, np.int32), np.array(idx, np.uint32)), want)
    ks = np.array([0, nw], np.int32)
    s = port.score_csr(want[0], want[1], want[0], want[1], ks)[0]
    assert self_score == s == db_score and common == nw and abs(s - 1.0) < 1e-12
