"""C++ drop-in classes (my-slam_b200/host): build, host-only logic, and loud failure without a GPU."""
import os
import random
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "my-slam_b200")
CLI = os.path.join(PKG, "orb_host_cli")


@pytest.fixture(scope="module")
def cli():
    subprocess.run(["make", "-s", "-f", os.path.join(PKG, "csrc", "Makefile")], check=True)
    subprocess.run(["make", "-s", "-f", os.path.join(PKG, "host", "Makefile")], check=True)
    assert os.path.exists(CLI) and os.path.exists(os.path.join(PKG, "liborbslam_frontend.so"))
    return CLI


def three_maxima_oracle(sizes):
    """Restatement of ORBmatcher::ComputeThreeMaxima, reference src/ORBmatcher.cc:1601-1642."""
    max1 = max2 = max3 = 0
    ind1 = ind2 = ind3 = -1
    for i, s in enumerate(sizes):
        if s > max1:
            max3, max2, max1 = max2, max1, s
            ind3, ind2, ind1 = ind2, ind1, i
        elif s > max2:
            max3, max2 = max2, s
            ind3, ind2 = ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if max2 < np.float32(0.1) * np.float32(max1):
        ind2 = ind3 = -1
    elif max3 < np.float32(0.1) * np.float32(max1):
        ind3 = -1
    return ind1, ind2, ind3


def test_three_maxima_matches_reference_semantics(cli):
    rnd = random.Random(7)
    cases = [[0] * 30, [5] * 30, [1, 0, 0, 50] + [0] * 26, [3, 5, 2, 9, 0, 1], [10, 1, 1], [100, 10, 9], [100, 9, 9]]
    for _ in range(60):
        cases.append([rnd.choice([0, 0, 1, 2, 3, 5, 8, 13, 40, 90]) for _ in range(30)])
    for sizes in cases:
        out = subprocess.run([cli, "maxima"], input=" ".join(map(str, sizes)), capture_output=True, text=True, check=True).stdout
        assert tuple(map(int, out.split())) == three_maxima_oracle(sizes), sizes


def test_frontend_library_exports_dropin_symbols(cli):
    out = subprocess.run(["nm", "-DC", "--defined-only", os.path.join(PKG, "liborbslam_frontend.so")],
                         capture_output=True, text=True, check=True).stdout
    for sym in ("ORB_SLAM2::ORBextractor::ORBextractor(int, float, int, int, int)",
                "ORB_SLAM2::ORBextractor::operator()(cv::_InputArray const&, cv::_InputArray const&, std::vector<cv::KeyPoint",
                "ORB_SLAM2::ORBmatcher::DescriptorDistance(cv::Mat const&, cv::Mat const&)",
                "ORB_SLAM2::ORBmatcher::ORBmatcher(float, bool)",
                "ORB_SLAM2::ORBmatcher::TH_LOW", "ORB_SLAM2::ORBmatcher::TH_HIGH", "ORB_SLAM2::ORBmatcher::HISTO_LENGTH",
                "ORB_SLAM2::ORBVocabulary::loadFromTextFile(std::",
                "ORB_SLAM2::ORBVocabulary::transform(std::vector<cv::Mat",
                "ORB_SLAM2::ORBVocabulary::score(DBoW2::BowVector const&, DBoW2::BowVector const&) const",
                "ORB_SLAM2::BowDatabase::add(DBoW2::BowVector const&)", "ORB_SLAM2::BowDatabase::query("):
        assert sym in out, sym


def test_extractor_fails_loudly_without_gpu(cli, pkg, tmp_path):
    if pkg.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    raw = tmp_path / "f.gray"
    raw.write_bytes(bytes(64 * 64))
    r = subprocess.run([cli, "extract", str(raw), "64", "64", "500", "1.2", "8", "20", "7", str(tmp_path / "o.bin")],
                       capture_output=True, text=True)
    assert r.returncode == 3 and "orbgpu_create failed" in r.stderr     # no CPU fallback


def test_vocabulary_fails_loudly_without_gpu(cli, pkg, tmp_path):
    if pkg.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    voc = tmp_path / "voc.txt"
    voc.write_text("10 1 0 0\n0 1 " + " ".join(["7"] * 32) + " 1.5")
    raw = tmp_path / "d.bin"
    raw.write_bytes(bytes(64))
    r = subprocess.run([cli, "bow", str(voc), str(raw), "2", "0", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 3 and "orbgpu_vocabulary_load_text" in r.stderr      # no CPU fallback
