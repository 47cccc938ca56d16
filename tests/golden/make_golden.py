"""Generates the committed golden fixtures in this directory.  Run in the dev container only:

    python tests/golden/make_golden.py

Sources of truth (nothing here comes from our own code):
  * cv2 (opencv-python-headless 4.13.0) for the pixel primitives the reference calls in OpenCV:
    resize INTER_LINEAR, GaussianBlur 7x7 sigma 2, FAST 9/16 with NMS, fastAtan2;
  * glibc's sincosf (via ctypes on libm) for the angle -> cos/sin step of the descriptor;
  * the UNMODIFIED reference extractor (oracle/_ref/liborbref.so = /root/reference/src/ORBextractor.cc
    compiled against oracle/cvshim, ascending-arena tie-break pin) for whole-frame results and for
    DistributeOctTree on random candidate sets.
Outputs: prims_golden.npz, extract_golden.json, octree_golden.npz, gray_golden.npz (`make_golden.py gray` alone).
"""
import ctypes
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g  # noqa: E402
import oracle_lib  # noqa: E402

synth = g.load_package().synth


def smooth_image(rng, w, h):
    import cv2
    a = rng.integers(0, 256, (h // 8 + 2, w // 8 + 2), dtype=np.uint8)
    a = cv2.resize(a, (w, h), interpolation=cv2.INTER_CUBIC)
    return np.ascontiguousarray(np.clip(a.astype(int) + rng.integers(-6, 7, (h, w)), 0, 255).astype(np.uint8))


def make_prims():
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(20261104)
    out = {}
    # resize: (src, dst size) pairs incl. the first level transitions of the four configs at small scale
    cases = [(160, 120, 133, 100), (133, 100, 111, 83), (311, 95, 259, 79), (97, 61, 81, 51), (64, 64, 53, 53)]
    for i, (sw, sh, dw, dh) in enumerate(cases):
        for kind in ("noise", "smooth"):
            src = rng.integers(0, 256, (sh, sw), dtype=np.uint8) if kind == "noise" else smooth_image(rng, sw, sh)
            out["resize_%d_%s_src" % (i, kind)] = src
            out["resize_%d_%s_dst" % (i, kind)] = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
    # blur
    for i, (w, h) in enumerate([(97, 61), (64, 64), (9, 8), (7, 7), (133, 40)]):
        for kind in ("noise", "smooth"):
            src = rng.integers(0, 256, (h, w), dtype=np.uint8) if kind == "noise" else smooth_image(rng, max(w, 16), max(h, 16))[:h, :w].copy()
            out["blur_%d_%s_src" % (i, kind)] = src
            out["blur_%d_%s_dst" % (i, kind)] = cv2.GaussianBlur(src, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    # FAST on cell-sized images at both thresholds
    k = 0
    for trial in range(40):
        w = int(rng.integers(7, 66)); h = int(rng.integers(7, 66))
        src = rng.integers(0, 256, (h, w), dtype=np.uint8) if trial % 2 else smooth_image(rng, max(w, 16), max(h, 16))[:h, :w].copy()
        for th in (7, 20):
            det = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kps = det.detect(src)
            out["fast_%d_src" % k] = src
            out["fast_%d_th" % k] = np.int32(th)
            out["fast_%d_kps" % k] = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in kps], np.int32).reshape(-1, 3)
            k += 1
    out["fast_count"] = np.int32(k)
    # fastAtan2 on moment-sized integer pairs
    ys = rng.integers(-1248480, 1248481, 4000).astype(np.float32); xs = rng.integers(-1248480, 1248481, 4000).astype(np.float32)
    ys[:8] = 0; xs[4:12] = 0
    out["atan2_y"] = ys; out["atan2_x"] = xs
    out["atan2_deg"] = np.array([cv2.fastAtan2(float(y), float(x)) for y, x in zip(ys, xs)], np.float32)
    # glibc sincosf over [0, 2*pi] incl. the tiny / pi/4 branch boundaries
    libm = ctypes.CDLL("libm.so.6")
    libm.sincosf.argtypes = [ctypes.c_float, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float)]
    ang = np.concatenate([rng.uniform(0, 6.2831855, 6000), rng.uniform(0, 1e-3, 500), rng.uniform(0.78, 0.79, 500),
                          np.array([0.0, 2.0 ** -12, 0.7853981, 0.7853982, 6.2831855])]).astype(np.float32)
    s = np.zeros(len(ang), np.float32); c = np.zeros(len(ang), np.float32)
    for i, a in enumerate(ang):
        fs, fc = ctypes.c_float(), ctypes.c_float()
        libm.sincosf(float(a), ctypes.byref(fs), ctypes.byref(fc))
        s[i] = fs.value; c[i] = fc.value
    out["sincos_x"] = ang; out["sincos_sin"] = s; out["sincos_cos"] = c
    np.savez_compressed(os.path.join(HERE, "prims_golden.npz"), **out)
    print("prims_golden.npz:", len(out), "arrays")


def frame_for(family, w, h, seed):
    if family == "synth": return synth.synth(w, h, seed)
    if family == "flat": return synth.flat(w, h)
    if family == "lowtex": return synth.lowtex(w, h, seed)
    if family == "noise": return synth.noise(w, h, seed)
    raise ValueError(family)


EXTRACT_CASES = (
    # (name, family, w, h, nfeatures, seeds)
    [("C1", "synth", 640, 480, 1000, [0x5EED0000 + i for i in range(8)]),
     ("C2", "synth", 1241, 376, 2000, [0x5EED1000 + i for i in range(4)]),
     ("C3", "synth", 752, 480, 1000, [0x5EED2000 + i for i in range(4)]),
     ("C4", "synth", 1920, 1080, 4000, [0x5EED3000 + i for i in range(2)]),
     ("mono_init", "synth", 640, 480, 2000, [0x5EED0000]),
     ("flat", "flat", 640, 480, 1000, [0]),
     ("lowtex", "lowtex", 640, 480, 1000, [7, 8]),
     ("noise", "noise", 640, 480, 1000, [9]),
     ("noise_small", "noise", 200, 160, 500, [10]),
     ("tiny100", "synth", 100, 100, 1000, [3]),
     ("tiny64", "synth", 64, 64, 300, [4]),
     ("odd", "synth", 333, 217, 700, [5])])


def make_extract():
    assert oracle_lib.have_ref(), "build oracle/_ref first (make -f oracle/Makefile ref)"
    cases = []
    for name, family, w, h, nf, seeds in EXTRACT_CASES:
        ref = oracle_lib.OracleRef(nf, 1.2, 8, 20, 7)
        for seed in seeds:
            img = frame_for(family, w, h, seed)
            kps, desc = ref.extract(img, arena_mode=1)
            kps2, desc2 = ref.extract(img, arena_mode=2)
            cases.append({
                "name": name, "family": family, "w": w, "h": h, "nfeatures": nf, "seed": seed,
                "image_sha256": hashlib.sha256(img.tobytes()).hexdigest(),
                "n": int(len(kps)),
                "per_level": np.bincount(kps["octave"], minlength=8).tolist() if len(kps) else [0] * 8,
                "kps_sha256": hashlib.sha256(kps.tobytes()).hexdigest(),
                "desc_sha256": hashlib.sha256(desc.tobytes()).hexdigest(),
                "tie_sensitive": bool(kps.tobytes() != kps2.tobytes() or desc.tobytes() != desc2.tobytes()),
            })
            print(name, seed, len(kps), cases[-1]["tie_sensitive"])
    ref = oracle_lib.OracleRef(1000, 1.2, 8, 20, 7)
    s4, feats, umax = ref.tables()
    doc = {"generator": "tests/golden/make_golden.py", "source": "unmodified reference src/ORBextractor.cc + oracle/cvshim, ascending arena",
           "params": {"scaleFactor": 1.2, "nlevels": 8, "iniThFAST": 20, "minThFAST": 7},
           "tables_nf1000": {"scale_hex": [float(x).hex() for x in s4[0]], "inv_scale_hex": [float(x).hex() for x in s4[1]],
                             "sigma2_hex": [float(x).hex() for x in s4[2]], "inv_sigma2_hex": [float(x).hex() for x in s4[3]],
                             "features_per_level": feats.tolist(), "umax": umax.tolist()},
           "cases": cases}
    with open(os.path.join(HERE, "extract_golden.json"), "w") as f:
        json.dump(doc, f, indent=1)
    print("extract_golden.json:", len(cases), "cases")


def make_octree():
    ref = oracle_lib.OracleRef(1000, 1.2, 8, 20, 7)
    rng = np.random.default_rng(7)
    out = {}
    k = 0
    for (W, H) in [(608, 448), (1209, 344), (147, 102), (40, 40), (300, 100)]:
        for n in (0, 1, 2, 17, 300, 2500):
            for N in (1, 5, 60, 217, 900):
                if n > W * H: continue
                flat_idx = rng.choice(W * H, size=n, replace=False)
                flat_idx.sort()                     # row-major order, as FAST emits within a cell
                if k % 3 == 0: rng.shuffle(flat_idx)
                xs = (flat_idx % W).astype(np.int32); ys = (flat_idx // W).astype(np.int32)
                resp = rng.integers(7, 60 if k % 2 else 255, n).astype(np.int32)
                sel = ref.octree(xs, ys, resp, W, H, N, arena_mode=1)
                out["oct_%d_in" % k] = np.stack([xs, ys, resp], 1).astype(np.int16)
                out["oct_%d_WHN" % k] = np.array([W, H, N], np.int32)
                out["oct_%d_out" % k] = sel.astype(np.int16)
                k += 1
    out["count"] = np.int32(k)
    np.savez_compressed(os.path.join(HERE, "octree_golden.npz"), **out)
    print("octree_golden.npz:", k, "cases")


def make_gray():
    """cv2.cvtColor to gray, the conversion Tracking::GrabImage* applies before the extractor (src/Tracking.cc:244-257)."""
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(77)
    c3 = rng.integers(0, 256, (45, 67, 3), dtype=np.uint8)
    c4 = rng.integers(0, 256, (33, 52, 4), dtype=np.uint8)
    # every value of each channel against extreme values of the others (rounding corners of the fixed point)
    ramp = np.zeros((6, 256, 3), np.uint8)
    for ch in range(3):
        ramp[2 * ch, :, ch] = np.arange(256); ramp[2 * ch + 1, :, ch] = np.arange(256)
        ramp[2 * ch + 1, :, (ch + 1) % 3] = 255; ramp[2 * ch + 1, :, (ch + 2) % 3] = 131
    out = {"c3": c3, "c4": c4, "ramp": ramp,
           "c3_bgr": cv2.cvtColor(c3, cv2.COLOR_BGR2GRAY), "c3_rgb": cv2.cvtColor(c3, cv2.COLOR_RGB2GRAY),
           "c4_bgra": cv2.cvtColor(c4, cv2.COLOR_BGRA2GRAY), "c4_rgba": cv2.cvtColor(c4, cv2.COLOR_RGBA2GRAY),
           "ramp_bgr": cv2.cvtColor(ramp, cv2.COLOR_BGR2GRAY), "ramp_rgb": cv2.cvtColor(ramp, cv2.COLOR_RGB2GRAY)}
    np.savez_compressed(os.path.join(HERE, "gray_golden.npz"), **out)
    print("gray_golden.npz written")


UNDISTORT_CAMERAS = [    # (fx, fy, cx, cy), (k1, k2, p1, p2[, k3]): Examples/Monocular/TUM1.yaml, EuRoC.yaml, TUM2.yaml, a strong barrel
    ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
    ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)),
    ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
    ((300.0, 305.0, 320.0, 240.0), (-0.9, 0.35, 0.01, -0.02, 1.9)),
]


def make_undistort():
    """cv2.undistortPoints(pts, K, D, None, K), what Frame::UndistortKeyPoints / ComputeImageBounds call (src/Frame.cc:404-463)."""
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(4041)
    out = {"count": np.int32(len(UNDISTORT_CAMERAS))}
    for c, (K, D) in enumerate(UNDISTORT_CAMERAS):
        pts = (rng.random((3000, 2)) * np.array([752.0, 480.0])).astype(np.float32)
        pts[:4] = [[0, 0], [752, 0], [0, 480], [752, 480]]                # the image-bounds corners
        pts[4:8] = [[K[2], K[3]], [-50.5, 20.25], [900.0, 700.0], [1e4, -1e4]]   # principal point, outside the image, far outside
        Km = np.array([[K[0], 0, K[2]], [0, K[1], K[3]], [0, 0, 1]], np.float32)
        Dm = np.array(D, np.float32).reshape(-1, 1)
        und = cv2.undistortPoints(pts.reshape(-1, 1, 2), Km, Dm, None, Km).reshape(-1, 2)
        out["K_%d" % c] = np.array(K, np.float32); out["D_%d" % c] = np.array(D, np.float32)
        out["pts_%d" % c] = pts; out["und_%d" % c] = und.astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "undistort_golden.npz"), **out)
    print("undistort_golden.npz written")


def make_gemm():
    """cv2.gemm on 3x3 * 3x1 CV_32F, the shapes of the pose arithmetic in ORBmatcher::SearchByProjection
    (src/ORBmatcher.cc:1338-1361: Rcw*x3Dw+tcw, -Rcw.t()*tcw): alpha*A*x + beta*c with and without GEMM_1_T."""
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(777)
    n = 4000
    A = rng.normal(0, 1, (n, 3, 3)).astype(np.float32)
    for i in range(0, n, 2):                                             # half of them near-rotations, as in a pose
        q, _ = np.linalg.qr(rng.normal(0, 1, (3, 3)))
        A[i] = (q + rng.normal(0, 1e-4, (3, 3))).astype(np.float32)
    x = (rng.normal(0, 1, (n, 3)) * rng.choice([0.01, 1.0, 30.0], (n, 1))).astype(np.float32)
    c = (rng.normal(0, 1, (n, 3)) * rng.choice([0.01, 1.0, 30.0], (n, 1))).astype(np.float32)
    out = {"A": A, "x": x, "c": c}
    for key, alpha, beta, flags in [("ax_plus_c", 1.0, 1.0, 0), ("neg_ax", -1.0, 0.0, 0), ("neg_atx", -1.0, 0.0, cv2.GEMM_1_T),
                                    ("atx_plus_c", 1.0, 1.0, cv2.GEMM_1_T)]:
        r = np.zeros((n, 3), np.float32)
        for i in range(n):
            cc = c[i].reshape(3, 1) if beta != 0 else None
            r[i] = cv2.gemm(A[i], x[i].reshape(3, 1), alpha, cc, beta, flags=flags).reshape(3)
        out[key] = r
    np.savez_compressed(os.path.join(HERE, "gemm_golden.npz"), **out)
    print("gemm_golden.npz written")


def make_scale():
    """cv::Mat::convertTo with a scale (what "M / s" and "s * M" lower to in the Sim3 routines, src/ORBmatcher.cc:301-302,
    :1118-1119), reached from Python through cv2.normalize(NORM_INF): scale = alpha / max|src| in double, then
    src.convertTo(dst, CV_32F, scale, 0).  With alpha = 1 and max|src| = s the call computes src / s."""
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(4242)
    n = 4000
    sc = rng.uniform(0.2, 5.0, n).astype(np.float32)
    src = np.zeros((n, 13), np.float32); dst = np.zeros((n, 13), np.float32)
    for i in range(n):
        v = np.clip((rng.uniform(-1, 1, 12) * sc[i]).astype(np.float32), -sc[i], sc[i])
        src[i] = np.concatenate([v, [sc[i]]])                             # the last element carries max|src| = s
        dst[i] = cv2.normalize(src[i].reshape(1, 13), None, 1.0, 0.0, cv2.NORM_INF, dtype=cv2.CV_32F).reshape(13)
    np.savez_compressed(os.path.join(HERE, "scale_golden.npz"), src=src, s=sc, dst=dst)
    print("scale_golden.npz written")


def make_scale_add():
    """cv::scaleAdd(v, 1 / s, acc) on CV_32F vectors: what "acc + v / s" lowers to in MapPoint::UpdateNormalAndDepth
    (src/MapPoint.cc:355, through MatOp_AddEx::assign).  s is a double (cv::norm's result there)."""
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(777)
    n = 4000
    v = rng.normal(0, 3, (n, 3)).astype(np.float32); acc = rng.normal(0, 2, (n, 3)).astype(np.float32)
    s = np.sqrt((v.astype(np.float64) ** 2).sum(1))
    s[::7] = rng.uniform(0.3, 9.0, len(s[::7]))
    out = np.zeros((n, 3), np.float32)
    for i in range(n):
        out[i] = cv2.scaleAdd(v[i].reshape(3, 1), 1.0 / s[i], acc[i].reshape(3, 1)).reshape(3)
    np.savez_compressed(os.path.join(HERE, "scale_add_golden.npz"), v=v, acc=acc, s=s, out=out)
    print("scale_add_golden.npz written")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "scale_add":
        make_scale_add()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "gemm":
        make_gemm()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "scale":
        make_scale()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "gray":
        make_gray()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "undistort":
        make_undistort()
        sys.exit(0)
    make_prims()
    make_extract()
    make_octree()
