"""ctypes bindings to the CPU checkers under oracle/ (test infrastructure only).

OraclePort  -> oracle/_build/liborbport.so  (our restatement, always buildable)
OracleRef   -> oracle/_ref/liborbref.so     (the unmodified reference extractor; built in the dev
                                             container where /root/reference exists, travels prebuilt)
OracleFrameRef -> oracle/_ref/libframeref.so (the unmodified reference Frame.cc / ORBmatcher.cc / KeyFrame.cc /
                                             MapPoint.cc behind oracle/ref_frame_wrap.cc; same methods as the
                                             frame-side restatements of OraclePort)
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE, "_build", "liborbport.so")
REF_SO = os.path.join(ORACLE, "_ref", "liborbref.so")
FRAME_REF_SO = os.path.join(ORACLE, "_ref", "libframeref.so")
# the same reference tree with src/ORBmatcher.cc replaced by the product's my-slam_b200/host/ORBmatcher_gpu.cc (calls liborbgpu.so)
FRAME_GPU_SO = os.path.join(ORACLE, "_ref", "libframegpu.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28


def build_oracle():
    subprocess.run(["make", "-s", "-f", os.path.join(ORACLE, "Makefile"), "all"], check=True)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _i32(n):
    return np.zeros(max(n, 1), np.int32)


class OraclePort:
    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        if not os.path.exists(PORT_SO):
            build_oracle()
        L = self.L = C.CDLL(PORT_SO)
        L.orbport_create.restype = C.c_void_p
        L.orbport_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orbport_destroy.argtypes = [C.c_void_p]
        L.orbport_tables.argtypes = [C.c_void_p] * 4
        L.orbport_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p,
                                      C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        L.orbport_pyramid.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t]
        L.orbport_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orbport_level_pixels.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.orbport_level_blurred.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        for f in (L.orbport_level_candidates, L.orbport_level_selected):
            f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orbport_octree.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_void_p, C.c_int]
        L.orbport_hamming256.argtypes = [C.c_void_p, C.c_void_p]
        L.orbport_match_bruteforce.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                               C.c_void_p, C.c_void_p]
        L.orbport_match_count.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float]
        L.orbport_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.frameport_features_in_area.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int,
                                                 C.c_int, C.c_void_p, C.c_int]
        L.frameport_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                          C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]
        L.frameport_search_by_projection.argtypes = [C.c_void_p, C.c_void_p]
        L.frameport_search_local_points.argtypes = [C.c_void_p, C.c_void_p]
        L.frameport_search_for_triangulation.argtypes = [C.c_void_p, C.c_void_p]
        L.frameport_search_local_map.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameport_fuse_search.argtypes = [C.c_void_p, C.c_void_p]
        L.frameport_fuse_search_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameport_fuse_search_sim3.restype = None
        L.frameport_search_by_projection_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameport_search_by_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameport_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.frameport_distinctive_descriptors.restype = None
        L.frameport_search_by_projection_keyframe.argtypes = [C.c_void_p, C.c_void_p]
        L.frameport_fuse_search.restype = None
        L.frameport_logf_array.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.frameport_logf_array.restype = None
        L.frameport_logf_port_mismatches.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
        L.frameport_logf_port_mismatches.restype = C.c_longlong
        L.frameport_search_by_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        self.nlevels = nlevels
        self.h = L.orbport_create(nfeatures, scale, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orbport_destroy(self.h)
            self.h = None

    def tables(self):
        s4 = np.zeros(4 * self.nlevels, np.float32)
        feats = _i32(self.nlevels)
        umax = _i32(16)
        self.L.orbport_tables(self.h, _ptr(s4), _ptr(feats), _ptr(umax))
        return s4.reshape(4, self.nlevels), feats, umax

    def extract(self, img, cap=None):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        cap = cap or 65536
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        rc = self.L.orbport_extract(self.h, _ptr(img), w, h, img.strides[0], _ptr(kps), _ptr(desc), cap, C.byref(n))
        if rc != 0:
            raise RuntimeError("orbport_extract rc=%d n=%d" % (rc, n.value))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def pyramid(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        self.L.orbport_pyramid(self.h, _ptr(img), img.shape[1], img.shape[0], img.strides[0])
        return [self.level_pixels(l) for l in range(self.nlevels)]

    def level_size(self, l):
        w, h = C.c_int(), C.c_int()
        self.L.orbport_level_size(self.h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_pixels(self, l, blurred=False):
        w, h = self.level_size(l)
        out = np.zeros((h, w), np.uint8)
        f = self.L.orbport_level_blurred if blurred else self.L.orbport_level_pixels
        rc = f(self.h, l, _ptr(out), w)
        return out if rc == 0 else None

    def _cands(self, f, l):
        cap = 1 << 20
        x, y, s = _i32(cap), _i32(cap), _i32(cap)
        n = f(self.h, l, _ptr(x), _ptr(y), _ptr(s), cap)
        return np.stack([x[:n], y[:n], s[:n]], 1)

    def level_candidates(self, l):
        return self._cands(self.L.orbport_level_candidates, l)

    def level_selected(self, l):
        return self._cands(self.L.orbport_level_selected, l)

    def octree(self, xs, ys, resp, W, H, N):
        xs = np.ascontiguousarray(xs, np.int32); ys = np.ascontiguousarray(ys, np.int32)
        resp = np.ascontiguousarray(resp, np.int32)
        out = _i32(len(xs) + 8)
        n = self.L.orbport_octree(_ptr(xs), _ptr(ys), _ptr(resp), len(xs), W, H, N, _ptr(out), len(out))
        if n < 0:
            raise RuntimeError("orbport_octree rc=%d" % n)
        return out[:n].copy()

    def hamming(self, a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.L.orbport_hamming256(_ptr(a), _ptr(b))

    def match_bruteforce(self, A, B):
        A = np.ascontiguousarray(A, np.uint8); B = np.ascontiguousarray(B, np.uint8)
        bi = np.zeros(len(A), np.int32); b1 = np.zeros(len(A), np.uint16); b2 = np.zeros(len(A), np.uint16)
        self.L.orbport_match_bruteforce(_ptr(A), len(A), _ptr(B), len(B), _ptr(bi), _ptr(b1), _ptr(b2))
        return bi, b1, b2

    def stereo_matches(self, kL, dL, kR, dR, lvlL, lvlR, bf, b):
        """Frame::ComputeStereoMatches restated (oracle/orb_port.h; pinned by tests/test_ref_frame.py) -> (mvuRight, mvDepth)."""
        nl = self.nlevels
        lvlL = [np.ascontiguousarray(a, np.uint8) for a in lvlL]; lvlR = [np.ascontiguousarray(a, np.uint8) for a in lvlR]
        pl = (C.c_void_p * nl)(*[a.ctypes.data for a in lvlL]); pr = (C.c_void_p * nl)(*[a.ctypes.data for a in lvlR])
        w = np.array([a.shape[1] for a in lvlL], np.int32); h = np.array([a.shape[0] for a in lvlL], np.int32)
        s4 = self.tables()[0]
        sc = np.ascontiguousarray(s4[0]); isc = np.ascontiguousarray(s4[1])
        kL = np.ascontiguousarray(kL); kR = np.ascontiguousarray(kR)
        dL = np.ascontiguousarray(dL, np.uint8); dR = np.ascontiguousarray(dR, np.uint8)
        u = np.zeros(max(len(kL), 1), np.float32); d = np.zeros(max(len(kL), 1), np.float32)
        self.L.orbport_stereo_matches(_ptr(kL), _ptr(dL), len(kL), _ptr(kR), _ptr(dR), len(kR), pl, pr, _ptr(w), _ptr(h),
                                      _ptr(sc), _ptr(isc), bf, b, _ptr(u), _ptr(d))
        return u[:len(kL)].copy(), d[:len(kL)].copy()

    # ---- frame_port.cc (restatements of src/Frame.cc / src/ORBmatcher.cc searches; pinned by tests/test_ref_frame.py) ----
    def features_in_area(self, keys_un, bounds, x, y, r, min_level=-1, max_level=-1):
        k = np.ascontiguousarray(keys_un); b4 = np.ascontiguousarray(bounds, np.float32)
        out = _i32(len(k))
        n = self.L.frameport_features_in_area(_ptr(k), len(k), _ptr(b4), x, y, r, min_level, max_level, _ptr(out), len(out))
        return out[:n].copy()

    def search_for_initialization(self, k1, d1, k2, d2, bounds, prev_matched, window_size=100, nnratio=0.9, check_orientation=True):
        k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
        d1 = np.ascontiguousarray(d1, np.uint8); d2 = np.ascontiguousarray(d2, np.uint8)
        b4 = np.ascontiguousarray(bounds, np.float32)
        pm = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
        m12 = _i32(len(k1))
        nm = self.L.frameport_search_for_initialization(_ptr(k1), _ptr(d1), len(k1), _ptr(k2), _ptr(d2), len(k2), _ptr(b4),
                                                        _ptr(pm), window_size, nnratio, int(check_orientation), _ptr(m12))
        return nm, m12[:len(k1)].copy(), pm

    def search_by_projection(self, args):
        """args: the product's ProjectionArgs ctypes struct (same layout as frameport_projection_args)."""
        cm = _i32(args.n_cur)
        nm = self.L.frameport_search_by_projection(C.byref(args), _ptr(cm))
        return nm, cm[:args.n_cur].copy()

    def search_local_points(self, args):
        """args: the product's LocalPointsArgs ctypes struct (same layout as frameport_local_points_args)."""
        cm = _i32(args.n_cur)
        nm = self.L.frameport_search_local_points(C.byref(args), _ptr(cm))
        return nm, cm[:args.n_cur].copy()

    def search_for_triangulation(self, args):
        """args: the product's TriangulationArgs ctypes struct (same layout as frameport_triangulation_args)."""
        m12 = _i32(args.kf1.n)
        nm = self.L.frameport_search_for_triangulation(C.byref(args), _ptr(m12))
        return nm, m12[:args.kf1.n].copy()

    def search_local_map(self, args, track_fields):
        """args: the product's LocalMapArgs; track_fields: its track_fields(n) factory -> (nmatches, cur_match, n_in_view, fields)."""
        t, arrs = track_fields(args.n_points)
        cm = _i32(args.n_cur); nv = C.c_int(0)
        nm = self.L.frameport_search_local_map(C.byref(args), C.byref(t), C.byref(nv), _ptr(cm))
        return nm, cm[:args.n_cur].copy(), nv.value, {k: v[:args.n_points] for k, v in arrs.items()}

    def search_by_projection_keyframe(self, args):
        cm = _i32(args.n_cur)
        nm = self.L.frameport_search_by_projection_keyframe(C.byref(args), _ptr(cm))
        return nm, cm[:args.n_cur].copy()

    def distinctive_descriptors(self, desc, group_start):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); gs = np.ascontiguousarray(group_start, np.int32)
        out = _i32(len(gs) - 1)
        self.L.frameport_distinctive_descriptors(_ptr(d), _ptr(gs), len(gs) - 1, _ptr(out))
        return out[:len(gs) - 1].copy()

    def fuse_search_sim3(self, args, Scw):
        S = np.ascontiguousarray(Scw, np.float32).reshape(16); best = _i32(args.n_points)
        self.L.frameport_fuse_search_sim3(C.byref(args), _ptr(S), _ptr(best))
        return best[:args.n_points].copy()

    def search_by_projection_sim3(self, args, Scw, matched_in):
        S = np.ascontiguousarray(Scw, np.float32).reshape(16); mi = np.ascontiguousarray(matched_in, np.uint8); out = _i32(args.n_kf)
        nm = self.L.frameport_search_by_projection_sim3(C.byref(args), _ptr(S), _ptr(mi), _ptr(out))
        return nm, out[:args.n_kf].copy()

    def search_by_sim3(self, a12, a21, s12, R12, t12):
        R = np.ascontiguousarray(R12, np.float32).reshape(9); t = np.ascontiguousarray(t12, np.float32).reshape(3); out = _i32(a12.n_points)
        nf = self.L.frameport_search_by_sim3(C.byref(a12), C.byref(a21), float(s12), _ptr(R), _ptr(t), _ptr(out))
        return nf, out[:a12.n_points].copy()

    def fuse_search(self, args):
        """args: the product's FuseArgs ctypes struct (same layout as frameport_fuse_args) -> best_idx[n_points]."""
        best = _i32(args.n_points)
        self.L.frameport_fuse_search(C.byref(args), _ptr(best))
        return best[:args.n_points].copy()

    def logf_port_mismatches(self, first, count, step):
        return int(self.L.frameport_logf_port_mismatches(first, count, step))

    def logf(self, x):
        """glibc logf (the call MapPoint::PredictScale makes) on a float32 array."""
        x = np.ascontiguousarray(x, np.float32); out = np.zeros_like(x)
        self.L.frameport_logf_array(_ptr(x), len(x), _ptr(out))
        return out

    def search_by_bow(self, side1, side2, nnratio=0.7, check_orientation=True, strict_th_low=False):
        """sides: the product's BowSide ctypes structs (same layout as frameport_bow_side)."""
        m12 = _i32(side1.n); m21 = _i32(side2.n)
        nm = self.L.frameport_search_by_bow(C.byref(side1), C.byref(side2), nnratio, int(check_orientation), int(strict_th_low),
                                            _ptr(m12), _ptr(m21))
        return nm, m12[:side1.n].copy(), m21[:side2.n].copy()

    def count_pairs_bench(self, desc, pairs, th_low=50, ratio=0.6, threads=1, popcnt=False):
        """All-pairs accepted-match counts on the CPU (timed): returns (counts int32[P], seconds)."""
        desc = np.ascontiguousarray(desc, np.uint8); pairs = np.ascontiguousarray(pairs, np.int32)
        counts = np.zeros(len(pairs), np.int32)
        f = self.L.orbport_count_pairs_bench
        f.restype = C.c_double
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_int, C.c_int]
        sec = f(_ptr(desc), desc.shape[1], _ptr(pairs), len(pairs), th_low, ratio, _ptr(counts), threads, 1 if popcnt else 0)
        return counts, sec

    def match_count(self, A, B, th_low=50, ratio=0.6):
        A = np.ascontiguousarray(A, np.uint8); B = np.ascontiguousarray(B, np.uint8)
        return self.L.orbport_match_count(_ptr(A), len(A), _ptr(B), len(B), th_low, ratio)


def have_ref():
    return os.path.exists(REF_SO)


class OracleRef:
    """The unmodified reference ORBextractor (src/ORBextractor.cc) behind a C wrapper."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        L = self.L = C.CDLL(REF_SO)
        L.orbref_create.restype = C.c_void_p
        L.orbref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orbref_destroy.argtypes = [C.c_void_p]
        L.orbref_tables.argtypes = [C.c_void_p] * 4
        L.orbref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p,
                                     C.c_int, C.POINTER(C.c_int), C.c_int]
        L.orbref_compute_pyramid.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t]
        L.orbref_pyramid_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int),
                                           C.POINTER(C.c_int)]
        L.orbref_pyramid_level_bordered.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
        L.orbref_octree.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.orbref_bench.restype = C.c_double
        L.orbref_bench.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_float, C.c_int,
                                   C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_long)]
        self.params = (nfeatures, scale, nlevels, ini_th, min_th)
        self.nlevels = nlevels
        self.h = L.orbref_create(nfeatures, scale, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orbref_destroy(self.h)
            self.h = None

    def tables(self):
        s4 = np.zeros(4 * self.nlevels, np.float32)
        feats = _i32(self.nlevels)
        umax = _i32(16)
        self.L.orbref_tables(self.h, _ptr(s4), _ptr(feats), _ptr(umax))
        return s4.reshape(4, self.nlevels), feats, umax

    def extract(self, img, arena_mode=1, cap=65536):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        rc = self.L.orbref_extract(self.h, _ptr(img), w, h, img.strides[0], _ptr(kps), _ptr(desc), cap,
                                   C.byref(n), arena_mode)
        if rc != 0:
            raise RuntimeError("orbref_extract rc=%d" % rc)
        return kps[:n.value].copy(), desc[:n.value].copy()

    def pyramid(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        self.L.orbref_compute_pyramid(self.h, _ptr(img), img.shape[1], img.shape[0], img.strides[0])
        out = []
        for l in range(self.nlevels):
            w, h = C.c_int(), C.c_int()
            self.L.orbref_pyramid_level(self.h, l, None, 0, C.byref(w), C.byref(h))
            a = np.zeros((h.value, w.value), np.uint8)
            self.L.orbref_pyramid_level(self.h, l, _ptr(a), w.value, C.byref(w), C.byref(h))
            out.append(a)
        return out

    def pyramid_level_bordered(self, l, w, h):
        a = np.zeros((h + 38, w + 38), np.uint8)
        self.L.orbref_pyramid_level_bordered(self.h, l, _ptr(a), w + 38)
        return a

    def octree(self, xs, ys, resp, W, H, N, arena_mode=1):
        """Returns (x, y, response) rows of the selected keys, in result order."""
        xs = np.ascontiguousarray(xs, np.float32); ys = np.ascontiguousarray(ys, np.float32)
        resp = np.ascontiguousarray(resp, np.float32)
        cap = len(xs) + 8
        ox, oy, orr = np.zeros(cap, np.float32), np.zeros(cap, np.float32), np.zeros(cap, np.float32)
        n = self.L.orbref_octree(self.h, _ptr(xs), _ptr(ys), _ptr(resp), len(xs), 0, W, 0, H, N,
                                 _ptr(ox), _ptr(oy), _ptr(orr), cap, arena_mode)
        if n < 0:
            raise RuntimeError("orbref_octree rc=%d" % n)
        return np.stack([ox[:n], oy[:n], orr[:n]], 1)

    def bench(self, frames, threads, reps=1):
        frames = np.ascontiguousarray(frames, np.uint8)
        n, h, w = frames.shape
        tot = C.c_long(0)
        nf, sc, nl, ini, mn = self.params
        sec = self.L.orbref_bench(_ptr(frames), n, w, h, frames.strides[0], nf, sc, nl, ini, mn, threads, reps,
                                  C.byref(tot))
        return sec, tot.value


def have_frame_ref():
    return os.path.exists(FRAME_REF_SO)


def have_frame_gpu():
    return os.path.exists(FRAME_GPU_SO) and os.path.exists(FRAME_REF_SO)


class OracleFrameRef:
    """The unmodified reference Frame / ORBmatcher code (oracle/ref_frame_wrap.cc).  Method names and results follow the
    frame-side methods of OraclePort so a test can run one case through both."""

    def __init__(self, scale_tables=None, so=None):
        L = self.L = C.CDLL(so or FRAME_REF_SO)
        L.frameref_gemm3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_void_p]
        L.frameref_neg_rt_t.argtypes = [C.c_void_p] * 4
        L.frameref_features_in_area.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int,
                                                C.c_int, C.c_void_p, C.c_int]
        L.frameref_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                                         C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_void_p]
        L.frameref_search_by_projection.argtypes = [C.c_void_p, C.c_void_p]
        L.frameref_search_local_points.argtypes = [C.c_void_p, C.c_void_p]
        L.frameref_search_for_triangulation.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameref_search_local_map.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameref_fuse_search.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameref_fuse_search_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameref_search_by_projection_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameref_search_by_sim3.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frameref_mat_div.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_void_p]
        L.frameref_mat_div.restype = None
        L.frameref_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.frameref_distinctive_descriptors.restype = None
        L.frameref_search_by_projection_keyframe.argtypes = [C.c_void_p, C.c_void_p]
        L.frameref_fuse_search.restype = None
        L.frameref_search_by_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.frameref_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                              C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.frameref_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                                   C.c_void_p, C.c_void_p]
        L.frameref_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p]
        L.frameref_three_maxima.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        self.scale_tables = scale_tables     # (scale, inv_scale) float32 arrays, for stereo_matches

    def descriptor_distance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.L.frameref_descriptor_distance(_ptr(a), _ptr(b))

    def three_maxima(self, sizes):
        s = np.ascontiguousarray(sizes, np.int32); out = _i32(3)
        self.L.frameref_three_maxima(_ptr(s), len(s), _ptr(out))
        return tuple(int(v) for v in out)

    def gemm3(self, A, x, c=None, alpha=1.0, beta=1.0, flags=0):
        A = np.ascontiguousarray(A, np.float32); x = np.ascontiguousarray(x, np.float32)
        c = None if c is None else np.ascontiguousarray(c, np.float32)
        d = np.zeros(3, np.float32)
        self.L.frameref_gemm3(_ptr(A), _ptr(x), None if c is None else _ptr(c), alpha, beta, flags, _ptr(d))
        return d

    def neg_rt_t(self, Tcw, Tlw):
        Tcw = np.ascontiguousarray(Tcw, np.float32); Tlw = np.ascontiguousarray(Tlw, np.float32)
        twc = np.zeros(3, np.float32); tlc = np.zeros(3, np.float32)
        self.L.frameref_neg_rt_t(_ptr(Tcw), _ptr(Tlw), _ptr(twc), _ptr(tlc))
        return twc, tlc

    def features_in_area(self, keys_un, bounds, x, y, r, min_level=-1, max_level=-1):
        k = np.ascontiguousarray(keys_un); b4 = np.ascontiguousarray(bounds, np.float32)
        out = _i32(len(k))
        n = self.L.frameref_features_in_area(_ptr(k), len(k), _ptr(b4), x, y, r, min_level, max_level, _ptr(out), len(out))
        return out[:n].copy()

    def search_for_initialization(self, k1, d1, k2, d2, bounds, prev_matched, window_size=100, nnratio=0.9, check_orientation=True):
        k1 = np.ascontiguousarray(k1); k2 = np.ascontiguousarray(k2)
        d1 = np.ascontiguousarray(d1, np.uint8); d2 = np.ascontiguousarray(d2, np.uint8)
        b4 = np.ascontiguousarray(bounds, np.float32)
        pm = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
        m12 = _i32(len(k1))
        nm = self.L.frameref_search_for_initialization(_ptr(k1), _ptr(d1), len(k1), _ptr(k2), _ptr(d2), len(k2), _ptr(b4),
                                                       _ptr(pm), window_size, nnratio, int(check_orientation), _ptr(m12))
        return nm, m12[:len(k1)].copy(), pm

    def search_by_projection(self, args):
        """-> (nmatches, cur_match) with cur_match[i2] = last-frame index or -1 (the restatement's -2 reads as -1 here)."""
        cm = _i32(args.n_cur)
        nm = self.L.frameref_search_by_projection(C.byref(args), _ptr(cm))
        return nm, cm[:args.n_cur].copy()

    def search_local_points(self, args):
        """ORBmatcher::SearchByProjection(F, vpMapPoints, th) executed by the reference -> (nmatches, cur_match)."""
        cm = _i32(args.n_cur)
        nm = self.L.frameref_search_local_points(C.byref(args), _ptr(cm))
        return nm, cm[:args.n_cur].copy()

    def search_for_triangulation(self, args, T1w):
        """ORBmatcher::SearchForTriangulation executed by the reference on KeyFrames posed at T1w / args.T2w
        -> (nmatches, vMatches12, the camera centre of pKF1 the reference derived from T1w)."""
        m12 = _i32(args.kf1.n); cw = np.zeros(3, np.float32)
        T1 = np.ascontiguousarray(T1w, np.float32)
        nm = self.L.frameref_search_for_triangulation(C.byref(args), _ptr(T1), _ptr(m12), _ptr(cw))
        return nm, m12[:args.kf1.n].copy(), cw

    def search_local_map(self, args, track_fields):
        """args: the product's LocalMapArgs; track_fields: its track_fields(n) factory -> (nmatches, cur_match, n_in_view, fields)."""
        t, arrs = track_fields(args.n_points)
        cm = _i32(args.n_cur); nv = C.c_int(0)
        nm = self.L.frameref_search_local_map(C.byref(args), C.byref(t), C.byref(nv), _ptr(cm))
        return nm, cm[:args.n_cur].copy(), nv.value, {k: v[:args.n_points] for k, v in arrs.items()}

    def search_by_projection_keyframe(self, args):
        """-> (nmatches, cur_match) with cur_match[i2] = keyframe feature index or -1 (the restatement's -2 reads as -1 here)."""
        cm = _i32(args.n_cur)
        nm = self.L.frameref_search_by_projection_keyframe(C.byref(args), _ptr(cm))
        return nm, cm[:args.n_cur].copy()

    def distinctive_descriptors(self, desc, group_start):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); gs = np.ascontiguousarray(group_start, np.int32)
        out = _i32(len(gs) - 1)
        self.L.frameref_distinctive_descriptors(_ptr(d), _ptr(gs), len(gs) - 1, _ptr(out))
        return out[:len(gs) - 1].copy()

    def mat_div(self, v, s):
        """cv::Mat(v) / s through the shim's MatExpr lowering."""
        v = np.ascontiguousarray(v, np.float32); out = np.zeros_like(v)
        self.L.frameref_mat_div(_ptr(v), len(v), float(s), _ptr(out))
        return out

    def mat_add_div(self, acc, v, s):
        """cv::Mat(acc) + cv::Mat(v) / s (s a double) through the shim's MatExpr lowering (cv::scaleAdd)."""
        acc = np.ascontiguousarray(acc, np.float32); v = np.ascontiguousarray(v, np.float32); out = np.zeros_like(v)
        self.L.frameref_mat_add_div.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p]
        self.L.frameref_mat_add_div.restype = None
        self.L.frameref_mat_add_div(_ptr(acc), _ptr(v), len(v), float(s), _ptr(out))
        return out

    def update_normal_and_depth(self, pos, obs_start, centers, ref_slot, ref_level, scale_factors, normal, max_dist, min_dist):
        """MapPoint::UpdateNormalAndDepth executed by the reference, point by point"""
        pos = np.ascontiguousarray(pos, np.float32); os_ = np.ascontiguousarray(obs_start, np.int32)
        c = np.ascontiguousarray(centers, np.float32); rs = np.ascontiguousarray(ref_slot, np.int32); rl = np.ascontiguousarray(ref_level, np.int32)
        sf = np.ascontiguousarray(scale_factors, np.float32)
        nv = np.array(normal, np.float32).copy(); mx = np.array(max_dist, np.float32).copy(); mn = np.array(min_dist, np.float32).copy()
        f = self.L.frameref_update_normal_and_depth
        f.argtypes = [C.c_void_p] * 6 + [C.c_int, C.c_int] + [C.c_void_p] * 3
        f.restype = None
        f(_ptr(pos), _ptr(os_), _ptr(c), _ptr(rs), _ptr(rl), _ptr(sf), len(sf), len(os_) - 1, _ptr(nv), _ptr(mx), _ptr(mn))
        return nv, mx, mn

    def fuse_search_sim3(self, args, Scw):
        S = np.ascontiguousarray(Scw, np.float32).reshape(16); best = _i32(args.n_points)
        self.L.frameref_fuse_search_sim3(C.byref(args), _ptr(S), _ptr(best))
        return best[:args.n_points].copy()

    def search_by_projection_sim3(self, args, Scw, matched_in):
        S = np.ascontiguousarray(Scw, np.float32).reshape(16); mi = np.ascontiguousarray(matched_in, np.uint8); out = _i32(args.n_kf)
        nm = self.L.frameref_search_by_projection_sim3(C.byref(args), _ptr(S), _ptr(mi), _ptr(out))
        return nm, out[:args.n_kf].copy()

    def search_by_sim3(self, a12, a21, s12, R12, t12):
        R = np.ascontiguousarray(R12, np.float32).reshape(9); t = np.ascontiguousarray(t12, np.float32).reshape(3); out = _i32(a12.n_points)
        nf = self.L.frameref_search_by_sim3(C.byref(a12), C.byref(a21), float(s12), _ptr(R), _ptr(t), _ptr(out))
        return nf, out[:a12.n_points].copy()

    def fuse_search(self, args):
        """The reference's ORBmatcher::Fuse, one map point per call -> (best_idx[n_points], the camera centre it used)."""
        best = _i32(args.n_points); ow = np.zeros(3, np.float32)
        self.L.frameref_fuse_search(C.byref(args), _ptr(best), _ptr(ow))
        return best[:args.n_points].copy(), ow

    def search_by_bow(self, side1, side2, nnratio=0.7, check_orientation=True, strict_th_low=False):
        m12 = _i32(side1.n); m21 = _i32(side2.n)
        nm = self.L.frameref_search_by_bow(C.byref(side1), C.byref(side2), nnratio, int(check_orientation), int(strict_th_low),
                                           _ptr(m12), _ptr(m21))
        return nm, m12[:side1.n].copy(), m21[:side2.n].copy()

    def search_by_bow_thrice(self, side1, side2, nnratio=0.7, check_orientation=True):
        """SearchByBoW(pKF1, pKF2) on the same two KeyFrame objects: before pKF2 has a FeatureVector, after, and after every
        third map point of pKF1 was erased -> [(nmatches, match12)] * 3."""
        f = self.L.frameref_search_by_bow_thrice
        f.restype = C.c_long
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p]
        m = _i32(3 * side1.n)
        packed = f(C.byref(side1), C.byref(side2), nnratio, int(check_orientation), _ptr(m))
        return [(int(packed // 1000 ** k % 1000), m[k * side1.n:(k + 1) * side1.n].copy()) for k in range(3)]

    def stereo_matches(self, kL, dL, kR, dR, lvlL, lvlR, bf, b):
        nl = len(lvlL)
        lvlL = [np.ascontiguousarray(a, np.uint8) for a in lvlL]; lvlR = [np.ascontiguousarray(a, np.uint8) for a in lvlR]
        pl = (C.c_void_p * nl)(*[a.ctypes.data for a in lvlL]); pr = (C.c_void_p * nl)(*[a.ctypes.data for a in lvlR])
        w = np.array([a.shape[1] for a in lvlL], np.int32); h = np.array([a.shape[0] for a in lvlL], np.int32)
        sc = np.ascontiguousarray(self.scale_tables[0], np.float32); isc = np.ascontiguousarray(self.scale_tables[1], np.float32)
        kL = np.ascontiguousarray(kL); kR = np.ascontiguousarray(kR)
        dL = np.ascontiguousarray(dL, np.uint8); dR = np.ascontiguousarray(dR, np.uint8)
        u = np.zeros(max(len(kL), 1), np.float32); d = np.zeros(max(len(kL), 1), np.float32)
        self.L.frameref_stereo_matches(_ptr(kL), _ptr(dL), len(kL), _ptr(kR), _ptr(dR), len(kR), pl, pr, _ptr(w), _ptr(h),
                                       _ptr(sc), _ptr(isc), nl, bf, b, _ptr(u), _ptr(d))
        return u[:len(kL)].copy(), d[:len(kL)].copy()

    def undistort_keypoints(self, keys, K4, dist, w=0, h=0):
        """Frame::UndistortKeyPoints (+ ComputeImageBounds when w, h > 0) -> (mvKeysUn, bounds or None)."""
        k = np.ascontiguousarray(keys); out = k.copy()
        K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
        b4 = np.zeros(4, np.float32)
        self.L.frameref_undistort_keypoints(_ptr(k), len(k), _ptr(K4), _ptr(dist), len(dist), w, h, _ptr(out),
                                            _ptr(b4) if w > 0 else None)
        return out, (b4 if w > 0 else None)


def _bow_outputs(n):
    m = max(n, 1)
    return (np.zeros(m, np.uint32), np.zeros(m, np.float64), np.zeros(m, np.uint32), np.zeros(m + 1, np.int32),
            np.zeros(m, np.uint32), np.zeros(3, np.int32))


def _bow_result(words, values, nodes, start, idx, counts):
    nw, nn, ni = (int(x) for x in counts)
    return words[:nw].copy(), values[:nw].copy(), nodes[:nn].copy(), start[:nn + 1].copy(), idx[:ni].copy()


class PortVocabulary:
    """oracle/vocab_port.cc: the restatement of DBoW2's transform / L1 score on flat arrays (same methods as
    my-slam_b200's ORBVocabulary)."""

    def __init__(self, k, L, parent, is_leaf, desc, weight, scoring=0, weighting=0):
        lib = self.L = C.CDLL(PORT_SO)
        lib.frameport_vocab_create.restype = C.c_void_p
        lib.frameport_vocab_create.argtypes = [C.c_int] * 5 + [C.c_void_p] * 4
        lib.frameport_vocab_free.argtypes = [C.c_void_p]
        lib.frameport_vocab_free.restype = None
        lib.frameport_vocab_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6
        lib.frameport_vocab_transform.restype = None
        lib.frameport_vocab_score.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        lib.frameport_vocab_score.restype = None
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); weight = np.ascontiguousarray(weight, np.float64)
        self.h = lib.frameport_vocab_create(k, L, scoring, weighting, len(parent), _ptr(parent), _ptr(is_leaf), _ptr(desc), _ptr(weight))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.frameport_vocab_free(self.h)
            self.h = None

    def transform(self, desc, levelsup=4):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        out = _bow_outputs(len(d))
        self.L.frameport_vocab_transform(self.h, _ptr(d), len(d), levelsup, *[_ptr(a) for a in out])
        return _bow_result(*out)

    def score_csr(self, qw, qv, kw, kv, kstart):
        qw = np.ascontiguousarray(qw, np.uint32); qv = np.ascontiguousarray(qv, np.float64)
        kw = np.ascontiguousarray(kw, np.uint32); kv = np.ascontiguousarray(kv, np.float64)
        ks = np.ascontiguousarray(kstart, np.int32)
        out = np.zeros(max(len(ks) - 1, 1), np.float64)
        self.L.frameport_vocab_score(_ptr(qw), _ptr(qv), len(qw), _ptr(kw), _ptr(kv), _ptr(ks), len(ks) - 1, _ptr(out))
        return out[:len(ks) - 1].copy()


class RefVocabulary:
    """The reference's own ORBVocabulary (oracle/ref_vocab_wrap.cc in oracle/_ref/libframeref.so), read from a text file by
    TemplatedVocabulary::loadFromTextFile."""

    def __init__(self, path):
        lib = self.L = C.CDLL(FRAME_REF_SO)
        lib.frameref_vocab_load.restype = C.c_void_p
        lib.frameref_vocab_load.argtypes = [C.c_char_p]
        lib.frameref_vocab_free.argtypes = [C.c_void_p]
        lib.frameref_vocab_free.restype = None
        lib.frameref_vocab_info.argtypes = [C.c_void_p, C.c_void_p]
        lib.frameref_vocab_info.restype = None
        for f in (lib.frameref_compute_bow, lib.frameref_vocab_transform):
            f.restype = None
        lib.frameref_compute_bow.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 6
        lib.frameref_vocab_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6
        lib.frameref_vocab_score.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        lib.frameref_vocab_score.restype = None
        self.h = lib.frameref_vocab_load(os.fsencode(path))
        if not self.h:
            raise RuntimeError("loadFromTextFile failed: " + str(path))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.frameref_vocab_free(self.h)
            self.h = None

    def info(self):
        out = np.zeros(6, np.int32)
        self.L.frameref_vocab_info(self.h, _ptr(out))
        return tuple(int(x) for x in out)

    def compute_bow(self, desc):
        """Frame::ComputeBoW on a frame holding these descriptors"""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        out = _bow_outputs(len(d))
        self.L.frameref_compute_bow(self.h, _ptr(d), len(d), *[_ptr(a) for a in out])
        return _bow_result(*out)

    def transform(self, desc, levelsup=4):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        out = _bow_outputs(len(d))
        self.L.frameref_vocab_transform(self.h, _ptr(d), len(d), levelsup, *[_ptr(a) for a in out])
        return _bow_result(*out)

    def score_csr(self, qw, qv, kw, kv, kstart):
        qw = np.ascontiguousarray(qw, np.uint32); qv = np.ascontiguousarray(qv, np.float64)
        kw = np.ascontiguousarray(kw, np.uint32); kv = np.ascontiguousarray(kv, np.float64)
        ks = np.ascontiguousarray(kstart, np.int32)
        out = np.zeros(max(len(ks) - 1, 1), np.float64)
        self.L.frameref_vocab_score(self.h, _ptr(qw), _ptr(qv), len(qw), _ptr(kw), _ptr(kv), _ptr(ks), len(ks) - 1, _ptr(out))
        return out[:len(ks) - 1].copy()


class DbArgs(C.Structure):
    """frameport_db_args (oracle/frame_port.h)"""
    _fields_ = [("n_kf", C.c_int), ("kw", C.c_void_p), ("kv", C.c_void_p), ("kstart", C.c_void_p), ("erased", C.c_void_p),
                ("qw", C.c_void_p), ("qv", C.c_void_p), ("nq", C.c_int), ("connected", C.c_void_p), ("neigh_start", C.c_void_p),
                ("neigh", C.c_void_p), ("prev_score", C.c_void_p), ("min_score", C.c_float), ("reloc", C.c_int)]


def db_args(kfs, erased, query, connected, neighbours, prev_score, min_score, reloc):
    """kfs: list of (words, values); neighbours: list of lists (ordered covisibility); -> (DbArgs, keep-alive)"""
    n = len(kfs)
    ks = np.zeros(n + 1, np.int32); ks[1:] = np.cumsum([len(k[0]) for k in kfs])
    kw = np.ascontiguousarray(np.concatenate([np.asarray(k[0], np.uint32) for k in kfs] + [np.zeros(0, np.uint32)]))
    kv = np.ascontiguousarray(np.concatenate([np.asarray(k[1], np.float64) for k in kfs] + [np.zeros(0, np.float64)]))
    er = np.ascontiguousarray(erased, np.uint8); co = np.ascontiguousarray(connected, np.uint8)
    qw = np.ascontiguousarray(query[0], np.uint32); qv = np.ascontiguousarray(query[1], np.float64)
    ns = np.zeros(n + 1, np.int32); ns[1:] = np.cumsum([len(x) for x in neighbours])
    ne = np.ascontiguousarray(np.concatenate([np.asarray(x, np.int32) for x in neighbours] + [np.zeros(0, np.int32)]).astype(np.int32))
    ps = np.ascontiguousarray(prev_score, np.float32)
    keep = (ks, kw, kv, er, co, qw, qv, ns, ne, ps)
    a = DbArgs(n, kw.ctypes.data, kv.ctypes.data, ks.ctypes.data, er.ctypes.data, qw.ctypes.data, qv.ctypes.data, len(qw),
               co.ctypes.data, ns.ctypes.data, ne.ctypes.data, ps.ctypes.data, float(min_score), int(reloc))
    return a, keep


def _db_detect(fn, n, *lead):
    marked = np.zeros(max(n, 1), np.int32); words = np.zeros(max(n, 1), np.int32); score = np.zeros(max(n, 1), np.float32)
    cand = np.zeros(max(n, 1), np.int32)
    nc = fn(*lead, _ptr(marked), _ptr(words), _ptr(score), _ptr(cand))
    return marked[:n].astype(bool), words[:n].copy(), score[:n].copy(), cand[:nc].copy()


def port_db_detect(args):
    lib = C.CDLL(PORT_SO)
    lib.frameport_db_detect.argtypes = [C.c_void_p] * 5
    return _db_detect(lib.frameport_db_detect, args.n_kf, C.byref(args))


def ref_db_detect(ref_vocab, args):
    """the reference's own KeyFrameDatabase.cc; ref_vocab: a RefVocabulary (sizes the inverted file, scores)"""
    lib = ref_vocab.L
    lib.frameref_db_detect.argtypes = [C.c_void_p] * 6
    return _db_detect(lib.frameref_db_detect, args.n_kf, ref_vocab.h, C.byref(args))


def port_update_normal_and_depth(pos, obs_start, centers, ref_center, ref_scale, last_scale, normal, max_dist, min_dist):
    """oracle/frame_port.cc:frameport_update_normal_and_depth"""
    lib = C.CDLL(PORT_SO)
    f = lib.frameport_update_normal_and_depth
    f.argtypes = [C.c_void_p] * 6 + [C.c_int] + [C.c_void_p] * 3
    f.restype = None
    pos = np.ascontiguousarray(pos, np.float32); os_ = np.ascontiguousarray(obs_start, np.int32)
    c = np.ascontiguousarray(centers, np.float32); rc = np.ascontiguousarray(ref_center, np.float32)
    rs = np.ascontiguousarray(ref_scale, np.float32); ls = np.ascontiguousarray(last_scale, np.float32)
    nv = np.array(normal, np.float32).copy(); mx = np.array(max_dist, np.float32).copy(); mn = np.array(min_dist, np.float32).copy()
    f(_ptr(pos), _ptr(os_), _ptr(c), _ptr(rc), _ptr(rs), _ptr(ls), len(os_) - 1, _ptr(nv), _ptr(mx), _ptr(mn))
    return nv, mx, mn
