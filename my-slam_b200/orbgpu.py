"""ctypes binding of include/orbgpu.h.  No CPU fallback: every compute call goes to liborbgpu.so and
raises OrbGpuError when the library or a CUDA device is missing."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
NUM_STAGES = 7
STAGE_NAMES = ("upload", "pyramid", "fast_cells", "octree", "blur", "describe", "download")


class OrbGpuError(RuntimeError):
    pass


class ProjectionArgs(C.Structure):
    """orbgpu_projection_args (include/orbgpu.h): the snapshot ORBmatcher::SearchByProjection(Frame, Frame) reads."""
    _fields_ = [("n_last", C.c_int32), ("world_pos", C.c_void_p), ("mp_desc", C.c_void_p), ("mp_valid", C.c_void_p),
                ("mp_observed", C.c_void_p), ("last_octave", C.c_void_p), ("last_angle", C.c_void_p), ("Tcw_last", C.c_void_p),
                ("n_cur", C.c_int32), ("cur_keys_un", C.c_void_p), ("cur_desc", C.c_void_p), ("cur_u_right", C.c_void_p),
                ("cur_occupied", C.c_void_p), ("Tcw_cur", C.c_void_p),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float), ("b", C.c_float),
                ("bounds", C.c_float * 4), ("scale_factors", C.c_void_p), ("nlevels", C.c_int32), ("th", C.c_float),
                ("mono", C.c_int32), ("check_orientation", C.c_int32)]


class LocalPointsArgs(C.Structure):
    """orbgpu_local_points_args (include/orbgpu.h): what ORBmatcher::SearchByProjection(Frame, vector<MapPoint*>, th) reads."""
    _fields_ = [("n_points", C.c_int32), ("in_view", C.c_void_p), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p),
                ("proj_xr", C.c_void_p), ("scale_level", C.c_void_p), ("view_cos", C.c_void_p), ("mp_desc", C.c_void_p),
                ("mp_observed", C.c_void_p), ("n_cur", C.c_int32), ("cur_keys_un", C.c_void_p), ("cur_desc", C.c_void_p),
                ("cur_u_right", C.c_void_p), ("cur_occupied", C.c_void_p), ("bounds", C.c_float * 4),
                ("scale_factors", C.c_void_p), ("nlevels", C.c_int32), ("th", C.c_float), ("nnratio", C.c_float)]


def local_points_args(in_view, proj_x, proj_y, proj_xr, scale_level, view_cos, mp_desc, mp_observed, cur_keys_un, cur_desc,
                      cur_u_right, cur_occupied, bounds, scale_factors, th, nnratio=0.8):
    """Builds a LocalPointsArgs; returns (struct, keepalive dict of the arrays it points into)."""
    f32, u8, i32 = np.float32, np.uint8, np.int32
    arrs = dict(in_view=np.ascontiguousarray(in_view, u8), proj_x=np.ascontiguousarray(proj_x, f32),
                proj_y=np.ascontiguousarray(proj_y, f32), proj_xr=np.ascontiguousarray(proj_xr, f32),
                scale_level=np.ascontiguousarray(scale_level, i32), view_cos=np.ascontiguousarray(view_cos, f32),
                mp_desc=np.ascontiguousarray(mp_desc, u8), mp_observed=np.ascontiguousarray(mp_observed, u8),
                cur_keys_un=np.ascontiguousarray(cur_keys_un), cur_desc=np.ascontiguousarray(cur_desc, u8),
                cur_u_right=np.ascontiguousarray(cur_u_right, f32), cur_occupied=np.ascontiguousarray(cur_occupied, u8),
                scale_factors=np.ascontiguousarray(scale_factors, f32))
    a = LocalPointsArgs()
    for k, v in arrs.items():
        setattr(a, k, v.ctypes.data)
    a.n_points = len(arrs["in_view"]); a.n_cur = len(arrs["cur_keys_un"])
    a.bounds = (C.c_float * 4)(*[float(v) for v in bounds])
    a.nlevels = len(arrs["scale_factors"]); a.th = float(th); a.nnratio = float(nnratio)
    return a, arrs


class TriangulationSide(C.Structure):
    """orbgpu_triangulation_side (include/orbgpu.h)."""
    _fields_ = [("keys_un", C.c_void_p), ("desc", C.c_void_p), ("n", C.c_int32), ("has_map_point", C.c_void_p),
                ("u_right", C.c_void_p), ("node_ids", C.c_void_p), ("node_start", C.c_void_p), ("indices", C.c_void_p),
                ("n_nodes", C.c_int32)]


class TriangulationArgs(C.Structure):
    """orbgpu_triangulation_args (include/orbgpu.h): what ORBmatcher::SearchForTriangulation reads."""
    _fields_ = [("kf1", TriangulationSide), ("kf2", TriangulationSide), ("F12", C.c_float * 9), ("Cw1", C.c_float * 3),
                ("T2w", C.c_float * 16), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("scale_factors2", C.c_void_p), ("level_sigma2", C.c_void_p), ("nlevels", C.c_int32),
                ("only_stereo", C.c_int32), ("check_orientation", C.c_int32)]


def triangulation_args(side1, side2, F12, Cw1, T2w, cam, scale_factors2, level_sigma2, only_stereo=False, check_orientation=True):
    """side = (keys_un, desc, has_map_point, u_right, node_ids, node_start, indices); returns (struct, keepalive list)."""
    f32, u8 = np.float32, np.uint8
    a = TriangulationArgs(); keep = []
    for dst, sd in ((a.kf1, side1), (a.kf2, side2)):
        arrs = [np.ascontiguousarray(sd[0]), np.ascontiguousarray(sd[1], u8), np.ascontiguousarray(sd[2], u8),
                np.ascontiguousarray(sd[3], f32), np.ascontiguousarray(sd[4], np.uint32), np.ascontiguousarray(sd[5], np.int32),
                np.ascontiguousarray(sd[6], np.uint32)]
        keep.append(arrs)
        dst.keys_un, dst.desc, dst.has_map_point, dst.u_right, dst.node_ids, dst.node_start, dst.indices = [v.ctypes.data for v in arrs]
        dst.n = len(arrs[0]); dst.n_nodes = len(arrs[4])
    a.F12 = (C.c_float * 9)(*[float(v) for v in np.asarray(F12, f32).reshape(9)])
    a.Cw1 = (C.c_float * 3)(*[float(v) for v in np.asarray(Cw1, f32).reshape(3)])
    a.T2w = (C.c_float * 16)(*[float(v) for v in np.asarray(T2w, f32).reshape(16)])
    a.fx, a.fy, a.cx, a.cy = [float(v) for v in cam]
    sc = np.ascontiguousarray(scale_factors2, f32); sg = np.ascontiguousarray(level_sigma2, f32)
    keep.append([sc, sg])
    a.scale_factors2 = sc.ctypes.data; a.level_sigma2 = sg.ctypes.data; a.nlevels = len(sc)
    a.only_stereo = int(only_stereo); a.check_orientation = int(check_orientation)
    return a, keep


class LocalMapArgs(C.Structure):
    """orbgpu_local_map_args (include/orbgpu.h): Frame::isInFrustum per candidate point + the local-map search."""
    _fields_ = [("n_points", C.c_int32), ("candidate", C.c_void_p), ("world_pos", C.c_void_p), ("normal", C.c_void_p),
                ("max_distance", C.c_void_p), ("min_distance", C.c_void_p), ("mp_desc", C.c_void_p), ("mp_observed", C.c_void_p),
                ("Tcw", C.c_void_p), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float),
                ("viewing_cos_limit", C.c_float), ("log_scale_factor", C.c_float),
                ("n_cur", C.c_int32), ("cur_keys_un", C.c_void_p), ("cur_desc", C.c_void_p), ("cur_u_right", C.c_void_p),
                ("cur_occupied", C.c_void_p), ("bounds", C.c_float * 4), ("scale_factors", C.c_void_p), ("nlevels", C.c_int32),
                ("th", C.c_float), ("nnratio", C.c_float)]


class TrackFields(C.Structure):
    """orbgpu_track_fields (include/orbgpu.h)."""
    _fields_ = [("in_view", C.c_void_p), ("proj_x", C.c_void_p), ("proj_y", C.c_void_p), ("proj_xr", C.c_void_p),
                ("scale_level", C.c_void_p), ("view_cos", C.c_void_p)]


def track_fields(n):
    """Output arrays for n map points -> (struct, dict of numpy arrays)."""
    n = max(int(n), 1)
    arrs = dict(in_view=np.zeros(n, np.uint8), proj_x=np.zeros(n, np.float32), proj_y=np.zeros(n, np.float32),
                proj_xr=np.zeros(n, np.float32), scale_level=np.zeros(n, np.int32), view_cos=np.zeros(n, np.float32))
    t = TrackFields()
    for k, v in arrs.items():
        setattr(t, k, v.ctypes.data)
    return t, arrs


def local_map_args(candidate, world_pos, normal, max_distance, min_distance, mp_desc, mp_observed, Tcw, cam, bf, cur_keys_un, cur_desc,
                   cur_u_right, cur_occupied, bounds, scale_factors, log_scale_factor, th, nnratio=0.8, viewing_cos_limit=0.5):
    """Builds a LocalMapArgs; returns (struct, keepalive dict of the arrays it points into)."""
    f32, u8 = np.float32, np.uint8
    arrs = dict(candidate=np.ascontiguousarray(candidate, u8), world_pos=np.ascontiguousarray(world_pos, f32),
                normal=np.ascontiguousarray(normal, f32), max_distance=np.ascontiguousarray(max_distance, f32),
                min_distance=np.ascontiguousarray(min_distance, f32), mp_desc=np.ascontiguousarray(mp_desc, u8),
                mp_observed=np.ascontiguousarray(mp_observed, u8), Tcw=np.ascontiguousarray(Tcw, f32),
                cur_keys_un=np.ascontiguousarray(cur_keys_un), cur_desc=np.ascontiguousarray(cur_desc, u8),
                cur_u_right=np.ascontiguousarray(cur_u_right, f32), cur_occupied=np.ascontiguousarray(cur_occupied, u8),
                scale_factors=np.ascontiguousarray(scale_factors, f32))
    a = LocalMapArgs()
    for k, v in arrs.items():
        setattr(a, k, v.ctypes.data)
    a.n_points = len(arrs["candidate"]); a.n_cur = len(arrs["cur_keys_un"])
    a.fx, a.fy, a.cx, a.cy = [float(v) for v in cam]
    a.bf = float(bf); a.viewing_cos_limit = float(viewing_cos_limit); a.log_scale_factor = float(log_scale_factor)
    a.bounds = (C.c_float * 4)(*[float(v) for v in bounds])
    a.nlevels = len(arrs["scale_factors"]); a.th = float(th); a.nnratio = float(nnratio)
    return a, arrs


class FuseArgs(C.Structure):
    """orbgpu_fuse_args (include/orbgpu.h): what the loop body of ORBmatcher::Fuse(pKF, vpMapPoints, th) reads."""
    _fields_ = [("n_points", C.c_int32), ("present", C.c_void_p), ("world_pos", C.c_void_p), ("normal", C.c_void_p),
                ("max_distance", C.c_void_p), ("min_distance", C.c_void_p), ("mp_desc", C.c_void_p), ("Tcw", C.c_void_p),
                ("Ow", C.c_float * 3), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("bf", C.c_float),
                ("log_scale_factor", C.c_float), ("n_kf", C.c_int32), ("keys_un", C.c_void_p), ("desc", C.c_void_p),
                ("u_right", C.c_void_p), ("bounds", C.c_float * 4), ("scale_factors", C.c_void_p), ("inv_level_sigma2", C.c_void_p),
                ("nlevels", C.c_int32), ("th", C.c_float)]


def fuse_args(present, world_pos, normal, max_distance, min_distance, mp_desc, Tcw, Ow, cam, bf, keys_un, desc, u_right, bounds,
              scale_factors, inv_level_sigma2, log_scale_factor, th=3.0):
    """Builds a FuseArgs; returns (struct, keepalive dict of the arrays it points into)."""
    f32, u8 = np.float32, np.uint8
    arrs = dict(present=np.ascontiguousarray(present, u8), world_pos=np.ascontiguousarray(world_pos, f32),
                normal=np.ascontiguousarray(normal, f32), max_distance=np.ascontiguousarray(max_distance, f32),
                min_distance=np.ascontiguousarray(min_distance, f32), mp_desc=np.ascontiguousarray(mp_desc, u8),
                Tcw=np.ascontiguousarray(Tcw, f32), keys_un=np.ascontiguousarray(keys_un), desc=np.ascontiguousarray(desc, u8),
                u_right=np.ascontiguousarray(u_right, f32), scale_factors=np.ascontiguousarray(scale_factors, f32),
                inv_level_sigma2=np.ascontiguousarray(inv_level_sigma2, f32))
    a = FuseArgs()
    for k, v in arrs.items():
        setattr(a, k, v.ctypes.data)
    a.n_points = len(arrs["present"]); a.n_kf = len(arrs["keys_un"])
    a.Ow = (C.c_float * 3)(*[float(v) for v in np.asarray(Ow, f32).reshape(3)])
    a.fx, a.fy, a.cx, a.cy = [float(v) for v in cam]
    a.bf = float(bf); a.log_scale_factor = float(log_scale_factor)
    a.bounds = (C.c_float * 4)(*[float(v) for v in bounds])
    a.nlevels = len(arrs["scale_factors"]); a.th = float(th)
    return a, arrs


class RelocProjectionArgs(C.Structure):
    """orbgpu_reloc_projection_args (include/orbgpu.h): SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)."""
    _fields_ = [("n_kf", C.c_int32), ("mp_valid", C.c_void_p), ("world_pos", C.c_void_p), ("max_distance", C.c_void_p),
                ("min_distance", C.c_void_p), ("mp_desc", C.c_void_p), ("kf_angle", C.c_void_p),
                ("n_cur", C.c_int32), ("cur_keys_un", C.c_void_p), ("cur_desc", C.c_void_p), ("cur_occupied", C.c_void_p),
                ("Tcw_cur", C.c_void_p), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("bounds", C.c_float * 4), ("scale_factors", C.c_void_p), ("log_scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("th", C.c_float), ("orb_dist", C.c_int32), ("check_orientation", C.c_int32)]


def reloc_projection_args(mp_valid, world_pos, max_distance, min_distance, mp_desc, kf_angle, cur_keys_un, cur_desc, cur_occupied,
                          Tcw_cur, cam, bounds, scale_factors, log_scale_factor, th, orb_dist, check_orientation=True):
    """Builds a RelocProjectionArgs; returns (struct, keepalive dict)."""
    f32, u8 = np.float32, np.uint8
    arrs = dict(mp_valid=np.ascontiguousarray(mp_valid, u8), world_pos=np.ascontiguousarray(world_pos, f32),
                max_distance=np.ascontiguousarray(max_distance, f32), min_distance=np.ascontiguousarray(min_distance, f32),
                mp_desc=np.ascontiguousarray(mp_desc, u8), kf_angle=np.ascontiguousarray(kf_angle, f32),
                cur_keys_un=np.ascontiguousarray(cur_keys_un), cur_desc=np.ascontiguousarray(cur_desc, u8),
                cur_occupied=np.ascontiguousarray(cur_occupied, u8), Tcw_cur=np.ascontiguousarray(Tcw_cur, f32),
                scale_factors=np.ascontiguousarray(scale_factors, f32))
    a = RelocProjectionArgs()
    for k, v in arrs.items():
        setattr(a, k, v.ctypes.data)
    a.n_kf = len(arrs["mp_valid"]); a.n_cur = len(arrs["cur_keys_un"])
    a.fx, a.fy, a.cx, a.cy = [float(v) for v in cam]
    a.bounds = (C.c_float * 4)(*[float(v) for v in bounds])
    a.log_scale_factor = float(log_scale_factor); a.nlevels = len(arrs["scale_factors"]); a.th = float(th)
    a.orb_dist = int(orb_dist); a.check_orientation = int(check_orientation)
    return a, arrs


class BowSide(C.Structure):
    """orbgpu_bow_side (include/orbgpu.h): descriptors, validity, angles and the flattened DBoW2 FeatureVector of one side."""
    _fields_ = [("desc", C.c_void_p), ("n", C.c_int32), ("valid", C.c_void_p), ("angle", C.c_void_p),
                ("node_ids", C.c_void_p), ("node_start", C.c_void_p), ("indices", C.c_void_p), ("n_nodes", C.c_int32)]


def bow_side(desc, valid, angle, node_ids, node_start, indices):
    """Builds a BowSide; returns (struct, keepalive dict)."""
    arrs = dict(desc=np.ascontiguousarray(desc, np.uint8), valid=np.ascontiguousarray(valid, np.uint8),
                angle=np.ascontiguousarray(angle, np.float32), node_ids=np.ascontiguousarray(node_ids, np.uint32),
                node_start=np.ascontiguousarray(node_start, np.int32), indices=np.ascontiguousarray(indices, np.uint32))
    s = BowSide()
    for k, v in arrs.items():
        setattr(s, k, v.ctypes.data)
    s.n = len(arrs["valid"]); s.n_nodes = len(arrs["node_ids"])
    return s, arrs


def search_by_bow(side1, side2, nnratio=0.7, check_orientation=True, strict_th_low=False, device=0):
    """ORBmatcher::SearchByBoW (src/ORBmatcher.cc:159-288, 522-655) -> (nmatches, match12[n1], match21[n2])."""
    m12 = np.zeros(max(side1.n, 1), np.int32); m21 = np.zeros(max(side2.n, 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_by_bow(device, C.byref(side1), C.byref(side2), float(nnratio), int(check_orientation),
                                      int(strict_th_low), _p(m12), _p(m21), C.byref(nm)))
    return nm.value, m12[:side1.n].copy(), m21[:side2.n].copy()


class KfData(C.Structure):
    """orbgpu_kf_data (include/orbgpu.h): what a KeyFrame never changes after construction / ComputeBoW."""
    _fields_ = [("n", C.c_int32), ("keys_un", C.c_void_p), ("desc", C.c_void_p), ("u_right", C.c_void_p),
                ("node_ids", C.c_void_p), ("node_start", C.c_void_p), ("indices", C.c_void_p), ("n_nodes", C.c_int32),
                ("scale_factors", C.c_void_p), ("level_sigma2", C.c_void_p), ("nlevels", C.c_int32),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float)]


class KfTriangulationJob(C.Structure):
    """orbgpu_kf_triangulation_job (include/orbgpu.h)."""
    _fields_ = [("id1", C.c_int32), ("id2", C.c_int32), ("has_map_point1", C.c_void_p), ("has_map_point2", C.c_void_p),
                ("F12", C.c_float * 9), ("Cw1", C.c_float * 3), ("T2w", C.c_float * 16), ("only_stereo", C.c_int32),
                ("check_orientation", C.c_int32)]


class KeyFrameStore:
    """Keyframes resident in HBM (orbgpu_kfstore): add once, search by id."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        self.device = device
        _check(lib().orbgpu_kfstore_create(C.byref(self._h), device))

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().orbgpu_kfstore_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __len__(self):
        return _check(lib().orbgpu_kfstore_size(self._h))

    def add(self, keys_un, desc, u_right, node_ids, node_start, indices, scale_factors, level_sigma2, cam=(0, 0, 0, 0)):
        """-> id.  u_right None = monocular."""
        f32 = np.float32
        k = np.ascontiguousarray(keys_un, KP_DTYPE); d = np.ascontiguousarray(desc, np.uint8)
        ur = None if u_right is None else np.ascontiguousarray(u_right, f32)
        ni = np.ascontiguousarray(node_ids, np.uint32); ns = np.ascontiguousarray(node_start, np.int32)
        ix = np.ascontiguousarray(indices, np.uint32)
        sc = np.ascontiguousarray(scale_factors, f32); sg = np.ascontiguousarray(level_sigma2, f32)
        a = KfData()
        a.n = len(k); a.keys_un = k.ctypes.data; a.desc = d.ctypes.data; a.u_right = None if ur is None else ur.ctypes.data
        a.node_ids = ni.ctypes.data; a.node_start = ns.ctypes.data; a.indices = ix.ctypes.data; a.n_nodes = len(ni)
        a.scale_factors = sc.ctypes.data; a.level_sigma2 = sg.ctypes.data; a.nlevels = len(sc)
        a.fx, a.fy, a.cx, a.cy = [float(v) for v in cam]
        kid = C.c_int32(-1)
        _check(lib().orbgpu_kf_add(self._h, C.byref(a), C.byref(kid)))
        if self._sizes is None:
            self._sizes = {}
        self._sizes[kid.value] = a.n
        return kid.value

    def erase(self, kid):
        _check(lib().orbgpu_kf_erase(self._h, int(kid)))

    def search_by_bow(self, id1, valid1, id2=-1, valid2=None, frame2=None, nnratio=0.7, check_orientation=True,
                      strict_th_low=False):
        """SearchByBoW with keyframe id1 against stored keyframe id2 (valid2) or a Frame's BowSide (frame2)
        -> (nmatches, match12[n1], match21[n2])."""
        v1 = np.ascontiguousarray(valid1, np.uint8)
        v2 = None if valid2 is None else np.ascontiguousarray(valid2, np.uint8)
        n2 = frame2.n if frame2 is not None else (0 if v2 is None else len(v2))
        m12 = np.zeros(max(len(v1), 1), np.int32); m21 = np.zeros(max(n2, 1), np.int32); nm = C.c_int(0)
        _check(lib().orbgpu_kf_search_by_bow(self._h, int(id1), _p(v1), int(id2), None if v2 is None else _p(v2),
                                             None if frame2 is None else C.addressof(frame2), float(nnratio),
                                             int(check_orientation), int(strict_th_low), _p(m12), _p(m21), C.byref(nm)))
        return nm.value, m12[:len(v1)].copy(), m21[:n2].copy()

    @staticmethod
    def triangulation_job(id1, has_mp1, id2, has_mp2, F12, Cw1, T2w, only_stereo=False, check_orientation=True):
        """-> (struct, keepalive)."""
        f32 = np.float32
        h1 = np.ascontiguousarray(has_mp1, np.uint8); h2 = np.ascontiguousarray(has_mp2, np.uint8)
        j = KfTriangulationJob()
        j.id1 = int(id1); j.id2 = int(id2); j.has_map_point1 = h1.ctypes.data; j.has_map_point2 = h2.ctypes.data
        j.F12 = (C.c_float * 9)(*[float(v) for v in np.asarray(F12, f32).reshape(9)])
        j.Cw1 = (C.c_float * 3)(*[float(v) for v in np.asarray(Cw1, f32).reshape(3)])
        j.T2w = (C.c_float * 16)(*[float(v) for v in np.asarray(T2w, f32).reshape(16)])
        j.only_stereo = int(only_stereo); j.check_orientation = int(check_orientation)
        return j, (h1, h2)

    def search_for_triangulation(self, job):
        """-> (nmatches, vMatches12[n1])."""
        m12 = np.empty(max(self._sizes[job.id1], 1), np.int32); nm = C.c_int(0)
        _check(lib().orbgpu_kf_search_for_triangulation(self._h, C.byref(job), _p(m12), C.byref(nm)))
        return nm.value, m12[:self._sizes[job.id1]]

    def search_for_triangulation_batch(self, jobs):
        n = len(jobs)
        if n == 0:
            return []
        arr = (KfTriangulationJob * n)(*jobs)
        sizes = [self._n_of(j) for j in jobs]
        m12 = np.zeros(max(sum(sizes), 1), np.int32); nm = np.zeros(n, np.int32)
        _check(lib().orbgpu_kf_search_for_triangulation_batch(self._h, C.cast(arr, C.c_void_p), n, _p(m12), _p(nm)))
        out, o = [], 0
        for k, sz in enumerate(sizes):
            out.append((int(nm[k]), m12[o:o + sz].copy())); o += sz
        return out

    def _n_of(self, job):
        return self._sizes[job.id1]

    _sizes = None                      # id -> N of the stored keyframe (output sizes of the triangulation search)


def projection_args(world_pos, mp_desc, mp_valid, mp_observed, last_octave, last_angle, Tcw_last, cur_keys_un, cur_desc,
                    cur_u_right, cur_occupied, Tcw_cur, cam, bf, b, bounds, scale_factors, th, mono, check_orientation=True):
    """Builds a ProjectionArgs; returns (struct, keepalive list of the arrays it points into)."""
    f32, u8, i32 = np.float32, np.uint8, np.int32
    arrs = dict(world_pos=np.ascontiguousarray(world_pos, f32), mp_desc=np.ascontiguousarray(mp_desc, u8),
                mp_valid=np.ascontiguousarray(mp_valid, u8), mp_observed=np.ascontiguousarray(mp_observed, u8),
                last_octave=np.ascontiguousarray(last_octave, i32), last_angle=np.ascontiguousarray(last_angle, f32),
                Tcw_last=np.ascontiguousarray(Tcw_last, f32), cur_keys_un=np.ascontiguousarray(cur_keys_un),
                cur_desc=np.ascontiguousarray(cur_desc, u8), cur_u_right=np.ascontiguousarray(cur_u_right, f32),
                cur_occupied=np.ascontiguousarray(cur_occupied, u8), Tcw_cur=np.ascontiguousarray(Tcw_cur, f32),
                scale_factors=np.ascontiguousarray(scale_factors, f32))
    a = ProjectionArgs()
    for k, v in arrs.items():
        setattr(a, k, v.ctypes.data)
    a.n_last = len(arrs["last_octave"]); a.n_cur = len(arrs["cur_keys_un"])
    a.fx, a.fy, a.cx, a.cy = [float(v) for v in cam]
    a.bf, a.b, a.th = float(bf), float(b), float(th)
    a.bounds = (C.c_float * 4)(*[float(v) for v in bounds])
    a.nlevels = len(arrs["scale_factors"]); a.mono = int(mono); a.check_orientation = int(check_orientation)
    return a, arrs


def lib_path():
    return os.path.join(_HERE, "liborbgpu.so")


_lib = None


def lib():
    """Loads liborbgpu.so (built in-tree by csrc/Makefile); fails loudly when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise OrbGpuError("%s not built: run `python -c 'import __graft_entry__ as g; g.build()'`" % p)
    L = C.CDLL(p)
    vp, i32, sz, f32, u64 = C.c_void_p, C.c_int, C.c_size_t, C.c_float, C.c_uint64
    pi = C.POINTER(C.c_int)
    sig = {
        "orbgpu_create": (i32, [C.POINTER(vp), i32, i32, f32, i32, i32, i32]),
        "orbgpu_destroy": (None, [vp]),
        "orbgpu_get_tables": (i32, [vp, vp, vp, vp]),
        "orbgpu_max_keypoints": (i32, [vp, i32, i32]),
        "orbgpu_extract": (i32, [vp, vp, i32, i32, sz, vp, vp, i32, pi]),
        "orbgpu_extract_batch": (i32, [vp, vp, i32, i32, i32, sz, sz, vp, vp, i32, vp]),
        "orbgpu_extract_batch_device": (i32, [vp, vp, i32, i32, i32, sz, sz, vp, vp, i32, vp]),
        "orbgpu_extract_batch_async": (i32, [vp, vp, i32, i32, i32, sz, sz, vp, vp, i32, vp]),
        "orbgpu_wait": (i32, [vp]),
        "orbgpu_host_alloc": (i32, [C.POINTER(vp), sz, i32]),
        "orbgpu_host_free": (None, [vp]),
        "orbgpu_extract_batch_multi": (i32, [vp, i32, vp, i32, i32, i32, sz, sz, vp, vp, i32, vp]),
        "orbgpu_stereo_match": (i32, [vp, vp, f32, f32, vp, vp, i32, pi]),
        "orbgpu_set_input_format": (i32, [vp, i32, i32]),
        "orbgpu_submit": (i32, [vp, vp, i32, i32, sz, C.POINTER(u64)]),
        "orbgpu_collect": (i32, [vp, u64, vp, vp, i32, pi]),
        "orbgpu_queue_depth": (i32, [vp]),
        "orbgpu_pyramid_level": (i32, [vp, i32, vp, sz, pi, pi]),
        "orbgpu_debug_level_blurred": (i32, [vp, i32, vp, sz]),
        "orbgpu_debug_level_candidates": (i32, [vp, i32, vp, vp, vp, i32]),
        "orbgpu_debug_level_selected": (i32, [vp, i32, vp, vp, vp, i32]),
        "orbgpu_octree": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, vp, i32]),
        "orbgpu_debug_atan2": (i32, [vp, vp, vp, i32, vp]),
        "orbgpu_debug_sincosf": (i32, [vp, vp, i32, vp, vp]),
        "orbgpu_debug_sincos_checksum": (i32, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(u64)]),
        "orbgpu_set_profiling": (i32, [vp, i32]),
        "orbgpu_get_stage_ms": (i32, [vp, vp, i32, i32]),
        "orbgpu_launch_count": (u64, [vp]),
        "orbgpu_last_batch_ms": (C.c_double, [vp]),
        "orbgpu_set_batching": (i32, [vp, i32, i32]),
        "orbgpu_set_graphs": (i32, [vp, i32]),
        "orbgpu_chunk_capacity": (i32, [vp, i32, i32]),
        "orbgpu_undistort_keypoints": (i32, [vp, vp, vp, i32, vp, i32, pi]),
        "orbgpu_undistort_points": (i32, [i32, vp, i32, vp, vp, i32, vp]),
        "orbgpu_extract_undistorted": (i32, [vp, vp, i32, i32, sz, vp, vp, i32, vp, vp, vp, i32, pi]),
        "orbgpu_search_for_initialization": (i32, [i32, vp, vp, i32, vp, vp, i32, vp, vp, i32, f32, i32, vp, pi]),
        "orbgpu_search_by_projection": (i32, [i32, vp, vp, pi]),
        "orbgpu_search_local_points": (i32, [i32, vp, vp, pi]),
        "orbgpu_search_by_bow": (i32, [i32, vp, vp, f32, i32, i32, vp, vp, pi]),
        "orbgpu_search_local_map": (i32, [i32, vp, vp, pi, vp, pi]),
        "orbgpu_fuse_search": (i32, [i32, vp, vp]),
        "orbgpu_fuse_search_sim3": (i32, [i32, vp, vp, vp]),
        "orbgpu_search_by_projection_sim3": (i32, [i32, vp, vp, vp, vp, pi]),
        "orbgpu_search_by_sim3": (i32, [i32, vp, vp, f32, vp, vp, vp, pi]),
        "orbgpu_search_by_projection_keyframe": (i32, [i32, vp, vp, pi]),
        "orbgpu_debug_logf": (i32, [i32, vp, i32, vp]),
        "orbgpu_search_for_triangulation": (i32, [i32, vp, vp, pi]),
        "orbgpu_search_for_triangulation_batch": (i32, [i32, vp, i32, vp, vp]),
        "orbgpu_fuse_search_batch": (i32, [i32, vp, i32, vp]),
        "orbgpu_kfstore_create": (i32, [C.POINTER(vp), i32]),
        "orbgpu_kfstore_destroy": (None, [vp]),
        "orbgpu_kfstore_size": (i32, [vp]),
        "orbgpu_kf_add": (i32, [vp, vp, C.POINTER(C.c_int32)]),
        "orbgpu_kf_erase": (i32, [vp, i32]),
        "orbgpu_kf_search_by_bow": (i32, [vp, i32, vp, i32, vp, vp, f32, i32, i32, vp, vp, pi]),
        "orbgpu_kf_search_for_triangulation": (i32, [vp, vp, vp, pi]),
        "orbgpu_kf_search_for_triangulation_batch": (i32, [vp, vp, i32, vp, vp]),
        "orbgpu_hamming256": (i32, [vp, vp]),
        "orbgpu_distinctive_descriptors": (i32, [i32, vp, vp, i32, vp]),
        "orbgpu_vocabulary_create": (i32, [C.POINTER(vp), i32, i32, i32, i32, i32, i32, vp, vp, vp, vp]),
        "orbgpu_vocabulary_load_text": (i32, [C.POINTER(vp), i32, C.c_char_p]),
        "orbgpu_vocabulary_destroy": (None, [vp]),
        "orbgpu_vocabulary_info": (i32, [vp, vp]),
        "orbgpu_vocabulary_transform": (i32, [vp, vp, i32, i32, vp, vp, vp, vp, vp, vp]),
        "orbgpu_vocabulary_transform_batch": (i32, [vp, vp, vp, i32, i32, vp, vp, vp, vp, vp, vp]),
        "orbgpu_vocabulary_score": (i32, [vp, vp, vp, i32, vp, vp, vp, i32, vp]),
        "orbgpu_bowdb_create": (i32, [C.POINTER(vp), i32]),
        "orbgpu_bowdb_destroy": (None, [vp]),
        "orbgpu_bowdb_add": (i32, [vp, vp, vp, i32, pi]),
        "orbgpu_bowdb_erase": (i32, [vp, i32]),
        "orbgpu_bowdb_clear": (i32, [vp]),
        "orbgpu_bowdb_size": (i32, [vp]),
        "orbgpu_bowdb_query": (i32, [vp, vp, vp, i32, vp, vp, vp]),
        "orbgpu_update_normals_and_depths": (i32, [i32, vp, vp, vp, vp, vp, vp, i32, vp, vp, vp]),
        "orbgpu_match_bruteforce": (i32, [i32, vp, i32, vp, i32, vp, vp, vp]),
        "orbgpu_match_bruteforce_noalloc": (i32, [i32, vp, i32, vp, i32, vp, vp, vp]),
        "orbgpu_debug_pipe_peak": (i32, [i32, i32, C.POINTER(C.c_double)]),
        "orbgpu_matcher_create": (i32, [C.POINTER(vp), i32]),
        "orbgpu_matcher_destroy": (None, [vp]),
        "orbgpu_matcher_set_keyframes": (i32, [vp, vp, i32, i32]),
        "orbgpu_matcher_set_keyframes_device": (i32, [vp, vp, i32, i32]),
        "orbgpu_matcher_count_pairs": (i32, [vp, vp, vp, i32, i32, f32, vp]),
        "orbgpu_matcher_count_pairs_device": (i32, [vp, vp, i32, i32, f32, vp]),
        "orbgpu_matcher_last_kernel_ms": (C.c_double, [vp]),
        "orbgpu_comm_unique_id": (i32, [vp]),
        "orbgpu_comm_init_rank": (i32, [C.POINTER(vp), i32, i32, i32, vp]),
        "orbgpu_comm_init_all": (i32, [vp, vp, i32]),
        "orbgpu_comm_destroy": (None, [vp]),
        "orbgpu_matcher_all_pairs_sharded": (i32, [vp, vp, i32, f32, vp, vp]),
        "orbgpu_matcher_all_pairs_multi": (i32, [vp, vp, i32, i32, f32, vp, vp]),
        "orbgpu_strerror": (C.c_char_p, [i32]),
        "orbgpu_last_cuda_error": (C.c_char_p, []),
        "orbgpu_device_count": (i32, []),
        "orbgpu_version": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)           # AttributeError here = header/library mismatch
        f.restype = res
        f.argtypes = args
    L._orbgpu_symbols = tuple(sig)
    _lib = L
    return L


def _check(rc):
    if rc < 0:
        L = lib()
        msg = L.orbgpu_strerror(rc).decode()
        if rc == -3:
            msg += ": " + L.orbgpu_last_cuda_error().decode()
        raise OrbGpuError("orbgpu error %d: %s" % (rc, msg))
    return rc


def device_count():
    return lib().orbgpu_device_count()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class ORBextractor:
    """Mirror of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-111): same constructor
    arguments, __call__(image) == operator()(image, mask, keypoints, descriptors), same getters."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=0):
        self._L = lib()
        self._h = C.c_void_p()
        _check(self._L.orbgpu_create(C.byref(self._h), device, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST))
        self.nlevels = nlevels
        self.device = device
        s4 = np.zeros(4 * nlevels, np.float32)
        self._feats = np.zeros(nlevels, np.int32)
        self._umax = np.zeros(16, np.int32)
        _check(self._L.orbgpu_get_tables(self._h, _p(s4), _p(self._feats), _p(self._umax)))
        self._s4 = s4.reshape(4, nlevels)
        self._scale_arg = np.float32(scaleFactor)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbgpu_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    # getters, include/ORBextractor.h:63-83
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return float(self._scale_arg)
    def GetScaleFactors(self): return self._s4[0].copy()
    def GetInverseScaleFactors(self): return self._s4[1].copy()
    def GetScaleSigmaSquares(self): return self._s4[2].copy()
    def GetInverseScaleSigmaSquares(self): return self._s4[3].copy()
    def features_per_level(self): return self._feats.copy()
    def umax(self): return self._umax.copy()

    def max_keypoints(self, w, h):
        return _check(self._L.orbgpu_max_keypoints(self._h, w, h))

    def __call__(self, image, mask=None):
        """Returns (keypoints[KP_DTYPE], descriptors[N,32] uint8); (None, None) for an empty image
        (the reference returns without touching its outputs, src/ORBextractor.cc:1048-1049)."""
        if image is None or image.size == 0:
            return None, None
        ch = getattr(self, "_channels", 1)
        if image.dtype != np.uint8 or image.ndim != (2 if ch == 1 else 3) or (ch != 1 and image.shape[2] != ch):
            raise ValueError("image must be CV_8UC%d" % ch)          # reference: assert, src/ORBextractor.cc:1052
        if image.strides[-1] != 1 or (ch != 1 and image.strides[1] != ch):
            image = np.ascontiguousarray(image)
        h, w = image.shape[:2]
        cap = self._last_cap = self.max_keypoints(w, h)
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        _check(self._L.orbgpu_extract(self._h, _p(image), w, h, image.strides[0], _p(kps), _p(desc), cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_batch(self, frames, kps=None, desc=None, n_out=None):
        """frames: [n,h,w] uint8 host array (pinned or pageable).  Returns (kps[n,cap], desc[n,cap,32], n_out[n])."""
        n, h, w = frames.shape
        cap = self.max_keypoints(w, h)
        if kps is None:
            kps = np.zeros((n, cap), KP_DTYPE); desc = np.zeros((n, cap, 32), np.uint8); n_out = np.zeros(n, np.int32)
        _check(self._L.orbgpu_extract_batch(self._h, _p(frames), n, w, h, frames.strides[1], frames.strides[0],
                                            _p(kps), _p(desc), kps.shape[1], _p(n_out)))
        return kps, desc, n_out

    def extract_batch_ptr(self, frames_ptr, n, w, h, row_stride, frame_stride, kps_ptr, desc_ptr, cap, nout_ptr, device=False):
        f = self._L.orbgpu_extract_batch_device if device else self._L.orbgpu_extract_batch
        _check(f(self._h, frames_ptr, n, w, h, row_stride, frame_stride, kps_ptr, desc_ptr, cap, nout_ptr))

    def extract_batch_async_ptr(self, frames_ptr, n, w, h, row_stride, frame_stride, kps_ptr, desc_ptr, cap, nout_ptr):
        _check(self._L.orbgpu_extract_batch_async(self._h, frames_ptr, n, w, h, row_stride, frame_stride, kps_ptr, desc_ptr, cap, nout_ptr))

    def set_input_format(self, channels, rgb_order=False):
        _check(self._L.orbgpu_set_input_format(self._h, int(channels), int(bool(rgb_order))))
        self._channels = int(channels)

    def wait(self):
        _check(self._L.orbgpu_wait(self._h))

    def submit(self, image):
        image = np.ascontiguousarray(image, np.uint8)
        t = C.c_uint64(0)
        _check(self._L.orbgpu_submit(self._h, _p(image), image.shape[1], image.shape[0], image.strides[0], C.byref(t)))
        return t.value

    def collect(self, ticket, w, h):
        cap = self.max_keypoints(w, h)
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = C.c_int(0)
        _check(self._L.orbgpu_collect(self._h, ticket, _p(kps), _p(desc), cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def queue_depth(self):
        return _check(self._L.orbgpu_queue_depth(self._h))

    # Frame::UndistortKeyPoints, src/Frame.cc:404-435
    def undistort_keypoints(self, K4, dist):
        K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
        cap = getattr(self, "_last_cap", 1 << 16)
        kps = np.empty(cap, KP_DTYPE); n = C.c_int(0)
        _check(self._L.orbgpu_undistort_keypoints(self._h, _p(K4), _p(dist), len(dist), _p(kps), cap, C.byref(n)))
        return kps[:n.value].copy()

    def extract_undistorted(self, image, K4, dist):
        """ExtractORB + UndistortKeyPoints of the Frame constructor in one device round trip (orbgpu_extract_undistorted)
        -> (mvKeys, mvKeysUn, mDescriptors); gray images only."""
        if image is None or image.size == 0:
            return None, None, None
        if image.strides[-1] != 1:
            image = np.ascontiguousarray(image)
        K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
        h, w = image.shape[:2]
        cap = self._last_cap = self.max_keypoints(w, h)
        kps = np.zeros(cap, KP_DTYPE); un = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8); n = C.c_int(0)
        _check(self._L.orbgpu_extract_undistorted(self._h, _p(image), w, h, image.strides[0], _p(K4), _p(dist), len(dist),
                                                  _p(kps), _p(un), _p(desc), cap, C.byref(n)))
        return kps[:n.value].copy(), un[:n.value].copy(), desc[:n.value].copy()

    # mvImagePyramid, include/ORBextractor.h:85
    def pyramid_level(self, level, bordered=False):
        w, h = C.c_int(), C.c_int()
        _check(self._L.orbgpu_pyramid_level(self._h, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        _check(self._L.orbgpu_pyramid_level(self._h, level, _p(out), w.value, C.byref(w), C.byref(h)))
        return np.pad(out, 19, mode="reflect") if bordered else out

    def debug_blurred(self, level):
        w, h = C.c_int(), C.c_int()
        _check(self._L.orbgpu_pyramid_level(self._h, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        _check(self._L.orbgpu_debug_level_blurred(self._h, level, _p(out), w.value))
        return out

    def _cands(self, fn, level):
        cap = 1 << 16
        while True:
            x = np.zeros(cap, np.int32); y = np.zeros(cap, np.int32); s = np.zeros(cap, np.int32)
            n = _check(fn(self._h, level, _p(x), _p(y), _p(s), cap))
            if n <= cap:
                return np.stack([x[:n], y[:n], s[:n]], 1)
            cap = n

    def debug_candidates(self, level): return self._cands(self._L.orbgpu_debug_level_candidates, level)
    def debug_selected(self, level): return self._cands(self._L.orbgpu_debug_level_selected, level)

    def octree(self, xs, ys, resp, W, H, N):
        xs = np.ascontiguousarray(xs, np.int32); ys = np.ascontiguousarray(ys, np.int32); resp = np.ascontiguousarray(resp, np.int32)
        out = np.zeros(len(xs) + 8, np.int32)
        n = _check(self._L.orbgpu_octree(self._h, _p(xs), _p(ys), _p(resp), len(xs), W, H, N, _p(out), len(out)))
        return out[:n].copy()

    def debug_atan2(self, y, x):
        y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
        out = np.zeros(len(y), np.float32)
        _check(self._L.orbgpu_debug_atan2(self._h, _p(y), _p(x), len(y), _p(out)))
        return out

    def debug_sincosf(self, rad):
        rad = np.ascontiguousarray(rad, np.float32)
        s = np.zeros(len(rad), np.float32); c = np.zeros(len(rad), np.float32)
        _check(self._L.orbgpu_debug_sincosf(self._h, _p(rad), len(rad), _p(s), _p(c)))
        return s, c

    def debug_sincos_checksum(self, first_bits, count, step=1):
        h = C.c_uint64(0)
        _check(self._L.orbgpu_debug_sincos_checksum(self._h, first_bits, count, step, C.byref(h)))
        return h.value

    def set_profiling(self, on): _check(self._L.orbgpu_set_profiling(self._h, int(on)))

    def stage_ms(self, reset=True):
        ms = np.zeros(NUM_STAGES, np.float64)
        _check(self._L.orbgpu_get_stage_ms(self._h, _p(ms), NUM_STAGES, int(reset)))
        return dict(zip(STAGE_NAMES, ms.tolist()))

    def launch_count(self): return int(self._L.orbgpu_launch_count(self._h))

    def set_batching(self, streams, max_chunk_frames): _check(self._L.orbgpu_set_batching(self._h, int(streams), int(max_chunk_frames)))

    def chunk_capacity(self, w, h): return _check(self._L.orbgpu_chunk_capacity(self._h, w, h))

    def set_graphs(self, on): _check(self._L.orbgpu_set_graphs(self._h, int(bool(on))))

    def last_batch_ms(self): return float(self._L.orbgpu_last_batch_ms(self._h))


class HostBuffer:
    """Page-locked host memory from orbgpu_host_alloc as a numpy array (`.array`); write_combined for CPU-write-only inputs."""

    def __init__(self, shape, dtype=np.uint8, write_combined=False):
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._p = C.c_void_p()
        _check(lib().orbgpu_host_alloc(C.byref(self._p), n, int(bool(write_combined))))
        buf = (C.c_ubyte * n).from_address(self._p.value)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)
        self.ptr = self._p.value

    def close(self):
        if getattr(self, "_p", None) and self._p.value:
            self.array = None
            lib().orbgpu_host_free(self._p)
            self._p = C.c_void_p()

    __del__ = close


def extract_batch_multi(extractors, frames):
    """frames [n,h,w] uint8 through several extractor handles (one per GPU) in one call: orbgpu_extract_batch_multi."""
    n, h, w = frames.shape
    cap = extractors[0].max_keypoints(w, h)
    for e in extractors[1:]:
        assert e.max_keypoints(w, h) == cap
    kps = np.zeros((n, cap), KP_DTYPE); desc = np.zeros((n, cap, 32), np.uint8); n_out = np.zeros(n, np.int32)
    hs = (C.c_void_p * len(extractors))(*[e._h for e in extractors])
    _check(lib().orbgpu_extract_batch_multi(hs, len(extractors), _p(frames), n, w, h, frames.strides[1], frames.strides[0],
                                            _p(kps), _p(desc), cap, _p(n_out)))
    return kps, desc, n_out


def undistort_points(xy, K4, dist, device=0):
    """cv::undistortPoints(xy, K, D, noArray(), K) (src/Frame.cc:404-463); K4 = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3])."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    K4 = np.ascontiguousarray(K4, np.float32); dist = np.ascontiguousarray(dist, np.float32)
    out = np.zeros_like(xy)
    _check(lib().orbgpu_undistort_points(device, _p(xy), len(xy), _p(K4), _p(dist), len(dist), _p(out)))
    return out


def search_for_initialization(keys1_un, desc1, keys2_un, desc2, bounds, prev_matched, window_size=100, nnratio=0.9,
                              check_orientation=True, device=0):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) -> (nmatches, vnMatches12, updated vbPrevMatched)."""
    k1 = np.ascontiguousarray(keys1_un); k2 = np.ascontiguousarray(keys2_un)
    d1 = np.ascontiguousarray(desc1, np.uint8); d2 = np.ascontiguousarray(desc2, np.uint8)
    b4 = np.ascontiguousarray(bounds, np.float32)
    pm = np.array(prev_matched, np.float32).reshape(-1, 2).copy()
    m12 = np.zeros(max(len(k1), 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_for_initialization(device, _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(b4), _p(pm),
                                                  int(window_size), float(nnratio), int(check_orientation), _p(m12), C.byref(nm)))
    return nm.value, m12[:len(k1)].copy(), pm


def search_by_projection(args, device=0):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1328-1470) on a
    ProjectionArgs snapshot -> (nmatches, cur_match[n_cur])."""
    cm = np.zeros(max(args.n_cur, 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_by_projection(device, C.byref(args), _p(cm), C.byref(nm)))
    return nm.value, cm[:args.n_cur].copy()


def search_local_points(args, device=0):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:44-125) on a LocalPointsArgs snapshot
    -> (nmatches, cur_match[n_cur]): cur_match[idx] = i where F.mvpMapPoints[idx] was set to vpMapPoints[i], else -1."""
    cm = np.zeros(max(args.n_cur, 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_local_points(device, C.byref(args), _p(cm), C.byref(nm)))
    return nm.value, cm[:args.n_cur].copy()


def search_for_triangulation(args, device=0):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-820) on a TriangulationArgs snapshot
    -> (nmatches, vMatches12[kf1.n])."""
    m12 = np.zeros(max(args.kf1.n, 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_for_triangulation(device, C.byref(args), _p(m12), C.byref(nm)))
    return nm.value, m12[:args.kf1.n].copy()


def search_for_triangulation_batch(args_list, device=0):
    """SearchForTriangulation for several keyframe pairs in one call (LocalMapping::CreateNewMapPoints: one per neighbour)
    -> list of (nmatches, vMatches12)."""
    n = len(args_list)
    if n == 0:
        return []
    arr = (TriangulationArgs * n)(*args_list)
    sizes = [a.kf1.n for a in args_list]
    m12 = np.zeros(max(sum(sizes), 1), np.int32); nm = np.zeros(n, np.int32)
    _check(lib().orbgpu_search_for_triangulation_batch(device, C.cast(arr, C.c_void_p), n, _p(m12), _p(nm)))
    out, o = [], 0
    for k, sz in enumerate(sizes):
        out.append((int(nm[k]), m12[o:o + sz].copy())); o += sz
    return out


def search_local_map(args, device=0):
    """Tracking::SearchLocalPoints (src/Tracking.cc:1150-1200): isInFrustum for the candidate points + the local-map search
    -> (nmatches, cur_match[n_cur], n_in_view, dict of the mTrack* arrays)."""
    t, arrs = track_fields(args.n_points)
    cm = np.zeros(max(args.n_cur, 1), np.int32); nm = C.c_int(0); nv = C.c_int(0)
    _check(lib().orbgpu_search_local_map(device, C.byref(args), C.byref(t), C.byref(nv), _p(cm), C.byref(nm)))
    return nm.value, cm[:args.n_cur].copy(), nv.value, {k: v[:args.n_points] for k, v in arrs.items()}


def debug_logf(x, device=0):
    """glibc logf as the device evaluates it in MapPoint::PredictScale (test hook)."""
    x = np.ascontiguousarray(x, np.float32); out = np.zeros_like(x)
    _check(lib().orbgpu_debug_logf(device, _p(x), len(x), _p(out)))
    return out


def fuse_search(args, device=0):
    """The search of ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:825-957) for every map point at once
    -> best_idx[n_points]: the keypoint of pKF each point fuses with, or -1."""
    best = np.zeros(max(args.n_points, 1), np.int32)
    _check(lib().orbgpu_fuse_search(device, C.byref(args), _p(best)))
    return best[:args.n_points].copy()


def fuse_search_batch(args_list, device=0):
    """The search of Fuse for several (keyframe, point set) problems in one call (LocalMapping::SearchInNeighbors)
    -> list of best_idx arrays."""
    n = len(args_list)
    if n == 0:
        return []
    arr = (FuseArgs * n)(*args_list)
    sizes = [a.n_points for a in args_list]
    best = np.zeros(max(sum(sizes), 1), np.int32)
    _check(lib().orbgpu_fuse_search_batch(device, C.cast(arr, C.c_void_p), n, _p(best)))
    out, o = [], 0
    for sz in sizes:
        out.append(best[o:o + sz].copy()); o += sz
    return out


def search_by_projection_keyframe(args, device=0):
    """ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1472-1597)
    -> (nmatches, cur_match[n_cur])."""
    cm = np.zeros(max(args.n_cur, 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_by_projection_keyframe(device, C.byref(args), _p(cm), C.byref(nm)))
    return nm.value, cm[:args.n_cur].copy()


def distinctive_descriptors(desc, group_start, device=0):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-303) for many map points at once: desc (total x 32),
    group_start (n_groups + 1) -> BestIdx per group (-1 for an empty group)."""
    d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); gs = np.ascontiguousarray(group_start, np.int32)
    out = np.zeros(max(len(gs) - 1, 1), np.int32)
    _check(lib().orbgpu_distinctive_descriptors(device, _p(d) if len(d) else None, _p(gs), len(gs) - 1, _p(out)))
    return out[:len(gs) - 1].copy()


def fuse_search_sim3(args, Scw, device=0):
    """The search of ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:969-1097) -> best_idx[n_points]."""
    S = np.ascontiguousarray(Scw, np.float32).reshape(16); best = np.zeros(max(args.n_points, 1), np.int32)
    _check(lib().orbgpu_fuse_search_sim3(device, C.byref(args), _p(S), _p(best)))
    return best[:args.n_points].copy()


def search_by_projection_sim3(args, Scw, matched_in, device=0):
    """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:290-403) -> (nmatches, matched[n_kf])."""
    S = np.ascontiguousarray(Scw, np.float32).reshape(16); mi = np.ascontiguousarray(matched_in, np.uint8)
    out = np.zeros(max(args.n_kf, 1), np.int32); nm = C.c_int(0)
    _check(lib().orbgpu_search_by_projection_sim3(device, C.byref(args), _p(S), _p(mi), _p(out), C.byref(nm)))
    return nm.value, out[:args.n_kf].copy()


def search_by_sim3(a12, a21, s12, R12, t12, device=0):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1102-1326) -> (nFound, match12[N1])."""
    R = np.ascontiguousarray(R12, np.float32).reshape(9); t = np.ascontiguousarray(t12, np.float32).reshape(3)
    out = np.zeros(max(a12.n_points, 1), np.int32); nf = C.c_int(0)
    _check(lib().orbgpu_search_by_sim3(device, C.byref(a12), C.byref(a21), float(s12), _p(R), _p(t), _p(out), C.byref(nf)))
    return nf.value, out[:a12.n_points].copy()


def stereo_match(left, right, bf, b):
    """Frame::ComputeStereoMatches between the last frames of two ORBextractor handles -> (mvuRight, mvDepth)."""
    cap = getattr(left, "_last_cap", 1 << 16)
    u = np.empty(cap, np.float32); d = np.empty(cap, np.float32); n = C.c_int(0)
    _check(lib().orbgpu_stereo_match(left._h, right._h, bf, b, _p(u), _p(d), cap, C.byref(n)))
    return u[:n.value].copy(), d[:n.value].copy()


class ORBmatcher:
    """The Hamming part of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:37-102): the static
    DescriptorDistance, the constants, and the batched best/second-best scan + all-pairs counts."""
    TH_LOW = 50        # src/ORBmatcher.cc:38
    TH_HIGH = 100      # :37
    HISTO_LENGTH = 30  # :39

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio = np.float32(nnratio)
        self.mbCheckOrientation = checkOri
        self.device = device
        self._L = lib()
        self._m = None

    @staticmethod
    def DescriptorDistance(a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return lib().orbgpu_hamming256(_p(a), _p(b))

    def match_bruteforce(self, A, B):
        A = np.ascontiguousarray(A, np.uint8); B = np.ascontiguousarray(B, np.uint8)
        bi = np.zeros(len(A), np.int32); b1 = np.zeros(len(A), np.uint16); b2 = np.zeros(len(A), np.uint16)
        _check(self._L.orbgpu_match_bruteforce(self.device, _p(A), len(A), _p(B), len(B), _p(bi), _p(b1), _p(b2)))
        return bi, b1, b2

    def _matcher(self):
        if self._m is None:
            self._m = C.c_void_p()
            _check(self._L.orbgpu_matcher_create(C.byref(self._m), self.device))
        return self._m

    def set_keyframes(self, desc):
        """desc: [K, n, 32] uint8 host array."""
        desc = np.ascontiguousarray(desc, np.uint8)
        _check(self._L.orbgpu_matcher_set_keyframes(self._matcher(), _p(desc), desc.shape[0], desc.shape[1]))

    def set_keyframes_device(self, ptr, K, n):
        _check(self._L.orbgpu_matcher_set_keyframes_device(self._matcher(), ptr, K, n))

    def count_pairs(self, kf_a, kf_b, th_low=None):
        a = np.ascontiguousarray(kf_a, np.int32); b = np.ascontiguousarray(kf_b, np.int32)
        out = np.zeros(len(a), np.int32)
        _check(self._L.orbgpu_matcher_count_pairs(self._matcher(), _p(a), _p(b), len(a),
                                                  self.TH_LOW if th_low is None else th_low, float(self.mfNNratio), _p(out)))
        return out

    def count_pairs_device(self, pairs_ptr, num_pairs, counts_ptr, th_low=None):
        _check(self._L.orbgpu_matcher_count_pairs_device(self._matcher(), pairs_ptr, num_pairs,
                                                         self.TH_LOW if th_low is None else th_low, float(self.mfNNratio), counts_ptr))

    def last_kernel_ms(self):
        return float(self._L.orbgpu_matcher_last_kernel_ms(self._matcher()))

    def all_pairs_sharded(self, comm, counts_ptr, th_low=None):
        """Collective over the ranks of `comm` (None = this GPU alone): counts of ALL K (K - 1) / 2 pairs land in the device
        buffer counts_ptr on every rank.  Returns (kernel ms, gather + scatter ms, total ms)."""
        ms = (C.c_double * 3)()
        _check(self._L.orbgpu_matcher_all_pairs_sharded(self._matcher(), comm._h if comm is not None else None,
                                                        self.TH_LOW if th_low is None else th_low, float(self.mfNNratio), counts_ptr, ms))
        return ms[0], ms[1], ms[2]

    def close(self):
        if self._m is not None:
            self._L.orbgpu_matcher_destroy(self._m)
            self._m = None

    __del__ = close


class Comm:
    """NCCL communicator created by the C ABI (orbgpu_comm_*): one rank of a one-process-per-GPU job."""

    @staticmethod
    def unique_id():
        buf = (C.c_ubyte * 128)()
        _check(lib().orbgpu_comm_unique_id(buf))
        return bytes(buf)

    def __init__(self, device, rank, world, id128):
        self._h = C.c_void_p()
        buf = (C.c_ubyte * 128).from_buffer_copy(id128)
        _check(lib().orbgpu_comm_init_rank(C.byref(self._h), device, rank, world, buf))
        self.rank, self.world, self.device = rank, world, device

    def close(self):
        if getattr(self, "_h", None):
            lib().orbgpu_comm_destroy(self._h)
            self._h = None

    __del__ = close


def all_pairs_multi(matchers, devices, counts_ptrs, th_low=50, ratio=0.6):
    """One process, several GPUs: matchers[r] on devices[r], one NCCL group gather (orbgpu_matcher_all_pairs_multi)."""
    L = lib()
    n = len(matchers)
    comms = (C.c_void_p * n)()
    devs = (C.c_int * n)(*devices)
    _check(L.orbgpu_comm_init_all(comms, devs, n))
    try:
        ms = (C.c_void_p * n)(*[m._matcher() for m in matchers])
        cp = (C.c_void_p * n)(*counts_ptrs)
        t = (C.c_double * 3)()
        _check(L.orbgpu_matcher_all_pairs_multi(ms, comms, n, th_low, ratio, cp, t))
        return t[0], t[1], t[2]
    finally:
        for i in range(n):
            if comms[i]:
                L.orbgpu_comm_destroy(comms[i])


class ORBVocabulary:
    """ORB_SLAM2::ORBVocabulary = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (include/ORBVocabulary.h:30-31) resident
    on the device: transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194) as Frame::ComputeBoW / KeyFrame::ComputeBoW
    call it (src/Frame.cc:395-402, src/KeyFrame.cc:59-68) and score (:1198-1203, L1Scoring) for batches of keyframes.
    Build it from flat arrays (node 0 = root) or, like loadFromTextFile (:1338-1425), from an ORBvoc.txt-format file."""

    TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3
    L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT = 0, 1, 2, 3, 4, 5

    def __init__(self, k, L, parent, is_leaf, desc, weight, scoring=0, weighting=0, device=0):
        parent = np.ascontiguousarray(parent, np.int32); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); weight = np.ascontiguousarray(weight, np.float64)
        if not (len(parent) == len(is_leaf) == len(desc) == len(weight)):
            raise ValueError("parent, is_leaf, desc and weight must have one entry per node (root included)")
        h = C.c_void_p()
        _check(lib().orbgpu_vocabulary_create(C.byref(h), device, k, L, scoring, weighting, len(parent), _p(parent), _p(is_leaf),
                                              _p(desc), _p(weight)))
        self._h = h

    @classmethod
    def load_from_text_file(cls, path, device=0):
        self = cls.__new__(cls)
        h = C.c_void_p()
        _check(lib().orbgpu_vocabulary_load_text(C.byref(h), device, os.fsencode(path)))
        self._h = h
        return self

    def close(self):
        if getattr(self, "_h", None):
            lib().orbgpu_vocabulary_destroy(self._h)
            self._h = None

    __del__ = close

    def info(self):
        """(k, L, scoring, weighting, number of nodes, number of words)"""
        out = np.zeros(6, np.int32)
        _check(lib().orbgpu_vocabulary_info(self._h, _p(out)))
        return tuple(int(x) for x in out)

    def transform(self, desc, levelsup=4):
        """-> (words, values, node_ids, start, idx): the BowVector as ascending (word id, value) arrays and the FeatureVector
        as ascending node ids with their feature indices idx[start[j]:start[j + 1]] (the layout ORBmatcher.search_by_bow takes)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d); m = max(n, 1)
        words = np.zeros(m, np.uint32); values = np.zeros(m, np.float64); nodes = np.zeros(m, np.uint32)
        start = np.zeros(m + 1, np.int32); idx = np.zeros(m, np.uint32); counts = np.zeros(3, np.int32)
        _check(lib().orbgpu_vocabulary_transform(self._h, _p(d) if n else None, n, levelsup, _p(words), _p(values), _p(nodes),
                                                 _p(start), _p(idx), _p(counts)))
        nw, nn, ni = (int(x) for x in counts)
        return words[:nw].copy(), values[:nw].copy(), nodes[:nn].copy(), start[:nn + 1].copy(), idx[:ni].copy()

    def transform_batch(self, descs, levelsup=4):
        """ComputeBoW for several frames in one call: descs = list of (n_f x 32) arrays -> list of transform() results"""
        fs = np.zeros(len(descs) + 1, np.int32)
        fs[1:] = np.cumsum([len(d) for d in descs])
        n = int(fs[-1]); m = max(n, 1); nf = len(descs)
        d = np.ascontiguousarray(np.concatenate([np.asarray(x, np.uint8).reshape(-1, 32) for x in descs] + [np.zeros((0, 32), np.uint8)]))
        words = np.zeros(m, np.uint32); values = np.zeros(m, np.float64); nodes = np.zeros(m, np.uint32)
        start = np.zeros(m + nf + 1, np.int32); idx = np.zeros(m, np.uint32); counts = np.zeros(3 * max(nf, 1), np.int32)
        _check(lib().orbgpu_vocabulary_transform_batch(self._h, _p(d) if n else None, _p(fs), nf, levelsup, _p(words), _p(values),
                                                       _p(nodes), _p(start), _p(idx), _p(counts)))
        out = []
        for f in range(nf):
            b = int(fs[f]); nw, nn, ni = (int(x) for x in counts[3 * f:3 * f + 3])
            out.append((words[b:b + nw].copy(), values[b:b + nw].copy(), nodes[b:b + nn].copy(), start[b + f:b + f + nn + 1].copy(),
                        idx[b:b + ni].copy()))
        return out

    def score(self, query, keyframes):
        """query = (words, values); keyframes = list of (words, values) -> scores (float64), one per keyframe."""
        qw = np.ascontiguousarray(query[0], np.uint32); qv = np.ascontiguousarray(query[1], np.float64)
        ks = np.zeros(len(keyframes) + 1, np.int32)
        ks[1:] = np.cumsum([len(k[0]) for k in keyframes])
        kw = np.ascontiguousarray(np.concatenate([np.asarray(k[0], np.uint32) for k in keyframes]) if keyframes else np.zeros(0, np.uint32))
        kv = np.ascontiguousarray(np.concatenate([np.asarray(k[1], np.float64) for k in keyframes]) if keyframes else np.zeros(0, np.float64))
        return self.score_csr(qw, qv, kw, kv, ks)

    def score_csr(self, qw, qv, kw, kv, kstart):
        out = np.zeros(max(len(kstart) - 1, 1), np.float64)
        _check(lib().orbgpu_vocabulary_score(self._h, _p(qw) if len(qw) else None, _p(qv) if len(qv) else None, len(qw),
                                             _p(kw) if len(kw) else None, _p(kv) if len(kv) else None, _p(kstart), len(kstart) - 1, _p(out)))
        return out[:len(kstart) - 1].copy()


def write_vocabulary_text(path, k, L, parent, is_leaf, desc, weight, scoring=0, weighting=0, trailing_newline=False):
    """TemplatedVocabulary::saveToTextFile's format (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1428-1449): what
    loadFromTextFile reads.  Weights are written with 17 significant digits so that they parse back to the same double."""
    with open(path, "w") as f:
        f.write("%d %d %d %d" % (k, L, scoring, weighting))
        for i in range(1, len(parent)):
            f.write("\n%d %d %s %.17g" % (parent[i], 1 if is_leaf[i] else 0, " ".join(str(int(b)) for b in desc[i]), weight[i]))
        if trailing_newline:
            f.write("\n")


class KeyFrameDatabase:
    """ORB_SLAM2::KeyFrameDatabase (include/KeyFrameDatabase.h:44-75, src/KeyFrameDatabase.cc) with the keyframes' BowVectors
    resident on the device.  Keyframes are known by the slot `add` returns.  The shared-word counts and L1 scores of a query
    against every keyframe come from one kernel launch (orbgpu_bowdb_query); the thresholds and the covisibility accumulation
    (:107-197, :224-309) follow on the host, in the reference's float arithmetic and list order.  loop_score / reloc_score keep
    mLoopScore / mRelocScore between queries, as the KeyFrame members do."""

    def __init__(self, device=0):
        h = C.c_void_p()
        _check(lib().orbgpu_bowdb_create(C.byref(h), device))
        self._h = h
        self.loop_score = []
        self.reloc_score = []

    def close(self):
        if getattr(self, "_h", None):
            lib().orbgpu_bowdb_destroy(self._h)
            self._h = None

    __del__ = close

    def add(self, words, values):
        w = np.ascontiguousarray(words, np.uint32); v = np.ascontiguousarray(values, np.float64)
        slot = C.c_int(-1)
        _check(lib().orbgpu_bowdb_add(self._h, _p(w) if len(w) else None, _p(v) if len(v) else None, len(w), C.byref(slot)))
        self.loop_score.append(np.float32(0)); self.reloc_score.append(np.float32(0))
        return slot.value

    def erase(self, slot):
        _check(lib().orbgpu_bowdb_erase(self._h, slot))

    def clear(self):
        _check(lib().orbgpu_bowdb_clear(self._h))
        self.loop_score = []; self.reloc_score = []

    def __len__(self):
        return lib().orbgpu_bowdb_size(self._h)

    def query(self, words, values):
        """-> (common_words, first_common, scores) for every slot"""
        w = np.ascontiguousarray(words, np.uint32); v = np.ascontiguousarray(values, np.float64)
        n = len(self); m = max(n, 1)
        common = np.zeros(m, np.int32); first = np.zeros(m, np.int32); scores = np.zeros(m, np.float64)
        _check(lib().orbgpu_bowdb_query(self._h, _p(w) if len(w) else None, _p(v) if len(v) else None, len(w), _p(common), _p(first), _p(scores)))
        return common[:n], first[:n], scores[:n]

    def _detect(self, query, reloc, connected, neighbours, min_score):
        common, first, scores = self.query(*query)
        n = len(common)
        state = self.reloc_score if reloc else self.loop_score
        marked = common > 0
        if not reloc and connected is not None:
            marked &= ~np.asarray(connected, bool)[:n]
        words = np.where(marked, common, 0)
        order = np.nonzero(marked)[0]
        order = order[np.lexsort((order, first[order]))]          # the inverted-file walk: by first shared word, then insertion
        if len(order) == 0:
            return marked, words, []
        min_common = int(np.float32(int(words[order].max())) * np.float32(0.8))
        scored = []
        for k in order:
            if words[k] > min_common:
                si = np.float32(scores[k])
                state[k] = si
                if reloc or si >= np.float32(min_score):
                    scored.append((si, int(k)))
        if not scored:
            return marked, words, []
        acc = []
        best_acc = np.float32(0) if reloc else np.float32(min_score)
        for si, k in scored:
            best, total, best_k = si, si, k
            for k2 in list(neighbours(k))[:10]:
                if not marked[k2] or (not reloc and not words[k2] > min_common):
                    continue
                total = np.float32(total + state[k2])
                if state[k2] > best:
                    best_k, best = k2, state[k2]
            acc.append((total, best_k))
            if total > best_acc:
                best_acc = total
        keep = np.float32(0.75) * best_acc
        out = []
        for total, k in acc:
            if total > keep and k not in out:
                out.append(k)
        return marked, words, out

    def detect_loop_candidates(self, query, connected, neighbours, min_score):
        """DetectLoopCandidates(pKF, minScore) (:76-197): query = pKF's (words, values); connected[slot] = member of
        pKF->GetConnectedKeyFrames(); neighbours(slot) = that keyframe's ordered covisibility list -> candidate slots"""
        return self._detect(query, False, connected, neighbours, min_score)[2]

    def detect_relocalization_candidates(self, query, neighbours):
        """DetectRelocalizationCandidates(F) (:199-309) -> candidate slots"""
        return self._detect(query, True, None, neighbours, 0.0)[2]


def update_normals_and_depths(pos, obs_start, centers, ref_center, ref_scale, last_scale, normal, max_dist, min_dist, device=0):
    """MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371) for many map points at once -> (normal (n x 3), max_dist, min_dist);
    the three last inputs are the values before the call (kept for points without observations)."""
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3); n = len(pos)
    os_ = np.ascontiguousarray(obs_start, np.int32); c = np.ascontiguousarray(centers, np.float32).reshape(-1, 3)
    rc = np.ascontiguousarray(ref_center, np.float32).reshape(-1, 3)
    rs = np.ascontiguousarray(ref_scale, np.float32); ls = np.ascontiguousarray(last_scale, np.float32)
    nv = np.array(normal, np.float32).reshape(-1, 3).copy(); mx = np.array(max_dist, np.float32).copy(); mn = np.array(min_dist, np.float32).copy()
    _check(lib().orbgpu_update_normals_and_depths(device, _p(pos), _p(os_), _p(c) if len(c) else None, _p(rc), _p(rs), _p(ls), n,
                                                  _p(nv), _p(mx), _p(mn)))
    return nv, mx, mn
