"""orbgpu: B200-native ORB front end (pyramid, FAST + octree, IC_Angle, rBRIEF, Hamming matching).

The product is the CUDA library liborbgpu.so (csrc/, C ABI in include/orbgpu.h) and the C++ drop-in
classes in host/.  This package is the thin ctypes binding used by the tests and bench.py; the
directory name has a hyphen, so import it through __graft_entry__.load_package() (module name
`my_slam_b200`).
"""
from .orbgpu import (KP_DTYPE, ORBextractor, ORBmatcher, OrbGpuError, device_count, lib, lib_path, stereo_match, undistort_points,
                     ProjectionArgs, projection_args, search_for_initialization, search_by_projection,
                     BowSide, bow_side, search_by_bow,
                     LocalPointsArgs, local_points_args, search_local_points,
                     TriangulationArgs, triangulation_args, search_for_triangulation, search_for_triangulation_batch,
                     LocalMapArgs, TrackFields, track_fields, local_map_args, search_local_map, debug_logf,
                     FuseArgs, fuse_args, fuse_search, fuse_search_batch,
                     RelocProjectionArgs, reloc_projection_args, search_by_projection_keyframe, distinctive_descriptors,
                     fuse_search_sim3, search_by_projection_sim3, search_by_sim3,
                     ORBVocabulary, write_vocabulary_text, KeyFrameDatabase, update_normals_and_depths, Comm, all_pairs_multi, extract_batch_multi, HostBuffer, KeyFrameStore, KfData, KfTriangulationJob)  # noqa: F401
from . import shard, synth  # noqa: F401
