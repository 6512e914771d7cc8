"""ctypes binding of libb200tsdf.so (include/b200tsdf.h).

This is the binding a maintainer of the reference would add.
There is no CPU fallback: if the library is missing or no CUDA device is
present, calls raise.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200TSDF_LIB: alternative build of the same library (kernel A/B experiments)
LIB_PATH = os.environ.get("B200TSDF_LIB") or os.path.join(_HERE, "libb200tsdf.so")

F32, F64 = 0, 1
E_INVALID, E_CUDA, E_CAPACITY, E_NOMEM = -1, -2, -3, -4

# every symbol include/b200tsdf.h declares (tests check the .so exports all)
SYMBOLS = [
    "b200tsdf_last_error", "b200tsdf_abi_version", "b200tsdf_create", "b200tsdf_destroy",
    "b200tsdf_reset", "b200tsdf_integrate_points", "b200tsdf_integrate_keyframes",
    "b200tsdf_integrate_keyframes_host", "b200tsdf_deintegrate_keyframes",
    "b200tsdf_integrate_keyframes_ex", "b200tsdf_selftest_div",
    "b200tsdf_keyframe_meta_words", "b200tsdf_keyframe_meta", "b200tsdf_integrate_keyframes_pull",
    "b200tsdf_ipc_open", "b200tsdf_ipc_close", "b200tsdf_peer_alloc", "b200tsdf_peer_free", "b200tsdf_counters", "b200tsdf_stats",
    "b200tsdf_query", "b200tsdf_export_voxels", "b200tsdf_export_blocks",
    "b200tsdf_extract_mesh", "b200tsdf_extract_mesh_async", "b200tsdf_extract_mesh_static", "b200tsdf_extract_mesh_indexed", "b200tsdf_weld_mesh",
    "b200tsdf_export_seam_blocks", "b200tsdf_import_halo_blocks",
    "b200tsdf_export_block_data", "b200tsdf_import_blocks",
    "b200tsdf_local_dims", "b200tsdf_local_build",
    "b200tsdf_local_raycast", "b200tsdf_quality_patch_grids", "b200tsdf_quality_classify",
    "b200tsdf_pose_normal_equations", "b200tsdf_pose_optimize", "b200tsdf_quality_batch",
    "b200tsdf_query_taps", "b200tsdf_query_finish", "b200tsdf_pose_normal_equations_taps",
]


class Config(C.Structure):
    _fields_ = [
        ("voxel_size", C.c_double), ("truncation", C.c_double),
        ("max_weight", C.c_double), ("min_weight", C.c_double),
        ("capacity_log2", C.c_int32), ("bounded", C.c_int32),
        ("key_min", C.c_int64 * 3), ("key_max", C.c_int64 * 3),
        ("device", C.c_int32), ("reserved", C.c_int32),
        ("slab_count", C.c_int32), ("slab_index", C.c_int32), ("slab_width", C.c_int64),
    ]


class KfBatch(C.Structure):
    """b200tsdf_kf_batch (include/b200tsdf.h)."""
    _fields_ = [
        ("X_dev", C.c_void_p), ("C_dev", C.c_void_p), ("q_grid_dev", C.c_void_p),
        ("pose_dev", C.c_void_p),
        ("B", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("gh", C.c_int32),
        ("gw", C.c_int32), ("seg_kf", C.c_int32), ("seg_stride", C.c_int64),
        ("min_conf", C.c_float), ("negate", C.c_int32), ("step_scale", C.c_double),
        ("max_points", C.c_int32), ("reserved", C.c_int32), ("seed", C.c_uint64),
        ("kf_ids_dev", C.c_void_p),
    ]


MAX_SEGMENTS = 16


class KfPull(C.Structure):
    """b200tsdf_kf_pull (include/b200tsdf.h)."""
    _fields_ = [
        ("X_seg_dev", C.c_void_p * MAX_SEGMENTS), ("C_seg_dev", C.c_void_p * MAX_SEGMENTS),
        ("meta_dev", C.c_void_p), ("meta_stride", C.c_int64),
        ("n_seg", C.c_int32), ("seg_kf", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("gh", C.c_int32), ("gw", C.c_int32), ("min_conf", C.c_float), ("negate", C.c_int32),
        ("step_scale", C.c_double),
    ]


class LocalCfg(C.Structure):
    _fields_ = [
        ("voxel_size", C.c_double), ("trunc_dist", C.c_double), ("min_confidence", C.c_double),
        ("max_displacement", C.c_double), ("max_grid_dim", C.c_int32), ("ray_samples", C.c_int32),
    ]


class B200TSDFError(RuntimeError):
    pass


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200TSDFError(
            f"{LIB_PATH} is missing: build it with "
            "`python -m mast3r_slam_quality_dualtsdf_b200.build` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    L.b200tsdf_last_error.restype = C.c_char_p
    L.b200tsdf_last_error.argtypes = []
    L.b200tsdf_abi_version.restype = C.c_int
    L.b200tsdf_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    L.b200tsdf_destroy.argtypes = [vp]
    L.b200tsdf_reset.argtypes = [vp, vp]
    L.b200tsdf_integrate_points.argtypes = [vp, vp, C.c_int, vp, vp, vp, i64, dbl, vp, vp]
    L.b200tsdf_integrate_keyframes.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32,
                                               C.c_float, dbl, vp]
    L.b200tsdf_integrate_keyframes_host.argtypes = L.b200tsdf_integrate_keyframes.argtypes
    L.b200tsdf_deintegrate_keyframes.argtypes = L.b200tsdf_integrate_keyframes.argtypes
    L.b200tsdf_integrate_keyframes_ex.argtypes = [vp, C.POINTER(KfBatch), vp]
    L.b200tsdf_selftest_div.argtypes = [i32, i64, C.c_uint64, vp, vp]
    L.b200tsdf_keyframe_meta_words.argtypes = [i32] * 5
    L.b200tsdf_keyframe_meta.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, C.c_float, vp, vp]
    L.b200tsdf_integrate_keyframes_pull.argtypes = [vp, C.POINTER(KfPull), vp]
    L.b200tsdf_ipc_open.argtypes = [i32, vp, i64, C.POINTER(vp)]
    L.b200tsdf_ipc_close.argtypes = [i32, vp, i64]
    L.b200tsdf_peer_alloc.argtypes = [i32, i64, C.POINTER(vp), vp]
    L.b200tsdf_peer_free.argtypes = [i32, vp]
    L.b200tsdf_counters.argtypes = [vp, vp, vp]
    L.b200tsdf_stats.argtypes = [vp, C.c_int, vp, vp, vp]
    L.b200tsdf_query.argtypes = [vp, C.c_int, vp, C.c_int, i64, vp, vp, vp, vp]
    L.b200tsdf_export_voxels.argtypes = [vp, C.c_int, vp, vp, vp, vp, i64, vp, vp]
    L.b200tsdf_export_blocks.argtypes = [vp, vp, i64, vp, vp]
    L.b200tsdf_extract_mesh.argtypes = [vp, C.c_int, vp, i64, vp, vp]
    L.b200tsdf_extract_mesh_async.argtypes = [vp, C.c_int, vp, i64, vp, vp]
    L.b200tsdf_extract_mesh_static.argtypes = [vp, C.c_int, vp, i64, vp, vp]
    L.b200tsdf_extract_mesh_indexed.argtypes = [vp, C.c_int, vp, i64, vp, i64, vp, vp, vp]
    L.b200tsdf_weld_mesh.argtypes = [i32, vp, i64, vp, i64, vp, vp, vp]
    L.b200tsdf_export_seam_blocks.argtypes = [vp, vp, vp, vp, i64, vp, vp]
    L.b200tsdf_import_halo_blocks.argtypes = [vp, vp, vp, vp, i64, vp]
    L.b200tsdf_export_block_data.argtypes = [vp, vp, vp, vp, i64, vp, vp]
    L.b200tsdf_quality_patch_grids.argtypes = [vp, vp, vp, vp, i32, i32, i32, C.c_double,
                                               C.c_double, vp, vp, vp, i32, vp]
    L.b200tsdf_pose_normal_equations.argtypes = [vp, i32, vp, vp, i64, vp, C.c_double, vp, vp, vp,
                                                 vp]
    L.b200tsdf_pose_optimize.argtypes = [vp, i32, vp, vp, i64, vp, C.c_double, C.c_double, i32, vp,
                                         vp, vp]
    L.b200tsdf_query_taps.argtypes = [vp, C.c_int, vp, C.c_int, i64, vp, vp, vp]
    L.b200tsdf_query_finish.argtypes = [vp, vp, i64, vp, vp, vp, vp]
    L.b200tsdf_pose_normal_equations_taps.argtypes = [vp, vp, vp, vp, i64, vp, C.c_double, vp, vp,
                                                      vp, vp]
    L.b200tsdf_quality_classify.argtypes = [vp, vp, vp, i32, C.c_double, C.c_double, C.c_double,
                                            vp, vp, vp, i32, vp]
    L.b200tsdf_quality_batch.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32] + \
        [C.c_double] * 6 + [vp] * 8 + [i32, vp]
    L.b200tsdf_import_blocks.argtypes = [vp, vp, vp, vp, i64, vp]
    L.b200tsdf_local_dims.argtypes = [vp, C.POINTER(LocalCfg), vp]
    L.b200tsdf_local_build.argtypes = [vp, vp, vp, vp, i64, C.POINTER(LocalCfg), vp, vp, vp, vp]
    L.b200tsdf_local_raycast.argtypes = [vp, vp, vp, vp, vp, i64, C.POINTER(LocalCfg), vp, vp, vp]
    for s in SYMBOLS:
        if s != "b200tsdf_last_error":
            getattr(L, s).restype = C.c_int
    L.b200tsdf_keyframe_meta_words.restype = C.c_int64
    _lib = L
    return L


def check(rc, what=""):
    if rc != 0:
        msg = lib().b200tsdf_last_error().decode("utf-8", "replace")
        raise B200TSDFError(f"{what} failed (rc={rc}): {msg}")
