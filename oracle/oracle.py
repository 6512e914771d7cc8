"""ctypes wrapper of oracle/tsdf_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.  Parity status: pinned for the
global volume by tests/golden (fixtures made from the unmodified reference).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")
_SRCS = ["tsdf_oracle.c", "local_oracle.c", "mc_oracle.c"]


def build(force=False):
    srcs = [os.path.join(_HERE, s) for s in _SRCS if os.path.exists(os.path.join(_HERE, s))]
    deps = srcs + [os.path.join(_HERE, "..", "mast3r_slam_quality_dualtsdf_b200", "csrc",
                                "mc_tables.h")]
    if not force and os.path.exists(_LIB) and all(
        os.path.getmtime(_LIB) >= os.path.getmtime(s) for s in deps
    ):
        return _LIB
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared",
           "-o", _LIB] + srcs + ["-lm"]
    subprocess.check_call(cmd)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_double] * 4
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_integrate_f32.restype = C.c_int64
        L.orc_integrate_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_int64, C.c_double]
        L.orc_integrate_f64.restype = C.c_int64
        L.orc_integrate_f64.argtypes = L.orc_integrate_f32.argtypes
        L.orc_num_voxels.restype = C.c_uint64
        L.orc_num_voxels.argtypes = [C.c_void_p]
        L.orc_num_events.restype = C.c_uint64
        L.orc_num_events.argtypes = [C.c_void_p]
        L.orc_export.restype = C.c_uint64
        L.orc_export.argtypes = [C.c_void_p] * 5
        L.orc_stats.argtypes = [C.c_void_p] * 3
        L.orc_query.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64,
                                C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_sim3_act_f32.argtypes = [C.c_void_p] * 3
        L.orc_stage_keyframe.restype = C.c_int64
        L.orc_stage_keyframe.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_marching_cubes.restype = C.c_int64
        L.orc_marching_cubes.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                         C.c_double, C.c_double, C.c_void_p, C.c_int64]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleVolume:
    """CPU restatement of reference TSDFVolume (global_volume.py:15-140)."""

    def __init__(self, voxel_size, truncation, max_weight=100.0, min_weight=1.0e-3):
        self.voxel_size = float(voxel_size)
        self.truncation = float(truncation)
        self.max_weight = float(max_weight)
        self.min_weight = float(min_weight)
        self._h = lib().orc_create(self.voxel_size, self.truncation, self.max_weight,
                                   self.min_weight)

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.orc_destroy(self._h)
            self._h = None

    def integrate(self, points_world, confidences, cam_origin, step_scale=0.5):
        pts = np.asarray(points_world)
        if pts.size == 0:
            return 0
        conf = np.ascontiguousarray(confidences, dtype=np.float64).reshape(-1)
        if pts.dtype == np.float32:
            pts = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 3)
            org = np.ascontiguousarray(cam_origin, dtype=np.float32).reshape(3)
            fn = lib().orc_integrate_f32
        else:
            pts = np.ascontiguousarray(pts, dtype=np.float64).reshape(-1, 3)
            org = np.ascontiguousarray(cam_origin, dtype=np.float64).reshape(3)
            fn = lib().orc_integrate_f64
        return int(fn(self._h, _p(pts), _p(conf), _p(org), pts.shape[0], float(step_scale)))

    def export(self):
        n = int(lib().orc_num_voxels(self._h))
        keys = np.zeros((n, 3), np.int64)
        tsdf = np.zeros(n, np.float64)
        w = np.zeros(n, np.float64)
        cnt = np.zeros(n, np.uint32)
        lib().orc_export(self._h, _p(keys), _p(tsdf), _p(w), _p(cnt))
        return keys, tsdf, w, cnt

    @property
    def events(self):
        return int(lib().orc_num_events(self._h))

    def stats(self):
        a = C.c_int64()
        b = C.c_int64()
        lib().orc_stats(self._h, C.byref(a), C.byref(b))
        return {"valid_voxels": a.value, "total_voxels": b.value}

    def query_batch(self, pts):
        pts = np.asarray(pts)
        f64 = pts.dtype != np.float32
        pts = np.ascontiguousarray(pts, dtype=np.float64 if f64 else np.float32).reshape(-1, 3)
        m = pts.shape[0]
        t = np.zeros(m, np.float64)
        g = np.zeros((m, 3), np.float64)
        fl = np.zeros(m, np.uint8)
        lib().orc_query(self._h, _p(pts), int(f64), m, _p(t), _p(g), _p(fl))
        return t, g, fl

    def query(self, point_world):
        t, g, fl = self.query_batch(np.asarray(point_world).reshape(1, 3))
        if not fl[0] & 1:
            return None, None
        return float(t[0]), (g[0].copy() if fl[0] & 2 else None)


def sim3_act(pose, x):
    pose = np.ascontiguousarray(pose, np.float32).reshape(8)
    x = np.ascontiguousarray(x, np.float32).reshape(-1, 3)
    y = np.zeros_like(x)
    for i in range(x.shape[0]):
        lib().orc_sim3_act_f32(_p(pose), _p(x[i]), _p(y[i]))
    return y


def stage_keyframe(X, Cf, pose, q_grid, H, W, min_conf, max_points=0, seed=0, kf_id=0):
    """Returns pts_world f32 (n,3), conf f64 (n,), conf_q f64 (n,), cam_origin f32 (3,).
    max_points > 0: only the pixels of `subsample_mask` are staged."""
    X = np.ascontiguousarray(X, np.float32).reshape(-1, 3)
    Cf = np.ascontiguousarray(Cf, np.float32).reshape(-1)
    if max_points > 0:
        Cf = np.where(subsample_mask(Cf, min_conf, max_points, seed, kf_id), Cf,
                      np.float32(0.0)).astype(np.float32)
    pose = np.ascontiguousarray(pose, np.float32).reshape(8)
    n = H * W
    pts = np.zeros((n, 3), np.float32)
    conf = np.zeros(n, np.float64)
    confq = np.zeros(n, np.float64)
    org = np.zeros(3, np.float32)
    gh = gw = 0
    qg = None
    if q_grid is not None:
        qg = np.ascontiguousarray(q_grid, np.float32)
        gh, gw = qg.shape
    k = int(lib().orc_stage_keyframe(_p(X), _p(Cf), _p(pose), _p(qg), gh, gw, H, W,
                                     float(min_conf), _p(pts), _p(conf), _p(confq), _p(org)))
    return pts[:k].copy(), conf[:k].copy(), confq[:k].copy(), org


def _fmix32(x):
    """murmur3 finaliser on uint32 arrays (a bijection of Z_2^32)."""
    x = np.asarray(x, np.uint64) & 0xFFFFFFFF
    x ^= x >> 16
    x = (x * 0x85EBCA6B) & 0xFFFFFFFF
    x ^= x >> 13
    x = (x * 0xC2B2AE35) & 0xFFFFFFFF
    x ^= x >> 16
    return x


def subsample_mask(Cf, min_conf, max_points, seed=0, kf_id=0):
    """The seeded stand-in for the reference's unseeded random choice of at most
    `max_points_per_kf` valid pixels (global_manager.py:92-97): boolean mask over
    the H*W pixels that keeps the max_points valid pixels (C > min_conf) with the
    smallest hash(pixel) -- the definition the product's
    subsample_threshold_kernel implements (csrc/integrate.cu: sub_hash, sub_salt)."""
    Cf = np.ascontiguousarray(Cf, np.float32).reshape(-1)
    valid = Cf > np.float32(min_conf)
    if max_points <= 0 or int(valid.sum()) <= max_points:
        return valid
    seed, kf_id = int(seed) & 0xFFFFFFFFFFFFFFFF, int(kf_id) & 0xFFFFFFFFFFFFFFFF
    a = int(_fmix32(((kf_id & 0xFFFFFFFF) + 0x9E3779B9 * (seed >> 32)) & 0xFFFFFFFF))
    b = int(_fmix32(((kf_id >> 32) + 0x7F4A7C15) & 0xFFFFFFFF))
    salt = int(_fmix32((seed & 0xFFFFFFFF) ^ a ^ b))
    pix = np.arange(Cf.shape[0], dtype=np.uint64)
    h = _fmix32((pix * 0x9E3779B1 + salt) & 0xFFFFFFFF)
    h_valid = np.where(valid, h, np.uint64(1 << 40))  # invalid pixels sort last
    keep = np.argsort(h_valid, kind="stable")[:max_points]
    mask = np.zeros(Cf.shape[0], bool)
    mask[keep] = True
    return mask & valid


def marching_cubes(keys, tsdf, weight, voxel_size, min_weight=1.0e-3):
    """CPU marching cubes (oracle/mc_oracle.c; parity unpinned -- the reference
    has no mesh path).  Returns verts float32 (T, 3, 3)."""
    keys = np.ascontiguousarray(keys, np.int64).reshape(-1, 3)
    tsdf = np.ascontiguousarray(tsdf, np.float32).reshape(-1)
    weight = np.ascontiguousarray(weight, np.float32).reshape(-1)
    n = keys.shape[0]
    nt = int(lib().orc_marching_cubes(_p(keys), _p(tsdf), _p(weight), n, float(voxel_size),
                                      float(min_weight), None, 0))
    verts = np.zeros((max(nt, 1), 3, 3), np.float32)
    if nt:
        lib().orc_marching_cubes(_p(keys), _p(tsdf), _p(weight), n, float(voxel_size),
                                 float(min_weight), _p(verts), nt)
    return verts[:nt]


# ---------------------------------------------------------------- local ROI
def _local_lib():
    L = lib()
    if not getattr(L, "_local_ready", False):
        L.orc_local_dims.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_int, C.c_void_p]
        L.orc_local_build.restype = C.c_int64
        L.orc_local_build.argtypes = [C.c_void_p] * 5 + [C.c_int64, C.c_double, C.c_double, C.c_int,
                                                         C.c_double, C.c_void_p, C.c_void_p,
                                                         C.c_void_p]
        L.orc_local_raycast.restype = C.c_int64
        L.orc_local_raycast.argtypes = [C.c_void_p] * 6 + [C.c_int64, C.c_int, C.c_double,
                                                           C.c_void_p, C.c_void_p]
        L._local_ready = True
    return L


def local_build(X_canon, Cf, pose, xyz_min, xyz_max, voxel_size, trunc_dist, max_grid_dim,
                min_confidence):
    """tsdf_refine.py:837-940 -> (tsdf (nz,ny,nx) f32, weights, fused)."""
    L = _local_lib()
    X = np.ascontiguousarray(X_canon, np.float32).reshape(-1, 3)
    Cf = np.ascontiguousarray(Cf, np.float32).reshape(-1)
    pose = np.ascontiguousarray(pose, np.float32).reshape(8)
    lo = np.ascontiguousarray(xyz_min, np.float32).reshape(3)
    hi = np.ascontiguousarray(xyz_max, np.float32).reshape(3)
    dims = np.zeros(3, np.int32)
    L.orc_local_dims(_p(lo), _p(hi), float(voxel_size), int(max_grid_dim), _p(dims))
    nx, ny, nz = (int(v) for v in dims)
    tsdf = np.zeros((nz, ny, nx), np.float32)
    w = np.zeros((nz, ny, nx), np.float32)
    fused = int(L.orc_local_build(_p(X), _p(Cf), _p(pose), _p(lo), _p(hi), X.shape[0],
                                  float(voxel_size), float(trunc_dist), int(max_grid_dim),
                                  float(min_confidence), _p(tsdf), _p(w), _p(dims)))
    return tsdf, w, fused


def local_raycast(tsdf, xyz_min, xyz_max, X_original, pix, n_samples, max_displacement):
    """tsdf_refine.py:942-1021 for an explicit pixel list -> (X_refined, hits u8 (m,))."""
    L = _local_lib()
    tsdf = np.ascontiguousarray(tsdf, np.float32)
    nz, ny, nx = tsdf.shape
    dims = np.array([nx, ny, nz], np.int32)
    lo = np.ascontiguousarray(xyz_min, np.float32).reshape(3)
    hi = np.ascontiguousarray(xyz_max, np.float32).reshape(3)
    X = np.ascontiguousarray(X_original, np.float32).reshape(-1, 3)
    pix = np.ascontiguousarray(pix, np.int64).reshape(-1)
    Xr = X.copy()
    hits = np.zeros(len(pix), np.uint8)
    L.orc_local_raycast(_p(tsdf), _p(dims), _p(lo), _p(hi), _p(X), _p(pix), len(pix),
                        int(n_samples), float(max_displacement), _p(Xr), _p(hits))
    return Xr, hits
