"""B200 drop-in for the reference `mast3r_slam/tsdf/global_volume.py`.

Same class name, constructor, `integrate` / `query` / `stats` signatures and
attributes as the reference TSDFVolume (global_volume.py:15-140); storage is a
device voxel-block hash owned by libb200tsdf.so, reached through the C ABI in
include/b200tsdf.h.  New on top of the reference: a second, quality-weighted
volume fused in the same pass (`confidences_q`, `integrate_keyframes`),
`query_batch`, `export`, `extract_mesh`.  No CPU fallback.
"""
import contextlib
import ctypes as C
import threading
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from .. import _ffi


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def packed_layout(seg_kf, H, W, gh=0, gw=0):
    """Float offsets of the four arrays inside one packed keyframe segment
    [X seg_kf*H*W*3 | C seg_kf*H*W | q seg_kf*gh*gw | pose seg_kf*8], padded to a
    multiple of 64 floats so every segment (and, with H*W % 4 == 0, every
    keyframe row) stays 16-byte aligned for the TMA staging."""
    n = H * W
    o_c = seg_kf * n * 3
    o_q = o_c + seg_kf * n
    o_p = o_q + seg_kf * gh * gw
    o_p = (o_p + 3) & ~3
    seg_len = (o_p + seg_kf * 8 + 63) & ~63
    return {"X": 0, "C": o_c, "q": o_q, "pose": o_p, "seg_len": seg_len}


def pack_keyframes(X, Cf, poses, q_grid=None, out=None):
    """One packed segment (see `packed_layout`) from the four keyframe arrays
    (numpy or torch, host or device).  `out`: preallocated 1-D float32 tensor
    (e.g. pinned host memory) of at least seg_len floats."""
    X = torch.as_tensor(X)
    k = int(X.shape[0])
    n = int(X.shape[1])
    q = None if q_grid is None else torch.as_tensor(q_grid).reshape(k, -1)
    gcells = 0 if q is None else int(q.shape[1])
    # H, W only enter through their product here
    lay = packed_layout(k, n, 1, gcells, 1 if gcells else 0)
    if out is None:
        out = torch.zeros(lay["seg_len"], dtype=torch.float32, device=X.device)
    out[lay["X"]:lay["X"] + k * n * 3].copy_(X.reshape(-1))
    out[lay["C"]:lay["C"] + k * n].copy_(torch.as_tensor(Cf).reshape(-1))
    if q is not None:
        out[lay["q"]:lay["q"] + k * gcells].copy_(q.reshape(-1))
    out[lay["pose"]:lay["pose"] + k * 8].copy_(torch.as_tensor(poses).reshape(-1))
    return out


def weld_mesh(tri_soup):
    """Exact weld of a device triangle soup (3T,3) / (T,3,3) float32 on the
    device (b200tsdf_weld_mesh): (verts (V,3) f32, faces (T,3) i32).  Used for
    the gathered per-rank meshes of a sharded volume, whose seam vertices agree
    bit for bit."""
    soup = tri_soup.reshape(-1, 3).contiguous()
    if not soup.is_cuda or soup.dtype != torch.float32:
        raise ValueError("weld_mesh needs a float32 device tensor")
    nt = soup.shape[0] // 3
    nv = C.c_int64()
    faces = torch.empty(max(nt, 1), 3, dtype=torch.int32, device=soup.device)
    verts = torch.empty(max(3 * nt, 1), 3, dtype=torch.float32, device=soup.device)
    _ffi.check(_ffi.lib().b200tsdf_weld_mesh(
        soup.device.index or 0, soup.data_ptr(), nt, verts.data_ptr(), 3 * nt, faces.data_ptr(),
        C.byref(nv), C.c_void_p(torch.cuda.current_stream(soup.device).cuda_stream)), "weld_mesh")
    return verts[: nv.value].clone(), faces[:nt]


class TSDFVolume:
    """Sparse dual TSDF volume (voxel-block hash on the GPU)."""

    def __init__(
        self,
        voxel_size: float,
        truncation: float,
        max_weight: float = 100.0,
        min_weight: float = 1.0e-3,
        *,
        capacity_log2: int = 17,
        bounds=None,
        slabs=None,
        device="cuda:0",
    ) -> None:
        self.voxel_size = float(voxel_size)
        self.truncation = float(truncation)
        self.max_weight = float(max_weight)
        self.min_weight = float(min_weight)
        self._lock = threading.RLock()  # reference: global_volume.py:30
        self.device = torch.device(device)
        if self.device.type != "cuda" or not torch.cuda.is_available():
            raise _ffi.B200TSDFError("TSDFVolume needs a CUDA device (no CPU fallback)")
        cfg = _ffi.Config()
        cfg.voxel_size, cfg.truncation = self.voxel_size, self.truncation
        cfg.max_weight, cfg.min_weight = self.max_weight, self.min_weight
        cfg.capacity_log2 = int(capacity_log2)
        cfg.bounded = 0
        if bounds is not None:  # ((x0,y0,z0),(x1,y1,z1)) integer voxel keys, half-open
            cfg.bounded = 1
            for i in range(3):
                cfg.key_min[i] = int(bounds[0][i])
                cfg.key_max[i] = int(bounds[1][i])
        if slabs is not None:  # (slab_count, slab_index, slab_width): interleaved z-slabs
            cfg.slab_count, cfg.slab_index, cfg.slab_width = (int(x) for x in slabs)
        cfg.device = self.device.index or 0
        self.bounds = bounds
        self.slabs = slabs
        self._h = C.c_void_p()
        _ffi.check(_ffi.lib().b200tsdf_create(C.byref(cfg), C.byref(self._h)), "b200tsdf_create")
        # device-side ordering of writers (fuse, import, reset) against readers (query,
        # normal equations, sweeps) that run on OTHER streams: the RLock only orders
        # the launches on the host.  A writer waits for the last reader and vice versa.
        self._write_ev = torch.cuda.Event()
        self._write_stream = None
        self._read_evs = {}

    @contextlib.contextmanager
    def _writing(self):
        """Host lock + device order for a call that MUTATES the volume."""
        with self._lock:
            st = torch.cuda.current_stream(self.device)
            for other, ev in self._read_evs.items():
                if other != st:
                    st.wait_event(ev)
            if self._write_stream is not None and self._write_stream != st:
                st.wait_event(self._write_ev)
            try:
                yield
            finally:
                self._read_evs.clear()  # ordered before this write; later writers wait on it
                self._write_ev.record(st)
                self._write_stream = st

    @contextlib.contextmanager
    def _reading(self):
        """Host lock + device order for a call that only READS the volume."""
        with self._lock:
            st = torch.cuda.current_stream(self.device)
            if self._write_stream is not None and self._write_stream != st:
                st.wait_event(self._write_ev)
            try:
                yield
            finally:
                ev = self._read_evs.get(st)
                if ev is None:
                    ev = self._read_evs[st] = torch.cuda.Event()
                ev.record(st)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            _ffi.lib().b200tsdf_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_reset(self._h, _stream_ptr(self.device)), "reset")

    # ------------------------------------------------------------------
    # Integration
    # ------------------------------------------------------------------
    def integrate(
        self,
        points_world: np.ndarray,
        confidences: np.ndarray,
        cam_origin: np.ndarray,
        step_scale: float = 0.5,
        confidences_q: Optional[np.ndarray] = None,
    ) -> int:
        """Reference signature (global_volume.py:35-41).  float32 points use
        float32 ray arithmetic, anything else float64 -- as numpy >= 2 does
        inside the reference."""
        points_world = np.asarray(points_world)
        if points_world.size == 0:  # :42-43
            return 0
        f32 = points_world.dtype == np.float32
        dt, tdt = (np.float32, _ffi.F32) if f32 else (np.float64, _ffi.F64)
        pts = torch.from_numpy(np.ascontiguousarray(points_world, dtype=dt).reshape(-1, 3))
        conf = torch.from_numpy(np.ascontiguousarray(confidences, dtype=np.float64).reshape(-1))
        org = np.ascontiguousarray(np.asarray(cam_origin).reshape(3), dtype=dt)
        m = pts.shape[0]
        if conf.shape[0] < m:  # zip() in the reference stops at the shorter one
            m = conf.shape[0]
        pts_d = pts.to(self.device)
        conf_d = conf.to(self.device)
        cq_d = None
        if confidences_q is not None:
            cq_d = torch.from_numpy(
                np.ascontiguousarray(confidences_q, dtype=np.float64).reshape(-1)).to(self.device)
        fused = C.c_int64(0)
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_integrate_points(
                self._h, pts_d.data_ptr(), tdt, conf_d.data_ptr(),
                cq_d.data_ptr() if cq_d is not None else None,
                org.ctypes.data_as(C.c_void_p), m, float(step_scale), C.byref(fused),
                _stream_ptr(self.device)), "integrate_points")
        return int(fused.value)

    def deintegrate_keyframes(self, X, Cf, poses, q_grid=None, min_conf=0.05, H=None, W=None,
                              step_scale=0.5, max_points=0, seed=0, kf_ids=None):
        """Exact inverse of `integrate_keyframes` for the same device tensors and
        subsample arguments (new: the reference never removes a stale
        contribution)."""
        return self.integrate_keyframes(X, Cf, poses, q_grid, min_conf, H, W, step_scale,
                                        _negate=True, max_points=max_points, seed=seed,
                                        kf_ids=kf_ids)

    def integrate_keyframes(self, X, Cf, poses, q_grid=None, min_conf=0.05, H=None, W=None,
                            step_scale=0.5, _negate=False, max_points=0, seed=0, kf_ids=None):
        """Fused staging + allocation + dual fold for a batch of keyframes
        (replaces global_manager.py:81-106 per keyframe).  Tensors may live on
        the device (zero-copy) or on the host (copied inside the call; the host
        buffers are free again when it returns).
        X (B,H*W,3) f32 camera-frame pointmap, Cf (B,H*W[,1]) f32, poses (B,8)
        f32 Sim3 [t,q,s], q_grid (B,gh,gw) f32 or None.
        max_points > 0: the reference's `max_points_per_kf` cap
        (global_manager.py:94-99) -- of the pixels with C > min_conf exactly
        min(n_valid, max_points) per keyframe are fused, chosen by a seeded
        bijective hash of the pixel index (`seed`, and the keyframe's id from
        `kf_ids`, default its position in the batch); the reference draws an
        unseeded randperm.  Device tensors only."""
        X = X if X.dim() == 3 else X.unsqueeze(0)
        B, N = X.shape[0], X.shape[1]
        if H is None or W is None:
            raise ValueError("integrate_keyframes needs H and W")
        if H * W != N:
            raise ValueError(f"H*W = {H * W} does not match the pointmap ({N} pixels)")
        gh = gw = 0
        if q_grid is not None:
            q_grid = q_grid.reshape(B, q_grid.shape[-2], q_grid.shape[-1])
            gh, gw = q_grid.shape[-2], q_grid.shape[-1]
        on_dev = X.is_cuda
        ts = [X, Cf.reshape(B, N), poses.reshape(B, 8)] + ([q_grid] if q_grid is not None else [])
        for t in ts:
            if t.dtype != torch.float32 or not t.is_contiguous() or t.is_cuda != on_dev:
                raise ValueError("keyframe tensors must be contiguous float32 on one device")
        Xc, Cc, Pc = ts[0], ts[1], ts[2]
        capped = max_points is not None and 0 < int(max_points) < N
        if (_negate or capped) and not on_dev:
            raise ValueError("deintegrate_keyframes / max_points need device tensors")
        if not on_dev:
            with self._writing():
                _ffi.check(_ffi.lib().b200tsdf_integrate_keyframes_host(
                    self._h, Xc.data_ptr(), Cc.data_ptr(),
                    q_grid.data_ptr() if q_grid is not None else None, Pc.data_ptr(),
                    B, H, W, gh, gw, float(min_conf), float(step_scale),
                    _stream_ptr(self.device)), "integrate_keyframes_host")
            return
        kb = _ffi.KfBatch()
        kb.X_dev, kb.C_dev, kb.pose_dev = Xc.data_ptr(), Cc.data_ptr(), Pc.data_ptr()
        kb.q_grid_dev = q_grid.data_ptr() if q_grid is not None else None
        kb.B, kb.H, kb.W, kb.gh, kb.gw = B, H, W, gh, gw
        kb.seg_kf, kb.seg_stride = B, 0
        kb.min_conf, kb.negate, kb.step_scale = float(min_conf), int(bool(_negate)), float(step_scale)
        ids = None
        if capped:
            kb.max_points, kb.seed = int(max_points), int(seed) & 0xFFFFFFFFFFFFFFFF
            if kf_ids is not None:
                ids = torch.as_tensor(kf_ids, dtype=torch.int64).reshape(-1).to(self.device)
                if ids.numel() != B:
                    raise ValueError("kf_ids must name every keyframe of the batch")
                kb.kf_ids_dev = ids.data_ptr()
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_integrate_keyframes_ex(
                self._h, C.byref(kb), _stream_ptr(self.device)), "integrate_keyframes")

    def integrate_keyframes_packed(self, buf, n_seg, seg_kf, H, W, gh=0, gw=0, min_conf=0.05,
                                   step_scale=0.5, negate=False):
        """Fuse a batch that arrives as `n_seg` packed segments of `seg_kf`
        keyframes each (one segment per source rank -- exactly what ONE
        all-gather of the ranks' packed shares delivers, see `packed_layout`):
        `buf` is a contiguous float32 device tensor of n_seg * seg_len floats.
        gh = gw = 0: no quality grid in the segments (q = 1)."""
        lay = packed_layout(seg_kf, H, W, gh, gw)
        if buf.dtype != torch.float32 or not buf.is_cuda or not buf.is_contiguous():
            raise ValueError("packed keyframes must be a contiguous float32 device tensor")
        if buf.numel() < n_seg * lay["seg_len"]:
            raise ValueError(f"packed buffer holds {buf.numel()} floats, {n_seg} segments need "
                             f"{n_seg * lay['seg_len']}")
        kb = _ffi.KfBatch()
        base = buf.data_ptr()
        kb.X_dev = base + 4 * lay["X"]
        kb.C_dev = base + 4 * lay["C"]
        kb.q_grid_dev = (base + 4 * lay["q"]) if gh and gw else None
        kb.pose_dev = base + 4 * lay["pose"]
        kb.B, kb.H, kb.W, kb.gh, kb.gw = n_seg * seg_kf, H, W, gh, gw
        kb.seg_kf, kb.seg_stride = seg_kf, lay["seg_len"] if n_seg > 1 else 0
        kb.min_conf, kb.negate, kb.step_scale = float(min_conf), int(bool(negate)), float(step_scale)
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_integrate_keyframes_ex(
                self._h, C.byref(kb), _stream_ptr(self.device)), "integrate_keyframes_ex")

    # -- multi-GPU peer pull (parallel.PeerKeyframes drives these) ---------------
    @staticmethod
    def keyframe_meta_words(seg_kf, H, W, gh=0, gw=0):
        return int(_ffi.lib().b200tsdf_keyframe_meta_words(seg_kf, H, W, gh, gw))

    def keyframe_meta(self, X, Cf, poses, q_grid=None, min_conf=0.05, H=None, W=None, out=None):
        """OWNER side of the peer pull: the meta record of the keyframes this rank
        holds -- per 32 x 8 tile the z range of its bands (voxel layers) and its
        ray count, then the poses and quality grids (int32 device tensor of
        `keyframe_meta_words` words; ~28 KB per keyframe instead of 3.1 MB)."""
        B, N = X.shape[0], X.shape[1]
        gh = gw = 0
        if q_grid is not None:
            gh, gw = int(q_grid.shape[-2]), int(q_grid.shape[-1])
        words = self.keyframe_meta_words(B, H, W, gh, gw)
        if out is None:
            out = torch.zeros(words, dtype=torch.int32, device=self.device)
        with self._lock:  # reads the volume's geometry only, not its voxels: no stream ordering
            _ffi.check(_ffi.lib().b200tsdf_keyframe_meta(
                self._h, X.data_ptr(), Cf.data_ptr(),
                q_grid.data_ptr() if q_grid is not None else None, poses.data_ptr(), B, H, W, gh,
                gw, float(min_conf), out.data_ptr(), _stream_ptr(self.device)), "keyframe_meta")
        return out

    def integrate_keyframes_pull(self, x_ptrs, c_ptrs, meta, seg_kf, H, W, gh=0, gw=0,
                                 min_conf=0.05, step_scale=0.5, negate=False):
        """Fuse a batch whose pointmaps / confidences stay where they are: segment
        s (seg_kf keyframes of source rank s) is read tile by tile from the device
        addresses x_ptrs[s] / c_ptrs[s] (ints: this GPU's memory or a peer buffer
        mapped over NVLink); `meta` = the gathered meta records (n_seg x
        keyframe_meta_words int32).  Only tiles whose band reaches this volume's
        slabs are loaded."""
        n_seg = len(x_ptrs)
        if n_seg != len(c_ptrs) or not 0 < n_seg <= _ffi.MAX_SEGMENTS:
            raise ValueError(f"1..{_ffi.MAX_SEGMENTS} segments, X and C pointers for each")
        stride = self.keyframe_meta_words(seg_kf, H, W, gh, gw)
        if meta.dtype != torch.int32 or not meta.is_cuda or meta.numel() < n_seg * stride:
            raise ValueError("meta must be an int32 device tensor of n_seg meta records")
        kp = _ffi.KfPull()
        for s_ in range(n_seg):
            kp.X_seg_dev[s_], kp.C_seg_dev[s_] = int(x_ptrs[s_]), int(c_ptrs[s_])
        kp.meta_dev, kp.meta_stride = meta.data_ptr(), stride
        kp.n_seg, kp.seg_kf, kp.H, kp.W, kp.gh, kp.gw = n_seg, seg_kf, H, W, gh, gw
        kp.min_conf, kp.negate, kp.step_scale = float(min_conf), int(bool(negate)), float(step_scale)
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_integrate_keyframes_pull(
                self._h, C.byref(kp), _stream_ptr(self.device)), "integrate_keyframes_pull")

    # ------------------------------------------------------------------
    # Query helpers
    # ------------------------------------------------------------------
    def query_batch(self, points, which: int = 0):
        """Batched query: returns (tsdf f64 (M,), grad f64 (M,3), flags u8 (M,))
        device tensors; flags bit0 = value present, bit1 = gradient present."""
        if isinstance(points, np.ndarray):
            dt = np.float32 if points.dtype == np.float32 else np.float64
            points = torch.from_numpy(np.ascontiguousarray(points, dtype=dt))
        pts = points.reshape(-1, 3).to(self.device).contiguous()
        if pts.dtype not in (torch.float32, torch.float64):
            pts = pts.double()
        m = pts.shape[0]
        t = torch.zeros(m, dtype=torch.float64, device=self.device)
        g = torch.zeros(m, 3, dtype=torch.float64, device=self.device)
        fl = torch.zeros(m, dtype=torch.uint8, device=self.device)
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_query(
                self._h, which, pts.data_ptr(), _ffi.F32 if pts.dtype == torch.float32 else _ffi.F64,
                m, t.data_ptr(), g.data_ptr(), fl.data_ptr(), _stream_ptr(self.device)), "query")
        return t, g, fl

    def query(self, point_world: np.ndarray) -> Tuple[Optional[float], Optional[np.ndarray]]:
        """Reference signature (global_volume.py:93-105)."""
        t, g, fl = self.query_batch(np.asarray(point_world).reshape(1, 3))
        f = int(fl.cpu()[0])
        if not f & 1:
            return None, None
        return float(t.cpu()[0]), (g.cpu().numpy()[0].copy() if f & 2 else None)

    # ------------------------------------------------------------------
    # Utility
    # ------------------------------------------------------------------
    def pose_normal_equations(self, pts_cam, conf, pose, lam, which: int = 0):
        """One Gauss-Newton linearisation of the TSDF pose optimiser in ONE
        launch (Sim3 act + query + Jacobian + 7x7 reduction, float64):
        returns (H (7,7) ndarray, b (7,) ndarray, used int).  pts_cam (n,3)
        float32 and conf (n,) float64 device tensors, pose 8 floats."""
        import numpy as np

        pts = pts_cam.to(self.device, torch.float32).reshape(-1, 3).contiguous()
        cf = conf.to(self.device, torch.float64).reshape(-1).contiguous()
        ps = np.ascontiguousarray(np.asarray(pose, np.float32).reshape(8))
        H = np.zeros((7, 7), np.float64)
        b = np.zeros(7, np.float64)
        used = C.c_int64()
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_pose_normal_equations(
                self._h, which, pts.data_ptr(), cf.data_ptr(), pts.shape[0],
                ps.ctypes.data_as(C.c_void_p), float(lam), H.ctypes.data_as(C.c_void_p),
                b.ctypes.data_as(C.c_void_p), C.byref(used), _stream_ptr(self.device)),
                "pose_normal_equations")
        return H, b, int(used.value)

    def pose_optimize(self, pts_cam, conf, pose, lam, damping, iterations, which: int = 0):
        """`iterations` Gauss-Newton iterations of the TSDF pose optimiser entirely
        on the device (linearise, 7x7 solve, Sim3 retraction; one launch each, one
        synchronisation at the end): returns (pose (8,) float64 ndarray, used,
        iterations_done).  pts_cam (n,3) float32 / conf (n,) float64 device
        tensors, pose 8 values [t, q_xyzw, s]."""
        pts = pts_cam.to(self.device, torch.float32).reshape(-1, 3).contiguous()
        cf = conf.to(self.device, torch.float64).reshape(-1).contiguous()
        ps = np.ascontiguousarray(np.asarray(pose, np.float64).reshape(8)).copy()
        used, done = C.c_int64(), C.c_int32()
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_pose_optimize(
                self._h, which, pts.data_ptr(), cf.data_ptr(), pts.shape[0],
                ps.ctypes.data_as(C.c_void_p), float(lam), float(damping), int(iterations),
                C.byref(used), C.byref(done), _stream_ptr(self.device)), "pose_optimize")
        return ps, int(used.value), int(done.value)

    # -- sharded query (multi-GPU): taps of this handle, merged by the caller --
    def query_taps(self, points, which: int = 0, pose=None):
        """Step 1 of a sharded query: (M, 16) float32 records of the 7 voxels
        of every point THIS volume owns (see include/b200tsdf.h).  With `pose`
        (8 floats) the points are camera-frame float32 and are moved by
        Sim3(pose) first.  Sum the records over the ranks, then `query_finish`
        or `pose_normal_equations_taps`."""
        if isinstance(points, np.ndarray):
            dt = np.float32 if points.dtype == np.float32 else np.float64
            points = torch.from_numpy(np.ascontiguousarray(points, dtype=dt))
        pts = points.reshape(-1, 3).to(self.device).contiguous()
        if pose is not None:
            pts = pts.to(torch.float32)
        elif pts.dtype not in (torch.float32, torch.float64):
            pts = pts.double()
        m = pts.shape[0]
        taps = torch.zeros(m, 16, dtype=torch.float32, device=self.device)
        ps = None if pose is None else np.ascontiguousarray(np.asarray(pose, np.float32).reshape(8))
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_query_taps(
                self._h, which, pts.data_ptr(),
                _ffi.F32 if pts.dtype == torch.float32 else _ffi.F64, m,
                None if ps is None else ps.ctypes.data_as(C.c_void_p), taps.data_ptr(),
                _stream_ptr(self.device)), "query_taps")
            if ps is not None:  # the pose copy reads host memory asynchronously
                torch.cuda.current_stream(self.device).synchronize()
        return taps

    def query_finish(self, taps):
        """Step 2: (tsdf, grad, flags) like `query_batch` from merged records."""
        taps = taps.to(self.device, torch.float32).reshape(-1, 16).contiguous()
        m = taps.shape[0]
        t = torch.zeros(m, dtype=torch.float64, device=self.device)
        g = torch.zeros(m, 3, dtype=torch.float64, device=self.device)
        fl = torch.zeros(m, dtype=torch.uint8, device=self.device)
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_query_finish(
                self._h, taps.data_ptr(), m, t.data_ptr(), g.data_ptr(), fl.data_ptr(),
                _stream_ptr(self.device)), "query_finish")
        return t, g, fl

    def pose_normal_equations_taps(self, taps, pts_cam, conf, pose, lam):
        """`pose_normal_equations` with the query replaced by merged records
        (taps = sum over ranks of `query_taps(pts_cam, which, pose)`)."""
        taps = taps.to(self.device, torch.float32).reshape(-1, 16).contiguous()
        pts = pts_cam.to(self.device, torch.float32).reshape(-1, 3).contiguous()
        cf = conf.to(self.device, torch.float64).reshape(-1).contiguous()
        ps = np.ascontiguousarray(np.asarray(pose, np.float32).reshape(8))
        H = np.zeros((7, 7), np.float64)
        b = np.zeros(7, np.float64)
        used = C.c_int64()
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_pose_normal_equations_taps(
                self._h, taps.data_ptr(), pts.data_ptr(), cf.data_ptr(), pts.shape[0],
                ps.ctypes.data_as(C.c_void_p), float(lam), H.ctypes.data_as(C.c_void_p),
                b.ctypes.data_as(C.c_void_p), C.byref(used), _stream_ptr(self.device)),
                "pose_normal_equations_taps")
        return H, b, int(used.value)

    def stats(self, which: int = 0) -> Dict[str, float]:
        """Reference signature (global_volume.py:136-140)."""
        a, b = C.c_int64(), C.c_int64()
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_stats(self._h, which, C.byref(a), C.byref(b),
                                                 _stream_ptr(self.device)), "stats")
        return {"valid_voxels": a.value, "total_voxels": b.value}

    def counters(self, check: bool = True) -> Dict[str, int]:
        """rays, applied updates per volume, blocks, dropped, overflow, culled rays
        since create / reset (synchronises).  Raises when the voxel-block hash
        overflowed (B200TSDF_E_CAPACITY: updates were lost) unless check=False."""
        out = (C.c_int64 * 8)()
        with self._reading():
            rc = _ffi.lib().b200tsdf_counters(self._h, out, _stream_ptr(self.device))
        if rc not in (0, _ffi.E_CAPACITY) or (rc == _ffi.E_CAPACITY and check):
            _ffi.check(rc, "counters")
        names = ["rays", "updates_primary", "updates_quality", "blocks", "dropped", "overflow",
                 "culled_rays"]
        return dict(zip(names, [int(x) for x in out]))

    def check(self) -> None:
        """Raise B200TSDFError if any update was lost to a full voxel-block hash
        (the reference dict grows without limit; here `capacity_log2` bounds it)."""
        self.counters(check=True)

    def export(self, which: int = 0):
        """All voxels of one volume, sorted by key: (keys i64 (n,3), tsdf f32,
        weight f32, count u32) numpy arrays -- what reference _voxels holds."""
        n = C.c_int64()
        L = _ffi.lib()
        with self._reading():
            _ffi.check(L.b200tsdf_export_voxels(self._h, which, None, None, None, None, 0,
                                                C.byref(n), _stream_ptr(self.device)), "export")
            cap = n.value
            keys = torch.zeros(max(cap, 1), 3, dtype=torch.int64, device=self.device)
            t = torch.zeros(max(cap, 1), dtype=torch.float32, device=self.device)
            w = torch.zeros(max(cap, 1), dtype=torch.float32, device=self.device)
            c = torch.zeros(max(cap, 1), dtype=torch.int32, device=self.device)
            if cap:
                _ffi.check(L.b200tsdf_export_voxels(
                    self._h, which, keys.data_ptr(), t.data_ptr(), w.data_ptr(), c.data_ptr(),
                    cap, C.byref(n), _stream_ptr(self.device)), "export")
        k = keys[:cap].cpu().numpy()
        order = np.lexsort((k[:, 2], k[:, 1], k[:, 0]))
        return (k[order], t[:cap].cpu().numpy()[order], w[:cap].cpu().numpy()[order],
                c[:cap].cpu().numpy().astype(np.uint32)[order])

    def block_keys(self):
        n = C.c_int64()
        L = _ffi.lib()
        with self._reading():
            _ffi.check(L.b200tsdf_export_blocks(self._h, None, 0, C.byref(n),
                                                _stream_ptr(self.device)), "export_blocks")
            bk = torch.zeros(max(n.value, 1), 3, dtype=torch.int64, device=self.device)
            if n.value:
                _ffi.check(L.b200tsdf_export_blocks(self._h, bk.data_ptr(), n.value, C.byref(n),
                                                    _stream_ptr(self.device)), "export_blocks")
        k = bk[:n.value].cpu().numpy()
        return k[np.lexsort((k[:, 2], k[:, 1], k[:, 0]))]

    def extract_mesh(self, which: int = 0, out=None):
        """Marching cubes; returns verts f32 (3T,3) device tensor (triangle soup).
        One optimistic write pass into a buffer sized from the block count (or
        into `out`, a float32 (cap*3, 3) tensor to reuse); only if the mesh
        does not fit is it repeated with the exact size the first pass
        returned."""
        n = C.c_int64()
        L = _ffi.lib()
        with self._reading():
            if out is None:
                nblk = C.c_int64()
                _ffi.check(L.b200tsdf_export_blocks(self._h, None, 0, C.byref(nblk),
                                                    _stream_ptr(self.device)), "export_blocks")
                cap = max(1024, nblk.value * 96, int(getattr(self, "_last_tris", 0) * 1.25))
                out = torch.empty(cap * 3, 3, dtype=torch.float32, device=self.device)
            cap = out.shape[0] // 3
            _ffi.check(L.b200tsdf_extract_mesh(self._h, which, out.data_ptr(), cap, C.byref(n),
                                               _stream_ptr(self.device)), "extract_mesh")
            if n.value > cap:  # did not fit: exact size, second pass
                out = torch.empty(n.value * 3, 3, dtype=torch.float32, device=self.device)
                _ffi.check(L.b200tsdf_extract_mesh(self._h, which, out.data_ptr(), n.value,
                                                   C.byref(n), _stream_ptr(self.device)),
                           "extract_mesh")
            self._last_tris = n.value
        return out[: n.value * 3]

    def extract_mesh_async(self, which: int, out, count=None):
        """Marching cubes into `out` (float32 (cap*3, 3) device tensor) without any
        host synchronisation: returns `count`, a 1-element int64 device tensor
        holding the triangle count (it can exceed the capacity: the triangles
        beyond it were not written).  For device consumers and for timing the
        pass itself; `extract_mesh` is this plus the read-back of the count."""
        if count is None:
            count = torch.empty(1, dtype=torch.int64, device=self.device)
        with self._reading():
            _ffi.check(_ffi.lib().b200tsdf_extract_mesh_async(
                self._h, which, out.data_ptr(), out.shape[0] // 3, count.data_ptr(),
                _stream_ptr(self.device)), "extract_mesh_async")
        return count

    def extract_mesh_indexed(self, which: int = 0):
        """Marching cubes + weld on the device: (verts f32 (V,3), faces i32 (T,3))
        device tensors.  Shared vertices are stored once; vertex
        order is the order of first occurrence in the triangle soup."""
        nv, nt = C.c_int64(), C.c_int64()
        L = _ffi.lib()
        with self._reading():
            _ffi.check(L.b200tsdf_extract_mesh_indexed(self._h, which, None, 0, None, 0,
                                                       C.byref(nv), C.byref(nt),
                                                       _stream_ptr(self.device)), "extract_mesh")
            verts = torch.empty(max(nv.value, 1), 3, dtype=torch.float32, device=self.device)
            faces = torch.empty(max(nt.value, 1), 3, dtype=torch.int32, device=self.device)
            if nt.value:
                _ffi.check(L.b200tsdf_extract_mesh_indexed(
                    self._h, which, verts.data_ptr(), nv.value, faces.data_ptr(), nt.value,
                    C.byref(nv), C.byref(nt), _stream_ptr(self.device)), "extract_mesh_indexed")
        return verts[: nv.value], faces[: nt.value]

    # ------------------------------------------------------------------
    # Multi-GPU seams (see parallel.ShardedTSDFVolume)
    # ------------------------------------------------------------------
    def _export_payload(self, fn, what):
        n = C.c_int64()
        with self._reading():
            _ffi.check(fn(self._h, None, None, None, 0, C.byref(n), _stream_ptr(self.device)), what)
            cap = n.value
            bk = torch.zeros(cap, 3, dtype=torch.int64, device=self.device)
            vox = torch.zeros(cap, 512, 4, dtype=torch.float32, device=self.device)
            cnt = torch.zeros(cap, 512, dtype=torch.int64, device=self.device)
            if cap:
                _ffi.check(fn(self._h, bk.data_ptr(), vox.data_ptr(), cnt.data_ptr(), cap,
                              C.byref(n), _stream_ptr(self.device)), what)
        return bk, vox, cnt

    def export_seam_blocks(self):
        """Voxel blocks of the first block layer of every z-slab this volume
        owns: (bkeys i64 (n,3), vox f32 (n,512,4), cnt i64 (n,512)) device
        tensors (cnt holds the raw u64 hit-counter words)."""
        return self._export_payload(_ffi.lib().b200tsdf_export_seam_blocks, "export_seam")

    def export_block_data(self):
        """Every voxel block this volume owns, same layout as export_seam_blocks:
        the complete state of the volume."""
        return self._export_payload(_ffi.lib().b200tsdf_export_block_data, "export_block_data")

    def import_blocks(self, bkeys, vox, cnt):
        """ADD blocks exported by `export_block_data` as owned blocks.  Sums and
        hit counters are additive: into an empty volume this restores a
        checkpoint, into a filled one it merges two partial fusions."""
        n = int(bkeys.shape[0])
        if n == 0:
            return
        bkeys = bkeys.to(self.device, torch.int64).contiguous()
        vox = vox.to(self.device, torch.float32).contiguous()
        cnt = cnt.to(self.device, torch.int64).contiguous()
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_import_blocks(
                self._h, bkeys.data_ptr(), vox.data_ptr(), cnt.data_ptr(), n,
                _stream_ptr(self.device)), "import_blocks")

    def save_checkpoint(self, path):
        """Flat .npz of (block keys, packed voxels, hit counters) + the volume
        parameters; `load_checkpoint` restores it bit-exactly."""
        import numpy as np

        bk, vox, cnt = self.export_block_data()
        order = np.lexsort(bk.cpu().numpy().T[::-1])  # deterministic file contents
        np.savez(path, block_keys=bk.cpu().numpy()[order], vox=vox.cpu().numpy()[order],
                 cnt=cnt.cpu().numpy()[order],
                 params=np.array([self.voxel_size, self.truncation, self.max_weight,
                                  self.min_weight], np.float64))
        return int(bk.shape[0])

    def load_checkpoint(self, path, check_params=True):
        import numpy as np

        with np.load(path) as z:
            if check_params:
                mine = np.array([self.voxel_size, self.truncation, self.max_weight,
                                 self.min_weight], np.float64)
                if not np.allclose(z["params"], mine, rtol=1e-12, atol=0.0):
                    raise ValueError(f"checkpoint parameters {z['params']} != volume {mine}")
            self.import_blocks(torch.from_numpy(z["block_keys"]), torch.from_numpy(z["vox"]),
                               torch.from_numpy(z["cnt"]))
            return int(z["block_keys"].shape[0])

    def import_halo_blocks(self, bkeys, vox, cnt):
        """Install another rank's seam blocks as read-only +z neighbours for
        extract_mesh."""
        n = int(bkeys.shape[0])
        if n == 0:
            return
        bkeys = bkeys.to(self.device, torch.int64).contiguous()
        vox = vox.to(self.device, torch.float32).contiguous()
        cnt = cnt.to(self.device, torch.int64).contiguous()
        with self._writing():
            _ffi.check(_ffi.lib().b200tsdf_import_halo_blocks(
                self._h, bkeys.data_ptr(), vox.data_ptr(), cnt.data_ptr(), n,
                _stream_ptr(self.device)), "import_halo")
