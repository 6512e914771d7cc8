"""B200 drop-in for the reference `mast3r_slam/tsdf/global_manager.py`.

Mirrored CONTRACT (what `main.py:80-90` and the SLAM loop touch; these lines
cannot look different): the three class names, their constructor arguments,
`start / shutdown / on_after_backend_solve / mark_pose_update / enqueue /
request_stop / run`, the attribute names `volume, optimizer, integrator,
opt_worker, next_idx, pending, reintegration_queue` and every `tsdf_global`
config key with the reference's code default (global_manager.py:16-229,
config/base.yaml:110-125).

Everything behind that contract is built for the device instead of
transcribed:

  * The integrator is a BATCHING worker.  The reference fuses one keyframe per
    loop turn (clone -> subsample -> Sim3 on device -> copy to host -> Python
    ray loop, :64-106).  Here a wake-up gathers ALL new keyframes and ALL dirty
    ones, snapshots them under `keyframes.lock` with one device gather per
    array (SharedKeyframes holds stacked CUDA buffers, frame.py:236-252),
    releases the lock, and fuses each group with ONE
    `b200tsdf_integrate_keyframes_ex` launch -- no per-keyframe launch, no
    per-keyframe synchronisation.  The fold is commutative, so a
    batch equals the reference's one-by-one order.
  * `max_points_per_kf` is honoured on the device: exactly
    min(n_valid, max_points) pixels per keyframe, chosen by a seeded bijective
    hash of the pixel index (`tsdf_global.seed`); the reference draws an
    unseeded `torch.randperm` (:94-96), which no implementation can bit-match.
  * With `pre_icp_iters > 0` (base.yaml ships 2) the reference refines the pose
    of keyframe i against the volume right after fusing it and before fusing
    i + 1 (:66-68).  That order is kept: new keyframes are then fused one per
    launch with `pre_refine` in between, unless `tsdf_global.batch_with_pre_icp`
    is set (refinement then sees the whole batch).  Dirty keyframes are always
    batched (the reference runs no pre-refinement on them, :71-79).
  * Hash overflow is an error, not a log line: after every wake-up the worker
    checks the overflow counter and raises `B200TSDFError` (the reference dict
    cannot overflow; `capacity_log2` bounds this volume).
  * `tsdf_global.num_gpus > 1` builds a `ShardedTSDFVolume` (z-slabs over the
    ranks of the default process group); keyframes reach every rank by one
    broadcast per batch from `tsdf_global.src_rank`.

New optional keys (old YAML still loads): `capacity_log2`, `bounds` (voxel-key
box) or `volume_dim` + `volume_origin` (metres), `quality_volume`, `num_gpus`,
`slab_width`, `src_rank`, `seed`, `parity_mode` (only `reference_scatter`),
`integrate_batch`, `batch_with_pre_icp`, `reintegration_mode`
(`additive` = reference | `replace`), `patch_size`, `pin`.
"""
import math
import queue
import threading
import time
from typing import Iterable, List, Optional

import torch

from .. import _ffi
from . import sim3
from .global_volume import TSDFVolume
from .tsdf_optimizer import TSDFPoseOptimizer


def quality_grid_from_result(result, gh, gw, device):
    """Per-patch quality weight q = 1 - priority from the
    dict `quality_core.pack_result` produces (quality_core.py:111-125); None
    where no usable result exists (q = 1)."""
    if result is None or "priority" not in result:
        return None
    pr = torch.as_tensor(result["priority"], dtype=torch.float32).reshape(-1)
    if pr.numel() != gh * gw:
        return None
    return (1.0 - pr.clamp(0.0, 1.0)).reshape(1, gh, gw).to(device).contiguous()


def bounds_from_cfg(cfg):
    """Voxel-key box of the volume: `bounds` ((x0,y0,z0),(x1,y1,z1)) as is, or
    `volume_dim` (int or 3 ints, voxels) with `volume_origin` (metres, the
    minimum corner; default: the box is centred on the world origin).  The
    minimum corner is rounded down to a multiple of 8 voxels so that voxel
    blocks (and z-slabs) are aligned with it.  None: unbounded like the
    reference dict."""
    b = cfg.get("bounds", None)
    if b is not None:
        return (tuple(int(v) for v in b[0]), tuple(int(v) for v in b[1]))
    dim = cfg.get("volume_dim", None)
    if dim is None:
        return None
    dims = [int(dim)] * 3 if isinstance(dim, (int, float)) else [int(d) for d in dim]
    vs = float(cfg.get("voxel_size", 0.03))
    org = cfg.get("volume_origin", None)
    if org is None:
        k0 = [-(d // 2) for d in dims]
    else:
        k0 = [int(math.floor(float(o) / vs)) for o in org]
    k0 = [(k >> 3) << 3 for k in k0]
    return (tuple(k0), tuple(k + d for k, d in zip(k0, dims)))


class _Snapshot:
    """Device copies of a group of keyframes taken under `keyframes.lock`."""

    __slots__ = ("idx", "X", "C", "pose", "H", "W", "fids")

    def __init__(self, idx, X, C, pose, H, W, fids):
        self.idx, self.X, self.C, self.pose, self.H, self.W, self.fids = idx, X, C, pose, H, W, fids


class TSDFGlobalIntegrator(threading.Thread):
    """Background worker that fuses keyframes into the global dual TSDF."""

    def __init__(self, volume, keyframes, cfg, optimizer: TSDFPoseOptimizer,
                 quality_service=None) -> None:
        super().__init__(daemon=True)
        self.volume = volume
        self.keyframes = keyframes
        self.cfg = cfg
        self.optimizer = optimizer
        self.quality_service = quality_service
        self.next_idx = 0
        self.sleep_interval = float(cfg.get("queue_check_interval", 0.1))
        self.stop_flag = threading.Event()
        self.reintegration_queue: "queue.Queue[int]" = queue.Queue(
            maxsize=int(cfg.get("reintegration_queue", 256)))
        self.pending = set()
        self.max_points = int(cfg.get("max_points_per_kf", 40000))
        self.min_conf = float(cfg.get("min_confidence", 0.05))
        self.log_interval = float(cfg.get("log_interval", 30.0))
        self.patch_size = int(cfg.get("patch_size", 16))
        self.seed = int(cfg.get("seed", 0))
        self.batch = max(1, int(cfg.get("integrate_batch", 64)))
        self.use_quality = bool(cfg.get("quality_volume", quality_service is not None))
        self.batch_with_pre_icp = bool(cfg.get("batch_with_pre_icp", False))
        # "additive" = the reference (old contribution stays, :71-79); "replace" =
        # de-integrate the keyframe as it was fused, then fuse it at the new pose
        self.reintegration_mode = str(cfg.get("reintegration_mode", "additive"))
        if self.reintegration_mode not in ("additive", "replace"):
            raise ValueError("tsdf_global.reintegration_mode must be 'additive' or 'replace'")
        self._last_log = time.time()
        self.integrated = 0
        self.launches = 0
        self._fused = {}  # replace mode: idx -> (X, C, pose, q, H, W) exactly as fused
        self._sharded = hasattr(volume, "vol")  # parallel.ShardedTSDFVolume
        self.src_rank = int(cfg.get("src_rank", 0))
        self.error = None

    # -- contract ------------------------------------------------------------
    def request_stop(self) -> None:
        self.stop_flag.set()

    def mark_pose_update(self, indices: Iterable[int]) -> None:
        """Queue keyframes whose pose the backend moved (global_manager.py:46-54):
        once each, silently dropping what the bounded queue cannot take."""
        for idx in indices:
            if idx in self.pending:
                continue
            try:
                self.reintegration_queue.put_nowait(idx)
            except queue.Full:
                break
            self.pending.add(idx)

    def run(self) -> None:
        while not self.stop_flag.is_set():
            try:
                self.step()
            except Exception as e:  # noqa: BLE001 -- surface it to the owner thread, then stop
                self.error = e
                raise
            time.sleep(self.sleep_interval)

    def step(self) -> None:
        """One wake-up (callable synchronously in tests and benches)."""
        fused = self._integrate_new_keyframes()
        fused += self._process_dirty_queue()
        if fused:
            self._check_capacity()
        self._maybe_log_stats()

    # -- gathering -------------------------------------------------------------
    def _frame_hw(self, n, shape=None):
        if shape is not None:
            hw = torch.as_tensor(shape).reshape(-1)
            if hw.numel() >= 2 and int(hw[0]) * int(hw[1]) == n:
                return int(hw[0]), int(hw[1])
        kh, kw = getattr(self.keyframes, "h", None), getattr(self.keyframes, "w", None)
        if kh and kw and int(kh) * int(kw) == n:
            return int(kh), int(kw)
        return 1, n  # no image shape: one row (the fused kernel only needs H*W = N)

    def _snapshot(self, indices: List[int]) -> Optional[_Snapshot]:
        """Copies of keyframes `indices` (those that exist) on the volume's
        device, taken under the keyframe lock like the reference's clones
        (:82-88): one gather per stacked SharedKeyframes buffer, or a stack of
        per-frame tensors for any other container."""
        dev = self.volume.device
        kfs = self.keyframes
        with kfs.lock:
            n_kf = len(kfs)
            idx = [i for i in indices if 0 <= i < n_kf]
            if not idx:
                return None
            Xb, Cb, Tb = (getattr(kfs, a, None) for a in ("X", "C", "T_WC"))
            stacked = all(torch.is_tensor(t) and t.dim() >= 2 for t in (Xb, Cb, Tb))
            if stacked:
                sel = torch.as_tensor(idx, dtype=torch.long, device=Xb.device)
                X = Xb.index_select(0, sel).reshape(len(idx), -1, 3)
                Cf = Cb.index_select(0, sel).reshape(len(idx), -1)
                pose = Tb.index_select(0, sel).reshape(len(idx), 8)
                shp = getattr(kfs, "img_shape", None)
                shape0 = shp[idx[0]] if torch.is_tensor(shp) else None
                did = getattr(kfs, "dataset_idx", None)
                fids = [int(v) for v in did.index_select(0, sel.to(did.device)).cpu()] \
                    if torch.is_tensor(did) else list(idx)
            else:
                frames = [kfs[i] for i in idx]
                X = torch.stack([f.X_canon.detach().reshape(-1, 3) for f in frames])
                Cf = torch.stack([f.C.detach().reshape(-1) for f in frames])
                pose = torch.stack([sim3.pose_data(f.T_WC).reshape(8) for f in frames])
                shape0 = getattr(frames[0], "img_shape", None)
                fids = [int(getattr(f, "frame_id", i)) for f, i in zip(frames, idx)]
            X = X.to(dev, torch.float32).contiguous()
            Cf = Cf.to(dev, torch.float32).contiguous()
            pose = pose.to(dev, torch.float32).contiguous()
            if not stacked:  # torch.stack already copied; gathers above are copies too
                pass
        H, W = self._frame_hw(X.shape[1], shape0)
        return _Snapshot(idx, X, Cf, pose, H, W, fids)

    def _quality_grids(self, snap: _Snapshot):
        """(B, gh, gw) per-patch weights q = 1 - priority for the snapshot (ones
        where the quality service has no result), or None when the second
        volume takes the plain confidences."""
        if not self.use_quality or self.quality_service is None:
            return None
        ps = self.patch_size
        if snap.H % ps or snap.W % ps:
            return None
        gh, gw = snap.H // ps, snap.W // ps
        dev = self.volume.device
        grids, any_q = [], False
        for fid in snap.fids:
            q = quality_grid_from_result(self.quality_service.get(fid), gh, gw, dev)
            any_q = any_q or q is not None
            grids.append(q if q is not None else torch.ones(1, gh, gw, device=dev))
        return torch.cat(grids, 0).contiguous() if any_q else None

    # -- fusing ------------------------------------------------------------------
    def _fuse(self, snap: _Snapshot) -> int:
        """ONE launch for the whole snapshot (two in replace mode)."""
        B = len(snap.idx)
        q = self._quality_grids(snap)
        cap = self.max_points if 0 < self.max_points < snap.H * snap.W else 0
        kw = dict(min_conf=self.min_conf, H=snap.H, W=snap.W)
        sub = dict(max_points=cap, seed=self.seed, kf_ids=snap.idx) if cap else {}
        vol = self.volume
        if self.reintegration_mode == "replace":
            old = [(i, self._fused.pop(i)) for i in snap.idx if i in self._fused]
            for i, (X, Cf, pose, oq, H, W) in old:  # shapes may differ: one call each
                self._deintegrate(X, Cf, pose, oq, H, W, cap, i)
            for b, i in enumerate(snap.idx):
                self._fused[i] = (snap.X[b:b + 1], snap.C[b:b + 1], snap.pose[b:b + 1],
                                  None if q is None else q[b:b + 1], snap.H, snap.W)
        if self._sharded:  # the batch lives on src_rank: one broadcast, every rank fuses its slabs
            vol.integrate_keyframes(snap.X, snap.C, snap.pose, q, self.min_conf, snap.H, snap.W,
                                    src=self.src_rank, broadcast=True, **sub)
        else:
            vol.integrate_keyframes(snap.X, snap.C, snap.pose, q, **kw, **sub)
        self.launches += 1
        self.integrated += B
        return B

    def _deintegrate(self, X, Cf, pose, q, H, W, cap, idx):
        sub = dict(max_points=cap, seed=self.seed, kf_ids=[idx]) if cap else {}
        target = self.volume.vol if self._sharded else self.volume
        target.deintegrate_keyframes(X, Cf, pose, q, self.min_conf, H, W, **sub)

    def _integrate_new_keyframes(self) -> int:
        total = len(self.keyframes)
        fused = 0
        ordered = self.optimizer.pre_icp_iters > 0 and not self.batch_with_pre_icp
        step = 1 if ordered else self.batch
        while self.next_idx < total and not self.stop_flag.is_set():
            chunk = list(range(self.next_idx, min(total, self.next_idx + step)))
            snap = self._snapshot(chunk)
            if snap is not None:
                fused += self._fuse(snap)
            for i in chunk:  # no-op unless pre_icp_iters > 0 (tsdf_optimizer.py:29-37)
                self.optimizer.pre_refine(i)
            self.next_idx = chunk[-1] + 1
        return fused

    def _process_dirty_queue(self) -> int:
        todo = []
        while not self.stop_flag.is_set():
            try:
                idx = self.reintegration_queue.get_nowait()
            except queue.Empty:
                break
            self.pending.discard(idx)
            todo.append(idx)
        fused = 0
        for a in range(0, len(todo), self.batch):
            snap = self._snapshot(todo[a:a + self.batch])
            if snap is not None:
                fused += self._fuse(snap)
        return fused

    def _check_capacity(self) -> None:
        """One synchronising read per wake-up: raises `B200TSDFError`
        (B200TSDF_E_CAPACITY) when the voxel-block hash lost updates."""
        target = self.volume.vol if self._sharded else self.volume
        target.check()

    def _maybe_log_stats(self) -> None:
        now = time.time()
        if now - self._last_log < self.log_interval:
            return
        self._last_log = now
        st = self.volume.stats()
        print(f"[TSDF-GLOBAL] voxels={st['valid_voxels']}/{st['total_voxels']}"
              f"\tnext_idx={self.next_idx}\tlaunches={self.launches}")


class TSDFGlobalOptThread(threading.Thread):
    """Pose refinement against the volume off the integrator's thread
    (global_manager.py:118-175): a bounded, de-duplicated index queue drained
    `max_opt_batch` keyframes at a time on a low-priority stream.  The volume
    orders its readers and writers on the device (TSDFVolume._reading /
    _writing), so a fuse queued by the integrator cannot overlap a query issued
    here."""

    def __init__(self, optimizer: TSDFPoseOptimizer, keyframes, cfg) -> None:
        super().__init__(daemon=True)
        self.optimizer = optimizer
        self.keyframes = keyframes
        self.cfg = cfg
        self.stop_flag = threading.Event()
        self.sleep_interval = float(cfg.get("queue_check_interval", 0.1))
        self.queue: "queue.Queue[int]" = queue.Queue(
            maxsize=int(cfg.get("opt_queue", cfg.get("reintegration_queue", 256))))
        self.pending = set()
        self.max_opt_batch = int(cfg.get("max_opt_batch", 1))
        self.cooldown_ms = float(cfg.get("opt_cooldown_ms", 50.0))
        self._stream = None
        dev = torch.device(self.optimizer.device)
        if dev.type == "cuda" and torch.cuda.is_available():
            self._stream = torch.cuda.Stream(device=dev, priority=1)  # :131-138

    def request_stop(self) -> None:
        self.stop_flag.set()

    def enqueue(self, indices: Iterable[int]) -> None:
        for idx in indices:
            if idx in self.pending:
                continue
            try:
                self.queue.put_nowait(idx)
            except queue.Full:
                break
            self.pending.add(idx)

    def drain_once(self) -> int:
        """One batch of the loop; returns the number of keyframes optimised."""
        batch = []
        while len(batch) < self.max_opt_batch and not self.stop_flag.is_set():
            try:
                idx = self.queue.get_nowait()
            except queue.Empty:
                break
            if idx not in batch:
                batch.append(idx)
        if not batch:
            return 0
        try:
            if self._stream is None:
                self.optimizer.optimize_keyframes(batch, context="async")
            else:
                with torch.cuda.stream(self._stream):
                    self.optimizer.optimize_keyframes(batch, context="async")
                self._stream.synchronize()
        finally:
            self.pending.difference_update(batch)
        return len(batch)

    def run(self) -> None:
        pause = self.sleep_interval + self.cooldown_ms / 1000.0
        while not self.stop_flag.is_set():
            self.drain_once()
            time.sleep(pause)


def _global_pin(cfg) -> int:
    """`config["local_opt"]["pin"]` of the reference's global config
    (global_manager.py:221); `tsdf_global.pin` when that module is absent."""
    try:
        from mast3r_slam.config import config  # the reference package, when present

        return int(config.get("local_opt", {}).get("pin", 1))
    except Exception:  # noqa: BLE001
        return int(cfg.get("pin", 1))


def build_volume(cfg, device):
    """The volume the `tsdf_global` section describes: a `TSDFVolume`, or for
    `num_gpus > 1` a `parallel.ShardedTSDFVolume` over the default process group."""
    mode = str(cfg.get("parity_mode", "reference_scatter"))
    if mode != "reference_scatter":
        raise ValueError(f"tsdf_global.parity_mode = {mode!r}: only 'reference_scatter' (the "
                         "reference's per-point ray-march scatter, global_volume.py:53-71) is built")
    common = dict(voxel_size=cfg.get("voxel_size", 0.03), truncation=cfg.get("trunc_dist", 0.12),
                  max_weight=cfg.get("max_weight", 100.0),
                  min_weight=cfg.get("min_tsdf_weight", 1.0e-3))
    bounds = bounds_from_cfg(cfg)
    cap = int(cfg.get("capacity_log2", 17))
    n = int(cfg.get("num_gpus", 1))
    if n <= 1:
        return TSDFVolume(**common, capacity_log2=cap, bounds=bounds, device=device)
    import torch.distributed as dist

    from ..parallel import ShardedTSDFVolume

    if not dist.is_initialized() or dist.get_world_size() != n:
        raise _ffi.B200TSDFError(
            f"tsdf_global.num_gpus = {n} needs torch.distributed initialised with {n} ranks "
            "(one process per GPU)")
    if bounds is None:
        raise ValueError("tsdf_global.num_gpus > 1 needs a bounded volume (volume_dim or bounds)")
    dim_z = bounds[1][2] - bounds[0][2]
    width = int(cfg.get("slab_width", 0)) or max(8, 1 << int(math.log2(max(8, dim_z // (4 * n)))))
    return ShardedTSDFVolume(common["voxel_size"], common["truncation"], common["max_weight"],
                             common["min_weight"], slab_width=width, bounds=bounds,
                             capacity_log2=cap, device=device)


class TSDFGlobalManager:
    """High-level orchestrator for global TSDF fusion and optimization."""

    def __init__(self, keyframes, cfg, use_calib: bool, device, quality_service=None) -> None:
        self.enabled = bool(cfg.get("enabled", False))
        self.keyframes = keyframes
        self.cfg = cfg
        self.volume = build_volume(cfg, device)
        self.optimizer = TSDFPoseOptimizer(self.volume, keyframes, cfg, use_calib, device)
        self.integrator = TSDFGlobalIntegrator(self.volume, keyframes, cfg, self.optimizer,
                                               quality_service)
        self.async_opt = bool(cfg.get("async_optimize", True))
        self.opt_worker = TSDFGlobalOptThread(self.optimizer, keyframes, cfg) \
            if self.async_opt else None

    def _workers(self):
        return [w for w in (self.integrator, self.opt_worker) if w is not None]

    def start(self) -> None:
        if not self.enabled:
            return
        for w in self._workers():
            if not w.is_alive():
                w.start()

    def shutdown(self) -> None:
        if not self.enabled:
            return
        for w in self._workers():
            w.request_stop()
        for w in self._workers():
            if w.is_alive():
                w.join(timeout=2.0)

    def on_after_backend_solve(self, factor_graph) -> None:
        """The backend's hook (global_manager.py:214-229): keyframes of the solve
        at or above `local_opt.pin` are queued for re-fusion and pose refinement."""
        if not self.enabled:
            return
        touched = getattr(factor_graph, "last_unique_kf_idx", None)
        if touched is None:
            return
        pin = _global_pin(self.cfg)
        indices = [i for i in (int(v) for v in torch.as_tensor(touched).cpu().tolist()) if i >= pin]
        if not indices:
            return
        self.integrator.mark_pose_update(indices)
        if self.opt_worker is not None:
            self.opt_worker.enqueue(indices)
        else:
            self.optimizer.optimize_keyframes(indices, context="factor")


class _BenchKeyframes:
    """SharedKeyframes-shaped container of device-resident synthetic keyframes
    (stacked buffers + an RLock), for `bench_manager`."""

    def __init__(self, X, C, pose, H, W):
        self.lock = threading.RLock()
        self.X, self.C = X, C.reshape(C.shape[0], -1, 1)
        self.T_WC = pose.reshape(-1, 1, 8)
        self.h, self.w = H, W
        self.n = 0

    def __len__(self):
        return self.n


def bench_manager(wl, device, n_kf, H, W, gh, gw, voxel, trunc, min_conf, gen_batch):
    """keyframes/s through `TSDFGlobalManager` -- the API `main.py:80-90` calls --
    with device-resident keyframes, every pixel fused (max_points_per_kf = 0) so
    that the work equals the raw kernel's, and with the reference's default cap."""
    from .. import synth

    lo, hi = synth.bounds_for(wl["scene"], voxel, wl["dim"])
    parts = [gen_batch(wl, l, 40) for l in range(n_kf // 40)]
    X = torch.cat([torch.from_numpy(p["X"]) for p in parts]).to(device)
    Cf = torch.cat([torch.from_numpy(p["C"]) for p in parts]).to(device)
    pose = torch.cat([torch.from_numpy(p["pose"]) for p in parts]).to(device)
    out = {}
    for name, cap in (("all_pixels", 0), ("max_points_40000", 40000)):
        kfs = _BenchKeyframes(X, Cf, pose, H, W)
        cfg = {"enabled": True, "voxel_size": voxel, "trunc_dist": trunc, "min_confidence": min_conf,
               "max_points_per_kf": cap, "capacity_log2": wl["cap_log2"], "bounds": (lo, hi),
               "pre_icp_iters": 0, "async_optimize": False, "integrate_batch": 40}
        mgr = TSDFGlobalManager(kfs, cfg, False, device)
        kfs.n = 40
        mgr.integrator.step()  # warm-up wake-up
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        kfs.n = n_kf
        mgr.integrator.step()  # one wake-up fuses the 560 pending keyframes in 14 launches
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        c = mgr.volume.counters()
        out[name] = {"keyframes_per_s": (n_kf - 40) / dt, "seconds": dt,
                     "launches": mgr.integrator.launches - 1,
                     "updates": c["updates_primary"] + c["updates_quality"]}
        # the launches themselves on the device (CUDA events): the orbit's 40-keyframe
        # batches through TSDFVolume.integrate_keyframes with the same cap
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        n_l = n_kf // 40
        ev[0].record()
        for l in range(n_l):
            sl = slice(40 * l, 40 * l + 40)
            mgr.volume.integrate_keyframes(X[sl], Cf[sl], pose[sl], None, min_conf, H, W,
                                           max_points=cap, seed=0)
        ev[1].record()
        torch.cuda.synchronize()
        out[name]["device_ms_per_launch"] = ev[0].elapsed_time(ev[1]) / n_l
        del mgr
    out["what"] = ("TSDFGlobalIntegrator.step(): snapshot under the keyframe lock (one device "
                   "gather per array) + one fused launch per 40 keyframes + one overflow check; "
                   "host wall clock incl. the final synchronise")
    return out
