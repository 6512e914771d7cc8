"""B200 mirror of the reference `mast3r_slam/tsdf/global_manager.py`.

Same three classes, constructor arguments, methods and `tsdf_global` config
keys (global_manager.py:16-229).  What changes on the hot path:
`_integrate_single` (:81-106) no longer clones, subsamples, transforms and
copies the keyframe to the host -- it hands the device tensors SharedKeyframes
already holds to ONE fused `b200tsdf_integrate_keyframes` launch (staging +
block allocation + dual fold).  Queueing, de-duplication, thread lifetime and
the additive re-integration after `global_opt` (:46-79, :214-229) are kept as
the reference has them.

New config keys (all optional): `capacity_log2`, `bounds`, `quality_service`
is passed as a constructor keyword, not through the config.
"""
import queue
import threading
import time
from typing import Iterable

import torch

from . import sim3
from .global_volume import TSDFVolume
from .tsdf_optimizer import TSDFPoseOptimizer


def quality_grid_from_result(result, gh, gw, device):
    """Per-patch quality weight q = 1 - priority from the
    dict `quality_core.pack_result` produces (quality_core.py:111-125); q = 1
    where no result exists."""
    if result is None or "priority" not in result:
        return None
    pr = torch.as_tensor(result["priority"], dtype=torch.float32).reshape(-1)
    if pr.numel() != gh * gw:
        return None
    return (1.0 - pr.clamp(0.0, 1.0)).reshape(1, gh, gw).to(device).contiguous()


class TSDFGlobalIntegrator(threading.Thread):
    """Background thread that fuses keyframes into the global TSDF."""

    def __init__(self, volume: TSDFVolume, keyframes, cfg, optimizer: TSDFPoseOptimizer,
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
        self.max_points = int(cfg.get("max_points_per_kf", 40000))  # kept for cfg parity; unused
        self.min_conf = float(cfg.get("min_confidence", 0.05))
        self.log_interval = float(cfg.get("log_interval", 30.0))
        self.patch_size = int(cfg.get("patch_size", 16))
        self._last_log = time.time()
        self.integrated = 0
        # "additive" = the reference (old contribution stays, global_manager.py:71-79);
        # "replace" = de-integrate the keyframe at the pose it was fused with, then fuse
        # it at the new pose
        self.reintegration_mode = str(cfg.get("reintegration_mode", "additive"))
        self._fused = {}  # idx -> (X, C, pose, q, H, W) exactly as fused

    def request_stop(self) -> None:
        self.stop_flag.set()

    def mark_pose_update(self, indices: Iterable[int]) -> None:  # :46-54
        for idx in indices:
            if idx in self.pending:
                continue
            try:
                self.reintegration_queue.put_nowait(idx)
                self.pending.add(idx)
            except queue.Full:
                break

    def run(self) -> None:  # :56-61
        while not self.stop_flag.is_set():
            self.step()
            time.sleep(self.sleep_interval)

    def step(self) -> None:
        """One pass of the thread loop (callable synchronously in tests)."""
        self._integrate_new_keyframes()
        self._process_dirty_queue()
        self._maybe_log_stats()

    # ------------------------------------------------------------------
    def _integrate_new_keyframes(self) -> None:  # :64-69
        total = len(self.keyframes)
        while self.next_idx < total and not self.stop_flag.is_set():
            self._integrate_single(self.next_idx)
            self.optimizer.pre_refine(self.next_idx)
            self.next_idx += 1

    def _process_dirty_queue(self) -> None:  # :71-79
        while not self.reintegration_queue.empty() and not self.stop_flag.is_set():
            try:
                idx = self.reintegration_queue.get_nowait()
            except queue.Empty:
                break
            self.pending.discard(idx)
            if idx < len(self.keyframes):
                self._integrate_single(idx)

    def _frame_hw(self, frame, n):
        shp = getattr(frame, "img_shape", None)
        if shp is not None:
            hw = torch.as_tensor(shp).reshape(-1)
            if hw.numel() >= 2 and int(hw[0]) * int(hw[1]) == n:
                return int(hw[0]), int(hw[1])
        return 1, n  # no image shape: one row (the fused kernel only needs H*W = N)

    def _integrate_single(self, idx: int) -> None:
        dev = self.volume.device
        with self.keyframes.lock:  # :82-88 (the call is stream-ordered: no clone needed)
            if idx >= len(self.keyframes):
                return
            frame = self.keyframes[idx]
            X = frame.X_canon.detach().reshape(1, -1, 3).to(dev, torch.float32).contiguous()
            Cf = frame.C.detach().reshape(1, -1).to(dev, torch.float32).contiguous()
            pose = sim3.pose_data(frame.T_WC).reshape(1, 8).to(dev, torch.float32).contiguous()
            n = X.shape[1]
            H, W = self._frame_hw(frame, n)
            q = None
            if self.quality_service is not None and H % self.patch_size == 0 \
                    and W % self.patch_size == 0:
                fid = getattr(frame, "frame_id", idx)
                q = quality_grid_from_result(self.quality_service.get(fid),
                                             H // self.patch_size, W // self.patch_size, dev)
            if self.reintegration_mode == "replace":
                old = self._fused.pop(idx, None)
                if old is not None:
                    self.volume.deintegrate_keyframes(old[0], old[1], old[2], old[3],
                                                      self.min_conf, old[4], old[5])
                # private copies: the removal must see exactly what was fused
                X, Cf, pose = X.clone(), Cf.clone(), pose.clone()
                self._fused[idx] = (X, Cf, pose, q, H, W)
            self.volume.integrate_keyframes(X, Cf, pose, q, self.min_conf, H, W)
            torch.cuda.current_stream(dev).synchronize()  # inputs may change after the lock
        self.integrated += 1

    def _maybe_log_stats(self) -> None:  # :108-115
        if time.time() - self._last_log < self.log_interval:
            return
        self._last_log = time.time()
        stats = self.volume.stats()
        print(f"[TSDF-GLOBAL] voxels={stats['valid_voxels']}/{stats['total_voxels']}"
              f"\tnext_idx={self.next_idx}")


class TSDFGlobalOptThread(threading.Thread):  # :118-175
    def __init__(self, optimizer: TSDFPoseOptimizer, keyframes, cfg) -> None:
        super().__init__(daemon=True)
        self.optimizer = optimizer
        self.keyframes = keyframes
        self.cfg = cfg
        self.stop_flag = threading.Event()
        self.sleep_interval = float(cfg.get("queue_check_interval", 0.1))
        max_q = int(cfg.get("opt_queue", cfg.get("reintegration_queue", 256)))
        self.queue: "queue.Queue[int]" = queue.Queue(maxsize=max_q)
        self.pending = set()
        self.max_opt_batch = int(cfg.get("max_opt_batch", 1))
        self.cooldown_ms = float(cfg.get("opt_cooldown_ms", 50.0))
        self._stream = None
        try:  # low-priority stream, as the reference does (:131-138)
            if torch.cuda.is_available() and str(self.optimizer.device).startswith("cuda"):
                self._stream = torch.cuda.Stream(device=self.optimizer.device, priority=1)
        except Exception:  # noqa: BLE001
            self._stream = None

    def request_stop(self) -> None:
        self.stop_flag.set()

    def enqueue(self, indices: Iterable[int]) -> None:
        for idx in indices:
            if idx in self.pending:
                continue
            try:
                self.queue.put_nowait(idx)
                self.pending.add(idx)
            except queue.Full:
                break

    def drain_once(self) -> int:
        """One batch of the thread loop; returns the number of keyframes optimised."""
        batch = []
        while not self.queue.empty() and not self.stop_flag.is_set() \
                and len(batch) < self.max_opt_batch:
            try:
                idx = self.queue.get_nowait()
            except queue.Empty:
                break
            if idx not in batch:
                batch.append(idx)
        if batch:
            try:
                if self._stream is not None:
                    self._stream.wait_stream(torch.cuda.current_stream(self.optimizer.device))
                    with torch.cuda.stream(self._stream):
                        self.optimizer.optimize_keyframes(batch, context="async")
                    self._stream.synchronize()
                else:
                    self.optimizer.optimize_keyframes(batch, context="async")
            finally:
                for i in batch:
                    self.pending.discard(i)
        return len(batch)

    def run(self) -> None:
        while not self.stop_flag.is_set():
            self.drain_once()
            time.sleep(self.sleep_interval + self.cooldown_ms / 1000.0)


def _global_pin(cfg) -> int:
    """`config["local_opt"]["pin"]` of the reference's global config
    (global_manager.py:221); `tsdf_global.pin` overrides when that module is
    not importable."""
    try:
        from mast3r_slam.config import config  # the reference package, when present

        return int(config.get("local_opt", {}).get("pin", 1))
    except Exception:  # noqa: BLE001
        return int(cfg.get("pin", 1))


class TSDFGlobalManager:
    """High-level orchestrator for global TSDF fusion and optimization."""

    def __init__(self, keyframes, cfg, use_calib: bool, device, quality_service=None) -> None:
        self.enabled = bool(cfg.get("enabled", False))
        self.keyframes = keyframes
        self.cfg = cfg
        self.volume = TSDFVolume(
            voxel_size=cfg.get("voxel_size", 0.03),
            truncation=cfg.get("trunc_dist", 0.12),
            max_weight=cfg.get("max_weight", 100.0),
            min_weight=cfg.get("min_tsdf_weight", 1.0e-3),
            capacity_log2=int(cfg.get("capacity_log2", 17)),
            bounds=cfg.get("bounds", None),
            device=device,
        )
        self.optimizer = TSDFPoseOptimizer(self.volume, keyframes, cfg, use_calib, device)
        self.integrator = TSDFGlobalIntegrator(self.volume, keyframes, cfg, self.optimizer,
                                               quality_service)
        self.async_opt = bool(cfg.get("async_optimize", True))
        self.opt_worker = TSDFGlobalOptThread(self.optimizer, keyframes, cfg) \
            if self.async_opt else None

    def start(self) -> None:  # :196-202
        if not self.enabled:
            return
        if not self.integrator.is_alive():
            self.integrator.start()
        if self.opt_worker is not None and not self.opt_worker.is_alive():
            self.opt_worker.start()

    def shutdown(self) -> None:  # :204-211
        if not self.enabled:
            return
        self.integrator.request_stop()
        if self.integrator.is_alive():
            self.integrator.join(timeout=2.0)
        if self.opt_worker is not None:
            self.opt_worker.request_stop()
            if self.opt_worker.is_alive():
                self.opt_worker.join(timeout=2.0)

    # ------------------------------------------------------------------
    def on_after_backend_solve(self, factor_graph) -> None:  # :214-229
        if not self.enabled:
            return
        idx_tensor = getattr(factor_graph, "last_unique_kf_idx", None)
        if idx_tensor is None:
            return
        indices = [int(i) for i in torch.as_tensor(idx_tensor).cpu().tolist()]
        pin = _global_pin(self.cfg)
        indices = [i for i in indices if i >= pin]
        if not indices:
            return
        self.integrator.mark_pose_update(indices)
        if self.async_opt and self.opt_worker is not None:
            self.opt_worker.enqueue(indices)
        else:
            self.optimizer.optimize_keyframes(indices, context="factor")
