"""B200 mirror of the reference `mast3r_slam/tsdf_refine.py`.

Same public surface (tsdf_refine.py:34-49, :51-137, :139-260): `BlockKey`,
`PatchBlock`, `RefineRegistry`, `TSDFRefiner(cfg, shared_keyframes,
quality_service, device)` -- a `threading.Thread` with `.queue`, `.stop_flag`,
`.stats`, `.maybe_schedule_sliding_window`, `.schedule_final_pass` -- and the
same `tsdf_refine` config keys (config/base.yaml:74-108).  The numeric methods
`_build_tsdf_robust` / `_extract_surface_safe` run as CUDA kernels through the
C ABI (local_tsdf.py); scheduling, registry and acceptance logic are host code
with the reference's semantics.  No CPU fallback for the numerics.
"""
import math
import queue
import threading
import time
from dataclasses import dataclass
from typing import List, Tuple

import numpy as np
import torch

from . import local_tsdf


@dataclass(frozen=True)
class BlockKey:  # tsdf_refine.py:33-38
    kf_id: int
    block_id: int


@dataclass
class PatchBlock:  # :40-49
    kf_id: int
    block_id: int
    patch_indices: List[Tuple[int, int]]
    pixel_mask: torch.Tensor
    depth_median: float
    priority: float
    depth_variance: float = 0.0


class RefineRegistry:
    """Per-block state machine IDLE -> QUEUED -> RUNNING -> COOLDOWN with an
    attempt history (:51-137)."""

    IDLE, QUEUED, RUNNING, COOLDOWN = 0, 1, 2, 3

    def __init__(self, cfg):
        self.lock = threading.Lock()
        self.state = {}
        self.history = {}
        self.cooldown_frames = cfg["cooldown_frames"]
        self.frame_count = 0
        self.success_rate = 0.0

    def tick(self):
        with self.lock:
            self.frame_count += 1

    def _totals(self):
        a = sum(h.get("attempts", 0) for h in self.history.values())
        s = sum(h.get("successes", 0) for h in self.history.values())
        return a, s

    def get_stats(self):
        with self.lock:
            a, s = self._totals()
            return {"active_blocks": sum(1 for st in self.state.values() if st != self.IDLE),
                    "total_attempts": a, "success_rate": s / max(1, a),
                    "frame_count": self.frame_count}

    def try_enqueue(self, key: BlockKey) -> bool:
        with self.lock:
            if self.state.get(key, self.IDLE) != self.IDLE:
                return False
            self.state[key] = self.QUEUED
            return True

    def begin_run(self, key: BlockKey) -> bool:
        with self.lock:
            if self.state.get(key) != self.QUEUED:
                return False
            self.state[key] = self.RUNNING
            return True

    def finish_run(self, key: BlockKey, success: bool, gain: float):
        with self.lock:
            h = self.history.setdefault(key, {"attempts": 0, "successes": 0, "last_gain": 0.0,
                                              "best_gain": 0.0})
            h["attempts"] += 1
            h["successes"] += 1 if success else 0
            h["last_gain"] = gain
            h["best_gain"] = max(h["best_gain"], gain)
            h["last_frame"] = self.frame_count
            self.state[key] = self.COOLDOWN
            a, s = self._totals()
            self.success_rate = s / max(1, a)

    def cooldown_expired(self, key: BlockKey) -> bool:
        with self.lock:
            h = self.history.get(key)
            if h and self.frame_count - h["last_frame"] > self.cooldown_frames:
                self.state[key] = self.IDLE
                return True
            return False


def _frame_hw(kf):
    return int(kf.img_shape[0, 0].item()), int(kf.img_shape[0, 1].item())


def _conf_hw(kf, H, W):
    return kf.C.reshape(H, W) if kf.C.ndim == 2 else kf.C.reshape(H, W, 1)[..., 0]


class TSDFRefiner(threading.Thread):
    REQUIRED = ("voxel_size", "trunc_dist", "max_grid_dim", "roi_size")  # :147-150

    def __init__(self, cfg, shared_keyframes, quality_service, device=None):
        super().__init__(daemon=True)
        self.cfg = dict(cfg) if cfg else {}
        for p in self.REQUIRED:
            if p not in self.cfg:
                raise ValueError(f"Missing required TSDF config parameter: {p}")
        self.device = torch.device(device or "cuda")
        if self.device.type != "cuda":
            raise RuntimeError("TSDFRefiner needs a CUDA device (no CPU fallback)")
        self.keyframes = shared_keyframes
        self.quality_service = quality_service
        self.registry = RefineRegistry({"cooldown_frames": int(self.cfg.get("cooldown_frames", 35))})
        self.queue = queue.Queue(maxsize=int(self.cfg.get("max_pending_tasks", 50)))
        self.stop_flag = threading.Event()
        self.quality_events = {}
        self.quality_events_lock = threading.Lock()
        if hasattr(self.quality_service, "register_callback"):  # :171-173
            self.quality_service.register_callback(self._on_quality_result)
        self.stats = {  # :176-187, read by main.py:403-440
            "total_blocks": 0, "successful_blocks": 0, "total_processing_time": 0.0,
            "last_stats_print": time.time(),
            "debug_info": {"tsdf_constructions": 0, "surface_extractions": 0,
                           "displacement_rejects": 0, "hit_ratio_rejects": 0},
        }
        self._pending_map = {}
        self.generator = None  # optional torch.Generator for the pixel choice (:965 is unseeded)
        self.pixel_order = None  # optional callable n -> permutation (overrides the generator)

    def _on_quality_result(self, result):  # :202-209
        kf_id = result.get("kf_id")
        if kf_id is not None:
            with self.quality_events_lock:
                if kf_id in self.quality_events:
                    self.quality_events[kf_id].set()

    # ------------------------------------------------------------ scheduling
    def schedule_final_pass(self, final_kf_id: int):  # :246-258
        self.maybe_schedule_sliding_window(final_kf_id, is_final_pass=True)
        time.sleep(float(self.cfg.get("final_pass_wait_s", 1.0)))
        self.maybe_schedule_sliding_window(final_kf_id, is_final_pass=True)

    def _retry_later(self, kf_id, ent, now, max_attempts):
        ent["attempts"] += 1
        ent["next_ts"] = now + 0.02 * (2 ** max(0, ent["attempts"] - 1))  # exponential back-off
        if ent["attempts"] <= max_attempts:
            self._pending_map[kf_id] = ent
        else:
            self._pending_map.pop(kf_id, None)

    def maybe_schedule_sliding_window(self, current_kf_id: int, is_final_pass: bool = False):
        """:260-346 -- schedule keyframe `current - window`, retry failed ones
        with back-off, bound the retry map."""
        if not self.cfg["enabled"]:
            return
        window = int(self.cfg.get("window_size", 3))
        retry_slack = int(self.cfg.get("retry_slack_frames", 2))
        max_pending = int(self.cfg.get("max_pending_kf", 64))
        max_attempts = int(self.cfg.get("max_retry_attempts_per_kf", 3))
        now = time.perf_counter()
        min_keep = max(0, current_kf_id - window - retry_slack)
        for k in [k for k in self._pending_map if k < min_keep]:
            self._pending_map.pop(k, None)
        if is_final_pass:
            for kf_id in range(max(0, current_kf_id - window + 1), current_kf_id + 1):
                if kf_id in self._pending_map:
                    continue
                if not self._schedule_refinement(kf_id):
                    self._pending_map[kf_id] = {"attempts": 1, "next_ts": now + 0.1,
                                                "final_pass": True}
            return
        if current_kf_id >= window:
            target = current_kf_id - window
            if not self._schedule_refinement(target):
                self._retry_later(target, self._pending_map.get(target, {"attempts": 0,
                                                                         "next_ts": now}),
                                  now, max_attempts)
        due = [(k, v) for k, v in self._pending_map.items()
               if k >= min_keep and v.get("next_ts", 0.0) <= now and not v.get("final_pass", False)]
        due.sort(key=lambda kv: (kv[1].get("attempts", 0), -kv[1].get("best_gain", 0.0)))
        for k, ent in due[:3]:
            if self._schedule_refinement(k):
                self._pending_map.pop(k, None)
            else:
                self._retry_later(k, ent, now, max_attempts)
        if len(self._pending_map) > max_pending:
            for k in sorted(self._pending_map)[: len(self._pending_map) - max_pending]:
                self._pending_map.pop(k, None)

    def _fallback_priority(self, kf_id):
        """:367-396 -- no quality result: low-confidence pixels become the
        priority map (per pixel, patch_size 16 as the reference states)."""
        with self.keyframes.lock:
            if kf_id >= len(self.keyframes):
                return None
            kf = self.keyframes[kf_id]
            H, W = _frame_hw(kf)
            Cm = _conf_hw(kf, H, W)
            low = (Cm < 0.3) & (Cm > 0.05)
            if int(low.sum()) < 100:
                return None
            pr = (0.3 - Cm) * low.float()
            return {"priority": (pr / (pr.max() + 1e-8)).cpu(), "patch_size": 16}

    def _schedule_refinement(self, kf_id: int) -> bool:  # :348-429
        result = None
        if self.quality_service is not None:
            try:
                with self.keyframes.lock:
                    if kf_id < len(self.keyframes):
                        result = self.quality_service.get(self.keyframes[kf_id].frame_id)
                        try:
                            self.quality_service.poll()
                        except Exception:  # noqa: BLE001
                            pass
            except Exception:  # noqa: BLE001
                result = None
        if result is None:
            try:
                result = self._fallback_priority(kf_id)
            except Exception:  # noqa: BLE001
                return False
        if result is None:
            return False
        try:
            blocks = self._select_blocks_enhanced(kf_id, result)
        except Exception:  # noqa: BLE001
            return False
        if not blocks:
            return True
        blocks.sort(key=lambda b: b.priority, reverse=True)
        for block in blocks[: int(self.cfg.get("max_rois_per_kf", 3))]:
            key = BlockKey(block.kf_id, block.block_id)
            if self.registry.try_enqueue(key):
                try:
                    self.queue.put_nowait((key, block))
                except queue.Full:
                    break
        return True

    # ------------------------------------------------------- block selection
    def _select_blocks_enhanced(self, kf_id: int, quality_result: dict) -> List[PatchBlock]:
        """:431-517 -- top-5 % quantile, top-K patches, per-patch depth stats."""
        priority = quality_result.get("priority")
        if priority is None:
            return []
        priority = torch.as_tensor(np.asarray(priority) if isinstance(priority, list) else priority,
                                   dtype=torch.float32).cpu()
        patch_size = int(quality_result.get("patch_size", 16))
        if priority.ndim != 2 or priority.numel() == 0:
            return []
        Gw = priority.shape[1]
        flat = priority.flatten()
        cand = torch.where(flat >= torch.quantile(flat, 0.95))[0]
        K = min(int(self.cfg.get("max_rois_per_kf", 3)), int(cand.numel()))
        if K == 0:
            return []
        vals, rel = torch.topk(flat[cand], k=K)
        top = cand[rel]
        with self.keyframes.lock:
            kf = self.keyframes[min(kf_id, len(self.keyframes) - 1)]
            H, W = _frame_hw(kf)
            Z = kf.X_canon.reshape(H, W, 3)[..., 2]
            Cm = _conf_hw(kf, H, W)
        coords, depths, variances, prios = [], [], [], []
        for v, idx in zip(vals.tolist(), top.tolist()):
            gh, gw = idx // Gw, idx % Gw
            y0, y1 = gh * patch_size, min((gh + 1) * patch_size, H)
            x0, x1 = gw * patch_size, min((gw + 1) * patch_size, W)
            z = Z[y0:y1, x0:x1][Cm[y0:y1, x0:x1] > 0.05]
            if z.numel() > 2:
                coords.append((gh, gw))
                depths.append(float(torch.median(z)))
                variances.append(float(torch.var(z)))
                prios.append(float(v))
        if not coords:
            return []
        return self._cluster_patches_enhanced(kf_id, coords, depths, variances, prios, H, W,
                                              patch_size)

    def _cluster_patches_enhanced(self, kf_id, patch_coords, patch_depths, patch_variances,
                                  patch_priorities, H, W, patch_size):
        """:519-605 -- greedy 8-connected growth from the highest-priority seed,
        at most max_block_edge^2 patches per block (default 1)."""
        z_abs = float(self.cfg.get("z_abs_m", 1e9))
        z_rel = float(self.cfg.get("z_rel", 1e9))
        limit = int(self.cfg.get("max_block_edge", 1)) ** 2
        used = [False] * len(patch_coords)
        blocks = []
        for seed in sorted(range(len(patch_coords)), key=lambda i: patch_priorities[i],
                           reverse=True):
            if used[seed] or not math.isfinite(patch_depths[seed]):
                continue
            cluster, todo = [seed], [seed]
            used[seed] = True
            while todo and len(cluster) < limit:
                cur = todo.pop(0)
                (g0, w0), z1 = patch_coords[cur], patch_depths[cur]
                for j, (g2, w2) in enumerate(patch_coords):
                    if used[j] or not math.isfinite(patch_depths[j]):
                        continue
                    if max(abs(g0 - g2), abs(w0 - w2)) > 1:
                        continue
                    z2 = patch_depths[j]
                    dz = abs(z1 - z2)
                    lo = min(abs(z1), abs(z2))
                    if dz <= z_abs or (dz / lo if lo > 0 else float("inf")) <= z_rel:
                        cluster.append(j)
                        used[j] = True
                        todo.append(j)
            mask = torch.zeros(H, W, dtype=torch.bool, device=self.device)
            for i in cluster:
                gh, gw = patch_coords[i]
                mask[gh * patch_size:min((gh + 1) * patch_size, H),
                     gw * patch_size:min((gw + 1) * patch_size, W)] = True
            zs = [patch_depths[i] for i in cluster if math.isfinite(patch_depths[i])]
            blocks.append(PatchBlock(
                kf_id=kf_id, block_id=len(blocks), patch_indices=[patch_coords[i] for i in cluster],
                pixel_mask=mask.reshape(-1), depth_median=float(np.median(zs)),
                priority=float(np.mean([patch_priorities[i] for i in cluster])),
                depth_variance=float(np.var(zs))))
        return blocks

    # ------------------------------------------------------------- the thread
    def run(self):  # :607-643
        while not self.stop_flag.is_set():
            try:
                key, block = self.queue.get(timeout=0.1)
            except queue.Empty:
                continue
            self.process(key, block)

    def process(self, key, block):
        """One dequeued block (the body of the thread loop)."""
        if not self.registry.begin_run(key):
            return None
        t0 = time.time()
        try:
            success, gain = self._refine_block_enhanced(block)
            self.stats["total_blocks"] += 1
            if success:
                self.stats["successful_blocks"] += 1
        except Exception as e:  # noqa: BLE001  (the reference prints and carries on, :632-636)
            print(f"[TSDF-ERROR] Refinement error for block {key.block_id}: {e}")
            success, gain = False, 0.0
        self.stats["total_processing_time"] += time.time() - t0
        self.registry.finish_run(key, success, gain)
        return success, gain

    # numeric methods: CUDA through the C ABI (reference :837-1064)
    def _build_tsdf_robust(self, X_canon, C_flat, K, xyz_min, xyz_max, H, W, T_WC):
        return local_tsdf.build_tsdf_robust(self.cfg, X_canon, C_flat, K, xyz_min, xyz_max, H, W,
                                            T_WC, device=self.device)

    def _extract_surface_safe(self, tsdf, xyz_min, xyz_max, K, mask, H, W, X_original):
        order = None
        if self.pixel_order is not None:  # tests: the permutation the reference drew
            order = self.pixel_order(int(mask.sum()))
        return local_tsdf.extract_surface_safe(self.cfg, tsdf, xyz_min, xyz_max, K, mask, H, W,
                                               X_original, generator=self.generator, order=order)

    def _refine_block_enhanced(self, block: PatchBlock) -> Tuple[bool, float]:
        """:667-835 -- snapshot under the lock with the keyframe version, ROI
        from the block's valid camera-frame points, local TSDF build + ray cast
        on the GPU, acceptance tests, confidence boost under a version check."""
        with self.keyframes.lock:  # :678-686
            kf = self.keyframes[block.kf_id]
            start_version = int(self.keyframes.version[block.kf_id].item())
            H, W = _frame_hw(kf)
            X = kf.X_canon.clone()
            Cf = kf.C.clone() if kf.C.ndim == 2 else kf.C[..., 0].clone()
            K = kf.K.clone() if kf.K is not None else None
            T_WC = kf.T_WC.clone()
        X = X.to(self.device)
        Cf = Cf.to(self.device) if Cf.ndim == 1 else Cf.squeeze(-1).to(self.device)
        K = K.to(self.device)  # reference quirk (Appendix A #9): K = None raises -> block fails
        mask = block.pixel_mask.to(self.device)
        Xb, Cb = X[mask], Cf[mask]
        valid = (Cb > 0.03) & torch.isfinite(Xb).all(dim=1) & (Xb[:, 2] > 0.03)  # :702
        total_pixels = int(mask.sum())
        if int(valid.sum()) < max(3, total_pixels * 0.05):
            return False, 0.0
        Xv = Xb[valid]
        margin = max(0.01, min(0.05, 0.1 * float(self.cfg.get("roi_size", 0.4))))  # :715-716
        xyz_min, xyz_max = Xv.min(dim=0)[0] - margin, Xv.max(dim=0)[0] + margin
        roi = xyz_max - xyz_min
        if bool(torch.any(roi <= 0)) or bool(torch.any(roi > 15.0)):  # :723
            return False, 0.0
        tsdf, weights = self._build_tsdf_robust(X, Cf, K, xyz_min, xyz_max, H, W, T_WC)
        dbg = self.stats["debug_info"]
        dbg["tsdf_constructions"] += 1
        if int((weights > float(self.cfg.get("min_weight_threshold", 0.01))).sum()) < 3:  # :733-738
            return False, 0.0
        X_refined, hits = self._extract_surface_safe(tsdf, xyz_min, xyz_max, K, mask, H, W, X)
        dbg["surface_extractions"] += 1
        hit_count = int(hits.sum())
        hit_ratio = hit_count / max(1, total_pixels)  # :746
        gain = 0.0
        mask_hits = mask.clone()
        mask_hits[mask] = hits
        if hit_count > 0:
            disp = (X_refined[mask_hits] - X[mask_hits]).norm(dim=1)
            gain = float(disp.mean())
            if float(disp.max()) > float(self.cfg.get("max_displacement", 0.015)):  # :768-772
                dbg["displacement_rejects"] += 1
                return False, hit_ratio
        min_hit = float(self.cfg.get("min_hit_rate", 0.05))
        if not (hit_ratio >= min_hit and hit_count >= 1):  # :775-784
            if hit_ratio < min_hit:
                dbg["hit_ratio_rejects"] += 1
            return False, hit_ratio
        with self.keyframes.lock:  # :787-826
            if int(self.keyframes.version[block.kf_id].item()) != start_version:
                return False, hit_ratio
            kfw = self.keyframes[block.kf_id]
            mh = mask_hits.to(kfw.C.device)
            boost = float(self.cfg.get("confidence_boost", 0.08))
            if kfw.C.ndim == 2:
                kfw.C[mh] += boost
            else:
                kfw.C[mh, 0] += boost
            kfw.C.clamp_(max=float(self.cfg.get("confidence_max", 1.3)))  # whole keyframe (#12)
            gw = float(self.cfg.get("geometric_weight", 0.0))
            if gw > 0:
                blended = (1 - gw) * X + gw * X_refined
                kfw.X_canon[mh] = blended[mask_hits].to(kfw.X_canon.device)
            self.keyframes.version[block.kf_id] += 1
        return True, max(hit_ratio, gain)
