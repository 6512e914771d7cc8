"""B200 mirror of the reference `mast3r_slam/tsdf/tsdf_optimizer.py`.

Same class, constructor, `pre_refine` / `optimize_keyframes` and config keys
(tsdf_optimizer.py:12-43).  What changes: the per-point Python loop over
`volume.query` (7 dict lookups per point, :94-105) and the 7x7 normal
equations (:107-116) are ONE `b200tsdf_pose_normal_equations` launch per
Gauss-Newton iteration (Sim3 act + query + Jacobian + float64 reduction in the
kernel); only H (7x7) and b (7) cross to the host for the solve (:82).  The Sim3 retraction `exp(delta) * pose` (:85-86) uses tsdf/sim3.py
instead of the missing lietorch submodule.
"""
from typing import Sequence

import numpy as np
import torch

from . import sim3


class TSDFPoseOptimizer:
    """Pose refinement using TSDF residuals."""

    def __init__(self, volume, keyframes, cfg, use_calib: bool, device) -> None:
        self.volume = volume
        self.keyframes = keyframes
        self.cfg = cfg
        self.use_calib = use_calib
        self.device = device if isinstance(device, torch.device) else torch.device(device)
        self.samples_per_kf = int(cfg.get("samples_per_kf", 2000))  # :21
        self.min_conf = float(cfg.get("min_confidence", 0.05))
        self.max_iterations = int(cfg.get("max_iterations", 3))
        self.lambda_tsdf = float(cfg.get("lambda", 0.1))
        self.damping = float(cfg.get("damping", 1.0e-4))
        self.pre_icp_iters = int(cfg.get("pre_icp_iters", 0))
        self.which = int(cfg.get("optimize_against", 0))  # 0 primary, 1 quality-weighted volume
        # False: linearise on the device, solve with numpy on the host (one sync per iteration)
        self.device_solve = bool(cfg.get("device_solve", True))
        self.last_info = {}

    # ------------------------------------------------------------------
    def pre_refine(self, kf_idx: int) -> None:  # :29-37
        if self.pre_icp_iters <= 0:
            return
        self._optimize_single(kf_idx, iterations=self.pre_icp_iters,
                              sample_override=min(self.samples_per_kf // 2, 1000),
                              log_prefix="[TSDF-ICP]")

    def optimize_keyframes(self, indices: Sequence[int], context: str = "factor") -> None:
        if not indices:  # :40-41
            return
        for idx in indices:
            self._optimize_single(idx, iterations=self.max_iterations,
                                  log_prefix=f"[TSDF-{context}]")

    # ------------------------------------------------------------------
    def _optimize_single(self, idx: int, iterations: int, sample_override: int = 0,
                         log_prefix: str = "[TSDF]", choice=None) -> None:
        with self.keyframes.lock:  # :53-59
            if idx >= len(self.keyframes):
                return
            frame = self.keyframes[idx]
            X = frame.X_canon.detach().clone()
            Cf = frame.C.detach().clone()
            pose0 = sim3.pose_data(frame.T_WC).clone()
        points = X.reshape(-1, 3)
        conf = Cf.reshape(-1)
        valid_idx = torch.nonzero(conf > self.min_conf).reshape(-1)
        if valid_idx.numel() == 0:
            return
        max_samples = sample_override if sample_override > 0 else self.samples_per_kf
        sample_count = min(max_samples, int(valid_idx.numel()))
        if choice is None:  # the reference draws an unseeded CPU permutation (:70)
            choice = valid_idx.cpu()[torch.randperm(int(valid_idx.numel()))[:sample_count]]
        choice = torch.as_tensor(choice).to(points.device)
        pts_cam = points[choice].to(self.device, torch.float32)
        conf_sel = conf[choice].to(self.device, torch.float64)
        pose = pose0.to(torch.float64).cpu()
        used = 0
        on_device = hasattr(self.volume, "pose_optimize") and self.device_solve
        if on_device:
            # :72-92 as `iterations` launches and ONE synchronisation: the 7x7 solve and the
            # Sim3 retraction run in the last block of every linearisation launch
            new_pose, used, _ = self.volume.pose_optimize(
                pts_cam, conf_sel, pose.numpy(), self.lambda_tsdf, self.damping, iterations,
                self.which)
            pose = torch.from_numpy(new_pose)
        for _ in range(0 if on_device else iterations):
            # :76 + :94-116 fused: float32 Sim3 act, query, Jacobian, 7x7 reduction
            H, b, used = self.volume.pose_normal_equations(
                pts_cam, conf_sel, pose.to(torch.float32).numpy(), self.lambda_tsdf, self.which)
            if used < 6:  # :78-79
                break
            try:
                delta = np.linalg.solve(H + self.damping * np.eye(7), -b)  # :82
            except np.linalg.LinAlgError:
                break
            # :84-86 the reference casts delta to the point dtype (float32) before exp
            d32 = torch.from_numpy(delta).to(torch.float32).to(torch.float64)
            pose = sim3.mul(sim3.exp(d32), pose)
        self.last_info = {"idx": idx, "used": used, "prefix": log_prefix}
        with self.keyframes.lock:  # :89-92
            if idx < len(self.keyframes):
                tgt = self.keyframes.T_WC[idx]
                tgt_data = tgt.data if hasattr(tgt, "data") else tgt
                self.keyframes.T_WC[idx] = pose.to(tgt_data.device, tgt_data.dtype).reshape(
                    tgt_data.shape)

    def _normal_equations(self, p_world, conf_sel):
        """Device-tensor statement of the same linearisation from world points
        (`b200tsdf_query` + torch reductions): kept as the cross-check of the
        fused kernel (tests) and for callers that hold world points.
        :94-116 batched: residual r = tsdf(p), J = [g, -p x g, p . g],
        weight lambda * C, H = sum (sqrt(w) J)(sqrt(w) J)^T, b = sum sqrt(w) J r sqrt(w)."""
        t, g, fl = self.volume.query_batch(p_world, self.which)
        ok = fl == 3  # value and gradient present (:99-100)
        p = p_world.to(torch.float64)
        J = torch.cat([g, -torch.linalg.cross(p, g), (p * g).sum(-1, keepdim=True)], dim=1)
        ok = ok & torch.isfinite(t) & torch.isfinite(J).all(dim=1)  # :111-112
        w = torch.clamp(self.lambda_tsdf * conf_sel, min=1.0e-6)  # sqrt(max(w, 1e-6))^2
        w = torch.where(ok, w, torch.zeros_like(w))
        J = torch.where(ok[:, None], J, torch.zeros_like(J))
        r = torch.where(ok, t, torch.zeros_like(t))
        H = (J * w[:, None]).T @ J
        b = (J * (w * r)[:, None]).sum(0)
        return H.cpu().numpy(), b.cpu().numpy(), int(ok.sum().item())
