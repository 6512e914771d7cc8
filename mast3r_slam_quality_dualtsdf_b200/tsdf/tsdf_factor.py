"""Unary TSDF factors for the backend's Gauss-Newton system.

The reference ships the pose refinement as a separate optimiser
(mast3r_slam/tsdf/tsdf_optimizer.py:46-124) and only SKETCHES the factor-graph
path in its roadmap (MASt3R_SLAM_TSDF_Integration.md:240-252, section 9.3.2):
residual r_i = TSDF(T_WC_k X_i), Jacobian J_i = grad TSDF(p_i) [I, -[p_i]x, p_i],
weight lambda * C_i, and `H = H_ray + lambda H_tsdf` inside `solve_GN_rays /
solve_GN_calib` (global_opt.py:123-223).  This class is that interface: per
keyframe it returns the 7x7 block and the 7-vector the factor adds to the
system -- computed by the same fused launch the optimiser uses
(`b200tsdf_pose_normal_equations`: Sim3 act + query + Jacobian + float64
reduction) -- and scatters them into a dense (7 n x 7 n) system in the
variable order of `FactorGraph.get_unique_kf_idx()[pin:]`.  The backend's own
Gauss-Newton CUDA (mast3r_slam/backend) stays the reference's; nothing here
links against it.
"""
from typing import Dict, Iterable, Optional, Sequence, Tuple

import numpy as np
import torch

from . import sim3


class TSDFFactor:
    """r = TSDF(T_WC X) for sampled points of a keyframe, as a unary Sim3 factor."""

    def __init__(self, volume, keyframes, cfg, device) -> None:
        self.volume = volume
        self.keyframes = keyframes
        self.device = device if isinstance(device, torch.device) else torch.device(device)
        self.samples_per_kf = int(cfg.get("samples_per_kf", 2000))   # tsdf_optimizer.py:21
        self.min_conf = float(cfg.get("min_confidence", 0.05))
        self.lambda_tsdf = float(cfg.get("lambda", 0.1))
        self.which = int(cfg.get("optimize_against", 0))
        self.seed = int(cfg.get("seed", 0))

    def sample(self, idx: int, choice=None) -> Optional[Tuple[torch.Tensor, torch.Tensor, np.ndarray]]:
        """Points of keyframe idx the factor is built on: `samples_per_kf` of the
        pixels with C > min_confidence (tsdf_optimizer.py:61-71), seeded per
        keyframe.  Returns (pts_cam f32 (n,3), conf f64 (n,), pose (8,)) or None."""
        with self.keyframes.lock:
            if idx >= len(self.keyframes):
                return None
            frame = self.keyframes[idx]
            X = frame.X_canon.detach().clone()
            Cf = frame.C.detach().clone()
            pose = sim3.pose_data(frame.T_WC).clone()
        points, conf = X.reshape(-1, 3), Cf.reshape(-1)
        valid = torch.nonzero(conf > self.min_conf).reshape(-1)
        if valid.numel() == 0:
            return None
        if choice is None:
            g = torch.Generator(device="cpu").manual_seed(self.seed * 1000003 + int(idx))
            n = min(self.samples_per_kf, int(valid.numel()))
            choice = valid.cpu()[torch.randperm(int(valid.numel()), generator=g)[:n]]
        choice = torch.as_tensor(choice).to(points.device)
        return (points[choice].to(self.device, torch.float32),
                conf[choice].to(self.device, torch.float64),
                pose.to(torch.float64).cpu().numpy().reshape(8))

    def linearize(self, indices: Iterable[int], poses: Optional[Dict[int, Sequence[float]]] = None,
                  choices: Optional[Dict[int, Sequence[int]]] = None):
        """{idx: (H (7,7), b (7,), used)} at the keyframes' current poses (or `poses[idx]`,
        the backend's working estimate).  One launch per keyframe; a keyframe with fewer
        than 6 usable points contributes nothing (tsdf_optimizer.py:78-79)."""
        out = {}
        for idx in indices:
            s = self.sample(int(idx), None if choices is None else choices.get(int(idx)))
            if s is None:
                continue
            pts, conf, pose = s
            if poses is not None and int(idx) in poses:
                pose = np.asarray(poses[int(idx)], np.float64).reshape(8)
            H, b, used = self.volume.pose_normal_equations(pts, conf, pose.astype(np.float32),
                                                           self.lambda_tsdf, self.which)
            if used >= 6:
                out[int(idx)] = (H, b, used)
        return out

    @staticmethod
    def add_to(H_full: np.ndarray, b_full: np.ndarray, blocks, order: Sequence[int]) -> int:
        """H_full (7n,7n) += H_k, b_full (7n,) += b_k on the diagonal block of every
        keyframe in `order` (the system's variable order: unique_kf_idx[pin:]); pinned
        or absent keyframes are skipped.  Returns the number of factors added."""
        pos = {int(k): n for n, k in enumerate(order)}
        added = 0
        for idx, (H, b, _used) in blocks.items():
            n = pos.get(int(idx))
            if n is None:
                continue
            H_full[7 * n:7 * n + 7, 7 * n:7 * n + 7] += H
            b_full[7 * n:7 * n + 7] += b
            added += 1
        return added
