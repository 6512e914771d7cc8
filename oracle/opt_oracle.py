"""CPU restatement of the reference TSDFPoseOptimizer._optimize_single
(mast3r_slam/tsdf/tsdf_optimizer.py:46-124) -- TEST INFRASTRUCTURE, NOT PRODUCT.

Pinned by tests/golden/opt_sphere.npz (frozen from the unmodified reference
with the lietorch stub of oracle/ref_harness.py; parity unpinned at the
lietorch boundary only).  `volume` is an oracle.OracleVolume (or anything with
the reference `query`); the sample choice is an input (the reference draws it
with an unseeded torch.randperm, :70).
"""
import math

import numpy as np

from .ref_harness import _quat_R, sim3_exp_np, sim3_mul_np


def _act_f32(pose, pts):
    """pose.act(pts) in float32, like lietorch on float32 tensors (:76)."""
    p = pose.astype(np.float32)
    R = _quat_R(p[3:7].astype(np.float64)).astype(np.float32)
    return (p[7] * (pts.astype(np.float32) @ R.T) + p[:3]).astype(np.float32)


def sim3_jacobian(p, g):  # :118-124
    J = np.zeros(7)
    J[:3] = g
    J[3:6] = -np.cross(p, g)
    J[6] = float(np.dot(p, g))
    return J


def normal_equations(volume, p_world, conf, lam):  # :94-116
    H, b, used = np.zeros((7, 7)), np.zeros(7), 0
    for p, w in zip(p_world, conf):
        t, g = volume.query(p)
        if t is None or g is None:
            continue
        used += 1
        J = sim3_jacobian(p.astype(np.float64), g)
        if not np.isfinite(t) or not np.all(np.isfinite(J)):
            continue
        sw = math.sqrt(max(lam * float(w), 1.0e-6))
        Jw = sw * J
        H += np.outer(Jw, Jw)
        b += Jw * t * sw
    return H, b, used


def optimize_single(volume, X, C, pose, choice, iterations=3, lam=0.1, damping=1.0e-4):
    pts_cam = X.reshape(-1, 3)[choice].astype(np.float32)
    conf = C.reshape(-1)[choice]
    pose = np.asarray(pose, np.float32).astype(np.float32)
    for _ in range(iterations):
        pw = _act_f32(pose, pts_cam)
        H, b, used = normal_equations(volume, pw, conf, lam)
        if used < 6:  # :78-79
            break
        try:
            delta = np.linalg.solve(H + damping * np.eye(7), -b)  # :82
        except np.linalg.LinAlgError:
            break
        d32 = delta.astype(np.float32).astype(np.float64)  # :84 delta cast to the point dtype
        pose = sim3_mul_np(sim3_exp_np(d32), pose.astype(np.float64)).astype(np.float32)
    return pose.astype(np.float64)
