"""Minimal Sim3 arithmetic on `[t(3), q_xyzw(4), s]` tensors: what the TSDF
path needs from `lietorch.Sim3` (act, exp, composition), which is a missing
submodule of the reference.  Data layout and `act` follow the
reference's own CUDA (mast3r_slam/backend/src/gn_kernels.cu:195-221, layout
:415-450); `exp` is the closed form of the Sim(3) exponential with tangent
order [tau, phi, sigma] (the order the reference Jacobian uses,
mast3r_slam/tsdf/tsdf_optimizer.py:118-124).
"""
import math

import torch


def quat_to_R(q):
    x, y, z, w = q.unbind(-1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
        2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
        2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], -1).reshape(
            q.shape[:-1] + (3, 3))


def quat_mul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([aw * bx + ax * bw + ay * bz - az * by,
                        aw * by - ax * bz + ay * bw + az * bx,
                        aw * bz + ax * by - ay * bx + az * bw,
                        aw * bw - ax * bx - ay * by - az * bz], -1)


def act(pose, pts):
    """s * R(q) p + t for pose (8,) and pts (M,3)."""
    pose = pose.reshape(8)
    return pose[7] * (pts @ quat_to_R(pose[3:7]).T) + pose[:3]


def act_kernel_order(pose, pts):
    """`act` in the float32 operation order of the CUDA kernels (csrc/sim3.cuh,
    the reference's actSO3 / actSim3): x + 2 w (q x x) + 2 q x (q x x), then
    scale, then translate.  Bit-identical to the device path on CPU tensors."""
    P = pose.reshape(8).to(torch.float32).cpu()
    x = pts.to(torch.float32).cpu()
    q, w, s, t = P[3:6], P[6], P[7], P[:3]
    uv = 2.0 * torch.stack([q[1] * x[:, 2] - q[2] * x[:, 1], q[2] * x[:, 0] - q[0] * x[:, 2],
                            q[0] * x[:, 1] - q[1] * x[:, 0]], dim=1)
    cr = torch.stack([q[1] * uv[:, 2] - q[2] * uv[:, 1], q[2] * uv[:, 0] - q[0] * uv[:, 2],
                      q[0] * uv[:, 1] - q[1] * uv[:, 0]], dim=1)
    a = (x + w * uv) + cr
    return (a * s + t).to(pts.device)


def exp(delta):
    """Sim3 exponential of a tangent vector (7,) -> pose (8,) (float64)."""
    d = delta.reshape(7).to(torch.float64)
    tau, phi, sigma = d[:3], d[3:6], float(d[6])
    theta = float(phi.norm())
    s = math.exp(sigma)
    eps = 1.0e-9
    if abs(sigma) < eps:
        C = 1.0
        if theta < eps:
            A, B = 0.5, 1.0 / 6.0
        else:
            A = (1 - math.cos(theta)) / theta ** 2
            B = (theta - math.sin(theta)) / theta ** 3
    else:
        C = (s - 1) / sigma
        if theta < eps:
            A = ((sigma - 1) * s + 1) / sigma ** 2
            B = (s * 0.5 * sigma ** 2 + s - 1 - sigma * s) / sigma ** 3
        else:
            a, b, c = s * math.sin(theta), s * math.cos(theta), theta ** 2 + sigma ** 2
            A = (a * sigma + (1 - b) * theta) / (theta * c)
            B = (C - ((b - 1) * sigma + a * theta) / c) / theta ** 2
    O = torch.zeros(3, 3, dtype=torch.float64)
    O[0, 1], O[0, 2], O[1, 0], O[1, 2], O[2, 0], O[2, 1] = -phi[2], phi[1], phi[2], -phi[0], \
        -phi[1], phi[0]
    Wm = A * O + B * (O @ O) + C * torch.eye(3, dtype=torch.float64)
    t = Wm @ tau
    if theta < eps:
        q = torch.cat([0.5 * phi, torch.ones(1, dtype=torch.float64)])
    else:
        q = torch.cat([math.sin(0.5 * theta) / theta * phi,
                       torch.tensor([math.cos(0.5 * theta)], dtype=torch.float64)])
    q = q / q.norm()
    return torch.cat([t, q, torch.tensor([s], dtype=torch.float64)])


def mul(a, b):
    """Composition a * b: x -> a(b(x))."""
    a, b = a.reshape(8).to(torch.float64), b.reshape(8).to(torch.float64)
    t = a[7] * (quat_to_R(a[3:7]) @ b[:3]) + a[:3]
    q = quat_mul(a[3:7], b[3:7])
    return torch.cat([t, q / q.norm(), (a[7] * b[7]).reshape(1)])


def pose_data(T_WC):
    """(8,) tensor from a lietorch-like object (`.data` (1,8)) or a tensor."""
    data = T_WC.data if hasattr(T_WC, "data") and not isinstance(T_WC, torch.Tensor) else T_WC
    return torch.as_tensor(data).detach().reshape(-1)[:8]
