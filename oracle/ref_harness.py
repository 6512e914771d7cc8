"""Load the UNMODIFIED reference TSDF code from /root/reference (build container
only; the GPU box has no /root/reference).  TEST INFRASTRUCTURE: used by
tests/golden/make_golden.py to freeze golden vectors and by the not-gpu tests
that re-validate the C oracle when the reference tree is present.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("TSDF_REFERENCE_ROOT", "/root/reference")


def available():
    return os.path.exists(os.path.join(REF_ROOT, "mast3r_slam", "tsdf", "global_volume.py"))


def load_global_volume():
    """reference mast3r_slam/tsdf/global_volume.py by file path (skips the
    package __init__, which would import lietorch)."""
    path = os.path.join(REF_ROOT, "mast3r_slam", "tsdf", "global_volume.py")
    spec = importlib.util.spec_from_file_location("_ref_global_volume", path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules["_ref_global_volume"] = mod  # dataclass needs the module registered
    spec.loader.exec_module(mod)
    return mod


class CountingVolume:
    """Reference TSDFVolume plus a per-voxel hit counter (wraps _update_voxel
    without changing its arithmetic)."""

    def __init__(self, *a, **k):
        mod = load_global_volume()
        self.vol = mod.TSDFVolume(*a, **k)
        self.counts = {}
        self.events = 0
        inner = self.vol._update_voxel

        def wrapped(point_world, tsdf, weight):
            if weight > 0.0:
                key = self.vol._world_to_voxel(point_world)
                self.counts[key] = self.counts.get(key, 0) + 1
                self.events += 1
            return inner(point_world, tsdf, weight)

        self.vol._update_voxel = wrapped

    def integrate(self, *a, **k):
        return self.vol.integrate(*a, **k)

    def export(self):
        keys = sorted(self.vol._voxels.keys())
        k = np.array(keys, np.int64).reshape(-1, 3)
        t = np.array([float(self.vol._voxels[q].tsdf) for q in keys], np.float64)
        w = np.array([float(self.vol._voxels[q].weight) for q in keys], np.float64)
        c = np.array([self.counts[q] for q in keys], np.uint32)
        return k, t, w, c


def _quat_R(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _R_quat(R):
    from scipy.spatial.transform import Rotation

    q = Rotation.from_matrix(R).as_quat()  # xyzw
    return q if q[3] >= 0 else -q


def sim3_exp_np(d):
    """exp of a Sim(3) tangent [tau, phi, sigma] as [t, q_xyzw, s] (float64)."""
    from scipy.linalg import expm

    phi = d[3:6]
    O = np.array([[0, -phi[2], phi[1]], [phi[2], 0, -phi[0]], [-phi[1], phi[0], 0]])
    M = np.zeros((4, 4))
    M[:3, :3] = O + d[6] * np.eye(3)
    M[:3, 3] = d[:3]
    E = expm(M)
    s = np.cbrt(np.linalg.det(E[:3, :3]))
    return np.concatenate([E[:3, 3], _R_quat(E[:3, :3] / s), [s]])


def sim3_mul_np(a, b):
    Ra, Rb = _quat_R(a[3:7]), _quat_R(b[3:7])
    t = a[7] * (Ra @ b[:3]) + a[:3]
    return np.concatenate([t, _R_quat(Ra @ Rb), [a[7] * b[7]]])


def _sim3_stub():
    """~30-line stand-in for the absent lietorch; only what
    tsdf_refine.py touches: Sim3(data).act/.matrix/.clone/.to/.data."""
    import torch

    class Sim3:
        embedded_dim = 8

        def __init__(self, data):
            self.data = data

        @staticmethod
        def Identity(n, device="cpu", dtype=None):
            d = torch.zeros(n, 8, device=device, dtype=dtype or torch.float32)
            d[:, 6] = 1
            d[:, 7] = 1
            return Sim3(d)

        def clone(self):
            return Sim3(self.data.clone())

        def to(self, *a, **k):
            return Sim3(self.data.to(*a, **k))

        def _parts(self):
            d = self.data.reshape(-1, 8)
            return d[:, :3], d[:, 3:7], d[:, 7:8]

        def act(self, X):
            t, q, s = self._parts()
            qv, qw = q[:, :3], q[:, 3:4]
            uv = 2.0 * torch.cross(qv.expand_as(X), X, dim=-1)
            Y = (X + qw * uv) + torch.cross(qv.expand_as(X), uv, dim=-1)
            return Y * s + t

        def matrix(self):
            t, q, s = self._parts()
            eye = torch.eye(3, dtype=self.data.dtype)
            R = torch.stack([self.act(e.reshape(1, 3))[0] - t[0] for e in eye], dim=1)
            M = torch.eye(4, dtype=self.data.dtype)
            M[:3, :3] = R
            M[:3, 3] = t[0]
            return M.unsqueeze(0)

        # exp / composition for tsdf_optimizer.py:85-86, through the 4x4 matrix
        # exponential (scipy) -- deliberately independent of the product's sim3.py
        @staticmethod
        def exp(delta):
            d = delta.detach().cpu().double().numpy().reshape(7)
            return Sim3(torch.from_numpy(sim3_exp_np(d)).to(delta.dtype).reshape(1, 8).to(
                delta.device))

        def __mul__(self, other):
            a = self.data.detach().cpu().double().numpy().reshape(8)
            b = other.data.detach().cpu().double().numpy().reshape(8)
            return Sim3(torch.from_numpy(sim3_mul_np(a, b)).to(self.data.dtype).reshape(1, 8).to(
                self.data.device))

    class SE3(Sim3):
        embedded_dim = 7

    m = types.ModuleType("lietorch")
    m.Sim3 = Sim3
    m.SE3 = SE3
    return m


def load_tsdf_refine():
    """reference mast3r_slam/tsdf_refine.py with the lietorch stub."""
    sys.modules["lietorch"] = _sim3_stub()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    # do not let a same-named product package shadow the reference
    for k in [k for k in sys.modules if k == "mast3r_slam" or k.startswith("mast3r_slam.")]:
        del sys.modules[k]
    import importlib

    mod = importlib.import_module("mast3r_slam.tsdf_refine")
    return mod, sys.modules["lietorch"]


def load_tsdf_optimizer():
    """reference mast3r_slam/tsdf/tsdf_optimizer.py by file path (it imports
    only numpy, torch and lietorch -> the stub)."""
    sys.modules["lietorch"] = _sim3_stub()
    path = os.path.join(REF_ROOT, "mast3r_slam", "tsdf", "tsdf_optimizer.py")
    spec = importlib.util.spec_from_file_location("_ref_tsdf_optimizer", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod, sys.modules["lietorch"]


class FakeKeyframes:
    """Duck-type of SharedKeyframes (frame.py:220-334) for the TSDF path."""

    class _Frame:
        pass

    def __init__(self, X, C, poses, lietorch_mod=None, img_shape=None, K=None):
        import threading

        import torch

        self.lock = threading.RLock()
        self.X = [torch.as_tensor(x) for x in X]
        self.C = [torch.as_tensor(c).reshape(-1, 1) for c in C]
        self.T_WC = [torch.as_tensor(p).reshape(1, 8).clone() for p in poses]
        self.lt = lietorch_mod
        self.img_shape = img_shape
        self.K = K
        self.version = torch.zeros(len(self.X), dtype=torch.long)
        self.device = "cpu"

    def __len__(self):
        return len(self.X)

    def __getitem__(self, idx):
        import torch

        f = self._Frame()
        f.X_canon, f.C, f.frame_id = self.X[idx], self.C[idx], idx
        f.T_WC = self.lt.Sim3(self.T_WC[idx]) if self.lt is not None else self.T_WC[idx]
        f.K = self.K
        if self.img_shape is not None:
            f.img_shape = torch.tensor([list(self.img_shape)])
        return f
