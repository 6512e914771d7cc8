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
