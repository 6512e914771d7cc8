"""CPU: properties of the derived marching-cubes tables / CPU MC oracle
(parity unpinned: the reference has no mesh extraction)."""
import numpy as np

from oracle import oracle as orc


def _sphere_grid(n=28, r=0.31, v=0.03):
    idx = np.stack(np.meshgrid(*[np.arange(-n // 2, n // 2)] * 3, indexing="ij"), -1).reshape(-1, 3)
    c = np.array([0.013, -0.007, 0.004])
    t = (np.linalg.norm(idx * v - c, axis=1) - r) / 0.12
    return idx.astype(np.int64), np.clip(t, -1, 1).astype(np.float32), c, r, v


def test_sphere_mesh_is_watertight_and_oriented():
    keys, t, c, r, v = _sphere_grid()
    tri = orc.marching_cubes(keys, t, np.ones(len(t), np.float32), v)
    assert len(tri) > 500
    # vertices on the sphere
    d = np.linalg.norm(tri.reshape(-1, 3) - c, axis=1)
    assert np.abs(d - r).max() < 0.15 * v
    # every directed edge appears once, its reverse once (closed, consistently oriented)
    e = np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]]).reshape(-1, 6)
    fwd = {tuple(x.tobytes() for x in (row[:3], row[3:])) for row in e}
    assert len(fwd) == len(e)
    for a, b in fwd:
        assert (b, a) in fwd
    # normals point from inside (tsdf < 0) to outside
    nrm = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    cen = tri.mean(axis=1) - c
    assert ((nrm * cen).sum(1) > 0).mean() > 0.999
    area = 0.5 * np.linalg.norm(nrm, axis=1).sum()
    assert abs(area - 4 * np.pi * r * r) / (4 * np.pi * r * r) < 0.02


def test_all_256_cases_close_locally():
    """Random sign fields: the mesh of a 3x3x3-cube lattice never leaves an
    open edge strictly inside the lattice (ambiguous faces are consistent)."""
    rng = np.random.default_rng(0)
    idx = np.stack(np.meshgrid(*[np.arange(4)] * 3, indexing="ij"), -1).reshape(-1, 3).astype(np.int64)
    for _ in range(200):
        t = rng.choice([-0.6, 0.4], size=len(idx)).astype(np.float32)
        tri = orc.marching_cubes(idx, t, np.ones(len(t), np.float32), 1.0)
        if not len(tri):
            continue
        e = np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]]).reshape(-1, 6)
        fwd = {}
        for row in e:
            fwd[(row[:3].tobytes(), row[3:].tobytes())] = row
        for (a, b), row in fwd.items():
            mid = (row[:3] + row[3:]) / 2
            interior = np.all((mid > 1e-6) & (mid < 3 - 1e-6))
            if interior:
                assert (b, a) in fwd


def test_invalid_corner_suppresses_cube():
    keys, t, c, r, v = _sphere_grid(12, 0.1)
    w = np.ones(len(t), np.float32)
    full = len(orc.marching_cubes(keys, t, w, v))
    w[:] = 1e-4  # below min_weight everywhere
    assert len(orc.marching_cubes(keys, t, w, v)) == 0
    assert full > 0
