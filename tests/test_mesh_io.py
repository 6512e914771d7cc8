"""CPU: PLY mesh writer round trip on a CPU marching-cubes mesh (oracle)."""
import numpy as np

from mast3r_slam_quality_dualtsdf_b200 import mesh_io
from tests.test_mc_oracle import _sphere_grid
from oracle import oracle as orc


def test_ply_round_trip_and_weld(tmp_path):
    keys, t, _c, _r, v = _sphere_grid()
    tri = orc.marching_cubes(keys, t, np.ones(len(t), np.float32), v)
    assert len(tri) > 100
    p = tmp_path / "m.ply"
    nv, nf = mesh_io.save_mesh_ply(str(p), tri)
    assert nf == len(tri) and nv < 3 * nf  # shared vertices are stored once
    # closed surface: Euler characteristic 2 (V - E + F with E = 3F/2)
    assert nv - 3 * nf // 2 + nf == 2
    verts, faces = mesh_io.load_mesh_ply(str(p))
    assert verts.shape == (nv, 3) and faces.shape == (nf, 3)
    assert np.array_equal(verts[faces].reshape(-1, 3, 3), np.asarray(tri, np.float32).reshape(-1, 3, 3))
    nv2, nf2 = mesh_io.save_mesh_ply(str(p), tri, weld_vertices=False)
    assert (nv2, nf2) == (3 * nf, nf)
    e, f = mesh_io.weld(np.zeros((0, 3, 3), np.float32))
    assert e.shape == (0, 3) and f.shape == (0, 3)
