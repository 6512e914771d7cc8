"""GPU: marching-cubes kernel vs the CPU MC oracle (parity unpinned: the
reference has no mesh path).  Triangle counts bit-exact; triangle soups equal
as sets (same float32 operation order => identical vertices)."""
import numpy as np
import pytest
import torch

from mast3r_slam_quality_dualtsdf_b200 import synth
from oracle import oracle as orc
from tests.util import oracle_dual

pytestmark = pytest.mark.gpu


def _canon(tri):
    t = np.ascontiguousarray(tri.reshape(-1, 9))
    return t[np.lexsort(t.T[::-1])]


@pytest.mark.parametrize("scene,voxel,n_kf", [("sphere", 0.03, 3), ("plane", 0.02, 2),
                                              ("room", 0.02, 2)])
def test_mesh_matches_cpu_marching_cubes(cuda_device, scene, voxel, n_kf):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    trunc = 4 * voxel
    kf = synth.make_keyframes(synth.SCENES[scene], n_kf, H=48, W=64, seed=4, f=50.0, total=12)
    vol = TSDFVolume(voxel, trunc, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            torch.from_numpy(kf["q"]).to(cuda_device), 0.05, 48, 64)
    for which in (0, 1):
        k, t, w, c = vol.export(which)
        want = orc.marching_cubes(k, t, w, voxel, vol.min_weight)
        got = vol.extract_mesh(which).cpu().numpy().reshape(-1, 3, 3)
        assert len(got) == len(want) and len(want) > 100
        assert np.array_equal(_canon(got), _canon(want))
    # against the mesh of the ORACLE volume (float64 fold): same surface up to
    # sign flips of |tsdf| < 1e-6 voxels
    vp, vq, _ = oracle_dual(kf, voxel, trunc)
    k, t, w, c = vp.export()
    ref = orc.marching_cubes(k, t.astype(np.float32), w.astype(np.float32), voxel, vol.min_weight)
    got = vol.extract_mesh(0)
    assert abs(len(ref) - got.shape[0] // 3) <= max(2, len(ref) // 500)


def test_mesh_empty_volume(cuda_device):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    vol = TSDFVolume(0.03, 0.12, device=cuda_device)
    assert vol.extract_mesh().shape == (0, 3)
