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
        # the async entry leaves the count on the device and never synchronises
        buf = torch.empty(len(want) * 3 + 30, 3, dtype=torch.float32, device=cuda_device)
        cnt = vol.extract_mesh_async(which, buf)
        small = torch.empty(30, 3, dtype=torch.float32, device=cuda_device)
        cnt_small = vol.extract_mesh_async(which, small)  # capacity 10 triangles: count still exact
        assert int(cnt.item()) == len(want) and int(cnt_small.item()) == len(want)
        assert np.array_equal(_canon(buf[: len(want) * 3].cpu().numpy().reshape(-1, 3, 3)), _canon(want))
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


def test_mesh_ply_export(cuda_device, tmp_path):
    """extract_mesh -> welded binary PLY: shared edge vertices are bit-identical
    on the device, so the weld is exact and every face survives the round trip."""
    from mast3r_slam_quality_dualtsdf_b200 import mesh_io
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kf = synth.make_keyframes(synth.SCENES["room"], 3, H=48, W=64, seed=9, f=50.0, total=12)
    vol = TSDFVolume(0.02, 0.08, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            None, 0.05, 48, 64)
    soup = vol.extract_mesh(0)
    nv, nf = mesh_io.save_mesh_ply(str(tmp_path / "m.ply"), soup)
    assert nf == soup.shape[0] // 3 and nf > 100 and nv < 2 * nf
    verts, faces = mesh_io.load_mesh_ply(str(tmp_path / "m.ply"))
    assert np.array_equal(verts[faces].reshape(-1, 3), soup.cpu().numpy())


def _same_soup(got, want, tol=1.0e-5):
    """Triangle soups equal as sets up to float32 rounding of the voxel sums
    (reductions land in a different order in every run)."""
    from scipy.spatial import cKDTree

    if len(got) != len(want):
        return False
    d, idx = cKDTree(want.reshape(-1, 9)).query(got.reshape(-1, 9))
    return bool(d.max(initial=0.0) <= tol and len(np.unique(idx)) == len(want))


@pytest.mark.parametrize("world,width", [(2, 8), (4, 8), (2, 16), (3, 8)])
def test_mesh_stitched_at_slab_seams(cuda_device, world, width):
    """`world` ranks emulated by `world` handles on one GPU.  Each
    fuses its interleaved z-slabs, seam blocks go to the rank below as halo
    blocks, and the union of the per-rank meshes is the single-volume mesh,
    triangle for triangle (none duplicated, none missing)."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    voxel, trunc = 0.03, 0.12
    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=48, W=64, seed=4, f=50.0, total=12)
    args = [torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")]
    q = torch.from_numpy(kf["q"]).to(cuda_device)
    single = TSDFVolume(voxel, trunc, device=cuda_device)
    single.integrate_keyframes(*args, q, 0.05, 48, 64)
    ranks = [TSDFVolume(voxel, trunc, device=cuda_device, slabs=(world, r, width))
             for r in range(world)]
    for v in ranks:
        v.integrate_keyframes(*args, q, 0.05, 48, 64)
    seams = [v.export_seam_blocks() for v in ranks]
    assert sum(s[0].shape[0] for s in seams) > 0
    for r, v in enumerate(ranks):
        v.import_halo_blocks(*seams[(r + 1) % world])
    for which in (0, 1):
        want = single.extract_mesh(which).cpu().numpy().reshape(-1, 3, 3)
        parts = [v.extract_mesh(which).cpu().numpy().reshape(-1, 3, 3) for v in ranks]
        got = np.concatenate(parts)
        assert len(want) > 100 and len(got) == len(want)
        assert _same_soup(got, want)
    # halo blocks are invisible to stats / export / block list
    tot = sum(v.stats(0)["total_voxels"] for v in ranks)
    assert tot == single.stats(0)["total_voxels"]
    assert sum(len(v.block_keys()) for v in ranks) == len(single.block_keys())
    # re-import after more fusion overwrites the halo copy (no double counting)
    for v in ranks:
        v.integrate_keyframes(*args, q, 0.05, 48, 64)
    single.integrate_keyframes(*args, q, 0.05, 48, 64)
    seams = [v.export_seam_blocks() for v in ranks]
    for r, v in enumerate(ranks):
        v.import_halo_blocks(*seams[(r + 1) % world])
    want = single.extract_mesh(0).cpu().numpy().reshape(-1, 3, 3)
    got = np.concatenate([v.extract_mesh(0).cpu().numpy().reshape(-1, 3, 3) for v in ranks])
    assert _same_soup(got, want)


def test_indexed_mesh_equals_host_weld_of_the_soup(cuda_device, tmp_path):
    """SURVEY 8(b) / VERDICT r1 #3: verts + int32 faces produced on the device
    (csrc/weld.cu).  The indexed mesh reconstructs the soup bit for bit, has
    exactly as many vertices as the host weld (np.unique on the 96-bit
    positions, mesh_io.weld) and the same vertex set; it is watertight-manifold
    where the soup is; and the PLY writer stores it unchanged."""
    from mast3r_slam_quality_dualtsdf_b200 import mesh_io
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume
    from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import weld_mesh

    kf = synth.make_keyframes(synth.SCENES["sphere"], 4, H=48, W=64, seed=4, f=50.0, total=12)
    vol = TSDFVolume(0.02, 0.08, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            torch.from_numpy(kf["q"]).to(cuda_device), 0.05, 48, 64)
    for which in (0, 1):
        soup = vol.extract_mesh(which)
        verts, faces = vol.extract_mesh_indexed(which)
        s = soup.cpu().numpy().reshape(-1, 3, 3)
        v, f = verts.cpu().numpy(), faces.cpu().numpy()
        assert f.shape == (len(s), 3) and f.dtype == np.int32 and len(s) > 1000
        assert f.min() >= 0 and f.max() == len(v) - 1
        hv, hf = mesh_io.weld(s)
        assert len(v) == len(hv) < 3 * len(s)
        # same triangle multiset (the two MC passes may order triangles differently)
        assert np.array_equal(_canon(v[f]), _canon(s))
        key = lambda a: np.sort(np.ascontiguousarray(a).view(np.dtype((np.void, 12))).reshape(-1))
        assert np.array_equal(key(v), key(hv))
        # the weld alone on the soup the caller holds: reconstructs it in order
        wv, wf = weld_mesh(soup)
        assert np.array_equal(wv.cpu().numpy()[wf.cpu().numpy()], s)
        first = wf.cpu().numpy().reshape(-1)
        # vertex order = first occurrence in the soup
        _, idx = np.unique(first, return_index=True)
        assert np.array_equal(first[np.sort(idx)], np.arange(len(wv)))
    path = str(tmp_path / "m.ply")
    nv, nf = mesh_io.save_mesh_ply(path, verts, faces=faces)
    rv, rf = mesh_io.load_mesh_ply(path)
    assert (nv, nf) == (len(v), len(f)) and np.array_equal(rv, v) and np.array_equal(rf, f)
    # empty volume: empty indexed mesh
    ev, ef = TSDFVolume(0.02, 0.08, device=cuda_device).extract_mesh_indexed(0)
    assert ev.shape == (0, 3) and ef.shape == (0, 3)


@pytest.mark.parametrize("generation", ["1", "2"])
def test_older_kernel_generations_stay_correct(cuda_device, generation):
    """B200TSDF_MC=1 (per-cube classification) and =2 (bit planes, CTA per block) are kept as
    the A/B baselines of mc_kernel3; the knob is read once per process, so the oracle
    comparison above is re-run in a child process."""
    import os
    import subprocess
    import sys

    env = dict(os.environ, B200TSDF_MC=generation)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_gpu_mesh.py", "-m", "gpu", "-q", "-x",
                        "-k", "matches_cpu_marching_cubes or stitched or empty"],
                       cwd=root, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
