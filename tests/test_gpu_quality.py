"""GPU: the device quality producer (csrc/quality.cu behind the C ABI) against
the vectors frozen from the unmodified reference and against the CPU oracle."""
import numpy as np
import pytest
import torch

from oracle import quality_oracle as qo
from tests.golden import make_golden_quality as mg
from tests.test_quality_oracle_golden import U_TOL, check_quality, job_numpy

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", mg.CASES, ids=[c[0] for c in mg.CASES])
def test_quality_producer_matches_reference(cuda_device, case):
    from mast3r_slam_quality_dualtsdf_b200 import quality_core as qc

    name, seed, H, W, with_prev, holes = case
    job = mg.make_job(seed, H, W, with_prev, holes)
    res = qc.compute_batch([dict(job)], device=cuda_device, **mg.PARAMS)[0]
    assert res["priority"].is_cuda and res["q_grid"].is_cuda  # nothing left the device
    got = {k: (v.cpu().numpy() if torch.is_tensor(v) else v) for k, v in res.items()}
    check_quality(got, name)
    # CUDA sqrt is IEEE like numpy's: against the oracle u is bit-exact
    want = qo.compute_job(job_numpy(job), **mg.PARAMS)
    assert np.array_equal(got["u"], want["u"], equal_nan=True)
    assert np.array_equal(got["r"], want["r"]) and np.array_equal(got["class_id"], want["class_id"])
    np.testing.assert_allclose(got["priority"], want["priority"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(got["q_grid"], 1.0 - np.clip(got["priority"], 0, 1), atol=0)
    assert U_TOL > 0


def test_quality_grid_feeds_the_fuse(cuda_device):
    """q = 1 - priority goes from the producer to integrate_keyframes without a
    host copy; the quality volume equals the one fused with the same grid from
    the reference-format dict (pack_result -> quality_grid_from_result)."""
    from mast3r_slam_quality_dualtsdf_b200 import quality_core as qc
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume
    from mast3r_slam_quality_dualtsdf_b200.tsdf.global_manager import quality_grid_from_result

    H, W = 48, 64
    kf = synth.make_keyframes(synth.SCENES["sphere"], 1, H=H, W=W, seed=5, f=50.0, total=8)
    job = mg.make_job(7, H, W, False, False)
    res = qc.compute_batch([dict(job)], device=cuda_device, **mg.PARAMS)[0]
    q_dev = qc.quality_grid(res)
    host = qc.pack_result(0, *(res[k] for k in ("delta_cov", "r", "u", "class_id", "priority",
                                                "cov_ewma")), to_host=True)
    q_ref = quality_grid_from_result(host, H // 16, W // 16, cuda_device)
    assert torch.equal(q_dev, q_ref) and float(q_dev.min()) < 1.0
    vols = []
    for q in (q_dev, q_ref):
        v = TSDFVolume(0.02, 0.08, device=cuda_device)
        v.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                              q, 0.05, H, W)
        vols.append(v.export(1))
    assert all(np.array_equal(a, b) for a, b in zip(*vols))


def test_quality_batch_equals_per_keyframe(cuda_device):
    """compute_batch runs every keyframe of one size through two launches
    (b200tsdf_quality_batch); each job's grids are bit-identical to the
    one-keyframe-at-a-time statement, for mixed sizes, holes and with / without
    a previous coverage EMA in the same batch."""
    from mast3r_slam_quality_dualtsdf_b200 import quality_core as qc

    specs = [(3, 96, 128, True, False), (4, 96, 128, False, True), (5, 64, 64, True, True),
             (6, 96, 128, False, False), (7, 64, 64, False, False), (8, 96, 128, True, True)]
    jobs = []
    for k, (seed, H, W, with_prev, holes) in enumerate(specs):
        jb = dict(mg.make_job(seed, H, W, with_prev, holes))
        jb["kf_id"] = k
        jobs.append(jb)
    got = qc.compute_batch([dict(j) for j in jobs], device=cuda_device, **mg.PARAMS)
    want = qc._compute_batch_per_job([dict(j) for j in jobs], device=cuda_device, **mg.PARAMS)
    assert [g["kf_id"] for g in got] == list(range(len(specs)))
    for g, w in zip(got, want):
        for key in ("delta_cov", "cov_ewma", "r", "u", "class_id", "priority", "q_grid"):
            a, b = g[key].cpu().numpy(), w[key].cpu().numpy()
            assert a.shape == b.shape and np.array_equal(a, b, equal_nan=True), key
    assert qc.compute_batch([], device=cuda_device, **mg.PARAMS) == []


def test_quality_argument_errors(cuda_device):
    from mast3r_slam_quality_dualtsdf_b200 import _ffi
    from mast3r_slam_quality_dualtsdf_b200 import quality_core as qc

    z = torch.zeros(64 * 64, device=cuda_device)
    with pytest.raises(_ffi.B200TSDFError):
        qc.patch_grids(z, z, z, z > 0, 64, 64, 64, 1.5, 1.5, cuda_device)  # patch_size > 32
    with pytest.raises(_ffi.B200TSDFError):
        qc.classify(torch.zeros(4000, device=cuda_device), torch.zeros(4000, device=cuda_device),
                    torch.zeros(4000, device=cuda_device))  # n > 3072
