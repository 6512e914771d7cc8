"""CPU: the numpy restatement of the reference quality producer against vectors
frozen from the unmodified reference (tests/golden/make_golden_quality.py)."""
import os

import numpy as np
import pytest

from oracle import quality_oracle as qo
from tests.golden import make_golden_quality as mg

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "quality_cases.npz"))


def job_numpy(job):
    return {k: (v.numpy() if hasattr(v, "numpy") else v) for k, v in job.items()}


@pytest.mark.parametrize("case", mg.CASES, ids=[c[0] for c in mg.CASES])
def test_quality_oracle_matches_reference(case):
    name, seed, H, W, with_prev, holes = case
    got = qo.compute_job(job_numpy(mg.make_job(seed, H, W, with_prev, holes)), **mg.PARAMS)
    check_quality(got, name)


# Tolerances: every quantity is bit-exact except `u` and what depends on it.
# torch's vectorised CPU sqrt (the build that froze the goldens) is 1 ulp below
# the IEEE result on ~0.5 % of inputs; numpy and CUDA sqrt are correctly rounded.
# u in [0, 1]: 1 ulp = 6e-8.  priority divides by the MAD of u: 2e-5.
U_TOL, PRI_TOL = 1.2e-7, 2.0e-5


def check_quality(got, name):
    for k in ("delta_cov", "r", "cov_ewma"):
        assert np.array_equal(np.asarray(got[k]), G[f"{name}/{k}"], equal_nan=True), k
    np.testing.assert_allclose(np.asarray(got["u"]), G[f"{name}/u"], rtol=0, atol=U_TOL,
                               equal_nan=True)
    assert np.array_equal(np.asarray(got["class_id"]), G[f"{name}/class_id"])
    np.testing.assert_allclose(np.asarray(got["priority"]), G[f"{name}/priority"], rtol=0,
                               atol=PRI_TOL)


def test_reference_helper_functions_are_kept():
    """every public function of the reference module exists in the
    mirror with the same results (live check when /root/reference is present;
    the names are checked always)."""
    import importlib.util
    import os

    import torch

    from mast3r_slam_quality_dualtsdf_b200 import quality_core as q

    names = ["_hw", "_grid", "reduce_grid", "upsample_to_hw", "view_weight", "ema_delta",
             "u_from_CQ", "r_from_scalar", "valid_grid", "robust_z", "classify", "pack_result",
             "compute_batch"]
    assert all(callable(getattr(q, n, None)) for n in names)
    path = "/root/reference/mast3r_slam/quality_core.py"
    if not os.path.exists(path):
        return
    spec = importlib.util.spec_from_file_location("refq", path)
    r = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(r)
    g = torch.Generator().manual_seed(0)
    x = torch.rand(48 * 64, generator=g)
    v = torch.rand(48 * 64, generator=g) > 0.4
    for m in ("median", "mean"):
        assert torch.equal(q.reduce_grid(x, 48, 64, 16, method=m), r.reduce_grid(x, 48, 64, 16, method=m))
        assert torch.allclose(q.reduce_grid(x, 48, 64, 16, valid=v, method=m),
                              r.reduce_grid(x, 48, 64, 16, valid=v, method=m))
    assert torch.equal(q.valid_grid(v, 48, 64, 16), r.valid_grid(v, 48, 64, 16))
    assert torch.equal(q.u_from_CQ(x * 3, x * 2, 1.5, 1.5, 48, 64, 16),
                       r.u_from_CQ(x * 3, x * 2, 1.5, 1.5, 48, 64, 16))
    assert torch.equal(q.robust_z(x), r.robust_z(x))
    assert torch.equal(q.r_from_scalar(x, 48, 64, 16, valid=v), r.r_from_scalar(x, 48, 64, 16, valid=v))
