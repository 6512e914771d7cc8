"""CPU: C restatement of the reference LOCAL TSDF (oracle/local_oracle.c)
against vectors frozen from the unmodified reference methods on CPU tensors."""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.golden import make_golden_local as mgl


@pytest.mark.parametrize("case", sorted(mgl.CASES))
def test_local_oracle_matches_reference_golden(case):
    g = np.load(os.path.join(mgl.HERE, case + ".npz"))
    X, Cf, pose, lo, hi, mask, order, H, W = mgl.case_inputs(case)
    c = mgl.CFG
    tsdf, w, fused = orc.local_build(X, Cf, pose, lo, hi, c["voxel_size"], c["trunc_dist"],
                                     c["max_grid_dim"], c["min_confidence"])
    assert tsdf.shape == g["tsdf"].shape and fused > 0
    assert np.array_equal(tsdf, g["tsdf"]), np.abs(tsdf - g["tsdf"]).max()
    assert np.array_equal(w, g["w"])
    pix = np.nonzero(mask)[0][order]
    Xr, hits = orc.local_raycast(g["tsdf"], lo, hi, X, pix, c["ray_samples"],
                                 c["max_displacement"])
    want_hits = g["hits"][order]
    assert np.array_equal(hits.astype(bool), want_hits)
    assert int(g["hits"].sum()) == int(want_hits.sum())
    assert np.array_equal(Xr, g["X_refined"])
