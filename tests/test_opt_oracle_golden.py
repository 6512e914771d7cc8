"""CPU: oracle restatement of the reference TSDFPoseOptimizer vs the golden
frozen from the unmodified reference (tests/golden/make_golden_opt.py)."""
import os

import numpy as np
import pytest

from oracle import opt_oracle
from oracle import oracle as orc
from tests.golden import make_golden_opt as mg

G = np.load(os.path.join(os.path.dirname(mg.__file__), "opt_sphere.npz"))


def oracle_volume(kf):
    vol = orc.OracleVolume(mg.VOXEL, mg.TRUNC)
    for b in range(3):
        pts, conf, _, org = orc.stage_keyframe(kf["X"][b], kf["C"][b], kf["pose"][b], None, mg.H,
                                               mg.W, mg.CFG["min_confidence"])
        vol.integrate(pts, conf, org)
    return vol


@pytest.mark.parametrize("case", sorted(mg.CASES))
def test_opt_oracle_matches_reference_golden(case):
    kf, bad = mg.scene()
    assert np.array_equal(bad, G["pose_in"])
    vol = oracle_volume(kf)
    c = mg.CASES[case]
    got = opt_oracle.optimize_single(vol, kf["X"][1], kf["C"][1], bad, G["choice"],
                                     iterations=c["iterations"], lam=mg.CFG["lambda"],
                                     damping=c["damping"])
    want = G["pose_out_" + case]
    # float32 pose storage on both sides: agreement to float32 rounding of a 7x7 solve
    assert np.allclose(got, want, rtol=2e-4, atol=2e-5), (got, want)
