"""CPU: the C oracle restatement against golden vectors frozen from the
UNMODIFIED reference TSDFVolume (tests/golden/make_golden.py).  Bit-exact:
keys, counts, tsdf and weight (float64 equality), stats, query."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle as orc
from oracle import ref_harness
from tests.golden import make_golden as mg

CASES = sorted(mg.CASES)


def _load(case):
    return np.load(os.path.join(mg.HERE, case + ".npz"))


@pytest.mark.parametrize("case", CASES)
def test_oracle_matches_reference_golden(case):
    g = _load(case)
    scene, n_kf, H, W, f, voxel, trunc, mw, dtype, seed = mg.CASES[case]
    kf, staged = mg.staged_inputs(case)
    assert mg.checksum(staged).encode() == g["checksum"].tobytes(), "synthetic inputs drifted"
    for name, col in (("p", 1), ("q", 2)):
        vol = orc.OracleVolume(voxel, trunc, mw)
        fused = sum(vol.integrate(s[0], s[col], s[3]) for s in staged)
        k, t, w, c = vol.export()
        assert fused == int(g[f"{name}_fused"]) and vol.events == int(g[f"{name}_events"])
        assert np.array_equal(k, g[f"{name}_keys"].astype(np.int64))
        assert np.array_equal(c, g[f"{name}_count"])
        assert np.array_equal(t, g[f"{name}_tsdf"]), np.abs(t - g[f"{name}_tsdf"]).max()
        assert np.array_equal(w, g[f"{name}_w"])
        st = vol.stats()
        assert [st["valid_voxels"], st["total_voxels"]] == list(g[f"{name}_stats"])
        qt, qg, fl = vol.query_batch(g[f"{name}_probe"])
        assert np.array_equal(fl, g[f"{name}_qfl"])
        has = fl & 1 > 0
        assert np.array_equal(qt[has], g[f"{name}_qt"][has])
        hg = fl == 3
        assert np.allclose(qg[hg], g[f"{name}_qg"][hg], rtol=0, atol=1e-15)


def test_all_fixtures_have_cases():
    names = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(mg.HERE, "*.npz"))
                   if not os.path.basename(p).startswith(("local_", "mc_", "opt_", "refine_", "quality_")))
    assert names == CASES


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree absent (GPU box)")
def test_oracle_matches_reference_live():
    """Re-run the unmodified reference here and compare with the oracle on a
    fresh random case (not a frozen one)."""
    rng = np.random.default_rng(123)
    pts = (rng.normal(size=(400, 3)) * 0.3 + [0.1, 0.0, 1.3]).astype(np.float32)
    conf = 0.2 + rng.random(400)
    org = np.array([0.02, -0.01, 0.03], np.float32)
    ref = ref_harness.CountingVolume(0.03, 0.12, 4.0)
    vo = orc.OracleVolume(0.03, 0.12, 4.0)
    for _ in range(2):
        assert ref.integrate(pts, conf, org) == vo.integrate(pts, conf, org)
    for a, b in zip(ref.export(), vo.export()):
        assert np.array_equal(a, b)
