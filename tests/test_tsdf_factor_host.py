"""CPU: host logic of tsdf.TSDFFactor (scatter of the per-keyframe blocks into the
backend's system; roadmap MASt3R_SLAM_TSDF_Integration.md 9.3.2).  The blocks
themselves come from the CUDA kernel and are checked in tests/test_gpu_manager.py."""
import numpy as np

from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFFactor


def test_add_to_scatters_on_the_diagonal_in_variable_order():
    rng = np.random.default_rng(0)
    blocks = {}
    for idx in (3, 5, 9):
        J = rng.standard_normal((20, 7))
        blocks[idx] = (J.T @ J, J.T @ rng.standard_normal(20), 20)
    order = [5, 9, 4]  # 3 is pinned / not in the window, 4 has no factor
    H = np.eye(21) * 0.5
    b = np.ones(21)
    assert TSDFFactor.add_to(H, b, blocks, order) == 2
    assert np.array_equal(H[0:7, 0:7], blocks[5][0] + 0.5 * np.eye(7))
    assert np.array_equal(H[7:14, 7:14], blocks[9][0] + 0.5 * np.eye(7))
    assert np.array_equal(H[14:, 14:], 0.5 * np.eye(7))
    assert not H[0:7, 7:].any() and not H[7:14, 14:].any()
    assert np.array_equal(b[0:7], 1 + blocks[5][1]) and np.array_equal(b[14:], np.ones(7))
    assert np.allclose(H, H.T)
    assert TSDFFactor.add_to(H, b, {}, order) == 0
