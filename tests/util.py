"""Shared helpers of the parity tests (oracle = checker only)."""
import numpy as np

from oracle import oracle as orc

TSDF_TOL = 1.0e-5   # normalised tsdf units = 1e-5 of the truncation distance (north_star)
W_TOL = 1.0e-5      # |dw| <= 1e-5 * max(1, w)
# Where the tolerance is relative (DESIGN.md 3 "error model"): the stored state is
# the pair of float32 sums (S, W); every red.add.f32 rounds the running sum to 2^-24
# relative, so after k folds |dS| <~ sqrt(k) 2^-24 |S| (worst case k 2^-24 |S|).  The
# reference's clamp quirk (global_volume.py:84-87: the running average is divided by
# the CLAMPED weight) makes tsdf = S / max_weight grow past 1 once W > max_weight
# (|tsdf| reaches 7.7 in BASELINE config 1), and the absolute error grows with it.
# Inside the meaningful range |tsdf| <= 1 the bound stays the absolute 1e-5.


def oracle_dual(kf, voxel, trunc, min_conf=0.05, max_weight=100.0, use_q=True, idx=None,
                max_points=0, seed=0, kf_ids=None):
    """Reference semantics for a keyframe batch: two oracle volumes, the second
    fed confidences C*q.  max_points: the seeded
    max_points_per_kf subsample (oracle.subsample_mask), keyframe b salted with
    kf_ids[b] (default b)."""
    vp = orc.OracleVolume(voxel, trunc, max_weight)
    vq = orc.OracleVolume(voxel, trunc, max_weight)
    rays = 0
    for b in (idx if idx is not None else range(kf["X"].shape[0])):
        pts, conf, confq, org = orc.stage_keyframe(
            kf["X"][b], kf["C"][b], kf["pose"][b], kf["q"][b] if use_q else None,
            kf["H"], kf["W"], min_conf, max_points, seed,
            b if kf_ids is None else kf_ids[b])
        rays += vp.integrate(pts, conf, org)
        vq.integrate(pts, confq, org)
    return vp, vq, rays


def assert_volume_matches(got, want, what=""):
    """got/want = (keys, tsdf, weight, count) sorted by key."""
    gk, gt, gw, gc = got
    wk, wt, ww, wc = want
    assert gk.shape == wk.shape, f"{what}: voxel key set size {gk.shape[0]} != {wk.shape[0]}"
    assert np.array_equal(gk, wk), f"{what}: voxel key sets differ"
    assert np.array_equal(gc.astype(np.int64), wc.astype(np.int64)), f"{what}: update counts differ"
    dt = np.abs(gt.astype(np.float64) - wt) / np.maximum(1.0, np.abs(wt))
    assert dt.max(initial=0.0) <= TSDF_TOL, f"{what}: max |dtsdf| / max(1, |tsdf|) = {dt.max()}"
    dw = np.abs(gw.astype(np.float64) - ww) / np.maximum(1.0, np.abs(ww))
    assert dw.max(initial=0.0) <= W_TOL, f"{what}: max rel |dw| = {dw.max()}"


def block_set(keys):
    return np.unique(keys >> 3, axis=0)
