"""Freeze golden vectors of the LOCAL (per-ROI) TSDF by running the UNMODIFIED
reference methods TSDFRefiner._build_tsdf_robust / _extract_surface_safe
(/root/reference/mast3r_slam/tsdf_refine.py:837-1021) on CPU tensors, with the
lietorch.Sim3 stand-in of oracle/ref_harness.py and torch.randperm pinned to a
seeded permutation.  Build container only:  python tests/golden/make_golden_local.py
"""
import contextlib
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from mast3r_slam_quality_dualtsdf_b200 import synth  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CFG = {"voxel_size": 0.02, "trunc_dist": 0.08, "max_grid_dim": 64, "roi_size": 0.4,
       "ray_samples": 64, "max_displacement": 0.015, "min_weight_threshold": 0.01,
       "min_confidence": 0.2}

CASES = {
    # name: (scene, H, W, f, seed, identity_pose, block (y0, x0, size))
    "local_plane_id": ("plane", 48, 64, 50.0, 5, True, (16, 24, 16)),
    "local_sphere_id": ("sphere", 48, 64, 50.0, 6, True, (16, 24, 16)),
    "local_plane_pose": ("plane", 48, 64, 50.0, 7, False, (8, 8, 16)),
}


def case_inputs(case):
    scene, H, W, f, seed, ident, (y0, x0, sz) = CASES[case]
    kf = synth.make_keyframes(synth.SCENES[scene], 1, H=H, W=W, seed=seed, f=f, total=12, start=3)
    X, Cf, pose = kf["X"][0].copy(), kf["C"][0].copy(), kf["pose"][0].copy()
    if ident:
        pose = np.array([0, 0, 0, 0, 0, 0, 1, 1], np.float32)
    else:
        pose[:3] *= 0.05  # small motion: the camera-frame ROI still overlaps the world points
        pose[3:6] *= 0.1
        pose[3:7] /= np.linalg.norm(pose[3:7].astype(np.float64))
        pose[7] = 1.0
    mask = np.zeros((H, W), bool)
    mask[y0:y0 + sz, x0:x0 + sz] = True
    mask = mask.reshape(-1)
    pts = X[mask & (Cf > 0.03)]
    margin = max(0.01, min(0.05, 0.1 * CFG["roi_size"]))  # tsdf_refine.py:715
    lo = (pts.min(0) - np.float32(margin)).astype(np.float32)
    hi = (pts.max(0) + np.float32(margin)).astype(np.float32)
    rng = np.random.default_rng(seed)
    order = rng.permutation(int(mask.sum()))[:100]
    return X, Cf, pose, lo, hi, mask, order, H, W


def main():
    import torch

    from oracle import ref_harness

    mod, lt = ref_harness.load_tsdf_refine()

    class KF:  # minimal duck type; only the constructor touches it
        lock = contextlib.nullcontext()

        def __len__(self):
            return 0

    with contextlib.redirect_stdout(io.StringIO()):
        ref = mod.TSDFRefiner(dict(CFG), KF(), None, device="cpu")
    for case in CASES:
        X, Cf, pose, lo, hi, mask, order, H, W = case_inputs(case)
        Xt, Ct = torch.from_numpy(X), torch.from_numpy(Cf)
        T = lt.Sim3(torch.from_numpy(pose).reshape(1, 8))
        lo_t, hi_t = torch.from_numpy(lo), torch.from_numpy(hi)
        with contextlib.redirect_stdout(io.StringIO()):
            tsdf, w = ref._build_tsdf_robust(Xt, Ct, None, lo_t, hi_t, H, W, T)
        real = torch.randperm
        torch.randperm = lambda n, **k: torch.from_numpy(
            np.concatenate([order, np.setdiff1d(np.arange(n), order)]).astype(np.int64))
        try:
            with contextlib.redirect_stdout(io.StringIO()):
                Xr, hits = ref._extract_surface_safe(tsdf, lo_t, hi_t, None,
                                                     torch.from_numpy(mask), H, W, Xt)
        finally:
            torch.randperm = real
        np.savez_compressed(os.path.join(HERE, case + ".npz"), tsdf=tsdf.numpy(), w=w.numpy(),
                            X_refined=Xr.numpy(), hits=hits.numpy())
        print(case, "grid", tuple(tsdf.shape), "valid voxels", int((w > 0.01).sum()),
              "hits", int(hits.sum()), "/", len(order))


if __name__ == "__main__":
    main()
