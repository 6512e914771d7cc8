"""Freeze golden vectors of the UNMODIFIED reference TSDFRefiner block flow
(_schedule_refinement -> _select_blocks_enhanced -> _refine_block_enhanced,
/root/reference/mast3r_slam/tsdf_refine.py:348-835) on CPU tensors.
Build container only:  python -m tests.golden.make_golden_refine
"""
import contextlib
import io
import os

import numpy as np
import torch

from mast3r_slam_quality_dualtsdf_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
H, W = 48, 64
CFG = {"enabled": True, "voxel_size": 0.02, "trunc_dist": 0.08, "max_grid_dim": 64,
       "roi_size": 0.4, "ray_samples": 64, "max_displacement": 0.015,
       "min_weight_threshold": 0.01, "min_confidence": 0.2, "min_hit_rate": 0.05,
       "confidence_boost": 0.08, "confidence_max": 1.3, "max_rois_per_kf": 3,
       "cooldown_frames": 35, "window_size": 3}
# three tied top patches pass the 95 % quantile (tsdf_refine.py:455); patch (1, 0) is made
# low-confidence so one block exercises a rejection path
PRIORITY = np.array([[0.1, 0.2, 0.9, 0.3], [0.9, 0.05, 0.4, 0.7], [0.0, 0.9, 0.1, 0.2]], np.float32)


def perm(n):
    """The pinned stand-in for the reference's unseeded torch.randperm (:965)."""
    return torch.from_numpy(np.random.default_rng(1000 + n).permutation(n).astype(np.int64))


class Quality:
    def get(self, frame_id):
        return {"priority": PRIORITY.copy(), "patch_size": 16}

    def poll(self):
        pass


def inputs():
    kf = synth.make_keyframes(synth.SCENES["plane"], 1, H=H, W=W, seed=5, f=50.0, total=12, start=3)
    pose = np.array([0, 0, 0, 0, 0, 0, 1, 1], np.float32)  # identity (Appendix A #7)
    K = np.array([[50.0, 0, 32], [0, 50.0, 24], [0, 0, 1]], np.float32)
    Cf = kf["C"][0].copy().reshape(H, W)
    Cf[16:32, 0:16] = 0.1
    return kf["X"][0].copy(), Cf.reshape(-1).copy(), pose, K


def run_flow(refiner, kfs):
    """Drive one keyframe through scheduling and every queued block."""
    ok = refiner._schedule_refinement(0)
    results = []
    while not refiner.queue.empty():
        key, block = refiner.queue.get_nowait()
        assert refiner.registry.begin_run(key)
        success, gain = refiner._refine_block_enhanced(block)
        refiner.registry.finish_run(key, success, gain)
        results.append((block.block_id, block.patch_indices[0][0], block.patch_indices[0][1],
                        float(block.priority), float(block.depth_median), int(success),
                        float(gain)))
    return ok, results


def main():
    from oracle import ref_harness as rh

    mod, lt = rh.load_tsdf_refine()
    X, Cf, pose, K = inputs()
    kfs = rh.FakeKeyframes([X], [Cf.copy()], [pose], lt, img_shape=(H, W), K=torch.from_numpy(K))
    real = torch.randperm
    torch.randperm = lambda n, **k: perm(n)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            ref = mod.TSDFRefiner(dict(CFG), kfs, Quality(), device="cpu")
            ok, results = run_flow(ref, kfs)
    finally:
        torch.randperm = real
    np.savez_compressed(os.path.join(HERE, "refine_plane.npz"), ok=ok,
                        results=np.array(results, np.float64), C_after=kfs.C[0].numpy(),
                        version=kfs.version.numpy(),
                        debug=np.array([ref.stats["debug_info"][k] for k in
                                        ("tsdf_constructions", "surface_extractions",
                                         "displacement_rejects", "hit_ratio_rejects")]))
    print(ok, results, kfs.version, ref.stats["debug_info"],
          "boosted", int((kfs.C[0].numpy().reshape(-1) != np.minimum(Cf, 1.3)).sum()))


if __name__ == "__main__":
    main()
