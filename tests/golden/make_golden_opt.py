"""Freeze golden vectors of the UNMODIFIED reference TSDFPoseOptimizer
(mast3r_slam/tsdf/tsdf_optimizer.py) for SURVEY 8(a) row A9.  Run in the build
container (needs /root/reference):  python -m tests.golden.make_golden_opt

lietorch is a missing submodule: Sim3.act/exp/mul come from the stub in
oracle/ref_harness.py (matrix exponential) -> parity is unpinned AT THE
LIETORCH BOUNDARY only; everything else is the reference's own code.
"""
import os

import numpy as np
import torch

from mast3r_slam_quality_dualtsdf_b200 import synth
from oracle import oracle as orc
from oracle import ref_harness as rh

HERE = os.path.dirname(os.path.abspath(__file__))
CFG = {"samples_per_kf": 300, "min_confidence": 0.05, "max_iterations": 3, "lambda": 0.15,
       "damping": 1.0e-4, "pre_icp_iters": 0}
VOXEL, TRUNC, H, W, SEED = 0.03, 0.12, 48, 64, 123


def scene():
    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=H, W=W, seed=9, f=50.0, total=12)
    bad = kf["pose"][1].copy()
    bad[:3] += np.array([0.012, -0.008, 0.01], np.float32)
    q = bad[3:7].astype(np.float64) + np.array([0.004, -0.003, 0.002, 0.0])
    bad[3:7] = (q / np.linalg.norm(q)).astype(np.float32)
    bad[7] *= 1.01
    return kf, bad


CASES = {"iter1": dict(iterations=1, damping=1.0e-4), "damped3": dict(iterations=3, damping=500.0)}


def main():
    mod, lt = rh.load_tsdf_optimizer()
    gv = rh.load_global_volume()
    kf, bad = scene()
    vol = gv.TSDFVolume(VOXEL, TRUNC)
    for b in range(3):
        pts, conf, _, org = orc.stage_keyframe(kf["X"][b], kf["C"][b], kf["pose"][b], None, H, W,
                                               CFG["min_confidence"])
        vol.integrate(pts, conf, org)
    out = {}
    for name, c in CASES.items():
        poses = [kf["pose"][0], bad, kf["pose"][2]]
        kfs = rh.FakeKeyframes(kf["X"], kf["C"], poses, lt)
        cfg = dict(CFG, damping=c["damping"], max_iterations=c["iterations"])
        opt = mod.TSDFPoseOptimizer(vol, kfs, cfg, False, "cpu")
        torch.manual_seed(SEED)
        opt.optimize_keyframes([1], "golden")
        out[name] = kfs.T_WC[1].detach().cpu().numpy().reshape(8).astype(np.float64)
        print(name, out[name])
    # the permutation the reference drew (same generator state)
    torch.manual_seed(SEED)
    conf = torch.from_numpy(kf["C"][1]).reshape(-1)
    valid_idx = torch.nonzero(conf > CFG["min_confidence"]).reshape(-1)
    choice = valid_idx[torch.randperm(valid_idx.numel())[:CFG["samples_per_kf"]]].numpy()
    np.savez_compressed(os.path.join(HERE, "opt_sphere.npz"), pose_in=bad, choice=choice,
                        pose_true=kf["pose"][1], **{"pose_out_" + k: v for k, v in out.items()})
    print("pose_in ", bad)


if __name__ == "__main__":
    main()
