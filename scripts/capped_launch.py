"""Device time of the 40-keyframe fuse launch with the reference's default
max_points_per_kf cap (config/base.yaml:116) on the Target volume; run under
`ncu --metrics gpu__time_duration.sum` for the threshold / march split.
  python scripts/capped_launch.py [--cap 40000] [--launches 6]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200 import synth  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cap", type=int, default=40000)
ap.add_argument("--launches", type=int, default=6)
a = ap.parse_args()
dev = torch.device("cuda:0")
wl = bench.TARGET
lo, hi = synth.bounds_for(wl["scene"], bench.VOXEL, wl["dim"])
vol = TSDFVolume(bench.VOXEL, bench.TRUNC, capacity_log2=wl["cap_log2"], bounds=(lo, hi), device=dev)
batches = []
for l in range(a.launches):
    kf = bench.gen_batch(wl, l, 40)
    batches.append({k: torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")})
out = {}
for name, cap in (("capped", a.cap), ("all_pixels", 0)):
    for rep in range(2):  # first repetition allocates the blocks
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ev[0].record()
        for b in batches:
            vol.integrate_keyframes(b["X"], b["C"], b["pose"], b["q"], bench.MIN_CONF, bench.H, bench.W,
                                    max_points=cap, seed=3)
        ev[1].record()
        torch.cuda.synchronize()
    out[name] = {"ms_per_launch": ev[0].elapsed_time(ev[1]) / a.launches}
c = vol.counters()
out["overflow"] = c.get("overflow", 0)
print(json.dumps(out))
