"""Per-rank kernel time of the sharded fuse, emulated on ONE GPU.

For R in {1,2,4,8} and every rank r, a volume that owns the interleaved
z-slabs of rank r fuses the same keyframe batches (the N>1 bench scene); the
step time a real R-GPU run would see is the max over r (plus the broadcast).
A tuning aid: python scripts/slab_rank_time.py [--kf 40] [--steps 4]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200 import synth  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kf", type=int, default=40)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--ranks", default="1,2,4,8")
    ap.add_argument("--width", default="0", help="slab widths to try, comma separated; 0: bench.slab_width_for(R)")
    ap.add_argument("--only-rank", type=int, default=-1)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    wl = bench.TARGET
    dim = wl["dim"]
    lo, hi = synth.bounds_for(wl["scene"], bench.VOXEL, dim)
    batches = []
    for s in range(a.warmup + a.steps):
        kf = bench.gen_batch(wl, 3 * s, a.kf)  # spread over the orbit
        batches.append(tuple(torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")))
    out = {}
    for R, width in ((int(x), int(w)) for x in a.ranks.split(",") for w in a.width.split(",")):
        per_rank = []
        for r in (range(R) if a.only_rank < 0 else [a.only_rank % R]):
            slabs = (R, r, width or bench.slab_width_for(R, dim)) if R > 1 else None
            vol = TSDFVolume(bench.VOXEL, bench.TRUNC, capacity_log2=wl["cap_log2"], bounds=(lo, hi),
                             slabs=slabs, device=dev)
            evs = []
            for i, (X, C, P, q) in enumerate(batches):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                vol.integrate_keyframes(X, C, P, q, 0.05, bench.H, bench.W)
                e1.record()
                torch.cuda.synchronize()
                if i >= a.warmup:
                    evs.append(e0.elapsed_time(e1))
            c = vol.counters()
            per_rank.append((sum(evs) / len(evs), int(c["updates_primary"] + c["updates_quality"]),
                             1.0 - c.get("culled_rays", 0) / max(1, c["rays"])))
            del vol
        ms = [p[0] for p in per_rank]
        upd = sum(p[1] for p in per_rank)
        out[(R, width)] = {"width": width or bench.slab_width_for(R, dim), "max_ms": max(ms), "mean_ms": sum(ms) / len(ms), "ms": [round(m, 4) for m in ms],
                  "updates": upd,
                  # share of the rays a rank marches / share of the updates it applies
                  "marched": [round(p[2], 4) for p in per_rank],
                  "applied": [round(p[1] / max(1, upd), 4) for p in per_rank]}
        print(R, json.dumps(out[(R, width)]), flush=True)


if __name__ == "__main__":
    main()
