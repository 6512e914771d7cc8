"""CPU model of the fuse kernel's reduction traffic by voxel layout (no GPU needed).

For warps of the fuse launch (8 x 4 pixel sub-tiles, lockstep march as in
fuse_fast.cuh: a lane folds the voxel it leaves, all lanes that leave a voxel in the
same iteration share ONE warp-level RED.128 + RED.64) it counts, per warp-level
reduction, the distinct 128-byte lines (L1TEX tag-stage requests) and 32-byte sectors
(L2 sector operations) the lanes touch, for the shipped block layout
(z*64 + y*8 + x: a 128-byte line = 8 voxels along x) and for brick layouts in which a
line holds a 2x2x2 / 4x2x1 group of voxels.  Sample positions follow
global_volume.py:58-71 in float64 (keys may differ from the float32 path in 1e-4 of
the samples; irrelevant for the statistics).
  python scripts/layout_study.py [--keyframes 3] [--stride 7]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def quat_R(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def line_ids(keys, layout):
    """128-byte line of the float4 voxel array / 64-byte half line of the u64 counters."""
    x, y, z = keys[..., 0], keys[..., 1], keys[..., 2]
    blk = ((z >> 3) * 4096 + (y >> 3)) * 4096 + (x >> 3)
    lx, ly, lz = x & 7, y & 7, z & 7
    if layout == "row":        # shipped: li = z*64 + y*8 + x, 8 voxels (one x row) per line
        sub = lz * 8 + ly
    elif layout == "brick222":  # a line = 2x2x2 voxels
        sub = ((lz >> 1) * 4 + (ly >> 1)) * 4 + (lx >> 1)
    elif layout == "brick421":  # a line = 4 (x) x 2 (y) x 1 (z)
        sub = (lz * 4 + (ly >> 1)) * 2 + (lx >> 2)
    else:
        raise ValueError(layout)
    return blk * 64 + sub


def sector_ids(keys, layout):
    """32-byte sector of the float4 voxel array: two voxels adjacent in x in every layout
    (x is the fastest index inside a line)."""
    return line_ids(keys, layout) * 4 + {
        "row": (keys[..., 0] & 7) >> 1,
        "brick222": ((keys[..., 2] & 1) * 2 + (keys[..., 1] & 1)),
        "brick421": ((keys[..., 1] & 1) * 2 + ((keys[..., 0] & 3) >> 1)),
    }[layout]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--keyframes", type=int, default=3)
    ap.add_argument("--stride", type=int, default=7, help="every stride-th tile of a keyframe")
    a = ap.parse_args()
    wl, vs, tr = bench.TARGET, bench.VOXEL, bench.TRUNC
    step = vs * 0.5
    H, W = bench.H, bench.W
    stats = {k: {"reds": 0, "lanes": 0, "lines": 0, "sectors": 0}
             for k in ("row", "brick222", "brick421")}
    iters = 0
    for kfi in range(a.keyframes):
        kf = bench.gen_batch(wl, (bench.ORBIT_TOTAL // a.keyframes) * kfi, 1)  # spread over the orbit
        X = kf["X"][0].reshape(H, W, 3).astype(np.float64)
        C = kf["C"][0].reshape(H, W)
        p = kf["pose"][0].astype(np.float64)
        R, t, s = quat_R(p[3:7]), p[:3], p[7]
        tiles = [(by, bx) for by in range(H // 8) for bx in range(W // 32)][::a.stride]
        for by, bx in tiles:
            for w in range(8):
                ys = by * 8 + (w >> 2) * 4 + (np.arange(32) >> 3)
                xs = bx * 32 + (w & 3) * 8 + (np.arange(32) & 7)
                ok = C[ys, xs] > bench.MIN_CONF
                pw = s * (X[ys, xs] @ R.T) + t
                ray = pw - t
                ln = np.linalg.norm(ray, axis=1)
                ok &= ln >= 1e-4
                d = ray / np.maximum(ln, 1e-9)[:, None]
                maxd = ln + tr
                n = np.maximum((maxd / step).astype(np.int64), 1)
                keys, cnt = [], np.zeros(32, np.int64)
                for l in range(32):  # in-band samples of lane l (global_volume.py:63-67)
                    if not ok[l]:
                        keys.append(np.zeros((0, 3), np.int64))
                        continue
                    dist = np.linspace(0.0, maxd[l], n[l])
                    band = np.abs(ln[l] - dist) <= tr
                    pos = t + dist[band, None] * d[l]
                    keys.append(np.floor(pos / vs).astype(np.int64))
                    cnt[l] = band.sum()
                J = int(cnt.max())
                iters += J
                for j in range(J):  # lockstep: lane folds at j when its next key differs / band ends
                    fold = [l for l in range(32) if j < cnt[l] and
                            (j + 1 == cnt[l] or np.any(keys[l][j + 1] != keys[l][j]))]
                    if not fold:
                        continue
                    kk = np.stack([keys[l][j] for l in fold])
                    for lay, st in stats.items():
                        st["reds"] += 1
                        st["lanes"] += len(fold)
                        st["lines"] += len(np.unique(line_ids(kk, lay)))
                        st["sectors"] += len(np.unique(sector_ids(kk, lay)))
    out = {"warp_iterations": iters}
    for lay, st in stats.items():
        out[lay] = {"warp_reds": st["reds"], "lanes_per_red": round(st["lanes"] / st["reds"], 2),
                    "lines_per_red_128": round(st["lines"] / st["reds"], 2),
                    "sectors_per_red_128": round(st["sectors"] / st["reds"], 2)}
    out["reds_per_warp_iteration"] = round(stats["row"]["reds"] / iters, 3)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
