"""Streaming kernels on a volume far larger than L2 (126 MB): the full 600-pose
orbit fused into the 1024^3 @ 5 mm scene on ONE GPU (~2 GB of voxel blocks),
then marching cubes, the stats sweep and the voxel export timed with CUDA
events against the HBM roofline.  python scripts/big_volume_sweeps.py [--kf 600]
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


def ms_of(fn, reps=5):
    best, out = None, None
    for _ in range(reps):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        out = fn()
        b.record()
        torch.cuda.synchronize()
        t = a.elapsed_time(b)
        best = t if best is None else min(best, t)
    return best, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kf", type=int, default=600)
    ap.add_argument("--batch", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    wl = bench.TARGET
    scene, dim, orbit_r = wl["scene"], wl["dim"], wl["orbit_r"]
    lo, hi = synth.bounds_for(scene, bench.VOXEL, dim)
    vol = TSDFVolume(bench.VOXEL, bench.TRUNC, capacity_log2=19, bounds=(lo, hi), device=dev)
    fuse_ms = 0.0
    for s in range(0, a.kf, a.batch):
        kf = synth.make_keyframes(scene, a.batch, bench.H, bench.W, orbit_radius=orbit_r,
                                  gh=bench.GH, gw=bench.GW, start=s, total=bench.ORBIT_TOTAL,
                                  up_axis="z")
        t = [torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")]
        ms, _ = ms_of(lambda: vol.integrate_keyframes(t[0], t[1], t[2], t[3], 0.05, bench.H,
                                                      bench.W), reps=1)
        fuse_ms += ms
    c = vol.counters()
    nblk = c["blocks"]
    vox_bytes = nblk * 512 * 24
    peak = 6551.4
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..",
                                                 "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:  # noqa: BLE001
        pass
    out = {"keyframes": a.kf, "blocks": nblk, "volume_mb": vox_bytes / 1e6,
           "fuse_ms_total": fuse_ms,
           "updates": c["updates_primary"] + c["updates_quality"], "peak_gbs": peak}
    verts = vol.extract_mesh(0)
    ntri = int(verts.shape[0] // 3)
    buf = torch.empty(max(ntri, 1) * 3, 3, dtype=torch.float32, device=dev)
    del verts
    api_ms, _ = ms_of(lambda: vol.extract_mesh(0, out=buf))
    cnt = torch.empty(1, dtype=torch.int64, device=dev)

    def passes():
        for _ in range(8):
            vol.extract_mesh_async(0, buf, cnt)
    ms, _ = ms_of(passes)
    ms /= 8
    b = vox_bytes + ntri * 36
    out["extract_mesh"] = {"ms": ms, "api_ms": api_ms, "triangles": ntri, "bytes": b,
                           "gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak,
                           "kernel": "mc_kernel%s" % os.environ.get("B200TSDF_MC", "3")}
    ms, (iv, jf) = ms_of(lambda: vol.extract_mesh_indexed(0), reps=2)
    out["extract_mesh_indexed"] = {"ms": ms, "vertices": int(iv.shape[0]),
                                   "faces": int(jf.shape[0]),
                                   "what": "count pass + write pass + device weld (hash + scan)"}
    ms, st = ms_of(lambda: vol.stats(0))
    out["stats"] = {"ms": ms, "bytes": vox_bytes, "gbs": vox_bytes / ms / 1e6,
                    "frac": vox_bytes / ms / 1e6 / peak, "valid_voxels": st["valid_voxels"]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
