"""Pose-only re-integration over real NCCL ranks:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
        --master-addr 127.0.0.1 --master-port 29631 scripts/sharded_reintegrate_check.py

Rank 0 holds the keyframes and, later, the new poses.  Every rank fuses the
broadcast batch into its z-slabs and keeps it in its store (`ids=`); after a
simulated backend solve only the poses are broadcast (`reintegrate`).  Checks on
every rank, against a fresh volume of the same slabs fused directly:
  replace  == keyframes fused once at the new poses (voxel key set and hit
              counters exact, tsdf / weight within the de-integration tolerance);
  additive == keyframes fused at the old AND the new poses (reference semantics).
Writes gpurun_out/sharded_reintegrate_n<N>.json."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_quality_dualtsdf_b200 import parallel, synth  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume  # noqa: E402

VOXEL, TRUNC, H, W = 0.02, 0.08, 48, 64


def same(a, b, tol):
    (gk, gt, gw, gc), (wk, wt, ww, wc) = a, b
    if gk.shape != wk.shape or not np.array_equal(gk, wk):
        return False, "keys"
    if not np.array_equal(gc.astype(np.int64), wc.astype(np.int64)):
        return False, "counts"
    if len(gt) and (np.abs(gt - wt).max() > tol or
                    (np.abs(gw - ww) / np.maximum(1.0, ww)).max() > tol):
        return False, "values"
    return True, ""


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    kf = synth.make_keyframes(synth.SCENES["sphere"], 4, H=H, W=W, seed=8, f=50.0, total=9)
    t = {k: torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")}
    new_pose = t["pose"].clone()
    new_pose[:, :3] += torch.tensor([0.03, -0.02, 0.025], device=dev)  # the backend moved them
    ids = [3, 5, 8, 13]
    sel, sel_ids = [2, 0, 3], [8, 3, 13]  # the solve touched three of the four
    res = {}
    for mode in ("replace", "additive"):
        sv = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=8, device=dev)
        src = {k: (v.clone() if rank == 0 else torch.zeros_like(v)) for k, v in t.items()}
        sv.integrate_keyframes(src["X"], src["C"], src["pose"], src["q"], 0.05, H, W, src=0,
                               ids=ids)
        p = new_pose[sel].contiguous() if rank == 0 else torch.zeros(len(sel), 8, device=dev)
        sv.reintegrate(sel_ids, p, src=0, mode=mode)
        want = TSDFVolume(VOXEL, TRUNC, device=dev, slabs=(world, rank, 8))
        if mode == "replace":
            pose = t["pose"].clone()
            pose[sel] = new_pose[sel]
            want.integrate_keyframes(t["X"], t["C"], pose, t["q"], 0.05, H, W)
        else:
            want.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)
            want.integrate_keyframes(t["X"][sel].contiguous(), t["C"][sel].contiguous(),
                                     new_pose[sel].contiguous(), t["q"][sel].contiguous(),
                                     0.05, H, W)
        oks = [same(sv.vol.export(w), want.export(w), 5e-5) for w in (0, 1)]
        store_ok = all(torch.equal(sv._store[i]["pose"], new_pose[b])
                       for b, i in zip(sel, sel_ids))
        ok = torch.tensor([float(all(o for o, _ in oks) and store_ok)], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        res[mode] = {"ok": bool(ok.item()), "why": [w for _, w in oks],
                     "voxels_rank0": int(sv.vol.stats(0)["total_voxels"])}
    good = all(r["ok"] for r in res.values())
    if rank == 0:
        out = {"n_gpus": world, "keyframes": len(ids), "re_fused": len(sel),
               "pose_bytes_broadcast": 32 * len(sel), **res}
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/sharded_reintegrate_n{world}.json", "w") as f:
            json.dump(out, f)
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if good else 1)


if __name__ == "__main__":
    main()
