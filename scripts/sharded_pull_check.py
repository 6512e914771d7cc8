"""Peer-pull ingest over real NCCL ranks and real peer memory:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port 29651 scripts/sharded_pull_check.py

Every rank owns 8 / N ... keyframes per slot in a `PeerKeyframes` ring (its own
HBM, mapped into the other ranks with CUDA IPC).  `integrate_slot` all-gathers
only the meta records; the fuse kernel of every rank TMA-loads the tiles that
reach its z-slabs from the owners' memory over NVLink.  Rank 0 merges the ranks'
voxel blocks and compares them with ONE volume fused directly: voxel key sets and
hit counters exact, tsdf / weight within 1e-5; rays / applied updates add up.
Writes gpurun_out/sharded_pull_n<N>.json."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_quality_dualtsdf_b200 import parallel, synth  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import pack_keyframes  # noqa: E402

VOXEL, TRUNC, H, W, GH, GW, DIM, N_KF, SLOTS = 0.01, 0.04, 96, 128, 6, 8, 512, 8, 2


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    scene = synth.SCENES["room"]
    lo, hi = synth.bounds_for(scene, VOXEL, DIM)
    share = N_KF // world
    width = max(8, DIM // (4 * world))
    sv = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=width, bounds=(lo, hi),
                                    capacity_log2=16, device=dev)
    peers = parallel.PeerKeyframes(SLOTS, share, H, W, GH, GW, device=dev)
    batches = []
    for s in range(SLOTS):
        kf = synth.make_keyframes(scene, N_KF, H=H, W=W, seed=5 + s, f=100.0, orbit_radius=0.6,
                                  gh=GH, gw=GW, start=s * N_KF, total=24, up_axis="z")
        batches.append(kf)
        sl = slice(rank * share, (rank + 1) * share)
        pack_keyframes(torch.from_numpy(kf["X"][sl]).to(dev), torch.from_numpy(kf["C"][sl]).to(dev),
                       torch.from_numpy(kf["pose"][sl]).to(dev),
                       torch.from_numpy(kf["q"][sl]).to(dev), out=peers.slot(s))
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    if os.environ.get("PULL_SERIAL") == "1":
        for s in range(SLOTS):
            sv.integrate_slot(peers, s, 0.05)
    else:  # software-pipelined: meta pass + all-gather of slot s+1 behind the fuse of slot s
        sv.integrate_slots(peers, range(SLOTS), 0.05)
    c = sv.counters()
    dt = time.perf_counter() - t0
    bk, vox, cnt = sv.vol.export_block_data()
    parts = [None] * world
    dist.all_gather_object(parts, (bk.cpu().numpy(), vox.cpu().numpy(), cnt.cpu().numpy()))
    ok = True
    res = {"world": world, "pipelined": os.environ.get("PULL_SERIAL") != "1", "slab_width": width, "keyframes": N_KF * SLOTS, "seconds": dt}
    if rank == 0:
        merged = TSDFVolume(VOXEL, TRUNC, capacity_log2=16, bounds=(lo, hi), device=dev)
        for pb, pv, pc in parts:
            merged.import_blocks(torch.from_numpy(pb), torch.from_numpy(pv), torch.from_numpy(pc))
        direct = TSDFVolume(VOXEL, TRUNC, capacity_log2=16, bounds=(lo, hi), device=dev)
        for kf in batches:
            t = {k: torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")}
            direct.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)
        dc = direct.counters()
        for which in (0, 1):
            a, b = merged.export(which), direct.export(which)
            same = a[0].shape == b[0].shape and np.array_equal(a[0], b[0]) and \
                np.array_equal(a[3], b[3])
            dmax = float(np.abs(a[1] - b[1]).max()) if same else None
            res[f"volume{which}"] = {"keys_and_hit_counts_exact": bool(same), "max_abs_dtsdf": dmax,
                                     "voxels": int(a[0].shape[0])}
            ok = ok and same and dmax <= 1e-5
        sums = c["updates_primary"] == dc["updates_primary"] and \
            c["updates_quality"] == dc["updates_quality"] and c["rays"] == dc["rays"]
        res["counters_add_up"] = bool(sums)
        res["culled_rays_all_ranks"] = c["culled_rays"]
        ok = ok and sums and a[0].shape[0] > 10000
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    peers.close()
    if rank == 0:
        res["ok"] = bool(ok)
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/sharded_pull_n{world}.json", "w") as f:
            json.dump(res, f, indent=1)
        print(json.dumps(res))
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
