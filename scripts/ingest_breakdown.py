"""Where the N > 1 end-to-end step goes (not yet run on the GPU pool: written at
the end of round 1, when `profiles/r01j_e2e_chunks.txt` showed the N = 8 step
at 1.42 ms against ~0.8 ms of PCIe + NVLink + fuse):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 \
        --master-addr 127.0.0.1 --master-port 29641 scripts/ingest_breakdown.py

Times, per step and as the max over ranks (CUDA events, barrier + synchronize
on both sides): the pinned-host -> device copy of the rank's share alone, the
four all-gathers alone, the fuse alone, the counter read-back alone, and the
serial composition the bench's e2e leg runs.  Writes
gpurun_out/ingest_breakdown_n<N>.json."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200 import synth  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    B, steps, warm = bench.KF_PER_STEP, 8, 3
    share = B // world
    scene, dim, orbit_r = bench.scene_for(world)
    lo, hi = synth.bounds_for(scene, bench.VOXEL, dim)
    vol = TSDFVolume(bench.VOXEL, bench.TRUNC, capacity_log2=19, bounds=(lo, hi),
                     slabs=(world, rank, bench.slab_width_for(world, dim)), device=dev)
    names = ("X", "C", "q", "pose")
    host = []
    for s in range(steps + warm):
        kf = bench.gen_batch(scene, orbit_r, s, B, up_axis="z", offset=rank * share, count=share)
        host.append({k: torch.from_numpy(kf[k]).pin_memory() for k in names})
    loc = {k: torch.empty_like(host[0][k], device=dev) for k in names}
    full = {k: torch.empty((B,) + tuple(host[0][k].shape[1:]), dtype=torch.float32, device=dev)
            for k in names}

    def h2d(s):
        for k in names:
            loc[k].copy_(host[s][k], non_blocking=True)

    def gather(s):
        for k in names:
            dist.all_gather_into_tensor(full[k], loc[k])

    def fuse(s):
        vol.integrate_keyframes(full["X"], full["C"], full["pose"], full["q"], bench.MIN_CONF,
                                bench.H, bench.W)

    def counters(s):
        vol.counters()

    def serial(s):
        h2d(s), gather(s), fuse(s), counters(s)

    def timed(fn):
        vol.reset()
        h2d(0), gather(0)  # every leg starts from valid buffers
        for s in range(warm):
            fn(s)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(warm, warm + steps):
            fn(s)
            torch.cuda.synchronize()  # each step starts from an idle GPU, like the e2e leg
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {"n_gpus": world, "keyframes_per_step": B, "share": share,
           "ms": {name: timed(fn) for name, fn in
                  (("h2d", h2d), ("all_gather_x4", gather), ("fuse", fuse),
                   ("counters", counters), ("serial_step", serial))}}
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/ingest_breakdown_n{world}.json", "w") as f:
            json.dump(out, f)
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
