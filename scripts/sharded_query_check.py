"""Sharded query / pose linearisation over real NCCL ranks (one process per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
        --master-addr 127.0.0.1 --master-port 29611 scripts/sharded_query_check.py

Every rank fuses the same keyframes into its z-slabs (broadcast from rank 0),
then all ranks run `ShardedTSDFVolume.query_batch` / `pose_normal_equations`
(per-rank taps + ONE all-reduce).  Checks: all ranks get identical results; the
results equal, bit for bit, the query of a single volume holding the exact
merge of the rank volumes (block data gathered on rank 0); timing of the
collective query is printed.  Writes gpurun_out/sharded_query_n<N>.json."""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_quality_dualtsdf_b200 import parallel, synth  # noqa: E402
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume, sim3  # noqa: E402

VOXEL, TRUNC, H, W = 0.01, 0.04, 96, 128


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    kf = synth.make_keyframes(synth.SCENES["sphere"], 4, H=H, W=W, seed=4, f=100.0, total=8)
    sv = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=8, device=dev)
    X, Cf, P, Q = (torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q"))
    if rank != 0:  # only rank 0 holds the batch
        X, Cf, P, Q = (torch.zeros_like(t) for t in (X, Cf, P, Q))
    sv.integrate_keyframes(X, Cf, P, Q, 0.05, H, W, src=0)
    # probes: surface points of keyframe 1 and jittered copies (same on every rank)
    g = torch.Generator().manual_seed(3)
    pw = sim3.act(P[1], X[1].reshape(-1, 3)).cpu()
    sel = pw[torch.randperm(pw.shape[0], generator=g)[:4000]]
    pts = torch.cat([sel, sel + (torch.rand(sel.shape, generator=g) - 0.5) * 0.03]).to(dev)
    t, gr, fl = sv.query_batch(pts)
    # every rank holds the same answer
    sig = torch.stack([t.sum(), gr.sum(), fl.double().sum()])
    sigs = [torch.zeros_like(sig) for _ in range(world)]
    dist.all_gather(sigs, sig)
    same = all(torch.equal(s, sigs[0]) for s in sigs)
    # pose linearisation over the seams
    choice = torch.randperm(H * W, generator=g)[:2000]
    pc = X[1].reshape(-1, 3)[choice.to(dev)].contiguous()
    cf = Cf[1].reshape(-1)[choice.to(dev)].double()
    pose = P[1].cpu().numpy().copy()
    pose[:3] += np.float32(0.002)
    Hs, bs, us = sv.pose_normal_equations(pc, cf, pose, 0.5)
    # timing of the collective query (taps kernel + all-reduce + finish kernel)
    for _ in range(3):
        sv.query_batch(pts)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    e0.record()
    for _ in range(20):
        sv.query_batch(pts)
    e1.record()
    torch.cuda.synchronize()
    us_per_query = 1e3 * e0.elapsed_time(e1) / 20
    # exact merge of the rank volumes on rank 0
    bk, vox, cnt = sv.vol.export_block_data()
    n = torch.tensor([bk.shape[0]], dtype=torch.int64, device=dev)
    ns = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(ns, n)
    ns = [int(x.item()) for x in ns]
    pad = max(ns + [1])

    def gather(x):
        buf = torch.zeros((pad,) + tuple(x.shape[1:]), dtype=x.dtype, device=dev)
        buf[: x.shape[0]] = x
        parts = [torch.zeros_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf)
        return [p[:c] for p, c in zip(parts, ns)]

    bks, voxs, cnts = gather(bk), gather(vox), gather(cnt)
    ok = True
    out = None
    if rank == 0:
        full = TSDFVolume(VOXEL, TRUNC, device=dev)
        for a, b, c in zip(bks, voxs, cnts):
            full.import_blocks(a, b, c)
        tf, gf, ff = full.query_batch(pts)
        Hf, bf, uf = full.pose_normal_equations(pc, cf, pose, 0.5)
        exact = torch.equal(ff, fl) and torch.equal(tf, t) and torch.equal(gf, gr)
        pose_ok = us == uf and np.allclose(Hs, Hf, rtol=1e-11, atol=1e-13 * np.abs(Hf).max()) \
            and np.allclose(bs, bf, rtol=1e-11, atol=1e-13 * np.abs(bf).max())
        ok = exact and pose_ok and same
        out = {"n_gpus": world, "points": int(pts.shape[0]), "blocks_per_rank": ns,
               "with_gradient": int((fl == 3).sum()), "no_value": int((fl == 0).sum()),
               "ranks_agree": same, "bit_equal_to_merged_single_volume": exact,
               "pose_normal_equations_equal": pose_ok, "pose_points_used": us,
               "us_per_collective_query": us_per_query,
               "allreduce_bytes": int(pts.shape[0]) * 64}
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/sharded_query_n{world}.json", "w") as f:
            json.dump(out, f)
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
