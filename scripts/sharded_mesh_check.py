"""Seam-stitched marching cubes over real NCCL ranks (north_star
"meshes are stitched at slab seams"; BASELINE configs[2] + [3]):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port 29641 scripts/sharded_mesh_check.py

Every rank fuses the same synthetic keyframes into its interleaved z-slabs
(broadcast from rank 0), the ranks exchange their seam block layers over NCCL
(`exchange_seams`: isend / irecv of the packed blocks) and extract the mesh of
BOTH volumes; rank 0 gathers the per-rank soups and compares the triangle
multiset -- bit for bit -- with the mesh of ONE volume holding exactly the
ranks' voxel data (their exported blocks merged by `import_blocks`; a second
fusion would differ in the last bits, float32 reductions are unordered), and
welds the gathered soup on the device (seam vertices must merge: same vertex
count as the single-volume indexed mesh).  The merged volume is also compared
with a volume fused directly on one GPU (keys and hit counts exact).
Writes gpurun_out/sharded_mesh_n<N>.json."""
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
from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import weld_mesh  # noqa: E402

VOXEL, TRUNC, H, W, DIM, N_KF = 0.01, 0.04, 96, 128, 512, 8


def canon(tri):
    t = np.ascontiguousarray(tri.reshape(-1, 9))
    return t[np.lexsort(t.T[::-1])]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    scene = synth.SCENES["room"]
    kf = synth.make_keyframes(scene, N_KF, H=H, W=W, seed=5, f=100.0, orbit_radius=0.6,
                              total=24, up_axis="z")
    lo, hi = synth.bounds_for(scene, VOXEL, DIM)
    t = {k: torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")}
    width = max(8, DIM // (4 * world))
    sv = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=width, bounds=(lo, hi),
                                    capacity_log2=16, device=dev)
    src = {k: (v.clone() if rank == 0 else torch.zeros_like(v)) for k, v in t.items()}
    sv.integrate_keyframes(src["X"], src["C"], src["pose"], src["q"], 0.05, H, W, src=0)
    res = {"world": world, "slab_width": width, "volumes": {}}
    ok = True
    # the ranks' voxel blocks, merged on rank 0 into ONE volume
    bk, vox, cnt = sv.vol.export_block_data()
    parts = [None] * world
    dist.all_gather_object(parts, (bk.cpu().numpy(), vox.cpu().numpy(), cnt.cpu().numpy()))
    one = None
    if rank == 0:
        one = TSDFVolume(VOXEL, TRUNC, capacity_log2=16, bounds=(lo, hi), device=dev)
        for pb, pv, pc in parts:
            one.import_blocks(torch.from_numpy(pb), torch.from_numpy(pv), torch.from_numpy(pc))
        direct = TSDFVolume(VOXEL, TRUNC, capacity_log2=16, bounds=(lo, hi), device=dev)
        direct.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)
        a, b = one.export(0), direct.export(0)
        same_keys = a[0].shape == b[0].shape and np.array_equal(a[0], b[0]) and \
            np.array_equal(a[3], b[3])
        res["union_equals_direct_fusion"] = {
            "keys_and_hit_counts_exact": bool(same_keys),
            "max_abs_dtsdf": float(np.abs(a[1] - b[1]).max()) if same_keys else None}
        ok = ok and same_keys
    for which, name in ((0, "primary"), (1, "quality")):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        soup = sv.extract_mesh(which, gather=True)  # incl. exchange_seams
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if rank == 0:
            want = one.extract_mesh(which)
            a, b = canon(soup.cpu().numpy()), canon(want.cpu().numpy())
            same = a.shape == b.shape and np.array_equal(a, b)
            wv, wf = weld_mesh(soup)
            iv, _ = one.extract_mesh_indexed(which)
            welded = int(wv.shape[0]) == int(iv.shape[0])
            res["volumes"][name] = {"triangles": int(soup.shape[0] // 3),
                                    "triangles_single_volume": int(want.shape[0] // 3),
                                    "multiset_bit_equal": bool(same),
                                    "welded_vertices": int(wv.shape[0]),
                                    "single_volume_vertices": int(iv.shape[0]),
                                    "seam_vertices_merge": bool(welded),
                                    "seconds_incl_seam_exchange_and_gather": dt}
            ok = ok and same and welded and soup.shape[0] > 3000
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    if rank == 0:
        res["ok"] = bool(ok)
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/sharded_mesh_n{world}.json", "w") as f:
            json.dump(res, f, indent=1)
        print(json.dumps(res))
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
