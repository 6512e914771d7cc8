"""Host logic of parallel.ShardedTSDFVolume on CPU: world_size-2 gloo group,
the CUDA volume replaced by a numpy stand-in through `volume_factory` (the
kernels themselves are covered by the gpu-marked slab / seam tests)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mast3r_slam_quality_dualtsdf_b200 import parallel


class FakeVolume:
    """Records what the sharded wrapper asks of a volume; 'fuses' one voxel
    per pixel at kz = round(100 * X.z) so ownership can be checked."""

    def __init__(self, capacity_log2, bounds, device, slabs):
        self.world, self.rank, self.width = slabs
        self.z0 = bounds[0][2] if bounds is not None else 0
        self.blocks = {}  # (bx,by,bz) -> payload value
        self.halo = {}
        self.rays = 0
        self.applied = 0
        self.culled = 0

    def integrate_keyframes(self, X, Cf, poses, q_grid, min_conf, H, W, step_scale):
        kz = torch.round(X[..., 2] * 100).to(torch.int64).reshape(-1)
        self.rays += kz.numel()
        own = parallel.slab_owner(kz, self.z0, self.width, self.world) == self.rank
        self.applied += int(own.sum())
        self.culled += int((~own).sum())
        for z in kz[own].tolist():
            key = (0, 0, z >> 3)
            self.blocks[key] = self.blocks.get(key, 0.0) + float(poses.sum())

    def deintegrate_keyframes(self, X, Cf, poses, q_grid, min_conf, H, W, step_scale):
        kz = torch.round(X[..., 2] * 100).to(torch.int64).reshape(-1)
        own = parallel.slab_owner(kz, self.z0, self.width, self.world) == self.rank
        self.removed = getattr(self, "removed", 0) + int(own.sum())
        for z in kz[own].tolist():
            self.blocks[(0, 0, z >> 3)] -= float(poses.sum())

    def counters(self):
        return {"rays": self.rays, "updates_primary": self.applied, "updates_quality": self.applied,
                "blocks": len(self.blocks), "dropped": 0, "overflow": 0,
                "culled_rays": self.culled}

    def stats(self, which=0):
        return {"valid_voxels": len(self.blocks), "total_voxels": len(self.blocks)}

    def export_seam_blocks(self):
        keys = [k for k in sorted(self.blocks) if (k[2] * 8 - self.z0) % self.width == 0]
        bk = torch.tensor(keys, dtype=torch.int64).reshape(-1, 3)
        vox = torch.zeros(len(keys), 512, 4)
        for i, k in enumerate(keys):
            vox[i] += self.blocks[k]
        return bk, vox, torch.full((len(keys), 512), self.rank, dtype=torch.int64)

    def import_halo_blocks(self, bk, vox, cnt):
        for i in range(bk.shape[0]):
            self.halo[tuple(int(x) for x in bk[i])] = (float(vox[i, 0, 0]), int(cnt[i, 0]))

    def extract_mesh(self, which=0):
        return torch.full((3 * (self.rank + 1), 3), float(self.rank))

    def save_checkpoint(self, path):
        keys = sorted(self.blocks)
        np.savez(path, block_keys=np.array(keys, np.int64).reshape(-1, 3),
                 vox=np.array([self.blocks[k] for k in keys], np.float32).reshape(-1, 1),
                 cnt=np.ones((len(keys), 1), np.int64), params=np.zeros(4))
        return len(keys)

    def import_blocks(self, bk, vox, cnt):
        for i in range(bk.shape[0]):
            k = tuple(int(x) for x in bk[i])
            self.blocks[k] = self.blocks.get(k, 0.0) + float(vox[i, 0])

    # sharded query stand-ins: point i is "owned" by rank i % world
    def query_taps(self, points, which=0, pose=None):
        pts = torch.as_tensor(points, dtype=torch.float32).reshape(-1, 3)
        taps = torch.zeros(pts.shape[0], 16)
        mine = torch.arange(pts.shape[0]) % self.world == self.rank
        taps[mine, 0] = pts[mine, 2] + (0.0 if pose is None else float(pose[0]))
        taps[mine, 14] = 1.0
        return taps

    def query_finish(self, taps):
        return taps[:, 0].double(), torch.zeros(taps.shape[0], 3, dtype=torch.float64), \
            taps[:, 14].to(torch.uint8)

    def pose_normal_equations_taps(self, taps, pts_cam, conf, pose, lam):
        return np.eye(7) * float(taps[:, 0].sum()), np.zeros(7), int(taps[:, 14].sum())


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out, ckpt_dir=None):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bounds = ((-64, -64, -64), (64, 64, 64))
        sv = parallel.ShardedTSDFVolume(0.01, 0.04, slab_width=16, bounds=bounds, device="cpu",
                                        volume_factory=lambda **kw: FakeVolume(**kw))
        assert sv.vol.world == world and sv.vol.rank == rank
        # rank 0 holds the keyframes; the others pass empty buffers of the same shape
        g = torch.Generator().manual_seed(7)
        X = torch.rand(2, 48, 3, generator=g) * 1.2 - 0.6
        Cf, P = torch.ones(2, 48), torch.arange(16.0).reshape(2, 8)
        if rank != 0:
            X, Cf, P = torch.zeros_like(X), torch.zeros_like(Cf), torch.zeros_like(P)
        sv.integrate_keyframes(X, Cf, P, None, 0.05, 6, 8)
        g = torch.Generator().manual_seed(7)
        want = torch.rand(2, 48, 3, generator=g) * 1.2 - 0.6
        assert torch.equal(X, want) and float(P.sum()) == 120.0  # broadcast delivered
        c = sv.counters()
        assert c["rays"] == 96 and c["updates_primary"] == 96  # every voxel fused exactly once
        assert c["culled_rays"] == 96 * (world - 1)
        # sharded ingest: every rank contributes one keyframe, all get both in rank order
        gx, gc, gp, gq = sv.gather_keyframes(want[rank:rank + 1].clone(), Cf[:1] * (rank + 1),
                                             P[rank:rank + 1].clone())
        assert torch.equal(gx, want) and gq is None and torch.equal(gp, P)
        assert gc[:, 0].tolist() == [1.0, 2.0]
        # the same ingest from host shares in one call (pipelined on NCCL, serial here)
        before = sv.counters()["rays"]
        sv.integrate_keyframes_from_host(want[rank:rank + 1].clone(), Cf[:1].clone(),
                                         P[rank:rank + 1].clone(), None, 0.05, 6, 8)
        assert sv.counters()["rays"] == before + 96
        # sharded query: every rank ends up with every point's merged record
        t, _, fl = sv.query_batch(want[0])
        assert fl.tolist() == [1] * 48 and torch.equal(t.float(), want[0][:, 2])
        tv, gv = sv.query(want[0][5].numpy())
        assert tv == float(want[0][5, 2]) and gv is None
        Hq, _, used = sv.pose_normal_equations(want[0], torch.ones(48), [2.0] + [0.0] * 7, 0.5)
        assert used == 48 and abs(Hq[0, 0] - float(want[0][:, 2].sum() + 96.0)) < 1e-3
        # keyframe store + pose-only re-integration (config 5): rank 0 alone knows the new poses
        def fake(**kw):
            return FakeVolume(**kw)
        sv2 = parallel.ShardedTSDFVolume(0.01, 0.04, slab_width=16, bounds=bounds, device="cpu",
                                         volume_factory=fake)
        X2, C2, P2 = want.clone(), torch.ones(2, 48), torch.arange(16.0).reshape(2, 8)
        if rank != 0:
            X2, C2, P2 = torch.zeros_like(X2), torch.zeros_like(C2), torch.zeros_like(P2)
        sv2.integrate_keyframes(X2, C2, P2, None, 0.05, 6, 8, ids=[10, 11])
        assert sv2.stored_ids() == [10, 11]
        new_p = torch.arange(16.0).reshape(2, 8) * 2 + 1 if rank == 0 else torch.zeros(2, 8)
        assert sv2.reintegrate([11, 10], new_p, src=0, mode="replace") == 2
        want_p = torch.arange(16.0).reshape(2, 8) * 2 + 1   # rows follow ids [11, 10]
        assert torch.equal(sv2._store[11]["pose"], want_p[0])
        assert torch.equal(sv2._store[10]["pose"], want_p[1])
        fresh = FakeVolume(17, bounds, "cpu", (world, rank, 16))
        fresh.integrate_keyframes(want, C2, want_p, None, 0.05, 6, 8, 0.5)
        assert sv2.vol.blocks.keys() == fresh.blocks.keys()
        assert all(abs(sv2.vol.blocks[k] - fresh.blocks[k]) < 1e-3 for k in fresh.blocks)
        assert sv2.vol.removed == sv2.vol.applied // 2   # every applied update taken out once
        assert sv2.reintegrate([10], want_p[1:], mode="additive") == 1   # reference semantics
        assert sv2.vol.rays == 96 * 2 + 48
        # the reference's hook: pin filter, dedup, unknown indices skipped
        rays0 = sv2.vol.rays
        hook_p = torch.zeros(5, 8) if rank else torch.arange(40.0).reshape(5, 8)
        assert sv2.on_after_backend_solve([0, 11, 11, 42, 10], hook_p, pin=1) == [11, 10]
        assert sv2.vol.rays == rays0 + 96
        assert torch.equal(sv2._store[11]["pose"], torch.arange(8.0) + 8)
        assert torch.equal(sv2._store[10]["pose"], torch.arange(8.0) + 32)
        assert sv2.on_after_backend_solve([0], hook_p[:1], pin=1) == []
        assert sv2.on_after_backend_solve([10, 11], hook_p[:2], pin=0, queue_cap=1) == [10]
        assert sv2.forget([10]) == 4 * (48 * 3 + 48 + 8) and sv2.stored_ids() == [11]
        assert sv2.on_after_backend_solve([10, 11], hook_p[:2], pin=0) == [11]
        try:
            sv2.reintegrate([99], want_p[:1])
            raise AssertionError("unknown keyframe id must raise")
        except KeyError:
            pass
        # checkpoint: every rank writes its slabs; reload onto another slab width
        if ckpt_dir is not None:
            wrote = sv.save_checkpoint(ckpt_dir)
            assert wrote == len(sv.vol.blocks)
            sv3 = parallel.ShardedTSDFVolume(0.01, 0.04, slab_width=32, bounds=bounds,
                                             device="cpu", volume_factory=fake)
            got = sv3.load_checkpoint(ckpt_dir)
            assert got == len(sv3.vol.blocks)
            assert all(parallel.slab_owner(k[2] * 8, -64, 32, world) == rank
                       for k in sv3.vol.blocks)
            tot = torch.tensor([got, wrote])
            dist.all_reduce(tot)
            assert tot[0] == tot[1]  # re-sharded, nothing lost or duplicated
        st = sv.stats()
        n_halo = sv.exchange_seams()
        # what I received is exactly what the rank above exported
        up = (rank + 1) % world
        assert all(v[1] == up for v in sv.vol.halo.values()) and len(sv.vol.halo) == n_halo
        for k in sv.vol.halo:
            assert parallel.slab_owner(k[2] * 8, -64, 16, world) == up
            assert (k[2] * 8 + 64) % 16 == 0
        mesh = sv.extract_mesh(gather=True)
        out[rank] = (st["total_voxels"], len(sv.vol.blocks), n_halo, tuple(mesh.shape),
                     sorted(sv.vol.halo))
    finally:
        dist.destroy_process_group()


def test_sharded_volume_host_logic_world2(tmp_path):
    world = 2
    ckpt = str(tmp_path / "ckpt")
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), out, ckpt), nprocs=world, join=True)
        res = dict(out)
    # resume on ONE rank (no process group): all blocks of both files, values kept
    one = parallel.ShardedTSDFVolume(0.01, 0.04, slab_width=16,
                                     bounds=((-64, -64, -64), (64, 64, 64)), device="cpu",
                                     volume_factory=lambda **kw: FakeVolume(**kw))
    assert one.load_checkpoint(ckpt) == res[0][1] + res[1][1] == len(one.vol.blocks)
    assert one.load_checkpoint(ckpt) == len(one.vol.blocks)  # additive: a second load merges
    with np.load(ckpt + "/slabs_rank000.npz") as z:
        k0 = tuple(int(x) for x in z["block_keys"][0])
        assert abs(one.vol.blocks[k0] - 2 * float(z["vox"][0, 0])) < 1e-4
    assert res[0][0] == res[1][0] == res[0][1] + res[1][1]  # stats are all-reduced
    assert res[0][3] == (3 * 1 + 3 * 2, 3) and res[1][3] == (6, 3)  # rank 0 gathers the mesh
    assert res[0][2] + res[1][2] > 0


def test_slab_owner_matches_kernel_convention():
    kz = np.arange(-70, 70)
    own = parallel.slab_owner(kz, -64, 16, 4)
    assert own[kz == -64][0] == 0 and own[kz == -49][0] == 0 and own[kz == -48][0] == 1
    assert own[kz == -65][0] == 3  # floor division below the origin
    assert set(np.unique(own)) == {0, 1, 2, 3}
