"""Sharded query / pose linearisation across z-slab seams: every
rank reports the taps it owns (`b200tsdf_query_taps`), the records are summed
(what the all-reduce of `parallel.ShardedTSDFVolume` does) and the reference
arithmetic of TSDFVolume.query (global_volume.py:93-128) runs on the merged
taps.  The ranks are emulated on one GPU; the result must equal the
query of the single volume holding the same voxels BIT FOR BIT."""
import numpy as np
import pytest
import torch

from mast3r_slam_quality_dualtsdf_b200 import synth

pytestmark = pytest.mark.gpu

VOXEL, TRUNC, H, W = 0.02, 0.08, 48, 64


def _vol(dev, **kw):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    return TSDFVolume(VOXEL, TRUNC, device=dev, **kw)


def _fuse(vol, kf, dev):
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose")),
                            torch.from_numpy(kf["q"]).to(dev), 0.05, H, W)


def _world_points(kf, dev):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import sim3

    return torch.cat([sim3.act(torch.from_numpy(kf["pose"][b]).to(dev),
                               torch.from_numpy(kf["X"][b]).reshape(-1, 3).to(dev))
                      for b in range(kf["X"].shape[0])])


def _scene(dev, count, width, bounded):
    kf = synth.make_keyframes(synth.SCENES["sphere"], 2, H=H, W=W, seed=4, f=50.0, total=8)
    full = _vol(dev)
    _fuse(full, kf, dev)
    keys = full.export(0)[0]
    lo = (keys.min(0) // 8) * 8 - 8
    hi = keys.max(0) + 9
    bounds = (tuple(int(x) for x in lo), tuple(int(x) for x in hi)) if bounded else None
    ranks = [_vol(dev, bounds=bounds, slabs=(count, r, width)) for r in range(count)]
    # float32 atomic sums depend on the order the updates land in, so two fusions
    # of the same keyframes differ in the last bits; the single volume the
    # sharded query is held against is therefore the exact MERGE of the rank
    # volumes (additive block import, disjoint voxels) -- same bits, one hash
    merged = _vol(dev)
    for v in ranks:
        _fuse(v, kf, dev)
        merged.import_blocks(*v.export_block_data())
    assert merged.stats()["total_voxels"] == full.stats()["total_voxels"]
    full = merged
    # halo copies of foreign seam blocks (installed before meshing) must not be
    # reported a second time
    for r, v in enumerate(ranks):
        bk, vox, cnt = ranks[(r + 1) % count].export_seam_blocks()
        v.import_halo_blocks(bk, vox, cnt)
    return kf, full, ranks


def _probes(kf, dev, dtype):
    g = torch.Generator().manual_seed(3)
    pw = _world_points(kf, dev).cpu()
    pw = pw[torch.isfinite(pw).all(1)]
    sel = pw[torch.randperm(pw.shape[0], generator=g)[:3000]]
    jit = sel + (torch.rand(sel.shape, generator=g) - 0.5) * 0.06  # inside and outside the band
    far = torch.tensor([[50.0, 50.0, 50.0], [float("nan"), 0.0, 0.0], [0.0, 0.0, 0.0]])
    return torch.cat([sel, jit, far]).to(dtype).to(dev)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("count,width,bounded", [(2, 8, True), (3, 16, False), (4, 8, False),
                                                 (2, 64, True)])
def test_sharded_query_equals_single_volume(cuda_device, count, width, bounded, dtype):
    kf, full, ranks = _scene(cuda_device, count, width, bounded)
    pts = _probes(kf, cuda_device, dtype)
    for which in (0, 1):
        t, g, fl = full.query_batch(pts, which)
        taps = [v.query_taps(pts, which) for v in ranks]
        masks = torch.stack([tp[:, 14] for tp in taps])
        # ownership is disjoint: no tap is reported by two ranks
        m_int = masks.to(torch.int64)
        acc = torch.zeros_like(m_int[0])
        for m in m_int:
            assert not (acc & m).any()
            acc |= m
        if width == 8:  # thin slabs: some point's 7 voxels really are split between ranks
            assert ((m_int > 0).sum(0) >= 2).any()
        merged = torch.stack(taps).sum(0)
        for v in ranks:
            ts, gs, fs = v.query_finish(merged)
            assert torch.equal(fs, fl)
            assert torch.equal(ts, t) and torch.equal(gs, g)  # bit for bit
        assert int((fl == 3).sum()) > 500 and int((fl == 0).sum()) > 3


def test_single_volume_taps_round_trip(cuda_device):
    """slab_count = 1: taps + finish is the plain query (same code path)."""
    kf = synth.make_keyframes(synth.SCENES["plane"], 1, H=H, W=W, seed=2, f=50.0, total=8)
    vol = _vol(cuda_device)
    _fuse(vol, kf, cuda_device)
    pts = _probes(kf, cuda_device, torch.float32)
    t, g, fl = vol.query_batch(pts)
    ts, gs, fs = vol.query_finish(vol.query_taps(pts))
    assert torch.equal(fs, fl) and torch.equal(ts, t) and torch.equal(gs, g)
    e = vol.query_taps(pts[:0])
    assert e.shape == (0, 16) and vol.query_finish(e)[0].numel() == 0


def test_sharded_pose_normal_equations(cuda_device):
    """The pose optimiser's linearisation (tsdf_optimizer.py:72-116) from merged
    taps equals the single-volume fused kernel (float64 sums, warp order)."""
    count, width = 3, 8
    kf, full, ranks = _scene(cuda_device, count, width, False)
    g = torch.Generator().manual_seed(5)
    ok = np.nonzero(kf["C"][1].reshape(-1) > 0.05)[0]
    choice = ok[torch.randperm(len(ok), generator=g)[:2000].numpy()]
    pts = torch.from_numpy(kf["X"][1].reshape(-1, 3)[choice]).to(cuda_device)
    conf = torch.from_numpy(kf["C"][1].reshape(-1)[choice]).to(cuda_device, torch.float64)
    pose = kf["pose"][1].copy()
    pose[:3] += np.float32(0.004)  # a slightly wrong pose: non-zero residuals
    Hf, bf, uf = full.pose_normal_equations(pts, conf, pose, 0.5)
    merged = torch.stack([v.query_taps(pts, 0, pose) for v in ranks]).sum(0)
    for v in ranks:
        H, b, u = v.pose_normal_equations_taps(merged, pts, conf, pose, 0.5)
        assert u == uf and u > 200
        assert np.allclose(H, Hf, rtol=1e-11, atol=1e-13 * np.abs(Hf).max())
        assert np.allclose(b, bf, rtol=1e-11, atol=1e-13 * np.abs(bf).max())
    # and the acted points of the taps kernel are the ones the query sees
    from mast3r_slam_quality_dualtsdf_b200.tsdf.sim3 import act_kernel_order

    pw = act_kernel_order(torch.from_numpy(pose), pts)
    t, gr, fl = full.query_batch(pw)
    ts, gs, fs = ranks[0].query_finish(merged)
    assert torch.equal(fs, fl) and torch.equal(ts, t) and torch.equal(gs, gr)


@pytest.mark.parametrize("mode", ["replace", "additive"])
def test_store_and_pose_only_reintegration(cuda_device, mode):
    """SURVEY 8(e) config 5 on one rank (the NCCL run is
    scripts/sharded_reintegrate_check.py): keyframes kept in the store at ingest
    are fused again from their new poses alone.  replace == fused once at the
    new poses; additive == the reference's re-fusion on top
    (global_manager.py:71-79)."""
    from mast3r_slam_quality_dualtsdf_b200 import parallel
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=H, W=W, seed=8, f=50.0, total=9)
    t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    new_pose = t["pose"].clone()
    new_pose[:, :3] += torch.tensor([0.03, -0.02, 0.025], device=cuda_device)
    sv = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=8, device=cuda_device)
    sv.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W, ids=[4, 7, 9])
    assert sv.stored_ids() == [4, 7, 9]
    assert sv.reintegrate([9, 4], new_pose[[2, 0]].contiguous(), mode=mode) == 2
    want = TSDFVolume(VOXEL, TRUNC, device=cuda_device)
    if mode == "replace":
        pose = t["pose"].clone()
        pose[[2, 0]] = new_pose[[2, 0]]
        want.integrate_keyframes(t["X"], t["C"], pose, t["q"], 0.05, H, W)
    else:
        want.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)
        want.integrate_keyframes(t["X"][[2, 0]].contiguous(), t["C"][[2, 0]].contiguous(),
                                 new_pose[[2, 0]].contiguous(), t["q"][[2, 0]].contiguous(),
                                 0.05, H, W)
    for which in (0, 1):
        gk, gt, gw, gc = sv.vol.export(which)
        wk, wt, ww, wc = want.export(which)
        assert np.array_equal(gk, wk) and np.array_equal(gc.astype(np.int64), wc.astype(np.int64))
        assert np.abs(gt - wt).max() <= 5e-5  # float32 cancellation of the removed sums
        assert (np.abs(gw - ww) / np.maximum(1.0, ww)).max() <= 5e-5
    assert torch.equal(sv._store[9]["pose"], new_pose[2])
    assert torch.equal(sv._store[7]["pose"], t["pose"][1])
    with pytest.raises(KeyError):
        sv.reintegrate([1], new_pose[:1])


def test_sharded_checkpoint_round_trip(cuda_device, tmp_path):
    """save_checkpoint / load_checkpoint of the sharded wrapper on the real
    volume (one rank; the re-sharding logic is in the world-2 gloo test):
    restoring is bit-exact, a second load merges (sums and counters double)."""
    from mast3r_slam_quality_dualtsdf_b200 import parallel

    kf = synth.make_keyframes(synth.SCENES["sphere"], 2, H=H, W=W, seed=8, f=50.0, total=9)
    a = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=8, device=cuda_device)
    _fuse(a.vol, kf, cuda_device)
    n = a.save_checkpoint(str(tmp_path / "ck"))
    b = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=16, device=cuda_device)
    assert b.load_checkpoint(str(tmp_path / "ck")) == n > 0
    for which in (0, 1):
        for x, y in zip(a.vol.export(which), b.vol.export(which)):
            assert np.array_equal(x, y)
    assert b.load_checkpoint(str(tmp_path / "ck")) == n
    ka, ta, wa, ca = a.vol.export(0)
    kb, tb, wb, cb = b.vol.export(0)
    assert np.array_equal(ka, kb) and np.array_equal(2 * ca.astype(np.int64), cb.astype(np.int64))
    c = parallel.ShardedTSDFVolume(VOXEL * 2, TRUNC, slab_width=8, device=cuda_device)
    with pytest.raises(ValueError):
        c.load_checkpoint(str(tmp_path / "ck"))
