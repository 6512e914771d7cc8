"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle on the same
seeded inputs.  Bit-exact: voxel key set, voxel-block key set, per-voxel update
counts, rays fused, update totals.  Tolerance: tsdf 1e-5 (normalised), weight
1e-5 relative."""
import numpy as np
import pytest
import torch

from mast3r_slam_quality_dualtsdf_b200 import synth
from oracle import oracle as orc
from tests.util import assert_volume_matches, block_set, oracle_dual

pytestmark = pytest.mark.gpu


def _vol(dev, voxel, trunc, **kw):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    return TSDFVolume(voxel, trunc, device=dev, **kw)


@pytest.mark.parametrize("scene,voxel", [("plane", 0.03), ("sphere", 0.03), ("room", 0.01),
                                         ("plane", 0.005), ("room_aligned", 0.005),
                                         ("room_aligned", 0.01)])
@pytest.mark.parametrize("n_kf", [1, 3])
def test_keyframes_match_oracle(cuda_device, scene, voxel, n_kf):
    trunc = 4 * voxel
    kf = synth.make_keyframes(synth.SCENES[scene], n_kf, H=48, W=64, seed=1, f=50.0, total=12)
    vp, vq, rays = oracle_dual(kf, voxel, trunc)
    vol = _vol(cuda_device, voxel, trunc)
    vol.integrate_keyframes(torch.from_numpy(kf["X"]).to(cuda_device),
                            torch.from_numpy(kf["C"]).to(cuda_device),
                            torch.from_numpy(kf["pose"]).to(cuda_device),
                            torch.from_numpy(kf["q"]).to(cuda_device), 0.05, 48, 64)
    c = vol.counters()
    assert c["rays"] == rays
    assert c["updates_primary"] == vp.events and c["updates_quality"] == vq.events
    assert c["dropped"] == 0
    assert_volume_matches(vol.export(0), vp.export(), "primary")
    assert_volume_matches(vol.export(1), vq.export(), "quality")
    assert np.array_equal(vol.block_keys(), block_set(vp.export()[0]))
    assert vol.stats(0) == vp.stats() and vol.stats(1) == vq.stats()


@pytest.mark.parametrize("n_kf,pinned", [(2, True), (9, True), (7, False), (1, True)])
def test_host_entry_matches_device_entry(cuda_device, n_kf, pinned):
    """b200tsdf_integrate_keyframes_host: the batch crosses PCIe in chunks through a ring
    of three staging slots (9 and 7 keyframes = 4 chunks of unequal size, the fourth reuses
    slot 0 behind the fuse of the first), quality grids and poses in one copy each; twice
    in a row on the same handle (the second call waits for the first call's kernels before
    it overwrites the slots).  Equal to the device entry fed the same tensors."""
    kf = synth.make_keyframes(synth.SCENES["sphere"], n_kf, H=48, W=64, seed=3, f=50.0,
                              total=max(n_kf, 8))
    a, b = _vol(cuda_device, 0.02, 0.08), _vol(cuda_device, 0.02, 0.08)
    t = {k: torch.from_numpy(kf[k]) for k in ("X", "C", "pose", "q")}
    th = {k: (v.pin_memory() if pinned else v) for k, v in t.items()}
    for _ in range(2):
        a.integrate_keyframes(th["X"], th["C"], th["pose"], th["q"], 0.05, 48, 64)
        b.integrate_keyframes(*(t[k].to(cuda_device) for k in ("X", "C", "pose")),
                              t["q"].to(cuda_device), 0.05, 48, 64)
    assert a.counters()["rays"] == b.counters()["rays"] > 0
    for which in (0, 1):
        assert_volume_matches(a.export(which), tuple(np.asarray(x, np.float64) if x.dtype.kind == "f"
                                                     else x for x in b.export(which)), "host")


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_integrate_points_reference_signature(cuda_device, dtype):
    """TSDFVolume.integrate(points_world, confidences, cam_origin) as the SLAM
    loop calls it (global_manager.py:106), float32 and float64 arithmetic."""
    rng = np.random.default_rng(7)
    pts = (rng.normal(size=(3000, 3)) * 0.4 + np.array([0.2, -0.1, 1.4])).astype(dtype)
    pts[5] = np.nan                  # non-finite ray: skipped (global_volume.py:56)
    pts[6] = [0.01, 0.02, 0.03]      # zero-length ray after origin subtraction
    org = np.array([0.01, 0.02, 0.03], dtype)
    conf = 0.5 + rng.random(3000)
    conf[10] = 0.0                   # weight <= 0: no voxel update but the ray counts
    conf[11] = -1.0
    vo = orc.OracleVolume(0.03, 0.12)
    n_ref = vo.integrate(pts, conf, org)
    vol = _vol(cuda_device, 0.03, 0.12)
    n = vol.integrate(pts, conf, org)
    assert n == n_ref
    assert vol.counters()["updates_primary"] == vo.events
    assert_volume_matches(vol.export(0), vo.export(), f"points {dtype.__name__}")
    assert vol.stats(1) == {"valid_voxels": 0, "total_voxels": 0}


def test_empty_and_all_filtered(cuda_device):
    vol = _vol(cuda_device, 0.03, 0.12)
    assert vol.integrate(np.zeros((0, 3), np.float32), np.zeros(0), np.zeros(3, np.float32)) == 0
    kf = synth.make_keyframes(synth.SCENES["plane"], 1, H=16, W=16, seed=0, f=20.0)
    vol.integrate_keyframes(torch.from_numpy(kf["X"]).to(cuda_device),
                            torch.zeros(1, 256, device=cuda_device),
                            torch.from_numpy(kf["pose"]).to(cuda_device), None, 0.05, 16, 16)
    assert vol.stats() == {"valid_voxels": 0, "total_voxels": 0}
    assert vol.counters()["rays"] == 0
    assert vol.query(np.zeros(3)) == (None, None)


def test_max_weight_clamp_and_order_quirk(cuda_device):
    """Reference quirk (global_volume.py:84-88): divide by the CLAMPED weight.
    Drive voxels far past max_weight and compare with the sequential oracle."""
    kf = synth.make_keyframes(synth.SCENES["plane"], 1, H=32, W=32, seed=5, f=40.0)
    kf["C"] *= 4.0
    vol = _vol(cuda_device, 0.03, 0.12, )
    vol2 = orc.OracleVolume(0.03, 0.12, 5.0)
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume
    vol = TSDFVolume(0.03, 0.12, 5.0, device=cuda_device)
    pts, conf, confq, org = orc.stage_keyframe(kf["X"][0], kf["C"][0], kf["pose"][0], None, 32, 32,
                                               0.05)
    for _ in range(6):
        vol2.integrate(pts, conf, org)
        vol.integrate(pts, conf, org)
    k, t, w, c = vol2.export()
    assert (w >= 5.0 - 1e-9).sum() > 100, "test must saturate many voxels"
    assert_volume_matches(vol.export(0), (k, t, w, c), "clamped")


def test_query_matches_oracle(cuda_device):
    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=48, W=64, seed=2, f=50.0, total=12)
    vp, vq, _ = oracle_dual(kf, 0.03, 0.12)
    vol = _vol(cuda_device, 0.03, 0.12)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            torch.from_numpy(kf["q"]).to(cuda_device), 0.05, 48, 64)
    keys = vp.export()[0]
    rng = np.random.default_rng(0)
    probe = (keys[rng.integers(0, len(keys), 4000)] + rng.random((4000, 3))) * 0.03
    probe = np.concatenate([probe, rng.normal(size=(500, 3)) * 3])
    for dtype in (np.float32, np.float64):
        p = probe.astype(dtype)
        for which, vo in ((0, vp), (1, vq)):
            t, g, fl = vo.query_batch(p)
            tt, gg, ff = (x.cpu().numpy() for x in vol.query_batch(p, which))
            assert np.array_equal(fl & 1, ff & 1)
            # gradient presence can flip only where tsdf differences sit at the
            # float32 rounding-noise level (reference: exact zeros)
            assert ((fl ^ ff) & 2).astype(bool).mean() <= 2e-3
            assert np.abs(t - tt).max() <= 1e-5
            both = (fl == 3) & (ff == 3)
            # unit gradients of tsdf differences: f32 storage => looser, still tight
            assert np.abs(g - gg)[both].max() <= 2e-3
    tv, gv = vol.query(probe[0])
    tr, gr = vp.query(probe[0])
    assert abs(tv - tr) <= 1e-5 and (gv is None) == (gr is None)
    # a non-finite coordinate in ANY axis: the reference returns (None, None)
    # (floor(nan) is no voxel's key) -- also for points whose other coordinates
    # would land in a stored voxel
    bad = np.repeat(probe[:6], 1, axis=0).copy()
    bad[0, 0] = np.nan
    bad[1, 1] = np.nan
    bad[2, 2] = np.nan
    bad[3, 1] = np.inf
    bad[4, 2] = -np.inf
    for dtype in (np.float32, np.float64):
        _, _, ff = (x.cpu().numpy() for x in vol.query_batch(bad.astype(dtype), 0))
        assert not ff[:5].any() and ff[5] & 1
        assert vol.query(bad[1].astype(dtype)) == (None, None)


def test_bounds_drop_outside(cuda_device):
    kf = synth.make_keyframes(synth.SCENES["plane"], 1, H=48, W=64, seed=1, f=50.0)
    vp, _, _ = oracle_dual(kf, 0.03, 0.12)
    keys, t, w, c = vp.export()
    lo, hi = keys.min(0), keys.max(0) + 1
    mid = int((lo[2] + hi[2]) // 2)
    parts = []
    for z0, z1 in ((int(lo[2]), mid), (mid, int(hi[2]))):
        vol = _vol(cuda_device, 0.03, 0.12, bounds=((int(lo[0]), int(lo[1]), z0),
                                                    (int(hi[0]), int(hi[1]), z1)))
        vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device)
                                  for k in ("X", "C", "pose")),
                                torch.from_numpy(kf["q"]).to(cuda_device), 0.05, 48, 64)
        parts.append(vol.export(0))
    gk = np.concatenate([p[0] for p in parts])
    order = np.lexsort((gk[:, 2], gk[:, 1], gk[:, 0]))
    got = tuple(np.concatenate([p[i] for p in parts])[order] for i in range(4))
    assert_volume_matches(got, (keys, t, w, c), "z-slab union")


@pytest.mark.parametrize("bounded", [True, False])
@pytest.mark.parametrize("count,width", [(2, 8), (4, 8), (3, 16), (2, 64)])
def test_interleaved_z_slabs_union_is_exact(cuda_device, bounded, count, width):
    """every rank fuses only its own z-slabs; the union of the
    per-rank voxel sets equals the single-volume result bit-exactly, applied
    update counters add up, and rays whose band misses a rank are culled."""
    kf = synth.make_keyframes(synth.SCENES["sphere"], 2, H=48, W=64, seed=4, f=50.0, total=8)
    vp, vq, rays = oracle_dual(kf, 0.02, 0.08)
    keys = vp.export()[0]
    lo = (keys.min(0) // 8) * 8 - 8
    hi = keys.max(0) + 9
    bounds = (tuple(int(x) for x in lo), tuple(int(x) for x in hi)) if bounded else None
    parts, partsq, ev_p, ev_q, culled = [], [], 0, 0, 0
    for r in range(count):
        vol = _vol(cuda_device, 0.02, 0.08, bounds=bounds, slabs=(count, r, width))
        vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device)
                                  for k in ("X", "C", "pose")),
                                torch.from_numpy(kf["q"]).to(cuda_device), 0.05, 48, 64)
        c = vol.counters()
        assert c["rays"] == rays
        ev_p += c["updates_primary"]
        ev_q += c["updates_quality"]
        culled += c["culled_rays"]
        parts.append(vol.export(0))
        partsq.append(vol.export(1))
        z0 = bounds[0][2] if bounded else 0
        if len(parts[-1][0]):
            assert np.all(((parts[-1][0][:, 2] - z0) // width) % count == r)
    assert ev_p == vp.events and ev_q == vq.events
    if width >= 64:
        assert culled > 0
    for ps, want in ((parts, vp), (partsq, vq)):
        gk = np.concatenate([p[0] for p in ps])
        order = np.lexsort((gk[:, 2], gk[:, 1], gk[:, 0]))
        got = tuple(np.concatenate([p[i] for p in ps])[order] for i in range(4))
        assert_volume_matches(got, want.export(), "interleaved z-slab union")


@pytest.mark.parametrize("count", [3, 4])
def test_slab_classify_path_ragged_and_degenerate(cuda_device, count):
    """3+ ranks run the classify pass + listed-tile fuse: ragged image (no TMA,
    partial tiles), zero-length and non-finite rays, a far-away keyframe whose
    tiles are all foreign for some ranks.  Rays are counted on every rank
    exactly as on one GPU and the union stays bit-exact."""
    H, W = 45, 60
    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=H, W=W, seed=11, f=50.0, total=8,
                              gh=9, gw=12)
    X = kf["X"].reshape(3, H * W, 3)
    Cf = kf["C"].reshape(3, H * W)
    X[0, np.nonzero(Cf[0] > 0.05)[0][:30]] = 0.0   # point = camera centre: len 0, skipped (:56)
    X[1, np.nonzero(Cf[1] > 0.05)[0][10:40], 1] = np.nan  # non-finite length: skipped
    X[2] *= 0.25                # a keyframe that only reaches a few slabs
    kf["X"] = X.reshape(kf["X"].shape)
    vp, vq, rays = oracle_dual(kf, 0.02, 0.08)
    assert rays < int((kf["C"] > 0.05).sum())
    parts, ev_p, culled = [], 0, 0
    for r in range(count):
        vol = _vol(cuda_device, 0.02, 0.08, slabs=(count, r, 8))
        vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device)
                                  for k in ("X", "C", "pose")),
                                torch.from_numpy(kf["q"]).to(cuda_device), 0.05, H, W)
        c = vol.counters()
        assert c["rays"] == rays and c["culled_rays"] <= rays
        ev_p += c["updates_primary"]
        culled += c["culled_rays"]
        parts.append(vol.export(0))
    assert ev_p == vp.events and culled > 0
    gk = np.concatenate([p[0] for p in parts])
    order = np.lexsort((gk[:, 2], gk[:, 1], gk[:, 0]))
    got = tuple(np.concatenate([p[i] for p in parts])[order] for i in range(4))
    assert_volume_matches(got, vp.export(), "classify-path union")


@pytest.mark.parametrize("case", ["plane_v30_f32", "sphere_v30_f32", "sphere_v30_f64",
                                  "room_v10_f32", "plane_clamp_f32"])
def test_against_reference_golden(cuda_device, case):
    """CUDA path vs vectors frozen from the UNMODIFIED reference TSDFVolume."""
    import os

    from tests.golden import make_golden as mg

    g = np.load(os.path.join(mg.HERE, case + ".npz"))
    scene, n_kf, H, W, f, voxel, trunc, mw, dtype, seed = mg.CASES[case]
    kf, staged = mg.staged_inputs(case)
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    vol = TSDFVolume(voxel, trunc, mw, device=cuda_device)
    fused = sum(vol.integrate(s[0], s[1], s[3], confidences_q=s[2]) for s in staged)
    assert fused == int(g["p_fused"])
    c = vol.counters()
    assert c["updates_primary"] == int(g["p_events"]) and c["updates_quality"] == int(g["q_events"])
    for which, name in ((0, "p"), (1, "q")):
        want = (g[f"{name}_keys"].astype(np.int64), g[f"{name}_tsdf"], g[f"{name}_w"],
                g[f"{name}_count"])
        assert_volume_matches(vol.export(which), want, f"{case}/{name}")
        st = vol.stats(which)
        assert [st["valid_voxels"], st["total_voxels"]] == list(g[f"{name}_stats"])
        t, gr, fl = (x.cpu().numpy() for x in vol.query_batch(g[f"{name}_probe"], which))
        assert np.array_equal(fl & 1, g[f"{name}_qfl"] & 1)
        has = (fl & 1) > 0
        assert np.abs(t[has] - g[f"{name}_qt"][has]).max() <= 1e-5


def test_peer_keyframes_ring_one_rank(cuda_device):
    """parallel.PeerKeyframes / ShardedTSDFVolume.integrate_slot with one rank:
    the ring buffer is libb200tsdf's own allocation seen by torch through
    __cuda_array_interface__; the pull path (meta record -> tile list -> fuse
    through per-segment pointers) equals the plain batch fuse bit for bit in
    keys and counts."""
    from mast3r_slam_quality_dualtsdf_b200 import parallel
    from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import pack_keyframes

    kf = synth.make_keyframes(synth.SCENES["sphere"], 4, H=48, W=64, seed=2, f=50.0, total=12)
    bounds = ((-128, -128, -64), (128, 128, 192))
    peers = parallel.PeerKeyframes(2, 2, 48, 64, 3, 4, device=cuda_device, local_only=True)
    for s in range(2):
        sl = slice(2 * s, 2 * s + 2)
        pack_keyframes(kf["X"][sl], kf["C"][sl], kf["pose"][sl], kf["q"][sl], out=peers.slot(s))
    X, Cf, q, P = peers.views(1)
    assert np.array_equal(X.cpu().numpy(), kf["X"][2:4]) and np.array_equal(q.cpu().numpy(), kf["q"][2:4])
    sv = parallel.ShardedTSDFVolume(0.03, 0.12, slab_width=8, bounds=bounds, device=cuda_device)
    for s in range(2):
        sv.integrate_slot(peers, s, 0.05)
    vp, vq, rays = oracle_dual(kf, 0.03, 0.12)
    assert sv.vol.counters()["rays"] == rays
    assert_volume_matches(sv.vol.export(0), vp.export(), "ring primary")
    assert_volume_matches(sv.vol.export(1), vq.export(), "ring quality")
    # the software-pipelined run of slots (meta + gather of slot j+1 on a side stream
    # behind the fuse of slot j) is the same fusion
    sv2 = parallel.ShardedTSDFVolume(0.03, 0.12, slab_width=8, bounds=bounds, device=cuda_device)
    ready = [torch.cuda.Event() for _ in range(2)]
    for e in ready:
        e.record()
    sv2.integrate_slots(peers, [0, 1], 0.05, ready=ready)
    bufs = sv2.integrate_slots(peers, [], 0.05)
    assert bufs is None and sv2.vol.counters()["rays"] == rays
    assert_volume_matches(sv2.vol.export(0), vp.export(), "pipelined ring primary")
    assert_volume_matches(sv2.vol.export(1), vq.export(), "pipelined ring quality")
    peers.close()
