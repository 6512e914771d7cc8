"""GPU: size-independent properties at BASELINE.json's full size (512x384
keyframes, 512^3 @ 5 mm, dual volume) where the sequential oracle would take
hours: checksum of hit counters == applied-update counters, batch == one by
one, keyframe-order invariance, additive re-fusion doubles every counter and
keeps tsdf, oracle parity on a sub-sampled keyframe through the points entry."""
import numpy as np
import pytest
import torch

from mast3r_slam_quality_dualtsdf_b200 import synth
from oracle import oracle as orc
from tests.util import TSDF_TOL, W_TOL, assert_volume_matches

pytestmark = pytest.mark.gpu
H, W, VOXEL, TRUNC = 384, 512, 0.005, 0.02


def _vol(dev):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    scene = synth.SCENES["room"]
    lo, hi = synth.bounds_for(scene, VOXEL, 512)
    return TSDFVolume(VOXEL, TRUNC, capacity_log2=17, bounds=(lo, hi), device=dev)


@pytest.fixture(scope="module")
def kf():
    return synth.make_keyframes(synth.SCENES["room"], 4, H, W, seed=0, orbit_radius=0.6, gh=24,
                                gw=32, start=100, total=600)


def _fuse(vol, kf, idx, dev):
    t = {k: torch.from_numpy(np.ascontiguousarray(kf[k][idx])).to(dev) for k in
         ("X", "C", "pose", "q")}
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)


def test_full_size_properties(cuda_device, kf):
    a = _vol(cuda_device)
    _fuse(a, kf, [0, 1, 2, 3], cuda_device)
    ca = a.counters()
    assert ca["rays"] == int((kf["C"] > 0.05).sum()) and ca["dropped"] == 0
    ka, ta, wa, na = a.export(0)
    kq, tq, wq, nq = a.export(1)
    # checksum of checksums: per-voxel hit counters add up to the applied updates
    assert int(na.astype(np.int64).sum()) == ca["updates_primary"]
    assert int(nq.astype(np.int64).sum()) == ca["updates_quality"]
    assert ca["updates_primary"] > 10_000_000 and ca["updates_quality"] <= ca["updates_primary"]
    assert len(np.unique(ka >> 3, axis=0)) == ca["blocks"] == len(a.block_keys())
    assert np.all(np.abs(ta) <= 1.0 + 1e-6) and np.all(wa > 0)
    st = a.stats(0)
    assert st["total_voxels"] == len(ka) and st["valid_voxels"] == int((wa >= 1e-3).sum())
    # one launch of 4 keyframes == 4 launches of one, in reverse order
    b = _vol(cuda_device)
    for i in (3, 2, 1, 0):
        _fuse(b, kf, [i], cuda_device)
    assert b.counters()["updates_primary"] == ca["updates_primary"]
    assert_volume_matches(b.export(0), (ka, ta.astype(np.float64), wa.astype(np.float64), na),
                          "batch vs reversed singles (primary)")
    assert_volume_matches(b.export(1), (kq, tq.astype(np.float64), wq.astype(np.float64), nq),
                          "batch vs reversed singles (quality)")
    # additive re-fusion (reference semantics): counters double, tsdf unchanged,
    # weight doubles up to the max_weight clamp
    _fuse(a, kf, [0, 1, 2, 3], cuda_device)
    k2, t2, w2, n2 = a.export(0)
    assert np.array_equal(k2, ka) and np.array_equal(n2, 2 * na)
    wa64, ta64 = wa.astype(np.float64), ta.astype(np.float64)
    free = 2.0 * wa64 <= 100.0
    assert free.sum() > 0.5 * len(free)
    assert np.abs(t2[free] - ta64[free]).max() <= TSDF_TOL  # same mean below the clamp
    # reference quirk: past max_weight the sum is divided by the CLAMPED weight
    assert np.abs(t2[~free] - ta64[~free] * 2.0 * wa64[~free] / 100.0).max(initial=0.0) <= 1e-4
    assert np.allclose(w2, np.minimum(2.0 * wa64, 100.0), rtol=1e-5)


def test_full_size_subsample_matches_oracle(cuda_device, kf):
    """Reference flow with its 40 000-point cap (global_manager.py:94-99): a
    fixed sub-sample of one full-size keyframe through TSDFVolume.integrate
    against the sequential oracle (bit-exact keys and counts)."""
    pts, conf, confq, org = orc.stage_keyframe(kf["X"][0], kf["C"][0], kf["pose"][0], kf["q"][0],
                                               H, W, 0.05)
    sel = np.random.default_rng(5).permutation(len(pts))[:40000]
    vol = _vol(cuda_device)
    fused = vol.integrate(pts[sel], conf[sel], org, confidences_q=confq[sel])
    op, oq = orc.OracleVolume(VOXEL, TRUNC), orc.OracleVolume(VOXEL, TRUNC)
    assert fused == op.integrate(pts[sel], conf[sel], org)
    oq.integrate(pts[sel], confq[sel], org)
    assert_volume_matches(vol.export(0), op.export(), "40k sub-sample primary")
    assert_volume_matches(vol.export(1), oq.export(), "40k sub-sample quality")


def test_full_size_keyframe_cap_matches_oracle(cuda_device, kf):
    """config/base.yaml:116 through the keyframe entry: max_points_per_kf = 40 000 of
    ~196 608 valid pixels per 512x384 keyframe (TMA-staged tiles, survivors compacted
    inside the CTA) against the sequential oracle fed the same seeded choice."""
    from tests.util import oracle_dual

    sub = {k: (v[:2] if isinstance(v, np.ndarray) and v.shape[:1] == kf["X"].shape[:1] else v)
           for k, v in kf.items()}
    vp, vq, rays = oracle_dual(sub, VOXEL, TRUNC, max_points=40000, seed=7)
    assert rays == 80000
    vol = _vol(cuda_device)
    t = {k: torch.from_numpy(np.ascontiguousarray(sub[k])).to(cuda_device)
         for k in ("X", "C", "pose", "q")}
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W, max_points=40000, seed=7)
    c = vol.counters()
    assert c["rays"] == rays and c["updates_primary"] == vp.events
    assert c["updates_quality"] == vq.events and c["dropped"] == 0
    assert_volume_matches(vol.export(0), vp.export(), "capped full-size primary")
    assert_volume_matches(vol.export(1), vq.export(), "capped full-size quality")


def _oracle_dual_threaded(kf, voxel, trunc):
    """tests.util.oracle_dual with the two volumes on one thread each (the C
    oracle releases the GIL): full-resolution keyframes take ~0.2 s each."""
    import threading

    vp, vq = orc.OracleVolume(voxel, trunc), orc.OracleVolume(voxel, trunc)
    staged = [orc.stage_keyframe(kf["X"][b], kf["C"][b], kf["pose"][b], kf["q"][b], H, W, 0.05)
              for b in range(kf["X"].shape[0])]
    rays = [0]

    def run(vol, which):
        for pts, conf, confq, org in staged:
            n = vol.integrate(pts, conf if which == 0 else confq, org)
            if which == 0:
                rays[0] += n

    th = [threading.Thread(target=run, args=(vp, 0)), threading.Thread(target=run, args=(vq, 1))]
    [t.start() for t in th]
    [t.join() for t in th]
    return vp, vq, rays[0]


@pytest.mark.parametrize("name,n_kf,voxel,dim", [("config2_prefix", 16, 0.005, 512),
                                                 ("config1", 16, 0.01, 256)])
def test_full_resolution_keyframes_match_oracle(cuda_device, name, n_kf, voxel, dim):
    """BASELINE config 2's 16-keyframe prefix (512x384 into 512^3 @
    5 mm) and config 1 (16 keyframes into 256^3 @ 10 mm), every pixel of every
    keyframe, dual volume, against the sequential oracle: voxel key sets and
    per-voxel update counts bit-exact, tsdf / weight within 1e-5, counters and
    stats equal."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    trunc = 4 * voxel
    scene = synth.SCENES["room"]
    kfs = synth.make_keyframes(scene, n_kf, H, W, seed=0, orbit_radius=0.6, gh=24, gw=32,
                               start=100, total=600)
    vp, vq, rays = _oracle_dual_threaded(kfs, voxel, trunc)
    lo, hi = synth.bounds_for(scene, voxel, dim)
    vol = TSDFVolume(voxel, trunc, capacity_log2=17, bounds=(lo, hi), device=cuda_device)
    t = {k: torch.from_numpy(kfs[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)
    c = vol.counters()
    assert c["rays"] == rays and c["dropped"] == 0 and c["overflow"] == 0
    assert c["updates_primary"] == vp.events and c["updates_quality"] == vq.events
    assert vp.events > 2_000_000 * n_kf
    assert_volume_matches(vol.export(0), vp.export(), f"{name} primary")
    assert_volume_matches(vol.export(1), vq.export(), f"{name} quality")
    assert vol.stats(0) == vp.stats() and vol.stats(1) == vq.stats()


# ---------------------------------------------------------------------------
# north_star Target: 1024^3 @ 5 mm (bench.py's workload at every N)
# ---------------------------------------------------------------------------
TARGET_SCENE = {"kind": "room", "half": 2.4874}
TARGET_DIM, TARGET_ORBIT = 1024, 1.2


@pytest.fixture(scope="module")
def target_case():
    """8 full-resolution keyframes of bench.py's Target scene (z-up orbit of
    radius 1.2 m in the 5 m room) and the two sequential oracle volumes."""
    kfs = synth.make_keyframes(TARGET_SCENE, 8, H, W, seed=0, orbit_radius=TARGET_ORBIT, gh=24,
                               gw=32, start=100, total=600, up_axis="z")
    vp, vq, rays = _oracle_dual_threaded(kfs, VOXEL, TRUNC)
    return kfs, vp, vq, rays


def test_target_1024_single_gpu_matches_oracle(cuda_device, target_case):
    """T1: the Target configuration on ONE GPU (capacity_log2 = 19 as in
    bench.py), every pixel of 8 keyframes, against the sequential oracle."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kfs, vp, vq, rays = target_case
    lo, hi = synth.bounds_for(TARGET_SCENE, VOXEL, TARGET_DIM)
    vol = TSDFVolume(VOXEL, TRUNC, capacity_log2=19, bounds=(lo, hi), device=cuda_device)
    t = {k: torch.from_numpy(kfs[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W)
    c = vol.counters()
    assert c["rays"] == rays and c["dropped"] == 0 and c["overflow"] == 0
    assert c["updates_primary"] == vp.events and c["updates_quality"] == vq.events
    assert_volume_matches(vol.export(0), vp.export(), "target primary")
    assert_volume_matches(vol.export(1), vq.export(), "target quality")
    assert vol.stats(0) == vp.stats() and vol.stats(1) == vq.stats()


@pytest.mark.parametrize("R", [2, 4, 8])
def test_target_1024_slab_union_matches_oracle(cuda_device, target_case, R):
    """The N > 1 bench configuration with one GPU emulating the R ranks in turn:
    interleaved z-slabs of 1024 / (4 R) layers, the batch arriving as R packed
    segments (what the all-gather delivers).  The union of the ranks' volumes
    is the sequential oracle's volume: keys and counts exact, values 1e-5."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume
    from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import pack_keyframes, packed_layout

    kfs, vp, vq, rays = target_case
    lo, hi = synth.bounds_for(TARGET_SCENE, VOXEL, TARGET_DIM)
    share = 8 // R
    lay = packed_layout(share, H, W, 24, 32)
    buf = torch.zeros(R * lay["seg_len"], dtype=torch.float32, device=cuda_device)
    for r in range(R):
        sl = slice(r * share, (r + 1) * share)
        pack_keyframes(kfs["X"][sl], kfs["C"][sl], kfs["pose"][sl], kfs["q"][sl],
                       out=buf[r * lay["seg_len"]:(r + 1) * lay["seg_len"]])
    width = TARGET_DIM // (4 * R)
    parts = {0: [], 1: []}
    upd_p = upd_q = 0
    for r in range(R):
        vol = TSDFVolume(VOXEL, TRUNC, capacity_log2=18, bounds=(lo, hi), slabs=(R, r, width),
                         device=cuda_device)
        vol.integrate_keyframes_packed(buf, R, share, H, W, 24, 32, 0.05)
        c = vol.counters()
        assert c["rays"] == rays and c["overflow"] == 0
        upd_p += c["updates_primary"]
        upd_q += c["updates_quality"]
        for which in (0, 1):
            parts[which].append(vol.export(which))
        vol.close()
    assert upd_p == vp.events and upd_q == vq.events
    for which, orc_vol, name in ((0, vp, "primary"), (1, vq, "quality")):
        k = np.concatenate([p[0] for p in parts[which]])
        order = np.lexsort((k[:, 2], k[:, 1], k[:, 0]))
        got = tuple(np.concatenate([p[i] for p in parts[which]])[order] for i in range(4))
        assert len(np.unique(got[0], axis=0)) == len(got[0]), "a voxel has two owners"
        assert_volume_matches(got, orc_vol.export(), f"R={R} slab union {name}")


@pytest.mark.parametrize("R", [1, 2, 8])
def test_target_1024_peer_pull_matches_oracle(cuda_device, target_case, R):
    """The peer-pull ingest (b200tsdf_keyframe_meta -> gathered meta records ->
    b200tsdf_integrate_keyframes_pull reading the tiles through per-segment
    pointers) with one GPU emulating the R ranks in turn: segments are separate
    buffers, every emulated rank lists and loads only the tiles whose band
    reaches its slabs.  Union == sequential oracle; ray / culled accounting
    equals the broadcast path's."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kfs, vp, vq, rays = target_case
    lo, hi = synth.bounds_for(TARGET_SCENE, VOXEL, TARGET_DIM)
    share = 8 // R
    seg = []
    for r in range(R):
        sl = slice(r * share, (r + 1) * share)
        seg.append({k: torch.from_numpy(np.ascontiguousarray(kfs[k][sl])).to(cuda_device)
                    for k in ("X", "C", "pose", "q")})
    width = max(8, TARGET_DIM // (4 * R))
    owner = TSDFVolume(VOXEL, TRUNC, capacity_log2=10, bounds=(lo, hi), device=cuda_device)
    words = owner.keyframe_meta_words(share, H, W, 24, 32)
    meta = torch.cat([owner.keyframe_meta(s["X"], s["C"], s["pose"], s["q"], 0.05, H, W)
                      for s in seg])
    assert meta.numel() == R * words
    xs, cs = [s["X"].data_ptr() for s in seg], [s["C"].data_ptr() for s in seg]
    parts = {0: [], 1: []}
    upd_p = upd_q = culled = 0
    for r in range(R):
        vol = TSDFVolume(VOXEL, TRUNC, capacity_log2=18, bounds=(lo, hi),
                         slabs=(R, r, width) if R > 1 else None, device=cuda_device)
        vol.integrate_keyframes_pull(xs, cs, meta, share, H, W, 24, 32, 0.05)
        c = vol.counters()
        assert c["rays"] == rays and c["overflow"] == 0
        upd_p += c["updates_primary"]
        upd_q += c["updates_quality"]
        culled += c["culled_rays"]
        for which in (0, 1):
            parts[which].append(vol.export(which))
        vol.close()
    assert upd_p == vp.events and upd_q == vq.events
    assert (culled > 0) == (R > 1)
    for which, orc_vol, name in ((0, vp, "primary"), (1, vq, "quality")):
        k = np.concatenate([p[0] for p in parts[which]])
        order = np.lexsort((k[:, 2], k[:, 1], k[:, 0]))
        got = tuple(np.concatenate([p[i] for p in parts[which]])[order] for i in range(4))
        assert len(np.unique(got[0], axis=0)) == len(got[0]), "a voxel has two owners"
        assert_volume_matches(got, orc_vol.export(), f"R={R} peer pull {name}")
