"""GPU: host mirrors of tsdf/global_manager.py and tsdf/tsdf_optimizer.py
 against the oracle and the reference goldens."""
import os
import types

import numpy as np
import pytest
import torch

from oracle import opt_oracle
from oracle import ref_harness as rh
from tests.golden import make_golden_opt as mg
from tests.test_opt_oracle_golden import G, oracle_volume
from tests.util import assert_volume_matches, oracle_dual

pytestmark = pytest.mark.gpu


def _canon_q(p):
    """q and -q are the same rotation."""
    p = np.array(p, np.float64)
    if p[6] < 0:
        p[3:7] = -p[3:7]
    return p


def _keyframes(kf, poses, dev):
    k = rh.FakeKeyframes([torch.from_numpy(x).to(dev) for x in kf["X"]],
                         [torch.from_numpy(c).to(dev) for c in kf["C"]],
                         [torch.from_numpy(np.asarray(p, np.float32)).to(dev) for p in poses],
                         None, img_shape=(kf["H"], kf["W"]))
    k.device = dev
    return k


@pytest.mark.parametrize("case", sorted(mg.CASES))
def test_pose_optimizer_matches_reference_golden(cuda_device, case):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFPoseOptimizer, TSDFVolume

    kf, bad = mg.scene()
    vol = TSDFVolume(mg.VOXEL, mg.TRUNC, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            None, mg.CFG["min_confidence"], mg.H, mg.W)
    c = mg.CASES[case]
    cfg = dict(mg.CFG, damping=c["damping"], max_iterations=c["iterations"])
    kfs = _keyframes(kf, [kf["pose"][0], bad, kf["pose"][2]], cuda_device)
    opt = TSDFPoseOptimizer(vol, kfs, cfg, False, cuda_device)
    opt._optimize_single(1, iterations=c["iterations"], choice=G["choice"])
    got = kfs.T_WC[1].cpu().numpy().reshape(8).astype(np.float64)
    want = G["pose_out_" + case]
    # iter1 is one undamped step of an ill-conditioned 7x7 system: float32 voxel
    # sums (1e-7 relative) are amplified; the damped case is tight
    tol = 2e-3 if case == "iter1" else 2e-4
    got, want = _canon_q(got), _canon_q(want)
    assert np.allclose(got, want, rtol=tol, atol=tol * 0.1), (got, want)
    # same statement against the oracle restatement
    ovol = oracle_volume(kf)
    ref = opt_oracle.optimize_single(ovol, kf["X"][1], kf["C"][1], bad, G["choice"],
                                     iterations=c["iterations"], lam=cfg["lambda"],
                                     damping=c["damping"])
    assert np.allclose(got, _canon_q(ref), rtol=tol, atol=tol * 0.1)


def test_tsdf_factor_blocks_match_oracle_and_optimizer_step(cuda_device):
    """SURVEY 8(f) rank 4 / roadmap 9.3.2 (MASt3R_SLAM_TSDF_Integration.md:240-252): the
    unary TSDF factor's 7x7 block and gradient per keyframe equal the oracle's normal
    equations of tsdf_optimizer.py:94-116 at the same pose and sample choice; scattered
    into a two-keyframe system (variable order as unique_kf_idx[pin:]) and solved with
    the optimiser's damping, the update of a keyframe is the optimiser's own first
    Gauss-Newton step (:82-86)."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFFactor, TSDFPoseOptimizer, TSDFVolume
    from mast3r_slam_quality_dualtsdf_b200.tsdf import sim3
    from oracle.opt_oracle import _act_f32, normal_equations

    kf, bad = mg.scene()
    vol = TSDFVolume(mg.VOXEL, mg.TRUNC, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            None, mg.CFG["min_confidence"], mg.H, mg.W)
    ovol = oracle_volume(kf)
    poses = [kf["pose"][0], bad, kf["pose"][2]]
    kfs = _keyframes(kf, poses, cuda_device)
    cfg = dict(mg.CFG, damping=1.0e-4)
    fac = TSDFFactor(vol, kfs, cfg, cuda_device)
    choice = {1: G["choice"], 2: G["choice"]}
    blocks = fac.linearize([1, 2, 7], choices=choice)  # 7: not a keyframe, skipped
    assert sorted(blocks) == [1, 2]
    for idx in (1, 2):
        pts = kf["X"][idx].reshape(-1, 3)[G["choice"]].astype(np.float32)
        conf = kf["C"][idx].reshape(-1)[G["choice"]]
        Ho, bo, uo = normal_equations(ovol, _act_f32(np.asarray(poses[idx], np.float32), pts), conf,
                                      cfg["lambda"])
        H, b, used = blocks[idx]
        # the probe set may differ by the few points whose gradient presence flips between
        # float32 sums and the reference's float64 dict: same bound as the
        # optimiser's golden test
        assert abs(used - uo) <= max(2, uo // 200)
        assert np.allclose(H, Ho, rtol=2e-3, atol=2e-3 * np.abs(Ho).max())
        assert np.allclose(b, bo, rtol=2e-3, atol=2e-3 * np.abs(bo).max())
        assert np.allclose(H, H.T)
    # variable order [2, 1] with keyframe 0 pinned: blocks land on the diagonal
    Hf, bf = np.zeros((14, 14)), np.zeros(14)
    assert TSDFFactor.add_to(Hf, bf, blocks, order=[2, 1]) == 2
    assert np.array_equal(Hf[:7, :7], blocks[2][0]) and np.array_equal(Hf[7:, 7:], blocks[1][0])
    assert not Hf[:7, 7:].any() and np.array_equal(bf[7:], blocks[1][1])
    delta = np.linalg.solve(Hf + cfg["damping"] * np.eye(14), -bf)
    want = sim3.mul(sim3.exp(torch.from_numpy(delta[7:]).to(torch.float32).to(torch.float64)),
                    torch.from_numpy(np.asarray(bad, np.float64))).numpy()
    opt = TSDFPoseOptimizer(vol, kfs, dict(cfg, device_solve=False), False, cuda_device)
    opt._optimize_single(1, iterations=1, choice=G["choice"])
    got = kfs.T_WC[1].cpu().numpy().reshape(8).astype(np.float64)
    assert np.allclose(_canon_q(got), _canon_q(want), rtol=1e-6, atol=1e-7)


def test_device_gauss_newton_equals_host_solve(cuda_device):
    """the whole loop of _optimize_single on the device
    (b200tsdf_pose_optimize: one launch per iteration -- linearise, 7x7 solve,
    Sim3 retraction -- and one synchronisation) against the same loop with the
    solve in numpy on the host after every linearisation."""
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFPoseOptimizer, TSDFVolume

    kf, bad = mg.scene()
    vol = TSDFVolume(mg.VOXEL, mg.TRUNC, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            None, mg.CFG["min_confidence"], mg.H, mg.W)
    out = {}
    for dev_solve in (True, False):
        kfs = _keyframes(kf, [kf["pose"][0], bad, kf["pose"][2]], cuda_device)
        opt = TSDFPoseOptimizer(vol, kfs, dict(mg.CFG, device_solve=dev_solve, max_iterations=3,
                                               damping=500.0), False, cuda_device)
        opt._optimize_single(1, iterations=3, choice=G["choice"])
        out[dev_solve] = (_canon_q(kfs.T_WC[1].cpu().numpy().reshape(8)), opt.last_info["used"])
    assert out[True][1] == out[False][1] > 50
    # same float64 normal equations, LU with partial pivoting on both sides
    assert np.allclose(out[True][0], out[False][0], rtol=1e-6, atol=1e-7), out
    # direct call: iteration count and the early stop of an empty volume (:78-79)
    pts = torch.from_numpy(kf["X"][1][G["choice"]]).to(cuda_device)
    conf = torch.from_numpy(kf["C"][1][G["choice"]]).to(cuda_device, torch.float64)
    pose, used, done = vol.pose_optimize(pts, conf, bad, mg.CFG["lambda"], 500.0, 3)
    assert done == 3 and used > 50 and np.isfinite(pose).all()
    # the undamped system of this scene runs away after one step: the loop stops like the
    # reference's `if used < 6: break` (:78-79), on the device as on the host
    _, used_u, done_u = vol.pose_optimize(pts, conf, bad, mg.CFG["lambda"], 1e-4, 3)
    assert done_u < 3 and used_u < 6
    empty = TSDFVolume(mg.VOXEL, mg.TRUNC, device=cuda_device)
    pose0, used0, done0 = empty.pose_optimize(pts, conf, bad, mg.CFG["lambda"], 500.0, 3)
    assert done0 == 0 and used0 == 0 and np.allclose(pose0, np.asarray(bad, np.float64))


def test_normal_equations_match_oracle(cuda_device):
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFPoseOptimizer, TSDFVolume
    from mast3r_slam_quality_dualtsdf_b200.tsdf import sim3

    kf, bad = mg.scene()
    vol = TSDFVolume(mg.VOXEL, mg.TRUNC, device=cuda_device)
    vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose")),
                            None, mg.CFG["min_confidence"], mg.H, mg.W)
    kfs = _keyframes(kf, kf["pose"], cuda_device)
    opt = TSDFPoseOptimizer(vol, kfs, mg.CFG, False, cuda_device)
    pts = torch.from_numpy(kf["X"][1][G["choice"]]).to(cuda_device)
    pw = sim3.act(torch.from_numpy(bad).to(cuda_device), pts)
    conf = torch.from_numpy(kf["C"][1][G["choice"]]).to(cuda_device, torch.float64)
    H, b, used = opt._normal_equations(pw, conf)
    Ho, bo, uo = opt_oracle.normal_equations(oracle_volume(kf), pw.cpu().numpy(),
                                             conf.cpu().numpy(), mg.CFG["lambda"])
    assert used == uo and used > 50
    assert np.allclose(H, Ho, rtol=1e-4, atol=1e-6 * np.abs(Ho).max())
    assert np.allclose(b, bo, rtol=1e-4, atol=1e-6 * np.abs(bo).max())
    # the fused kernel (Sim3 act + query + Jacobian + 7x7 reduction in one launch,
    # b200tsdf_pose_normal_equations) states the same linearisation: float64 sums
    # in a different order, world points by the kernel's own float32 Sim3 act
    Hf, bf, uf = vol.pose_normal_equations(pts, conf, bad, mg.CFG["lambda"])
    assert abs(uf - used) <= 2  # a point on a voxel face may round to the other side
    assert np.allclose(Hf, H, rtol=2e-2, atol=2e-2 * np.abs(H).max())
    assert np.allclose(Hf, Hf.T) and np.all(np.linalg.eigvalsh(Hf) > -1e-9 * np.abs(Hf).max())
    pw_k = torch.from_numpy(kf["X"][1][G["choice"]]).to(cuda_device)
    # exact cross-check: feed the torch statement the SAME world points the kernel forms
    from mast3r_slam_quality_dualtsdf_b200.tsdf.sim3 import act_kernel_order
    Hk, bk, uk = opt._normal_equations(act_kernel_order(torch.from_numpy(bad), pw_k), conf)
    assert uk == uf
    assert np.allclose(Hf, Hk, rtol=1e-10, atol=1e-12 * np.abs(Hk).max())
    assert np.allclose(bf, bk, rtol=1e-10, atol=1e-12 * np.abs(bk).max())
    # no point, empty volume: zeros
    H0, b0, u0 = TSDFVolume(mg.VOXEL, mg.TRUNC, device=cuda_device).pose_normal_equations(
        pts, conf, bad, mg.CFG["lambda"])
    assert u0 == 0 and not H0.any() and not b0.any()


def test_manager_integrates_and_reintegrates_like_reference(cuda_device):
    """global_manager.py:56-106, :214-229: new keyframes are fused once, in
    order; after a backend solve the listed keyframes (minus the pinned ones)
    are fused AGAIN at their current pose (additive, no de-integration)."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFGlobalManager

    kf = synth.make_keyframes(synth.SCENES["sphere"], 4, H=48, W=64, seed=3, f=50.0, total=12)
    kfs = _keyframes(kf, kf["pose"], cuda_device)
    cfg = {"enabled": True, "voxel_size": 0.03, "trunc_dist": 0.12, "min_confidence": 0.05,
           "async_optimize": True, "pre_icp_iters": 0, "pin": 1, "reintegration_queue": 3}
    mgr = TSDFGlobalManager(kfs, cfg, False, cuda_device)
    assert mgr.enabled and mgr.opt_worker is not None
    mgr.integrator.step()  # what the thread does each wake-up
    assert mgr.integrator.next_idx == 4
    vp, _, rays = oracle_dual(kf, 0.03, 0.12, use_q=False)
    assert_volume_matches(mgr.volume.export(0), vp.export(), "manager first pass")
    assert mgr.volume.stats() == vp.stats()
    # backend solve touched keyframes 0,1,2,2,3: 0 is pinned, 2 is de-duplicated,
    # the queue holds 3 -> {1, 2, 3}
    fg = types.SimpleNamespace(last_unique_kf_idx=torch.tensor([0, 1, 2, 2, 3]))
    mgr.on_after_backend_solve(fg)
    assert sorted(mgr.integrator.pending) == [1, 2, 3]
    assert mgr.opt_worker.queue.qsize() == 3
    mgr.integrator.step()
    assert not mgr.integrator.pending
    vp2, _, _ = oracle_dual(kf, 0.03, 0.12, use_q=False)
    # additive re-fusion = the same keyframes integrated twice
    from oracle import oracle as orc
    for b in (1, 2, 3):
        pts, conf, _, org = orc.stage_keyframe(kf["X"][b], kf["C"][b], kf["pose"][b], None, 48, 64,
                                               0.05)
        vp2.integrate(pts, conf, org)
    assert_volume_matches(mgr.volume.export(0), vp2.export(), "manager after re-integration")
    # disabled manager is inert (global_manager.py:197, :215)
    off = TSDFGlobalManager(kfs, dict(cfg, enabled=False), False, cuda_device)
    off.start()
    off.on_after_backend_solve(fg)
    assert off.integrator.reintegration_queue.empty() and not off.integrator.is_alive()
    off.shutdown()


def test_manager_threads_start_and_stop(cuda_device):
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFGlobalManager
    import time

    kf = synth.make_keyframes(synth.SCENES["plane"], 2, H=32, W=32, seed=3, f=40.0)
    kfs = _keyframes(kf, kf["pose"], cuda_device)
    cfg = {"enabled": True, "voxel_size": 0.03, "trunc_dist": 0.12, "queue_check_interval": 0.01,
           "opt_cooldown_ms": 1.0, "damping": 500.0, "samples_per_kf": 100}
    mgr = TSDFGlobalManager(kfs, cfg, False, cuda_device)
    mgr.start()
    t0 = time.time()
    while mgr.integrator.next_idx < 2 and time.time() - t0 < 20:
        time.sleep(0.01)
    assert mgr.integrator.next_idx == 2
    before = kfs.T_WC[1].clone()
    mgr.on_after_backend_solve(types.SimpleNamespace(last_unique_kf_idx=torch.tensor([0, 1])))
    t0 = time.time()
    while (mgr.opt_worker.pending or mgr.integrator.pending) and time.time() - t0 < 20:
        time.sleep(0.01)
    mgr.shutdown()
    assert not mgr.integrator.is_alive() and not mgr.opt_worker.is_alive()
    assert mgr.integrator.integrated == 3  # 2 new + 1 re-integration (keyframe 0 is pinned)
    assert torch.isfinite(kfs.T_WC[1]).all() and kfs.T_WC[1].shape == before.shape


def test_deintegrate_is_the_exact_inverse(cuda_device):
    """integrate A, B, C then remove B == integrate A, C
    (voxel key set and hit counters exact, tsdf / weight within tolerance)."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=48, W=64, seed=8, f=50.0, total=9)
    t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    vol = TSDFVolume(0.03, 0.12, device=cuda_device)
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, 48, 64)
    sl = slice(1, 2)
    vol.deintegrate_keyframes(t["X"][sl].contiguous(), t["C"][sl].contiguous(),
                              t["pose"][sl].contiguous(), t["q"][sl].contiguous(), 0.05, 48, 64)
    vp, vq, _ = oracle_dual(kf, 0.03, 0.12, idx=[0, 2])
    for which, want in ((0, vp), (1, vq)):
        gk, gt, gw, gc = vol.export(which)
        wk, wt, ww, wc = want.export()
        assert np.array_equal(gk, wk) and np.array_equal(gc.astype(np.int64), wc.astype(np.int64))
        assert np.abs(gt - wt).max() <= 5e-5  # float32 cancellation of the removed sums
        assert (np.abs(gw - ww) / np.maximum(1.0, ww)).max() <= 5e-5
    assert vol.stats(0) == vp.stats()


def test_manager_replace_mode_removes_the_stale_pose(cuda_device):
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFGlobalManager

    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=48, W=64, seed=3, f=50.0, total=12)
    kfs = _keyframes(kf, kf["pose"], cuda_device)
    cfg = {"enabled": True, "voxel_size": 0.03, "trunc_dist": 0.12, "min_confidence": 0.05,
           "async_optimize": False, "max_iterations": 0, "pin": 1,
           "reintegration_mode": "replace"}
    mgr = TSDFGlobalManager(kfs, cfg, False, cuda_device)
    mgr.integrator.step()
    # the backend moves keyframe 2; the manager replaces its contribution
    new_pose = kf["pose"][2].copy()
    new_pose[:3] += np.array([0.02, -0.01, 0.015], np.float32)
    kfs.T_WC[2] = torch.from_numpy(new_pose).to(cuda_device).reshape(1, 8)
    mgr.on_after_backend_solve(types.SimpleNamespace(last_unique_kf_idx=torch.tensor([0, 2])))
    mgr.integrator.step()
    moved = dict(kf)
    moved["pose"] = kf["pose"].copy()
    moved["pose"][2] = new_pose
    vp, _, _ = oracle_dual(moved, 0.03, 0.12, use_q=False)
    gk, gt, gw, gc = mgr.volume.export(0)
    wk, wt, ww, wc = vp.export()
    assert np.array_equal(gk, wk) and np.array_equal(gc.astype(np.int64), wc.astype(np.int64))
    assert np.abs(gt - wt).max() <= 5e-5


def test_checkpoint_round_trip_and_merge(cuda_device, tmp_path):
    """the volume state is a flat (block keys, voxels,
    counters) file; restoring is bit-exact, and because the fold state is
    additive, importing one partial fusion into another equals fusing all
    keyframes into one volume (counts exact, sums up to float32 order)."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    kf = synth.make_keyframes(synth.SCENES["sphere"], 4, H=48, W=64, seed=3, f=50.0, total=8)
    t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}

    def fused(idx, **kw):
        v = TSDFVolume(0.02, 0.08, device=cuda_device, **kw)
        v.integrate_keyframes(t["X"][idx], t["C"][idx], t["pose"][idx], t["q"][idx], 0.05, 48, 64)
        return v

    whole = fused(slice(0, 4))
    path = str(tmp_path / "vol.npz")
    nblk = whole.save_checkpoint(path)
    assert nblk == whole.counters()["blocks"] > 0
    # the oracle is the checker of what comes back from disk, not only the volume that
    # wrote it: the reference dict fused with the same keyframes
    vp, vq, _ = oracle_dual(kf, 0.02, 0.08)
    for kw in ({}, {"bounds": ((-64, -64, 0), (64, 64, 128))}):  # hash-only and directory volumes
        back = TSDFVolume(0.02, 0.08, device=cuda_device, **kw)
        assert back.load_checkpoint(path) == nblk
        for which in (0, 1):
            a, b = whole.export(which), back.export(which)
            assert all(np.array_equal(x, y) for x, y in zip(a, b))
        assert back.stats() == whole.stats()
        assert_volume_matches(back.export(0), vp.export(), "restored primary vs oracle")
        assert_volume_matches(back.export(1), vq.export(), "restored quality vs oracle")
        assert back.stats(0) == vp.stats() and back.stats(1) == vq.stats()
        assert np.array_equal(back.block_keys(), np.unique(vp.export()[0] >> 3, axis=0))
        # the restored volume keeps fusing like the original
        back.integrate_keyframes(t["X"][:1], t["C"][:1], t["pose"][:1], t["q"][:1], 0.05, 48, 64)
        assert back.counters()["blocks"] >= nblk
    # merge: {0,1} + {2,3} == {0,1,2,3}
    left, right = fused(slice(0, 2)), fused(slice(2, 4))
    left.import_blocks(*right.export_block_data())
    for which in (0, 1):
        gk, gt, gw, gc = left.export(which)
        wk, wt, ww, wc = whole.export(which)
        assert np.array_equal(gk, wk) and np.array_equal(gc, wc)
        assert np.abs(gt - wt).max() <= 1e-5 and np.abs(gw - ww).max() <= 1e-4 * max(1.0, ww.max())
    assert_volume_matches(left.export(0), vp.export(), "merged primary vs oracle")
    assert_volume_matches(left.export(1), vq.export(), "merged quality vs oracle")
    # a checkpoint loaded on top of a fused volume = the reference fusing both sets
    # (additive sums): {0,1} from disk onto {2,3}
    half = str(tmp_path / "half.npz")
    fused(slice(0, 2)).save_checkpoint(half)
    onto = fused(slice(2, 4))
    onto.load_checkpoint(half)
    assert_volume_matches(onto.export(0), vp.export(), "checkpoint merged onto a volume vs oracle")
    with pytest.raises(ValueError):
        TSDFVolume(0.03, 0.08, device=cuda_device).load_checkpoint(path)


# ---------------------------------------------------------------------------
# the batching integrator, max_points_per_kf, overflow, config keys
# ---------------------------------------------------------------------------
class _StackedKeyframes:
    """SharedKeyframes-shaped container (frame.py:236-252): stacked device
    buffers X (n, HW, 3), C (n, HW, 1), T_WC (n, 1, 8) behind an RLock."""

    def __init__(self, kf, dev, n=None):
        import threading

        self.lock = threading.RLock()
        self.X = torch.from_numpy(kf["X"]).to(dev)
        self.C = torch.from_numpy(kf["C"]).to(dev).reshape(kf["C"].shape[0], -1, 1)
        self.T_WC = torch.from_numpy(kf["pose"]).to(dev).reshape(-1, 1, 8)
        self.h, self.w = kf["H"], kf["W"]
        self.n = self.X.shape[0] if n is None else n

    def __len__(self):
        return self.n

    def __getitem__(self, idx):
        f = types.SimpleNamespace()
        f.X_canon, f.C, f.T_WC, f.frame_id = self.X[idx], self.C[idx], self.T_WC[idx], idx
        f.img_shape = torch.tensor([[self.h, self.w]])
        return f


def test_integrator_batches_a_wakeup_into_one_launch(cuda_device):
    """the drop-in class must reach the batched kernel --
    one wake-up = one snapshot + ONE launch for all pending keyframes, a second
    wake-up one launch for all dirty ones; results equal the sequential oracle."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFGlobalManager

    kf = synth.make_keyframes(synth.SCENES["sphere"], 6, H=48, W=64, seed=4, f=50.0, total=12)
    kfs = _StackedKeyframes(kf, cuda_device)
    cfg = {"enabled": True, "voxel_size": 0.03, "trunc_dist": 0.12, "min_confidence": 0.05,
           "async_optimize": False, "max_iterations": 0, "pre_icp_iters": 0, "pin": 1,
           "volume_dim": 256, "capacity_log2": 14}
    mgr = TSDFGlobalManager(kfs, cfg, False, cuda_device)
    assert mgr.volume.bounds == ((-128, -128, -128), (128, 128, 128))
    mgr.integrator.step()
    assert mgr.integrator.next_idx == 6 and mgr.integrator.launches == 1
    vp, _, _ = oracle_dual(kf, 0.03, 0.12, use_q=False)
    assert_volume_matches(mgr.volume.export(0), vp.export(), "batched wake-up")
    mgr.on_after_backend_solve(types.SimpleNamespace(last_unique_kf_idx=torch.tensor([1, 3, 5])))
    mgr.integrator.step()
    assert mgr.integrator.launches == 2 and mgr.integrator.integrated == 9
    # with pre-ICP the reference order (fuse i, refine i, fuse i+1) is kept: one launch each
    kfs2 = _StackedKeyframes(kf, cuda_device)
    m2 = TSDFGlobalManager(kfs2, dict(cfg, pre_icp_iters=1, samples_per_kf=64, damping=1.0e3),
                           False, cuda_device)
    m2.integrator.step()
    assert m2.integrator.launches == 6
    m3 = TSDFGlobalManager(_StackedKeyframes(kf, cuda_device),
                           dict(cfg, pre_icp_iters=1, samples_per_kf=64, damping=1.0e3,
                                batch_with_pre_icp=True), False, cuda_device)
    m3.integrator.step()
    assert m3.integrator.launches == 1


@pytest.mark.parametrize("bounded", [False, True])
def test_max_points_per_kf_is_honoured(cuda_device, bounded):
    """global_manager.py:94-99 / config/base.yaml:116: at most max_points valid
    pixels per keyframe are fused.  The choice is the seeded hash subsample of
    oracle.subsample_mask; volumes equal the oracle fused with the same choice."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFGlobalManager, TSDFVolume

    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=48, W=64, seed=6, f=50.0, total=9)
    cap, seed = 700, 1234
    n_valid = (kf["C"] > 0.05).sum(axis=1)
    assert (n_valid > cap).all()
    vp, vq, rays = oracle_dual(kf, 0.03, 0.12, max_points=cap, seed=seed)
    assert rays == 3 * cap
    kw = {"bounds": ((-128, -128, -64), (128, 128, 192))} if bounded else {}
    vol = TSDFVolume(0.03, 0.12, device=cuda_device, **kw)
    t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, 48, 64, max_points=cap,
                            seed=seed)
    c = vol.counters()
    assert c["rays"] == rays and c["updates_primary"] == vp.events
    assert_volume_matches(vol.export(0), vp.export(), "capped primary")
    assert_volume_matches(vol.export(1), vq.export(), "capped quality")
    # the same selection removes what it added
    vol.deintegrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, 48, 64, max_points=cap,
                              seed=seed)
    assert vol.stats(0)["total_voxels"] == 0
    # through the manager: keyframe b is salted with its index b
    kfs = _StackedKeyframes(kf, cuda_device)
    cfg = {"enabled": True, "voxel_size": 0.03, "trunc_dist": 0.12, "min_confidence": 0.05,
           "async_optimize": False, "pre_icp_iters": 0, "max_points_per_kf": cap, "seed": seed}
    mgr = TSDFGlobalManager(kfs, cfg, False, cuda_device)
    mgr.integrator.step()
    assert_volume_matches(mgr.volume.export(0), vp.export(), "capped manager")
    # a different seed is a different choice of the same size
    other = TSDFVolume(0.03, 0.12, device=cuda_device)
    other.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, 48, 64, max_points=cap,
                              seed=seed + 1)
    assert other.counters()["rays"] == rays
    ka, kb = other.export(0)[0], vp.export()[0]
    assert ka.shape != kb.shape or not np.array_equal(ka, kb)


@pytest.mark.parametrize("ranks", [1, 2, 4])
def test_capped_rays_are_compacted_ragged_and_sharded(cuda_device, ranks):
    """The max_points_per_kf survivors of a tile are compacted onto the CTA's first
    warps (csrc/integrate.cu tile_consume).  Ragged image (45x60: partial tiles, the
    non-TMA loads), degenerate rays among the survivors, one volume and the z-slab
    paths (2 ranks: direct launch; 4 ranks: classify pass + listed tiles): the same
    rays as the oracle's seeded choice are fused, counters add up, union bit-exact."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    H, W = 45, 60
    kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=H, W=W, seed=21, f=50.0, total=8,
                              gh=9, gw=12)
    X = kf["X"].reshape(3, H * W, 3)
    Cf = kf["C"].reshape(3, H * W)
    X[0, np.nonzero(Cf[0] > 0.05)[0][::7]] = 0.0          # len 0: skipped (global_volume.py:56)
    X[1, np.nonzero(Cf[1] > 0.05)[0][3::11], 2] = np.inf  # non-finite: skipped
    kf["X"] = X.reshape(kf["X"].shape)
    cap, seed, ids = 500, 99, [5, 17, 2]
    assert ((kf["C"] > 0.05).reshape(3, -1).sum(axis=1) > cap).all()
    vp, vq, rays = oracle_dual(kf, 0.02, 0.08, max_points=cap, seed=seed, kf_ids=ids)
    assert 0 < rays < 3 * cap  # some of the chosen pixels are not rays
    t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    kid = torch.tensor(ids, dtype=torch.int64, device=cuda_device)
    parts, partsq, ev_p, ev_q = [], [], 0, 0
    for r in range(ranks):
        kw = {"slabs": (ranks, r, 8)} if ranks > 1 else {}
        vol = TSDFVolume(0.02, 0.08, device=cuda_device, **kw)
        vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W, max_points=cap,
                                seed=seed, kf_ids=kid)
        c = vol.counters()
        assert c["rays"] == rays
        ev_p += c["updates_primary"]
        ev_q += c["updates_quality"]
        parts.append(vol.export(0))
        partsq.append(vol.export(1))
    assert ev_p == vp.events and ev_q == vq.events
    for ps, want in ((parts, vp), (partsq, vq)):
        gk = np.concatenate([p[0] for p in ps])
        order = np.lexsort((gk[:, 2], gk[:, 1], gk[:, 0]))
        got = tuple(np.concatenate([p[i] for p in ps])[order] for i in range(4))
        assert_volume_matches(got, want.export(), f"capped, {ranks} rank(s)")


def test_threshold_select_list_and_radix_paths_agree(cuda_device, monkeypatch):
    """subsample_threshold_kernel picks the need-th smallest hash either from the list of
    the threshold bin's hashes (default) or, for bins larger than the list, by two more
    radix passes; B200TSDF_SUB_LIST_CAP = 0 forces the latter.  Same selection either way:
    equal to the oracle's argsort choice, on small and on full-size keyframes."""
    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    for (H, W, voxel, cap) in ((48, 64, 0.03, 700), (384, 512, 0.02, 40000)):
        scene = synth.SCENES["sphere"] if H == 48 else synth.SCENES["room"]
        kw = {"f": 50.0} if H == 48 else {"orbit_radius": 0.6, "gh": 24, "gw": 32}
        kf = synth.make_keyframes(scene, 2, H=H, W=W, seed=8, total=9, **kw)
        vp, vq, rays = oracle_dual(kf, voxel, 4 * voxel, max_points=cap, seed=5)
        assert rays == 2 * cap
        t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
        for knob in (None, "0", "3"):
            if knob is None:
                monkeypatch.delenv("B200TSDF_SUB_LIST_CAP", raising=False)
            else:
                monkeypatch.setenv("B200TSDF_SUB_LIST_CAP", knob)
            vol = TSDFVolume(voxel, 4 * voxel, capacity_log2=17, device=cuda_device)
            vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, H, W, max_points=cap,
                                    seed=5)
            c = vol.counters()
            assert c["rays"] == rays and c["updates_primary"] == vp.events, f"list cap {knob}"
            assert_volume_matches(vol.export(0), vp.export(), f"{H}x{W}, list cap {knob}")
            assert_volume_matches(vol.export(1), vq.export(), f"{H}x{W} quality, list cap {knob}")


def test_hash_overflow_is_an_error(cuda_device):
    """B200TSDF_E_CAPACITY is returned, and the drop-in
    integrator raises instead of losing updates silently."""
    from mast3r_slam_quality_dualtsdf_b200 import _ffi, synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFGlobalManager, TSDFVolume

    kf = synth.make_keyframes(synth.SCENES["sphere"], 2, H=48, W=64, seed=3, f=50.0, total=8)
    t = {k: torch.from_numpy(kf[k]).to(cuda_device) for k in ("X", "C", "pose", "q")}
    vol = TSDFVolume(0.01, 0.04, capacity_log2=4, device=cuda_device)  # 16 blocks only
    vol.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], 0.05, 48, 64)
    with pytest.raises(_ffi.B200TSDFError, match="capacity_log2"):
        vol.counters()
    assert vol.counters(check=False)["overflow"] > 0
    with pytest.raises(_ffi.B200TSDFError):
        vol.check()
    cfg = {"enabled": True, "voxel_size": 0.01, "trunc_dist": 0.04, "capacity_log2": 4,
           "async_optimize": False, "pre_icp_iters": 0}
    mgr = TSDFGlobalManager(_StackedKeyframes(kf, cuda_device), cfg, False, cuda_device)
    with pytest.raises(_ffi.B200TSDFError):
        mgr.integrator.step()


def test_config_keys_of_the_survey(cuda_device):
    """volume_dim / volume_origin / quality_volume / seed /
    parity_mode / num_gpus load from the tsdf_global section."""
    from mast3r_slam_quality_dualtsdf_b200 import _ffi
    from mast3r_slam_quality_dualtsdf_b200.tsdf import global_manager as gm

    assert gm.bounds_from_cfg({}) is None
    assert gm.bounds_from_cfg({"volume_dim": 512, "voxel_size": 0.005}) == \
        ((-256, -256, -256), (256, 256, 256))
    b = gm.bounds_from_cfg({"volume_dim": [64, 32, 16], "voxel_size": 0.01,
                            "volume_origin": [-0.105, 0.0, 0.33]})
    assert b == ((-16, 0, 32), (48, 32, 48))  # origin floored to a multiple of 8 voxels
    assert gm.bounds_from_cfg({"bounds": [[0, 0, 0], [8, 8, 8]]}) == ((0, 0, 0), (8, 8, 8))
    with pytest.raises(ValueError, match="parity_mode"):
        gm.build_volume({"parity_mode": "projective"}, cuda_device)
    with pytest.raises(_ffi.B200TSDFError, match="num_gpus"):
        gm.build_volume({"num_gpus": 2, "volume_dim": 64}, cuda_device)

    class Svc:
        def get(self, fid):
            return {"priority": np.full(12, 0.25, np.float32)}

    from mast3r_slam_quality_dualtsdf_b200 import synth

    kf = synth.make_keyframes(synth.SCENES["plane"], 2, H=48, W=64, seed=3, f=50.0)
    cfg = {"enabled": True, "voxel_size": 0.03, "trunc_dist": 0.12, "async_optimize": False,
           "pre_icp_iters": 0}
    on = gm.TSDFGlobalManager(_StackedKeyframes(kf, cuda_device), cfg, False, cuda_device, Svc())
    on.integrator.step()
    off = gm.TSDFGlobalManager(_StackedKeyframes(kf, cuda_device), dict(cfg, quality_volume=False),
                               False, cuda_device, Svc())
    off.integrator.step()
    # q = 1 - 0.25 scales the quality volume's weights; quality_volume: false leaves them plain
    w_on, w_off = on.volume.export(1)[2], off.volume.export(1)[2]
    assert np.allclose(w_on, 0.75 * w_off, rtol=1e-5)
    assert np.allclose(off.volume.export(0)[2], w_off, rtol=1e-6)
