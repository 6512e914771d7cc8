"""GPU: TSDFRefiner mirror against the golden frozen from the
unmodified reference block flow (tests/golden/make_golden_refine.py)."""
import os
import queue

import numpy as np
import pytest
import torch

from oracle import ref_harness as rh
from tests.golden import make_golden_refine as mg

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(mg.__file__), "refine_plane.npz"))


def _setup(dev, quality=True):
    from mast3r_slam_quality_dualtsdf_b200.tsdf_refine import TSDFRefiner

    X, Cf, pose, K = mg.inputs()
    kfs = rh.FakeKeyframes([torch.from_numpy(X).to(dev)], [torch.from_numpy(Cf.copy()).to(dev)],
                           [torch.from_numpy(pose).to(dev)], None, img_shape=(mg.H, mg.W),
                           K=torch.from_numpy(K))
    r = TSDFRefiner(dict(mg.CFG), kfs, mg.Quality() if quality else None, device=dev)
    r.pixel_order = lambda n: mg.perm(n).to(dev)
    return r, kfs, Cf


def test_block_flow_matches_reference_golden(cuda_device):
    r, kfs, Cf = _setup(cuda_device)
    ok, results = mg.run_flow(r, kfs)
    want = G["results"]
    assert ok == bool(G["ok"]) and len(results) == len(want)
    got = np.array(results, np.float64)
    assert np.array_equal(got[:, :3], want[:, :3])  # block ids and patch coordinates
    assert np.allclose(got[:, 3:5], want[:, 3:5], atol=1e-6)  # priority, median depth
    assert np.array_equal(got[:, 5], want[:, 5])  # success flags
    assert np.allclose(got[:, 6], want[:, 6], atol=1e-6)  # gain = max(hit ratio, mean displacement)
    assert np.array_equal(kfs.version.numpy(), G["version"])
    assert np.allclose(kfs.C[0].cpu().numpy(), G["C_after"], atol=1e-6)
    dbg = r.stats["debug_info"]
    assert [dbg[k] for k in ("tsdf_constructions", "surface_extractions", "displacement_rejects",
                             "hit_ratio_rejects")] == list(G["debug"])


def test_thread_registry_and_sliding_window(cuda_device):
    import time

    r, kfs, Cf = _setup(cuda_device)
    # not enough keyframes behind the window: nothing scheduled (:307-308)
    r.maybe_schedule_sliding_window(0)
    assert r.queue.qsize() == 0
    r.cfg["window_size"] = 0
    r.maybe_schedule_sliding_window(0)
    n = r.queue.qsize()
    assert n == 3
    r.maybe_schedule_sliding_window(0)  # same blocks are QUEUED: de-duplicated by the registry
    assert r.queue.qsize() == n
    r.start()
    t0 = time.time()
    while r.stats["total_blocks"] < n and time.time() - t0 < 30:
        time.sleep(0.01)
    r.stop_flag.set()
    r.join(timeout=5)
    assert not r.is_alive()
    assert r.stats["total_blocks"] == 3 and r.stats["successful_blocks"] == 2
    st = r.registry.get_stats()
    assert st["total_attempts"] == 3 and abs(st["success_rate"] - 2 / 3) < 1e-9
    assert all(s == r.registry.COOLDOWN for s in r.registry.state.values())
    r.registry.frame_count = 100
    key = next(iter(r.registry.state))
    assert r.registry.cooldown_expired(key) and r.registry.state[key] == r.registry.IDLE


def test_missing_config_and_missing_K(cuda_device):
    from mast3r_slam_quality_dualtsdf_b200.tsdf_refine import TSDFRefiner

    with pytest.raises(ValueError):
        TSDFRefiner({"voxel_size": 0.02}, None, None, device=cuda_device)
    # reference quirk: K = None -> K.to() raises -> the block fails
    r, kfs, _ = _setup(cuda_device)
    kfs.K = None
    assert r._schedule_refinement(0)
    key, block = r.queue.get_nowait()
    assert r.process(key, block) == (False, 0.0)
    assert r.stats["total_blocks"] == 0 and r.registry.history[key]["attempts"] == 1


def test_fallback_priority_without_quality_service(cuda_device):
    """:367-396 -- without a quality result the low-confidence pixels drive the
    priority; fewer than 100 of them -> nothing to do (False = retry later)."""
    r, kfs, Cf = _setup(cuda_device, quality=False)
    assert r._schedule_refinement(0) is True  # 256 low-confidence pixels (patch (1,0) at C = 0.1)
    kfs.C[0][:] = 1.0
    assert r._schedule_refinement(0) is False
    assert r._schedule_refinement(5) is False  # no such keyframe
