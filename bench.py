#!/usr/bin/env python
"""bench.py -- dual-TSDF fusion throughput on B200 (BASELINE.json metric).

Workload (the same at every N, so the 1 -> 8 curve divides like by like): the
north_star Target -- a TUM fr1_room-shaped synthetic orbit of 600 512x384
keyframes fused into a 1024^3 @ 5 mm dual volume (primary + quality-weighted),
truncation 20 mm.  A STEP = one pass over the 600-keyframe sequence, issued as
15 launches of 40 keyframes (every launch reads a fresh 126 MB batch, the whole
pass 1.9 GB, so nothing is served from L2 twice).

  N = 1 : one volume on one GPU.
  N > 1 : the volume is cut into interleaved z-slabs; every rank holds 1/N of
          each launch's keyframes in its own HBM and the launch's keyframes are
          NCCL all-gathered (= N simultaneous broadcasts over NVLink) INSIDE
          the timed region -- the collective is part of the step.

  value : voxel updates/s (primary + quality volume), inputs resident in HBM
          when the timed region starts
  e2e   : same metric through the reference-facing TSDFVolume API from HOST
          (pinned) keyframe buffers: H2D (+ all-gather) inside the timed
          region, one D2H read of the counters per step
  roofline / cpu_baseline : see DESIGN.md "Measurement"

`--impl reference` times the CPU oracle port of the reference path (the
reference itself is Python and does not travel to the GPU box) on the same
workload, each step a bounded sample of the step's keyframes.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, GH, GW = 384, 512, 24, 32
VOXEL, TRUNC = 0.005, 0.02
MIN_CONF = 0.05
KF_PER_LAUNCH = 40
ORBIT_TOTAL = 600
# algorithmic bytes per keyframe input: X + C + pose + quality grid
KF_BYTES = H * W * 12 + H * W * 4 + 32 + GH * GW * 4

# the Target workload (north_star): every N runs this scene
TARGET = {"scene": {"kind": "room", "half": 2.4874}, "dim": 1024, "orbit_r": 1.2, "up": "z",
          "cap_log2": 19}
# BASELINE.json configs[1] (512^3 @ 5 mm): reported as a second section at N = 1
CONFIG2 = {"scene": {"kind": "room", "half": 1.2437}, "dim": 512, "orbit_r": 0.6, "up": "y",
           "cap_log2": 18}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kf-per-launch", type=int, default=KF_PER_LAUNCH)
    ap.add_argument("--launches-per-step", type=int, default=ORBIT_TOTAL // KF_PER_LAUNCH)
    ap.add_argument("--passes-per-step", type=int, default=4,
                    help="passes over the 600-keyframe orbit per step (every pass fuses all "
                         "launches again): keeps the timed region >= 0.2 s at 8 ranks")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ingest", default="pull", choices=["pull", "gather"],
                    help="N > 1: how a launch's keyframes reach the ranks -- pull: ONE small "
                         "all-gather of meta records, the fuse kernel TMA-loads the tiles it needs "
                         "from the owners' HBM over NVLink; gather: ONE all-gather of the packed "
                         "keyframes (= N broadcasts), double-buffered against the fuse")
    ap.add_argument("--no-pipeline", action="store_true",
                    help="N > 1, pull ingest: meta pass, all-gather and fuse of a launch in series "
                         "on one stream (A/B against the default software pipeline)")
    ap.add_argument("--e2e-per-launch", action="store_true",
                    help="N = 1: one *_host call per 40-keyframe launch instead of one per pass (A/B)")
    ap.add_argument("--no-sections", action="store_true",
                    help="skip the secondary sections (config2, extract, sweep, quality, manager, "
                         "reintegrate): headline value / e2e / roofline only")
    return ap.parse_args()


def slab_width_for(n_gpus, dim=1024):
    """Four slabs per rank: wide enough that few rays straddle a seam, interleaved
    enough that every view loads all ranks (measured: scripts/slab_rank_time.py)."""
    return max(8, dim // (4 * max(1, n_gpus)))


def gen_batch(wl, launch_idx, kf_per_launch, seed=0, offset=0, count=None):
    """Keyframes [offset, offset + count) of launch `launch_idx`'s batch (default: all)."""
    from mast3r_slam_quality_dualtsdf_b200 import synth

    start = (launch_idx * kf_per_launch) % ORBIT_TOTAL + offset
    return synth.make_keyframes(wl["scene"], count or kf_per_launch, H, W, seed=seed,
                                orbit_radius=wl["orbit_r"], gh=GH, gw=GW, start=start,
                                total=ORBIT_TOTAL, up_axis=wl["up"])


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region through NVML
    from a thread (~1 kHz)."""

    def __init__(self, index=0):
        self.sm, self.mx, self.reasons = [], None, set()
        self.stop_flag = threading.Event()
        self.err = None
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv = nv
            self.h = nv.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.t = threading.Thread(target=self._run, daemon=True)
            self.t.start()
        except Exception as e:  # noqa: BLE001
            self.err = f"nvml unavailable: {e}"
            self.t = None

    def _poll(self):
        nv = self.nv
        self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) \
            if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
            else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
        for name, bit in (("hw_slowdown", 0x8), ("sw_power_cap", 0x4),
                          ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20)):
            if r & bit:
                self.reasons.add(name)

    def _run(self):
        while not self.stop_flag.is_set():
            try:
                self._poll()
            except Exception as e:  # noqa: BLE001
                self.err = str(e)
                return
            time.sleep(0.0005)

    def stop(self):
        if self.t is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [self.err], "samples": 0}
        self.stop_flag.set()
        self.t.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ---------------------------------------------------------------- CPU legs
def host_threads():
    """Threads the CPU legs use: every host core, in (primary, quality) pairs."""
    return max(1, min(os.cpu_count() or 1, 64))


def cpu_fuse(kf, idx, threads=2, voxel=VOXEL, trunc=TRUNC):
    """Reference semantics on the host: the C oracle port of
    TSDFVolume.integrate.  The reference is one Python thread under an RLock;
    to use every host core the port runs threads // 2 independent volume
    replicas (like that many reference processes), each fusing its share of the
    keyframes with the primary and the quality volume (C*q) on one thread each
    (ctypes releases the GIL).  Returns (events, seconds)."""
    from oracle import oracle as orc

    idx = list(idx)
    n_rep = max(1, min(threads // 2, len(idx)))
    staged = [orc.stage_keyframe(kf["X"][b], kf["C"][b], kf["pose"][b], kf["q"][b], H, W,
                                 MIN_CONF) for b in idx]
    vols = [(orc.OracleVolume(voxel, trunc), orc.OracleVolume(voxel, trunc))
            for _ in range(n_rep)]

    def run(r, which):
        for pts, conf, confq, org in staged[r::n_rep]:
            vols[r][which].integrate(pts, conf if which == 0 else confq, org)

    t0 = time.perf_counter()
    if threads >= 2:
        th = [threading.Thread(target=run, args=(r, w)) for r in range(n_rep) for w in (0, 1)]
        [t.start() for t in th]
        [t.join() for t in th]
    else:
        run(0, 0)
        run(0, 1)
    dt = time.perf_counter() - t0
    return sum(vp.events + vq.events for vp, vq in vols), dt


def cpu_sample_text(n_kf, threads, per_step):
    return (f"{n_kf} of the step's {per_step} keyframes per step; C port of "
            "TSDFVolume.integrate (oracle/tsdf_oracle.c); the reference is single-threaded, so "
            f"the host cores run {max(1, threads // 2)} independent volume replicas, one keyframe "
            "each, primary and quality volume on one thread each")


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the Python
    reference cannot travel to the GPU box) on the same workload and config,
    each step a bounded sample of the step's keyframes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as orc

    orc.build()
    threads = host_threads()
    sample_kf = max(2, threads // 2)
    per_step = args.kf_per_launch * args.launches_per_step * args.passes_per_step
    ev_total, t_total = 0, 0.0
    for s in range(args.warmup + args.steps):
        # a different launch of the pass every step, so the sample walks the orbit
        kf = gen_batch(TARGET, s % args.launches_per_step, sample_kf)
        ev, dt = cpu_fuse(kf, range(sample_kf), threads)
        if s >= args.warmup:
            ev_total += ev
            t_total += dt
    value = ev_total / t_total
    line = {
        "impl": "reference", "metric": "tsdf_voxel_updates_per_s", "value": value,
        "unit": "voxel updates/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "keyframes_per_s": sample_kf * args.steps / t_total,
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": "voxel updates/s", "cores": threads,
                         "kind": "port", "sample": cpu_sample_text(sample_kf, threads, per_step)},
        "e2e": {"value": value, "unit": "voxel updates/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "host_cores": os.cpu_count(),
    }
    emit(line)


def workload_config(args):
    """The SAME dict on both arms and at every N (the partition of the volume over the
    ranks is reported under `partition`, outside `config`)."""
    return {
        "workload": ("north_star Target: TUM fr1_room-shaped synthetic orbit of "
                     f"{args.kf_per_launch * args.launches_per_step} keyframes 512x384, fused "
                     f"{args.passes_per_step} times per step, "
                     "dual TSDF (primary + quality-weighted), 1024^3 @ 5 mm, trunc 20 mm"),
        "keyframes_per_step": args.kf_per_launch * args.launches_per_step * args.passes_per_step,
        "keyframes_per_launch": args.kf_per_launch,
        "launches_per_step": args.launches_per_step * args.passes_per_step,
        "passes_per_step": args.passes_per_step,
        "volume_dim": TARGET["dim"], "voxel_size": VOXEL, "truncation": TRUNC,
        "orbit_keyframes": ORBIT_TOTAL,
        "l2": "every launch reads a fresh 126 MB keyframe batch and a pass 1.9 GB "
              "(L2 = 126 MB): inputs larger than L2, no flush needed",
    }


# ---------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist

    from mast3r_slam_quality_dualtsdf_b200 import build, synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume
    from mast3r_slam_quality_dualtsdf_b200.tsdf.global_volume import pack_keyframes, packed_layout

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    if not os.path.exists(build.LIB):
        build.build()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime

        # a short collective timeout: a rank that dies must not hold the others for 10 minutes
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=240))
    wl = TARGET
    lo, hi = synth.bounds_for(wl["scene"], VOXEL, wl["dim"])
    bounds = (lo, hi)
    # N > 1: interleaved z-slabs (slab s -> rank s mod N); z is the vertical axis of the
    # orbit so the slabs cut every view (DESIGN.md "Multi-GPU")
    slabs = (world, rank, slab_width_for(world, wl["dim"])) if world > 1 else None
    vol = TSDFVolume(VOXEL, TRUNC, capacity_log2=wl["cap_log2"], bounds=bounds, slabs=slabs,
                     device=dev)
    scratch = TSDFVolume(VOXEL, TRUNC, capacity_log2=16, bounds=bounds, slabs=slabs, device=dev)
    B, L, P = args.kf_per_launch, args.launches_per_step, max(1, args.passes_per_step)
    if B % world:
        raise SystemExit(f"bench.py: --kf-per-launch {B} must be a multiple of the {world} ranks")
    share = B // world
    lay = packed_layout(share, H, W, GH, GW)
    seg_len = lay["seg_len"]

    # every rank holds ITS share of every launch: a packed segment in pinned host memory
    # (e2e source) and the same segment resident in its HBM (value source).  N = 1 also
    # keeps the four separate pinned arrays the reference-facing host entry takes.
    host_seg, host_sep, res_seg = [], [], []
    for l in range(L):
        kf = gen_batch(wl, l, B, offset=rank * share, count=share)
        seg = torch.zeros(seg_len, dtype=torch.float32).pin_memory()
        pack_keyframes(kf["X"], kf["C"], kf["pose"], kf["q"], out=seg)
        host_seg.append(seg)
        res_seg.append(seg.to(dev, non_blocking=True))
        if world == 1:
            # the four separate arrays the reference-facing host entry takes: the whole pass
            # in one pinned allocation per array, host_sep[l] = the views of launch l
            if not host_sep:
                host_all = {k: torch.empty((L * B,) + tuple(kf[k].shape[1:]),
                                           dtype=torch.float32).pin_memory()
                            for k in ("X", "C", "pose", "q")}
            for k in ("X", "C", "pose", "q"):
                host_all[k][l * B:(l + 1) * B].copy_(torch.from_numpy(kf[k]))
            host_sep.append({k: host_all[k][l * B:(l + 1) * B] for k in ("X", "C", "pose", "q")})
    torch.cuda.synchronize()

    main = torch.cuda.current_stream(dev)
    pull = world > 1 and args.ingest == "pull"
    peers = sh_pull = None
    if pull:
        from mast3r_slam_quality_dualtsdf_b200 import parallel

        # every rank's shares of the L launches in ONE buffer the other ranks map (CUDA IPC)
        peers = parallel.PeerKeyframes(L, share, H, W, GH, GW, device=dev)
        for l in range(L):
            peers.slot(l).copy_(res_seg[l])
        res_seg = [peers.slot(l) for l in range(L)]
        torch.cuda.synchronize()
        dist.barrier()
        sh_pull = {}

    def pull_volume(v):
        """ShardedTSDFVolume view of the TSDFVolume v (integrate_slot lives there)."""
        if id(v) not in sh_pull:
            from mast3r_slam_quality_dualtsdf_b200 import parallel

            sh_pull[id(v)] = parallel.ShardedTSDFVolume(
                VOXEL, TRUNC, slab_width=slabs[2], bounds=bounds, device=dev,
                volume_factory=lambda **kw: v)
        return sh_pull[id(v)]

    words = TSDFVolume.keyframe_meta_words(share, H, W, GH, GW) if pull else 0
    meta_bufs = [(torch.empty(words, dtype=torch.int32, device=dev),
                  torch.empty(words * world, dtype=torch.int32, device=dev)) for _ in range(2)] \
        if pull else None

    def fuse_packed(v, buf, n_seg):
        v.integrate_keyframes_packed(buf, n_seg, share, H, W, GH, GW, MIN_CONF)

    # N > 1: two all-gather targets, gather of launch l+1 overlaps the fuse of launch l
    full = [torch.empty(world * seg_len, dtype=torch.float32, device=dev) for _ in range(2)] \
        if world > 1 else None
    stage_seg = [torch.empty(seg_len, dtype=torch.float32, device=dev) for _ in range(2)] \
        if world > 1 else None
    gs = torch.cuda.Stream(dev) if world > 1 else None
    cs = torch.cuda.Stream(dev) if world > 1 else None
    fuse_done = [torch.cuda.Event() for _ in range(2)]
    copy_done = [torch.cuda.Event() for _ in range(2)]
    gath_done = [torch.cuda.Event() for _ in range(2)]
    slot_ready = [torch.cuda.Event() for _ in range(L)]

    def step_value(v):
        """One pass over the orbit from HBM-resident shares."""
        if world == 1:
            for l in range(L):
                fuse_packed(v, res_seg[l], 1)
            return
        if pull:  # meta record + ONE small all-gather + fuse pulling its tiles from the owners
            sv = pull_volume(v)
            if args.no_pipeline:
                for l in range(L):
                    sv.integrate_slot(peers, l, MIN_CONF, meta_bufs=meta_bufs[l & 1])
            else:  # meta + all-gather of launch l+1 behind the fuse of launch l
                sv.integrate_slots(peers, range(L), MIN_CONF, meta_bufs=meta_bufs)
            return
        for l in range(L + 1):
            b = l & 1
            if l < L:
                with torch.cuda.stream(gs):
                    gs.wait_event(fuse_done[b])  # launch l-2 has consumed full[b]
                    dist.all_gather_into_tensor(full[b], res_seg[l])
                    gath_done[b].record(gs)
            if l > 0:
                main.wait_event(gath_done[b ^ 1])
                fuse_packed(v, full[b ^ 1], world)
                fuse_done[b ^ 1].record(main)

    def step_e2e(v):
        """The same pass from pinned HOST buffers."""
        if world == 1 and args.e2e_per_launch:
            for l in range(L):  # one *_host call per 40-keyframe launch (A/B)
                t = host_sep[l]
                v.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], MIN_CONF, H, W)
        elif world == 1:
            # ONE *_host call for the pass: inside the C ABI call the pointmaps cross PCIe in
            # 40-keyframe chunks through a ring of staging slots, each chunk fused by its own
            # launch while the next one is copied; the call returns when the copies are done
            t = host_all
            v.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], MIN_CONF, H, W)
        elif pull:  # one H2D of my share per launch into its ring slot, then as step_value
            sv = pull_volume(v)
            if args.no_pipeline:
                for l in range(L):
                    with torch.cuda.stream(cs):
                        peers.slot(l).copy_(host_seg[l], non_blocking=True)
                        copy_done[l & 1].record(cs)
                    main.wait_event(copy_done[l & 1])
                    sv.integrate_slot(peers, l, MIN_CONF, meta_bufs=meta_bufs[l & 1])
            else:
                cs.wait_stream(main)
                with torch.cuda.stream(cs):
                    for l in range(L):
                        peers.slot(l).copy_(host_seg[l], non_blocking=True)
                        slot_ready[l].record(cs)
                sv.integrate_slots(peers, range(L), MIN_CONF, meta_bufs=meta_bufs,
                                   ready=slot_ready)
            v.counters()
            dist.barrier()  # every rank has pulled from my slots: the next step may overwrite them
            return
        else:  # one H2D + one all-gather + one fuse per launch, three-stage pipeline
            for l in range(L + 2):
                if l < L:
                    b = l & 1
                    with torch.cuda.stream(cs):
                        cs.wait_event(gath_done[b])  # gather l-2 has read stage_seg[b]
                        stage_seg[b].copy_(host_seg[l], non_blocking=True)
                        copy_done[b].record(cs)
                if 1 <= l <= L:
                    b = (l - 1) & 1
                    with torch.cuda.stream(gs):
                        gs.wait_event(copy_done[b])
                        gs.wait_event(fuse_done[b])
                        dist.all_gather_into_tensor(full[b], stage_seg[b])
                        gath_done[b].record(gs)
                if l >= 2:
                    b = l & 1
                    main.wait_event(gath_done[b])
                    fuse_packed(v, full[b], world)
                    fuse_done[b].record(main)
        v.counters()  # the step's result (8 x int64) read back; synchronises

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, v, steps, warmup, sample_clocks=True):
        v.reset()
        def step_fn(vv, one_pass=step_fn):
            for _ in range(P):
                one_pass(vv)
        for _ in range(warmup):
            step_fn(v)
        # started BEFORE the barrier: its start-up (NVML + a thread) on rank 0 would otherwise
        # sit inside the other ranks' timed region (they wait for rank 0 at the first collective)
        sampler = ClockSampler(local) if (rank == 0 and sample_clocks) else None
        c0 = v.counters()
        barrier()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        ev[0].record()
        for i in range(steps):
            step_fn(v)
            ev[i + 1].record()
        barrier()
        clocks = sampler.stop() if sampler else None
        c1 = v.counters()
        ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
        total_ms = ev[0].elapsed_time(ev[-1])
        upd = (c1["updates_primary"] - c0["updates_primary"]) + \
              (c1["updates_quality"] - c0["updates_quality"])
        tt = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        uu = torch.tensor([float(upd)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(uu, op=dist.ReduceOp.SUM)
        return float(tt.item()), float(uu.item()), ms, clocks, c1

    # ---- value: device-resident inputs (N > 1: incl. the NCCL keyframe all-gather)
    total_ms, updates, step_ms, clocks, ctr = timed(step_value, vol, args.steps, args.warmup)
    value = updates / (total_ms * 1e-3)
    # ---- e2e: host buffers -> (H2D, all-gather) -> fuse -> counters back
    e2e_ms, e2e_updates, _, _, _ = timed(step_e2e, vol, args.steps, args.warmup,
                                         sample_clocks=False)
    e2e_value = e2e_updates / (e2e_ms * 1e-3)
    n_kf_step = B * L * P

    # ---- algorithmic bytes per launch: voxels / blocks a launch's batch touches on this
    # rank (fused into an empty scratch volume)
    touched_v = touched_b = 0
    for l in range(L):
        scratch.reset()
        if world == 1:
            fuse_packed(scratch, res_seg[l], 1)
        elif pull:
            pull_volume(scratch).integrate_slot(peers, l, MIN_CONF, meta_bufs=meta_bufs[0])
        else:
            dist.all_gather_into_tensor(full[0], res_seg[l])
            fuse_packed(scratch, full[0], world)
        touched_v += scratch.stats(0)["total_voxels"]
        touched_b += scratch.counters()["blocks"]
    del scratch
    # the launch on this rank reads the whole batch and writes its slabs; with the peer pull
    # it reads its own share (meta pass), the gathered meta records and the tiles it lists
    # (bounded from above by the share of the batch's rays it does not cull)
    kf_read = B * KF_BYTES
    if pull:
        frac = 1.0 - ctr["culled_rays"] / max(1, ctr["rays"])
        kf_read = share * KF_BYTES + world * words * 4 + frac * B * KF_BYTES
    alg_bytes_launch = kf_read + (touched_v * 32 + touched_b * 8) / L
    launch_ms = float(np.mean(step_ms)) / (L * P)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes_launch / (launch_ms * 1e-3) / 1e9
    # DRAM traffic and issue statistics of the same kernel from the committed ncu --set full
    # capture (profiles/ncu_traffic.json, written by scripts/ncu_summary.py)
    traffic, issue = None, None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        k = tr["integrate_keyframes_kernel"]
        if B == KF_PER_LAUNCH and world == 1 and k.get("workload", "target") == "target":
            traffic = float(k["dram_bytes_per_launch"])
            if "warp_inst_per_launch" in k:
                issue = {"warp_inst_per_launch": k["warp_inst_per_launch"],
                         "issue_active_pct": k.get("issue_active_pct"),
                         "inst_per_update": k["warp_inst_per_launch"] * 32.0 /
                         (updates / (args.steps * L * P)),
                         "source": k.get("source")}
    except Exception:
        pass

    sections = {}
    if not args.no_sections:
        sections = secondary_sections(args, vol, dev, rank, world, peak, bounds, slabs)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": "tsdf_voxel_updates_per_s", "value": value, "unit": "voxel updates/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "keyframes_per_s": n_kf_step * args.steps / (total_ms * 1e-3),
        "timed_region_s": total_ms * 1e-3,
        "config": workload_config(args),
        "partition": "single volume" if world == 1 else
        f"interleaved z-slabs of {slab_width_for(world, wl['dim'])} voxel layers over {world} "
        f"ranks; every rank holds {share} keyframes of each launch; " +
        ("peer pull: per launch ONE NCCL all-gather of the meta records (tile z ranges, poses, "
         "quality grids: ~28 KB per keyframe) inside the timed region, then every rank's fuse "
         "kernel TMA-loads the tiles that reach its slabs from the owners' HBM over NVLink "
         "(CUDA IPC mapped buffers)" if pull else
         f"ONE packed NCCL all-gather (= {world} broadcasts over NVLink) per launch inside the "
         "timed region, double-buffered against the fuse"),
        "e2e": {"value": e2e_value, "unit": "voxel updates/s",
                "h2d_bytes_per_step": n_kf_step * KF_BYTES, "d2h_bytes_per_step": 64 * P,
                "keyframes_per_s": n_kf_step * args.steps / (e2e_ms * 1e-3),
                "ms_per_step": e2e_ms / args.steps,
                "ingest": (("b200tsdf_integrate_keyframes_host per launch (chunked H2D overlapped "
                            "with the fuse), one counters read per step" if args.e2e_per_launch else
                            "ONE b200tsdf_integrate_keyframes_host call per 600-keyframe pass: "
                            "inside it 40-keyframe chunks cross PCIe through a ring of three "
                            "staging slots, each fused by its own launch behind the next copy; "
                            "one counters read per step") if world == 1 else
                           ("per launch ONE packed H2D of the rank's share into its ring slot, "
                            "meta record + ONE small all-gather, fuse pulling its tiles from the "
                            "owners' slots; one counters read + one barrier per pass" if pull else
                            "per launch ONE packed H2D of the rank's share, ONE all-gather, one "
                            "fuse (classify + listed-tile march); copy / gather / fuse of "
                            "consecutive launches overlap on three streams; one counters read "
                            "per step"))},
        # kernels of this library inside the timed `value` region, all ranks: one fused launch
        # per 40 keyframes and rank; from 3 ranks up the classify pass + the listed-tile fuse
        "gpu_launches": args.steps * L * P * world * (3 if pull else 2 if world > 2 else 1),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "kernel": "integrate_keyframes_kernel" if world <= 2 and not pull else
                     "keyframe_meta_kernel + tiles_from_meta_kernel + "
                     "integrate_keyframes_listed_kernel" if pull else
                     "slab_classify_kernel + integrate_keyframes_listed_kernel",
                     "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                     "algorithmic_bytes_per_launch": alg_bytes_launch,
                     "launch_ms": launch_ms,
                     "binding": "instruction issue, not HBM (DESIGN.md 5): see `issue`",
                     "issue": issue},
        "clocks": clocks,
        "volume": {"blocks": ctr["blocks"], "dropped_updates": ctr["dropped"],
                   "dropped_means": "samples outside the bounds or, at N > 1, marched by a rank "
                                    "that does not own their voxel (the owner applies them)"},
        "host_cores": os.cpu_count(),
    }
    line.update(sections)
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as orc

        orc.build()
        threads = host_threads()
        n_cpu_kf = max(2, threads // 2)
        kf = gen_batch(wl, 0, n_cpu_kf)
        ev_c, dt_c = 0, 0.0
        while dt_c < 10.0:
            e, d = cpu_fuse(kf, range(n_cpu_kf), threads)
            ev_c += e
            dt_c += d
        line["cpu_baseline"] = {
            "value": ev_c / dt_c, "unit": "voxel updates/s", "cores": threads, "kind": "port",
            "sample": cpu_sample_text(n_cpu_kf, threads, n_kf_step) + ", repeated for >= 10 s; "
            "the reference's own single-threaded Python loop measured 12.9 k updates/s "
            "(BASELINE.md)"}
        # BASELINE.json configs[0]: the reference's own CPU-runnable case
        from mast3r_slam_quality_dualtsdf_b200 import synth as sy

        k1 = sy.make_keyframes(sy.SCENES["room"], 16, H, W, seed=0, orbit_radius=0.6, gh=GH, gw=GW,
                               start=100, total=ORBIT_TOTAL)
        e1, d1 = cpu_fuse(k1, range(16), threads, voxel=0.01, trunc=0.04)
        line["cpu_baseline"]["config1"] = {
            "workload": "BASELINE configs[0]: 16 keyframes 512x384 into 256^3 @ 1 cm, dual volume",
            "value": e1 / d1, "unit": "voxel updates/s", "seconds": d1, "cores": threads}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def secondary_sections(args, vol, dev, rank, world, peak, bounds, slabs):
    """Kernels beside the headline, on the volume the e2e loop left behind (this rank's
    slabs): marching cubes of both volumes (north_star kernel 4; N > 1: incl. the seam
    exchange over NCCL), the stats sweep, the quality producer, BASELINE config 2 at N = 1,
    re-integration after a pose update (config 5) and the drop-in manager."""
    import torch
    import torch.distributed as dist

    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    out = {}

    def dev_ms(fn, reps=3, collective=True):
        """Best-of-reps device time of fn(); collective: every rank calls it (barrier
        before each repetition, max over ranks); else a rank-local measurement."""
        best = None
        res = None
        for _ in range(reps):
            torch.cuda.synchronize()
            if world > 1 and collective:
                dist.barrier()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            res = fn()
            b.record()
            torch.cuda.synchronize()
            t = a.elapsed_time(b)
            best = t if best is None else min(best, t)
        if world > 1 and collective:
            tt = torch.tensor([best], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            best = float(tt.item())
        return best, res

    def allsum(x):
        if world == 1:
            return x
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        return float(t.item())

    # ---- extract: BASELINE configs[2] -- marching cubes of BOTH volumes
    sharded = None
    if world > 1:
        from mast3r_slam_quality_dualtsdf_b200 import parallel

        sharded = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=slabs[2], bounds=bounds,
                                             device=dev, volume_factory=lambda **kw: vol)
    nblk = vol.counters()["blocks"]
    seam_ms = 0.0
    if sharded is not None:
        seam_ms, halo = dev_ms(sharded.exchange_seams, reps=2)
    extract = {"kernel": "mc_neighbors_kernel + mc_kernel3<write> (one warp per voxel block), one "
                         "optimistic write pass per volume",
               "ms_means": "device time of one pass (CUDA events around %d passes queued back to "
                           "back through b200tsdf_extract_mesh_async, count left on the device); "
                           "api_ms = one synchronous b200tsdf_extract_mesh call incl. the read-back "
                           "of the triangle count and the host wake-up" % 8, "volumes": {}}
    tot_ms = 0.0
    for which, name in ((0, "primary"), (1, "quality")):
        verts = vol.extract_mesh(which)
        n_tris = int(verts.shape[0] // 3)
        buf = torch.empty(max(n_tris, 1) * 3, 3, dtype=torch.float32, device=dev)
        del verts
        api_ms, _ = dev_ms(lambda w=which, b=buf: vol.extract_mesh(w, out=b))
        cnt = torch.empty(1, dtype=torch.int64, device=dev)

        def passes(w=which, b=buf, c=cnt):
            for _ in range(8):
                vol.extract_mesh_async(w, b, c)
        mc_ms, _ = dev_ms(passes)
        mc_ms /= 8
        assert int(cnt.item()) == n_tris
        tot_ms += mc_ms
        vox_bytes = nblk * 512 * 24  # packed voxel (16 B) + hit counters (8 B)
        alg = vox_bytes + n_tris * 36
        extract["volumes"][name] = {
            "ms": mc_ms, "api_ms": api_ms, "triangles": allsum(n_tris), "blocks": allsum(nblk),
            "triangles_per_s": allsum(n_tris) / (mc_ms * 1e-3),
            "algorithmic_bytes_this_rank": alg,
            "achieved_gbs": alg / (mc_ms * 1e-3) / 1e9, "frac": alg / (mc_ms * 1e-3) / 1e9 / peak}
        del buf
    extract["ms_both_volumes"] = tot_ms
    extract["l2"] = "volume of this rank %.0f MB vs 126 MB L2" % (nblk * 512 * 24 / 1e6)
    if sharded is not None:
        extract["seam_exchange_ms"] = seam_ms
        extract["seam_exchange"] = ("export_seam_blocks -> NCCL isend/irecv to the rank below -> "
                                    "import_halo_blocks; max over ranks")
    out["extract"] = extract
    st_ms, st = dev_ms(lambda: vol.stats(0))
    vox_bytes = nblk * 512 * 24
    out["sweep"] = {"kernel": "stats_kernel", "ms": st_ms, "algorithmic_bytes_this_rank": vox_bytes,
                    "achieved_gbs": vox_bytes / (st_ms * 1e-3) / 1e9,
                    "frac": vox_bytes / (st_ms * 1e-3) / 1e9 / peak,
                    "valid_voxels": allsum(st["valid_voxels"])}

    # ---- reintegrate: BASELINE configs[4] -- after a backend solve every stored keyframe of
    # the solve window is fused again at its new pose (global_manager.py:214-229 -> :71-79)
    try:
        out["reintegrate"] = reintegrate_section(args, dev, rank, world, bounds, slabs, dev_ms)
    except Exception as e:  # noqa: BLE001
        out["reintegrate"] = {"error": repr(e)}

    if rank != 0:
        return out
    # ---- local ROI path of the refiner
    try:
        out["local"] = local_section(args, dev, lambda fn, reps=3: dev_ms(fn, reps, False))
    except Exception as e:  # noqa: BLE001
        out["local"] = {"error": repr(e)}
    # ---- quality producer: per-patch grids + classify on the device
    try:
        out["quality"] = quality_section(args, dev, lambda fn, reps=3: dev_ms(fn, reps, False))
    except Exception as e:  # noqa: BLE001
        out["quality"] = {"error": repr(e)}
    if world == 1:
        try:
            out["config2"] = config2_section(args, dev)
        except Exception as e:  # noqa: BLE001
            out["config2"] = {"error": repr(e)}
        try:
            out["manager"] = manager_section(args, dev)
        except Exception as e:  # noqa: BLE001
            out["manager"] = {"error": repr(e)}
    return out


def reintegrate_section(args, dev, rank, world, bounds, slabs, dev_ms):
    """Config 5: 120 keyframes are ingested with ids (every rank keeps them in its store),
    then a backend solve moves all of them: only the 32-byte poses travel (one broadcast),
    every rank re-fuses its slabs from its own copy -- additive (the reference's semantics)
    and replace (de-integrate at the old pose first)."""
    import torch

    from mast3r_slam_quality_dualtsdf_b200 import parallel

    n_kf = 120
    sh = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=(slabs[2] if slabs else 256),
                                    bounds=bounds, capacity_log2=TARGET["cap_log2"], device=dev)
    for a in range(0, n_kf, 40):
        kf = gen_batch(TARGET, a // 40, 40)
        t = {k: torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")}
        sh.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], MIN_CONF, H, W,
                               broadcast=False, ids=list(range(a, a + 40)))
    ids = list(range(n_kf))
    poses = torch.stack([sh._store[i]["pose"] for i in ids]).clone()
    poses[:, :3] += 0.004  # the solve shifted every keyframe by 4 mm
    res = {}
    for mode in ("additive", "replace"):
        ms, _ = dev_ms(lambda m=mode: sh.reintegrate(ids, poses, src=0, mode=m), reps=2)
        res[mode] = {"ms": ms, "keyframes_per_s": n_kf / (ms * 1e-3)}
    res["keyframes"] = n_kf
    res["what"] = ("ShardedTSDFVolume.reintegrate of all stored keyframes after a pose update: "
                   "32 B per keyframe broadcast, pointmaps from every rank's own store")
    del sh
    return res


def local_section(args, dev, dev_ms):
    """the refiner's per-ROI dense TSDF (tsdf_refine.py:837-1064) --
    `build_tsdf_robust` + `extract_surface_safe` for 64x64-pixel blocks of one full-size
    Target keyframe (ROI = the block's points + the refiner's 4 cm margin, 2 cm voxels,
    <= 64^3 grid, 100 ray-cast pixels per block as the reference).  Latency-bound by
    construction: reported as blocks/s through the Python mirror."""
    import numpy as np
    import torch

    from mast3r_slam_quality_dualtsdf_b200 import local_tsdf
    from mast3r_slam_quality_dualtsdf_b200.tsdf import sim3

    cfg = {"voxel_size": 0.02, "trunc_dist": 0.08, "max_grid_dim": 64, "roi_size": 0.4,
           "ray_samples": 64, "max_displacement": 0.015, "min_weight_threshold": 0.01,
           "min_confidence": 0.2}
    kf = gen_batch(TARGET, 3, 1)
    X = torch.from_numpy(kf["X"][0]).to(dev)
    Cf = torch.from_numpy(kf["C"][0]).to(dev)
    # the refiner takes the ROI from the block's camera-frame points (+ margin,
    # tsdf_refine.py:715-716) and hands the kernels the keyframe's pose; with the identity
    # pose the two frames coincide and every block has its surface inside its ROI
    pose = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float32, device=dev)
    Xw = sim3.act(pose, X).reshape(H, W, 3)
    T = type("Pose", (), {"data": pose.reshape(1, 8).cpu()})()
    blocks = []
    for y0 in range(32, H - 64, 96):
        for x0 in range(32, W - 64, 96):
            pts = Xw[y0:y0 + 64, x0:x0 + 64].reshape(-1, 3)
            lo, hi = pts.min(0).values - 0.04, pts.max(0).values + 0.04
            mask = torch.zeros(H, W, dtype=torch.bool, device=dev)
            mask[y0:y0 + 64, x0:x0 + 64] = True
            blocks.append((lo.cpu(), hi.cpu(), mask.reshape(-1)))
    order = torch.arange(100)
    hits_total = [0]

    def run():
        hits_total[0] = 0
        for lo, hi, mask in blocks:
            tsdf, w = local_tsdf.build_tsdf_robust(cfg, X, Cf.unsqueeze(-1), None, lo, hi, H, W, T)
            _, hits = local_tsdf.extract_surface_safe(cfg, tsdf, lo, hi, None, mask, H, W, X,
                                                      order=order)
            hits_total[0] = hits_total[0] + hits.sum()
        return hits_total[0]

    run()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    hits = run()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    ms, _ = dev_ms(run, reps=2)
    dims = local_tsdf.grid_dims(cfg, blocks[0][0], blocks[0][1])
    return {"kernels": "local_accumulate_kernel + local_finalize_kernel + local_raycast_kernel",
            "blocks": len(blocks), "blocks_per_s": len(blocks) / dt,
            "us_per_block_wall": dt / len(blocks) * 1e6, "us_per_block_device": ms * 1e3 / len(blocks),
            "grid_of_first_block": list(dims), "raycast_hits": int(hits),
            "what": "TSDFRefiner numerics per 64x64-pixel block of a 512x384 keyframe through "
                    "local_tsdf.build_tsdf_robust + extract_surface_safe (196 608 points "
                    "tested against the ROI per block, <= 64^3 voxels, 100 rays); wall clock "
                    "incl. the Python mirror, device = CUDA events around the same calls"}


def quality_section(args, dev, dev_ms):
    import torch

    from mast3r_slam_quality_dualtsdf_b200 import quality_core as qc

    B = args.kf_per_launch
    gq = torch.Generator(device="cpu").manual_seed(1)
    qjob = {"H": H, "W": W, "kf_id": 0, "t_norm": torch.tensor(0.05), "theta": torch.tensor(0.1),
            "valid_kf": (torch.rand(H * W, generator=gq) > 0.3).to(dev),
            "r_pix": torch.rand(H * W, generator=gq).to(dev),
            "Ck": (torch.rand(H * W, generator=gq) * 3).to(dev),
            "Qk": (torch.rand(H * W, generator=gq) * 2).to(dev)}
    qargs = dict(ps=16, alpha=0.8, b0=0.15, theta0=0.1745, C_thr=1.5, Q_thr=1.5, thr_zr=1.0,
                 thr_zu=1.0, thr_dc=0.02, device=dev)
    q_ms1, _ = dev_ms(lambda: [qc.compute_batch([qjob], **qargs) for _ in range(20)])
    qjobs = [dict(qjob, kf_id=i) for i in range(B)]
    qc.compute_batch(qjobs, **qargs)
    q_ms, _ = dev_ms(lambda: [qc.compute_batch(qjobs, **qargs) for _ in range(5)])
    st = {k: torch.stack([j[k] for j in qjobs]) for k in ("Ck", "Qk", "r_pix")}
    st["valid"] = torch.stack([j["valid_kf"] for j in qjobs]).to(torch.uint8)
    tn, th = (torch.stack([j[k] for j in qjobs]).to(dev) for k in ("t_norm", "theta"))
    sargs = {k: v for k, v in qargs.items() if k != "device"}
    run_st = lambda: qc.compute_stacked(st["Ck"], st["Qk"], st["r_pix"], st["valid"], tn, th,  # noqa: E731
                                        None, H, W, device=dev, **sargs)
    run_st()
    q_ms_st, _ = dev_ms(lambda: [run_st() for _ in range(5)])
    q_bytes = H * W * 13  # C, Q, r_pix float32 + valid u8, read once
    return {"kernels": "quality_patch_kernel + quality_finish_batch_kernel (2 launches per batch)",
            "keyframes_per_batch": B,
            "us_per_keyframe": 1e3 * q_ms / 5 / B,
            "us_per_keyframe_batch_of_1": 1e3 * q_ms1 / 20,
            "us_per_keyframe_stacked_inputs": 1e3 * q_ms_st / 5 / B,
            "stacked_gbs": q_bytes * B * 5 / (q_ms_st * 1e-3) / 1e9,
            "algorithmic_bytes": q_bytes,
            "note": "through quality_core.compute_batch incl. the stacking of the job tensors; a "
                    "batch of 1 is launch/latency-bound"}


def config2_section(args, dev):
    """BASELINE configs[1]: the same 600-keyframe pass into 512^3 @ 5 mm (the round-1
    headline workload), device-resident inputs, 3 timed passes."""
    import torch

    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

    wl = CONFIG2
    lo, hi = synth.bounds_for(wl["scene"], VOXEL, wl["dim"])
    v = TSDFVolume(VOXEL, TRUNC, capacity_log2=wl["cap_log2"], bounds=(lo, hi), device=dev)
    B, L = args.kf_per_launch, args.launches_per_step
    res = []
    for l in range(L):
        kf = gen_batch(wl, l, B)
        res.append({k: torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose", "q")})

    def one_pass():
        for t in res:
            v.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], MIN_CONF, H, W)

    one_pass()
    torch.cuda.synchronize()
    c0 = v.counters()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3):
        one_pass()
    b.record()
    torch.cuda.synchronize()
    c1 = v.counters()
    ms = a.elapsed_time(b)
    upd = (c1["updates_primary"] - c0["updates_primary"]) + \
          (c1["updates_quality"] - c0["updates_quality"])
    return {"workload": "BASELINE configs[1]: 600 keyframes 512x384 into 512^3 @ 5 mm, 1 B200",
            "value": upd / (ms * 1e-3), "unit": "voxel updates/s",
            "keyframes_per_s": 3 * B * L / (ms * 1e-3), "ms_per_launch": ms / (3 * L),
            "blocks": c1["blocks"]}


def manager_section(args, dev):
    """keyframes/s through the drop-in TSDFGlobalManager (the API main.py:80-90 calls) with
    device-resident keyframes, against the raw kernel rate of the line."""
    import torch

    from mast3r_slam_quality_dualtsdf_b200 import synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import global_manager as gm

    if not hasattr(gm, "bench_manager"):
        return {"skipped": "global_manager.bench_manager not built yet"}
    return gm.bench_manager(TARGET, dev, n_kf=ORBIT_TOTAL, H=H, W=W, gh=GH, gw=GW, voxel=VOXEL,
                            trunc=TRUNC, min_conf=MIN_CONF, gen_batch=gen_batch)


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else this process
    or its libraries print (NCCL writes its version banner to fd 1) goes to
    stderr."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1

if __name__ == "__main__":
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
