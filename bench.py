#!/usr/bin/env python
"""bench.py -- dual-TSDF fusion throughput on B200 (BASELINE.json metric).

A step = one launch of the hot path over one batch of KF_PER_STEP synthetic
512x384 keyframes (TUM fr1_room-shaped orbit inside a room box) fused into a
512^3 @ 5 mm dual volume (N=1, BASELINE.json configs[1]) or into the z-slabs
of a 1024^3 @ 5 mm volume (N>1, configs[3]; rank 0 NCCL-broadcasts each batch).

  value : voxel updates/s (primary + quality volume), inputs resident in HBM
  e2e   : same metric through the reference-facing TSDFVolume API with HOST
          (pinned) keyframe buffers: H2D inside the timed region + D2H of the
          step's counters
  roofline / cpu_baseline : see DESIGN.md "Measurement"

`--impl reference` times the CPU oracle port of the reference path (the
reference itself is Python and does not travel to the GPU box).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, GH, GW = 384, 512, 24, 32
VOXEL, TRUNC = 0.005, 0.02
MIN_CONF = 0.05
KF_PER_STEP = 40
ORBIT_TOTAL = 600
# algorithmic bytes per keyframe input: X + C + pose + quality grid
KF_BYTES = H * W * 12 + H * W * 4 + 32 + GH * GW * 4


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=15)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kf-per-step", type=int, default=KF_PER_STEP)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=-1,
                    help="N > 1: pieces the per-rank share is cut into for the pipelined "
                         "PCIe -> all-gather -> fuse ingest (0: serial copy, gather, fuse; "
                         "-1: 2 pieces when a rank ingests >= 16 keyframes per step, else serial "
                         "-- small shares are launch-latency-bound, profiles/r01j_e2e_chunks.txt)")
    return ap.parse_args()


def scene_for(n_gpus):
    from mast3r_slam_quality_dualtsdf_b200 import synth

    if n_gpus == 1:
        return dict(synth.SCENES["room"]), 512, 0.6
    return {"kind": "room", "half": 2.4874}, 1024, 1.2


SLAB_WIDTH = 32  # voxel layers per interleaved z-slab at N = 8: 16 cm at 5 mm


def slab_width_for(n_gpus, dim=1024):
    """Four slabs per rank: wide enough that few rays straddle a seam, interleaved
    enough that every view loads all ranks (measured: scripts/slab_rank_time.py)."""
    return max(8, dim // (4 * max(1, n_gpus)))


def gen_batch(scene, orbit_r, step_idx, kf_per_step, seed=0, up_axis="y", offset=0, count=None):
    """Keyframes [offset, offset + count) of step `step_idx`'s batch (default: all)."""
    from mast3r_slam_quality_dualtsdf_b200 import synth

    start = (step_idx * kf_per_step) % ORBIT_TOTAL + offset
    return synth.make_keyframes(scene, count or kf_per_step, H, W, seed=seed,
                                orbit_radius=orbit_r, gh=GH, gw=GW, start=start,
                                total=ORBIT_TOTAL, up_axis=up_axis)


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region.  The region
    lasts milliseconds, so NVML is polled from a thread (~1 kHz); nvidia-smi
    -lms 100 would return no sample at all."""

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


def cpu_fuse(kf, idx, threads=2):
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
    vols = [(orc.OracleVolume(VOXEL, TRUNC), orc.OracleVolume(VOXEL, TRUNC))
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


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the Python
    reference cannot travel to the GPU box) on the same workload, each step a
    bounded sample of the step's keyframes."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as orc

    orc.build()
    scene, dim, orbit_r = scene_for(args.gpus)
    threads = host_threads()
    sample_kf = max(2, threads // 2)
    ev_total, t_total = 0, 0.0
    for s in range(args.warmup + args.steps):
        kf = gen_batch(scene, orbit_r, s, sample_kf, up_axis="z" if args.gpus > 1 else "y")
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
        "config": workload_config(args.gpus, dim, sample_kf),
        "cpu_baseline": {"value": value, "unit": "voxel updates/s", "cores": threads,
                         "kind": "port",
                         "sample": f"{sample_kf} keyframes per step (of {args.kf_per_step}), "
                                   "C port of TSDFVolume.integrate; the reference is "
                                   "single-threaded, so the host cores run "
                                   f"{max(1, threads // 2)} independent volume replicas, one "
                                   "keyframe each, primary and quality volume on one thread "
                                   "each"},
        "e2e": {"value": value, "unit": "voxel updates/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "host_cores": os.cpu_count(),
    }
    emit(line)


def workload_config(n_gpus, dim, kf_per_step):
    return {
        "workload": ("TUM fr1_room-shaped synthetic orbit, 512x384 keyframes, dual TSDF "
                     f"(primary + quality-weighted), {dim}^3 @ 5 mm, trunc 20 mm"),
        "keyframes_per_step": kf_per_step, "volume_dim": dim, "voxel_size": VOXEL,
        "truncation": TRUNC, "orbit_keyframes": ORBIT_TOTAL,
        "partition": "single volume" if n_gpus == 1 else
        f"interleaved z-slabs of {slab_width_for(n_gpus, dim)} voxel layers over {n_gpus} ranks; every rank ingests "
        f"{kf_per_step // max(1, n_gpus)} keyframes per step, NCCL all-gather (= {n_gpus} broadcasts) over NVLink",
        "l2": "each step reads a fresh keyframe batch larger than L2 (157 MB > 126 MB)",
    }


# ---------------------------------------------------------------- GPU arm
def run_b200(args):
    import torch
    import torch.distributed as dist

    from mast3r_slam_quality_dualtsdf_b200 import build, synth
    from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume

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
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world
    scene, dim, orbit_r = scene_for(n_gpus)
    lo, hi = synth.bounds_for(scene, VOXEL, dim)
    # N > 1: interleaved z-slabs (slab s -> rank s mod N); z is the vertical axis of the
    # orbit so the slabs cut every view (DESIGN.md "Multi-GPU")
    bounds = (lo, hi)
    slabs = (world, rank, slab_width_for(world, dim)) if world > 1 else None
    up_axis = "z" if world > 1 else "y"
    cap = 18 if world == 1 else 19
    vol = TSDFVolume(VOXEL, TRUNC, capacity_log2=cap, bounds=bounds, slabs=slabs, device=dev)
    scratch = TSDFVolume(VOXEL, TRUNC, capacity_log2=16, bounds=bounds, slabs=slabs, device=dev)
    B = args.kf_per_step
    n_steps = args.warmup + args.steps
    names = ("X", "C", "q", "pose")
    shapes = {"X": (B, H * W, 3), "C": (B, H * W), "q": (B, GH, GW), "pose": (B, 8)}

    # N = 1: the whole batch in pinned host memory.  N > 1: sharded ingest -- every rank
    # holds B / N keyframes of each batch (pinned host memory, its own PCIe link); the N
    # simultaneous NCCL broadcasts run as one all-gather over NVLink
    if B % world:
        raise SystemExit(f"bench.py: --kf-per-step {B} must be a multiple of the {world} ranks")
    share = B // world
    host = []
    for s in range(n_steps):
        kf = gen_batch(scene, orbit_r, s, B, up_axis=up_axis, offset=rank * share, count=share)
        host.append({k: torch.from_numpy(kf[k]).pin_memory() for k in names})
    local_shapes = {k: (share,) + shapes[k][1:] for k in names}

    def fuse(v, t):
        v.integrate_keyframes(t["X"], t["C"], t["pose"], t["q"], MIN_CONF, H, W)

    def gather(full, part):
        for k in names:
            dist.all_gather_into_tensor(full[k], part[k])

    # resident copies on every rank (the `value` loop starts from HBM)
    resident = []
    for s in range(n_steps):
        t = {k: torch.empty(shapes[k], dtype=torch.float32, device=dev) for k in names}
        if world == 1:
            for k in names:
                t[k].copy_(host[s][k], non_blocking=True)
        else:
            part = {k: host[s][k].to(dev, non_blocking=True) for k in names}
            gather(t, part)
        resident.append(t)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(loop_body, reset):
        reset()
        for s in range(args.warmup):
            loop_body(s)
        barrier()
        c0 = vol.counters()
        sampler = ClockSampler(local) if rank == 0 else None
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
        ev[0].record()
        for i in range(args.steps):
            loop_body(args.warmup + i)
            ev[i + 1].record()
        barrier()
        clocks = sampler.stop() if sampler else None
        c1 = vol.counters()
        ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
        total_ms = ev[0].elapsed_time(ev[-1])
        upd = (c1["updates_primary"] - c0["updates_primary"]) + \
              (c1["updates_quality"] - c0["updates_quality"])
        tt = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        uu = torch.tensor([float(upd)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(uu, op=dist.ReduceOp.SUM)
        return float(tt.item()), float(uu.item()), ms, clocks, c1

    # ---- value: device-resident inputs, one launch per step
    total_ms, updates, step_ms, clocks, ctr = timed(lambda s: fuse(vol, resident[s]), vol.reset)
    value = updates / (total_ms * 1e-3)

    # ---- e2e: host buffers -> (H2D, all-gather) -> fuse -> counters back
    sharded = None
    if args.e2e_chunks < 0:
        args.e2e_chunks = 2 if share >= 16 else 0
    if world > 1:
        from mast3r_slam_quality_dualtsdf_b200 import parallel

        sharded = parallel.ShardedTSDFVolume(VOXEL, TRUNC, slab_width=slabs[2], bounds=bounds,
                                             device=dev, volume_factory=lambda **kw: vol)
    stage = {k: torch.empty(shapes[k], dtype=torch.float32, device=dev) for k in names}
    stage_local = {k: torch.empty(local_shapes[k], dtype=torch.float32, device=dev) for k in names}
    d2h = [0]

    def e2e_body(s):
        if world == 1:
            fuse(vol, host[s])  # *_host entry: H2D inside the C ABI call
        else:
            if args.e2e_chunks > 0:  # pipelined PCIe -> NVLink all-gather -> fuse
                h = host[s]
                sharded.integrate_keyframes_from_host(h["X"], h["C"], h["pose"], h["q"], MIN_CONF,
                                                      H, W, chunks=args.e2e_chunks)
            else:
                for k in names:
                    stage_local[k].copy_(host[s][k], non_blocking=True)
                gather(stage, stage_local)
                fuse(vol, stage)
        vol.counters()  # the step's result (8 x int64) read back, synchronises
        d2h[0] = 64

    e2e_ms, e2e_updates, _, _, _ = timed(e2e_body, vol.reset)
    e2e_value = e2e_updates / (e2e_ms * 1e-3)

    # ---- algorithmic bytes per launch: voxels / blocks a step's batch touches
    touched_v, touched_b = [], []
    for i in range(args.steps):
        scratch.reset()
        fuse(scratch, resident[args.warmup + i])
        st = scratch.stats(0)
        cc = scratch.counters()
        touched_v.append(st["total_voxels"])
        touched_b.append(cc["blocks"])
    tv = torch.tensor([float(sum(touched_v)), float(sum(touched_b))], dtype=torch.float64,
                      device=dev)
    if world > 1:
        dist.all_reduce(tv, op=dist.ReduceOp.SUM)
    alg_bytes_step = (B * KF_BYTES * world + (tv[0].item() * 32 + tv[1].item() * 8) / args.steps)
    # per-rank kernel: the launch on this rank reads the whole batch, writes its slab
    alg_bytes_launch = B * KF_BYTES + (sum(touched_v) * 32 + sum(touched_b) * 8) / args.steps
    launch_ms = float(np.mean(step_ms))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes_launch / (launch_ms * 1e-3) / 1e9
    # DRAM traffic of the same kernel from the committed ncu --set full capture
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if B == KF_PER_STEP and world == 1:
            traffic = float(tr["integrate_keyframes_kernel"]["dram_bytes_per_launch"])
    except Exception:
        pass

    # ---- secondary kernels on the volume the e2e loop left behind (this rank's slabs):
    # marching cubes (north_star kernel 4) and the stats sweep, both streaming / HBM-bound
    def dev_ms(fn, reps=3):
        best = None
        for _ in range(reps):
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            out = fn()
            b.record()
            torch.cuda.synchronize()
            t = a.elapsed_time(b)
            best = t if best is None else min(best, t)
        return best, out

    nblk = vol.counters()["blocks"]
    verts = vol.extract_mesh(0)
    n_tris = int(verts.shape[0] // 3)
    buf = torch.empty(max(n_tris, 1) * 3, 3, dtype=torch.float32, device=dev)
    del verts
    mc_ms, _ = dev_ms(lambda: vol.extract_mesh(0, out=buf))  # one write pass, buffer reused
    st_ms, st = dev_ms(lambda: vol.stats(0))
    vox_bytes = nblk * 512 * 24  # packed voxel (16 B) + hit counters (8 B)
    extract = {
        "kernel": "mc_kernel<write> (single pass, incl. counter memset + D2H of the count)", "blocks": nblk, "triangles": n_tris,
        "ms": mc_ms, "voxels_per_s": nblk * 512 / (mc_ms * 1e-3),
        "triangles_per_s": n_tris / (mc_ms * 1e-3),
        "algorithmic_bytes": vox_bytes + n_tris * 36,
        "achieved_gbs": (vox_bytes + n_tris * 36) / (mc_ms * 1e-3) / 1e9,
        "frac": (vox_bytes + n_tris * 36) / (mc_ms * 1e-3) / 1e9 / peak,
        "l2": "volume %.0f MB vs 126 MB L2" % (vox_bytes / 1e6),
    }
    sweep = {"kernel": "stats_kernel", "ms": st_ms, "algorithmic_bytes": vox_bytes,
             "achieved_gbs": vox_bytes / (st_ms * 1e-3) / 1e9,
             "frac": vox_bytes / (st_ms * 1e-3) / 1e9 / peak, "valid_voxels": st["valid_voxels"]}

    # ---- quality producer: per-patch grids + classify on the device
    quality = None
    if rank == 0:
        from mast3r_slam_quality_dualtsdf_b200 import quality_core as qc

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
        quality = {"kernels": "quality_patch_kernel + quality_finish_batch_kernel (2 launches per batch)",
                   "keyframes_per_batch": B,
                   "us_per_keyframe": 1e3 * q_ms / 5 / B,
                   "us_per_keyframe_batch_of_1": 1e3 * q_ms1 / 20,
                   "us_per_keyframe_stacked_inputs": 1e3 * q_ms_st / 5 / B,
                   "stacked_gbs": q_bytes * B * 5 / (q_ms_st * 1e-3) / 1e9,
                   "algorithmic_bytes": q_bytes,
                   "note": "through quality_core.compute_batch incl. the stacking of the job "
                           "tensors; a batch of 1 is launch/latency-bound"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": "tsdf_voxel_updates_per_s", "value": value, "unit": "voxel updates/s",
        "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "keyframes_per_s": B * args.steps / (total_ms * 1e-3),
        "config": workload_config(n_gpus, dim, B),
        "e2e": {"value": e2e_value, "unit": "voxel updates/s",
                "h2d_bytes_per_step": B * KF_BYTES, "d2h_bytes_per_step": d2h[0],
                "keyframes_per_s": B * args.steps / (e2e_ms * 1e-3),
                "ms_per_step": e2e_ms / args.steps,
                "ingest": ("b200tsdf_integrate_keyframes_host (chunked H2D overlapped with the "
                           "fuse)" if world == 1 else
                           f"pipelined, {args.e2e_chunks} pieces (H2D || all-gather || fuse)"
                           if args.e2e_chunks > 0 else "serial H2D, all-gather, fuse")},
        # kernels of this library inside the timed `value` region, all ranks: one fused launch
        # per step and rank; from 3 ranks up the classify pass + the listed-tile fuse
        "gpu_launches": args.steps * world * (2 if world > 2 else 1),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic,
                     "kernel": "integrate_keyframes_kernel" if world <= 2 else
                     "slab_classify_kernel + integrate_keyframes_listed_kernel",
                     "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback",
                     "algorithmic_bytes_per_launch": alg_bytes_launch,
                     "launch_ms": launch_ms},
        "clocks": clocks,
        "extract": extract, "sweep": sweep, "quality": quality,
        "volume": {"blocks": ctr["blocks"], "dropped_updates": ctr["dropped"],
                   "dropped_means": "samples outside the bounds or, at N > 1, marched by a rank "
                                    "that does not own their voxel (the owner applies them)"},
        "host_cores": os.cpu_count(),
    }
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as orc

        orc.build()
        threads = host_threads()
        n_cpu_kf = max(2, threads // 2)
        kf = gen_batch(scene, orbit_r, args.warmup, n_cpu_kf)
        ev_c, dt_c = 0, 0.0
        while dt_c < 10.0:
            e, d = cpu_fuse(kf, range(n_cpu_kf), threads)
            ev_c += e
            dt_c += d
        line["cpu_baseline"] = {
            "value": ev_c / dt_c, "unit": "voxel updates/s", "cores": threads, "kind": "port",
            "sample": f"{n_cpu_kf} keyframes of the first timed step, repeated for >= 10 s; C "
                      "port of TSDFVolume.integrate (oracle/tsdf_oracle.c) as "
                      f"{max(1, threads // 2)} independent volume replicas (one keyframe each), "
                      "primary and quality volume on one thread each; the reference's own "
                      "single-threaded Python loop measured 12.9 k updates/s (BASELINE.md)"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


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
