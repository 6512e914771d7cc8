"""One process per GPU: the dual TSDF volume sharded in interleaved z-slabs
(BASELINE north_star "z-slabs across the GPUs of one box").

No reference counterpart: the reference is single-process Python.  Layout:
  * world-z is cut into slabs of `slab_width` voxel layers; slab s belongs to
    rank s mod world_size (slab_width = dim / world gives the classic
    contiguous z-slabs; a smaller power of two deals slabs round-robin and
    balances surfaces that are not uniform in z);
  * every keyframe batch (pointmaps, confidences, quality grid, poses) is
    broadcast from the rank that holds it (NCCL over NVLink), or all-gathered
    when every rank ingests a share of the batch; every rank runs
    the staging of the full batch, culls the rays whose +-truncation band
    misses its slabs and fuses the rest into its own voxel-block hash -- there
    is no other data-path collective;
  * meshes are stitched at slab seams: before marching cubes every rank sends
    the first block layer of each of its slabs to the rank below, which
    installs it as read-only halo blocks; the lower rank emits the seam cubes,
    so no triangle is duplicated or missing;
  * queries (TSDFVolume.query, the pose optimiser's normal equations) read a
    voxel and its six axis neighbours, which a seam can split between ranks:
    every rank writes the taps it owns (`query_taps`), ONE float all-reduce
    merges them (ownership is disjoint, so the sum is exact), and the reference
    arithmetic runs on the merged taps -- bit-identical to the single volume;
  * re-integration after a backend solve (`reintegrate`): the keyframes are
    replicated in every rank's store at first ingest, so a pose update moves
    32 bytes per keyframe (one broadcast of the new poses);
  * counters / stats are sums over ranks (all_reduce).
`torch.distributed` is plumbing only; the group may be NCCL (GPU) or gloo
(host-logic tests with an injected volume factory).
"""
import torch
import torch.distributed as dist


def slab_of(kz, z0, slab_width):
    """Slab index of voxel layer kz (floor division, like the kernels)."""
    return (kz - z0) // slab_width


def slab_owner(kz, z0, slab_width, world):
    return slab_of(kz, z0, slab_width) % world


class PeerKeyframes:
    """Keyframe batches that are NOT broadcast (DESIGN.md 6 "peer
    pull").  Every rank keeps ITS share of each batch -- `seg_kf` keyframes as one
    packed segment (`tsdf.global_volume.packed_layout`) per slot -- in one device
    buffer of `n_slots` segments, and maps the buffers of all other ranks into
    its address space with CUDA IPC (one process per GPU; NVSwitch gives every
    peer full NVLink bandwidth).  `ShardedTSDFVolume.integrate_slot` then moves
    ~28 KB of meta per keyframe through ONE all-gather and lets the fuse kernel
    TMA-load just the tiles that reach this rank's slabs from the owners' HBM.
    With one rank (or `local_only`) nothing is mapped: the same code path runs on
    local pointers (tests)."""

    def __init__(self, n_slots, seg_kf, H, W, gh=0, gw=0, device=None, group=None,
                 local_only=False):
        from . import _ffi
        from .tsdf.global_volume import packed_layout

        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() and not local_only else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() and not local_only else 1
        self.n_slots, self.seg_kf, self.H, self.W, self.gh, self.gw = n_slots, seg_kf, H, W, gh, gw
        self.lay = packed_layout(seg_kf, H, W, gh, gw)
        self.seg_len = self.lay["seg_len"]
        import ctypes as C

        self.device = torch.device(device)
        self._dev_index = self.device.index if self.device.index is not None else \
            torch.cuda.current_device()
        n = n_slots * self.seg_len
        # the buffer is its own cudaMalloc allocation owned by libb200tsdf (offset 0 in its
        # IPC handle); torch sees it through __cuda_array_interface__ (zero copy)
        ptr, hnd = C.c_void_p(), (C.c_ubyte * 64)()
        _ffi.check(_ffi.lib().b200tsdf_peer_alloc(self._dev_index, n * 4, C.byref(ptr), hnd),
                   "peer_alloc")
        self._own = int(ptr.value)

        class _Raw:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "version": 2,
                                        "data": (self._own, False)}

        self.buf = torch.as_tensor(_Raw(), device=self.device)
        assert self.buf.data_ptr() == self._own
        self._opened = []
        self.base = [0] * self.world  # device address of rank r's buffer in THIS process
        self.base[self.rank] = self._own
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(hnd), group=group)
            for r, hb in enumerate(handles):
                if r == self.rank:
                    continue
                p2 = C.c_void_p()
                _ffi.check(_ffi.lib().b200tsdf_ipc_open(self._dev_index, hb, 0, C.byref(p2)),
                           "ipc_open")
                self.base[r] = int(p2.value)
                self._opened.append((int(p2.value), 0))
            dist.barrier(group=group)

    def slot(self, i):
        """This rank's packed segment of slot i (1-D float32 view)."""
        return self.buf[i * self.seg_len:(i + 1) * self.seg_len]

    def views(self, i):
        """(X (k,HW,3), C (k,HW), q (k,gh,gw) or None, pose (k,8)) views of my slot i."""
        sg, lay, k, n = self.slot(i), self.lay, self.seg_kf, self.H * self.W
        X = sg[lay["X"]:lay["X"] + k * n * 3].view(k, n, 3)
        Cf = sg[lay["C"]:lay["C"] + k * n].view(k, n)
        q = sg[lay["q"]:lay["q"] + k * self.gh * self.gw].view(k, self.gh, self.gw) \
            if self.gh and self.gw else None
        P = sg[lay["pose"]:lay["pose"] + k * 8].view(k, 8)
        return X, Cf, q, P

    def pointers(self, i):
        """Per source rank: device addresses of its slot-i pointmaps / confidences."""
        off = i * self.seg_len * 4
        xs = [b + off + 4 * self.lay["X"] for b in self.base]
        cs = [b + off + 4 * self.lay["C"] for b in self.base]
        return xs, cs

    def close(self):
        from . import _ffi

        for ptr, off in self._opened:
            _ffi.lib().b200tsdf_ipc_close(self._dev_index, ptr, off)
        self._opened = []
        if dist.is_initialized() and self.world > 1:
            dist.barrier(group=self.group)  # nobody reads my buffer any more
        if self._own:
            self.buf = None
            _ffi.lib().b200tsdf_peer_free(self._dev_index, self._own)
            self._own = 0

    def __del__(self):
        try:
            if getattr(self, "_own", 0) and not self._opened:
                from . import _ffi

                self.buf = None
                _ffi.lib().b200tsdf_peer_free(self._dev_index, self._own)
                self._own = 0
        except Exception:  # noqa: BLE001
            pass


class ShardedTSDFVolume:
    def __init__(self, voxel_size, truncation, max_weight=100.0, min_weight=1.0e-3, *,
                 slab_width, bounds=None, capacity_log2=17, device=None, group=None,
                 volume_factory=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.slab_width = int(slab_width)
        if self.slab_width < 8 or self.slab_width & (self.slab_width - 1):
            raise ValueError("slab_width must be a power of two >= 8")
        self.z0 = int(bounds[0][2]) if bounds is not None else 0
        self.device = device
        if volume_factory is None:
            from .tsdf.global_volume import TSDFVolume

            def volume_factory(**kw):
                return TSDFVolume(voxel_size, truncation, max_weight, min_weight, **kw)
        self.vol = volume_factory(capacity_log2=capacity_log2, bounds=bounds, device=device,
                                  slabs=(self.world, self.rank, self.slab_width))
        self._store = {}  # keyframe id -> tensors of the keyframe, replicated on every rank

    # -- keyframes ---------------------------------------------------------
    def broadcast_keyframes(self, X, Cf, poses, q_grid=None, src=0):
        """In-place broadcast of one batch from rank `src`; every rank passes
        tensors of the batch's shape (contents matter on `src` only)."""
        if self.world > 1:
            for t in (X, Cf, poses) + ((q_grid,) if q_grid is not None else ()):
                dist.broadcast(t, src, group=self.group)
        return X, Cf, poses, q_grid

    def gather_keyframes(self, X, Cf, poses, q_grid=None, out=None):
        """Sharded ingest: every rank contributes the SAME number of keyframes
        (its share of the batch, e.g. the cameras it serves); all ranks end up
        with the whole batch in rank order.  This is the world_size
        simultaneous broadcasts of `broadcast_keyframes` issued as one
        all-gather, so the host->device copies run on every rank's own PCIe
        link.  `out` = optional preallocated (X, Cf, poses, q_grid) of the full
        batch shape."""
        parts = (X, Cf, poses) + ((q_grid,) if q_grid is not None else ())
        if self.world == 1:
            return X, Cf, poses, q_grid
        res = []
        for i, t in enumerate(parts):
            t = t.contiguous()
            full = out[i] if out is not None else \
                torch.empty((t.shape[0] * self.world,) + tuple(t.shape[1:]), dtype=t.dtype,
                            device=t.device)
            if dist.get_backend(self.group) == "nccl":
                dist.all_gather_into_tensor(full, t, group=self.group)
            else:
                dist.all_gather(list(full.chunk(self.world)), t, group=self.group)
            res.append(full)
        if q_grid is None:
            res.append(None)
        return tuple(res)

    def integrate_keyframes(self, X, Cf, poses, q_grid=None, min_conf=0.05, H=None, W=None,
                            step_scale=0.5, src=0, broadcast=True, ids=None, max_points=0,
                            seed=0, kf_ids=None):
        """Fuse one batch on every rank.  With `ids` (one int per keyframe, the
        same list on every rank) the batch is also kept in this rank's
        keyframe store, so that `reintegrate` later needs the new poses only."""
        if broadcast:
            self.broadcast_keyframes(X, Cf, poses, q_grid, src)
        sub = dict(max_points=max_points, seed=seed, kf_ids=kf_ids) if max_points else {}
        self.vol.integrate_keyframes(X, Cf, poses, q_grid, min_conf, H, W, step_scale, **sub)
        if ids is not None:
            ids = [int(i) for i in ids]
            if len(ids) != int(X.shape[0]):
                raise ValueError("ids must name every keyframe of the batch")
            # one clone of the batch (not one per keyframe); the entries are views of it and
            # remember their run, so that re-integration of a stored run needs no re-stacking
            # (the batch's memory is released when its last keyframe is forgotten)
            run = {"X": X.clone(), "C": Cf.clone(),
                   "q": None if q_grid is None else q_grid.clone()}
            pc = poses.reshape(-1, 8).clone()
            for b, i in enumerate(ids):
                self._store[i] = {
                    "X": run["X"][b], "C": run["C"][b], "pose": pc[b].clone(),
                    "q": None if q_grid is None else run["q"][b], "run": (run, b),
                    "args": (float(min_conf), int(H), int(W), float(step_scale))}

    def integrate_slot(self, peers, i, min_conf=0.05, step_scale=0.5, meta_bufs=None):
        """Fuse slot i of a `PeerKeyframes` ring on every rank WITHOUT moving the
        pointmaps: meta record of my share (one kernel), ONE all-gather of the
        records (tile z ranges, poses, quality grids: ~28 KB per keyframe), then
        the fuse kernel pulls the tiles that reach my slabs from the owners' HBM.
        Everything is queued on the current stream; the all-gather also orders the
        owners' writes to slot i (queued before it) ahead of the peers' reads.
        meta_bufs: optional (share, full) int32 tensors to reuse."""
        X, Cf, q, P = peers.views(i)
        words = self.vol.keyframe_meta_words(peers.seg_kf, peers.H, peers.W, peers.gh, peers.gw)
        if meta_bufs is None:
            meta_bufs = (torch.empty(words, dtype=torch.int32, device=self.device),
                         torch.empty(words * self.world, dtype=torch.int32, device=self.device))
        share, full = meta_bufs
        self.vol.keyframe_meta(X, Cf, P, q, min_conf, peers.H, peers.W, out=share)
        if self.world > 1:
            dist.all_gather_into_tensor(full, share, group=self.group)
        else:
            full = share
        xs, cs = peers.pointers(i)
        self.vol.integrate_keyframes_pull(xs, cs, full, peers.seg_kf, peers.H, peers.W, peers.gh,
                                          peers.gw, min_conf, step_scale)
        return meta_bufs

    def integrate_slots(self, peers, slots, min_conf=0.05, step_scale=0.5, meta_bufs=None,
                        ready=None):
        """`integrate_slot` for a run of ring slots, software-pipelined: the meta
        pass and the all-gather of slot j+1 run on a side stream while the fuse of
        slot j runs on the current stream, so the collective's latency (tens of
        microseconds per launch, 1/3 of a step at 8 ranks) is hidden behind the
        fuse.  ready: optional events, one per slot, after which the slot's
        content is complete (host->device copies on another stream); without it
        the side stream starts behind everything queued on the current stream.
        meta_bufs: optional two (share, full) pairs to reuse."""
        slots = list(slots)
        if not slots:
            return meta_bufs
        main = torch.cuda.current_stream(self.device)
        side = getattr(self, "_side", None)
        if side is None:
            side = self._side = torch.cuda.Stream(self.device)
            self._gathered = [torch.cuda.Event(), torch.cuda.Event()]
            self._consumed = [torch.cuda.Event(), torch.cuda.Event()]
        words = self.vol.keyframe_meta_words(peers.seg_kf, peers.H, peers.W, peers.gh, peers.gw)
        if meta_bufs is None:
            meta_bufs = [(torch.empty(words, dtype=torch.int32, device=self.device),
                          torch.empty(words * self.world, dtype=torch.int32, device=self.device))
                         for _ in range(2)]
        side.wait_stream(main)

        def stage(j):  # meta record of my share of slot j + the all-gather, on the side stream
            share, full = meta_bufs[j & 1]
            X, Cf, q, P = peers.views(slots[j])
            with torch.cuda.stream(side):
                if j >= 2:
                    side.wait_event(self._consumed[j & 1])  # fuse j-2 has read this record
                if ready is not None:
                    side.wait_event(ready[j])
                self.vol.keyframe_meta(X, Cf, P, q, min_conf, peers.H, peers.W, out=share)
                if self.world > 1:
                    dist.all_gather_into_tensor(full, share, group=self.group)
                self._gathered[j & 1].record(side)

        stage(0)
        for j, i in enumerate(slots):
            if j + 1 < len(slots):
                stage(j + 1)
            share, full = meta_bufs[j & 1]
            main.wait_event(self._gathered[j & 1])
            xs, cs = peers.pointers(i)
            self.vol.integrate_keyframes_pull(xs, cs, full if self.world > 1 else share,
                                              peers.seg_kf, peers.H, peers.W, peers.gh, peers.gw,
                                              min_conf, step_scale)
            self._consumed[j & 1].record(main)
        side.wait_stream(main)  # the buffers may be reused by the next call
        return meta_bufs

    # -- re-integration after a backend solve ------------
    def stored_ids(self):
        return sorted(self._store)

    def forget(self, ids=None):
        """Drop keyframes from the store (all of them with ids=None): their
        contribution stays in the volume, they can no longer be re-fused.
        Returns the bytes the dropped entries held on this rank (entries are views
        of one clone per ingested batch: the allocator gets the memory back when
        the last keyframe of a batch is forgotten)."""
        ids = list(self._store) if ids is None else [int(i) for i in ids]
        freed = 0
        for i in ids:
            e = self._store.pop(i, None)
            if e is not None:
                freed += sum(t.numel() * t.element_size() for t in (e["X"], e["C"], e["pose"])
                             + ((e["q"],) if e["q"] is not None else ()))
        return freed

    def reintegrate(self, ids, poses, src=0, mode="additive", batch=40):
        """`TSDFGlobalManager.on_after_backend_solve` -> `_process_dirty_queue`
        (global_manager.py:214-229, :71-79) over ranks: the keyframes `ids`
        (same list on every rank) are fused again at their new poses.  Only
        `poses` (len(ids) x 8 float32, 32 bytes per keyframe) travels -- it is
        broadcast from `src`; pointmaps, confidences and quality grids come
        from the store every rank filled at first ingest.
        mode "additive": the reference's semantics (the keyframe is fused on
        top of its earlier contribution); "replace": the contribution at the
        stored pose is de-integrated first (exact inverse), so no ghost stays.
        Returns the number of keyframes fused."""
        if mode not in ("additive", "replace"):
            raise ValueError("mode must be 'additive' or 'replace'")
        ids = [int(i) for i in ids]
        missing = [i for i in ids if i not in self._store]
        if missing:
            raise KeyError(f"keyframes {missing} are not in the store (pass ids= at ingest)")
        if not ids:
            return 0
        first = self._store[ids[0]]
        poses = poses.reshape(len(ids), 8).to(first["pose"].device, torch.float32).contiguous()
        if self.world > 1:
            dist.broadcast(poses, src, group=self.group)
        # group by ingest arguments / shapes, keep the caller's order inside a group
        groups = {}
        for n, i in enumerate(ids):
            e = self._store[i]
            key = (e["args"], tuple(e["X"].shape), None if e["q"] is None else tuple(e["q"].shape))
            groups.setdefault(key, []).append((n, i))
        for (args, _, qshape), members in groups.items():
            min_conf, H, W, step_scale = args
            for a in range(0, len(members), batch):
                part = members[a:a + batch]
                es = [self._store[i] for _, i in part]
                run, b0 = es[0].get("run", (None, 0))
                if run is not None and all(e.get("run", (None, 0))[0] is run and
                                           e["run"][1] == b0 + j for j, e in enumerate(es)):
                    # consecutive keyframes of one stored batch: zero-copy slices
                    X, Cf = run["X"][b0:b0 + len(es)], run["C"][b0:b0 + len(es)]
                    Q = None if qshape is None else run["q"][b0:b0 + len(es)]
                else:
                    X = torch.stack([e["X"] for e in es])
                    Cf = torch.stack([e["C"] for e in es])
                    Q = None if qshape is None else torch.stack([e["q"] for e in es])
                if mode == "replace":
                    old = torch.stack([e["pose"] for e in es])
                    self.vol.deintegrate_keyframes(X, Cf, old, Q, min_conf, H, W, step_scale)
                new = poses[[n for n, _ in part]].contiguous()
                self.vol.integrate_keyframes(X, Cf, new, Q, min_conf, H, W, step_scale)
                for (n, _), e in zip(part, es):
                    e["pose"] = poses[n].clone()
        return len(ids)

    def integrate_keyframes_from_host(self, X, Cf, poses, q_grid=None, min_conf=0.05, H=None,
                                      W=None, step_scale=0.5, chunks=None):
        """Sharded ingest from HOST memory, pipelined: every rank passes ITS
        share of the batch (the same number of keyframes on every rank, pinned
        host tensors).  The share is cut into `chunks` pieces; piece c+1 crosses
        PCIe (copy stream) while piece c is all-gathered over NVLink (NCCL
        stream) and piece c-1 is fused (current stream).  One fuse launch per
        piece over the R x piece keyframes of all ranks; the fold is
        commutative, so the piece order does not matter.  chunks=None: 2 pieces
        for shares of >= 16 keyframes, else 1 (small shares are launch-latency-
        bound: at 5 keyframes per rank 2 pieces cost 1.59 vs 1.42 ms per step,
        at 20 they gain 2.08 -> 2.02 ms; profiles/r01j_e2e_chunks.txt)."""
        parts = [X, Cf, poses] + ([q_grid] if q_grid is not None else [])
        n = int(X.shape[0])
        nccl = self.world > 1 and dist.get_backend(self.group) == "nccl"
        if not nccl:  # single rank / host-logic tests: no streams to overlap
            dev_parts = [t.to(self.device) for t in parts]
            full = self.gather_keyframes(*dev_parts[:3],
                                         dev_parts[3] if q_grid is not None else None)
            self.vol.integrate_keyframes(full[0], full[1], full[2], full[3], min_conf, H, W,
                                         step_scale)
            return
        if chunks is None:
            chunks = 2 if n >= 16 else 1
        chunks = max(1, min(int(chunks), n))
        cuts = [(n * c) // chunks for c in range(chunks + 1)]
        key = (tuple(tuple(t.shape) for t in parts), chunks)
        if getattr(self, "_ingest_key", None) != key:
            self._ingest_key = key
            self._copy_stream = torch.cuda.Stream(self.device)
            self._ingest_bufs = []
            for c in range(chunks):
                m = cuts[c + 1] - cuts[c]
                loc = [torch.empty((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=self.device)
                       for t in parts]
                full = [torch.empty((m * self.world,) + tuple(t.shape[1:]), dtype=t.dtype,
                                    device=self.device) for t in parts]
                self._ingest_bufs.append((loc, full, torch.cuda.Event()))
        main = torch.cuda.current_stream(self.device)
        self._copy_stream.wait_stream(main)  # the previous batch's fuse still reads the buffers
        with torch.cuda.stream(self._copy_stream):
            for c in range(chunks):
                loc, _, ev = self._ingest_bufs[c]
                for dst, src in zip(loc, parts):
                    dst.copy_(src[cuts[c]:cuts[c + 1]], non_blocking=True)
                ev.record(self._copy_stream)
        works = []
        for c in range(chunks):
            loc, full, ev = self._ingest_bufs[c]
            main.wait_event(ev)
            works.append([dist.all_gather_into_tensor(f, l, group=self.group, async_op=True)
                          for f, l in zip(full, loc)])
        for c in range(chunks):
            for w in works[c]:
                w.wait()  # stream-level wait: the fuse below is ordered after the gather
            full = self._ingest_bufs[c][1]
            self.vol.integrate_keyframes(full[0], full[1], full[2],
                                         full[3] if q_grid is not None else None, min_conf, H, W,
                                         step_scale)

    def on_after_backend_solve(self, indices, poses, pin=1, src=0, mode="additive",
                               queue_cap=256):
        """The reference's hook with its filters (global_manager.py:214-229,
        :46-54, :71-79), over ranks: of the solve's keyframe `indices` (same
        list on every rank; `poses` row n belongs to indices[n] and matters on
        `src` only) keep those >= `pin` (`local_opt.pin`), once each, at most
        `queue_cap` (the bounded re-integration queue drops the rest), skip the
        ones never ingested (`idx < len(keyframes)` there), and re-fuse them.
        Returns the list of re-fused indices."""
        keep, rows, seen = [], [], set()
        for n, i in enumerate(int(i) for i in indices):
            if i < pin or i in seen:
                continue
            if len(seen) >= queue_cap:
                break
            seen.add(i)
            if i in self._store:
                keep.append(i)
                rows.append(n)
        if keep:
            poses = poses.reshape(-1, 8)
            self.reintegrate(keep, poses[rows].contiguous(), src=src, mode=mode)
        return keep

    # -- checkpoint / resume ---------------------------------------------------
    def save_checkpoint(self, directory):
        """Collective: every rank writes its slabs as
        `<directory>/slabs_rank<r>.npz` (`TSDFVolume.save_checkpoint` format:
        block keys, packed voxels, hit counters, volume parameters); rank 0 adds
        `meta.json`.  Returns the number of blocks this rank wrote."""
        import json
        import os

        os.makedirs(directory, exist_ok=True)
        n = self.vol.save_checkpoint(os.path.join(directory, f"slabs_rank{self.rank:03d}.npz"))
        if self.rank == 0:
            with open(os.path.join(directory, "meta.json"), "w") as f:
                json.dump({"world": self.world, "slab_width": self.slab_width, "z0": self.z0,
                           "format": "b200tsdf-slabs-1"}, f)
        if self.world > 1:
            dist.barrier(group=self.group)
        return n

    def load_checkpoint(self, directory):
        """Restore a `save_checkpoint` directory -- also onto a DIFFERENT number
        of ranks or slab width: every rank reads all the slab files and imports
        the blocks whose slab it owns now.  A voxel block is 8 layers and slabs
        are multiples of 8 layers measured from an origin that is a multiple of
        8, so a block never straddles a seam; the stored sums are additive, so
        loading into a filled volume merges.  Returns the blocks imported."""
        import json
        import os

        import numpy as np

        with open(os.path.join(directory, "meta.json")) as f:
            meta = json.load(f)
        if meta.get("format") != "b200tsdf-slabs-1":
            raise ValueError(f"{directory}: not a sharded TSDF checkpoint")
        if (self.z0 - int(meta["z0"])) % 8:
            raise ValueError("slab origin of the checkpoint is not block-aligned with this volume")
        mine = None
        if all(hasattr(self.vol, a) for a in ("voxel_size", "truncation", "max_weight",
                                              "min_weight")):
            mine = np.array([self.vol.voxel_size, self.vol.truncation, self.vol.max_weight,
                             self.vol.min_weight], np.float64)
        total = 0
        for r in range(int(meta["world"])):
            with np.load(os.path.join(directory, f"slabs_rank{r:03d}.npz")) as z:
                if mine is not None and not np.allclose(z["params"], mine, rtol=1e-12, atol=0.0):
                    raise ValueError(f"checkpoint parameters {z['params']} != volume {mine}")
                bk = z["block_keys"].reshape(-1, 3)
                own = slab_owner(bk[:, 2] * 8, self.z0, self.slab_width, self.world) == self.rank
                if own.any():
                    self.vol.import_blocks(torch.from_numpy(bk[own]),
                                           torch.from_numpy(z["vox"][own]),
                                           torch.from_numpy(z["cnt"][own]))
                    total += int(own.sum())
        if self.world > 1:
            dist.barrier(group=self.group)
        return total

    # -- reductions --------------------------------------------------------
    def _sum(self, values):
        t = torch.tensor(values, dtype=torch.int64,
                         device=self.device if dist.is_initialized() and
                         dist.get_backend(self.group) == "nccl" else "cpu")
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
        return [int(x) for x in t.cpu()]

    def counters(self):
        c = self.vol.counters()
        keys = ["updates_primary", "updates_quality", "blocks", "dropped", "culled_rays"]
        out = dict(zip(keys, self._sum([c[k] for k in keys])))
        out["rays"] = c["rays"]  # every rank stages the full batch
        out["local"] = c
        return out

    def stats(self, which=0):
        s = self.vol.stats(which)
        v = self._sum([s["valid_voxels"], s["total_voxels"]])
        return {"valid_voxels": v[0], "total_voxels": v[1]}

    # -- queries across seams --------------------------------------------------
    def _merged_taps(self, points, which, pose=None):
        taps = self.vol.query_taps(points, which, pose)
        if self.world > 1:
            if dist.get_backend(self.group) == "nccl":
                dist.all_reduce(taps, group=self.group)
            else:  # host-logic tests: gloo reduces host tensors
                host = taps.cpu()
                dist.all_reduce(host, group=self.group)
                taps = host.to(taps.device)
        return taps

    def query_batch(self, points, which=0):
        """`TSDFVolume.query_batch` over all ranks: every rank passes the same
        points and gets the same (tsdf, grad, flags)."""
        return self.vol.query_finish(self._merged_taps(points, which))

    def query(self, point_world):
        """Reference signature (global_volume.py:93-105); collective."""
        import numpy as np

        t, g, fl = self.query_batch(np.asarray(point_world).reshape(1, 3))
        f = int(fl.cpu()[0])
        if not f & 1:
            return None, None
        return float(t.cpu()[0]), (g.cpu().numpy()[0].copy() if f & 2 else None)

    def pose_normal_equations(self, pts_cam, conf, pose, lam, which=0):
        """`TSDFVolume.pose_normal_equations` over all ranks (collective; every
        rank passes the same samples and gets the same H, b, used), so a
        `TSDFPoseOptimizer` can be handed a sharded volume."""
        taps = self._merged_taps(pts_cam, which, pose)
        return self.vol.pose_normal_equations_taps(taps, pts_cam, conf, pose, lam)

    # -- seams / mesh --------------------------------------------------------
    def exchange_seams(self):
        """Send my seam blocks to the rank below, install the ones of the rank
        above as halo blocks.  Returns the number of halo blocks installed."""
        if self.world == 1:
            return 0
        bk, vox, cnt = self.vol.export_seam_blocks()
        down, up = (self.rank - 1) % self.world, (self.rank + 1) % self.world
        n_mine = torch.tensor([bk.shape[0]], dtype=torch.int64, device=bk.device)
        counts = [torch.zeros_like(n_mine) for _ in range(self.world)]
        dist.all_gather(counts, n_mine, group=self.group)
        n_up = int(counts[up].item())
        rbk = torch.zeros(n_up, 3, dtype=torch.int64, device=bk.device)
        rvox = torch.zeros(n_up, 512, 4, dtype=torch.float32, device=bk.device)
        rcnt = torch.zeros(n_up, 512, dtype=torch.int64, device=bk.device)
        ops = []
        for s_t, r_t in ((bk, rbk), (vox, rvox), (cnt, rcnt)):
            if s_t.shape[0]:
                ops.append(dist.P2POp(dist.isend, s_t, down, group=self.group))
            if r_t.shape[0]:
                ops.append(dist.P2POp(dist.irecv, r_t, up, group=self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        self.vol.import_halo_blocks(rbk, rvox, rcnt)
        return n_up

    def extract_mesh(self, which=0, gather=False):
        """Marching cubes of this rank's slabs incl. their top seams.  With
        gather=True rank 0 receives the whole mesh (others get their part)."""
        self.exchange_seams()
        verts = self.vol.extract_mesh(which)
        if not gather or self.world == 1:
            return verts
        n = torch.tensor([verts.shape[0]], dtype=torch.int64, device=verts.device)
        counts = [torch.zeros_like(n) for _ in range(self.world)]
        dist.all_gather(counts, n, group=self.group)
        counts = [int(c.item()) for c in counts]
        pad = max(counts + [1])
        buf = torch.zeros(pad, 3, dtype=torch.float32, device=verts.device)
        buf[: verts.shape[0]] = verts
        parts = [torch.zeros_like(buf) for _ in range(self.world)]
        dist.all_gather(parts, buf, group=self.group)
        if self.rank != 0:
            return verts
        return torch.cat([p[:c] for p, c in zip(parts, counts)], dim=0)
