"""B200 mirror of the reference `mast3r_slam/quality_core.py`.

Every public function of the reference keeps its name and arguments
(quality_core.py:4-140).  The per-keyframe path -- `compute_batch`, `classify`,
`u_from_CQ`, `r_from_scalar`, `valid_grid` -- runs on CUDA kernels
(`csrc/quality.cu` through the C ABI) and its grids STAY on the device: the
reference moves each keyframe's maps tensor -> numpy -> queue -> tensor ->
numpy (tracker.py:129-141, quality_async.py:12-46) before the TSDF path reads
them.  Two deliberate differences, both switchable: `compute_batch` /
`pack_result` return device tensors unless `to_host=True` (then the
reference's numpy dict), and `quality_grid(result)` is new (the per-patch
weight q = 1 - priority that `TSDFVolume.integrate_keyframes` takes as
`q_grid`).  The small tensor helpers `_hw`, `_grid`, `reduce_grid`,
`upsample_to_hw`, `robust_z` are plain torch as in the reference (they are
off the keyframe path and work on any device).
"""
import ctypes as C

import torch

from . import _ffi


def _dev_index(device):
    d = torch.device(device)
    if d.type != "cuda":
        raise _ffi.B200TSDFError("quality_core: CUDA device required (there is no CPU fallback)")
    return d.index if d.index is not None else torch.cuda.current_device()


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _f32(x, device, n):
    return torch.as_tensor(x).to(device, torch.float32).reshape(-1)[:n].contiguous()


def patch_grids(Ck, Qk, r_pix, valid, h, w, ps, C_thr, Q_thr, device):
    """u (u_from_CQ :43-49), r (r_from_scalar :51-52 with `valid`), valid grid
    (:54-56) of one keyframe in ONE launch: (gh, gw) float32 device tensors."""
    n = h * w
    Ct, Qt, rt = _f32(Ck, device, n), _f32(Qk, device, n), _f32(r_pix, device, n)
    # any non-zero value is valid, like the reference's .to(torch.bool) (:17)
    vt = (torch.as_tensor(valid).to(device).reshape(-1)[:n] != 0).to(torch.uint8).contiguous()
    gh, gw = h // ps, w // ps
    u, r, v = (torch.empty(gh, gw, dtype=torch.float32, device=device) for _ in range(3))
    _ffi.check(_ffi.lib().b200tsdf_quality_patch_grids(
        Ct.data_ptr(), Qt.data_ptr(), rt.data_ptr(), vt.data_ptr(), h, w, ps, float(C_thr),
        float(Q_thr), u.data_ptr(), r.data_ptr(), v.data_ptr(), _dev_index(device),
        _stream(device)), "quality_patch_grids")
    return u, r, v


# ---- tensor helpers of the reference (:4-32, :58-61), off the keyframe path ---
def _hw(x, h, w):
    """(h*w,) or (h*w, 1) -> (h, w); anything else unchanged (:4-7)."""
    if x is not None and (x.ndim == 1 or (x.ndim == 2 and x.shape[-1] == 1)):
        x = x.view(h, w)
    return x


def _grid(x, ps):
    """(H, W) -> (gh, gw, ps*ps) patch pixels, with gh, gw (:9-13)."""
    gh, gw = x.shape[-2] // ps, x.shape[-1] // ps
    return x[:gh * ps, :gw * ps].unfold(0, ps, ps).unfold(1, ps, ps).reshape(gh, gw, ps * ps), gh, gw


def reduce_grid(x, h, w, ps, valid=None, method="median"):
    """Per-patch lower median / mean, over the valid pixels when a mask is given
    (patches without a valid pixel give 0) (:15-27)."""
    X, _, _ = _grid(_hw(x, h, w), ps)
    if valid is not None:
        M, _, _ = _grid(_hw(valid, h, w).to(torch.bool).float(), ps)
        X = X.masked_fill(M < 0.5, float("nan"))
        v = torch.nanmedian(X, dim=-1).values if method == "median" else torch.nanmean(X, dim=-1)
        return torch.nan_to_num(v, nan=0.0)
    return torch.nanmedian(X, dim=-1).values if method == "median" else X.mean(-1)


def upsample_to_hw(g, h, w, mode="bilinear"):  # :29-32
    kw = {} if mode == "nearest" else {"align_corners": False}
    return torch.nn.functional.interpolate(g[None, None], (h, w), mode=mode, **kw)[0, 0]


def robust_z(x, eps=1e-6):  # :58-61
    m = torch.median(x)
    return (x - m) / (torch.median(torch.abs(x - m)) + eps)


def u_from_CQ(C, Q, C_thr, Q_thr, h, w, ps, quant=0.5):
    """Reference signature (:43-49); on a CUDA tensor the fused patch kernel."""
    C = torch.as_tensor(C)
    if C.is_cuda:
        ones = torch.ones(h * w, dtype=torch.uint8, device=C.device)
        return patch_grids(C, Q, torch.zeros(h * w, device=C.device), ones, h, w, ps, C_thr, Q_thr,
                           C.device)[0]
    Cn = torch.clamp(_hw(C, h, w) / (C_thr + 1e-8), 0, 1)
    Qn = torch.clamp(_hw(torch.as_tensor(Q), h, w) / (Q_thr + 1e-8), 0, 1)
    return reduce_grid(1 - torch.sqrt(torch.clamp(Cn * Qn, 0, 1)), h, w, ps, method="median")


def r_from_scalar(r, h, w, ps, valid=None):  # :51-52
    return reduce_grid(r, h, w, ps, valid=valid, method="median")


def valid_grid(valid, h, w, ps):  # :54-56
    return (reduce_grid(torch.as_tensor(valid).float(), h, w, ps, method="mean") > 0).float()


def view_weight(t_norm, theta, b0, theta0, device):  # :34-37
    t = torch.clamp(t_norm / b0, 0, 1) if b0 > 0 else torch.ones((), device=device)
    r = torch.clamp(theta / theta0, 0, 1) if theta0 > 0 else torch.ones((), device=device)
    return 0.5 * (t + r)


def ema_delta(prev, inc, alpha):  # :39-41
    new = alpha * prev + (1 - alpha) * inc
    return new, new - prev


def classify(delta_cov, r, u, thr_zr=1.0, thr_zu=1.0, thr_dc=0.02, return_q=False):
    """:63-109 on the device: (class_id int64, priority float32[, q = 1 - priority])."""
    shape = delta_cov.shape
    dev = delta_cov.device
    dc = delta_cov.to(torch.float32).reshape(-1).contiguous()
    rr = r.to(dev, torch.float32).reshape(-1).contiguous()
    uu = u.to(dev, torch.float32).reshape(-1).contiguous()
    cls = torch.empty(dc.numel(), dtype=torch.int64, device=dev)
    pri = torch.empty(dc.numel(), dtype=torch.float32, device=dev)
    q = torch.empty(dc.numel(), dtype=torch.float32, device=dev)
    _ffi.check(_ffi.lib().b200tsdf_quality_classify(
        dc.data_ptr(), rr.data_ptr(), uu.data_ptr(), dc.numel(), float(thr_zr), float(thr_zu),
        float(thr_dc), cls.data_ptr(), pri.data_ptr(), q.data_ptr(), _dev_index(dev),
        _stream(dev)), "quality_classify")
    out = (cls.reshape(shape), pri.reshape(shape))
    return out + (q.reshape(shape),) if return_q else out


def pack_result(kfid, dc, r, u, cls, pri, ewma, q=None, to_host=False):  # :111-125
    conv = (lambda x: x.cpu().numpy()) if to_host else (lambda x: x)
    out = {"kf_id": int(kfid), "delta_cov": conv(dc), "r": conv(r), "u": conv(u),
           "class_id": conv(cls), "priority": conv(pri), "cov_ewma": conv(ewma)}
    if q is not None:
        out["q_grid"] = conv(q)
    return out


def _compute_batch_per_job(batch, ps, alpha, b0, theta0, C_thr, Q_thr, thr_zr, thr_zu, thr_dc,
                           device, to_host=False):
    """compute_batch one keyframe at a time (two launches + ~15 tensor ops per
    keyframe): the statement the batched path is checked against."""
    outs = []
    for jb in batch:
        h, w = jb["H"], jb["W"]
        u, r, vg = patch_grids(jb["Ck"], jb["Qk"], jb["r_pix"], jb["valid_kf"], h, w, ps, C_thr,
                               Q_thr, device)
        inc = vg * view_weight(torch.as_tensor(jb["t_norm"]).to(device),
                               torch.as_tensor(jb["theta"]).to(device), b0, theta0, device)
        prev = jb.get("cov_ewma", None)
        prev = torch.zeros_like(inc) if prev is None else torch.as_tensor(prev).to(device)
        ew, dc = ema_delta(prev, inc, alpha)
        cls, pri, q = classify(dc, r, u, thr_zr, thr_zu, thr_dc, return_q=True)
        outs.append(pack_result(jb["kf_id"], dc, r, u, cls, pri, ew, q, to_host))
    return outs


def _stack(batch, key, device, dtype, n):
    """(B, n) device tensor from the jobs' `key`: one stack (on the device the
    tensors live on), then ONE device / dtype conversion for the whole batch."""
    ts = [torch.as_tensor(jb[key]).reshape(-1)[:n] for jb in batch]
    if len({t.device for t in ts}) > 1:
        ts = [t.to(device) for t in ts]
    return torch.stack(ts).to(device=device, dtype=dtype).contiguous()


def compute_stacked(Ck, Qk, r_pix, valid, t_norm, theta, prev_ewma, h, w, ps, alpha, b0, theta0,
                    C_thr, Q_thr, thr_zr, thr_zu, thr_dc, device):
    """compute_batch (:127-140) on already stacked device tensors -- maps
    (B, h*w) float32 (valid uint8), t_norm / theta (B,), prev_ewma (B, gh*gw) or
    None -- in TWO launches (`b200tsdf_quality_batch`).  Returns the dict of
    (B, gh, gw) device grids: delta_cov, r, u, class_id, priority, cov_ewma, q_grid."""
    B, gh, gw = Ck.shape[0], h // ps, w // ps
    vw = view_weight(t_norm, theta, b0, theta0, device).to(torch.float32).expand(B).contiguous()
    f32 = dict(dtype=torch.float32, device=device)
    u, r, vg, ew, dc, pri, q = (torch.empty(B, gh, gw, **f32) for _ in range(7))
    cls = torch.empty(B, gh, gw, dtype=torch.int64, device=device)
    _ffi.check(_ffi.lib().b200tsdf_quality_batch(
        Ck.data_ptr(), Qk.data_ptr(), r_pix.data_ptr(), valid.data_ptr(), vw.data_ptr(),
        prev_ewma.data_ptr() if prev_ewma is not None else None, B, h, w, ps, float(C_thr),
        float(Q_thr), float(alpha), float(thr_zr), float(thr_zu), float(thr_dc), u.data_ptr(),
        r.data_ptr(), vg.data_ptr(), ew.data_ptr(), dc.data_ptr(), cls.data_ptr(), pri.data_ptr(),
        q.data_ptr(), _dev_index(device), _stream(device)), "quality_batch")
    return {"delta_cov": dc, "r": r, "u": u, "class_id": cls, "priority": pri, "cov_ewma": ew,
            "q_grid": q}


def compute_batch(batch, ps, alpha, b0, theta0, C_thr, Q_thr, thr_zr, thr_zu, thr_dc, device,
                  to_host=False):  # :127-140
    """All keyframes of one size go through TWO launches (`compute_stacked`:
    per-patch medians of every keyframe, then one CTA per keyframe for the
    coverage EMA and classify); the reference loops over the jobs.  Results are
    per job, in order, views into the (B, gh, gw) device grids."""
    outs = [None] * len(batch)
    groups = {}
    for i, jb in enumerate(batch):
        groups.setdefault((int(jb["H"]), int(jb["W"])), []).append(i)
    for (h, w), idx in groups.items():
        jobs = [batch[i] for i in idx]
        B, n, npx = len(jobs), (h // ps) * (w // ps), h * w
        prev = None
        if any(jb.get("cov_ewma", None) is not None for jb in jobs):
            zero = torch.zeros(n, dtype=torch.float32)
            prev = _stack([{"p": zero if jb.get("cov_ewma", None) is None else jb["cov_ewma"]}
                           for jb in jobs], "p", device, torch.float32, n)
        g = compute_stacked(_stack(jobs, "Ck", device, torch.float32, npx),
                            _stack(jobs, "Qk", device, torch.float32, npx),
                            _stack(jobs, "r_pix", device, torch.float32, npx),
                            (_stack(jobs, "valid_kf", device, torch.float32, npx) != 0)
                            .to(torch.uint8),
                            _stack(jobs, "t_norm", device, torch.float32, 1).reshape(B),
                            _stack(jobs, "theta", device, torch.float32, 1).reshape(B), prev, h, w,
                            ps, alpha, b0, theta0, C_thr, Q_thr, thr_zr, thr_zu, thr_dc, device)
        for k, i in enumerate(idx):
            outs[i] = pack_result(batch[i]["kf_id"], g["delta_cov"][k], g["r"][k], g["u"][k],
                                  g["class_id"][k], g["priority"][k], g["cov_ewma"][k],
                                  g["q_grid"][k], to_host)
    return outs


def quality_grid(result):
    """(1, gh, gw) device tensor q = 1 - priority for `integrate_keyframes`."""
    q = result.get("q_grid")
    if q is None:
        q = 1.0 - torch.as_tensor(result["priority"], dtype=torch.float32).clamp(0.0, 1.0)
    return q.reshape(1, *q.shape[-2:]).contiguous()
