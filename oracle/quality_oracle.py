"""CPU restatement (numpy, float32) of the reference quality producer
`mast3r_slam/quality_core.py` -- TEST INFRASTRUCTURE ONLY (tests/, smoke()).

Pinned by tests/golden/quality_cases.npz, frozen from the unmodified reference
(tests/golden/make_golden_quality.py).  Function names follow the reference;
each cites the lines it restates.
"""
import numpy as np

f32 = np.float32


def _grid(x, ps):  # quality_core.py:9-13
    h, w = x.shape
    gh, gw = h // ps, w // ps
    return x[:gh * ps, :gw * ps].reshape(gh, ps, gw, ps).transpose(0, 2, 1, 3).reshape(gh, gw, ps * ps)


def _nanmedian_last(X):
    """torch.nanmedian(dim=-1): lower median of the non-NaN entries, NaN if none."""
    out = np.full(X.shape[:-1], np.nan, f32)
    for idx in np.ndindex(*X.shape[:-1]):
        v = X[idx]
        v = np.sort(v[~np.isnan(v)])
        if v.size:
            out[idx] = v[(v.size - 1) // 2]
    return out


def reduce_grid(x, ps, valid=None):  # :15-27, method="median"
    X = _grid(np.asarray(x, f32), ps)
    if valid is None:
        return _nanmedian_last(X)
    M = _grid(np.asarray(valid, f32), ps)
    X = np.where(M < 0.5, f32(np.nan), X)
    return np.nan_to_num(_nanmedian_last(X), nan=0.0).astype(f32)


def view_weight(t_norm, theta, b0, theta0):  # :34-37
    t = np.clip(f32(t_norm) / f32(b0), 0, 1).astype(f32) if b0 > 0 else f32(1)
    r = np.clip(f32(theta) / f32(theta0), 0, 1).astype(f32) if theta0 > 0 else f32(1)
    return f32(0.5) * (t + r)


def u_from_CQ(C, Q, C_thr, Q_thr, ps):  # :43-49
    Cn = np.clip(np.asarray(C, f32) / f32(C_thr + 1e-8), 0, 1).astype(f32)
    Qn = np.clip(np.asarray(Q, f32) / f32(Q_thr + 1e-8), 0, 1).astype(f32)
    U = f32(1) - np.sqrt(np.clip(Cn * Qn, 0, 1).astype(f32))
    return reduce_grid(U.astype(f32), ps)


def valid_grid(valid, ps):  # :54-56
    return (_grid(np.asarray(valid, f32), ps).mean(-1) > 0).astype(f32)


def _median_all(x):
    """torch.median(x): lower median; NaN if any element is NaN."""
    if np.isnan(x).any():
        return f32(np.nan)
    return np.sort(x)[(x.size - 1) // 2]


def robust_z(x, eps=1e-6):  # :58-61
    m = _median_all(x)
    mad = _median_all(np.abs(x - m).astype(f32)) + f32(eps)
    return ((x - m) / mad).astype(f32)


def classify(delta_cov, r, u, thr_zr=1.0, thr_zu=1.0, thr_dc=0.02):  # :63-109
    shape = delta_cov.shape
    dc, r, u = (np.asarray(a, f32).reshape(-1) for a in (delta_cov, r, u))
    zr, zu = robust_z(r), robust_z(u)
    with np.errstate(invalid="ignore"):
        c1 = (dc < f32(thr_dc)) & (zu > f32(thr_zu))
        c2 = (dc >= f32(thr_dc)) & (zr > f32(thr_zr)) & (zu > f32(thr_zu))
        c3 = (zr > f32(thr_zr)) & (zu <= f32(thr_zu))
    cls = np.zeros(r.shape, np.int64)
    cls[c1] = 1
    cls[c2] = 2
    cls[c3] = 3
    p = np.zeros(r.shape, f32)
    p1 = (f32(1) - np.clip(dc, 0, 1)) + np.maximum(zu, 0)
    p2 = np.maximum(zr, 0) + np.maximum(zu, 0)
    p3 = np.maximum(zr, 0) + np.maximum(f32(1) - u, 0)
    p[cls == 1] = p1[cls == 1]
    p[cls == 2] = p2[cls == 2]
    p[cls == 3] = p3[cls == 3]
    p = (p / (p.max() + f32(1e-6))).astype(f32)
    return cls.reshape(shape), p.reshape(shape)


def compute_job(jb, ps, alpha, b0, theta0, C_thr, Q_thr, thr_zr, thr_zu, thr_dc):  # :127-140
    h, w = jb["H"], jb["W"]
    arr = lambda k: np.asarray(jb[k], f32).reshape(h, w)  # noqa: E731
    valid = np.asarray(jb["valid_kf"]).reshape(h, w).astype(bool)
    inc = valid_grid(valid, ps) * view_weight(float(jb["t_norm"]), float(jb["theta"]), b0, theta0)
    prev = jb.get("cov_ewma")
    prev = np.zeros_like(inc) if prev is None else np.asarray(prev, f32)
    ew = (f32(alpha) * prev + f32(1 - alpha) * inc).astype(f32)  # :39-41
    dc = (ew - prev).astype(f32)
    r = reduce_grid(arr("r_pix"), ps, valid=valid)
    u = u_from_CQ(arr("Ck"), arr("Qk"), C_thr, Q_thr, ps)
    cls, pri = classify(dc, r, u, thr_zr, thr_zu, thr_dc)
    return {"delta_cov": dc, "r": r, "u": u, "class_id": cls, "priority": pri, "cov_ewma": ew}
