"""Synthetic keyframes for tests and bench: analytic plane /
sphere / room scenes rendered to H x W pointmaps in the CAMERA frame
(row-major pixel order idx = y*W + x, like frame.py:242 of the reference),
summed confidences C = 1 + 0.5*U(0,1), a gh x gw quality grid in 1/16 steps
and Sim3 poses [t, q_xyzw, s] on a smooth orbit.  Seeded numpy only.
"""
import numpy as np


def quat_from_R(R):
    """xyzw quaternion of a rotation matrix (float64)."""
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = [(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s]
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k]) * 2
        q = [0, 0, 0, 0]
        q[i] = 0.25 * s
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
        q[3] = (R[k, j] - R[j, k]) / s
    q = np.array(q, np.float64)
    return q / np.linalg.norm(q)


def R_from_quat(q):
    x, y, z, w = [float(v) for v in q]
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]], np.float64)


def look_at(eye, target, up=(0.0, 1.0, 0.0)):
    """Camera-to-world rotation with +z forward, +x right, +y down."""
    f = np.asarray(target, np.float64) - np.asarray(eye, np.float64)
    f /= np.linalg.norm(f)
    r = np.cross(f, np.asarray(up, np.float64))
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    return np.stack([r, d, f], axis=1)


def pixel_dirs(H, W, f=400.0):
    cx, cy = W / 2.0, H / 2.0
    v, u = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64),
                       indexing="ij")
    d = np.stack([(u + 0.5 - cx) / f, (v + 0.5 - cy) / f, np.ones_like(u)], axis=-1)
    return d.reshape(-1, 3)


def _hit_plane(o, d, n, dist):
    den = d @ n
    lam = (dist - o @ n) / np.where(np.abs(den) < 1e-12, 1e-12, den)
    return np.where((np.abs(den) > 1e-9) & (lam > 0.05), lam, np.inf)


def _hit_sphere(o, d, c, r):
    oc = o - c
    a = np.sum(d * d, axis=1)
    b = 2.0 * (d @ oc)
    cc = oc @ oc - r * r
    disc = b * b - 4 * a * cc
    sq = np.sqrt(np.maximum(disc, 0.0))
    l0, l1 = (-b - sq) / (2 * a), (-b + sq) / (2 * a)
    lam = np.where(l0 > 0.05, l0, l1)
    return np.where((disc > 0) & (lam > 0.05), lam, np.inf)


def _hit_room(o, d, half):
    """Inside of the axis-aligned box [-half, half]^3: nearest exit wall."""
    with np.errstate(divide="ignore", invalid="ignore"):
        lp = (half - o) / d
        ln = (-half - o) / d
    lam = np.where(d > 0, lp, ln)
    lam = np.where(np.abs(d) < 1e-12, np.inf, lam)
    return lam.min(axis=1)


def render(scene, R, t, s, H, W, f=400.0):
    """Pointmap (H*W,3) float32 in the camera frame + hit mask, such that
    s*R*X + t lies on the scene surface."""
    dc = pixel_dirs(H, W, f)
    dw = dc @ R.T
    kind = scene["kind"]
    if kind == "plane":
        lam = _hit_plane(t, dw, np.asarray(scene["n"], np.float64), float(scene["d"]))
    elif kind == "sphere":
        lam = _hit_sphere(t, dw, np.asarray(scene["c"], np.float64), float(scene["r"]))
    elif kind == "room":
        lam = _hit_room(t, dw, float(scene["half"]))
    else:
        raise ValueError(kind)
    hit = np.isfinite(lam)
    lam = np.where(hit, lam, 1.0)
    X = dc * (lam / s)[:, None]
    return X.astype(np.float32), hit


def make_keyframes(scene, n_kf, H=384, W=512, seed=0, f=400.0, orbit_radius=0.6, gh=None,
                   gw=None, start=0, total=None, up_axis="y"):
    """n_kf keyframes on a closed orbit of `total` poses (default n_kf),
    starting at pose index `start`.  Returns dict of float32 arrays:
    X (B,H*W,3), C (B,H*W), q (B,gh,gw), pose (B,8)."""
    total = total or n_kf
    gh = gh or max(1, H // 16)
    gw = gw or max(1, W // 16)
    X = np.zeros((n_kf, H * W, 3), np.float32)
    Cf = np.zeros((n_kf, H * W), np.float32)
    Q = np.zeros((n_kf, gh, gw), np.float32)
    P = np.zeros((n_kf, 8), np.float32)
    for b in range(n_kf):
        i = start + b
        rng = np.random.default_rng([seed, i])
        ang = 2 * np.pi * i / total
        if scene["kind"] == "room":
            eye = np.array([orbit_radius * np.cos(ang), 0.1 * np.sin(3 * ang),
                            orbit_radius * np.sin(ang)])
            tgt = eye + np.array([np.cos(ang), 0.25 * np.sin(2 * ang), np.sin(ang)])
            if up_axis == "z":  # orbit in the xy plane, z vertical (z-slabs cut every view)
                eye, tgt = eye[[0, 2, 1]], tgt[[0, 2, 1]]
        else:
            eye = np.array([0.15 * np.cos(ang), 0.1 * np.sin(ang), 0.05 * np.sin(2 * ang)])
            tgt = np.array([0.05 * np.sin(ang), 0.05 * np.cos(ang), 1.5])
        R = look_at(eye, tgt, (0.0, 0.0, 1.0) if up_axis == "z" else (0.0, 1.0, 0.0))
        s = 1.0 + 0.05 * np.sin(ang + 0.3)
        q = quat_from_R(R)
        Rq = R_from_quat(q.astype(np.float32).astype(np.float64))
        pose = np.concatenate([eye, q, [s]]).astype(np.float32)
        Xb, hit = render(scene, Rq, pose[:3].astype(np.float64), float(pose[7]), H, W, f)
        X[b] = Xb
        Cf[b] = np.where(hit, 1.0 + 0.5 * rng.random(H * W), 0.0).astype(np.float32)
        Q[b] = (np.round(rng.random((gh, gw)) * 16) / 16).astype(np.float32)
        P[b] = pose
    return {"X": X, "C": Cf, "q": Q, "pose": P, "H": H, "W": W, "gh": gh, "gw": gw}


SCENES = {
    "plane": {"kind": "plane", "n": [0.1, -0.2, 1.0], "d": 1.2},
    "sphere": {"kind": "sphere", "c": [0.0, 0.0, 1.5], "r": 0.5},
    "room": {"kind": "room", "half": 1.2437},
    # walls exactly on voxel faces for 5 mm / 1 cm voxels: stresses floor() exactness
    "room_aligned": {"kind": "room", "half": 1.25},
}


def bounds_for(scene, voxel_size, dim):
    """Half-open integer voxel-key bounds of a dim^3 grid centred on the scene."""
    if scene["kind"] == "room":
        c = np.zeros(3)
    elif scene["kind"] == "sphere":
        c = np.asarray(scene["c"], np.float64)
    else:
        c = np.array([0.0, 0.0, float(scene["d"])])
    k0 = np.floor(c / voxel_size).astype(np.int64) - dim // 2
    return (tuple(int(v) for v in k0), tuple(int(v) + dim for v in k0))
