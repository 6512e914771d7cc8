"""Freeze golden vectors by running the UNMODIFIED reference TSDFVolume
(/root/reference/mast3r_slam/tsdf/global_volume.py) on seeded synthetic scenes.

Run in the build container only (the GPU box has no /root/reference):
    python tests/golden/make_golden.py
Each fixture holds the reference's outputs: sorted voxel keys, tsdf, weight,
per-voxel _update_voxel hit counts (CountingVolume wraps, never alters, the
reference arithmetic), stats() and query() on a probe set.  Inputs are
regenerated from the seeded generator (synth.py) by the tests; a checksum of
the staged inputs is stored to catch generator drift.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from mast3r_slam_quality_dualtsdf_b200 import synth  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from oracle import ref_harness  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (scene, n_kf, H, W, f, voxel, trunc, max_weight, dtype, seed)
    "plane_v30_f32": ("plane", 3, 24, 32, 25.0, 0.03, 0.12, 100.0, "f32", 1),
    "sphere_v30_f32": ("sphere", 3, 24, 32, 25.0, 0.03, 0.12, 100.0, "f32", 2),
    "sphere_v30_f64": ("sphere", 2, 24, 32, 25.0, 0.03, 0.12, 100.0, "f64", 2),
    "room_v10_f32": ("room", 2, 24, 32, 25.0, 0.01, 0.04, 100.0, "f32", 3),
    "plane_clamp_f32": ("plane", 3, 16, 16, 20.0, 0.03, 0.12, 3.0, "f32", 4),
}


def staged_inputs(case):
    scene, n_kf, H, W, f, voxel, trunc, mw, dtype, seed = CASES[case]
    kf = synth.make_keyframes(synth.SCENES[scene], n_kf, H=H, W=W, seed=seed, f=f, total=12)
    out = []
    for b in range(n_kf):
        pts, conf, confq, org = orc.stage_keyframe(kf["X"][b], kf["C"][b], kf["pose"][b],
                                                   kf["q"][b], H, W, 0.05)
        if dtype == "f64":
            pts, org = pts.astype(np.float64), org.astype(np.float64)
        out.append((pts, conf, confq, org))
    return kf, out


def checksum(staged):
    h = hashlib.sha256()
    for pts, conf, confq, org in staged:
        for a in (pts, conf, confq, org):
            h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def probes(keys, voxel, seed):
    rng = np.random.default_rng(seed)
    p = (keys[rng.integers(0, len(keys), 300)] + rng.random((300, 3))) * voxel
    return np.concatenate([p, rng.normal(size=(30, 3)) * 3.0])


def main():
    assert ref_harness.available(), "reference tree not found"
    for case, spec in CASES.items():
        scene, n_kf, H, W, f, voxel, trunc, mw, dtype, seed = spec
        kf, staged = staged_inputs(case)
        out = {"checksum": np.frombuffer(checksum(staged).encode(), np.uint8)}
        for name, col in (("p", 1), ("q", 2)):
            vol = ref_harness.CountingVolume(voxel, trunc, mw)
            fused = 0
            for s in staged:
                fused += vol.integrate(s[0], s[col], s[3])
            k, t, w, c = vol.export()
            st = vol.vol.stats()
            pr = probes(k, voxel, seed)
            if dtype == "f32":
                pr = pr.astype(np.float32)
            qt = np.full(len(pr), np.nan)
            qg = np.full((len(pr), 3), np.nan)
            fl = np.zeros(len(pr), np.uint8)
            for i, p in enumerate(pr):
                tv, gv = vol.vol.query(p)
                if tv is not None:
                    qt[i], fl[i] = float(tv), 1
                if gv is not None:
                    qg[i], fl[i] = gv, 3
            out.update({f"{name}_keys": k.astype(np.int32), f"{name}_tsdf": t, f"{name}_w": w,
                        f"{name}_count": c, f"{name}_fused": np.int64(fused),
                        f"{name}_events": np.int64(vol.events),
                        f"{name}_stats": np.array([st["valid_voxels"], st["total_voxels"]]),
                        f"{name}_probe": pr, f"{name}_qt": qt, f"{name}_qg": qg, f"{name}_qfl": fl})
        np.savez_compressed(os.path.join(HERE, case + ".npz"), **out)
        print(case, "voxels", len(out["p_keys"]), len(out["q_keys"]), "events",
              int(out["p_events"]), int(out["q_events"]))


if __name__ == "__main__":
    main()
