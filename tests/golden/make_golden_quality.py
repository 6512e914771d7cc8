"""Freeze outputs of the UNMODIFIED reference quality producer
(`mast3r_slam/quality_core.py:127-140 compute_batch`, pure torch, imported from
/root/reference in the build container) on seeded inputs.

    python tests/golden/make_golden_quality.py   ->  tests/golden/quality_cases.npz
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/mast3r_slam/quality_core.py"
PARAMS = dict(ps=16, alpha=0.8, b0=0.15, theta0=float(np.deg2rad(10.0)), C_thr=1.5, Q_thr=1.5,
              thr_zr=1.0, thr_zu=1.0, thr_dc=0.02)  # config/base.yaml:60-72


def make_job(seed, H, W, with_prev, holes):
    g = torch.Generator().manual_seed(seed)
    n = H * W
    ps = PARAMS["ps"]
    valid = torch.rand(n, generator=g) > 0.35
    r_pix = torch.rand(n, generator=g) ** 2 * 0.05
    Ck = torch.rand(n, 1, generator=g) * 3.0
    Qk = torch.rand(n, 1, generator=g) * 2.5
    if holes:  # patches without a valid pixel, NaN inputs, a saturated region
        v2 = valid.view(H, W)
        v2[:ps, :2 * ps] = False
        v2[2 * ps:3 * ps, ps:2 * ps] = False
        r_pix.view(H, W)[ps:2 * ps, :ps] = float("nan")
        Ck.view(H, W)[ps:2 * ps, ps:2 * ps] = float("nan")
        Ck.view(H, W)[:, -ps:] = 10.0
        Qk.view(H, W)[:, -ps:] = 10.0
    job = {"H": H, "W": W, "kf_id": seed, "valid_kf": valid, "r_pix": r_pix, "Ck": Ck, "Qk": Qk,
           "t_norm": torch.tensor(0.02 + 0.03 * seed), "theta": torch.tensor(0.05 * (seed + 1))}
    if with_prev:
        job["cov_ewma"] = torch.rand(H // ps, W // ps, generator=g)
    return job


CASES = [("small", 0, 48, 64, False, False), ("holes", 1, 64, 80, True, True),
         ("ragged", 2, 57, 70, True, True), ("full", 3, 384, 512, True, False)]


def main():
    spec = importlib.util.spec_from_file_location("ref_quality_core", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    out = {}
    for name, seed, H, W, with_prev, holes in CASES:
        job = make_job(seed, H, W, with_prev, holes)
        res = ref.compute_batch([dict(job)], device="cpu", **PARAMS)[0]
        for k in ("delta_cov", "r", "u", "class_id", "priority", "cov_ewma"):
            out[f"{name}/{k}"] = np.asarray(res[k])
    np.savez_compressed(os.path.join(HERE, "quality_cases.npz"), **out)
    print({k: v.shape for k, v in out.items() if k.endswith("priority")})


if __name__ == "__main__":
    main()
