import sys, numpy as np, torch
sys.path.insert(0, ".")
from mast3r_slam_quality_dualtsdf_b200 import synth
from mast3r_slam_quality_dualtsdf_b200.tsdf import TSDFVolume
from tests.util import oracle_dual
dev = torch.device("cuda:0")
kf = synth.make_keyframes(synth.SCENES["sphere"], 3, H=48, W=64, seed=2, f=50.0, total=12)
vp, vq, _ = oracle_dual(kf, 0.03, 0.12)
vol = TSDFVolume(0.03, 0.12, device=dev)
vol.integrate_keyframes(*(torch.from_numpy(kf[k]).to(dev) for k in ("X", "C", "pose")),
                        torch.from_numpy(kf["q"]).to(dev), 0.05, 48, 64)
keys = vp.export()[0]
rng = np.random.default_rng(0)
probe = (keys[rng.integers(0, len(keys), 4000)] + rng.random((4000, 3))) * 0.03
for dtype in (np.float32, np.float64):
    p = probe.astype(dtype)
    for which, vo in ((0, vp), (1, vq)):
        t, g, fl = vo.query_batch(p)
        tt, gg, ff = (x.cpu().numpy() for x in vol.query_batch(p, which))
        mis = np.nonzero(fl != ff)[0]
        print(dtype.__name__, which, "flag mismatches", len(mis), fl[mis][:10], ff[mis][:10])
        both = (fl == 3) & (ff == 3)
        err = np.abs(g - gg).max(axis=1)[both]
        print("  t err", np.abs(t - tt).max(), "grad err max", err.max(), "q99", np.quantile(err, 0.99))
        for i in mis[:5]:
            print("  ", p[i], t[i], tt[i], g[i], gg[i])
