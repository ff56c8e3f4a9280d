"""fp32 error budget on the C2 configuration (largest detector offsets): pW vs K_RJ, worst detector / sample."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import maria_b200 as M
from maria_b200 import configs
from oracle import atm_oracle as O

cfg = configs.c2_act_ces(n_per_array=24, duration=120.0)
np.random.seed(17)
rec = {}
res = O.simulate(cfg["instrument"], cfg["plan"], cfg["site"], cfg["weather"], cfg["spectrum"], np.random, dtype=np.float64, record=rec)
ref_k, ref_p = res["tod_K_RJ"], res["loading_pW"]
np.random.seed(17)
sim = M.Simulation(cfg["instrument"], cfg["plan"], cfg["site"], atmosphere="2d", dtype=np.float32,
                   atmosphere_kwargs=dict(weather=cfg["weather"], spectrum=cfg["spectrum"], geometry="exact"))
tod = sim.run(units="K_RJ")[0]
k = tod.data["atmosphere"].astype(np.float64)
p = tod.to("pW").data["atmosphere"].astype(np.float64)
obs = sim.obs_list[0]; atm = obs.atmosphere
pw_direct = sim._compute_atmospheric_loading(obs, units="pW", apply_gain=True).cpu().numpy().astype(np.float64)
off = cfg["instrument"].dets.offsets
r = np.degrees(np.hypot(off[:, 0], off[:, 1]))
for name, dev, ref in [("K_RJ", k, ref_k), ("pW (TOD.to)", p, ref_p), ("pW direct", pw_direct, ref_p)]:
    e = np.abs(dev - ref) / np.abs(ref).max()
    d, t = np.unravel_index(e.argmax(), e.shape)
    print(f"{name:12s} relmax {e.max():.3e} at det {d} (r = {r[d]:.2f} deg, band {cfg['instrument'].dets.band_name[d]}) sample {t}; "
          f"median-over-det of max-over-t {np.median(e.max(axis=1)):.2e}; mean rel bias {np.mean((dev - ref) / ref):.2e}")
zp = atm.zenith_scaled_pwv
print("zenith pwv (atm rate) relmax", np.abs(zp - res["atmosphere"].zenith_scaled_pwv).max() / np.abs(zp).max())
ratio = (k / ref_k) / (pw_direct / ref_p)
print("calibration factor dev/ref - 1: max", np.abs(ratio - 1).max(), "mean", (ratio - 1).mean())
