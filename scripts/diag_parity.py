"""Stage-by-stage parity diagnostics against the golden fixtures and the oracle (prints max relative differences)."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import maria_b200 as M
from common import load_case, relmax, small_case
from oracle import atm_oracle as O

def run(case, dtype):
    np.random.seed(case["seed"])
    sim = M.Simulation(case["instrument"], case["plan"], case["site"], atmosphere="2d", dtype=dtype,
                       atmosphere_kwargs=dict(weather=case["weather"], spectrum=case["spectrum"], geometry="exact"))
    return sim, sim.run(units="K_RJ")[0]

for name in ["m2_daisy", "1deg_ces"]:
    case = load_case(name); fx = case["fx"]
    np.random.seed(case["seed"]); rec = {}
    res = O.simulate(case["instrument"], case["plan"], case["site"], case["weather"], case["spectrum"], np.random, dtype=np.float64, record=rec)
    for dtype in [np.float64, np.float32]:
        sim, tod = run(case, dtype)
        obs = sim.obs_list[0]; atm = obs.atmosphere
        print(f"== {name} {np.dtype(dtype).name} n_t_atm={len(atm.boresight.t)}")
        print("  values   ", max(relmax(p.values, fx[f"ref_values_{k}"]) for k, p in atm.processes.items()))
        print("  smoothed ", max(relmax(atm.smoothed_layer(li), sm) for li, sm in rec["smoothed"].items()))
        print("  zpwv     ", relmax(atm.zenith_scaled_pwv, fx["ref_zenith_scaled_pwv"]))
        load = sim._compute_atmospheric_loading(obs).cpu().numpy()
        ol = res["loading_pW"] / rec["gain_error"][:, None]
        err = np.abs(load - ol).max(axis=0) / np.abs(ol).max()
        print("  loading  ", err.max(), "argmax t", err.argmax(), "of", len(err), "first/last", err[:3], err[-3:])
        # at knots only
        tk = atm.boresight.t
        print("  loading atm-rate (oracle) vs device at nearest samples: skip")
        sel = fx["det_sel"]
        k = tod.data["atmosphere"]
        e2 = np.abs(k[sel] - fx["ref_tod_K_RJ_sel"]).max(axis=0) / np.abs(fx["ref_tod_K_RJ_sel"]).max()
        print("  K_RJ     ", e2.max(), "argmax", e2.argmax(), e2[:3], e2[-3:])
        ref = fx["ref_tod_K_RJ_sel"]; fl = ref - ref.mean(axis=1, keepdims=True)
        print("  K_RJ fluct err", np.abs((k[sel] - ref)).max() / np.abs(fl).max())
