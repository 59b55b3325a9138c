"""Experiment: where does the fp32 build's flux error vs the fp64 oracle come from?"""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut
from helpers import oracle_lw, oracle_sw, as_f32_valued
F32IN = os.environ.get('F32IN') == '1'

for name, atm in (("rce-cloudy", rce_atmosphere(256, 64, cloudy=True)), ("rce128", rce_atmosphere(128, 128, cloudy=True)),
                  ("allsky", allsky_atmosphere(24, 72))):
    for kind in ("sw", "lw"):
        kd, lut = synthetic_kdist(kind), synthetic_cloud_lut(kind)
        for cloudy in (False, True):
            ref = (oracle_sw if kind == "sw" else oracle_lw)(kd, as_f32_valued(atm) if F32IN else atm, lut if cloudy else None)
            S = api.Radiation_solver_shortwave if kind == "sw" else api.Radiation_solver_longwave
            for dtype in (torch.float32,):
                sol = S(kd, atm.vmr.keys(), lut, dtype)
                out = sol.solve_gpu(api.DeviceColumns(atm, dtype), switch_cloud_optics=cloudy)
                errs = {k: np.abs(out[k].numpy() - ref[k]) for k in ("flux_up", "flux_dn")}
                e = max(v.max() for v in errs.values())
                col = int(np.argmax(errs["flux_dn"].max(axis=1)))
                cl = bool(atm.lwp is not None and (atm.lwp[col].sum() + atm.iwp[col].sum()) > 0)
                print(f"{name:10s} {kind} cloudy={cloudy!s:5s} {str(dtype)[6:]:8s} max|err|={e:.3e} worst col {col} (has cloud {cl}, mu0 {atm.mu0[col]:.2f})")
