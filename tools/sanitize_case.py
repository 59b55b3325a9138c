"""Small fused LW+SW case (clouds + aerosols, both kernel families, host pipeline) for compute-sanitizer runs:
  compute-sanitizer --tool {memcheck,racecheck,initcheck,synccheck} python tools/sanitize_case.py"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, add_aerosols
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut

atm = add_aerosols(rce_atmosphere(int(os.environ.get("NCOL", 45)), int(os.environ.get("NLAY", 24)), cloudy=True))
for dtype in (torch.float32, torch.float64):
    lw = api.Radiation_solver_longwave(synthetic_kdist("lw"), atm.vmr.keys(), synthetic_cloud_lut("lw"), dtype)
    sw = api.Radiation_solver_shortwave(synthetic_kdist("sw"), atm.vmr.keys(), synthetic_cloud_lut("sw"), dtype,
                                        aerosol_lut=synthetic_aerosol_lut())
    cols = api.DeviceColumns(atm, dtype)
    ref = {}
    for path in (0, 1):
        lw.set_fused_path(path); sw.set_fused_path(path)
        a = lw.solve_gpu(cols, switch_cloud_optics=True, switch_output_bnd_fluxes=True, heating=True)["flux_up"].numpy().copy()
        b = sw.solve_gpu(cols, switch_cloud_optics=True, switch_aerosol_optics=True, switch_output_bnd_fluxes=True)["flux_dn"].numpy().copy()
        torch.cuda.synchronize()
        if path == 0:
            ref = dict(a=a, b=b)
        else:
            assert np.max(np.abs(a - ref["a"])) < 1e-2 and np.max(np.abs(b - ref["b"])) < 1e-2
    lw.set_fused_path(0); sw.set_fused_path(0)
    ol, os_ = api.solve_host(lw, sw, api.HostColumns(atm, dtype), switch_cloud_optics=True, switch_aerosol_optics=True, chunk_columns=16)
    assert np.array_equal(ol["flux_up"].numpy(), ref["a"]) and np.array_equal(os_["flux_dn"].numpy(), ref["b"])
print("sanitize case ok")
