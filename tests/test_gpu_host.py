"""Host-buffer entry (rrtmgp_b200_fluxes_host): pinned host arrays in, pinned host arrays out,
chunked H2D -> fused LW/SW -> D2H pipeline.  Columns are independent, so the result must be
bit-identical to the device-resident call whatever the chunking, and within tolerance of the oracle."""
import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut
from helpers import oracle_lw, oracle_sw

pytestmark = pytest.mark.gpu
LWK = ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net", "heating_rate")
SWK = LWK + ("flux_dn_dir", "bnd_flux_dn_dir")


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("chunk", [0, 13, 64])
def test_host_pipeline_equals_device_call(dtype, chunk):
    atm = rce_atmosphere(70, 40, night_fraction=0.2)
    atm.tsi_scaling[:] = np.linspace(0.9, 1.1, atm.ncol)
    lw = api.Radiation_solver_longwave(synthetic_kdist("lw"), atm.vmr.keys(), dtype=dtype)
    sw = api.Radiation_solver_shortwave(synthetic_kdist("sw"), atm.vmr.keys(), dtype=dtype)
    dev = api.DeviceColumns(atm, dtype)
    ref_lw = lw.solve_gpu(dev, switch_output_bnd_fluxes=True, heating=True)
    ref_sw = sw.solve_gpu(dev, switch_output_bnd_fluxes=True, heating=True)
    torch.cuda.synchronize()
    o_lw, o_sw = api.solve_host(lw, sw, api.HostColumns(atm, dtype), switch_output_bnd_fluxes=True, heating=True, chunk_columns=chunk)
    for k in LWK:
        assert not o_lw[k].t.is_cuda and o_lw[k].t.is_pinned()
        assert np.array_equal(o_lw[k].numpy(), ref_lw[k].numpy()), k
    for k in SWK:
        assert np.array_equal(o_sw[k].numpy(), ref_sw[k].numpy()), k


@pytest.mark.parametrize("dtype,tol", [(torch.float64, 1e-5), (torch.float32, 0.05)])
def test_host_pipeline_allsky_vs_oracle(dtype, tol):
    atm = allsky_atmosphere(24, 72)
    kl, ks, ll, ls = synthetic_kdist("lw"), synthetic_kdist("sw"), synthetic_cloud_lut("lw"), synthetic_cloud_lut("sw")
    lw = api.Radiation_solver_longwave(kl, atm.vmr.keys(), ll, dtype)
    sw = api.Radiation_solver_shortwave(ks, atm.vmr.keys(), ls, dtype)
    o_lw, o_sw = api.solve_host(lw, sw, api.HostColumns(atm, dtype), switch_cloud_optics=True, chunk_columns=10)
    r_lw, r_sw = oracle_lw(kl, atm, ll), oracle_sw(ks, atm, ls)
    for k in ("flux_up", "flux_dn", "flux_net"):
        assert np.max(np.abs(o_lw[k].numpy() - r_lw[k])) <= tol, k
    for k in ("flux_up", "flux_dn", "flux_net", "flux_dn_dir"):
        assert np.max(np.abs(o_sw[k].numpy() - r_sw[k])) <= tol, k


def test_host_pipeline_single_solver_and_repeat():
    atm = rce_atmosphere(33, 24, top_at_1=True)
    sw = api.Radiation_solver_shortwave(synthetic_kdist("sw"), atm.vmr.keys(), dtype=torch.float64)
    ref = oracle_sw(synthetic_kdist("sw"), atm)
    h = api.HostColumns(atm, torch.float64)
    out = {}
    for _ in range(3):      # buffers of the pipeline are reused across calls
        _, o = api.solve_host(None, sw, h, out_sw=out, chunk_columns=8)
        assert np.max(np.abs(o["flux_dn"].numpy() - ref["flux_dn"])) <= 1e-5
    with pytest.raises(RuntimeError):
        api.solve_host(None, None, h)
