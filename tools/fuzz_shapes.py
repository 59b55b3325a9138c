"""Randomised shape sweep on a GPU box: random (columns, layers, orientation, particulates, Gauss angles, precision, fused path,
tile size, paired or separate calls) against the CPU oracle at the GPU tests' tolerances.  Not part of the suite (its cases are
random by design); `python tools/fuzz_shapes.py [cases] [seed]` prints one line per case and exits non-zero on the first miss."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import oracle_lw, oracle_sw  # noqa: E402
from rte_rrtmgp_cpp_b200 import api  # noqa: E402
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, add_aerosols  # noqa: E402
from rte_rrtmgp_cpp_b200.kdist import (synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut,  # noqa: E402
                                       REDUCED_LW_128, REDUCED_SW_112)

TOL = {torch.float64: 1e-5, torch.float32: 0.05}
ncases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 20241105)
worst = {torch.float64: 0.0, torch.float32: 0.0}
t0 = time.time()
for case in range(ncases):
    ncol = int(rng.choice([1, 2, 3, 31, 32, 33, 64, 97, 128, 200, 257]))
    nlay = int(rng.choice([1, 2, 3, 7, 16, 33, 60, 64, 72, 127, 128, 129, 140]))
    top = bool(rng.integers(2)); cloudy = bool(rng.integers(2)); aero = cloudy and bool(rng.integers(2))
    reduced = bool(rng.integers(4) == 0); dtype = (torch.float64, torch.float32)[int(rng.integers(2))]
    n_ang = int(rng.choice([1, 1, 2, 3, 4])); path = int(rng.choice([0, 0, 1, 2])); pair = bool(rng.integers(2))
    tile = int(rng.choice([0, 0, 32, 96])); bnd = bool(rng.integers(2)); seed = int(rng.integers(1 << 30))
    if path == 1 and n_ang > 1:
        path = 0                                   # column-resident longwave: one angle
    atm = rce_atmosphere(ncol, nlay, seed=seed, top_at_1=top, cloudy=cloudy, night_fraction=0.2)
    if aero:
        atm = add_aerosols(atm, seed=seed)
    kd_lw = synthetic_kdist("lw", REDUCED_LW_128) if reduced else synthetic_kdist("lw")
    kd_sw = synthetic_kdist("sw", REDUCED_SW_112) if reduced else synthetic_kdist("sw")
    lut_lw = synthetic_cloud_lut("lw") if cloudy else None
    lut_sw = synthetic_cloud_lut("sw") if cloudy else None
    alut = synthetic_aerosol_lut() if aero else None
    ref_lw = oracle_lw(kd_lw, atm, lut_lw, n_ang=n_ang)
    ref_sw = oracle_sw(kd_sw, atm, lut_sw, aerosol_lut=alut)
    lw = api.Radiation_solver_longwave(kd_lw, atm.vmr.keys(), lut_lw, dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, atm.vmr.keys(), lut_sw, dtype, aerosol_lut=alut)
    for s_ in (lw, sw):
        s_.set_fused_path(path)
        if tile:
            s_.set_tile_columns(tile)
        s_.set_debug(1)                            # NaN-poisoned workspace + canaries: an unwritten or overrun piece shows
    lw.n_gauss_angles = n_ang
    cols = api.DeviceColumns(atm, dtype)
    kw = dict(switch_cloud_optics=cloudy, switch_output_bnd_fluxes=bnd)
    host = bool(rng.integers(4) == 0)
    if host:                                       # the host-buffer pipeline with a chunk size that does not divide the columns
        chunk = int(rng.choice([0, 32, 50, 96]))
        o_lw, o_sw = api.solve_host(lw, sw, api.HostColumns(atm, dtype), chunk_columns=chunk, switch_aerosol_optics=aero, **kw)
    elif pair:
        o_lw, o_sw = api.solve_pair(lw, sw, cols, switch_aerosol_optics=aero, **kw)
    else:
        o_lw, o_sw = lw.solve_gpu(cols, **kw), sw.solve_gpu(cols, switch_aerosol_optics=aero, **kw)
    torch.cuda.synchronize()
    d = 0.0
    keys = ["flux_up", "flux_dn", "flux_net"] + (["bnd_flux_up", "bnd_flux_dn"] if bnd else [])
    for out, ref, extra in ((o_lw, ref_lw, []), (o_sw, ref_sw, ["flux_dn_dir"])):
        for k in keys + extra:
            v = out[k].numpy()
            assert np.isfinite(v).all(), (case, k, "not finite")
            d = max(d, float(np.max(np.abs(v - ref[k]))))
    worst[dtype] = max(worst[dtype], d)
    desc = (f"case {case:3d} ncol {ncol:3d} nlay {nlay:3d} top {int(top)} cloud {int(cloudy)} aero {int(aero)} reduced {int(reduced)} "
            f"{str(dtype)[6:]:7s} angles {n_ang} path {path} pair {int(pair)} host {int(host)} tile {tile:2d} bnd {int(bnd)}  max|d| {d:.2e}")
    print(desc, flush=True)
    if d > TOL[dtype]:
        print("MISS", desc)
        sys.exit(1)
print(f"{ncases} cases ok in {time.time() - t0:.0f} s; worst fp64 {worst[torch.float64]:.2e}, fp32 {worst[torch.float32]:.2e} W/m2")
