"""Longwave fused call with 1..4 Gauss angles on a large domain: ms per call, event-timed.  RB200_ANGLES=launches in the
environment selects the one-pass-per-angle path for comparison (set before the library loads).
Usage: python tools/time_angles.py [ncol] [nlay]"""
import os
import sys
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from rte_rrtmgp_cpp_b200 import api  # noqa: E402
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere  # noqa: E402
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist  # noqa: E402

ncol = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
nlay = int(sys.argv[2]) if len(sys.argv) > 2 else 128
for dtype in (torch.float32, torch.float64):
    kd = synthetic_kdist("lw")
    atm = rce_atmosphere(min(ncol, bench.NBASE), nlay)
    lw = api.Radiation_solver_longwave(kd, atm.vmr.keys(), None, dtype)
    lw.set_fused_path(0)
    cols = api.DeviceColumns(atm, dtype)
    if ncol > bench.NBASE:
        cols = bench.tile_columns(cols, ncol, shift=0)
    out = {}
    res = {}
    for nang in (1, 2, 3, 4):
        lw.n_gauss_angles = nang
        for _ in range(2):
            lw.solve_gpu(cols, out=out)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            lw.solve_gpu(cols, out=out)
        e1.record()
        torch.cuda.synchronize()
        res[nang] = e0.elapsed_time(e1) / 3
    print(str(dtype)[6:], os.environ.get("RB200_ANGLES", "in-kernel"), f"ncol {ncol} nlay {nlay}",
          {k: f"{v:.1f} ms ({v / res[1]:.2f}x)" for k, v in res.items()})
