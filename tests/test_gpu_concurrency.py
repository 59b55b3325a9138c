"""Calls on ONE handle from different CUDA streams and host threads (include/rrtmgp_b200.h, "Concurrency"): every call re-carves
the handle's workspace, so the library serialises them — a per-handle mutex on the host, an event chain across streams.  Without
that the second call overwrites the tile state and scratch strips the first call's kernels are still reading."""
import threading

import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import api
from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut

pytestmark = pytest.mark.gpu
KEYS = ("flux_up", "flux_dn", "flux_net")


def _setup(path):
    kd, lut = synthetic_kdist("lw"), synthetic_cloud_lut("lw")
    atm_a = rce_atmosphere(3000, 40, cloudy=True)
    atm_b = rce_atmosphere(1777, 52, cloudy=True, top_at_1=True)
    sol = api.Radiation_solver_longwave(kd, atm_a.vmr.keys(), lut, torch.float32)
    sol.set_fused_path(path)
    cols = [api.DeviceColumns(atm_a, torch.float32), api.DeviceColumns(atm_b, torch.float32)]
    ref = []
    for c in cols:
        out = sol.solve_gpu(c, switch_cloud_optics=True, out={})
        torch.cuda.synchronize()
        ref.append({k: out[k].numpy().copy() for k in KEYS})
    return sol, cols, ref


@pytest.mark.parametrize("path", [0, 1])
def test_two_streams_on_one_handle(path):
    sol, cols, ref = _setup(path)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [{}, {}]
    torch.cuda.synchronize()
    for rep in range(6):                 # back to back, no synchronisation between the calls
        for i in (0, 1):
            with torch.cuda.stream(streams[i]):
                sol.solve_gpu(cols[i], switch_cloud_optics=True, out=outs[i])
    torch.cuda.synchronize()
    for i in (0, 1):
        for k in KEYS:
            assert np.array_equal(outs[i][k].numpy(), ref[i][k]), (i, k)


def test_two_host_threads_on_one_handle():
    sol, cols, ref = _setup(0)
    outs = [{}, {}]
    errors = []

    def work(i):
        try:
            torch.cuda.set_device(0)
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for rep in range(6):
                    sol.solve_gpu(cols[i], switch_cloud_optics=True, out=outs[i])
            st.synchronize()
        except Exception as ex:          # surfaced below: an exception in a thread would otherwise pass silently
            errors.append(repr(ex))

    threads = [threading.Thread(target=work, args=(i,)) for i in (0, 1)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    torch.cuda.synchronize()
    assert not errors, errors
    for i in (0, 1):
        for k in KEYS:
            assert np.array_equal(outs[i][k].numpy(), ref[i][k]), (i, k)
