"""rrtmgp_b200_sw_solver_2stream_noalloc (include/rrtmgp_b200.h): the allocation-free variant of the shortwave stage
solver, scratch in a static device array of the library.  No counterpart in the reference.
Called through the C ABI and compared with the existing entry point rrtmgp_b200_sw_solver_2stream on
the same device inputs, twice in a row (the second call on a different shape, so stale scratch of the first call would
show), in both precisions; the size limit is refused with a clear message.

This file sorts after every other test file on purpose (`pytest --collect-only`)."""
import ctypes as C

import numpy as np
import pytest
import torch

from rte_rrtmgp_cpp_b200 import capi

pytestmark = pytest.mark.gpu
TOL = {torch.float64: 1e-5, torch.float32: 0.05}          # W/m2, the bars of the other GPU tests (BASELINE.json)


def _inputs(ncol, nlay, ngpt, dtype, seed):
    """(ncol, nlay, ngpt) column-major == torch (ngpt, nlay, ncol) contiguous."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    r = lambda *s: torch.rand(*s, device="cuda", dtype=torch.float64, generator=g)
    tau = (10.0 ** (r(ngpt, nlay, ncol) * 4.0 - 3.0)).to(dtype)          # 1e-3 .. 10
    ssa = (r(ngpt, nlay, ncol) * 0.999).to(dtype)
    asy = (r(ngpt, nlay, ncol) * 0.85).to(dtype)
    mu0 = (r(ncol) * 0.95 + 0.05).to(dtype)
    mu0[::7] = 0.0                                                       # night columns
    alb_dir = (r(ngpt, ncol) * 0.3 + 0.05).to(dtype)
    alb_dif = (r(ngpt, ncol) * 0.3 + 0.05).to(dtype)
    inc_dir = (r(ngpt, ncol) * 5.0 + 1.0).to(dtype)
    inc_dif = (r(ngpt, ncol) * 0.5).to(dtype)
    return tau, ssa, asy, mu0, alb_dir, alb_dif, inc_dir, inc_dif


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _both(ncol, nlay, ngpt, top_at_1, dtype, seed, with_dif):
    L = capi.lib()
    code = capi.F32 if dtype == torch.float32 else capi.F64
    ins = list(_inputs(ncol, nlay, ngpt, dtype, seed))
    if not with_dif:
        ins[-1] = None
    out = lambda: [torch.full((ngpt, nlay + 1, ncol), float("nan"), device="cuda", dtype=dtype) for _ in range(3)]
    ref, got = out(), out()
    torch.cuda.synchronize()
    capi.check(L.rrtmgp_b200_sw_solver_2stream(code, ncol, nlay, ngpt, int(top_at_1), *[_p(x) for x in ins],
                                               *[_p(x) for x in ref], 0, None))
    capi.check(L.rrtmgp_b200_sw_solver_2stream_noalloc(code, ncol, nlay, ngpt, int(top_at_1), *[_p(x) for x in ins],
                                                       *[_p(x) for x in got], None))
    torch.cuda.synchronize()
    return ref, got


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_noalloc_matches_the_allocating_entry_point_twice_in_a_row(dtype):
    # two calls in a row on one process: different shapes, orientation and boundary condition, so the second call
    # finds the first one's values in the shared scratch
    for call, (ncol, nlay, ngpt, top, dif) in enumerate([(150, 33, 16, True, True), (97, 60, 9, False, False)]):
        ref, got = _both(ncol, nlay, ngpt, top, dtype, seed=1234 + call, with_dif=dif)
        for name, a, b in zip(("flux_up", "flux_dn", "flux_dir"), ref, got):
            assert torch.isfinite(a).all() and torch.isfinite(b).all(), (call, name)
            d = float((a.double() - b.double()).abs().max())
            assert d <= TOL[dtype], f"call {call} {name}: max abs diff {d}"
            assert float(a.abs().max()) > 0.1                                  # a real signal was compared


def test_noalloc_refuses_sizes_beyond_the_static_scratch():
    L = capi.lib()
    one = torch.zeros(8, device="cuda")
    p = _p(one)
    # 2 097 152 cells is the fp32 limit (32 MiB / (4 arrays * 4 bytes)); arguments are checked before any launch
    rc = L.rrtmgp_b200_sw_solver_2stream_noalloc(capi.F32, 1024, 129, 16, 1, p, p, p, p, p, p, p, None, p, p, p, None)
    assert rc == capi.E_ARG
    msg = L.rrtmgp_b200_last_error().decode()
    assert "exceeds the static scratch" in msg and "2097152" in msg
    rc = L.rrtmgp_b200_sw_solver_2stream_noalloc(capi.F64, 1024, 65, 16, 1, p, p, p, p, p, p, p, None, p, p, p, None)
    assert rc == capi.E_ARG and "1048576" in L.rrtmgp_b200_last_error().decode()
    # exactly at the limit is accepted (tiny columns, many of them: tau = 0 everywhere is a valid atmosphere)
    ncol, nlay, ngpt = 8192, 16, 16
    z = lambda *s: torch.zeros(*s, device="cuda", dtype=torch.float32)
    tau, mu0, a, inc = z(ngpt, nlay, ncol), z(ncol) + 0.5, z(ngpt, ncol) + 0.1, z(ngpt, ncol) + 1.0
    up, dn, dr = (z(ngpt, nlay + 1, ncol) for _ in range(3))
    capi.check(L.rrtmgp_b200_sw_solver_2stream_noalloc(capi.F32, ncol, nlay, ngpt, 1, _p(tau), _p(tau), _p(tau), _p(mu0), _p(a), _p(a),
                                                       _p(inc), None, _p(up), _p(dn), _p(dr), None))
    torch.cuda.synchronize()
    assert torch.allclose(dr, torch.full_like(dr, 0.5)) and torch.allclose(up, torch.full_like(up, 0.05))
