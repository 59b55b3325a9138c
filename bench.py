#!/usr/bin/env python
"""bench.py — LW+SW column-g-points per second on the named BASELINE.json workload.

  python bench.py --gpus N --steps K --warmup W            (this build; torchrun launches N ranks)
  python bench.py --impl reference --gpus N --steps K --warmup W   (reference CPU path, rank 0 only)

One step = one LW pass + one SW pass (gas optics -> solver -> broadband reduction) over the whole
1024x1024-column x 128-layer domain, column-sharded over the N ranks, plus the gather of the
broadband fluxes to rank 0 (NCCL) when N > 1.  One column-g-point = the complete vertical
calculation of one column for one g-point; a step is ncol*(256+224) of them.
`value`: inputs resident in HBM, CUDA-event timed, max over ranks.
`e2e`  : the same call with host (pinned) buffers: H2D of every input + D2H of every flux per step.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NCOL_DEFAULT = 1024 * 1024
NLAY_DEFAULT = 128
NBASE = 16384          # distinct seeded S-RCE columns, tiled over the domain
GASES = ["h2o", "co2", "o3", "n2o", "ch4", "n2", "o2"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ncol", type=int, default=NCOL_DEFAULT, help="columns of the whole domain")
    ap.add_argument("--nlay", type=int, default=NLAY_DEFAULT)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--cloudy", action="store_true", help="all-sky: cloud optics LUT + delta scaling")
    ap.add_argument("--reduced", action="store_true", help="reduced g-point sets LW 128 / SW 112 (BASELINE config 5)")
    ap.add_argument("--tile", type=int, default=0, help="columns per tile of the fused path (0 = auto)")
    ap.add_argument("--chunk", type=int, default=0, help="columns per chunk of the host-buffer pipeline (0 = auto)")
    ap.add_argument("--no-pair", action="store_true", help="longwave and shortwave as two calls instead of the paired call (rrtmgp_b200_lwsw_fluxes)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip timing the reference's own CUDA path (oracle/_ref/librefcuda_*.so)")
    ap.add_argument("--ref-cuda-cols", type=int, default=65536)
    ap.add_argument("--cpu-cols-per-core", type=int, default=2048)
    return ap.parse_args()


# ------------------------------------------------------------------------------------------
# CPU baseline: the reference's own class layer (oracle/_ref, built from /root/reference/src over
# the restated kernels) when it travelled with the repo, else the oracle port.  One instance per
# host core over disjoint column blocks (the reference CPU path itself is single-threaded).
# ------------------------------------------------------------------------------------------
def _cpu_worker(job):
    seed, ncols, nlay, block, cloudy = job
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["ORACLE_REF_FAST"] = "1"          # the -O3 build of the reference classes (oracle/Makefile ref_fast): timing only
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
    from oracle import ref as R
    from oracle import oracle as O
    atm = rce_atmosphere(ncols, nlay, seed=seed)
    kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    use_ref = R.available()
    if use_ref:
        glw, gsw = R.RefGasOptics(kd_lw, GASES), R.RefGasOptics(kd_sw, GASES)
    else:
        glw, gsw = O.GasOptics(kd_lw, GASES), O.GasOptics(kd_sw, GASES)
    t0 = time.perf_counter()
    for c0 in range(0, ncols, block):
        a = atm.subset(c0, min(c0 + block, ncols))
        if use_ref:
            cd = glw.get_col_dry(a.vmr["h2o"], a.plev)
            st = glw.gas_optics_lw(a.play, a.plev, a.tlay, a.tsfc, a.vmr, cd, a.tlev)
            glw.rte_lw(st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], a.emis_sfc, a.top_at_1, 1, True)
            ss = gsw.gas_optics_sw(a.play, a.plev, a.tlay, a.vmr, cd)
            gsw.rte_sw(ss["tau"], ss["ssa"], ss["g"], a.mu0, ss["toa_src"], a.sfc_alb_dir, a.sfc_alb_dif, a.top_at_1, True)
        else:
            st = glw.gas_optics_lw(a.play, a.plev, a.tlay, a.tsfc, a.vmr, None, a.tlev)
            O.rte_lw(glw, st["tau"], st["lay_src"], st["lev_src"], st["sfc_src"], a.emis_sfc, a.top_at_1, broadband=True)
            ss = gsw.gas_optics_sw(a.play, a.plev, a.tlay, a.vmr)
            O.rte_sw(gsw, ss["tau"], ss["ssa"], ss["g"], a.mu0, ss["toa_src"], a.sfc_alb_dir, a.sfc_alb_dif, a.top_at_1, broadband=True)
    return time.perf_counter() - t0, use_ref, (R.FLAGS["fast" if R.fast_selected() else "parity"] if use_ref else "gcc -O2 -ffp-contract=off (port)")


def cpu_baseline(nlay: int, cols_per_core: int, cores: int | None = None):
    """-> dict(value, unit, cores, kind, sample).  Runs in spawned processes (no CUDA state)."""
    import multiprocessing as mp
    cores = cores or min(os.cpu_count() or 1, 64)
    block = 12                                   # the reference's CPU column block (Radiation_solver.cpp:409)
    jobs = [(1000 + i, cols_per_core, nlay, block, False) for i in range(cores)]
    ctx = mp.get_context("spawn")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    tmax = max(r[0] for r in res)
    use_ref = all(r[1] for r in res)
    units = cores * cols_per_core * (256 + 224)
    return dict(value=units / tmax, unit="column-gpts/s", cores=cores,
                kind="reference" if use_ref else "port", build_flags=res[0][2],
                sample=f"{cores * cols_per_core} seeded S-RCE columns x {nlay} layers, LW g256 + SW g224, fp64, "
                       f"{cores} processes x {cols_per_core} columns in blocks of {block}; compute time max over "
                       f"processes {tmax:.1f} s (pool wall {wall:.1f} s)",
                seconds=tmax)


# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        sm, smax, reasons, power = [], None, set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t0 or ts > t1 + 0.2:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); smax = float(f[1]); power.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
def tile_columns(cols, ncol: int, shift: int):
    """Replicate the NBASE distinct columns of `cols` (a DeviceColumns) to `ncol` columns on the
    device, rolled by `shift` so ranks do not hold identical shards."""
    import torch
    from rte_rrtmgp_cpp_b200.api import Array_gpu
    nb = cols.ncol
    reps = (ncol + nb - 1) // nb

    def t2(a):        # (nb, n2) column-major -> (ncol, n2)
        n2 = a.dims[1]
        v = torch.roll(a.t.view(n2, nb), shift, dims=1).repeat(1, reps)[:, :ncol].contiguous()
        return Array_gpu((ncol, n2), a.t.dtype, data=v.view(-1))

    def tb(a):        # (nbnd, nb) column-major -> (nbnd, ncol)
        n1 = a.dims[0]
        v = torch.roll(a.t.view(nb, n1), shift, dims=0).repeat(reps, 1)[:ncol].contiguous()
        return Array_gpu((n1, ncol), a.t.dtype, data=v.view(-1))

    def t1(a):
        v = torch.roll(a.t, shift).repeat(reps)[:ncol].contiguous()
        return Array_gpu((ncol,), a.t.dtype, data=v)

    for k in ("play", "plev", "tlay", "tlev", "lwp", "iwp", "rel", "dei"):
        if getattr(cols, k) is not None:
            setattr(cols, k, t2(getattr(cols, k)))
    for k in ("emis_sfc", "sfc_alb_dir", "sfc_alb_dif"):
        setattr(cols, k, tb(getattr(cols, k)))
    for k in ("tsfc", "mu0", "tsi_scaling"):
        setattr(cols, k, t1(getattr(cols, k)))
    for name, a in list(cols.gas_concs.gas_concs_map.items()):
        if a.dims[0] > 1:
            cols.gas_concs.gas_concs_map[name] = t2(a)
    cols.ncol = ncol
    return cols


def input_arrays(cols, sw_only=False):
    from rte_rrtmgp_cpp_b200.api import Array_gpu
    arrs = [v for v in vars(cols).values() if isinstance(v, Array_gpu)]
    arrs += list(cols.gas_concs.gas_concs_map.values())
    return arrs


def run_b200(args):
    import torch
    import torch.distributed as dist
    from rte_rrtmgp_cpp_b200 import api, capi
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with torchrun (one process per GPU)")
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args.nlay, args.cpu_cols_per_core)       # before any CUDA state exists
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: there is no CPU fallback"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dtype = torch.float32 if args.dtype == "f32" else torch.float64
    esz = 4 if args.dtype == "f32" else 8

    # column shard of this rank: contiguous [c0, c1) of the x-fastest flattened domain
    from rte_rrtmgp_cpp_b200 import shard
    c0, c1 = shard.shard_bounds(args.ncol, world, rank)
    ncol = c1 - c0
    nlay, nlev = args.nlay, args.nlay + 1
    base = rce_atmosphere(min(NBASE, ncol), nlay, cloudy=args.cloudy)
    cols = tile_columns(api.DeviceColumns(base, dtype), ncol, shift=37 * rank)
    if args.reduced:
        from rte_rrtmgp_cpp_b200.kdist import REDUCED_LW_128, REDUCED_SW_112
        kd_lw, kd_sw = synthetic_kdist("lw", REDUCED_LW_128), synthetic_kdist("sw", REDUCED_SW_112)
    else:
        kd_lw, kd_sw = synthetic_kdist("lw"), synthetic_kdist("sw")
    g_lw, g_sw = kd_lw.ngpt, kd_sw.ngpt
    lw = api.Radiation_solver_longwave(kd_lw, GASES, synthetic_cloud_lut("lw") if args.cloudy else None, dtype)
    sw = api.Radiation_solver_shortwave(kd_sw, GASES, synthetic_cloud_lut("sw") if args.cloudy else None, dtype)
    if args.tile:
        lw.set_tile_columns(args.tile); sw.set_tile_columns(args.tile)
    ngpt = lw.kdist.get_ngpt() + sw.kdist.get_ngpt()
    # col_dry is an input of the class API (Radiation_solver.cu:458); computed once, outside the step
    cols.col_dry = api.Array_gpu((ncol, nlay), dtype)
    lw.kdist.get_col_dry(cols.col_dry, cols.gas_concs.get_vmr("h2o"), cols.plev)

    out_lw, out_sw = {}, {}
    flux_names = [("lw", k) for k in ("flux_up", "flux_dn", "flux_net")] + \
                 [("sw", k) for k in ("flux_up", "flux_dn", "flux_dn_dir", "flux_net")]
    gather_buf = None

    def step():
        if args.no_pair:
            lw.solve_gpu(cols, switch_cloud_optics=args.cloudy, out=out_lw)
            sw.solve_gpu(cols, switch_cloud_optics=args.cloudy, out=out_sw)
        else:       # one call: longwave and shortwave CTAs co-resident, band pair by band pair
            api.solve_pair(lw, sw, cols, switch_cloud_optics=args.cloudy, out_lw=out_lw, out_sw=out_sw)
        if world > 1:
            # final flux gather to rank 0 over NCCL: (rank, field, level, column-in-shard)
            nonlocal gather_buf
            packed = shard.pack_fields([(out_lw if w == "lw" else out_sw)[k].t for w, k in flux_names])
            gather_buf = shard.gather_fields(packed, args.ncol, nlev, dst=0, recv=gather_buf)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    L = capi.lib()
    L.rrtmgp_b200_profile(1, None, None, None, None)
    launches0 = L.rrtmgp_b200_launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    barrier()
    t_wall1 = time.time()
    ms = ev0.elapsed_time(ev1)
    launches = L.rrtmgp_b200_launch_count() - launches0
    lw_ms, sw_ms = C.c_double(), C.c_double(); lw_n, sw_n = C.c_longlong(), C.c_longlong()
    L.rrtmgp_b200_profile(0, C.byref(lw_ms), C.byref(lw_n), C.byref(sw_ms), C.byref(sw_n))
    pair_ms, pair_n = C.c_double(), C.c_longlong()
    L.rrtmgp_b200_profile_pair(C.byref(pair_ms), C.byref(pair_n))
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    units_per_step = args.ncol * ngpt
    value = units_per_step / (ms_per_step * 1e-3)

    # ---- e2e: the host-buffer entry of the C ABI (rrtmgp_b200_fluxes_host): pinned host arrays in,
    # pinned host arrays out; H2D, col_dry, LW+SW and D2H all inside the timed region, per rank on
    # its own column shard (the host model keeps its own decomposition: no gather).
    e2e = None
    if not args.no_e2e:
        import copy
        def to_host(a):
            h = torch.empty(a.t.shape, dtype=a.t.dtype, pin_memory=True); h.copy_(a.t)
            return api.Array_gpu(a.dims, a.t.dtype, "cpu", h)
        hcols = copy.copy(cols)
        hcols.col_dry = None
        h2d = 0
        for k, v in list(vars(hcols).items()):
            if isinstance(v, api.Array_gpu):
                used = args.cloudy or k not in ("lwp", "iwp", "rel", "dei")
                setattr(hcols, k, to_host(v))
                h2d += v.t.numel() * v.t.element_size() if used else 0
        hcols.gas_concs = api.Gas_concs_gpu(dtype, "cpu")
        for name, a in cols.gas_concs.gas_concs_map.items():
            hcols.gas_concs.gas_concs_map[name] = to_host(a)
            h2d += a.t.numel() * a.t.element_size()
        torch.cuda.synchronize()
        h_lw, h_sw = {}, {}

        def e2e_step():
            api.solve_host(lw, sw, hcols, switch_cloud_optics=args.cloudy, out_lw=h_lw, out_sw=h_sw, chunk_columns=args.chunk)

        e2e_step(); e2e_step(); barrier()
        d2h = sum(a.t.numel() * a.t.element_size() for o in (h_lw, h_sw) for a in o.values())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        for _ in range(args.steps):
            e2e_step()
        e1.record(); barrier()
        ems = max(e0.elapsed_time(e1), (time.time() - t0) * 1e3)     # the calls block: wall clock and events agree; keep the larger
        if world > 1:
            t = torch.tensor([ems], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX); ems = float(t.item())
            hb = torch.tensor([h2d, d2h], device="cuda", dtype=torch.float64)
            dist.all_reduce(hb); h2d, d2h = int(hb[0].item()), int(hb[1].item())
        # the device-resident and the host-buffer paths run the same kernels: results must agree
        same = all(torch.equal(h_lw[k].t, out_lw[k].t.cpu()) for k in ("flux_up", "flux_dn")) and \
               all(torch.equal(h_sw[k].t, out_sw[k].t.cpu()) for k in ("flux_up", "flux_dn", "flux_dn_dir"))
        e2e = {"value": units_per_step / (ems / args.steps * 1e-3), "unit": "column-gpts/s",
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": ems / args.steps,
               "api": "rrtmgp_b200_fluxes_host (pinned host buffers, chunked copy/compute pipeline)",
               "matches_device_path": bool(same)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (SURVEY 8d byte model; launches = one band of one tile)
    L_, V_ = nlay, nlev
    b_lw, b_sw = (4 * L_ + 6 * V_) * esz, (6 * L_ + 6 * V_) * esz
    paired = pair_ms.value > 0
    if paired:
        # one launch = one longwave band + one shortwave band of a tile (plus the few unpaired single-role launches of the
        # same loops): algorithmic bytes of both
        dom, k_ms, k_n = "lwsw_band_kernel", pair_ms.value, pair_n.value
        alg_bytes_total = ncol * (g_lw * b_lw + g_sw * b_sw) * args.steps
    else:
        dom = "sw_band_kernel" if sw_ms.value >= lw_ms.value else "lw_band_kernel"
        k_ms, k_n, k_b, k_g = (sw_ms.value, sw_n.value, b_sw, sw.kdist.get_ngpt()) if dom == "sw_band_kernel" else \
                              (lw_ms.value, lw_n.value, b_lw, lw.kdist.get_ngpt())
        alg_bytes_total = ncol * k_g * k_b * args.steps           # this rank, all launches of that kernel
    achieved = alg_bytes_total / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak, peak_src = float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        pass
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"{dom}_{args.dtype}")
    except Exception:
        pass
    # ---- the reference's own CUDA implementation of the path on this GPU (secondary baseline):
    # separate process (its memory pool, tuner file and any fault stay out of this one), after our timing
    ref_cuda = None
    if world == 1 and not args.no_ref_cuda and not args.cloudy and not args.reduced:
        try:
            from oracle import refcuda
            if refcuda.available(args.dtype):
                torch.cuda.empty_cache()
                r = subprocess.run([sys.executable, "-m", "oracle.refcuda", "--ncol", str(min(args.ref_cuda_cols, args.ncol)),
                                    "--nlay", str(nlay), "--dtype", args.dtype, "--reps", "3"], cwd=ROOT,
                                   capture_output=True, text=True, timeout=600)
                for ln in r.stdout.splitlines():
                    if ln.startswith("{"):
                        ref_cuda = json.loads(ln)
                if ref_cuda is None:
                    ref_cuda = {"unavailable": (r.stderr or r.stdout).strip().splitlines()[-1][:200] if (r.stderr or r.stdout).strip() else "no output"}
                else:
                    ref_cuda["note"] = ("unmodified reference src_cuda + src_kernels_cuda built for sm_100a, driven like "
                                        "Radiation_solver::solve_gpu; its fluxes are NOT the parity target (SURVEY 2.1 quirks)")
        except Exception as ex:      # noqa: BLE001 - a baseline failure must not lose the measurement
            ref_cuda = {"unavailable": str(ex)[:200]}
    whole = (ncol * (g_lw * b_lw + g_sw * b_sw) * args.steps) / ((lw_ms.value + sw_ms.value + pair_ms.value) * 1e-3) / 1e9
    line = {
        "metric": "LW+SW column-gpts/s", "value": value, "unit": "column-gpts/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
        "data": "synthetic (seeded S-RCE columns, synthetic k-distribution tables of the g256/g224 shape: "
                "the rrtmgp-gas-*.nc files are absent from the reference mount)",
        "config": {"workload": f"{args.ncol} columns x {nlay} layers, LW {g_lw} + SW {g_sw} g-points, "
                               f"{'all-sky (cloud LUT + delta scaling)' if args.cloudy else 'clear-sky'}, fused path, "
                               f"column-sharded over {world} GPU(s) + NCCL flux gather to rank 0",
                   "ncol": args.ncol, "nlay": nlay, "ngpt_lw": g_lw, "ngpt_sw": g_sw,
                   "distinct_columns": min(NBASE, ncol), "l2_policy": "inputs and scratch larger than L2 (GBs per step)",
                   "tile_columns": args.tile or "auto (one resident wave)"},
        "clocks": clocks, "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_column_gpt": {"lw": b_lw, "sw": b_sw},
                     "avg_launch_ms": k_ms / max(k_n, 1), "launches_timed": int(k_n),
                     "lw_band_ms_total": lw_ms.value, "sw_band_ms_total": sw_ms.value, "paired_band_ms_total": pair_ms.value,
                     "lw_plus_sw_effective_GBs": whole, "lw_plus_sw_frac": whole / peak,
                     "note": "effective bandwidth on the materialised-API byte model; the fused kernels move far fewer DRAM bytes"},
        "cpu_baseline": cpu, "e2e": e2e, "reference_cuda": ref_cuda,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals, samples = [], None
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(args.nlay, max(4, args.cpu_cols_per_core // 4))
        if i >= args.warmup:
            vals.append(r)
    v = float(np.mean([r["value"] for r in vals]))
    sec = float(np.mean([r["seconds"] for r in vals]))
    r = vals[-1]
    line = {"impl": "reference", "metric": "LW+SW column-gpts/s", "value": v, "unit": "column-gpts/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"bounded sample of the {args.ncol} columns x {args.nlay} layers LW 256 + SW 224 clear-sky workload: "
                                   + r["sample"], "ncol": args.ncol, "nlay": args.nlay},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")} | {"value": v},
            "e2e": {"value": v, "unit": "column-gpts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
