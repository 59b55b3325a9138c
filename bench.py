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

# BASELINE.json `configs`, in its order.  `big` (config 4) is the one the metric is quoted on and the default.
CONFIGS = {
    "rfmip":   dict(n=1, ncol=1800, nlay=60, cloudy=False, reduced=False, calls=18,
                    what="RFMIP clear-sky: 18 calls x 100 sites x 60 layers per step, as rfmip/rfmip_run.py drives the reference "
                         "(the 100 real sites of experiments 0 and 3 of the committed fixture, alternating)"),
    "allsky":  dict(n=2, ncol=16384, nlay=72, cloudy=True, reduced=False, calls=1,
                    what="all-sky: 16384 S-ALLSKY columns x 72 layers, liquid + ice cloud optics LUT, delta scaling"),
    "rce512":  dict(n=3, ncol=512 * 512, nlay=128, cloudy=True, reduced=False, calls=1,
                    what="RCEMIP-style slab 512 x 512 columns x 128 layers, all-sky (cloudy superset of the clear-sky path)"),
    "big":     dict(n=4, ncol=1024 * 1024, nlay=128, cloudy=False, reduced=False, calls=1,
                    what="1024 x 1024 columns x 128 layers, clear-sky"),
    "reduced": dict(n=5, ncol=1024 * 1024, nlay=128, cloudy=False, reduced=True, calls=1,
                    what="1024 x 1024 columns x 128 layers, reduced g-point sets LW 128 / SW 112"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="big", choices=list(CONFIGS), help="BASELINE.json configuration (default: big = config 4)")
    ap.add_argument("--ncol", type=int, default=None, help="columns of the whole domain (overrides the configuration's)")
    ap.add_argument("--nlay", type=int, default=None)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--cloudy", action="store_true", help="all-sky: cloud optics LUT + delta scaling")
    ap.add_argument("--reduced", action="store_true", help="reduced g-point sets LW 128 / SW 112 (BASELINE config 5)")
    ap.add_argument("--parts", type=int, default=1, help="N > 1: process the shard in this many column blocks, each gathered on a side stream")
    ap.add_argument("--nccl-gather", action="store_true", help="N > 1: rrtmgp_b200_gather_fluxes without the peer-memory registration")
    ap.add_argument("--torch-gather", action="store_true", help="N > 1: gather with torch.distributed.gather instead of rrtmgp_b200_gather_fluxes")
    ap.add_argument("--tile", type=int, default=0, help="columns per tile of the fused path (0 = auto)")
    ap.add_argument("--chunk", type=int, default=0, help="columns per chunk of the host-buffer pipeline (0 = auto)")
    ap.add_argument("--no-pair", action="store_true", help="longwave and shortwave as two calls instead of the paired call (rrtmgp_b200_lwsw_fluxes)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-affinity", action="store_true", help="multi-GPU runs: do not bind a rank's host threads to its GPU's NUMA node")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip timing the reference's own CUDA path (oracle/_ref/librefcuda_*.so)")
    ap.add_argument("--ref-cuda-cols", type=int, default=65536)
    ap.add_argument("--cpu-cols-per-core", type=int, default=2048)
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    a.ncol = cfg["ncol"] if a.ncol is None else a.ncol
    a.nlay = cfg["nlay"] if a.nlay is None else a.nlay
    a.cloudy = a.cloudy or cfg["cloudy"]
    a.reduced = a.reduced or cfg["reduced"]
    a.calls = cfg["calls"] if a.ncol == cfg["ncol"] else 1
    return a


# ------------------------------------------------------------------------------------------
# CPU baseline: the reference's own class layer (oracle/_ref, built from /root/reference/src over
# the restated kernels) when it travelled with the repo, else the oracle port.  One instance per
# host core over disjoint column blocks (the reference CPU path itself is single-threaded).
# ------------------------------------------------------------------------------------------
def _cpu_worker(job):
    """One process = one core: the reference's column loop (blocks of 12 columns) over its share of the sample."""
    kind, seed, ncols, nlay, block, cloudy, reduced = job
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["ORACLE_REF_FAST"] = "1"          # the -O3 build of the reference classes (oracle/Makefile ref_fast): timing only
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, REDUCED_LW_128, REDUCED_SW_112
    from oracle import ref as R
    from oracle import oracle as O
    if kind == "rfmip":                           # `seed` = experiment index (0 or 3 of the committed fixture)
        import real_data as RD
        atm = RD.rfmip_atmosphere(RD.fixture(), seed, tsi_ref=RD.default_tsi(synthetic_kdist("sw")))
        atm.vmr = {k: v for k, v in atm.vmr.items() if k in GASES}
        ncols = atm.ncol
    elif kind == "allsky":
        atm = allsky_atmosphere(ncols, nlay, seed=seed, random_clouds=True)
    else:
        atm = rce_atmosphere(ncols, nlay, seed=seed, cloudy=cloudy)
    kd_lw = synthetic_kdist("lw", REDUCED_LW_128) if reduced else synthetic_kdist("lw")
    kd_sw = synthetic_kdist("sw", REDUCED_SW_112) if reduced else synthetic_kdist("sw")
    lut_lw, lut_sw = (synthetic_cloud_lut("lw"), synthetic_cloud_lut("sw")) if cloudy else (None, None)
    use_ref = R.available()
    if use_ref:
        glw, gsw = R.RefGasOptics(kd_lw, GASES), R.RefGasOptics(kd_sw, GASES)
    else:
        glw, gsw = O.GasOptics(kd_lw, GASES), O.GasOptics(kd_sw, GASES)
        clw, csw = (O.CloudOptics(lut_lw), O.CloudOptics(lut_sw)) if cloudy else (None, None)
    t0 = time.perf_counter()
    for c0 in range(0, ncols, block):
        a = atm.subset(c0, min(c0 + block, ncols))
        if use_ref:
            cd = glw.get_col_dry(a.vmr["h2o"], a.plev)
            st = glw.gas_optics_lw(a.play, a.plev, a.tlay, a.tsfc, a.vmr, cd, a.tlev)
            tau = glw.cloud_add(lut_lw, 1, False, a.lwp, a.iwp, a.rel, a.dei, st["tau"])["tau"] if cloudy else st["tau"]
            glw.rte_lw(tau, st["lay_src"], st["lev_src"], st["sfc_src"], a.emis_sfc, a.top_at_1, 1, True)
            ss = gsw.gas_optics_sw(a.play, a.plev, a.tlay, a.vmr, cd)
            if cloudy:
                ss.update(gsw.cloud_add(lut_sw, 2, True, a.lwp, a.iwp, a.rel, a.dei, ss["tau"], ss["ssa"], ss["g"]))
            gsw.rte_sw(ss["tau"], ss["ssa"], ss["g"], a.mu0, ss["toa_src"], a.sfc_alb_dir, a.sfc_alb_dif, a.top_at_1, True)
        else:
            st = glw.gas_optics_lw(a.play, a.plev, a.tlay, a.tsfc, a.vmr, None, a.tlev)
            tau = O.add_to_1scl(glw, st["tau"], clw.cloud_optics_1scl(a.lwp, a.iwp, a.rel, a.dei)) if cloudy else st["tau"]
            O.rte_lw(glw, tau, st["lay_src"], st["lev_src"], st["sfc_src"], a.emis_sfc, a.top_at_1, broadband=True)
            ss = gsw.gas_optics_sw(a.play, a.plev, a.tlay, a.vmr)
            t3 = (ss["tau"], ss["ssa"], ss["g"])
            if cloudy:
                t3 = O.add_to_2str(gsw, *t3, *O.delta_scale(*csw.cloud_optics_2str(a.lwp, a.iwp, a.rel, a.dei)))
            O.rte_sw(gsw, *t3, a.mu0, ss["toa_src"], a.sfc_alb_dir, a.sfc_alb_dif, a.top_at_1, broadband=True)
    flags = (R.FLAGS["fast" if R.fast_selected() else "parity"]) if use_ref else "gcc -O2 -ffp-contract=off (port)"
    return time.perf_counter() - t0, use_ref, flags, ncols, kd_lw.ngpt + kd_sw.ngpt


def cpu_baseline(args, cols_per_core: int, cores: int | None = None):
    """The reference's CPU path on the configuration's workload -> dict(value, unit, cores, kind, sample).  RFMIP: the 1800
    columns in full (18 experiments over the cores); the large configurations: a bounded sample of seeded columns of the same
    kind, one share per core, scaled linearly (columns are independent).  Runs in spawned processes (no CUDA state)."""
    import multiprocessing as mp
    cores = cores or min(os.cpu_count() or 1, 64)
    block = 12                                   # the reference's CPU column block (Radiation_solver.cpp:409)
    if args.config == "rfmip" and args.calls > 1:
        jobs = [("rfmip", (0, 3)[i % 2], 100, args.nlay, block, False, args.reduced) for i in range(18)]
        what = "all 1800 RFMIP columns (18 experiments x 100 sites), no extrapolation"
    else:
        kind = "allsky" if args.config == "allsky" else "rce"
        jobs = [(kind, 1000 + i, cols_per_core, args.nlay, block, args.cloudy, args.reduced) for i in range(cores)]
        what = f"{cores * cols_per_core} seeded {'S-ALLSKY' if kind == 'allsky' else 'S-RCE'} columns{' with clouds' if args.cloudy else ''}, scaled linearly to the domain"
    ctx = mp.get_context("spawn")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        res = pool.map(_cpu_worker, jobs)
    wall = time.perf_counter() - t0
    use_ref = all(r[1] for r in res)
    units = sum(r[3] * r[4] for r in res)
    # the time the sample takes with `cores` workers, excluding interpreter start-up and table set-up: one job per core ->
    # the slowest worker's compute time; RFMIP (18 jobs) -> summed compute time / cores, i.e. perfectly balanced
    tsec = sum(r[0] for r in res) / cores if len(jobs) > cores else max(r[0] for r in res)
    return dict(value=units / tsec, unit="column-gpts/s", cores=cores,
                kind="reference" if use_ref else "port", build_flags=res[0][2],
                sample=f"{what}; {args.nlay} layers, {res[0][4]} g-points, fp64, {min(cores, len(jobs))} processes, column blocks of {block}; "
                       f"{tsec:.1f} s (pool wall {wall:.1f} s)",
                seconds=tsec)


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


def build_columns(args, api, dtype, ncol, rank):
    """Device-resident inputs of this rank: a list of (DeviceColumns, ncol) calls making up one step."""
    from rte_rrtmgp_cpp_b200.atmos import rce_atmosphere, allsky_atmosphere
    if args.config == "rfmip" and args.calls > 1:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import real_data as RD
        from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist
        z = RD.fixture()
        out = []
        for i in range(args.calls):
            atm = RD.rfmip_atmosphere(z, (0, 3)[i % 2], tsi_ref=RD.default_tsi(synthetic_kdist("sw")))
            atm.vmr = {k: v for k, v in atm.vmr.items() if k in GASES}
            out.append(api.DeviceColumns(atm, dtype))
        return out
    nparts = max(args.parts, 1)
    from rte_rrtmgp_cpp_b200 import shard
    out = []
    for part in range(nparts):
        p0, p1 = shard.shard_bounds(ncol, nparts, part)
        n = p1 - p0
        nb = min(NBASE, n)
        base = allsky_atmosphere(nb, args.nlay, random_clouds=True) if args.config == "allsky" else rce_atmosphere(nb, args.nlay, cloudy=args.cloudy)
        out.append(tile_columns(api.DeviceColumns(base, dtype), n, shift=37 * rank + 11 * part))
    return out


def bind_to_gpu_numa_node(local: int):
    """Several ranks on one node: run this rank's host threads on the CPUs NVML reports as local to its GPU, so the pinned host
    buffers of the e2e leg are first-touched on that NUMA node and the PCIe copies do not cross sockets (what `numactl` in a
    launch script does).  Returns the number of CPUs bound, or None when NVML has no answer."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    from rte_rrtmgp_cpp_b200 import api, capi
    from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 with torchrun (one process per GPU)")
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu = cpu_baseline(args, args.cpu_cols_per_core)       # before any CUDA state exists
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: there is no CPU fallback"
    torch.cuda.set_device(local)
    affinity = bind_to_gpu_numa_node(local) if world > 1 and not args.no_affinity else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dtype = torch.float32 if args.dtype == "f32" else torch.float64
    esz = 4 if args.dtype == "f32" else 8

    # column shard of this rank: contiguous [c0, c1) of the x-fastest flattened domain
    from rte_rrtmgp_cpp_b200 import shard
    c0, c1 = shard.shard_bounds(args.ncol, world, rank)
    ncol = c1 - c0
    nlay, nlev = args.nlay, args.nlay + 1
    calls = build_columns(args, api, dtype, ncol, rank)
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
    for cols in calls:
        cols.col_dry = api.Array_gpu((cols.ncol, nlay), dtype)
        lw.kdist.get_col_dry(cols.col_dry, cols.gas_concs.get_vmr("h2o"), cols.plev)

    outs = [({}, {}) for _ in calls]                 # (longwave, shortwave) outputs of every call of a step
    # what rank 0 assembles at N > 1: up / down (/ direct) travel, the two net fluxes are formed on the root (dn - up)
    sent = [("lw", "flux_up"), ("lw", "flux_dn"), ("sw", "flux_up"), ("sw", "flux_dn"), ("sw", "flux_dn_dir")]
    plan = full = net = None
    gather_mode = None
    gather_buf = None
    side = torch.cuda.Stream() if world > 1 else None
    ev_g = []
    if world > 1 and not args.torch_gather:
        plan = shard.ShardPlan(args.ncol, world, rank)
        if rank == 0:
            full = [torch.empty(nlev * args.ncol, device="cuda", dtype=dtype) for _ in sent]
            net = [torch.empty(nlev * args.ncol, device="cuda", dtype=dtype) for _ in range(2)]
        # peer-memory mode: every rank writes its slabs straight into rank 0's arrays over NVLink (CUDA IPC); falls back to the
        # NCCL path when the arrays cannot be mapped
        gather_mode = "nccl (one message per field and peer, layout copies on the root)"
        if not args.nccl_gather and plan.register(full, len(sent)) == len(sent):
            gather_mode = "peer-memory writes over NVLink (CUDA IPC) + one-element completion messages"
    L = capi.lib()
    code = capi.F32 if args.dtype == "f32" else capi.F64

    def step():
        nonlocal gather_buf
        for i, cols in enumerate(calls):
            o_lw, o_sw = outs[i]
            if args.no_pair:
                lw.solve_gpu(cols, switch_cloud_optics=args.cloudy, out=o_lw)
                sw.solve_gpu(cols, switch_cloud_optics=args.cloudy, out=o_sw)
            else:       # one call: longwave and shortwave CTAs co-resident, band pair by band pair
                api.solve_pair(lw, sw, cols, switch_cloud_optics=args.cloudy, out_lw=o_lw, out_sw=o_sw)
            if plan is not None:
                # this block's slabs go to rank 0 on a side stream (rrtmgp_b200_gather_fluxes: grouped ncclSend/Recv straight
                # into the reference-layout arrays) while the next block computes
                fields = [(o_lw if w == "lw" else o_sw)[k].t for w, k in sent]
                side.wait_stream(torch.cuda.current_stream())
                if rank == 0 and len(ev_g) < 2 * 64:
                    e = torch.cuda.Event(enable_timing=True); e.record(side); ev_g.append(e)
                plan.gather(fields, full, nlev, i, len(calls), root=0, stream=side)
                if rank == 0 and len(ev_g) < 2 * 64:
                    e = torch.cuda.Event(enable_timing=True); e.record(side); ev_g.append(e)
        if plan is not None:
            torch.cuda.current_stream().wait_stream(side)
            if rank == 0:
                n = nlev * args.ncol
                capi.check(L.rrtmgp_b200_net_flux(code, n, full[1].data_ptr(), full[0].data_ptr(), net[0].data_ptr(), None))
                capi.check(L.rrtmgp_b200_net_flux(code, n, full[3].data_ptr(), full[2].data_ptr(), net[1].data_ptr(), None))
        elif world > 1:
            # torch.distributed.gather of a packed buffer (the round-1 path): (rank, field, level, column-in-shard)
            o_lw, o_sw = outs[0]
            names7 = [("lw", k) for k in ("flux_up", "flux_dn", "flux_net")] + [("sw", k) for k in ("flux_up", "flux_dn", "flux_dn_dir", "flux_net")]
            packed = shard.pack_fields([(o_lw if w == "lw" else o_sw)[k].t for w, k in names7])
            gather_buf = shard.gather_fields(packed, args.ncol, nlev, dst=0, recv=gather_buf)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    ev_g.clear()
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
    gather_ms = sum(ev_g[i].elapsed_time(ev_g[i + 1]) for i in range(0, len(ev_g) - 1, 2)) / max(len(ev_g) // 2, 1) * len(calls) if ev_g else 0.0
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    units_per_step = args.ncol * ngpt
    value = units_per_step / (ms_per_step * 1e-3)
    cols = calls[0]
    out_lw, out_sw = outs[0]

    # ---- e2e: the host-buffer entry of the C ABI (rrtmgp_b200_fluxes_host): pinned host arrays in,
    # pinned host arrays out; H2D, col_dry, LW+SW and D2H all inside the timed region, per rank on
    # its own column shard (the host model keeps its own decomposition: no gather).
    e2e = None
    if not args.no_e2e:
        import copy
        def to_host(a):
            h = torch.empty(a.t.shape, dtype=a.t.dtype, pin_memory=True); h.copy_(a.t)
            return api.Array_gpu(a.dims, a.t.dtype, "cpu", h)
        h2d = 0
        hcalls = []
        for cols in calls:
            hcols = copy.copy(cols)
            hcols.col_dry = None
            for k, v in list(vars(hcols).items()):
                if isinstance(v, api.Array_gpu):
                    used = args.cloudy or k not in ("lwp", "iwp", "rel", "dei")
                    setattr(hcols, k, to_host(v))
                    h2d += v.t.numel() * v.t.element_size() if used else 0
            hcols.gas_concs = api.Gas_concs_gpu(dtype, "cpu")
            for name, a in cols.gas_concs.gas_concs_map.items():
                hcols.gas_concs.gas_concs_map[name] = to_host(a)
                h2d += a.t.numel() * a.t.element_size()
            hcalls.append((hcols, {}, {}))
        torch.cuda.synchronize()

        def e2e_step():
            for hcols, h_lw, h_sw in hcalls:
                api.solve_host(lw, sw, hcols, switch_cloud_optics=args.cloudy, out_lw=h_lw, out_sw=h_sw, chunk_columns=args.chunk)

        e2e_step(); e2e_step(); barrier()
        d2h = sum(a.t.numel() * a.t.element_size() for _, h_lw, h_sw in hcalls for o in (h_lw, h_sw) for a in o.values())
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
        same = all(torch.equal(h_lw[k].t, o_lw[k].t.cpu()) for (_, h_lw, _), (o_lw, _) in zip(hcalls, outs) for k in ("flux_up", "flux_dn")) and \
               all(torch.equal(h_sw[k].t, o_sw[k].t.cpu()) for (_, _, h_sw), (_, o_sw) in zip(hcalls, outs) for k in ("flux_up", "flux_dn", "flux_dn_dir"))
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
    # un-pairable calls (fp64, the reduced sets' narrow bands) run the longwave and the shortwave call on two streams: their own
    # band times overlap, and the third counter holds the wall time from fork to join (both calls whole, prologues included)
    two_streams = pair_ms.value > 0 and (lw_ms.value > 0 or sw_ms.value > 0)
    band_ms_total = pair_ms.value if pair_ms.value > 0 else lw_ms.value + sw_ms.value
    small_calls = args.ncol // len(calls) <= 6144       # column-resident calls, also on two streams, but timed per call only
    paired = pair_ms.value > 0 and not two_streams
    if two_streams:
        dom, k_ms, k_n = "lw_band_kernel || sw_band_kernel (two streams)", pair_ms.value, lw_n.value + sw_n.value
        alg_bytes_total = ncol * (g_lw * b_lw + g_sw * b_sw) * args.steps
    elif paired:
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
    # DRAM bytes per launch from the committed ncu capture of the same kernel at the `big` size (profiles/traffic.json):
    # a static figure, NOT measured in this run (ncu cannot run inside a timed bench)
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(f"{dom}_{args.dtype}") if args.config == "big" and args.ncol == CONFIGS["big"]["ncol"] and world == 1 else None
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
                                    "--nlay", str(nlay), "--dtype", args.dtype, "--reps", "3", "--check"], cwd=ROOT,
                                   capture_output=True, text=True, timeout=600)
                for ln in r.stdout.splitlines():
                    if ln.startswith("{"):
                        ref_cuda = json.loads(ln)
                if ref_cuda is None:
                    ref_cuda = {"unavailable": (r.stderr or r.stdout).strip().splitlines()[-1][:200] if (r.stderr or r.stdout).strip() else "no output"}
                else:
                    ref_cuda["note"] = ("unmodified reference src_cuda + src_kernels_cuda built for sm_100a, driven like "
                                        "Radiation_solver::solve_gpu; its fluxes are NOT the parity target (SURVEY 2.1 quirks: "
                                        "max_abs_diff_vs_cpu_oracle_Wm2 is its distance from the CPU oracle on 256 of these columns; "
                                        "stage by stage the two are compared in tests/test_oracle_vs_refcuda.py)")
        except Exception as ex:      # noqa: BLE001 - a baseline failure must not lose the measurement
            ref_cuda = {"unavailable": str(ex)[:200]}
    whole = (ncol * (g_lw * b_lw + g_sw * b_sw) * args.steps) / (band_ms_total * 1e-3) / 1e9
    line = {
        "metric": "LW+SW column-gpts/s", "value": value, "unit": "column-gpts/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": args.dtype,
        "data": "synthetic (seeded columns" + (", the 100 real RFMIP sites" if args.config == "rfmip" else "") + ", synthetic k-distribution "
                "tables of the g256/g224 shape: the rrtmgp-gas-*.nc files are absent from the reference mount)",
        "config": {"workload": f"BASELINE config {CONFIGS[args.config]['n']} ({args.config}): {CONFIGS[args.config]['what']}; "
                               f"{args.ncol} columns x {nlay} layers, LW {g_lw} + SW {g_sw} g-points, "
                               f"{'all-sky (cloud LUT + delta scaling)' if args.cloudy else 'clear-sky'}, fused path"
                               + (", longwave + shortwave CTAs co-resident (rrtmgp_b200_lwsw_fluxes)" if paired else "")
                               + (f", column-sharded over {world} GPUs, flux gather to rank 0 "
                                  + ("(torch.distributed.gather)" if args.torch_gather else f"(rrtmgp_b200_gather_fluxes: {gather_mode})") if world > 1 else ""),
                   "name": args.config, "ncol": args.ncol, "nlay": nlay, "ngpt_lw": g_lw, "ngpt_sw": g_sw, "calls_per_step": len(calls),
                   "distinct_columns": min(NBASE, ncol), "l2_policy": "inputs and scratch larger than L2 (GBs per step)" if args.ncol >= 65536 else
                                       "inputs smaller than L2 (a small-call workload by definition; nothing is flushed between calls)",
                   "tile_columns": args.tile or "auto: equal tiles of at most 16 waves of CTAs (8 when the tile state would exceed a quarter of free memory)"},
        "breakdown_ms_per_step": {"band_kernels": band_ms_total / args.steps, "gather": gather_ms,
                                  "other": max(0.0, ms_per_step - band_ms_total / args.steps - (gather_ms if len(calls) == 1 else 0.0)),
                                  "note": "rank 0; other = tile prologues (interpolation state), zeroing, net flux; with several blocks the gather of "
                                          "block i overlaps the kernels of block i+1"
                                          + ("; the longwave and the shortwave call run on two streams: band_kernels is the wall time from fork to "
                                             "join and contains their prologues, lw/sw_band_ms_total are each call's own (overlapping) band times" if two_streams else "")
                                          + ("; small calls: the longwave and the shortwave call of a pair run on two streams, so band_kernels "
                                             "(each kernel timed on its own stream) overlaps itself and can exceed the wall time" if small_calls else "")},
        "clocks": clocks, "gpu_launches": int(launches), "host_cpus_bound_to_gpu_node": affinity,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_source": "static: profiles/traffic.json (ncu capture of this kernel, not measured in this run)" if traffic else None,
                     # the DRAM bytes the kernel really moves (ncu) over this run's launch time: how busy the memory system is, next to
                     # `frac`, which counts the bytes of the materialised pipeline the kernel replaces
                     "dram_achieved": traffic / (k_ms / max(k_n, 1) * 1e-3) / 1e9 if traffic and k_ms > 0 else None,
                     "dram_frac": traffic / (k_ms / max(k_n, 1) * 1e-3) / 1e9 / peak if traffic and k_ms > 0 else None,
                     "peak_source": peak_src,
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
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(args, max(4, args.cpu_cols_per_core // 4))
        if i >= args.warmup:
            vals.append(r)
    v = float(np.mean([r["value"] for r in vals]))
    sec = float(np.mean([r["seconds"] for r in vals]))
    r = vals[-1]
    g = 240 if args.reduced else 480
    line = {"impl": "reference", "metric": "LW+SW column-gpts/s", "value": v, "unit": "column-gpts/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"BASELINE config {CONFIGS[args.config]['n']} ({args.config}): {CONFIGS[args.config]['what']}; reference CPU classes on "
                                   + r["sample"], "name": args.config, "ncol": args.ncol, "nlay": args.nlay, "ngpt": g},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "build_flags")} | {"value": v},
            "e2e": {"value": v, "unit": "column-gpts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
