"""The command-line driver (cli/test_rte_rrtmgp.cpp) and its file layer.

CPU: the netCDF-classic reader/writer and the coefficient loaders against scipy.io and the packed
table structs; the driver's option handling; loud failure without a GPU (no output file).
GPU: the reference's file contract end to end — coefficient files + rte_rrtmgp_input.nc in, the
executable run in that directory, rte_rrtmgp_output.nc compared with the CPU oracle
(the way the reference's own rfmip/allsky scripts use test_rte_rrtmgp)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest
import torch
from scipy.io import netcdf_file

from rte_rrtmgp_cpp_b200 import capi, ctables, ncfiles
from rte_rrtmgp_cpp_b200.atmos import add_aerosols, allsky_atmosphere, rce_atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_aerosol_lut, synthetic_cloud_lut, synthetic_kdist
from helpers import oracle_lw, oracle_sw, as_f32_valued

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "cpp"))
sys.path.insert(0, os.path.join(ROOT, "cli"))
import build_driver  # noqa: E402
import build as cli_build  # noqa: E402


def _nc():
    build_driver.build_nc()
    d = C.CDLL(build_driver.nc_path())
    d.nc_last_error.restype = C.c_char_p
    return d


def _exe(tag):
    capi.lib()
    cli_build.build()
    return cli_build.path(tag)


@pytest.mark.parametrize("kind", ["lw", "sw"])
def test_coefficient_files_load_identically(kind, tmp_path):
    """coefficients.hpp reads a gas-optics / cloud file into the same structs ctables.py packs."""
    d = _nc()
    kd, lut = synthetic_kdist(kind), synthetic_cloud_lut(kind)
    p = str(tmp_path / f"coefficients_{kind}.nc")
    ncfiles.write_kdist(kd, p)
    ref, keep = ctables.pack_kdist(kd)
    rc = d.nc_kdist_equals(p.encode(), C.byref(ref))
    assert rc == 0, (rc, d.nc_last_error())
    p2 = str(tmp_path / f"cloud_coefficients_{kind}.nc")
    ncfiles.write_cloud_lut(lut, p2)
    cref, keep2 = ctables.pack_cloud(lut)
    rc = d.nc_cloud_equals(p2.encode(), C.byref(cref))
    assert rc == 0, (rc, d.nc_last_error())
    # a longwave file is not a cloud file: loud error, not garbage
    assert d.nc_cloud_equals(p.encode(), C.byref(cref)) == -1 and b"nband" in d.nc_last_error()
    if kind == "sw":
        alut = synthetic_aerosol_lut()
        p3 = str(tmp_path / "aerosol_optics.nc")
        ncfiles.write_aerosol_lut(alut, p3)
        aref, keep3 = ctables.pack_aerosol(alut)
        rc = d.nc_aerosol_equals(p3.encode(), C.byref(aref))
        assert rc == 0, (rc, d.nc_last_error())
        assert d.nc_aerosol_equals(p2.encode(), C.byref(aref)) == -1 and b"band_sw" in d.nc_last_error()


@pytest.mark.skipif(not os.path.isdir("/root/reference/rrtmgp-data"), reason="reference mount absent (GPU box)")
@pytest.mark.parametrize("kind", ["lw", "sw"])
def test_cpp_loader_reads_the_real_netcdf4_cloud_lut(kind):
    """rrtmgp-data's rrtmgp-clouds-{lw,sw}.nc (netCDF-4/HDF5, deflate-compressed chunks) straight through the C++
    coefficient loader the command-line driver uses: equal to the committed fixture read by the python reader."""
    import real_data as RD
    d = _nc()
    cref, keep = ctables.pack_cloud(RD.cloud_lut(RD.fixture(), kind))
    rc = d.nc_cloud_equals(f"/root/reference/rrtmgp-data/rrtmgp-clouds-{kind}.nc".encode(), C.byref(cref))
    assert rc == 0, (rc, d.nc_last_error())


def test_netcdf_writer_is_readable_by_scipy(tmp_path):
    d = _nc()
    p = str(tmp_path / "sample.nc")
    assert d.nc_write_sample(p.encode(), 5, 3) == 0, d.nc_last_error()
    f = netcdf_file(p, "r", mmap=False)
    assert f.version_byte == 2 and f.dimensions == {"x": 5, "lev": 3, "pair": 2}
    k = np.arange(15)
    assert np.array_equal(f.variables["flux_d"][:], (0.125 * k - 3.0).reshape(3, 5))
    assert f.variables["flux_f"][:].dtype == np.dtype(">f4") and np.array_equal(f.variables["flux_f"][:], (1.5 * k).reshape(3, 5))
    assert np.array_equal(f.variables["lims"][:], (np.arange(10) * 7 - 5).reshape(5, 2))
    assert f.variables["scalar"].getValue() == 42.5 and np.all(f.variables["never_inserted"][:] == 0)
    f.close()


def test_netcdf_reader_rejects_broken_hdf5_and_garbage(tmp_path):
    d = _nc()
    ref, keep = ctables.pack_kdist(synthetic_kdist("lw"))
    p = tmp_path / "nc4.nc"
    p.write_bytes(b"\x89HDF\r\n\x1a\n" + b"\0" * 64)            # HDF5 signature, nothing behind it
    assert d.nc_kdist_equals(str(p).encode(), C.byref(ref)) == -1
    assert d.nc_last_error().startswith(b"hdf5:")
    p.write_bytes(b"not a netcdf file")
    assert d.nc_kdist_equals(str(p).encode(), C.byref(ref)) == -1 and b"not a netCDF classic" in d.nc_last_error()
    assert d.nc_kdist_equals(str(tmp_path / "absent.nc").encode(), C.byref(ref)) == -1 and b"cannot open" in d.nc_last_error()


def test_input_file_layout(tmp_path):
    """(lay, y, x) with x fastest is the class API's (ncol, nlay) column-major array."""
    atm = rce_atmosphere(6, 5)
    p = str(tmp_path / "rte_rrtmgp_input.nc")
    ncfiles.write_input(atm, p, 3, 2, tsi_ref=1360.0)
    f = netcdf_file(p, "r", mmap=False)
    assert f.variables["p_lay"].dimensions == ("lay", "y", "x") and f.variables["emis_sfc"].dimensions == ("y", "x", "band_lw")
    assert np.array_equal(f.variables["p_lay"][:].reshape(5, 6), atm.play.T)
    assert np.array_equal(f.variables["emis_sfc"][:].reshape(6, -1), atm.emis_sfc.T)
    assert np.allclose(f.variables["tsi"][:].ravel(), atm.tsi_scaling * 1360.0)
    assert f.variables["vmr_co2"].dimensions == () and f.variables["vmr_h2o"].dimensions == ("lay", "y", "x")
    f.close()


def test_cli_options_and_loud_failure(tmp_path):
    exe = _exe("f64")
    r = subprocess.run([exe, "--help"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0 and "--output-bnd-fluxes" in r.stdout and "--timings" in r.stdout
    r = subprocess.run([exe, "--frobnicate"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 1 and "EXCEPTION: frobnicate is an illegal command line option." in r.stdout
    r = subprocess.run([exe, "shortwave"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 1 and "illegal command line option" in r.stdout
    assert "--aerosol-optics" in subprocess.run([exe, "--help"], capture_output=True, text=True, cwd=tmp_path).stdout
    r = subprocess.run([exe], capture_output=True, text=True, cwd=tmp_path)        # no input file
    assert r.returncode == 1 and "cannot open rte_rrtmgp_input.nc" in r.stdout
    if not torch.cuda.is_available():
        # with inputs but without a GPU the driver must stop with an exception and write nothing
        ncfiles.write_input(rce_atmosphere(4, 8), str(tmp_path / "rte_rrtmgp_input.nc"), 2, 2)
        ncfiles.write_kdist(synthetic_kdist("lw"), str(tmp_path / "coefficients_lw.nc"))
        r = subprocess.run([exe, "--no-shortwave"], capture_output=True, text=True, cwd=tmp_path)
        assert r.returncode == 1 and "EXCEPTION" in r.stdout
        assert not os.path.exists(tmp_path / "rte_rrtmgp_output.nc")


def _stage(tmp_path, atm, nx, ny, kind="f8", tsi_ref=None):
    kl, ks = synthetic_kdist("lw"), synthetic_kdist("sw")
    ll, ls = synthetic_cloud_lut("lw"), synthetic_cloud_lut("sw")
    ncfiles.write_kdist(kl, str(tmp_path / "coefficients_lw.nc")); ncfiles.write_kdist(ks, str(tmp_path / "coefficients_sw.nc"))
    ncfiles.write_cloud_lut(ll, str(tmp_path / "cloud_coefficients_lw.nc")); ncfiles.write_cloud_lut(ls, str(tmp_path / "cloud_coefficients_sw.nc"))
    ncfiles.write_input(atm, str(tmp_path / "rte_rrtmgp_input.nc"), nx, ny, kind, tsi_ref)
    return kl, ks, ll, ls


TOL = {"f64": 1e-5, "f32": 0.05}


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_cli_allsky_band_fluxes(tag, tmp_path):
    """--cloud-optics --output-bnd-fluxes: fused path, LW + SW, output file against the oracle."""
    atm = allsky_atmosphere(24, 36)
    if tag == "f32":
        atm = as_f32_valued(atm)
    kl, ks, ll, ls = _stage(tmp_path, atm, 6, 4, "f8" if tag == "f64" else "f4")
    r = subprocess.run([_exe(tag), "--cloud-optics", "--output-bnd-fluxes"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    out = ncfiles.read_output(str(tmp_path / "rte_rrtmgp_output.nc"))
    ref_lw, ref_sw = oracle_lw(kl, atm, ll), oracle_sw(ks, atm, ls)
    for k in ("flux_up", "flux_dn", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_net"):
        assert np.abs(out["lw_" + k] - ref_lw[k]).max() <= TOL[tag], k
    for k in ("flux_up", "flux_dn", "flux_dn_dir", "flux_net", "bnd_flux_up", "bnd_flux_dn", "bnd_flux_dn_dir", "bnd_flux_net"):
        assert np.abs(out["sw_" + k] - ref_sw[k]).max() <= TOL[tag], k
    assert np.array_equal(out["p_lev"], atm.plev.astype(out["p_lev"].dtype))
    assert np.allclose(out["lw_band_lims_wvn"], kl.bnd_limits_wavenumber) and out["sw_band_lims_wvn"].shape == (ks.nbnd, 2)
    assert "lw_tau" not in out and "toa_source" not in out


@pytest.mark.gpu
def test_cli_output_optical_and_tsi(tmp_path):
    """--output-optical: staged class path; tau/ssa/g/sources/toa_source written with the reference's names;
    per-column `tsi` in the input becomes the tsi scaling (src_test/test_rte_rrtmgp.cu:561-568)."""
    atm = rce_atmosphere(10, 24)
    kl, ks, ll, ls = _stage(tmp_path, atm, 5, 2)
    # the reference divides tsi by the table's own tsi: write tsi = scaling * tsi(table) using the oracle's value
    ref_sw = oracle_sw(ks, atm, None)
    ncfiles.write_input(atm, str(tmp_path / "rte_rrtmgp_input.nc"), 5, 2, tsi_ref=ref_sw["go"].get_tsi())
    r = subprocess.run([_exe("f64"), "--output-optical", "--timings"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("Duration longwave solver") == 11 and r.stdout.count("Duration shortwave solver") == 11
    out = ncfiles.read_output(str(tmp_path / "rte_rrtmgp_output.nc"))
    ref_lw = oracle_lw(kl, atm, None)
    st = ref_lw["state"]
    for a, ko in ((ref_lw["tau"], "lw_tau"), (st["lay_src"], "lay_source"), (st["lev_src"], "lev_source"), (st["sfc_src"], "sfc_source")):
        assert out[ko].shape == a.shape and np.allclose(out[ko], a, rtol=1e-10, atol=1e-12), ko
    for a, ko in ((ref_sw["tau"], "sw_tau"), (ref_sw["ssa"], "ssa"), (ref_sw["g"], "g"), (ref_sw["toa"], "toa_source")):
        assert out[ko].shape == a.shape and np.allclose(out[ko], a, rtol=1e-10, atol=1e-12), ko
    assert np.array_equal(out["lw_band_lims_gpt"], kl.bnd_limits_gpt) and np.array_equal(out["sw_band_lims_gpt"], ks.bnd_limits_gpt)
    for k in ("flux_up", "flux_dn", "flux_net"):
        assert np.abs(out["lw_" + k] - ref_lw[k]).max() <= 1e-5 and np.abs(out["sw_" + k] - ref_sw[k]).max() <= 1e-5, k


@pytest.mark.gpu
def test_cli_switches(tmp_path):
    """--no-longwave / --no-fluxes leave the corresponding variables out, as the reference does."""
    atm = rce_atmosphere(4, 16)
    _stage(tmp_path, atm, 2, 2)
    r = subprocess.run([_exe("f64"), "--no-longwave"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stdout
    out = ncfiles.read_output(str(tmp_path / "rte_rrtmgp_output.nc"))
    assert "sw_flux_up" in out and "lw_flux_up" not in out and "sw_bnd_flux_up" not in out
    r = subprocess.run([_exe("f64"), "--no-shortwave", "--no-fluxes", "--output-optical"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stdout
    out = ncfiles.read_output(str(tmp_path / "rte_rrtmgp_output.nc"))
    assert "lw_tau" in out and "lw_flux_up" not in out and "sw_tau" not in out


@pytest.mark.gpu
@pytest.mark.parametrize("optical", [False, True])
def test_cli_aerosol_optics(optical, tmp_path):
    """--aerosol-optics [--delta-aerosol] with aerosol_optics.nc, rh and aermr01..11 in the input
    (src_test/test_rte_rrtmgp.cpp:274-293): fused path, and with --output-optical the staged class path."""
    atm = add_aerosols(rce_atmosphere(12, 20, cloudy=True))
    atm.aermr["aermr10"] = np.broadcast_to(atm.aermr["aermr10"], (1, atm.nlay)).copy()   # the file contract has no scalar form
    kl, ks, ll, ls = _stage(tmp_path, atm, 4, 3)
    alut = synthetic_aerosol_lut()
    ncfiles.write_aerosol_lut(alut, str(tmp_path / "aerosol_optics.nc"))
    flags = ["--no-longwave", "--cloud-optics", "--aerosol-optics", "--delta-aerosol"] + (["--output-optical"] if optical else [])
    r = subprocess.run([_exe("f64")] + flags, capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    out = ncfiles.read_output(str(tmp_path / "rte_rrtmgp_output.nc"))
    ref = oracle_sw(ks, atm, ls, aerosol_lut=alut, delta_aerosol=True)
    for k in ("flux_up", "flux_dn", "flux_dn_dir", "flux_net"):
        assert np.abs(out["sw_" + k] - ref[k]).max() <= 1e-5, k
    if optical:
        for a, ko in ((ref["tau"], "sw_tau"), (ref["ssa"], "ssa"), (ref["g"], "g")):
            assert np.allclose(out[ko], a, rtol=1e-10, atol=1e-12), ko
    # a missing species is the reference's error
    os.remove(tmp_path / "rte_rrtmgp_output.nc")
    atm.aermr.pop("aermr05")
    ncfiles.write_input(atm, str(tmp_path / "rte_rrtmgp_input.nc"), 4, 3)
    r = subprocess.run([_exe("f64")] + flags, capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 1 and 'Aerosol type "aermr05" not available in input file.' in r.stdout
