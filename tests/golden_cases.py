"""Loader of tests/golden/ref_fluxes.npz (fluxes of the reference's own compiled class layer; made by
tests/golden/make_golden.py) — shared by the CPU test (oracle vs golden) and the GPU test (CUDA vs golden)."""
import importlib.util
import os

import numpy as np

from rte_rrtmgp_cpp_b200.atmos import Atmosphere
from rte_rrtmgp_cpp_b200.kdist import synthetic_kdist, synthetic_cloud_lut, synthetic_aerosol_lut

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))


def _generator():
    """tests/golden/make_golden.py as a module (for CASES and table_hash); it imports oracle.ref lazily enough
    that the compiled reference is NOT needed to read the fixture."""
    m = importlib.util.module_from_spec(_spec)
    _spec.loader.exec_module(m)
    return m


def load():
    z = np.load(os.path.join(HERE, "golden", "ref_fluxes.npz"))
    gen = _generator()
    for kind in ("lw", "sw"):
        assert str(z[f"sha/kdist_{kind}"]) == gen.table_hash(synthetic_kdist(kind)), \
            "synthetic k-distribution changed: regenerate tests/golden (python tests/golden/make_golden.py)"
        assert str(z[f"sha/cloud_{kind}"]) == gen.table_hash(synthetic_cloud_lut(kind)), "cloud LUT changed: regenerate"
    assert str(z["sha/aerosol"]) == gen.table_hash(synthetic_aerosol_lut()), "aerosol LUT changed: regenerate"
    return z, gen.CASES


def atmosphere(z, name) -> Atmosphere:
    F = lambda a: np.asfortranarray(a)
    pick = lambda grp: {k.split("/")[2]: F(z[k]) for k in z.files if k.startswith(f"{name}/{grp}/")}
    inp = pick("in")
    return Atmosphere(vmr=pick("vmr"), aermr=(pick("aermr") or None), **inp)


def fluxes(z, name):
    return {k.split("/")[2]: z[k] for k in z.files if k.startswith(f"{name}/out/")}
