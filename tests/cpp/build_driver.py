"""Build the C++ class-layer test drivers (tests/cpp/classes_driver.cpp) against the in-tree
librrtmgp_b200.so: one with Float=double, one with -DRTE_USE_SP.  g++ only (the header has no
device code); links cudart for the Array_gpu allocations."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
PKG = os.path.join(ROOT, "rte_rrtmgp_cpp_b200")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def path(tag: str) -> str:
    return os.path.join(HERE, f"libclasses_driver_{tag}.so")


def ctor_path(tag: str) -> str:
    return os.path.join(HERE, f"libctor_driver_{tag}.so")


def build(force: bool = False) -> None:
    _build("classes_driver.cpp", path, force)
    _build("ctor_driver.cpp", ctor_path, force)


def _build(source: str, path, force: bool) -> None:
    src = os.path.join(HERE, source)
    hdrs = [os.path.join(ROOT, "include", "rrtmgp_b200", "classes.hpp"), os.path.join(ROOT, "include", "rrtmgp_b200", "array.hpp"),
            os.path.join(ROOT, "include", "rrtmgp_b200.h"), src]
    newest = max(os.path.getmtime(h) for h in hdrs)
    for tag, defs in (("f64", []), ("f32", ["-DRTE_USE_SP"])):
        out = path(tag)
        if not force and os.path.exists(out) and os.path.getmtime(out) >= newest:
            continue
        cmd = ["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-Wall", *defs, "-I" + os.path.join(ROOT, "include"),
               "-I" + os.path.join(CUDA, "include"), src, "-o", out, "-L" + PKG, "-lrrtmgp_b200",
               "-L" + os.path.join(CUDA, "lib64"), "-lcudart", "-Wl,-rpath,$ORIGIN/../../rte_rrtmgp_cpp_b200"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ failed for the class-layer driver:\n" + r.stdout + r.stderr)


def nc_path() -> str:
    return os.path.join(HERE, "libnc_driver.so")


def build_nc(force: bool = False) -> None:
    """Host-only checker of include/rrtmgp_b200/{netcdf_classic,coefficients}.hpp (no CUDA, no library)."""
    src = os.path.join(HERE, "nc_driver.cpp")
    inc = os.path.join(ROOT, "include", "rrtmgp_b200")
    newest = max(os.path.getmtime(f) for f in (src, os.path.join(inc, "netcdf_classic.hpp"), os.path.join(inc, "hdf5_lite.hpp"),
                                               os.path.join(inc, "coefficients.hpp"),
                                               os.path.join(ROOT, "include", "rrtmgp_b200_tables.h")))
    out = nc_path()
    if not force and os.path.exists(out) and os.path.getmtime(out) >= newest:
        return
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-fPIC", "-shared", "-Wall", "-I" + os.path.join(ROOT, "include"), src, "-o", out, "-lz"],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed for the netCDF driver:\n" + r.stdout + r.stderr)


if __name__ == "__main__":
    build_nc(force=True)
    build(force=True)
