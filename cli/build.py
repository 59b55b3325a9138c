"""Build the command-line driver (cli/test_rte_rrtmgp.cpp) against the in-tree librrtmgp_b200.so:
`test_rte_rrtmgp_b200` (Float = double) and `test_rte_rrtmgp_b200_sp` (-DRTE_USE_SP), the two
precisions the reference builds with USESP on/off.  g++ only: the driver holds no device code."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = os.path.join(ROOT, "rte_rrtmgp_cpp_b200")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def path(tag: str) -> str:
    return os.path.join(HERE, "test_rte_rrtmgp_b200" + ("_sp" if tag == "f32" else ""))


def build(force: bool = False) -> None:
    src = os.path.join(HERE, "test_rte_rrtmgp.cpp")
    inc = os.path.join(ROOT, "include")
    deps = [src, os.path.join(inc, "rrtmgp_b200.h"), os.path.join(inc, "rrtmgp_b200_tables.h")]
    deps += [os.path.join(inc, "rrtmgp_b200", f) for f in os.listdir(os.path.join(inc, "rrtmgp_b200"))]
    newest = max(os.path.getmtime(d) for d in deps)
    for tag, defs in (("f64", []), ("f32", ["-DRTE_USE_SP"])):
        out = path(tag)
        if not force and os.path.exists(out) and os.path.getmtime(out) >= newest:
            continue
        cmd = ["g++", "-std=c++17", "-O2", "-Wall", *defs, "-I" + inc, "-I" + os.path.join(CUDA, "include"), src, "-o", out,
               "-L" + PKG, "-lrrtmgp_b200", "-L" + os.path.join(CUDA, "lib64"), "-lcudart", "-lz",
               "-Wl,-rpath,$ORIGIN/../rte_rrtmgp_cpp_b200"]      # zlib: deflate chunks of netCDF-4 files (hdf5_lite.hpp)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("g++ failed for the command-line driver:\n" + r.stdout + r.stderr)


if __name__ == "__main__":
    build(force=True)
