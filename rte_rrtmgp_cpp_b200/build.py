"""Build librrtmgp_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.environ.get("RB200_LIB", os.path.join(HERE, "librrtmgp_b200.so"))
OBJ = os.path.join(HERE, "csrc", "_obj" + os.environ.get("RB200_TAG", ""))
EXTRA = os.environ.get("RB200_DEFS", "").split()
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
# k1_interp.cu carries the index arithmetic: no FMA contraction there (bit-exact indices)
SOURCES = {"cabi.cu": [], "k1_interp.cu": ["--fmad=false"], "fused_lw.cu": [], "fused_lw_ma.cu": [], "fused_sw.cu": [], "fused_pair.cu": [], "colres.cu": [],
           "stages.cu": [], "gasopt.cu": [], "arrays.cu": []}


def _newest_dep() -> float:
    return max(os.path.getmtime(os.path.join(CSRC, f)) for f in os.listdir(CSRC)
               if f.endswith((".cu", ".cuh", ".hpp"))) if os.path.isdir(CSRC) else 0.0


def build(force: bool = False, verbose: bool = False) -> str:
    inc = os.path.join(os.path.dirname(HERE), "include")
    dep = max(_newest_dep(), max(os.path.getmtime(os.path.join(inc, f)) for f in os.listdir(inc)))
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= dep:
        return LIB
    os.makedirs(OBJ, exist_ok=True)

    def cc(item):
        src, extra = item
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = ["nvcc", *ARCH, *COMMON, *EXTRA, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=6) as ex:
        objs = list(ex.map(cc, SOURCES.items()))
    r = subprocess.run(["nvcc", *ARCH, "-shared", "-o", LIB, *objs, "-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
