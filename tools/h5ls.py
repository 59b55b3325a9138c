"""List the variables of a netCDF-4/HDF5 or netCDF classic file the way `ncdump -h` would (name, dimensions, type),
through the repo's own readers (no libnetcdf/libhdf5 needed):   python tools/h5ls.py file.nc [variable]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rte_rrtmgp_cpp_b200 import h5lite, ncfiles          # noqa: E402


def main(path, var=None):
    r = ncfiles._Reader(path)
    if r.h5 is not None:
        f = r.h5
        for n in f.names():
            if f.is_pure_dimension(n) or (var and n != var):
                continue
            try:
                dims = ", ".join(f"{d}={s}" for d, s in zip(f.dims(n), f.shape(n)))
            except h5lite.H5Error:
                dims = ", ".join(str(s) for s in f.shape(n))
            a = f.attrs(n)
            units = f" [{a['units']}]" if isinstance(a.get("units"), str) else ""
            print(f"{n}({dims}){units}")
            if var:
                with np.printoptions(threshold=40, linewidth=140):
                    print(f.get(n))
    else:
        for n, v in r.nc.variables.items():
            if var and n != var:
                continue
            print(f"{n}({', '.join(f'{d}={s}' for d, s in zip(v.dimensions, v.shape))})")
            if var:
                with np.printoptions(threshold=40, linewidth=140):
                    print(r.get(n))
    r.close()


if __name__ == "__main__":
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    main(*sys.argv[1:3])
