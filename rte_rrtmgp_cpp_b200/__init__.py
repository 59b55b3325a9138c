"""rte_rrtmgp_cpp_b200 — B200-native (sm_100a) column radiation hot path of rte-rrtmgp-cpp.

(The repository asks for a package called ``rte-rrtmgp-cpp_b200``; Python package names cannot
contain '-', so the importable directory is ``rte_rrtmgp_cpp_b200``.)

Layout: csrc/ = hand-written CUDA kernels + the C ABI (include/rrtmgp_b200.h);
api.py = host-side mirror of the reference's class API over that ABI; kdist.py / atmos.py =
coefficient-file container, synthetic tables and seeded synthetic columns.
"""
__version__ = "0.2.0"
