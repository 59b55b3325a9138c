"""rte_rrtmgp_cpp_b200 — B200-native (sm_100a) column radiation hot path of rte-rrtmgp-cpp.

(The repository asks for a package called ``rte-rrtmgp-cpp_b200``; Python package names cannot
contain '-', so the importable directory is ``rte_rrtmgp_cpp_b200``.)
"""
__version__ = "0.1.0"
