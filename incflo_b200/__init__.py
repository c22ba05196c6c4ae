"""incflo_b200 — B200-native nodal projection (the MLMG hot path of asalmgren/incflo).

The product is libnodalproj.so (incflo_b200/csrc, C ABI in include/nodalproj.h); this package holds the
Python-side binding used by the tests and bench.py, the slab partition helper and the synthetic inputs.
"""
