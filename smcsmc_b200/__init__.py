"""smcsmc_b200 -- B200-native particle-filter engine for SMCSMC (luntergroup/smcsmc), behind a C ABI.

The product is `lib/libsmcgpu.so` (hand-written sm_100a CUDA, include/smcsmc_b200.h) and the C++ `smcsmc` driver in
`host/`.  This Python package is a thin ctypes binding used by the tests, bench.py and the smoke check; it contains no
compute and has NO fallback: loading fails loudly when the CUDA library has not been built.
"""
from ._abi import lib, lib_path, SmcError, SmcConfig  # noqa: F401
from .engine import (ParticleFilter, shard_unique_id, sweep_many, site_likelihood, branch_lengths, systematic_offspring, resample_counts,  # noqa: F401
                     normalize_log, fastexp, calibrate_survival, terminal_branch_quantiles, measure_fp64_fma)
from .segdata import read_seg, SegRows  # noqa: F401

__version__ = "0.1"
