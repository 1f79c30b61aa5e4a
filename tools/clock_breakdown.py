#!/usr/bin/env python
"""Per-phase cycle breakdown of smc_row_kernel (profiling tool): run the bench workload with the -DSMC_CLOCKS build
(smcsmc_b200/lib/libsmcgpu_clk.so: thread 0 of every block adds the cycles of its phases to device counters) and print
the share of a block's life spent in each phase.

  nvcc <flags of __graft_entry__> -DSMC_CLOCKS -o smcsmc_b200/lib/libsmcgpu_clk.so smcsmc_b200/csrc/smc_engine.cu
  SMCSMC_B200_LIB=$PWD/smcsmc_b200/lib/libsmcgpu_clk.so python tools/clock_breakdown.py [--chunks 16] [--len-mb 16]"""
import argparse
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import smcsmc_b200 as S  # noqa: E402
from smcsmc_b200._abi import lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--chunks", type=int, default=16)
ap.add_argument("--len-mb", type=float, default=16.0)
ap.add_argument("--particles", type=int, default=10000)
a = ap.parse_args()
rows, t_start, inv2N, seqlen = bench.workload(a.len_mb, a.chunks)
pfs = []
for c, r in enumerate(rows):
    pf = S.ParticleFilter(bench.NSAM, a.particles, t_start, inv2N, bench.MU, bench.RHO, seqlen, bench.ZIGZAG_LAGS, seed=3, chunk=c)
    pf.upload_rows(r.x_end, r.alleles, r.is_site, r.max_epoch_to_record(bench.ZIGZAG_LAGS))
    pf.init_particles()
    pfs.append(pf)
L = lib()
out = (ctypes.c_ulonglong * 16)()
L.smc_debug_clocks(out, 1)
S.sweep_many(pfs, 1)
for pf in pfs:
    pf.finish()
L.smc_debug_clocks(out, 0)
v = list(out)
names = ["phase 0 (tallies due)", "phase 1 (load / pull)", "phase 2 total", "  first stretch", "  move rounds", "  flush + site likelihood",
         "phase 3 (store + partials)"]
blocks = v[10]
tot = v[0] + v[1] + v[2] + v[6]
print("blocks %d, cycles per block %.0f, move rounds per block %.2f, tasks per round %.1f" %
      (blocks, tot / blocks, v[8] / blocks, v[9] / max(v[8], 1)))
for k, nm in zip((0, 1, 2, 3, 4, 5, 6), names):
    print("%-32s %9.0f cycles / block  %5.1f %%" % (nm, v[k] / blocks, 100.0 * v[k] / tot))
print("cycles per move round %.0f" % (v[4] / max(v[8], 1)))
