#!/bin/bash
# BASELINE.json configs[3] in small: particle-count sweep at 4 haplotypes (10 Mb of tests/data, 8 epochs) through the
# drop-in driver on one GPU; wall time of the whole process (.seg parse, device lag calibration, sweep, .out) and the
# device time of the sweep as the driver reports it.
gunzip -c tests/data/const_n4_10Mb_seed11.seg.gz > /tmp/c0.seg
M="-nsam 4 -N0 10000 -t 10000 -r 4000 10000000 -eN 0 1 -eN 0.01 1 -eN 0.02 1 -eN 0.04 1 -eN 0.08 1 -eN 0.16 1 -eN 0.32 1 -eN 0.64 1 -EM 0 -seed 1 -seg /tmp/c0.seg"
for NP in 1000 10000 100000 1000000; do
  for C in 1 16; do
    if [ $NP -ge 100000 ] && [ $C -eq 16 ]; then continue; fi
    EXTRA=""; if [ $C -gt 1 ]; then EXTRA="-chunks $C"; fi
    T0=$(date +%s%N)
    smcsmc_b200/bin/smcsmc $M -Np $NP $EXTRA -o /tmp/o_${NP}_$C > /tmp/o.stdout 2> /tmp/o.stderr
    T1=$(date +%s%N)
    echo "N=$NP chunks=$C wall $(( (T1 - T0) / 1000000 )) ms;$(grep 'Device time' /tmp/o.stderr); LogL $(grep LogL /tmp/o_${NP}_$C.out | awk '{print $9}'); $(grep -o 'resampled [0-9]*' /tmp/o.stderr)"
  done
done
