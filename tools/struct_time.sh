#!/bin/bash
# time the structured sweep (configs[2] full size) through the driver with a given library
LIB=$1
T=$(mktemp -d)
python - "$T" <<'PY'
import gzip, shutil, sys
with gzip.open("bench_data/split_mig_n8_10Mb_seed3.seg.gz","rb") as f, open(sys.argv[1]+"/c2.seg","wb") as g: shutil.copyfileobj(f,g)
PY
LD_LIBRARY_PATH=$LIB ./smcsmc_b200/bin/smcsmc -nsam 8 -Np 50000 -N0 10000 -t 10000 -r 4000 10000000 -I 2 4 4 -em 0 2 1 1 -eN 0.1 1 -ej 0.5 2 1 -EM 0 -seed 1 -seg $T/c2.seg -o $T/c2 2>&1 | grep -i "device time"
