#!/usr/bin/env python
"""Generate synthetic scrm-simulated .seg inputs (test / bench data tooling, not product code).

Runs the reference's own `scrm` (built by oracle/build_ref.sh into oracle/_ref/scrm) and converts its output to
the .seg format with the rules of the reference's smcsmc/populationmodels.py:502-577 (convert_scrm_to_seg):
positions are int(pos*L+0.5), a leading position 1, rows "start<TAB>len<TAB>T<TAB>F<TAB>1<TAB>genotype", and a
final all-missing row up to L.  Defaults N0=1e4, mu=2.5e-8, rho=1e-8 follow populationmodels.py:14-21.
"""
import argparse
import gzip
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SCRM = os.path.join(HERE, "..", "oracle", "_ref", "scrm")


def scrm_to_seg(scrm_lines, nsam, L, out, phased=True):
    data, positions = None, None
    for line in scrm_lines:
        if line.startswith("positions:"):
            positions = [float(x) for x in line.split()[1:]]
            data = []
        elif data is not None and len(data) < nsam:
            s = line.strip()
            if s:
                data.append(s)
    if data is None:  # no segregating sites at all
        data, positions = ["" for _ in range(nsam)], []
    assert len(data) == nsam, "unexpected scrm output"
    pos = [1] + [int(p * L + 0.5) for p in positions]
    row = "{}\t{}\tT\tF\t1\t{}\n"
    for i in range(len(pos) - 1):
        alleles = [seq[i] for seq in data]
        if not phased:
            for h in range(nsam // 2):
                if alleles[2 * h] != alleles[2 * h + 1]:
                    alleles[2 * h] = alleles[2 * h + 1] = "/"
        out.write(row.format(pos[i], pos[i + 1] - pos[i], "".join(alleles)))
    out.write(row.format(pos[-1], int(L + 0.5) - pos[-1], "." * nsam))


def simulate(nsam, L, demography, seed, out_path, N0=1e4, mu=2.5e-8, rho=1e-8, phased=True, scrm=SCRM):
    theta = 4 * N0 * mu * L
    R = 4 * N0 * rho * L
    cmd = [scrm, str(nsam), "1", "-N0", repr(N0), "-t", repr(theta), "-r", repr(R), str(int(L))]
    cmd += demography.split() + ["-seed", str(seed), "-p", "10", "-l", "300000"]
    res = subprocess.run(cmd, check=True, capture_output=True, text=True)
    opener = gzip.open if out_path.endswith(".gz") else open
    with opener(out_path, "wt") as f:
        scrm_to_seg(res.stdout.splitlines(), nsam, L, f, phased)
    return " ".join(cmd)


ZIGZAG_TIMES = [0, .0025, .005, .01, .02, .04, .08, .16, .32, .64, 1.28]
ZIGZAG_SIZES = [1, .2, 1, .2, 1, .2, 1, .2, 1, .2, 1]


def zigzag_demography():
    """11 epochs, relative sizes alternating 1 / 0.2."""
    return " ".join("-eN {} {}".format(t, s) for t, s in zip(ZIGZAG_TIMES, ZIGZAG_SIZES))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--nsam", type=int, default=4)
    ap.add_argument("--len", type=float, default=1e6)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--demography", default="", help="scrm demography flags, or 'zigzag'")
    ap.add_argument("--unphased", action="store_true")
    ap.add_argument("-o", "--out", required=True)
    a = ap.parse_args()
    dem = zigzag_demography() if a.demography == "zigzag" else a.demography
    print(simulate(a.nsam, a.len, dem, a.seed, a.out, phased=not a.unphased), file=sys.stderr)
