"""GPU ensemble parity on the configurations that are benchmarked (BASELINE.json configs[1], [2], [3]), against committed
ensembles of the reference binary (oracle/_ref/smcsmc_fixed, see tests/test_em_parity.py; generators in
tests/golden/make_golden.py):

  configs[1] in medium  8 haplotypes, first 10 Mb of the zigzag bench data, the 11 zigzag epochs from all-ones sizes,
                        1000 particles, 5 EM iterations on sizes and recombination rate, 10 seeds
  configs[2] in medium  4 + 4 haplotypes with split and directional migration, first 5 Mb of the structured bench data,
                        2000 particles, 2 EM iterations on sizes, migration and recombination rates, 10 seeds
  configs[3], one point 4 haplotypes x 10 Mb, ONE sweep with 10^5 particles against the 10^3-particle reference ensemble

RNG streams differ by design: seed ensembles are compared row by row of the .out file (Welch z, relative difference of
the means), the template being the reference's own test/old/newtests/test_generic.py:313-398."""
import gzip
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

from test_em_parity import parse_out

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "smcsmc_b200", "bin", "smcsmc")
G = os.path.join(ROOT, "tests", "golden")


def cut_seg(gz_path, out_path, length):
    with gzip.open(gz_path, "rt") as f, open(out_path, "w") as g:
        for line in f:
            if not line.startswith("#") and int(line.split("\t", 1)[0]) > length:
                break
            g.write(line)


def _run_all(ref, seg, tmp, extra=()):
    procs = []
    for seed in ref["seeds"]:          # the runs share the GPU (time-sliced); launch them together
        out = os.path.join(tmp, "s%d" % seed)
        cmd = [BIN] + ref["base"].format(seed=seed, seg=seg, out=out).split() + list(extra)
        procs.append((out, subprocess.Popen(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)))
    mine = []
    for out, p in procs:
        err = p.communicate()[1]
        assert p.returncode == 0, err[-2000:]
        mine.append(parse_out(out + ".out")[1])
    return mine


def _ensemble(ref_runs, mine, cols=("Opp", "Count")):
    """(row index, column) -> (reference mean, our mean, Welch z, relative difference)"""
    rep = {}
    for k, row in enumerate(ref_runs[0]):
        for col in cols:
            a = np.array([float(r[k][col]) for r in ref_runs])
            b = np.array([float(r[k][col]) for r in mine])
            se = np.sqrt(a.var(ddof=1) / len(a) + b.var(ddof=1) / len(b)) + 1e-300
            rep[(k, col)] = (a.mean(), b.mean(), (b.mean() - a.mean()) / se, b.mean() / a.mean() - 1 if a.mean() else 0.0)
    return rep


def _report(ref_runs, rep):
    return "\n".join("iter %s epoch %2s %-6s %2s->%2s %-5s ref %.6g ours %.6g z %+.2f rel %+.3f" %
                     (ref_runs[0][k]["Iter"], ref_runs[0][k]["Epoch"], ref_runs[0][k]["Type"], ref_runs[0][k]["From"],
                      ref_runs[0][k]["To"], col, ma, mb, z, rel) for (k, col), (ma, mb, z, rel) in rep.items())


def _check(ref_runs, rep, msg, tol, tol_recent, small=30.0):
    for (k, col), (ma, mb, z, rel) in rep.items():
        r = ref_runs[0][k]
        typ = r["Type"]
        if typ == "Delay" or (typ in ("Resamp", "LogL") and col == "Opp"):
            continue
        assert abs(z) < 4.5, msg
        # the relative bound applies where the two ensembles resolve it; a handful of events per run is the z test's job
        if col == "Count" and abs(ma) < small:
            continue
        bound = tol_recent if (typ in ("Coal", "Migr") and r["Epoch"] in ("0", "1", "2")) else tol
        resolved = abs(z) < 1e-12 or 3 * abs(rel / z) < bound
        if resolved:
            assert abs(rel) < bound, msg


def test_configs1_medium_zigzag(tmp_path):
    """every Coal / Recomb row of all six sweeps within Monte-Carlo error of the reference ensemble, and the final size
    estimates of the two programs within 10 % of each other per epoch, with the same first trace of the simulated
    alternation (10 000 / 2 000)"""
    ref = json.load(open(os.path.join(G, "configs1_medium_reference.json")))
    seg = str(tmp_path / "c1.seg")
    cut_seg(os.path.join(ROOT, ref["data"]), seg, ref["length"])
    mine = _run_all(ref, seg, str(tmp_path))
    rep = _ensemble(ref["runs"], mine)
    msg = _report(ref["runs"], rep)
    print(msg)
    # Each program calibrates its own lags: the device's are 15-35 % shorter in epoch 0 and in the deepest epochs
    # (DESIGN.md 7a), which moves those epochs' tallies by a few per cent -- systematically, so the z test is not applied
    # to them here (it is in test_configs1_first_sweep_with_the_reference_lags, where both use the same lags)
    cal = {k for k in rep if ref["runs"][0][k[0]]["Type"] == "Coal" and ref["runs"][0][k[0]]["Epoch"] in ("0", "8", "9", "10")}
    for k in cal:
        ma, mb, z, rel = rep[k]
        row = ref["runs"][0][k[0]]
        # first sweep: a few per cent; later sweeps: the EM feeds the difference back into the sizes of the deepest epoch
        # (~30 coalescences per sweep there), whose tallies then differ by up to 30 % while the size estimate itself --
        # opportunity / count, compared below -- stays within 10 %
        bound = 0.30 if row["Epoch"] == "0" else (0.15 if row["Iter"] == "0" else 0.35)
        assert abs(rel) < bound or abs(ma) < 30, msg
    _check(ref["runs"], {k: v for k, v in rep.items() if k not in cal}, msg, tol=0.10, tol_recent=0.25)
    # the zigzag: relative sizes 1, .2, 1, .2, ... per epoch.  Final-iteration Ne per epoch, both programs
    last = max(int(r["Iter"]) for r in ref["runs"][0])
    idx = [k for k, r in enumerate(ref["runs"][0]) if r["Type"] == "Coal" and int(r["Iter"]) == last]
    assert len(idx) == 11
    ne_ref = np.array([[float(run[k]["Ne"]) for k in idx] for run in ref["runs"]]).mean(axis=0)
    ne_mine = np.array([[float(run[k]["Ne"]) for k in idx] for run in mine]).mean(axis=0)
    print("final Ne  ref ", np.round(ne_ref), "\nfinal Ne  ours", np.round(ne_mine))
    for ne in (ne_ref, ne_mine):
        # Five EM iterations from all-ones sizes on 10 Mb do not resolve the 10 000 / 2 000 alternation -- the sizes of
        # BOTH programs are still drifting down from the start value towards its harmonic mean (3 300) -- but the
        # strongest bottleneck the data see (epoch 7, 6 400 - 12 800 generations) already shows as a dip below both
        # neighbours in both, and every epoch with data sits between the two simulated sizes
        assert ne[7] < ne[6] and ne[7] < ne[8], ne
        assert np.all(ne[4:] > 2000) and np.all(ne[4:] < 10000), ne
    for e in range(4, 11):
        assert ne_mine[e] == pytest.approx(ne_ref[e], rel=0.10), (e, ne_mine[e], ne_ref[e])
    # the recombination rate of the last M-step, both near the simulated 1e-8
    kr = [k for k, r in enumerate(ref["runs"][0]) if r["Type"] == "Recomb" and int(r["Iter"]) == last][0]
    rates = [np.mean([float(r[kr]["Rate"]) for r in runs]) for runs in (ref["runs"], mine)]
    assert rates[1] == pytest.approx(rates[0], rel=0.05), rates
    for x in rates:       # both programs sit ~16 % under the simulated 1e-8 after five iterations on this window
        assert x == pytest.approx(1e-8, rel=0.25), rates


def test_configs1_first_sweep_with_the_reference_lags(tmp_path):
    """The first sweep (all-ones sizes) run with the lags the reference itself calibrated for it (-lags, an extension of
    the drop-in driver): with the schedule of the lagged tallies equal in both programs every Coal / Recomb row agrees
    within Monte-Carlo error, the two deepest epochs included -- the few per cent by which they differ under each
    program's own calibration (test_configs1_medium_zigzag) are the calibration's, not the engine's."""
    ref = json.load(open(os.path.join(G, "configs1_medium_reference.json")))
    seg = str(tmp_path / "c1.seg")
    cut_seg(os.path.join(ROOT, ref["data"]), seg, ref["length"])
    one = dict(ref, base=ref["base"].replace("-EM 5", "-EM 0"))
    mine = _run_all(one, seg, str(tmp_path), extra=["-lags", ",".join(repr(x) for x in ref["lags_iter0"])])
    first = [[r for r in run if r["Iter"] == "0"] for run in ref["runs"]]
    rep = _ensemble(first, mine)
    msg = _report(first, rep)
    print(msg)
    _check(first, rep, msg, tol=0.06, tol_recent=0.25)


def test_configs2_medium_split_migration(tmp_path):
    """per-population Coal rows, Migr rows (cross-coalescence is read off these) and the Recomb row of all three sweeps"""
    ref = json.load(open(os.path.join(G, "configs2_medium_reference.json")))
    seg = str(tmp_path / "c2.seg")
    cut_seg(os.path.join(ROOT, ref["data"]), seg, ref["length"])
    mine = _run_all(ref, seg, str(tmp_path))
    rep = _ensemble(ref["runs"], mine)
    msg = _report(ref["runs"], rep)
    print(msg)
    _check(ref["runs"], rep, msg, tol=0.12, tol_recent=0.25)
    last = max(int(r["Iter"]) for r in ref["runs"][0])
    idx = {(r["Type"], r["Epoch"], r["From"], r["To"]): k for k, r in enumerate(ref["runs"][0]) if int(r["Iter"]) == last}

    def mean(runs, key, col):
        return np.mean([float(run[idx[key]][col]) for run in runs])
    for runs in (ref["runs"], mine):
        assert 1.0e-5 < mean(runs, ("Migr", "1", "1", "0"), "Rate") < 5.0e-5        # simulated 2.5e-5, 2 -> 1 only
        assert mean(runs, ("Migr", "1", "0", "1"), "Rate") == 0.0
        for pop in ("0", "1"):
            assert 7000 < mean(runs, ("Coal", "1", pop, "-1"), "Ne") < 14000


def test_configs3_point_100k_particles(tmp_path):
    """one sweep of configs[0]'s data with 10^5 particles: every tally within the spread the 10^3-particle reference
    ensemble allows for its own mean (a larger population estimates the same posterior expectations with less noise)"""
    ref = json.load(open(os.path.join(G, "configs0_reference.json")))
    seg = str(tmp_path / "c0.seg")
    with gzip.open(os.path.join(ROOT, "tests", "data", "const_n4_10Mb_seed11.seg.gz"), "rb") as f, open(seg, "wb") as g:
        shutil.copyfileobj(f, g)
    out = str(tmp_path / "big")
    cmd = [BIN] + ref["base"].format(seed=1, seg=seg, out=out).replace("-Np 1000", "-Np 100000").replace("-EM 3", "-EM 0").split()
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    rows = parse_out(out + ".out")[1]
    lines = []
    for k, row in enumerate(ref["runs"][0]):
        if row["Iter"] != "0" or row["Type"] not in ("Coal", "Recomb"):      # (the ln-likelihood estimate is biased in 1/N)
            continue
        for col in ("Opp", "Count"):
            a = np.array([float(run[k][col]) for run in ref["runs"]])
            b = float(rows[k][col])
            sd = a.std(ddof=1)
            z = (b - a.mean()) / (sd * np.sqrt(1 / len(a) + 1 / 100.0) + 1e-300)
            lines.append((row["Epoch"], row["Type"], col, a.mean(), b, z))
    msg = "\n".join("epoch %2s %-6s %-5s ref mean %.6g ours(1e5) %.6g z %+.2f" % x for x in lines)
    print(msg)
    # a particle filter's posterior expectations carry a bias of order 1/N, so the two population sizes differ by more
    # than their seed-to-seed noise (a few per cent at 10^3 particles); the bound is relative, the z column is printed
    for ep, typ, col, ma, b, z in lines:
        if abs(z) < 3.0:
            continue                                   # inside the reference ensemble's own spread
        if abs(ma) > 100:
            assert b == pytest.approx(ma, rel=0.04), msg
        else:
            assert abs(z) < 6.0 and abs(b - ma) < 5.0, msg     # a dozen events per sweep
