"""GPU end-to-end test of the drop-in `smcsmc` driver under a STRUCTURED model against the reference binary
(oracle/_ref/smcsmc_fixed: the reference with its uninitialised ForestState::max_epoch_to_record_ initialised, see
tests/test_em_parity.py and DESIGN.md 7a):
BASELINE.json configs[2] in small (two populations of 4 haplotypes, migration 2 -> 1 backwards in time, joined at
0.5 * 4N0; 2 Mb simulated by the reference's scrm under that model; inference epochs {0, 0.1, 0.5}; 500 particles),
same command lines, 12 seeds each.  RNG streams differ by design, so seed ensembles are compared (Welch z, relative
difference of the means): per-population coalescent opportunity and counts, migration opportunity and counts
("cross-coalescence" is read off these rows), recombination opportunity and count, resampling events and the
log-likelihood -- for one sweep, for three in-binary EM iterations on population sizes and migration rates
(CountModel::reset_Ne / reset_mig_rate, count.cpp:267-352) and for three on sizes, migration and recombination rate.  Reference ensemble: tests/golden/em_struct_reference.json
(tests/golden/make_golden.py::golden_em_struct, oracle/_ref/smcsmc)."""
import json
import os
import subprocess

import numpy as np
import pytest

from test_em_parity import parse_out

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "smcsmc_b200", "bin", "smcsmc")
SEG = os.path.join(ROOT, "tests", "data", "split_mig_n8_2Mb_seed13.seg")
REF = os.path.join(ROOT, "tests", "golden", "em_struct_reference.json")


@pytest.fixture(scope="module")
def runs(tmp_path_factory):
    ref = json.load(open(REF))
    tmp = tmp_path_factory.mktemp("ems")
    mine = {}
    for name, extra in ref["variants"].items():
        mine[name] = []
        for seed in ref["seeds"]:
            out = str(tmp / ("%s_s%d" % (name, seed)))
            cmd = [BIN] + ref["base"].format(extra=extra, seed=seed, seg=SEG, out=out).split()
            r = subprocess.run(cmd, capture_output=True, text=True)
            assert r.returncode == 0, r.stderr[-2000:]
            mine[name].append((out, r))
    return ref, mine


def _ensemble(ref_runs, mine_rows):
    """per .out row: (row, column) -> (reference mean, our mean, z, relative difference)"""
    rep = {}
    for k, row in enumerate(ref_runs[0]):
        for col in ("Opp", "Count"):
            a = np.array([float(r[k][col]) for r in ref_runs])
            b = np.array([float(r[k][col]) for r in mine_rows])
            se = np.sqrt(a.var(ddof=1) / len(a) + b.var(ddof=1) / len(b)) + 1e-300
            rep[(k, col)] = (a.mean(), b.mean(), (b.mean() - a.mean()) / se, b.mean() / a.mean() - 1 if a.mean() else 0.0)
    return rep


def _report(ref_runs, rep):
    lines = []
    for (k, col), (ma, mb, z, rel) in rep.items():
        r = ref_runs[0][k]
        lines.append("iter %s epoch %2s %-6s %2s->%2s %-5s ref %.6g ours %.6g z %+.2f rel %+.3f" %
                     (r["Iter"], r["Epoch"], r["Type"], r["From"], r["To"], col, ma, mb, z, rel))
    return "\n".join(lines)


def test_layout_matches_reference(runs):
    """same rows in the same order: Coal per epoch x population, one Recomb row, Migr per epoch x ordered pair, Delay,
    Resamp, LogL (CountModel::log_counts, count.cpp:66-158), identical Iter/Epoch/Start/End/Type/From/To columns"""
    ref, mine = runs
    for name in ref["variants"]:
        rows = parse_out(mine[name][0][0] + ".out")[1]
        ref_rows = ref["runs"][name][0]
        assert len(rows) == len(ref_rows)
        for a, b in zip(rows, ref_rows):
            for k in ("Iter", "Epoch", "Start", "End", "Type", "From", "To"):
                assert a[k] == b[k], (k, a, b)
    # cells that hold nothing but the pseudo-count are identical in both programs: population 2 after the join, and
    # migration 1 -> 2, which the model does not have
    rows = parse_out(mine["sweep"][0][0] + ".out")[1]
    for a, b in zip(rows, ref["runs"]["sweep"][0]):
        if (a["Type"], a["Epoch"], a["From"]) in (("Coal", "2", "1"), ("Migr", "2", "1")):
            assert (a["Opp"], a["Count"], a["Rate"]) == (b["Opp"], b["Count"], b["Rate"]), (a, b)
        if a["Type"] == "Migr" and a["From"] == "0":
            assert float(a["Count"]) == 0.0 and float(b["Count"]) == 0.0


def test_structured_post_lag_ess_column(runs):
    """the ESS column with populations (every cell has a weight-squared twin, count.cpp:519-526): pooled per ancestor
    slot as in the single-population engine; ensemble means within [0.7, 1.8] of the reference's on every Coal, Recomb
    and Migr row that holds more than the pseudo-count"""
    ref, mine = runs
    ref_runs = ref["runs"]["sweep"]
    mine_rows = [parse_out(out + ".out")[1] for out, _ in mine["sweep"]]
    checked = 0
    for k, row in enumerate(ref_runs[0]):
        if row["Type"] not in ("Coal", "Recomb", "Migr"):
            continue
        a = np.mean([float(r[k]["ESS"]) for r in mine_rows])
        b = np.mean([float(r[k]["ESS"]) for r in ref_runs])
        print(row["Type"], row["Epoch"], row["From"], row["To"], "ESS ours %.2f ref %.2f ratio %.2f" % (a, b, a / b))
        assert 0.7 < a / b < 1.8, (row, a, b)
        checked += 1
    assert checked == 13


def test_structured_sweep_within_monte_carlo_error(runs):
    ref, mine = runs
    ref_runs = ref["runs"]["sweep"]
    mine_rows = [parse_out(out + ".out")[1] for out, _ in mine["sweep"]]
    rep = _ensemble(ref_runs, mine_rows)
    msg = _report(ref_runs, rep)
    print(msg)
    for (k, col), (ma, mb, z, rel) in rep.items():
        r = ref_runs[0][k]
        typ = r["Type"]
        if typ == "Delay" or (typ in ("Resamp", "LogL") and col == "Opp"):
            continue
        assert abs(z) < 4.5, msg
        small = col == "Count" and abs(ma) < 30          # a handful of events per run: the z test is the test
        if not small:
            assert abs(rel) < (0.25 if r["Epoch"] == "0" else 0.12), msg


@pytest.mark.parametrize("variant", ["em", "em_all"])
def test_structured_em_within_monte_carlo_error(runs, variant):
    """three EM iterations, M-step on N_e per population and on the migration rates ("em"; recombination excluded from
    the EM with -xr, so its rows hold pseudo-counts only) and on the recombination rate as well ("em_all")"""
    ref, mine = runs
    ref_runs = ref["runs"][variant]
    mine_rows = [parse_out(out + ".out")[1] for out, _ in mine[variant]]
    rep = _ensemble(ref_runs, mine_rows)
    msg = _report(ref_runs, rep)
    print(msg)
    for (k, col), (ma, mb, z, rel) in rep.items():
        r = ref_runs[0][k]
        typ = r["Type"]
        if typ == "Delay" or (typ == "Recomb" and variant == "em") or (typ in ("Resamp", "LogL") and col == "Opp"):
            continue
        assert abs(z) < 4.5, msg
        small = col == "Count" and abs(ma) < 30
        if not small:
            assert abs(rel) < (0.30 if r["Epoch"] == "0" else 0.15), msg
    # the estimates of the last iteration sit near the simulated truth: N_e = 10 000 in both populations before the
    # join, migration 2 -> 1 at M = 1 (2.5e-5 per generation), none the other way
    last = [r for r in mine_rows[0] if r["Iter"] == "2"]
    idx = {(r["Type"], r["Epoch"], r["From"], r["To"]): i for i, r in enumerate(mine_rows[0]) if r["Iter"] == "2"}
    assert len(last) == len(ref_runs[0]) // 3

    def mean(key, col):
        return np.mean([float(rows[idx[key]][col]) for rows in mine_rows])
    for ep in ("1",):
        for pop in ("0", "1"):
            assert 7000 < mean(("Coal", ep, pop, "-1"), "Ne") < 14000
    assert 1.0e-5 < mean(("Migr", "1", "1", "0"), "Rate") < 5.0e-5
    assert mean(("Migr", "1", "0", "1"), "Rate") == 0.0


def test_mark_pool_grows_on_overflow(tmp_path):
    """migration marks live in a fixed pool per particle; when a particle needs more the engine raises a flag and the
    driver repeats the sweep with twice the room -- the result is the one of a run that had enough room from the start"""
    ref = json.load(open(REF))
    outs = {}
    for name, env in (("tiny", {"SMCSMC_MARKS_BASE": "1"}), ("ample", {"SMCSMC_MAX_MARKS": "60"})):
        out = str(tmp_path / name)
        cmd = [BIN] + ref["base"].format(extra="-EM 0 -lag 20000", seed=3, seg=SEG, out=out).split()
        r = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, **env))
        assert r.returncode == 0, r.stderr[-2000:]
        outs[name] = (parse_out(out + ".out")[1], r.stderr)
    assert "repeating the sweep" in outs["tiny"][1] and "repeating the sweep" not in outs["ample"][1]
    for a, b in zip(outs["tiny"][0], outs["ample"][0]):
        assert (a["Opp"], a["Count"]) == (b["Opp"], b["Count"]), (a, b)

