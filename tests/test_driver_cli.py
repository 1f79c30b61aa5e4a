"""CPU tests of the drop-in `smcsmc` driver's host logic (smcsmc_b200/host): command-line parsing with the reference's
messages and exit codes (pfparam.cpp:51-180, smcsmc.cpp:46-103), rejection of what the engine does not support yet,
.seg validation (segdata.cpp:55-166), and the loud failure without a GPU (no CPU fallback)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "smcsmc_b200", "bin", "smcsmc")
SEG = os.path.join(ROOT, "tests", "data", "const_n4_2Mb_seed7.seg")


@pytest.fixture(scope="module", autouse=True)
def built():
    if not os.path.exists(BIN):
        import __graft_entry__ as g
        g.build()
    assert os.path.exists(BIN)


def run(*args):
    return subprocess.run([BIN] + list(args), capture_output=True, text=True)


def test_help_and_version():
    r = run("-v")
    assert r.returncode == 0 and "smcsmc version" in r.stdout and "B200 engine" in r.stdout
    for a in ((), ("-h",), ("-help",)):
        r = run(*a)
        assert r.returncode == 0
        for flag in ("-Np", "-ESS", "-EM", "-seg", "-lag", "-calibrate_lag", "-xr", "-xc", "-cap", "-tmax", "-chunks"):
            assert flag in r.stdout, flag


@pytest.mark.parametrize("args, message", [
    (["-Np", "10", "-t", "10"], "The option -r is required"),
    (["-Np", "10", "-r", "10", "1000"], "The option -t is required"),
    (["-Np"], "Not enough parameters when parsing options: -Np"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-ESS", "1.5"], "Out of range: -ESS"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-startpos", "0"], "Out of range: -startpos"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-calibrate_lag", "-1"], "Out of range: -calibrate_lag"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-chunks", "0"], "Out of range: -chunks"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-xr", "3-1"], "Out of range: -xr/-xc"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-xr", "5"], "epochs specified in -xr/-xc options out of range"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-bogus"], "unknown/unexpected argument: -bogus"),
    (["-Np", "10", "-t", "10", "-r", "10", "1", ], "Locus length must be at least two"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-eN", "0.1", "1", "-vb"], "-vb must occur before"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-nsam", "40"], "-nsam must be between 2 and 32"),
    (["-Np", "10", "-t", "10", "-r", "10", "1000", "-N0", "5000"], "(-N0) must be set before other population-related"),
])
def test_parse_errors(args, message):
    r = run(*args)
    assert r.returncode == 1, (r.stdout, r.stderr)
    assert r.stderr.strip().splitlines()[-1].startswith("Error: "), r.stderr
    assert message in r.stderr


@pytest.mark.parametrize("args", [
    ["-eI", "0.1", "1"], ["-eG", "0", "1"],
    ["-guide", "x.gz"], ["-bias_heights", "400"], ["-apf", "1"], ["-arg"], ["-sr", "100", "2"],
])
def test_unsupported_features_are_rejected_not_ignored(args):
    r = run("-nsam", "4", "-Np", "10", "-t", "10", "-r", "10", "1000", "-seg", SEG, *args)
    assert r.returncode == 1 and "is not supported by the B200 engine yet" in r.stderr, r.stderr


def _dump(*args):
    r = run(*args, "-dump_model")
    assert r.returncode == 0, r.stderr
    d = {"rows": []}
    for line in r.stdout.splitlines():
        w = line.split()
        if w[0] == "epoch":
            i, j, k = w.index("inv2N"), w.index("mig"), w.index("single_mig")
            d["rows"].append((float(w[3]), [float(x) for x in w[i + 1:j]], [float(x) for x in w[j + 1:k]],
                                [float(x) for x in w[k + 1:]]))
        elif w[0] == "leaf_pop":
            d["leaf_pop"] = [int(x) for x in w[1:]]
        else:
            d[w[0]] = float(w[1])
    return d


def test_structured_models_parse_like_scrm():
    """-I / -en / -eM / -em / -ej: the flattened tables the driver hands to the engine against the tables of the
    reference's own parsed Model (scrm Param::parse + Model::finalize), stored with the structured fixtures
    (tests/golden/struct_moments.npz, make_golden.py::golden_struct_moments, same command lines)"""
    import numpy as np
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    G = np.load(os.path.join(ROOT, "tests", "golden", "struct_moments.npz"))
    models = {   # the flags of make_golden.py::STRUCT_MODELS
        "island2": (4, "-I 2 2 2 1.0"),
        "split2": (4, "-I 2 2 2 -ej 0.5 2 1"),
        "config3": (8, "-I 2 4 4 -em 0 2 1 1 -ej 0.5 2 1"),
        "island2_epochs": (6, "-I 2 3 3 0.5 -en 0.05 1 0.3 -en 0.05 2 2 -eM 0.2 3.0 -eN 0.6 1.5"),
        "three": (6, "-I 3 2 2 2 -em 0.1 2 1 2.0 -ej 0.3 3 2 -ej 0.8 2 1"),
        "admix3": (6, "-I 2 3 3 0.5 -es 0.1 1 0.3 -ej 0.4 3 2 -ej 0.9 2 1"),
    }
    L = 100000000
    for name, (n, flags) in models.items():
        d = _dump("-nsam", str(n), "-Np", "10", "-t", repr(4e4 * 2.5e-8 * L), "-r", repr(4e4 * 1e-8 * L), str(L),
                  "-seed", "1", *flags.split())
        ts, inv, mig, single, leaf = (G[name + k] for k in ("_t_start", "_inv2N", "_mig", "_single", "_leaf_pop"))
        E, P = inv.shape
        assert d["npop"] == P and d["epochs"] == E and len(d["rows"]) == E, name
        assert d["leaf_pop"] == list(leaf), name
        np.testing.assert_allclose([e[0] for e in d["rows"]], ts, rtol=1e-14)
        np.testing.assert_allclose([e[1] for e in d["rows"]], inv, rtol=1e-14)
        np.testing.assert_allclose(np.array([e[2] for e in d["rows"]]).reshape(E, P, P), mig, rtol=1e-14)
        np.testing.assert_allclose(np.array([e[3] for e in d["rows"]]).reshape(E, P, P), single, rtol=1e-14)
        assert d["mutation_rate"] == float(G[name + "_mu"]) and d["recombination_rate"] == float(G[name + "_rho"])


def test_smc2_frontend_command_lines_parse_like_the_reference():
    """the command lines the reference's Python frontend emits for its own model classes (populationmodels.py
    core_command_line, EM and -vb form: -I, -eN with event counts, -ema with diagonal placeholders and counts, -eM,
    -ej) give the driver the tables the reference's own parser builds from them (tests/golden/smc2_cmdlines.json,
    make_golden.py::golden_smc2_cmdlines)"""
    import json
    import numpy as np
    cases = json.load(open(os.path.join(ROOT, "tests", "golden", "smc2_cmdlines.json")))
    assert len(cases) >= 23 and any(c["vb"] and len(c["inv2N"][0]) == 2 for c in cases)
    for c in cases:
        tmax = [] if c.get("tmax_in_cmd") else ["-tmax", "4"]
        d = _dump("-nsam", str(c["nsam"]), "-Np", "10", *tmax, *c["cmd"].split(), "-seed", "1")
        E, P = np.array(c["inv2N"]).shape
        assert d["npop"] == P and d["epochs"] == E, c["model"]
        assert d["leaf_pop"] == c["leaf_pop"], c["model"]
        np.testing.assert_allclose([e[0] for e in d["rows"]], c["t_start"], rtol=1e-14)
        np.testing.assert_allclose([e[1] for e in d["rows"]], c["inv2N"], rtol=1e-14)
        np.testing.assert_allclose(np.array([e[2] for e in d["rows"]]).reshape(E, P, P), c["mig"], rtol=1e-14)
        np.testing.assert_allclose(np.array([e[3] for e in d["rows"]]).reshape(E, P, P), c["single_mig"], rtol=1e-14)
        assert d["mutation_rate"] == c["mu"] and d["recombination_rate"] == c["rho"]      # to the last bit


@pytest.mark.parametrize("args, message", [
    (["-I", "2", "2", "2"], "Model has multiple populations but no migration"),
    (["-I", "2", "2", "3", "1.0"], "do not add up to -nsam"),
    (["-I", "2", "2", "2", "-ej", "0.5", "2", "1", "-em", "0", "2", "1", "1"], "sort both arguments by time"),
    (["-I", "2", "2", "2", "1.0", "-en", "0.1", "3", "1"], "Invalid population"),
    (["-I", "5", "1", "1", "1", "1", "0", "1.0"], "1 to 4 populations"),
    (["-I", "3", "2", "1", "1", "1.0", "-es", "0.1", "1", "0.5", "-es", "0.2", "2", "0.5"], "1 to 4 populations"),
    (["-I", "2", "2", "2", "1.0", "-es", "0.1", "1", "1.5"], "greater than one"),
])
def test_structured_parse_errors(args, message):
    r = run("-nsam", "4", "-Np", "10", "-t", "10", "-r", "10", "1000", *args)
    assert r.returncode == 1 and message in r.stderr, r.stderr


def test_seg_validation(tmp_path):
    base = ["-nsam", "4", "-Np", "10", "-t", "10", "-r", "10", "100000"]
    r = run(*base, "-seg", "/nonexistent.seg")
    assert r.returncode == 1 and "Invalid input file" in r.stderr
    r = run(*base)
    assert r.returncode == 1 and "a .seg file is required" in r.stderr
    cases = {
        "1\t10\tT\tF\t1\t01\n": "Number of variant site is wrong",
        "1\t10\tT\tF\t1\t01x0\n": "Unknown character found in .seg file",
        "1\t10\tT\tF\t1\t0.00\n": "inconsistent unphased heterozygous marks",
        "1\t10\tT\tF\t1\t0100\n20\t5\tT\tF\t1\t0100\n": "Segments are not consecutive",
        "1\t10\n": "Require 3 or 6 columns",
        "1\t10\tT\tX\t1\t0100\n": "Expected T or F in .seg file column 3 and 4",
        "200000\t10\tT\tF\t1\t0100\n": "No data found in file",
    }
    for k, (text, msg) in enumerate(cases.items()):
        f = tmp_path / ("c%d.seg" % k)
        f.write_text(text)
        r = run(*base, "-seg", str(f))
        assert r.returncode == 1 and msg in r.stderr, (text, r.stderr)


def test_fails_loudly_without_a_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = run("-nsam", "4", "-Np", "10", "-t", "2000", "-r", "800", "2000000", "-seg", SEG, "-o", str(tmp_path / "x"))
    assert r.returncode == 1
    assert "no CUDA device available" in r.stderr and "no CPU fallback" in r.stderr
