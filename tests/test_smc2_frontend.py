"""The reference's Python frontend, unchanged, driving the drop-in executable (BASELINE north_star: "the smc2 Python
frontend runs unchanged"): smcsmc.run_smcsmc with -smcsmcpath smcsmc_b200/bin/smcsmc, two chunks, two EM iterations.
The frontend builds the command lines, launches one process per chunk, parses every chunk's .out, adds them up, runs its
own M-step (reference smcsmc/model.py:989-1100) and writes result.out; the same run with the reference binary is the
committed fixture tests/golden/smc2_frontend_reference.out (tests/golden/make_golden.py::golden_smc2_frontend)."""
import os
import subprocess
import sys

import numpy as np
import pytest

import smc2_helper
from test_em_parity import parse_out

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "smcsmc_b200", "bin", "smcsmc")
SEG = os.path.join(ROOT, "tests", "data", "const_n4_2Mb_seed7.seg")
GOLD = os.path.join(ROOT, "tests", "golden", "smc2_frontend_reference.out")


def test_smc2_runs_unchanged_against_the_drop_in_executable(tmp_path):
    if not smc2_helper.available():
        pytest.skip("baseline/_ref/smcsmc (pip install of the reference's frontend) not present")
    out = str(tmp_path / "run")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "smc2_helper.py"), BIN, out, SEG], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
    log = open(os.path.join(out, "result.log")).read()
    assert log.count("Found all chunks; calculating sufficient statistics") == 2      # both E-steps parsed
    assert "Setting population size (epoch 4 population 0)" in log                    # its M-step ran on our tallies
    for it in (0, 1):
        for c in (0, 1):
            assert os.path.getsize(os.path.join(out, "emiter%d" % it, "chunk%d.out" % c)) > 0
            assert "smcsmc_b200" in open(os.path.join(out, "emiter%d" % it, "chunk%d.log" % c)).read()
    hdr, rows = parse_out(os.path.join(out, "result.out"))
    ghdr, gold = parse_out(GOLD)
    assert hdr == ghdr
    assert [(x["Iter"], x["Epoch"], x["Type"], x["From"], x["To"]) for x in rows] == \
           [(x["Iter"], x["Epoch"], x["Type"], x["From"], x["To"]) for x in gold]
    for a, b in zip(rows, gold):
        if a["Type"] == "Coal" and float(b["Count"]) > 100:      # one seed each: ~1/sqrt(count) noise plus EM feedback
            assert float(a["Opp"]) == pytest.approx(float(b["Opp"]), rel=0.20), (a, b)
            assert float(a["Count"]) == pytest.approx(float(b["Count"]), rel=0.20), (a, b)
            if a["Iter"] == "1":
                assert 6000 < float(a["Ne"]) < 16000, a               # simulated: 10 000
        if a["Type"] == "LogL":
            assert float(a["Count"]) == pytest.approx(float(b["Count"]), rel=0.01), (a, b)
