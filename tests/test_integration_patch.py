"""INTEGRATION.md, patch B, made executable: the reference's OWN driver -- its PfParam, scrm Param / Model, Segment,
CountModel with the M-step and the .out writer, all unmodified -- with the particle filter inside pfARG_core replaced by
the B200 engine (smcsmc_b200/host/reference_patch_body.inc, applied by tools/patch_reference_driver.py to a temporary copy
of src/smcsmc.cpp; compiled and linked by oracle/build_patched_ref.sh against the reference's objects and libsmcgpu.so).

CPU: the patch applies, the result compiles against the reference's headers and links, the binary parses a command
line with the reference's parser and reaches the engine, which refuses to run without a GPU.
GPU: the patched reference driver and the repo's drop-in driver, given the same command line and seed, hand the engine
the same model, rows and schedule -- their .out files agree number for number (both programs format with their own
copy of FormatDouble), for a single population and for a structured model, through one EM update."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCHED = os.path.join(ROOT, "oracle", "_ref", "smcsmc_patched")
BIN = os.path.join(ROOT, "smcsmc_b200", "bin", "smcsmc")
SEG1 = os.path.join(ROOT, "tests", "data", "const_n4_2Mb_seed7.seg")
SEG2 = os.path.join(ROOT, "tests", "data", "split_mig_n8_2Mb_seed13.seg")
HAVE_REF = os.path.isdir("/root/reference/src")
BASE1 = "-nsam 4 -Np 300 -N0 10000 -t 2000 -r 800 2000000 -eN 0 1 -eN 0.05 1 -eN 0.2 1 -eN 0.6 1"


@pytest.mark.skipif(not HAVE_REF, reason="needs the reference sources (/root/reference) to build the patched driver")
def test_patch_applies_compiles_and_links(tmp_path):
    r = subprocess.run([os.path.join(ROOT, "oracle", "build_patched_ref.sh")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert os.path.exists(PATCHED)
    # the patched file keeps the reference's code outside the particle filter: none of ParticleContainer is left in
    # pfARG_core, the calibration / M-step / writer calls are all still there
    out = str(tmp_path / "patched.cpp")
    subprocess.run(["python", os.path.join(ROOT, "tools", "patch_reference_driver.py"),
                    "/root/reference/src/smcsmc.cpp", out], check=True)
    text = open(out).read()
    core = text.split("void pfARG_core(PfParam &pfARG_para,")[-1]        # the definition (a declaration precedes main)
    for kept in ("calculate_median_survival_distances", "countNe->init()", "countNe->reset_lag", "dump_local_recomb_logs",
                 "countNe->reset_model_parameters", "countNe->log_counts", "smc_sweep(", "smc_set_populations("):
        assert kept in core, kept
    assert "update_state_to_data" not in core and "current_states" not in core
    r = subprocess.run([PATCHED, "-v"], capture_output=True, text=True)
    assert r.returncode == 0 and "smcsmc version" in r.stdout
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([PATCHED, "-nsam", "4", "-Np", "10", "-t", "2000", "-r", "800", "2000000", "-lag", "20000",
                            "-seg", SEG1, "-o", str(tmp_path / "x")], capture_output=True, text=True)
        assert r.returncode == 1 and "no CUDA device available" in r.stderr      # parsed, calibrated, reached the engine


def _rows(path):
    rows = []
    with open(path) as f:
        hdr = f.readline().split()
        for line in f:
            rows.append(dict(zip(hdr, line.split())))
    return rows


@pytest.mark.gpu
@pytest.mark.parametrize("name, seg, model", [
    ("single", SEG1, "-nsam 4 -Np 400 -N0 10000 -t 2000 -r 800 2000000 -eN 0 1 -eN 0.05 1 -eN 0.2 1 -eN 0.6 1"),
    ("structured", SEG2, "-nsam 8 -Np 300 -N0 10000 -t 2000 -r 800 2000000 -I 2 4 4 -em 0 2 1 1 -eN 0.1 1 -ej 0.5 2 1"),
    # -vb in the form the smc2 frontend writes it: an event count after every size and every migration rate
    ("single_vb", SEG1, "-nsam 4 -Np 400 -N0 10000 -t 2000 -r 800 2000000 -vb -eN 0 1 40 -eN 0.05 1 60 -eN 0.2 1 25 "
                        "-eN 0.6 1 10000000000.0"),
    ("structured_vb", SEG2, "-nsam 8 -Np 300 -N0 10000 -t 2000 -r 800 2000000 -I 2 4 4 -vb -eN 0 1 50 "
                            "-ema 0 0 1e10 0 1e10 1 30 0 1e10 -eN 0.1 1 80 -ema 0.1 0 1e10 0 1e10 1 45 0 1e10 "
                            "-eN 0.5 1 1e10 -ema 0.5 0 1e10 0 1e10 1 1e10 0 1e10 -ej 0.5 2 1"),
    # the other flags of the boundary that reach the engine or the M-step
    ("aa_masks", SEG1, BASE1 + " -ancestral_aware -xr 1-2 -xc 0"),
    ("cap_ess", SEG1, BASE1 + " -cap 9800 -ESS 0.9"),
    ("pattern", SEG1, "-nsam 4 -Np 300 -N0 10000 -t 2000 -r 800 2000000 -p 1*2+3*2+1 -tmax 3"),
    ("window", SEG1, "-nsam 4 -Np 300 -N0 10000 -t 1000 -r 400 1000000 -eN 0 1 -eN 0.1 1 -startpos 500001"),
    ("dephase", SEG2, "-nsam 8 -Np 200 -N0 10000 -t 2000 -r 800 2000000 -eN 0 1 -eN 0.1 1 -dephase"),
    ("admixture", SEG2, "-nsam 8 -Np 200 -N0 10000 -t 2000 -r 800 2000000 -I 2 4 4 0.5 -es 0.05 1 0.4 -ej 0.3 3 2 "
                        "-ej 0.8 2 1"),
])
def test_patched_reference_driver_equals_drop_in_driver(tmp_path, name, seg, model):
    if not os.path.exists(PATCHED):
        pytest.skip("oracle/_ref/smcsmc_patched was not built (needs /root/reference at build time)")
    outs = {}
    for who, exe in (("patched", PATCHED), ("dropin", BIN)):
        out = str(tmp_path / who)
        cmd = [exe] + model.split() + ["-lag", "20000", "-EM", "1", "-seed", "5", "-seg", seg, "-o", out]
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, (who, r.stderr[-2000:])
        outs[who] = _rows(out + ".out")
    a, b = outs["patched"], outs["dropin"]
    assert len(a) == len(b) and len(a) > 10
    for x, y in zip(a, b):
        for k in ("Iter", "Epoch", "Start", "End", "Type", "From", "To"):
            assert x[k] == y[k], (x, y)
        for k in ("Opp", "Count", "Rate", "Ne"):
            assert float(x[k]) == pytest.approx(float(y[k]), rel=1e-6, abs=1e-12), (k, x, y)
