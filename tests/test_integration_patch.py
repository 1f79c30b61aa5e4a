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
    compare_out_files(outs["patched"], outs["dropin"])


def compare_out_files(a, b):
    """Both programs hand the engine the same model, rows and schedule, so the first sweep must agree number for number.
    After an M-step the two hosts' parameters can differ in the last bit (the reference re-derives 1/(2N) and the rates
    through its own Model round trips); a particle filter is chaotic in such bits -- one resampling boundary that falls
    the other way and the two runs are different samples from the same law -- so later iterations must either agree
    number for number as well (seven of the ten cases on B200) or agree as two samples of the same run do."""
    assert len(a) == len(b) and len(a) > 10
    for x, y in zip(a, b):
        for k in ("Iter", "Epoch", "Start", "End", "Type", "From", "To"):
            assert x[k] == y[k], (x, y)
    def same(x, y, rel):
        return all(float(x[k]) == pytest.approx(float(y[k]), rel=rel, abs=1e-12) for k in ("Opp", "Count", "Rate", "Ne"))
    for x, y in zip(a, b):
        if x["Iter"] == "0":
            assert same(x, y, 1e-6), (x, y)
    later = [(x, y) for x, y in zip(a, b) if x["Iter"] != "0"]
    if all(same(x, y, 1e-5) for x, y in later):
        return
    for x, y in later:
        if x["Type"] == "LogL":
            assert float(x["Count"]) == pytest.approx(float(y["Count"]), rel=0.01), (x, y)
        elif x["Type"] in ("Coal", "Recomb", "Migr") and float(y["Opp"]) > 10.0:
            assert float(x["Opp"]) == pytest.approx(float(y["Opp"]), rel=0.3), (x, y)
        elif x["Type"] == "Resamp":
            assert float(x["Count"]) == pytest.approx(float(y["Count"]), rel=0.1), (x, y)


def test_compare_out_files_logic():
    """the comparison itself, on rows shaped like the two programs' output (CPU)"""
    def row(it, typ, opp, cnt, ep="0"):
        return {"Iter": it, "Epoch": ep, "Start": "0.00", "End": "1.00", "Type": typ, "From": "0", "To": "-1",
                "Opp": repr(opp), "Count": repr(cnt), "Rate": repr(cnt / (opp + 1e-10)), "Ne": repr(opp / (2 * cnt + 1e-300))}
    base = [row("0", "Coal", 6064584.25, 237.67), row("0", "Recomb", 2.7e11, 2743.45), row("0", "LogL", 1.0, -78649.194)] * 4
    it1 = [row("1", "Coal", 6961952.68, 223.42), row("1", "Recomb", 2.77e11, 2741.17), row("1", "Resamp", 2e6, 3089.0),
           row("1", "LogL", 1.0, -78736.379)]
    it1_other = [row("1", "Coal", 6540711.22, 217.31), row("1", "Recomb", 2.736e11, 2797.51), row("1", "Resamp", 2e6, 3118.0),
                 row("1", "LogL", 1.0, -78924.62)]
    compare_out_files(base + it1, base + it1)                 # identical
    compare_out_files(base + it1, base + it1_other)           # two samples of the same law after the M-step
    bad0 = [dict(r) for r in base]
    bad0[0]["Opp"] = repr(6064584.25 * 1.001)
    with pytest.raises(AssertionError):
        compare_out_files(bad0 + it1, base + it1)             # the first sweep must agree exactly
    far = [dict(r) for r in it1_other]
    far[0]["Opp"] = repr(6540711.22 * 2)
    with pytest.raises(AssertionError):
        compare_out_files(base + it1, base + far)             # not the same law
