#!/usr/bin/env python
"""Apply INTEGRATION.md's patch B to the reference's own driver (test tooling, not product code).

Reads <reference>/src/smcsmc.cpp, replaces the particle-filter part of pfARG_core by the engine block
smcsmc_b200/host/reference_patch_body.inc and writes the result to the given path (outside the repo: no reference source
is copied into the tree).  What changes, by anchor text rather than by line number:
  * `#include "smcsmc_b200.h"` after the reference's own includes;
  * the ParticleContainer constructor call is dropped (the engine draws the prior trees);
  * everything from "Go through seg data" up to the "Inference step completed" message -- the row loop, the final
    normalisation and the end-of-data flush -- becomes the engine block;
  * the LogL row and message take the engine's ln normaliser; the posterior-ARG dump (resample + printTrees) and
    current_states.clear() go (they need the particles themselves; `-arg` is row f4 of the scope table).
Every other line -- calibration, CountModel::init / reset_lag, the recombination-map dump, the M-step, log_counts,
PfParam, Segment, main() -- is the reference's, unchanged."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
BODY = os.path.join(HERE, "..", "smcsmc_b200", "host", "reference_patch_body.inc")


def find(lines, text, start=0):
    for i in range(start, len(lines)):
        if text in lines[i]:
            return i
    raise SystemExit("anchor not found in the reference driver: %r" % text)


def patch(src_path, out_path):
    lines = open(src_path).read().split("\n")
    body = open(BODY).read().rstrip("\n").split("\n")
    inc = max(i for i, l in enumerate(lines[:find(lines, "int main(")]) if l.startswith("#include"))   # last include
    ctor0 = find(lines, "ParticleContainer current_states(model, rg, &pfARG_para,")
    ctor1 = find(lines, "Segfile->allelic_state_at_Segment_end);", ctor0)
    loop0 = find(lines, "/*! Go through seg data */", ctor1)
    done = find(lines, 'clog << " Inference step completed." << endl;', loop0)
    out = lines[:inc + 1] + ['#include <cstring>', '#include <cstdint>', '#include "smcsmc_b200.h"',
                             '#include "smcsmc_b200_reference_access.h"   // oracle/: Segment::segment_state() is private']
    out += lines[inc + 1:ctor0] + lines[ctor1 + 1:loop0] + body + lines[done:]
    text = "\n".join(out)
    assert text.count("current_states.ln_normalization_factor()") == 2
    text = text.replace("current_states.ln_normalization_factor()", "smc_b200_lnZ")
    kept = []
    for l in text.split("\n"):
        if "current_states.resample(" in l or "current_states.printTrees(" in l or "current_states.clear();" in l:
            continue
        kept.append(l)
    text = "\n".join(kept)
    assert "current_states" not in text.split("void pfARG_core(PfParam &pfARG_para,")[-1].split("\n}\n")[0], \
        "a use of the particle container survived the patch"
    open(out_path, "w").write(text)


if __name__ == "__main__":
    if len(sys.argv) != 3:
        raise SystemExit("usage: patch_reference_driver.py <reference>/src/smcsmc.cpp <out.cpp>")
    patch(sys.argv[1], sys.argv[2])
