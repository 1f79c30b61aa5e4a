#!/usr/bin/env python
"""Write a copy of the reference's particle.cpp with ONE defect removed (test tooling, not product code).

The defect (located with oracle/ref_harness.cpp::ref_spawn_probe): neither constructor of ForestState
initialises `max_epoch_to_record_` (particle.hpp:254; constructors particle.cpp:80-107 and :113-136).  The copy
constructor is what `spawn()` runs when a particle with multiplicity > 1 recombines (particle.cpp:843); the copy then
calls resample_recombination_position() (:852) BEFORE ParticleContainer::extend_ARGs assigns the field
(particleContainer.cpp:123), and that function's guard `epoch <= max_epoch_to_record_` (:402) reads whatever the heap
block held.  Epochs above the garbage value keep their recombination-opportunity records un-extended, so the stretch
from the spawn point to the copy's next recombination is never tallied for them: with resampling active the unmodified
binary reports 15-20 % less recombination opportunity than sum_i w_i * (local tree length)_i * bp, and the figure
changes with the size of the environment block.  The prior trees' first records (ParticleContainer ctor ->
record_recomb_extension, particleContainer.cpp:54) read the same uninitialised field.

The patch, by anchor text: the copy constructor copies the field, the main constructor sets it to INT_MAX (record every
epoch until extend_ARGs says otherwise).  Two added lines; everything else is the reference's, byte for byte.  The result
is written OUTSIDE the repo (oracle/build_fixed_ref.sh gives a temporary directory): no reference source enters the tree."""
import sys

COPY_ANCHOR = "    _current_seq_idx = copied_state._current_seq_idx;\n    setMultiplicity( 1 );\n"
MAIN_ANCHOR = "    owning_model_and_random_generator = own_model_and_random_generator;\n"


def patch(src_path, out_path):
    text = open(src_path).read()
    for anchor in (COPY_ANCHOR, MAIN_ANCHOR):
        if text.count(anchor) != 1:
            raise SystemExit("anchor not found exactly once in the reference's particle.cpp: %r" % anchor)
    if "max_epoch_to_record_ =" in text:
        raise SystemExit("the reference already initialises max_epoch_to_record_; nothing to patch")
    text = text.replace(COPY_ANCHOR, COPY_ANCHOR + "    max_epoch_to_record_ = copied_state.max_epoch_to_record_;   // b200 fix\n")
    text = text.replace(MAIN_ANCHOR, MAIN_ANCHOR + "    max_epoch_to_record_ = INT_MAX;   // b200 fix\n")
    open(out_path, "w").write(text)


if __name__ == "__main__":
    if len(sys.argv) != 3:
        raise SystemExit("usage: patch_reference_uninit.py <reference>/src/particle.cpp <out.cpp>")
    patch(sys.argv[1], sys.argv[2])
