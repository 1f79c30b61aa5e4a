#!/bin/bash
# TEST INFRASTRUCTURE ONLY.  The reference with its uninitialised ForestState::max_epoch_to_record_ initialised
# (tools/patch_reference_uninit.py: two added lines in a temporary copy of particle.cpp) ->
# oracle/_ref/smcsmc_fixed (git-ignored, travels with gpurun).  Every other object is the unmodified reference's, as
# oracle/build_ref.sh compiled it.  The end-to-end fixtures under tests/golden/ come from this binary: the unmodified
# one loses 15-20 % of the recombination opportunity through undefined behaviour and is not reproducible.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SMC_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
ROOT="$HERE/.."
if [ ! -d "$REF/src" ]; then echo "build_fixed_ref.sh: $REF/src not present; keeping the prebuilt file" >&2; exit 0; fi
[ -f "$OUT/obj/count.o" ] || "$HERE/build_ref.sh" > /dev/null
if [ -f "$OUT/smcsmc_fixed" ] && [ "$OUT/smcsmc_fixed" -nt "$ROOT/tools/patch_reference_uninit.py" ] && \
   [ "$OUT/smcsmc_fixed" -nt "$OUT/obj/main_smcsmc.o" ]; then echo "oracle/_ref/smcsmc_fixed up to date"; exit 0; fi
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
python "$ROOT/tools/patch_reference_uninit.py" "$REF/src/particle.cpp" "$TMP/particle_fixed.cpp"
R="$REF/src"; S="$R/scrm/src"
DEFS='-DNDEBUG -DLANE_MAX_SIZE=50 -DVERSION=\"ref\" -DSMCSMCVERSION=\"ref+fix\" -DSCRMVERSION=\"ref\" -DCOMPILEDATE=\"ref\"'
eval g++ -std=c++11 -O3 -g -fPIC -w $DEFS -I"$HERE/boost_shim" -I"$R" -I"$S" -I"$S/random" \
  -c "$TMP/particle_fixed.cpp" -o "$TMP/particle_fixed.o"
objs=""
for o in "$OUT"/obj/*.o; do
  case "$(basename "$o")" in main_scrm.o|particle.o) ;; *) objs="$objs $o" ;; esac
done
g++ -o "$OUT/smcsmc_fixed" "$TMP/particle_fixed.o" $objs -lz
echo "built: $OUT/smcsmc_fixed"
