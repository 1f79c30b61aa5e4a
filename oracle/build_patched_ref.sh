#!/bin/bash
# TEST INFRASTRUCTURE ONLY.  INTEGRATION.md, patch B, made executable: the reference's own driver (PfParam, scrm
# Param/Model, Segment, CountModel, .out writer, M-step -- all its own code) with the particle filter replaced by the
# B200 engine.  tools/patch_reference_driver.py writes the patched main file to a temporary directory (no reference
# source enters the repo), this script compiles it against the reference's headers and links it with the objects
# oracle/build_ref.sh made plus libsmcgpu.so -> oracle/_ref/smcsmc_patched (git-ignored, travels with gpurun).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SMC_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
ROOT="$HERE/.."
if [ ! -d "$REF/src" ]; then echo "build_patched_ref.sh: $REF/src not present; keeping the prebuilt file" >&2; exit 0; fi
[ -f "$OUT/obj/count.o" ] || "$HERE/build_ref.sh" > /dev/null
TMP="$(mktemp -d)"
trap 'rm -rf "$TMP"' EXIT
python "$ROOT/tools/patch_reference_driver.py" "$REF/src/smcsmc.cpp" "$TMP/smcsmc_patched.cpp"
R="$REF/src"; S="$R/scrm/src"
DEFS='-DNDEBUG -DLANE_MAX_SIZE=50 -DVERSION="ref+b200" -DSMCSMCVERSION="ref+b200" -DSCRMVERSION="ref" -DCOMPILEDATE="ref"'
objs=""
for o in "$OUT"/obj/*.o; do
  case "$(basename "$o")" in main_scrm.o|main_smcsmc.o) ;; *) objs="$objs $o" ;; esac
done
eval g++ -std=c++11 -O2 -w $DEFS -include "$HERE/integration_prelude.h" -I"$HERE/boost_shim" -I"$R" -I"$S" -I"$S/random" \
  -I"$ROOT/include" -I"$HERE" -c "$TMP/smcsmc_patched.cpp" -o "$TMP/main_patched.o"
g++ -o "$OUT/smcsmc_patched" "$TMP/main_patched.o" $objs -L"$ROOT/smcsmc_b200/lib" -lsmcgpu -lz \
  -Wl,-rpath,'$ORIGIN/../../smcsmc_b200/lib'
echo "built: $OUT/smcsmc_patched"
