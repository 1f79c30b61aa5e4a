#!/bin/bash
# TEST INFRASTRUCTURE ONLY.  Compiles the UNMODIFIED reference (luntergroup/smcsmc + its vendored scrm fork)
# from the sources where they lie under /root/reference into oracle/_ref/ (git-ignored, travels with gpurun):
#   oracle/_ref/scrm          data generator (src/scrm/src/scrm.cc + CMakeLists.txt:13-36 file list)
#   oracle/_ref/smcsmc        the reference CPU particle filter (CMakeLists.txt:38-48), timed as cpu_baseline
#   oracle/_ref/libsmcref.so  function-level oracle: oracle/ref_harness.cpp + the same objects
# Boost (gzip streams + ends_with only) is replaced by oracle/boost_shim (zlib-backed).
# Nothing under the product path may use these files; only tests/, smoke() and bench.py's cpu_baseline leg do.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${SMC_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then
  echo "build_ref.sh: $REF/src not present; keeping prebuilt files in $OUT" >&2
  exit 0
fi
mkdir -p "$OUT/obj"
R="$REF/src"; S="$R/scrm/src"
DEFS='-DNDEBUG -DLANE_MAX_SIZE=50 -DVERSION=\"ref\" -DSMCSMCVERSION=\"ref\" -DSCRMVERSION=\"ref\" -DCOMPILEDATE=\"ref\"'
INC="-I$HERE/boost_shim -I$R -I$S -I$S/random"
CXXFLAGS="-std=c++11 -O3 -g -fPIC -w"
SCRM_SRCS="$S/node.cc $S/node_container.cc $S/time_interval.cc $S/tree_point.cc $S/forest.cc $S/forest-debug.cc \
 $S/random/random_generator.cc $S/random/mersenne_twister.cc $S/random/fastfunc.cc $S/param.cc $S/model.cc \
 $(ls $S/summary_statistics/*.cc)"
SMC_SRCS="$R/coalevent.cpp $R/particle.cpp $R/particleContainer.cpp $R/count.cpp $R/count-semi-debug.cpp \
 $R/pattern.cpp $R/pfparam.cpp $R/segdata.cpp $R/arena.cpp $R/model_summary.cpp"
stamp="$OUT/obj/.stamp"
need=0
for f in "$OUT/scrm" "$OUT/smcsmc" "$OUT/libsmcref.so"; do [ -f "$f" ] || need=1; done
[ "$HERE/ref_harness.cpp" -nt "$OUT/libsmcref.so" ] && need=1
[ "$need" = 0 ] && { echo "oracle/_ref up to date"; exit 0; }
objs=""
pids=""
compile() { # src obj
  if [ ! -f "$2" ] || [ "$1" -nt "$2" ]; then
    eval g++ $CXXFLAGS $DEFS $INC -c "$1" -o "$2" &
    pids="$pids $!"
  fi
}
for f in $SCRM_SRCS $SMC_SRCS; do
  o="$OUT/obj/$(basename "$f" | sed 's/\.[a-z]*$//').o"
  objs="$objs $o"
  compile "$f" "$o"
done
compile "$S/scrm.cc" "$OUT/obj/main_scrm.o"
compile "$R/smcsmc.cpp" "$OUT/obj/main_smcsmc.o"
for p in $pids; do wait $p; done
scrm_objs=""
for f in $SCRM_SRCS; do scrm_objs="$scrm_objs $OUT/obj/$(basename "$f" | sed 's/\.[a-z]*$//').o"; done
g++ -o "$OUT/scrm" "$OUT/obj/main_scrm.o" $scrm_objs
g++ -o "$OUT/smcsmc" "$OUT/obj/main_smcsmc.o" $objs -lz
if [ -f "$HERE/ref_harness.cpp" ]; then
  eval g++ $CXXFLAGS $DEFS $INC -shared -o "$OUT/libsmcref.so" "$HERE/ref_harness.cpp" $objs -lz
fi
echo "built: $(ls $OUT | tr '\n' ' ')"
