#!/usr/bin/env bash
# Build the UNMODIFIED reference (biklooost/PROPhet) CPU code for use as the parity oracle.
#
# Test infrastructure only.  Nothing under prophet_b200/ may link or run what this script builds.
#
# The reference sources are compiled from where they lie (default /root/reference/src); no source
# file is copied.  Two missing-return undefined-behaviour sites crash g++ >= 8 at -O2/-O3
# (SURVEY.md section 8c): they are patched *in the preprocessed stream* between `g++ -E` and the
# compile step, changing no arithmetic:
#   * NN_node.h:103      `inline int set_index(int)` falls off its end
#   * xml_reader.cpp:160 `get_node_by_name(..., xml_node)` falls off its end when nothing matches
# Flags follow src/Makefile.in:8,88 (-O3 -DNOT_MD, no MPI -> Parallel.h serial stubs).
# Outputs go only to oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).
set -euo pipefail
REF=${PROPHET_REFERENCE:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
SRC=$REF/src
if [ ! -d "$SRC" ]; then
  echo "build_ref.sh: reference sources not found at $SRC (set PROPHET_REFERENCE)" >&2
  exit 3
fi
mkdir -p "$OUT/obj"
CXX=${CXX:-g++}
CXXFLAGS="-O3 -DNOT_MD -w -std=gnu++14"
# src/Makefile.in:26-29 source list (the shipped stale Makefile omits CUSTOM.cpp)
SOURCES="Structure Error Tables Functional Potential System_data Setup Parallel Grid_data System VASP QE FHIAIMS CUSTOM Optimizer Functional_params xml_reader pugixml NN NN_node NF NF_node Atom Analysis"

patch_stream() {
  perl -0pe '
    s/(int\s+set_index\s*\(\s*int\s+new_index\s*\)\s*\{\s*my_index\s*=\s*new_index\s*;)/$1 return my_index;/g;
    s/(child\s*=\s*child\.next_sibling\(\)\s*;\s*\}\s*)(\})/$1 return pugi::xml_node(); $2/g;
  '
}

compile_one() {
  local name=$1 extra=${2:-}
  local obj=$OUT/obj/$name.o
  if [ "$obj" -nt "$SRC/$name.cpp" ] && [ "$obj" -nt "$0" ]; then return 0; fi
  $CXX $CXXFLAGS $extra -I"$SRC" -E "$SRC/$name.cpp" | patch_stream | \
    $CXX $CXXFLAGS $extra -x c++-cpp-output -c - -o "$obj"
}

pids=()
for s in $SOURCES; do compile_one "$s" & pids+=($!); done
compile_one main & pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done

OBJS=""
for s in $SOURCES; do OBJS="$OBJS $OUT/obj/$s.o"; done
# the reference CLI itself (src/main.cpp) -- used by bench.py --impl reference and cpu_baseline
$CXX -O3 $OBJS "$OUT/obj/main.o" -o "$OUT/PROPhet" -lm
# dump driver: our own code, linked against the reference objects
$CXX $CXXFLAGS -I"$SRC" -E "$HERE/ref_driver.cpp" | patch_stream | \
  $CXX $CXXFLAGS -x c++-cpp-output -c - -o "$OUT/obj/ref_driver.o"
$CXX -O3 $OBJS "$OUT/obj/ref_driver.o" -o "$OUT/ref_driver" -lm
echo "built $OUT/PROPhet and $OUT/ref_driver"
