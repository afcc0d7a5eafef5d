#!/bin/bash
# TEST INFRASTRUCTURE — builds the UNMODIFIED reference (Cantera 4.0.0a2) from the sources where
# they lie under /root/reference into oracle/_ref/ (git-ignored, travels to the GPU box).
# Nothing is copied into the repository history; the reference's own build system (SCons) is not
# run (scons and Boost are absent from this image): this is a direct g++/gcc recipe following
# ext/SConscript:64-117 and src/SConscript, with a ~150-line Boost stand-in (boost_shim/) for the
# string/rootfinding helpers — none of which does hot-path arithmetic.
#
# Outputs: oracle/_ref/libcantera_ref.a, oracle/_ref/inc (header tree: symlinks, container only),
#          oracle/_ref/data/*.yaml (mechanism DATA files the reference arm needs on the GPU box).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
R="${CANTERA_REF:-/root/reference}"
OUT="$ROOT/oracle/_ref"
JOBS="${JOBS:-$(nproc)}"
if [ ! -d "$R/src" ]; then
    echo "build_ref: $R not present (GPU box?) — using prebuilt oracle/_ref as is"; exit 0
fi
mkdir -p "$OUT/obj" "$OUT/inc/cantera" "$OUT/data" "$OUT/testdata"

# ---- header tree (SConscript: include/cantera/ext gets fmt, yaml-cpp, Eigen, sundials headers)
INC="$OUT/inc"
for d in "$R"/include/cantera/*; do ln -sfn "$d" "$INC/cantera/$(basename "$d")"; done
# base/ must be a real dir because config.h is generated into it
rm -f "$INC/cantera/base"; mkdir -p "$INC/cantera/base"
for f in "$R"/include/cantera/base/*; do ln -sfn "$f" "$INC/cantera/base/$(basename "$f")"; done
rm -f "$INC/cantera/ext"; mkdir -p "$INC/cantera/ext"
ln -sfn "$R/ext/fmt/include/fmt" "$INC/cantera/ext/fmt"
ln -sfn "$R/ext/yaml-cpp/include/yaml-cpp" "$INC/cantera/ext/yaml-cpp"
ln -sfn "$R/ext/eigen/Eigen" "$INC/cantera/ext/Eigen"
for sub in sundials sundials/priv nvector cvodes idas sunmatrix sunlinsol sunnonlinsol; do
    mkdir -p "$INC/cantera/ext/$sub"
    for h in "$R/ext/sundials/include/$sub"/*.h; do ln -sfn "$h" "$INC/cantera/ext/$sub/$(basename "$h")"; done
done
ln -sfn "$R/ext/sundials_export.h" "$INC/cantera/ext/sundials/sundials_export.h"
python3 - "$R" "$INC" <<'PY'
import sys, re
R, INC = sys.argv[1], sys.argv[2]
cfg = {
 "CANTERA_VERSION": '#define CANTERA_VERSION "4.0.0a2"',
 "CANTERA_SHORT_VERSION": '#define CANTERA_SHORT_VERSION "4.0"',
 "FTN_TRAILING_UNDERSCORE": "#define FTN_TRAILING_UNDERSCORE 1",
 "LAPACK_FTN_STRING_LEN_AT_END": "#define LAPACK_FTN_STRING_LEN_AT_END 1",
 "LAPACK_FTN_TRAILING_UNDERSCORE": "#define LAPACK_FTN_TRAILING_UNDERSCORE 1",
 "CT_USE_LAPACK": "#define CT_USE_LAPACK 0",
 "CT_USE_SYSTEM_FMT": "#define CT_USE_SYSTEM_FMT 0",
 "CT_USE_SYSTEM_YAMLCPP": "#define CT_USE_SYSTEM_YAMLCPP 0",
 "CT_SUNDIALS_USE_LAPACK": "#define CT_SUNDIALS_USE_LAPACK 0",
 "SUNDIALS_USE_GENERIC_MATH": "#define SUNDIALS_USE_GENERIC_MATH 1",
}
def fill(src, dst):
    txt = open(src).read()
    txt = re.sub(r"\{(\w+)!s\}", lambda m: cfg.get(m.group(1), "/* #undef %s */" % m.group(1)), txt)
    open(dst, "w").write(txt)
fill(R + "/include/cantera/base/config.h.in", INC + "/cantera/base/config.h")
fill(R + "/ext/sundials_config.h.in", INC + "/cantera/ext/sundials/sundials_config.h")
PY

# ---- mechanism data (YAML inputs, not source) for run time on the GPU box
cp -f "$R"/data/{h2o2,gri30,nDodecane_Reitz,air,h2o2_noch}.yaml "$OUT/data/" 2>/dev/null || true
cp -f "$R"/data/{h2o2,gri30,nDodecane_Reitz}.yaml "$OUT/data/"
cp -f "$R"/test/data/{pdep-test,kineticsfromscratch,sri-falloff,tsang-falloff,jacobian-tests,chemically-activated-reaction,explicit-third-bodies,reaction-orders}.yaml "$OUT/testdata/" 2>/dev/null || true

# ---- compile
CXXFLAGS="-std=c++20 -O2 -DNDEBUG -fPIC -w -I$INC -isystem $INC/cantera/ext -I$R/src -isystem $HERE/boost_shim \
 -I$R/ext/yaml-cpp/include -I$R/ext/fmt/include -DGIT_COMMIT=\"11a2381\""
CFLAGS="-O2 -DNDEBUG -fPIC -w -I$INC/cantera/ext -I$R/ext/sundials/src/sundials -I$R/ext/sundials/src"
export CXXFLAGS CFLAGS R OUT
{
  ls "$R"/src/{base,thermo,tpx,equil,numerics,kinetics,transport,oneD,zeroD}/*.cpp
  ls "$R"/src/extensions/canteraStatic.cpp
  ls "$R"/ext/yaml-cpp/src/*.cpp "$R"/ext/fmt/src/format.cc "$R"/ext/fmt/src/os.cc
} > "$OUT/cxx.list"
{
  ls "$R"/src/equil/*.c 2>/dev/null || true
  for d in sundials sundials/sundatanode nvector/serial cvodes idas sunmatrix/band sunmatrix/dense \
           sunmatrix/sparse sunlinsol/dense sunlinsol/band sunlinsol/spgmr sunnonlinsol/newton; do
    ls "$R"/ext/sundials/src/$d/*.c | grep -v '/fsun\|_mpi_\|sundials_profiler.c\|sundials_xbraid.c'
  done
} > "$OUT/c.list"
objname() { echo "$OUT/obj/$(echo "${1#$R/}" | tr '/' '_').o"; }
comp_cxx() { o=$(objname "$1"); [ -f "$o" ] && [ "$o" -nt "$1" ] || g++ $CXXFLAGS -c "$1" -o "$o" || { echo "FAIL $1"; exit 255; }; }
comp_c()   { o=$(objname "$1"); [ -f "$o" ] && [ "$o" -nt "$1" ] || gcc $CFLAGS -c "$1" -o "$o" || { echo "FAIL $1"; exit 255; }; }
export -f objname comp_cxx comp_c
xargs -P "$JOBS" -I{} bash -c 'comp_cxx {}' < "$OUT/cxx.list"
xargs -P "$JOBS" -I{} bash -c 'comp_c {}' < "$OUT/c.list"
rm -f "$OUT/libcantera_ref.a"
ar rcs "$OUT/libcantera_ref.a" "$OUT"/obj/*.o
echo "build_ref: OK $(ls "$OUT"/obj/*.o | wc -l) objects -> $OUT/libcantera_ref.a"
