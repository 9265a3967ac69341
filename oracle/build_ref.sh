#!/usr/bin/env bash
# TEST INFRASTRUCTURE -- not part of the product path.
#
# Builds oracle/_ref/libngsxref.so: the reference's OWN cut-quadrature / FE
# sources compiled where they lie under /root/reference (ngsxfem
# cutint/straightcutrule.cpp, cutint/spacetimecutrule.cpp, spacetime/SpaceTimeFE.cpp + the NGSolve fem / netgen ngcore files they need),
# plus our thin extern "C" driver oracle/ref_driver.cpp.  No reference source is
# copied into this repository; outputs go to oracle/_ref/ only (git-ignored).
#
# The reference's own CMake build is NOT used (it needs a network for
# submodules, Tcl/Tk, pybind11 fork): this is the short hand-written recipe the
# task allows.  Usage:  bash oracle/build_ref.sh [REF_ROOT]
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${1:-/root/reference}"
NGS="$REF/external_dependencies/ngsolve"
NG="$NGS/external_dependencies/netgen"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
if [ ! -d "$REF/cutint" ]; then
  echo "build_ref.sh: reference tree not found at $REF (expected on the GPU box: prebuilt _ref travels)"; exit 0
fi
mkdir -p "$OBJ"
PYSP="$(python3 -c 'import site;print(site.getsitepackages()[0])')"
OB="$PYSP/opencv_python_headless.libs"
OBLIB="$(ls "$OB"/libopenblasp-r0-*.so 2>/dev/null | head -1 || true)"

# MapPatchIntegrationPoint<D> (xfem/symboliccutbfi.cpp): the one template of that file the driver calls.  The file as a whole
# needs NGSolve's comp layer + SymbolicBilinearFormIntegrator (do not link here), so the template is lifted out at build time
# into the git-ignored output directory -- from "template<int D>" before "void MapPatchIntegrationPoint(" to its closing brace.
awk '/^  template<int D>$/ { held = $0; next }
     held != "" { if ($0 ~ /void MapPatchIntegrationPoint\(/) { print held; on = 1 } held = "" }
     on { print }
     on && /^  }$/ { exit }' "$REF/xfem/symboliccutbfi.cpp" > "$OUT/mappatch_extract.hpp"
grep -q "MapPatchIntegrationPoint" "$OUT/mappatch_extract.hpp" || { echo "build_ref.sh: could not extract MapPatchIntegrationPoint"; exit 1; }

INC="-I$OUT -I$NGS/include -I$NGS/ngstd -I$NGS/basiclinalg -I$NGS/fem -I$NGS/comp -I$NGS/linalg -I$NGS/solve \
 -I$NGS/multigrid -I$NGS/parallel -I$NG/libsrc/include -I$NG/libsrc -I$NG/libsrc/core -I$NG/libsrc/visualization \
 -I$REF/cutint -I$REF/utils -I$REF/xfem -I$REF -DELEMENT_MAXPOINTS=20"
CXX="g++ -std=c++17 -O2 -fPIC -w -Wno-psabi -fopenmp"

# netgen's version.hpp wants a generated netgen_version.hpp
if [ ! -f "$OUT/netgen_version.hpp" ]; then
cat > "$OUT/netgen_version.hpp" <<'EOF'
#ifndef NETGEN_VERSION_HPP_INCLUDED
#define NETGEN_VERSION_HPP_INCLUDED
#define NETGEN_VERSION "6.2-ref"
#define NETGEN_VERSION_MAJOR 6
#define NETGEN_VERSION_MINOR 2
#define NETGEN_VERSION_PATCH 0
#define NETGEN_VERSION_TWEAK 0
#define NETGEN_VERSION_HASH "ec40e605"
#define NETGEN_VERSION_GIT "v6.2-ref"
#endif
EOF
fi

compile() { # src obj extra
  local src="$1" obj="$2"; shift 2
  if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ]; then
    $CXX $INC "$@" -c "$src" -o "$obj"
  fi
}

pids=()
par() { "$@" & pids+=($!); if [ ${#pids[@]} -ge 8 ]; then wait "${pids[0]}"; pids=("${pids[@]:1}"); fi; }
waitall() { for p in "${pids[@]}"; do wait "$p"; done; pids=(); }

CORE_OBJS=()
for f in "$NG"/libsrc/core/*.cpp; do
  b="$(basename "$f" .cpp)"
  case "$b" in python_ngcore*) continue;; esac
  CORE_OBJS+=("$OBJ/core_$b.o"); par compile "$f" "$OBJ/core_$b.o"
done
STD_OBJS=()
for b in blockalloc evalfunc stringops bspline; do
  STD_OBJS+=("$OBJ/std_$b.o"); par compile "$NGS/ngstd/$b.cpp" "$OBJ/std_$b.o"
done
waitall

# generated SIMD kernels header needed by ngblas.cpp
if [ ! -f "$OUT/matkernel.hpp" ]; then
  compile "$NGS/basiclinalg/generate_mat_kernels.cpp" "$OBJ/genmk.o"
  g++ -fopenmp -o "$OUT/genmk" "${CORE_OBJS[@]}" "$OBJ/genmk.o" -lpthread
  (cd "$OUT" && ./genmk >/dev/null)
fi

BLA_OBJS=()
for b in calcinverse vecmat avector; do
  BLA_OBJS+=("$OBJ/bla_$b.o"); par compile "$NGS/basiclinalg/$b.cpp" "$OBJ/bla_$b.o"
done
BLA_OBJS+=("$OBJ/bla_ngblas.o"); par compile "$NGS/basiclinalg/ngblas.cpp" "$OBJ/bla_ngblas.o" -DLAPACK
FEM_OBJS=()
for b in recursive_pol elementtopology finiteelement scalarfe intrule elementtransformation h1lofe h1hofe h1hofe_trig h1hofe_segm \
         l2hofe l2hofe_segm l2hofe_trig l2hofe_tet diffop integrator; do
  FEM_OBJS+=("$OBJ/fem_$b.o"); par compile "$NGS/fem/$b.cpp" "$OBJ/fem_$b.o"
done
X_OBJS=("$OBJ/x_xdecompose.o" "$OBJ/x_straightcutrule.o" "$OBJ/x_SpaceTimeFE.o" "$OBJ/x_spacetimecutrule.o" "$OBJ/x_xfiniteelement.o" "$OBJ/x_xfemdiffops.o")
par compile "$REF/xfem/xfiniteelement.cpp" "$OBJ/x_xfiniteelement.o"
par compile "$REF/xfem/xfemdiffops.cpp" "$OBJ/x_xfemdiffops.o"
par compile "$REF/spacetime/SpaceTimeFE.cpp" "$OBJ/x_SpaceTimeFE.o"
par compile "$REF/cutint/spacetimecutrule.cpp" "$OBJ/x_spacetimecutrule.o"
par compile "$REF/cutint/xdecompose.cpp" "$OBJ/x_xdecompose.o"
par compile "$REF/cutint/straightcutrule.cpp" "$OBJ/x_straightcutrule.o"
if [ "$OUT/mappatch_extract.hpp" -nt "$OBJ/ref_driver.o" ]; then rm -f "$OBJ/ref_driver.o"; fi
par compile "$HERE/ref_driver.cpp" "$OBJ/ref_driver.o"
waitall

# link order matters (static-init order): ngcore, ngstd, bla, fem, ngsxfem, driver
LAPACK_LINK=()
if [ -n "$OBLIB" ]; then LAPACK_LINK=(-L"$OB" -l:"$(basename "$OBLIB")" -Wl,-rpath,"$OB"); fi
g++ -shared -fopenmp -o "$OUT/libngsxref.so" "${CORE_OBJS[@]}" "${STD_OBJS[@]}" "${BLA_OBJS[@]}" "${FEM_OBJS[@]}" \
    "${X_OBJS[@]}" "$OBJ/ref_driver.o" "${LAPACK_LINK[@]}" -lpthread
echo "built $OUT/libngsxref.so"
