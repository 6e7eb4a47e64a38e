#!/bin/bash
# oracle/build_ref_ci.sh -- TEST INFRASTRUCTURE.
# Compiles CubicInterpolationCUDA's own CPU reference (the plain-C++ files of examples/referenceCubicTexture3D and
# the host/device headers they include) for the host, from where they lie under /root/reference, into
# oracle/_ref/libphoton_ref_ci.so (git-ignored).  Nothing is copied into the repo.  On-the-fly edits: "#include"
# lines are dropped (streamed in order instead) and each file sits in its own namespace, because
# cubicFilter3D_kernel.hpp and bspline_kernel.cu both define bspline().  cutil's timer calls (cutil.h would pull GL
# headers) are no-ops.  This is the known-answer test of the product's cubic mode (tests/test_cubic_mode.py).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
CI=${PHOTON_REFERENCE:-/root/reference}/cuda_codes/CubicInterpolationCUDA
OUT="$HERE/_ref"
[ -d "$CI" ] || { echo "reference sources not found at $CI"; exit 2; }
mkdir -p "$OUT"
CXX=/usr/bin/g++; [ -x $CXX ] || CXX=g++
strip_inc() { sed -E 's/^[[:space:]]*#[[:space:]]*include.*$//' "$1"; }
{
  echo '#include <stdio.h>'
  echo '#include <math.h>'                       # what cubicPrefilter3D.hpp:44-45 includes: C++ <math.h> = sqrt(float) stays float
  echo "#include \"$HERE/ref_shim/cuda_on_host.h\""
  echo 'typedef unsigned int uint; typedef unsigned char uchar;'
  echo '#define CUT_SAFE_CALL(x) (x)'
  echo 'inline int cutCreateTimer(uint *t) { *t = 0; return 1; } inline int cutResetTimer(uint) { return 1; }'
  echo 'inline int cutStartTimer(uint) { return 1; } inline int cutStopTimer(uint) { return 1; } inline double cutGetTimerValue(uint) { return 0.0; }'
  echo '#line 1 "cutil_math_bugfixes.h"'
  strip_inc "$CI/code/internal/cutil_math_bugfixes.h"
  echo '#line 1 "math_func.cu"'
  strip_inc "$CI/code/internal/math_func.cu"
  echo 'namespace ci_prefilter {'
  echo '#line 1 "cubicPrefilter_kernel.cu"'
  strip_inc "$CI/code/internal/cubicPrefilter_kernel.cu"
  echo '#line 1 "cubicPrefilter3D.hpp"'
  strip_inc "$CI/examples/referenceCubicTexture3D/cubicPrefilter3D.hpp" | sed -E 's/^extern "C"//'
  echo '}'
  echo 'namespace ci_filter {'
  echo '#line 1 "cubicFilter3D_kernel.hpp"'
  strip_inc "$CI/examples/referenceCubicTexture3D/cubicFilter3D_kernel.hpp"
  echo '}'
  echo 'namespace ci_bspline {'
  echo '#line 1 "bspline_kernel.cu"'
  strip_inc "$CI/code/internal/bspline_kernel.cu"
  echo '}'
  echo "#include \"$HERE/ref_shim/ci_exports.h\""
  echo "#include \"$HERE/ref_shim/exports.h\""
} | $CXX -x c++ -std=c++14 -O2 -ffp-contract=off -fPIC -shared -D__CUDACC__ -w -fpermissive \
        -o "$OUT/libphoton_ref_ci.so" - -lm
echo "built $OUT/libphoton_ref_ci.so"
