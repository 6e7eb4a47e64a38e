#!/bin/bash
# oracle/build_ref.sh -- TEST INFRASTRUCTURE.
# Compiles the reference's OWN sources for the host, from where they lie under /root/reference,
# into oracle/_ref/libphoton_ref_cpu.so (git-ignored).  Nothing is copied into the repo: the
# sources are streamed through sed into g++.  The only edits made on the fly are mechanical:
#   * "#include" lines are dropped (the shim header provides the std/CUDA/teem declarations, the
#     reference's local headers are streamed in order instead);
#   * kernel<<<grid,block>>>(args) becomes shim::launch(grid, block, [&](auto... a) { kernel(a...); }, args);
#   * the hard-coded /scratch path of setData's debug dump of the gradient field (trace.h:1840-1843) becomes
#     shim::gradient_dump_path(): $PHOTON_REF_GRADIENT_DUMP if set (the field-ingest tests read the reference's
#     own float4 texels from it), else /dev/null.
# See oracle/ref_shim/cuda_on_host.h for what is emulated (CUDA runtime, texture unit, cuRAND, teem).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${PHOTON_REFERENCE:-/root/reference}/cuda_codes
OUT="$HERE/_ref"
[ -d "$REF" ] || { echo "reference sources not found at $REF"; exit 2; }
mkdir -p "$OUT"
CXX=/usr/bin/g++; [ -x $CXX ] || CXX=g++
strip_inc() { sed -E '/^[[:space:]]*#[[:space:]]*include/d' "$1"; }
{
  echo "#include \"$HERE/ref_shim/cuda_on_host.h\""
  echo "#include \"$HERE/ref_shim/teem_shim.h\""
  echo '#line 1 "cutil_math_bugfixes.h"'
  strip_inc "$REF/CubicInterpolationCUDA/code/internal/cutil_math_bugfixes.h"
  echo '#line 1 "parallel_ray_tracing.h"'
  strip_inc "$REF/parallel_ray_tracing.h"
  echo '#line 1 "float3_operators.h"'
  strip_inc "$REF/float3_operators.h"
  echo "#include \"$HERE/ref_shim/tricubic_stubs.h\""
  echo '#line 1 "trace_rays_through_density_gradients.h"'
  strip_inc "$REF/trace_rays_through_density_gradients.h" | sed -E 's#"/scratch/[^"]*gradient_central.bin"#shim::gradient_dump_path()#'
  echo '#line 1 "parallel_ray_tracing.cu"'
  strip_inc "$REF/parallel_ray_tracing.cu" | sed -E 's/([A-Za-z_][A-Za-z_0-9]*)[[:space:]]*<<<([^>]*)>>>[[:space:]]*\(/shim::launch(\2, [\&](auto... a_) { \1(a_...); }, /'
  echo "#include \"$HERE/ref_shim/exports.h\""
} | $CXX -x c++ -std=c++14 -O3 -fopenmp -ffp-contract=off -fPIC -shared -D__CUDACC__ -w -fpermissive \
        -o "$OUT/libphoton_ref_cpu.so" - -lz -lm
echo "built $OUT/libphoton_ref_cpu.so"
