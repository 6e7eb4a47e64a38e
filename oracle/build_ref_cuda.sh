#!/bin/bash
# oracle/build_ref_cuda.sh -- TEST INFRASTRUCTURE ("Oracle B": the reference's OWN CUDA path on the B200).
# Compiles the reference's CUDA sources with nvcc for sm_100, from where they lie under /root/reference, into
# oracle/_ref/libphoton_ref_cuda.so (git-ignored, travels to the GPU box as a prebuilt file).  Nothing is copied
# into the repo: the sources are streamed through sed into nvcc.  Edits made on the fly, all mechanical:
#   * "#include" lines are dropped (cuda_texobj_patch.h provides the CUDA/std/teem declarations, the reference's
#     local headers are streamed in order instead; cutil.h would pull GL headers, helper_cuda.h CUDA-10 enums);
#   * the texture-REFERENCE API removed in CUDA 12 becomes a texture OBJECT (the patch BASELINE.json allows):
#       texture<float4, 3> tex_data;                 -> shimtex::Ref tex_data; __device__ cudaTextureObject_t tex_data_obj;
#       tex3D(tex_data, x, y, z)                     -> tex3D<float4>(tex_data_obj, x, y, z)
#       cudaBindTextureToArray(tex_data, arr, desc)  -> shimtex::bind(tex_data, tex_data_obj, arr)
#       cudaUnbindTexture(tex_data)                  -> shimtex::unbind(tex_data)
#     and the same for the (never executed) coeffs3D texture; the unused 2-D Mie texture declaration is dropped;
#   * the debug ofstream of the whole gradient field to a hard-coded /scratch path (trace.h:1840-1843) is sent
#     to /dev/null.
# Flags: -O3 -fmad=false.  The reference's own build is -G -O0 for sm_30 (Debug/subdir.mk): no FMA contraction
# there either, and -G on sm_100 would only make the checker slower.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${PHOTON_REFERENCE:-/root/reference}/cuda_codes
OUT="$HERE/_ref"
[ -d "$REF" ] || { echo "reference sources not found at $REF"; exit 2; }
mkdir -p "$OUT"
NVCC=${NVCC:-$(command -v nvcc || echo /usr/local/cuda/bin/nvcc)}
strip_inc() { sed -E 's/^[[:space:]]*#[[:space:]]*include.*$//' "$1"; }     # blanked, so line numbers stay the reference's
texobj() {
  sed -E \
    -e 's/^texture<float4, 3> tex_data;/shimtex::Ref tex_data; __device__ cudaTextureObject_t tex_data_obj;/' \
    -e 's/^texture<float4, 3, cudaReadModeElementType> coeffs3D;.*/shimtex::Ref coeffs3D; __device__ cudaTextureObject_t coeffs3D_obj;/' \
    -e 's/(cubicTex3D[A-Za-z_0-9]*)\(coeffs3D,/\1(coeffs3D_obj,/g' \
    -e '/^texture<float, 2> mie_scattering_irradiance;/d' \
    -e 's/tex3D\(tex_data,/tex3D<float4>(tex_data_obj,/g' \
    -e 's/cudaBindTextureToArray\(tex_data, data_array, channelDesc\)/shimtex::bind(tex_data, tex_data_obj, data_array)/' \
    -e 's/cudaUnbindTexture\((tex_data|coeffs3D)\)/shimtex::unbind(\1)/' \
    -e 's#"/scratch/[^"]*"#"/dev/null"#'
}
TMP="$(mktemp -d)"; trap 'rm -rf "$TMP"' EXIT
{
  echo "#include \"$HERE/ref_shim/cuda_texobj_patch.h\""
  echo "#include \"$HERE/ref_shim/teem_shim.h\""
  echo "namespace shim { thread_local std::string t_biff; }"
  echo '#line 1 "cutil_math_bugfixes.h"'
  strip_inc "$REF/CubicInterpolationCUDA/code/internal/cutil_math_bugfixes.h"
  echo '#line 1 "parallel_ray_tracing.h"'
  strip_inc "$REF/parallel_ray_tracing.h"
  echo '#line 1 "float3_operators.h"'
  strip_inc "$REF/float3_operators.h"
  echo "#include \"$HERE/ref_shim/tricubic_stubs_cuda.h\""
  echo '#line 1 "trace_rays_through_density_gradients.h"'
  strip_inc "$REF/trace_rays_through_density_gradients.h" | texobj
  echo '#line 1 "parallel_ray_tracing.cu"'
  strip_inc "$REF/parallel_ray_tracing.cu" | texobj
} > "$TMP/ref_cuda.cu"
"$NVCC" -gencode arch=compute_100,code=sm_100 -O3 -fmad=false -std=c++14 -w -Xcompiler -fPIC,-fpermissive -shared \
        -o "$OUT/libphoton_ref_cuda.so" "$TMP/ref_cuda.cu" -lz -lm ${PHOTON_REF_CUDA_KEEP:+--keep --keep-dir "$TMP"}
echo "built $OUT/libphoton_ref_cuda.so"
