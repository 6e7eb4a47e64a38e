/*
 * cuda_texobj_patch.h -- TEST INFRASTRUCTURE, NOT PRODUCT ("Oracle B").
 *
 * What the reference's CUDA sources need from outside /root/reference to compile with a CUDA 12 nvcc
 * (oracle/build_ref_cuda.sh streams them through sed into nvcc; nothing is copied into the repo):
 *   - the headers the stripped #include lines named (CUDA runtime, cuRAND device API, std);
 *   - helper_cuda.h's checkCudaErrors / cutil.h's CUDA_SAFE_CALL (CUDA-samples macros);
 *   - the texture-REFERENCE API that CUDA 12 removed, as the minimal mechanical patch BASELINE.json allows:
 *     the build recipe turns `texture<float4, 3> tex_data;` into a host-side settings record (shimtex::Ref,
 *     so that the reference's own `tex_data.addressMode[..] = ..; tex_data.filterMode = ..;` lines compile and
 *     are honoured) plus a __device__ cudaTextureObject_t, `tex3D(tex_data, ...)` into
 *     `tex3D<float4>(tex_data_obj, ...)`, `cudaBindTextureToArray` / `cudaUnbindTexture` into
 *     shimtex::bind / shimtex::unbind below.  The hardware that filters is the same unit either way.
 *   - teem's nrrdLoad (teem_shim.h, shared with the host build).
 * The reference's kernels, device functions and host driver are otherwise compiled as they are.
 */
#ifndef PHOTON_REF_CUDA_TEXOBJ_PATCH_H_
#define PHOTON_REF_CUDA_TEXOBJ_PATCH_H_

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <numeric>
#include <sstream>
#include <string>
#include <vector>

#include <cuda.h>
#include <cuda_runtime.h>
#include <curand.h>
#include <curand_kernel.h>
#include <vector_functions.h>
#include <vector_types.h>

typedef unsigned int uint;

namespace shim {
inline cudaError_t check(cudaError_t e, const char *what, int line) {
    if (e != cudaSuccess) fprintf(stderr, "photon_ref_cuda: %s -> %s (line %d)\n", what, cudaGetErrorString(e), line);
    return e;
}
}  // namespace shim
#define checkCudaErrors(x) shim::check((x), #x, __LINE__)
#define CUDA_SAFE_CALL(x) shim::check((x), #x, __LINE__)

namespace shimtex {
// the fields of the old `struct textureReference` that the reference sets (trace.h:1628-1631); defaults as the
// CUDA 10 `texture<>` constructor had them: not normalised, point filter, clamp on every axis
struct Ref {
    int normalized = 0;
    cudaTextureFilterMode filterMode = cudaFilterModePoint;
    cudaTextureAddressMode addressMode[3] = {cudaAddressModeClamp, cudaAddressModeClamp, cudaAddressModeClamp};
    cudaTextureObject_t obj = 0;
};
template <class Symbol>
inline cudaError_t bind(Ref &r, const Symbol &device_symbol, cudaArray *arr) {
    cudaResourceDesc rd; memset(&rd, 0, sizeof rd);
    rd.resType = cudaResourceTypeArray; rd.res.array.array = arr;
    cudaTextureDesc td; memset(&td, 0, sizeof td);
    for (int i = 0; i < 3; i++) td.addressMode[i] = r.addressMode[i];
    td.filterMode = r.filterMode; td.readMode = cudaReadModeElementType; td.normalizedCoords = r.normalized;
    if (r.obj) { cudaDestroyTextureObject(r.obj); r.obj = 0; }
    cudaError_t e = cudaCreateTextureObject(&r.obj, &rd, &td, nullptr);
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbol(device_symbol, &r.obj, sizeof r.obj);
}
inline cudaError_t unbind(Ref &r) {
    cudaError_t e = r.obj ? cudaDestroyTextureObject(r.obj) : cudaSuccess;
    r.obj = 0;
    return e;
}
}  // namespace shimtex
#endif
