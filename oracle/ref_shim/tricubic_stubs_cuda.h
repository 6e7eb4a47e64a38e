/*
 * tricubic_stubs_cuda.h -- TEST INFRASTRUCTURE.  The CUDA-build twin of tricubic_stubs.h: photon forces
 * interpolation_scheme = 1 (parallel_ray_tracing.cu:3330), so the CubicInterpolationCUDA entry points that
 * trace_rays_through_density_gradients.h names are never executed; they only have to exist.
 */
#ifndef PHOTON_REF_TRICUBIC_STUBS_CUDA_H_
#define PHOTON_REF_TRICUBIC_STUBS_CUDA_H_
namespace shim { inline void dead(const char *what) { fprintf(stderr, "photon_ref_cuda: %s reached (dead path in the reference)\n", what); abort(); } }
template <class TEX> __device__ inline float4 cubicTex3D(const TEX &, float3) { return make_float4(0, 0, 0, 0); }
template <class TEX> __device__ inline float4 cubicTex3D_1st_derivative_x(const TEX &, float3) { return make_float4(0, 0, 0, 0); }
template <class TEX> __device__ inline float4 cubicTex3D_1st_derivative_y(const TEX &, float3) { return make_float4(0, 0, 0, 0); }
template <class TEX> __device__ inline float4 cubicTex3D_1st_derivative_z(const TEX &, float3) { return make_float4(0, 0, 0, 0); }
template <class T> inline cudaPitchedPtr CopyVolumeHostToDevice(const T *, uint, uint, uint) { shim::dead("CopyVolumeHostToDevice"); return cudaPitchedPtr(); }
template <class T> inline void CubicBSplinePrefilter3DTimer(T *, uint, uint, uint, uint) { shim::dead("CubicBSplinePrefilter3DTimer"); }
template <class T> inline void CreateTextureFromVolume(T *, cudaArray **, cudaPitchedPtr, cudaExtent, bool) { shim::dead("CreateTextureFromVolume"); }
#endif
