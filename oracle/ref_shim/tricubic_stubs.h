/*
 * tricubic_stubs.h -- TEST INFRASTRUCTURE.  photon forces interpolation_scheme = 1 (trilinear,
 * parallel_ray_tracing.cu:3330), so the CubicInterpolationCUDA entry points that
 * trace_rays_through_density_gradients.h names (:290-298, 912, 1216-1262, 1654-1655) are never
 * executed.  They only have to exist for the file to compile; reaching one aborts.
 */
#ifndef PHOTON_REF_TRICUBIC_STUBS_H_
#define PHOTON_REF_TRICUBIC_STUBS_H_
namespace shim { [[noreturn]] inline void dead(const char *what) { fprintf(stderr, "photon_ref_cpu: %s reached (dead path in the reference)\n", what); abort(); } }
template <class TEX> inline float4 cubicTex3D(const TEX &, float3) { shim::dead("cubicTex3D"); }
template <class TEX> inline float4 cubicTex3D_1st_derivative_x(const TEX &, float3) { shim::dead("cubicTex3D_1st_derivative_x"); }
template <class TEX> inline float4 cubicTex3D_1st_derivative_y(const TEX &, float3) { shim::dead("cubicTex3D_1st_derivative_y"); }
template <class TEX> inline float4 cubicTex3D_1st_derivative_z(const TEX &, float3) { shim::dead("cubicTex3D_1st_derivative_z"); }
template <class T> inline cudaPitchedPtr CopyVolumeHostToDevice(const T *, uint, uint, uint) { shim::dead("CopyVolumeHostToDevice"); }
template <class T> inline void CubicBSplinePrefilter3DTimer(T *, uint, uint, uint, uint) { shim::dead("CubicBSplinePrefilter3DTimer"); }
template <class T> inline void CreateTextureFromVolume(T *, cudaArray **, cudaPitchedPtr, cudaExtent, bool) { shim::dead("CreateTextureFromVolume"); }
#endif
