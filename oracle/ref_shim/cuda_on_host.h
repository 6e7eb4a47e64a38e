/*
 * cuda_on_host.h -- TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * A minimal emulation of the CUDA pieces that photon's reference sources use, so that
 * cuda_codes/parallel_ray_tracing.cu and trace_rays_through_density_gradients.h can be
 * compiled by g++ FROM WHERE THEY LIE under /root/reference (oracle/build_ref.sh) and run
 * on the host cores: one CUDA thread = one loop iteration, one CUDA block = one OpenMP task.
 * The reference's device code (ray generation, RK4/Euler integration, lens, sensor, splat)
 * and its host driver (start_ray_tracing, loadNRRD, setData, Host_Init) run UNMODIFIED.
 * Emulated here, because they live outside /root/reference:
 *   - the CUDA runtime calls the driver makes (cudaMalloc/cudaMemcpy/... -> malloc/memcpy);
 *   - kernel launches (<<<grid,block>>> is rewritten to shim::launch by the build recipe);
 *   - the texture unit: tex3D on a float4 3-D array, unnormalised coordinates, clamp
 *     addressing, linear filtering as the CUDA programming guide states it
 *     (texel centres at +0.5; PHOTON_REF_TEXFILTER=fixed8 quantises the fractions to 8 bits
 *     like the hardware, default = exact float weights);
 *   - cuRAND (curand_init / curand_normal2 / curand_uniform): a counter-based generator
 *     (the reference seeds with time(NULL); only statistical parity is possible);
 *   - teem's nrrdLoad (teem/nrrd.h is a separate shim).
 */
#ifndef PHOTON_REF_CUDA_ON_HOST_H_
#define PHOTON_REF_CUDA_ON_HOST_H_

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <chrono>
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
#ifdef _OPENMP
#include <omp.h>
#endif

#define __device__
#define __global__
#define __host__
#define __constant__
#define __shared__ static thread_local
#define __forceinline__ inline
#define __restrict__

/* ---- vector types -------------------------------------------------------------- */
struct float2 { float x, y; };
struct float3 { float x, y, z; };
struct float4 { float x, y, z, w; };
struct double2 { double x, y; };
struct double3 { double x, y, z; };
struct double4 { double x, y, z, w; };
struct int2 { int x, y; };
struct int3 { int x, y, z; };
struct int4 { int x, y, z, w; };
struct uint2 { unsigned int x, y; };
struct uint3 { unsigned int x, y, z; };
struct uint4 { unsigned int x, y, z, w; };
inline float2 make_float2(float x, float y) { return float2{x, y}; }
inline float3 make_float3(float x, float y, float z) { return float3{x, y, z}; }
inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
inline double4 make_double4(double x, double y, double z, double w) { return double4{x, y, z, w}; }
inline int2 make_int2(int x, int y) { return int2{x, y}; }
inline int3 make_int3(int x, int y, int z) { return int3{x, y, z}; }
inline int4 make_int4(int x, int y, int z, int w) { return int4{x, y, z, w}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint3 make_uint3(unsigned x, unsigned y, unsigned z) { return uint3{x, y, z}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }

struct dim3 {
    unsigned int x, y, z;
    dim3(unsigned int x_ = 1, unsigned int y_ = 1, unsigned int z_ = 1) : x(x_), y(y_), z(z_) {}
};
extern thread_local uint3 threadIdx, blockIdx;
extern thread_local dim3 blockDim, gridDim;
inline void __syncthreads() {}

/* ---- device intrinsics and the math overloads CUDA puts in the global namespace -- */
inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
inline double __longlong_as_double(long long i) { double d; memcpy(&d, &i, 8); return d; }
inline int max(int a, int b) { return a > b ? a : b; }
inline int min(int a, int b) { return a < b ? a : b; }
inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
inline float atomicAdd(float *addr, float v) {
    float old;
#pragma omp atomic capture
    { old = *addr; *addr += v; }
    return old;
}

/* ---- runtime ---------------------------------------------------------------------- */
typedef int cudaError_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
inline const char *cudaGetErrorString(cudaError_t) { return "cuda_on_host"; }
inline cudaError_t cudaGetLastError() { return cudaSuccess; }
inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
inline cudaError_t cudaDeviceReset() { return cudaSuccess; }
/* Every "device" allocation gets a 64 KiB zeroed guard zone in front: the reference's bilinear
 * splat writes image[(ii-1)*nx + jj-1] with ii or jj == 0 (parallel_ray_tracing.cu:2223-2228),
 * up to nx+1 floats BEFORE the image.  On a GPU that lands in a neighbouring allocation
 * unnoticed; on the host it would corrupt the malloc heap. */
enum { SHIM_GUARD = 65536 };
template <class T> inline cudaError_t cudaMalloc(T **p, size_t n) {
    char *raw = (char *)calloc(1, SHIM_GUARD + (n ? n : 1));
    *p = raw ? (T *)(raw + SHIM_GUARD) : nullptr;
    return raw ? cudaSuccess : 2;
}
inline cudaError_t cudaFree(void *p) { if (p) free((char *)p - SHIM_GUARD); return cudaSuccess; }
inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { if (n) memcpy(d, s, n); return cudaSuccess; }
inline cudaError_t cudaMemset(void *d, int v, size_t n) { memset(d, v, n); return cudaSuccess; }
#define checkCudaErrors(x) do { (void)(x); } while (0)
#define CUDA_SAFE_CALL(x) do { (void)(x); } while (0)

struct cudaExtent { size_t width, height, depth; };
inline cudaExtent make_cudaExtent(size_t w, size_t h, size_t d) { return cudaExtent{w, h, d}; }
struct cudaPitchedPtr { void *ptr; size_t pitch, xsize, ysize; };
inline cudaPitchedPtr make_cudaPitchedPtr(void *p, size_t pitch, size_t xs, size_t ys) { return cudaPitchedPtr{p, pitch, xs, ys}; }
struct cudaChannelFormatDesc { int x, y, z, w; };
template <class T> inline cudaChannelFormatDesc cudaCreateChannelDesc() { return cudaChannelFormatDesc{32, 32, 32, 32}; }
struct cudaArray { float4 *texels; size_t w, h, d; };
typedef cudaArray *cudaArray_t;
inline cudaError_t cudaMalloc3DArray(cudaArray **a, const cudaChannelFormatDesc *, cudaExtent e) {
    *a = new cudaArray{(float4 *)malloc(sizeof(float4) * e.width * e.height * e.depth), e.width, e.height, e.depth};
    return cudaSuccess;
}
inline cudaError_t cudaFreeArray(cudaArray *a) { if (a) { free(a->texels); delete a; } return cudaSuccess; }
struct cudaMemcpy3DParms {
    cudaArray *srcArray; cudaPitchedPtr srcPtr; cudaArray *dstArray; cudaPitchedPtr dstPtr; cudaExtent extent; cudaMemcpyKind kind;
};
inline cudaError_t cudaMemcpy3D(const cudaMemcpy3DParms *p) {
    const size_t W = p->extent.width, H = p->extent.height, D = p->extent.depth;
    for (size_t z = 0; z < D; z++)
        for (size_t y = 0; y < H; y++)
            memcpy(p->dstArray->texels + (z * H + y) * W, (const char *)p->srcPtr.ptr + (z * p->srcPtr.ysize + y) * p->srcPtr.pitch,
                   W * sizeof(float4));
    return cudaSuccess;
}

/* ---- texture references ------------------------------------------------------------ */
enum cudaTextureReadMode { cudaReadModeElementType, cudaReadModeNormalizedFloat };
enum cudaTextureFilterMode { cudaFilterModePoint, cudaFilterModeLinear };
enum cudaTextureAddressMode { cudaAddressModeWrap, cudaAddressModeClamp, cudaAddressModeMirror, cudaAddressModeBorder };
template <class T, int DIM = 1, cudaTextureReadMode MODE = cudaReadModeElementType>
struct texture {
    int normalized = 0;
    cudaTextureFilterMode filterMode = cudaFilterModePoint;
    cudaTextureAddressMode addressMode[3] = {cudaAddressModeClamp, cudaAddressModeClamp, cudaAddressModeClamp};
    const cudaArray *bound = nullptr;
};
template <class T, int DIM, cudaTextureReadMode MODE>
inline cudaError_t cudaBindTextureToArray(texture<T, DIM, MODE> &t, const cudaArray *a, const cudaChannelFormatDesc &) { t.bound = a; return cudaSuccess; }
template <class T, int DIM, cudaTextureReadMode MODE>
inline cudaError_t cudaUnbindTexture(texture<T, DIM, MODE> &t) { t.bound = nullptr; return cudaSuccess; }

namespace shim {
/* where setData's debug dump of the gradient field goes (the reference hard-codes a /scratch path, trace.h:1840) */
inline std::string gradient_dump_path() { const char *p = getenv("PHOTON_REF_GRADIENT_DUMP"); return p && *p ? p : "/dev/null"; }
extern int g_tex_fixed8;                       /* PHOTON_REF_TEXFILTER=fixed8 */
extern thread_local long long t_tex_fetches;   /* per host thread, summed after each launch */
inline int clampi(int v, int n) { return v < 0 ? 0 : (v > n - 1 ? n - 1 : v); }
inline float quant8(float f) { return g_tex_fixed8 ? floorf(f * 256.0f + 0.5f) * (1.0f / 256.0f) : f; }
}  // namespace shim

/* tex(x,y,z) = sum over the 8 neighbours of T[i+di, j+dj, k+dk] * wx * wy * wz with
 * xB = x - 0.5, i = floor(xB), alpha = frac(xB) (CUDA C Programming Guide, "Linear Filtering"). */
inline float4 tex3D(const texture<float4, 3, cudaReadModeElementType> &t, float x, float y, float z) {
    const cudaArray *A = t.bound;
    shim::t_tex_fetches++;
    const int W = (int)A->w, H = (int)A->h, D = (int)A->d;
    float xb = x - 0.5f, yb = y - 0.5f, zb = z - 0.5f;
    float fx = floorf(xb), fy = floorf(yb), fz = floorf(zb);
    float a = shim::quant8(xb - fx), b = shim::quant8(yb - fy), c = shim::quant8(zb - fz);
    int i = (int)fx, j = (int)fy, k = (int)fz;
    int i0 = shim::clampi(i, W), i1 = shim::clampi(i + 1, W), j0 = shim::clampi(j, H), j1 = shim::clampi(j + 1, H);
    int k0 = shim::clampi(k, D), k1 = shim::clampi(k + 1, D);
    const float4 *T = A->texels;
    const size_t sW = (size_t)W, sWH = (size_t)W * H;
    const float4 &c000 = T[k0 * sWH + j0 * sW + i0], &c100 = T[k0 * sWH + j0 * sW + i1];
    const float4 &c010 = T[k0 * sWH + j1 * sW + i0], &c110 = T[k0 * sWH + j1 * sW + i1];
    const float4 &c001 = T[k1 * sWH + j0 * sW + i0], &c101 = T[k1 * sWH + j0 * sW + i1];
    const float4 &c011 = T[k1 * sWH + j1 * sW + i0], &c111 = T[k1 * sWH + j1 * sW + i1];
    float w000 = (1 - a) * (1 - b) * (1 - c), w100 = a * (1 - b) * (1 - c), w010 = (1 - a) * b * (1 - c), w110 = a * b * (1 - c);
    float w001 = (1 - a) * (1 - b) * c, w101 = a * (1 - b) * c, w011 = (1 - a) * b * c, w111 = a * b * c;
    float4 o;
    o.x = w000 * c000.x + w100 * c100.x + w010 * c010.x + w110 * c110.x + w001 * c001.x + w101 * c101.x + w011 * c011.x + w111 * c111.x;
    o.y = w000 * c000.y + w100 * c100.y + w010 * c010.y + w110 * c110.y + w001 * c001.y + w101 * c101.y + w011 * c011.y + w111 * c111.y;
    o.z = w000 * c000.z + w100 * c100.z + w010 * c010.z + w110 * c110.z + w001 * c001.z + w101 * c101.z + w011 * c011.z + w111 * c111.z;
    o.w = w000 * c000.w + w100 * c100.w + w010 * c010.w + w110 * c110.w + w001 * c001.w + w101 * c101.w + w011 * c011.w + w111 * c111.w;
    return o;
}
template <class T> inline T tex2D(const texture<T, 2, cudaReadModeElementType> &, float, float) { return T(); }

/* ---- cuRAND ---------------------------------------------------------------------------- */
struct curandStateXORWOW { unsigned long long key, ctr; };
typedef curandStateXORWOW curandState;
typedef curandStateXORWOW curandState_t;
namespace shim {
inline unsigned long long mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ULL; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; return z ^ (z >> 31);
}
}
inline void curand_init(unsigned long long seed, unsigned long long seq, unsigned long long off, curandState *s) {
    s->key = shim::mix64(seed ^ shim::mix64(seq)); s->ctr = off;
}
inline float curand_uniform(curandState *s) {
    unsigned long long r = shim::mix64(s->key + (s->ctr++) * 0xD1342543DE82EF95ULL);
    return ((float)(r >> 40) + 1.0f) * (1.0f / 16777216.0f);
}
inline float2 curand_normal2(curandState *s) {
    float u = curand_uniform(s), v = curand_uniform(s);
    float r = sqrtf(-2.0f * logf(u));
    return float2{r * cosf(6.2831853f * v), r * sinf(6.2831853f * v)};
}

/* ---- kernel launch ----------------------------------------------------------------------- */
namespace shim {
extern double g_kernel_seconds;                /* time inside the ray-tracing launches of the last call */
extern long long g_tex_fetches;                /* tex3D fetches of the last call */
extern long long g_threads_launched;
extern int g_host_threads;
template <class K, class... A>
inline void launch(dim3 grid, dim3 block, K kernel, A... args) {
    auto t0 = std::chrono::steady_clock::now();
    const long long nblocks = (long long)grid.x * grid.y * grid.z;
    long long fetches = 0;
#pragma omp parallel reduction(+ : fetches)
    {
#ifdef _OPENMP
#pragma omp single
        g_host_threads = omp_get_num_threads();
#endif
        t_tex_fetches = 0;
        gridDim = grid; blockDim = block;
#pragma omp for schedule(dynamic, 1)
        for (long long b = 0; b < nblocks; b++) {
            blockIdx.x = (unsigned)(b % grid.x); blockIdx.y = (unsigned)((b / grid.x) % grid.y); blockIdx.z = (unsigned)(b / ((long long)grid.x * grid.y));
            for (unsigned tz = 0; tz < block.z; tz++)
                for (unsigned ty = 0; ty < block.y; ty++)
                    for (unsigned tx = 0; tx < block.x; tx++) {      /* thread 0 first: it fills __shared__ */
                        threadIdx.x = tx; threadIdx.y = ty; threadIdx.z = tz;
                        kernel(args...);
                    }
        }
        fetches += t_tex_fetches;
    }
    g_tex_fetches += fetches;
    g_threads_launched += nblocks * block.x * block.y * block.z;
    g_kernel_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
}  // namespace shim

#endif /* PHOTON_REF_CUDA_ON_HOST_H_ */
