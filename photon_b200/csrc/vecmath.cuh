// vecmath.cuh -- small float3 layer for the ray kernels.
//
// The operation order is the one the reference's operator layer fixes
// (CubicInterpolationCUDA/code/internal/cutil_math_bugfixes.h:302-409: a/s = a*(1/s),
// normalize(v) = v*(1/sqrt(dot)), dot = x*x+y*y+z*z left to right), because sensor
// positions are compared to 1e-3 px.  FMA contraction is allowed only inside the integrator
// loop (type q3 below); everything else rounds each operation on its own (type f3).
//
// Everything is __host__ __device__ so tests/hostcheck can run the same code on the CPU.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define PH_HD __host__ __device__ __forceinline__
#define PH_D __device__ __forceinline__
#else
#define PH_HD inline
#define PH_D inline
#endif

namespace ph {

// Exactly rounded single operations that the compiler may not fuse into an FMA.
PH_HD float mul_rn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    volatile float r = a * b; return r;
#endif
}
PH_HD float add_rn(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    volatile float r = a + b; return r;
#endif
}

// Two flavours of the same 3-vector:
//   f3 = v3<1>  every multiply and add is rounded on its own (no FMA contraction).  Used by the
//               per-ray stages -- generation, frame rotations, lens, sensor -- whose float32
//               arithmetic at |coordinates| ~ 1e6 um cancels catastrophically (ray-sphere
//               quadratic prt.cu:271-276, plane hits), so that a fused multiply-add moves the
//               sensor hit by more than the 1e-3 px parity tolerance.  The reference's debug
//               build (-G -O0) does not contract either.
//   q3 = v3<0>  plain operators that nvcc contracts into FFMA.  Used only inside the integrator
//               loop (rk4_march / euler_march and the trilinear fetch), where the FP32 issue
//               rate is the bound and the contraction error is ~1e-7 relative per step.
template <int STRICT> struct v3 { float x, y, z; };
typedef v3<1> f3;
typedef v3<0> q3;
struct f4 { float x, y, z, w; };

template <int S> PH_HD float pmul(float a, float b) { return S ? mul_rn(a, b) : a * b; }
template <int S> PH_HD float padd(float a, float b) { return S ? add_rn(a, b) : a + b; }

template <int S> PH_HD v3<S> mkv(float x, float y, float z) { v3<S> r; r.x = x; r.y = y; r.z = z; return r; }
PH_HD f3 mk3(float x, float y, float z) { return mkv<1>(x, y, z); }
PH_HD q3 relax(f3 a) { return mkv<0>(a.x, a.y, a.z); }
PH_HD f3 strict(q3 a) { return mkv<1>(a.x, a.y, a.z); }
template <int S> PH_HD v3<S> operator+(v3<S> a, v3<S> b) { return mkv<S>(padd<S>(a.x, b.x), padd<S>(a.y, b.y), padd<S>(a.z, b.z)); }
template <int S> PH_HD v3<S> operator-(v3<S> a, v3<S> b) { return mkv<S>(padd<S>(a.x, -b.x), padd<S>(a.y, -b.y), padd<S>(a.z, -b.z)); }
template <int S> PH_HD v3<S> operator-(v3<S> a) { return mkv<S>(-a.x, -a.y, -a.z); }
template <int S> PH_HD v3<S> operator*(v3<S> a, float s) { return mkv<S>(pmul<S>(a.x, s), pmul<S>(a.y, s), pmul<S>(a.z, s)); }
template <int S> PH_HD v3<S> operator*(float s, v3<S> a) { return mkv<S>(pmul<S>(a.x, s), pmul<S>(a.y, s), pmul<S>(a.z, s)); }
template <int S> PH_HD v3<S> operator/(v3<S> a, float s) { float inv = 1.0f / s; return a * inv; }
template <int S> PH_HD float dot(v3<S> a, v3<S> b) { return padd<S>(padd<S>(pmul<S>(a.x, b.x), pmul<S>(a.y, b.y)), pmul<S>(a.z, b.z)); }
template <int S> PH_HD v3<S> normalize(v3<S> v) { float inv = 1.0f / sqrtf(dot(v, v)); return v * inv; }
PH_HD float dot_rn(f3 a, f3 b) { return dot(a, b); }
PH_HD float qnanf() {
#if defined(__CUDA_ARCH__)
    return __int_as_float(0x7fffffff);
#else
    return nanf("");
#endif
}
PH_HD f3 nan3() { float n = qnanf(); return mk3(n, n, n); }
PH_HD bool is_nan(float a) { return a != a; }
PH_HD bool any_nan(f3 a) { return is_nan(a.x) || is_nan(a.y) || is_nan(a.z); }
// row-major 3x3 times vector, one dot per row (prt.cu:2061-2076)
PH_HD f3 mat_mul(const float *m, f3 v) {
    return mk3(dot(mk3(m[0], m[1], m[2]), v), dot(mk3(m[3], m[4], m[5]), v), dot(mk3(m[6], m[7], m[8]), v));
}

// ---------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11).  Counter = (ray index lo, hi, stream, 0),
// key = (seed, domain tag).  Used for PHOTON_RNG=philox lens samples and for the
// position / gradient noise that the reference draws from time-seeded XORWOW.
// ---------------------------------------------------------------------------
PH_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}
PH_HD void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t h0 = mulhi32(M0, c[0]), l0 = M0 * c[0];
        uint32_t h1 = mulhi32(M1, c[2]), l1 = M1 * c[2];
        uint32_t n0 = h1 ^ c[1] ^ k0, n2 = h0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = l1; c[2] = n2; c[3] = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
PH_HD float u01_24(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

struct NormalStream {           // counter-based N(0,1) pairs per ray
    uint64_t ray; uint32_t ctr; uint32_t seed;
    PH_HD void next2(float &a, float &b) {
        uint32_t c[4] = {(uint32_t)ray, (uint32_t)(ray >> 32), ctr++, 0x6e6f6973u};
        philox4x32_10(c, seed, 0x70686f74u);
        float u = ((float)(c[0] >> 8) + 0.5f) * (1.0f / 16777216.0f);
        float v = ((float)(c[1] >> 8) + 0.5f) * (1.0f / 16777216.0f);
        float r = sqrtf(-2.0f * logf(u));
        a = r * cosf(6.2831853f * v); b = r * sinf(6.2831853f * v);
    }
};

}  // namespace ph
