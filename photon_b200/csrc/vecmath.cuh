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
#define PH_COLD __host__ __device__ __noinline__     // rare paths kept out of the hot loop's code
#else
#define PH_HD inline
#define PH_D inline
#define PH_COLD inline
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

PH_HD float mad_rn(float a, float b, float c) { return add_rn(mul_rn(a, b), c); }   // a*b + c, two roundings

// f3: every multiply and add is rounded on its own (no FMA contraction), whatever the compiler
// flags.  The per-ray stages -- generation, frame rotations, lens, sensor -- work in float32 at
// |coordinates| ~ 1e6 um where sums cancel catastrophically (ray-sphere quadratic
// prt.cu:271-276, plane hits), and the reference's debug build (-G -O0) does not contract.
// The library is also compiled with -fmad=false so that the scalar expressions of those
// stages stay unfused; the integrator loop asks for its fused multiply-adds explicitly
// through ph::fma_ (below), which is the only place where FMA is wanted.
struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };

PH_HD f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
PH_HD f3 operator+(f3 a, f3 b) { return mk3(add_rn(a.x, b.x), add_rn(a.y, b.y), add_rn(a.z, b.z)); }
PH_HD f3 operator-(f3 a, f3 b) { return mk3(add_rn(a.x, -b.x), add_rn(a.y, -b.y), add_rn(a.z, -b.z)); }
PH_HD f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
PH_HD f3 operator*(f3 a, float s) { return mk3(mul_rn(a.x, s), mul_rn(a.y, s), mul_rn(a.z, s)); }
PH_HD f3 operator*(float s, f3 a) { return mk3(mul_rn(a.x, s), mul_rn(a.y, s), mul_rn(a.z, s)); }
PH_HD f3 operator/(f3 a, float s) { float inv = 1.0f / s; return a * inv; }
PH_HD float dot(f3 a, f3 b) { return add_rn(add_rn(mul_rn(a.x, b.x), mul_rn(a.y, b.y)), mul_rn(a.z, b.z)); }
PH_HD f3 normalize(f3 v) { float inv = 1.0f / sqrtf(dot(v, v)); return v * inv; }
PH_HD float dot_rn(f3 a, f3 b) { return dot(a, b); }

// Integrator-loop arithmetic.  fma_(a,b,c) = a*b+c: one FFMA on the GPU.  With PH_NO_FMA
// (tests/hostcheck) it is the unfused product and sum, which reproduces the reference's
// operation order bit for bit on the CPU; the fused form differs by <= 1/2 ulp per operation.
PH_HD float fma_(float a, float b, float c) {
#if defined(PH_NO_FMA)
    return add_rn(mul_rn(a, b), c);
#else
    return fmaf(a, b, c);
#endif
}
PH_HD f3 axpy(float s, f3 v, f3 w) { return mk3(fma_(s, v.x, w.x), fma_(s, v.y, w.y), fma_(s, v.z, w.z)); }   // s*v + w
PH_HD f3 scale3(float s, f3 v) { return mk3(s * v.x, s * v.y, s * v.z); }
// floor(x) as float plus the integer, for |x| < 2^22, without the conversion pipe: adding
// 1.5*2^23 with round-down leaves floor(x) in the low mantissa bits (FADD.RM + FADD + IADD).
PH_HD float floor_split(float x, int &i) {
#if defined(__CUDA_ARCH__)
    float t = __fadd_rd(x, 12582912.0f);
    i = __float_as_int(t) - 0x4B400000;
    return __fadd_rn(t, -12582912.0f);
#else
    float f = floorf(x); i = (int)f; return f;
#endif
}
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
