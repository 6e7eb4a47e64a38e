// Microbenchmark: FP32 throughput of mixes of scalar FFMA and packed FFMA2 (and IMAD) on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp32_mix_probe tools/fp32_mix_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int N2, int N1, int NI>
__global__ void __launch_bounds__(256) mix(float *out, int iters, float a, float b, int ia) {
    float t = threadIdx.x * 1e-3f;
    float2 p[N2 > 0 ? N2 : 1]; float s[N1 > 0 ? N1 : 1]; int q[NI > 0 ? NI : 1];
#pragma unroll
    for (int i = 0; i < N2; i++) p[i] = make_float2(t + i, t - i);
#pragma unroll
    for (int i = 0; i < N1; i++) s[i] = t + 0.5f * i;
#pragma unroll
    for (int i = 0; i < NI; i++) q[i] = threadIdx.x + i;
    const float2 aa = make_float2(a, a), bb = make_float2(b, b);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int i = 0; i < N2; i++) p[i] = __ffma2_rn(p[i], aa, bb);
#pragma unroll
            for (int i = 0; i < N1; i++) s[i] = fmaf(s[i], a, b);
#pragma unroll
            for (int i = 0; i < NI; i++) q[i] = q[i] * ia + 7;
        }
    }
    float r = 0; for (int i = 0; i < N2; i++) r += p[i].x + p[i].y; for (int i = 0; i < N1; i++) r += s[i]; for (int i = 0; i < NI; i++) r += q[i];
    if (r == 123.456f) out[0] = r;
}
template <int N2, int N1, int NI> void run(const char *name) {
    float *d; cudaMalloc(&d, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 4096, blocks = 148 * 8; double best = 1e30;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0); mix<N2, N1, NI><<<blocks, 256>>>(d, iters, 0.999f, 0.001f, 3); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (r > 0 && ms < best) best = ms;
    }
    double inst = (double)(N2 + N1 + NI) * 8 * iters * 256.0 / 32 * blocks;      // warp instructions
    double fma = (double)(2 * N2 + N1) * 8 * iters * 256.0 * blocks;
    double cyc = best * 1e-3 * 1.965e9 * 148 * 4;                                 // SMSP-cycles
    printf("%-28s %7.3f ms  %6.1f TFLOP/s fp32  %.3f warp-inst/clk/SMSP\n", name, best, 2 * fma / (best * 1e-3) / 1e12, inst / cyc);
    cudaFree(d);
}
int main() {
    run<0, 16, 0>("16 FFMA");
    run<8, 0, 0>("8 FFMA2");
    run<8, 8, 0>("8 FFMA2 + 8 FFMA");
    run<8, 4, 0>("8 FFMA2 + 4 FFMA");
    run<4, 8, 0>("4 FFMA2 + 8 FFMA");
    run<4, 12, 0>("4 FFMA2 + 12 FFMA");
    run<0, 8, 8>("8 FFMA + 8 IMAD");
    run<0, 12, 4>("12 FFMA + 4 IMAD");
    run<8, 0, 4>("8 FFMA2 + 4 IMAD");
    run<0, 0, 16>("16 IMAD");
    return 0;
}
