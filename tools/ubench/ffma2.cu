// Microbenchmark: does fma.rn.f32x2 (FFMA2) save issue slots on sm_100a when mixed with ALU work?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2 ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    ra = *reinterpret_cast<unsigned long long*>(&a); rb = *reinterpret_cast<unsigned long long*>(&b); rc = *reinterpret_cast<unsigned long long*>(&c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
template <int MODE> __global__ void k(float* out, int iters, float s, unsigned m) {
    float a[8]; unsigned u[8];
    for (int i = 0; i < 8; ++i) { a[i] = threadIdx.x * 0.001f + i; u[i] = threadIdx.x + i; }
    float2 p[4]; for (int i = 0; i < 4; ++i) p[i] = make_float2(a[2 * i], a[2 * i + 1]);
    float2 s2 = make_float2(s, s * 1.0001f), c2 = make_float2(0.5f, 0.25f);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], s, 0.5f + i);
        }
        if (MODE == 1 || MODE == 3) {
#pragma unroll
            for (int i = 0; i < 4; ++i) p[i] = fma2(p[i], s2, c2);
        }
        if (MODE >= 2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) u[i] = (u[i] ^ m) + (u[i] >> 3);
        }
    }
    float r = 0; for (int i = 0; i < 8; ++i) r += a[i] + u[i]; for (int i = 0; i < 4; ++i) r += p[i].x + p[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> __global__ void mix(float* out, int iters, float s, unsigned m) {
    float a[8]; unsigned u[4];
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 4; ++i) u[i] = threadIdx.x + i;
    float2 p[4]; for (int i = 0; i < 4; ++i) p[i] = make_float2(a[2 * i], a[2 * i + 1]);
    float2 s2 = make_float2(s, s), c2 = make_float2(0.5f, 0.5f);
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], s, 0.5f);
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) p[i] = fma2(p[i], s2, c2);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) u[i] = __funnelshift_l(u[i], u[i], 5) ^ (m + it);   // SHF + LOP3 per element
    }
    float r = 0; for (int i = 0; i < 8; ++i) r += a[i]; for (int i = 0; i < 4; ++i) r += p[i].x + p[i].y + u[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void runmix(const char* name, int warps_per_smsp) {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 200000, threads = warps_per_smsp * 4 * 32;
    mix<MODE><<<148, threads>>>(d, 1000, 0.999f, 0x5bd1e995u);
    cudaEventRecord(e0); mix<MODE><<<148, threads>>>(d, iters, 0.999f, 0x5bd1e995u); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s warps/SMSP %d: %.2f cycles per warp-iteration / warps (SMSP cycles per iteration of one warp)\n", name, warps_per_smsp,
           ms * 1e-3 * 1.965e9 / iters / warps_per_smsp);
    cudaFree(d);
}
template <int MODE> void run(const char* name) {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 200000;
    k<MODE><<<148 * 8, 256>>>(d, 1000, 0.999f, 0x5bd1e995u);
    cudaEventRecord(e0); k<MODE><<<148 * 8, 256>>>(d, iters, 0.999f, 0x5bd1e995u); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fma = 8.0 * iters * 148 * 8 * 256;
    printf("%-28s %8.3f ms  %.2f TFLOP/s (fma flops only)  cycles/iter/SMSP at 1.965GHz: %.2f\n", name, ms, 2 * fma / ms * 1e-9,
           ms * 1e-3 * 1.965e9 / iters / 16.0);   // 8 blocks*8 warps/SM = 64 warps / 4 SMSP = 16 warps per SMSP
    cudaFree(d);
}
template <int MODE> __global__ void lat(float* out, int iters, float s) {
    float a = threadIdx.x; float2 p = make_float2(a, a + 1), s2 = make_float2(s, s), c2 = make_float2(0.5f, 0.25f);
    float m = s;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (MODE == 0) a = fmaf(a, m, a);
            if (MODE == 1) p = fma2(p, s2, p);
            if (MODE == 2) { a = fmaf(a, m, a); a = __int_as_float(__float_as_int(a) ^ 0x100); }   // FFMA -> LOP3 cross-pipe
            if (MODE == 3) { a = a * m; }                                                         // FMUL chain
            if (MODE == 4) { a = __sinf(a); }                                                     // FMUL + MUFU chain
        }
    }
    out[threadIdx.x] = a + p.x + p.y;
}
template <int MODE> void runlat(const char* name, int per) {
    float* d; cudaMalloc(&d, 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 100000;
    lat<MODE><<<1, 32>>>(d, 100, 0.999f);
    cudaEventRecord(e0); lat<MODE><<<1, 32>>>(d, iters, 0.999f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s %.2f cycles per dependent op (at 1.965 GHz)\n", name, ms * 1e-3 * 1.965e9 / iters / 16.0 / per);
    cudaFree(d);
}
int main() {
    for (int w = 1; w <= 4; ++w) { runmix<0>("8 FFMA + 4x(ALU)", w); runmix<1>("4 FFMA2 + 4x(ALU)", w); }
    runlat<0>("lat FFMA", 1); runlat<1>("lat FFMA2", 1); runlat<2>("lat FFMA+LOP3", 2); runlat<3>("lat FMUL", 1); runlat<4>("lat FMUL+MUFU.SIN", 1);
    run<0>("8 FFMA"); run<1>("4 FFMA2"); run<2>("8 FFMA + 16 ALU"); run<3>("4 FFMA2 + 16 ALU");
    return 0;
}
