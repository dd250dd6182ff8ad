// Register-read bandwidth of FFMA / FFMA2 operand forms on sm_100a: cycles per instruction per SMSP with 4 warps per SMSP.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
template <int MODE> __global__ void k(float* out, const float* in, int iters) {
    float a[8], b[8], c[8];
    for (int i = 0; i < 8; ++i) { a[i] = in[threadIdx.x + i]; b[i] = in[threadIdx.x + 8 + i]; c[i] = in[threadIdx.x + 16 + i]; }
    u64 p[8], q[8], r[8];
    for (int i = 0; i < 8; ++i) { p[i] = pk(a[i], b[i]); q[i] = pk(b[i], c[i]); r[i] = pk(c[i], a[i]); }
    const float s = in[64];
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = fmaf(a[i], s, 0.5f);            // R, R, imm
            if (MODE == 1) a[i] = fmaf(a[i], b[i], c[i]);          // R, R, R (distinct)
            if (MODE == 2) a[i] = fmaf(a[i], s, c[i]);             // R, R(shared), R
            if (MODE == 3) p[i] = fma2(p[i], pk(s, s), pk(0.5f, 0.5f));   // R2, R.F32, imm
            if (MODE == 4) p[i] = fma2(p[i], pk(s, s), r[i]);      // R2, R.F32, R2
            if (MODE == 5) p[i] = fma2(p[i], q[i], r[i]);          // R2, R2, R2
            if (MODE == 6) p[i] = fma2(p[i], pk(0.999f, 0.999f), r[i]);   // R2, imm, R2
            if (MODE == 7) a[i] = a[i] * b[i];                     // FMUL R, R
            if (MODE == 8) a[i] = fmaf(a[i], a[i], c[i]);          // R, same R, R
        }
    }
    float t = 0; for (int i = 0; i < 8; ++i) t += a[i] + (float)(p[i] & 0xffff);
    out[blockIdx.x * blockDim.x + threadIdx.x] = t;
}
template <int MODE> void run(const char* name, float* d, float* in) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 100000;
    for (int w = 1; w <= 4; w *= 2) {
        k<MODE><<<148, w * 128>>>(d, in, 1000);
        cudaEventRecord(e0); k<MODE><<<148, w * 128>>>(d, in, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("%-28s warps/SMSP %d: %.2f SMSP cycles per instruction\n", name, w, ms * 1e-3 * 1.965e9 / iters / 8.0 / w);
    }
}
int main() {
    float *d, *in; cudaMalloc(&d, 148 * 512 * 4); cudaMalloc(&in, 4096); cudaMemset(in, 0, 4096);
    run<0>("FFMA  R,R,imm", d, in); run<1>("FFMA  R,R,R", d, in); run<2>("FFMA  R,Rshared,R", d, in); run<8>("FFMA  R,sameR,R", d, in); run<7>("FMUL  R,R", d, in);
    run<3>("FFMA2 R2,R.F32,imm", d, in); run<4>("FFMA2 R2,R.F32,R2", d, in); run<5>("FFMA2 R2,R2,R2", d, in); run<6>("FFMA2 R2,imm,R2", d, in);
    return 0;
}
