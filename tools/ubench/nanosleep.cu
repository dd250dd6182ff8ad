// How long does __nanosleep(t) sleep on sm_100a, and what does a polling ld.acquire loop cost?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(unsigned ns, int iters, long long* out, const int* flag) {
    long long t0 = clock64();
    int acc = 0;
    for (int i = 0; i < iters; ++i) { __nanosleep(ns); acc += i; }
    long long t1 = clock64();
    for (int i = 0; i < iters; ++i) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory"); acc += v; }
    long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t1; out[2] = acc;
}
int main() {
    long long* d; int* f; cudaMalloc(&d, 64); cudaMalloc(&f, 64); cudaMemset(f, 0, 64);
    for (unsigned ns : {0u, 32u, 64u, 250u, 500u, 1000u, 4000u}) {
        k<<<1, 1>>>(ns, 1000, d, f); long long h[3]; cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
        printf("nanosleep(%u): %.1f cycles per call (%.1f ns at 1.965 GHz); ld.acquire.gpu: %.1f cycles per poll\n", ns, h[0] / 1000.0, h[0] / 1000.0 / 1.965, h[1] / 1000.0);
    }
    return 0;
}
