// Microbenchmark: issue rate of packed fp32 (fma.rn.f32x2 -> FFMA2) against scalar FMUL/FADD on sm_100a, alone and
// mixed with integer work, to decide whether packing the SAT support scans can pay.  Not part of the product.
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long pk(float a, float b) {
    unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d;
}

template <int MODE>
__global__ void k(float* out, int iters, float one, float nzero, long long* clk) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    unsigned long long p0 = pk(a0, a1), p1 = pk(a2, a3), p2 = pk(a4, a5), p3 = pk(a6, a7);
    const unsigned long long ones = pk(one, one), nz = pk(nzero, nzero), m = pk(1.0001f, 0.9999f);
    int i0 = threadIdx.x, i1 = i0 + 1, i2 = i0 + 2, i3 = i0 + 3;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        if (MODE == 0) {  // 8 scalar FMUL (independent chains)
            a0 = __fmul_rn(a0, 1.0001f); a1 = __fmul_rn(a1, 0.9999f); a2 = __fmul_rn(a2, 1.0001f); a3 = __fmul_rn(a3, 0.9999f);
            a4 = __fmul_rn(a4, 1.0001f); a5 = __fmul_rn(a5, 0.9999f); a6 = __fmul_rn(a6, 1.0001f); a7 = __fmul_rn(a7, 0.9999f);
        } else if (MODE == 1) {  // 4 packed multiplies = the same 8 products
            p0 = fma2(p0, m, nz); p1 = fma2(p1, m, nz); p2 = fma2(p2, m, nz); p3 = fma2(p3, m, nz);
        } else if (MODE == 2) {  // 8 scalar FMUL + 8 integer ops
            a0 = __fmul_rn(a0, 1.0001f); a1 = __fmul_rn(a1, 0.9999f); a2 = __fmul_rn(a2, 1.0001f); a3 = __fmul_rn(a3, 0.9999f);
            a4 = __fmul_rn(a4, 1.0001f); a5 = __fmul_rn(a5, 0.9999f); a6 = __fmul_rn(a6, 1.0001f); a7 = __fmul_rn(a7, 0.9999f);
            i0 = (i0 ^ i1) + 3; i1 = (i1 ^ i2) + 5; i2 = (i2 ^ i3) + 7; i3 = (i3 ^ i0) + 9;
        } else if (MODE == 3) {  // 4 packed + the same 8 integer ops
            p0 = fma2(p0, m, nz); p1 = fma2(p1, m, nz); p2 = fma2(p2, m, nz); p3 = fma2(p3, m, nz);
            i0 = (i0 ^ i1) + 3; i1 = (i1 ^ i2) + 5; i2 = (i2 ^ i3) + 7; i3 = (i3 ^ i0) + 9;
        } else if (MODE == 4) {  // packed add form x*1+y
            p0 = fma2(p0, ones, m); p1 = fma2(p1, ones, m); p2 = fma2(p2, ones, m); p3 = fma2(p3, ones, m);
        }
    }
    long long t1 = clock64();
    float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + __int_as_float((int)(p0 ^ p1 ^ p2 ^ p3)) + __int_as_float((int)((p0 ^ p1 ^ p2 ^ p3) >> 32)) + (i0 + i1 + i2 + i3);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}

template <int MODE>
void run(const char* name, int warps_per_sm) {
    float* out; long long* clk; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 8);
    const int iters = 100000;
    k<MODE><<<148, warps_per_sm * 32>>>(out, 10, 1.0f, -0.0f, clk);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<148, warps_per_sm * 32>>>(out, iters, 1.0f, -0.0f, clk);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    // per SMSP: warps_per_sm/4 warps each executing `iters` loop bodies
    printf("%-34s warps/SM %2d  %7.3f ms  cycles/iter/warp-on-SMSP %6.2f  (clock64 %lld)  %s\n", name, warps_per_sm, ms,
           (double)c / iters / (warps_per_sm / 4.0), c, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(clk);
}

int main() {
    for (int w : {16, 32}) {
        if (w == 16) { run<0>("8 FMUL", 16); run<1>("4 FFMA2 (mul form)", 16); run<4>("4 FFMA2 (add form)", 16); run<2>("8 FMUL + 8 int", 16); run<3>("4 FFMA2 + 8 int", 16); }
        else { run<0>("8 FMUL", 32); run<1>("4 FFMA2 (mul form)", 32); run<4>("4 FFMA2 (add form)", 32); run<2>("8 FMUL + 8 int", 32); run<3>("4 FFMA2 + 8 int", 32); }
    }
    return 0;
}
