// Microbenchmark (diagnostic): issue/pipe throughput of the FP32 instructions the plane tests use,
// scalar vs packed (f32x2), on sm_100a.  Prints warp-instructions per cycle per SM sub-partition.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -o gpurun_out/pipes profiles/micro/pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define ITERS 2048
#define CHAINS 8
template <int OP>
__global__ void k(float* out, u64* cycles, float seed) {
    float a[CHAINS], b[CHAINS];
    u64 p[CHAINS], q[CHAINS];
    for (int i = 0; i < CHAINS; ++i) {
        a[i] = seed + threadIdx.x + i; b[i] = seed * 0.5f + i;
        p[i] = ((u64)__float_as_uint(a[i]) << 32) | __float_as_uint(b[i]); q[i] = p[i] ^ 0x0010000000100000ull;
    }
    const float c1 = seed * 1.0001f, c2 = seed * 0.9999f;
    const u64 cc = ((u64)__float_as_uint(c1) << 32) | __float_as_uint(c2);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) {
            if (OP == 0) a[i] = __fadd_rn(a[i], c1);
            if (OP == 1) a[i] = __fmul_rn(a[i], c1);
            if (OP == 2) a[i] = __fmaf_rn(a[i], c1, c2);
            if (OP == 3) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cc));
            if (OP == 4) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(cc), "l"(q[i]));
            if (OP == 5) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cc));
            if (OP == 6) a[i] = fmaxf(a[i], b[i] + 0.f), b[i] = __int_as_float(__float_as_int(b[i]) + 1);  // FMNMX + IADD
            if (OP == 7) { a[i] = __fadd_rn(a[i], c1); asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cc)); }  // 1 FADD : 1 FMUL2
            if (OP == 8) { a[i] = __fadd_rn(a[i], c1); b[i] = __fadd_rn(b[i], c2); asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cc)); }  // 2 FADD : 1 FMUL2
            if (OP == 9) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(q[i]) : "l"(cc), "l"(q[i])); asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(cc)); }  // 1 FFMA2 : 1 FMUL2
            if (OP == 10) { a[i] = __fadd_rn(a[i], c1); b[i] = fmaxf(b[i], a[i]); }  // FADD + FMNMX
            if (OP == 11) { a[i] = __fmul_rn(a[i], c1); b[i] = __fadd_rn(b[i], c2); }  // FMUL + FADD
        }
    }
    const long long t1 = clock64();
    float s = 0; u64 x = 0;
    for (int i = 0; i < CHAINS; ++i) { s += a[i] + b[i]; x ^= p[i] ^ q[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)x;
    if (threadIdx.x == 0) cycles[blockIdx.x] = (u64)(t1 - t0);
}
template <int OP>
void run(const char* name, int per_iter, int warps) {
    float* out; u64* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    k<OP><<<148, warps * 32>>>(out, cyc, 1.0f); cudaDeviceSynchronize();
    k<OP><<<148, warps * 32>>>(out, cyc, 1.0f); cudaDeviceSynchronize();
    u64 h[148]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    const double inst = (double)ITERS * CHAINS * per_iter * warps / 4.0;  // warp-instructions per SMSP
    printf("%-22s warps/SM %2d: %.3f warp-inst/cycle/SMSP (%.0f cycles)\n", name, warps, inst / avg, avg);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    for (int w : {4, 8, 16}) {
        run<0>("FADD", 1, w); run<1>("FMUL", 1, w); run<2>("FFMA", 1, w); run<3>("FMUL2", 1, w); run<4>("FFMA2", 1, w);
        run<5>("FADD2", 1, w); run<6>("FMNMX+IADD", 2, w); run<7>("FADD+FMUL2", 2, w); run<8>("2FADD+FMUL2", 3, w);
        run<9>("FFMA2+FMUL2", 2, w); run<10>("FADD+FMNMX", 2, w); run<11>("FMUL+FADD", 2, w);
    }
    return 0;
}
