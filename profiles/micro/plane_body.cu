// Microbenchmark (diagnostic): cycles per (plane, entity pair) of candidate plane-test bodies of the frame kernel,
// registers only, 6 planes unrolled per loop trip, 4 warps per SM sub-partition (16 per SM) and 8 per sub-partition.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -O3 -o profiles/micro/plane_body profiles/micro/plane_body.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
#define ITERS 512
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*(u64*)&a), "l"(*(u64*)&b)); return *(float2*)&r; }
__device__ __forceinline__ float2 mul2z(float2 a, float2 b) { u64 r; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*(u64*)&a), "l"(*(u64*)&b)); return *(float2*)&r; }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*(u64*)&a), "l"(*(u64*)&b)); return *(float2*)&r; }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(*(u64*)&a), "l"(*(u64*)&b), "l"(*(u64*)&c)); return *(float2*)&r; }

template <int BODY>
__global__ void k(const float4* planes, float* out, u64* cycles, float seed) {
    // two entity pairs per thread, as in the kernel (kItems = 4)
    float2 hx[2], hy[2], hz[2], cx[2], cy[2], cz[2];
    for (int i = 0; i < 2; ++i) {
        hx[i] = make_float2(seed + threadIdx.x, seed * 2 + i); hy[i] = make_float2(seed * 3, seed + i); hz[i] = make_float2(seed * 0.5f, seed * 0.25f + i);
        cx[i] = make_float2(-seed - threadIdx.x, seed * 7 + i); cy[i] = make_float2(seed * 5, -seed + i); cz[i] = make_float2(seed * 1.5f, seed * 2.25f + i);
    }
    __shared__ float4 sp[5 * 6 * 4];  // five "views": the trip picks one, so the loads and every chain stay in the loop
    if (threadIdx.x < 120) sp[threadIdx.x] = planes[threadIdx.x % 24];
    __syncthreads();
    bool outside[4] = {false, false, false, false};
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int p = 0; p < 6; ++p) {
            const float4* pp = sp + (it % 5) * 24 + p * 4;
            const float4 q0 = pp[0], q1 = pp[1], q2 = pp[2], q3 = pp[3];
            const float2 ax = make_float2(q0.x, q0.y), ay = make_float2(q0.z, q0.w), az = make_float2(q1.x, q1.y);
            const float2 px = make_float2(q1.z, q1.w), py = make_float2(q2.x, q2.y), pz = make_float2(q2.z, q2.w), pw2 = make_float2(q3.x, q3.y);
            const float pw = q3.x;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                if (BODY == 0) {  // all packed: FMUL2.FTZ + FADD2
                    const float2 e = add2(add2(mul2z(hx[i], ax), mul2z(hy[i], ay)), mul2z(hz[i], az));
                    const float2 a = add2(mul2z(cx[i], px), mul2z(cz[i], pz)), b = add2(mul2z(cy[i], py), pw2);
                    const float2 s = add2(a, b);
                    outside[2 * i] |= s.x > e.x; outside[2 * i + 1] |= s.y > e.y;
                } else if (BODY == 1) {  // round-1 body: packed products, scalar sums, one packed sum
                    const float2 m1 = mul2(hx[i], ax), m2 = mul2(hy[i], ay), m3 = mul2(hz[i], az);
                    const float e0 = __fadd_rn(__fadd_rn(m1.x, m2.x), m3.x), e1 = __fadd_rn(__fadd_rn(m1.y, m2.y), m3.y);
                    const float2 p1 = mul2(cx[i], px), p3 = mul2(cz[i], pz), p2 = mul2(cy[i], py);
                    const float2 a = make_float2(__fadd_rn(p1.x, p3.x), __fadd_rn(p1.y, p3.y)), b = make_float2(__fadd_rn(p2.x, pw), __fadd_rn(p2.y, pw));
                    const float2 s = add2(a, b);
                    outside[2 * i] |= s.x > e0; outside[2 * i + 1] |= s.y > e1;
                } else if (BODY == 2) {  // all scalar
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float HX = h ? hx[i].y : hx[i].x, HY = h ? hy[i].y : hy[i].x, HZ = h ? hz[i].y : hz[i].x;
                        const float CX = h ? cx[i].y : cx[i].x, CY = h ? cy[i].y : cy[i].x, CZ = h ? cz[i].y : cz[i].x;
                        const float e = __fadd_rn(__fadd_rn(__fmul_rn(HX, ax.x), __fmul_rn(HY, ay.x)), __fmul_rn(HZ, az.x));
                        const float s = __fadd_rn(__fadd_rn(__fmul_rn(CX, px.x), __fmul_rn(CZ, pz.x)), __fadd_rn(__fmul_rn(CY, py.x), pw));
                        outside[2 * i + h] |= s > e;
                    }
                } else if (BODY == 3) {  // scalar products, packed sums
                    const float2 m1 = make_float2(__fmul_rn(hx[i].x, ax.x), __fmul_rn(hx[i].y, ax.x)), m2 = make_float2(__fmul_rn(hy[i].x, ay.x), __fmul_rn(hy[i].y, ay.x)),
                                 m3 = make_float2(__fmul_rn(hz[i].x, az.x), __fmul_rn(hz[i].y, az.x));
                    const float2 p1 = make_float2(__fmul_rn(cx[i].x, px.x), __fmul_rn(cx[i].y, px.x)), p2 = make_float2(__fmul_rn(cy[i].x, py.x), __fmul_rn(cy[i].y, py.x)),
                                 p3 = make_float2(__fmul_rn(cz[i].x, pz.x), __fmul_rn(cz[i].y, pz.x));
                    const float2 e = add2(add2(m1, m2), m3), s = add2(add2(p1, p3), add2(p2, pw2));
                    outside[2 * i] |= s.x > e.x; outside[2 * i + 1] |= s.y > e.y;
                } else if (BODY == 4) {  // NOT exact (reference point): contracted FFMA2 chains
                    const float2 e = fma2(hz[i], az, fma2(hy[i], ay, mul2(hx[i], ax)));
                    const float2 s = fma2(cz[i], pz, fma2(cy[i], py, fma2(cx[i], px, pw2)));
                    outside[2 * i] |= s.x > e.x; outside[2 * i + 1] |= s.y > e.y;
                } else if (BODY == 5) {  // NOT exact: scalar FFMA chains
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const float HX = h ? hx[i].y : hx[i].x, HY = h ? hy[i].y : hy[i].x, HZ = h ? hz[i].y : hz[i].x;
                        const float CX = h ? cx[i].y : cx[i].x, CY = h ? cy[i].y : cy[i].x, CZ = h ? cz[i].y : cz[i].x;
                        const float e = __fmaf_rn(HZ, az.x, __fmaf_rn(HY, ay.x, __fmul_rn(HX, ax.x)));
                        const float s = __fmaf_rn(CZ, pz.x, __fmaf_rn(CY, py.x, __fmaf_rn(CX, px.x, pw)));
                        outside[2 * i + h] |= s > e;
                    }
                }
            }
        }
        // keep the loop from being hoisted: perturb one input per trip
        cx[0].x = __int_as_float(__float_as_int(cx[0].x) ^ (it & 1));
        hx[1].y = __int_as_float(__float_as_int(hx[1].y) ^ (it & 2));
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = (outside[0] ? 1.f : 0.f) + (outside[1] ? 2.f : 0.f) + (outside[2] ? 4.f : 0.f) + (outside[3] ? 8.f : 0.f) + cx[0].x + hx[1].y;
    if (threadIdx.x == 0) cycles[blockIdx.x] = (u64)(t1 - t0);
}
template <int BODY>
void run(const char* name, const float4* planes, int warps) {
    float* out; u64* cyc;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
    for (int r = 0; r < 2; ++r) { k<BODY><<<148, warps * 32>>>(planes, out, cyc, 1.37f); cudaDeviceSynchronize(); }
    u64 h[148]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    const double pair_planes = (double)ITERS * 6 * 2 * warps / 4.0;  // per SM sub-partition
    printf("%-44s warps/SM %2d: %6.2f cycles per (plane, entity pair) per sub-partition\n", name, warps, avg / pair_planes);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    float4 hp[24]; for (int i = 0; i < 24; ++i) hp[i] = make_float4(0.3f + i, 0.3f + i, -0.7f + i, -0.7f + i);
    float4* dp; cudaMalloc(&dp, sizeof hp); cudaMemcpy(dp, hp, sizeof hp, cudaMemcpyHostToDevice);
    for (int w : {4, 16, 32}) {
        run<0>("packed products (.ftz) + packed sums", dp, w);
        run<1>("packed products + scalar sums (round 1)", dp, w);
        run<2>("scalar products + scalar sums", dp, w);
        run<3>("scalar products + packed sums", dp, w);
        run<4>("[inexact] FFMA2 chains", dp, w);
        run<5>("[inexact] scalar FFMA chains", dp, w);
    }
    return 0;
}
