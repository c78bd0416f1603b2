// rcull_device.cuh -- device-side primitives for librcull (sm_100a only).
//
// PTX wrappers for the TMA bulk-copy / mbarrier input pipeline and the
// release/acquire tile-status words of the decoupled-lookback scan, plus the
// per-entity arithmetic pinned to the reference's operation order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rcull {

constexpr uint32_t kFullMask = 0xffffffffu;

// ---- mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP) ------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// Orders earlier generic-proxy smem accesses before later async-proxy (TMA) writes.
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// try_wait with a suspend-time hint: the hardware parks the warp until the phase completes
// or the hint (ns) runs out, so a waiting warp issues one instruction per time-out instead
// of polling.
__device__ __forceinline__ bool mbar_try_wait_hint(uint64_t* bar, uint32_t parity, uint32_t hint_ns) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity), "r"(hint_ns)
        : "memory");
    return ok != 0;
}
#ifdef RCULL_WAIT_NANOSLEEP  // the previous scheme: poll, backing off with nanosleep
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    uint32_t ns = 32;
    do {
        __nanosleep(ns);
        if (ns < 256) ns *= 2;
    } while (!mbar_try_wait(bar, parity));
}
#else
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait_hint(bar, parity, 4000u)) {
    }
}
#endif
// non-blocking: has the phase with this parity completed?
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-B aligned.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// plain arrive (count 1) with release.cta semantics
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 2-D tiled TMA store shared -> global through a tensor map (SASS UTMASTG), bulk-group completion.
__device__ __forceinline__ void tma_store_2d(const void* tmap, const void* src_smem, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(tmap),
                 "r"(smem_u32(src_smem)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all of this thread's bulk groups have finished READING their shared-memory source
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// ---- tile-status words (decoupled lookback) --------------------------------------
// One 16-byte descriptor per (view, tile), read and written with single 128-bit
// accesses (the same single-transaction property CUB's 16-byte tile descriptors
// rely on), so count and index sum are always seen together and no fence is needed.
__device__ __forceinline__ ulonglong2 ld_relaxed_v2(const ulonglong2* p) {
    ulonglong2 v;
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_v2(ulonglong2* p, uint64_t a, uint64_t b) {
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_relaxed(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_release(uint64_t* p, uint64_t v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// explicit shared-window store: a generic pointer would make ptxas rebuild the window base
__device__ __forceinline__ void sts_u8(uint32_t addr, uint32_t v) {
    asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// ---- warp helpers ----------------------------------------------------------------
__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
    return v;
}
__device__ __forceinline__ uint32_t warp_incl_scan_u32(uint32_t v, uint32_t lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(kFullMask, v, o);
        if (lane >= (uint32_t)o) v += t;
    }
    return v;
}

// ---- packed FP32 (Blackwell FMUL2 / FADD2): two IEEE-rounded operations per instruction ----
__device__ __forceinline__ float2 mul2_rn(float2 a, float2 b) {
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %2;"
        : "=l"(r)
        : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
    return *reinterpret_cast<float2*>(&r);
}
// The same product with .ftz.  ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even with an explicit
// .rn and --fmad=false (the scalar forms are left alone), which would round once where the reference rounds
// twice; it does NOT contract when the two instructions disagree on .ftz.  A flushing product followed by a
// non-flushing sum therefore stays FMUL2.FTZ + FADD2, and equals the IEEE product whenever both factors are
// zero or at least 2^-63 in magnitude (no subnormal goes in, none can come out) -- callers check that.
__device__ __forceinline__ float2 mul2_ftz(float2 a, float2 b) {
    unsigned long long r;
    asm("mul.rn.ftz.f32x2 %0, %1, %2;"
        : "=l"(r)
        : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
    return *reinterpret_cast<float2*>(&r);
}
constexpr float kFtzSafe = 1.0842021724855044e-19f;  // 2^-63: |a|, |b| >= this (or zero) => a * b is zero or a normal number
__device__ __forceinline__ float2 add2_rn(float2 a, float2 b) {
    unsigned long long r;
    asm("add.rn.f32x2 %0, %1, %2;"
        : "=l"(r)
        : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
    return *reinterpret_cast<float2*>(&r);
}

// ---- per-entity arithmetic (operation order = SURVEY 8a; compiled -fmad=false) ----
struct Mat34 {  // upper 3 rows of the column-major model matrix; row 3 is (0,0,0,1)
    float m00, m10, m20;  // column 0
    float m01, m11, m21;  // column 1
    float m02, m12, m22;  // column 2
    float tx, ty, tz;     // column 3
};

// A1: glm::translation(pos) * rot.to_homogeneous() * glm::scaling(repeat(scale))
// (src/ecs.rs:61-62).  The generic 4x4 products only add exact zeros around
// fl(R_rc * s) and t, so the result is computed directly.
__device__ __forceinline__ Mat34 model_matrix(float4 ps, float4 q) {
    const float i = q.x, j = q.y, k = q.z, w = q.w, s = ps.w;
    const float ww = __fmul_rn(w, w), ii = __fmul_rn(i, i), jj = __fmul_rn(j, j), kk = __fmul_rn(k, k);
    const float ij = __fmul_rn(__fmul_rn(i, j), 2.0f), wk = __fmul_rn(__fmul_rn(w, k), 2.0f);
    const float wj = __fmul_rn(__fmul_rn(w, j), 2.0f), ik = __fmul_rn(__fmul_rn(i, k), 2.0f);
    const float jk = __fmul_rn(__fmul_rn(j, k), 2.0f), wi = __fmul_rn(__fmul_rn(w, i), 2.0f);
    Mat34 m;
    m.m00 = __fmul_rn(__fsub_rn(__fsub_rn(__fadd_rn(ww, ii), jj), kk), s);
    m.m01 = __fmul_rn(__fsub_rn(ij, wk), s);
    m.m02 = __fmul_rn(__fadd_rn(wj, ik), s);
    m.m10 = __fmul_rn(__fadd_rn(wk, ij), s);
    m.m11 = __fmul_rn(__fsub_rn(__fadd_rn(__fsub_rn(ww, ii), jj), kk), s);
    m.m12 = __fmul_rn(__fsub_rn(jk, wi), s);
    m.m20 = __fmul_rn(__fsub_rn(ik, wj), s);
    m.m21 = __fmul_rn(__fadd_rn(wi, jk), s);
    m.m22 = __fmul_rn(__fadd_rn(__fsub_rn(__fsub_rn(ww, ii), jj), kk), s);
    m.tx = ps.x;
    m.ty = ps.y;
    m.tz = ps.z;
    return m;
}

// A2 for one row of M: min/max over the 8 transformed corners.  Each corner is
// fl(fl(fl(a*X + b*Y) + c*Z) + t) (nalgebra gemv order) and rounded addition is
// monotone in each argument, so the extremum over {min,max}^3 is reached at the
// per-term extrema: 6 products and 6 sums instead of 8 full transforms.
__device__ __forceinline__ void row_extent(float a, float b, float c, float t, float4 mn, float4 mx, float& lo,
                                           float& hi) {
    const float ax0 = __fmul_rn(a, mn.x), ax1 = __fmul_rn(a, mx.x);
    const float by0 = __fmul_rn(b, mn.y), by1 = __fmul_rn(b, mx.y);
    const float cz0 = __fmul_rn(c, mn.z), cz1 = __fmul_rn(c, mx.z);
    lo = __fadd_rn(t, __fadd_rn(fminf(cz0, cz1), __fadd_rn(fminf(by0, by1), fminf(ax0, ax1))));
    hi = __fadd_rn(t, __fadd_rn(fmaxf(cz0, cz1), __fadd_rn(fmaxf(by0, by1), fmaxf(ax0, ax1))));
}

// A2 tail (src/ecs.rs:173-178) + ncollide3d AABB::from_half_extents.
__device__ __forceinline__ void aabb_store_form(float lo, float hi, float& mins, float& maxs) {
    const float c = __fmul_rn(__fadd_rn(hi, lo), 0.5f);  // (max+min)/2.0, exact same rounding as *0.5
    const float h = __fmul_rn(__fsub_rn(hi, lo), 0.5f);
    mins = __fsub_rn(c, h);
    maxs = __fadd_rn(c, h);
}

}  // namespace rcull
