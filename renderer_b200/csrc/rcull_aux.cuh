// rcull_aux.cuh -- small streaming kernels around the frame kernel: the
// range-shard merge step (firstIndex rebasing into the merged draw list) and
// the counter-based synthetic scene generator used by the harness.
#pragma once
#include "rcull_device.cuh"

namespace rcull {

// copy this rank's dense draw list into its slot of the
// merged list, adding the exclusive prefix of earlier ranks' index totals to
// firstIndex.  Word-granular so loads and stores stay coalesced: word w of the
// 5-word command stream is firstIndex iff w % 5 == 2.
__global__ void __launch_bounds__(256) rebase_into(const uint32_t* __restrict__ src_cmds,
                                                   const uint32_t* __restrict__ src_visible,
                                                   const uint32_t* __restrict__ count_ptr, uint32_t cap,
                                                   uint32_t index_prefix, uint32_t* __restrict__ dst_cmds,
                                                   uint32_t* __restrict__ dst_visible) {
    uint32_t count = *count_ptr;
    if (count > cap) count = cap;
    const size_t words = (size_t)count * 5;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += stride) {
        uint32_t v = src_cmds[w];
        if (w % 5 == 2) v += index_prefix;
        dst_cmds[w] = v;
    }
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        dst_visible[i] = src_visible[i];
}

// ---- synthetic scene (twin of renderer_b200/scene.py and oracle_synth_entities) ----
__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t chan(uint64_t seed, uint64_t i, uint64_t c) {
    return mix64(seed + (i * 16 + c + 1) * 0x9E3779B97F4A7C15ull);
}
__device__ __forceinline__ float unit24(uint64_t h) {
    return __fmul_rn((float)(uint32_t)(h >> 40), 5.9604644775390625e-08f);
}

__global__ void __launch_bounds__(256) synth_entities(uint64_t seed, uint64_t first_global, uint32_t n,
                                                      float half_box, uint32_t n_meshes, float4* pos_scale,
                                                      float4* rot, uint32_t* mesh) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint64_t i = first_global + k;
    float4 ps;
    ps.x = __fmul_rn(__fsub_rn(__fmul_rn(unit24(chan(seed, i, 0)), 2.0f), 1.0f), half_box);
    ps.y = __fmul_rn(__fsub_rn(__fmul_rn(unit24(chan(seed, i, 1)), 2.0f), 1.0f), half_box);
    ps.z = __fmul_rn(__fsub_rn(__fmul_rn(unit24(chan(seed, i, 2)), 2.0f), 1.0f), half_box);
    ps.w = __fadd_rn(__fmul_rn(unit24(chan(seed, i, 3)), 3.75f), 0.25f);
    float c0 = __fsub_rn(__fmul_rn(unit24(chan(seed, i, 4)), 2.0f), 1.0f);
    float c1 = __fsub_rn(__fmul_rn(unit24(chan(seed, i, 5)), 2.0f), 1.0f);
    float c2 = __fsub_rn(__fmul_rn(unit24(chan(seed, i, 6)), 2.0f), 1.0f);
    float c3 = __fsub_rn(__fmul_rn(unit24(chan(seed, i, 7)), 2.0f), 1.0f);
    const float n2 = __fadd_rn(
        __fadd_rn(__fadd_rn(__fmul_rn(c0, c0), __fmul_rn(c1, c1)), __fmul_rn(c2, c2)), __fmul_rn(c3, c3));
    if (n2 < 1e-4f) {
        c0 = c1 = c2 = 0.0f;
        c3 = 1.0f;
    } else {
        const float nn = __fsqrt_rn(n2);
        c0 = __fdiv_rn(c0, nn);
        c1 = __fdiv_rn(c1, nn);
        c2 = __fdiv_rn(c2, nn);
        c3 = __fdiv_rn(c3, nn);
    }
    // UnitQuaternion::new_normalize: q / sqrt((q0^2 + q2^2) + (q1^2 + q3^2))
    const float nn = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(c0, c0), __fmul_rn(c2, c2)),
                                          __fadd_rn(__fmul_rn(c1, c1), __fmul_rn(c3, c3))));
    rot[k] = make_float4(__fdiv_rn(c0, nn), __fdiv_rn(c1, nn), __fdiv_rn(c2, nn), __fdiv_rn(c3, nn));
    pos_scale[k] = ps;
    mesh[k] = (uint32_t)(((chan(seed, i, 8) >> 32) * (uint64_t)n_meshes) >> 32);
}

}  // namespace rcull
