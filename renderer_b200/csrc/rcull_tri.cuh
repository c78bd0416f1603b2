// rcull_tri.cuh -- stage 2 of the cull pipeline (SURVEY 8a row A8 + row A7 on its output):
// per-triangle cull of the camera view's survivors and compaction of the commands that kept
// a triangle.  Replaces the reference's one-vkCmdDispatch-per-visible-entity loop
// (cull_pipeline.rs:534-577 driving src/shaders/generate_work.comp) and the
// compact_draw_stream dispatch (cull_pipeline.rs:589-605) with two launches per frame.
//
// Arithmetic is pinned to the GLSL source: `mat4 * vec4` is the left-to-right combination
// of the columns with every product and sum rounded (no contraction: the library is built
// with -fmad=false and uses explicit _rn intrinsics), `/` is IEEE division.
#pragma once
#include "rcull_device.cuh"

namespace rcull {

constexpr int kTriThreads = 128;
constexpr int kTriItems = 2;  // triangles per thread per step: 256 triangles of one command per barrier
constexpr int kCompactThreads = 256;
constexpr int kCompactItems = 4;
constexpr int kCompactTile = kCompactThreads * kCompactItems;

struct Mat4 {
    float m[16];  // column-major
};
// The same matrix with every entry stored twice: the multiplier of a packed (FMUL2) product
// whose two halves belong to the two triangles a thread handles per step.
struct Mat4x2 {
    float2 m[16];
};

constexpr int kTriVertCap = 2048;  // clip-space vertices of one chunk kept in shared memory (x, y, w: 24 KB)
constexpr uint32_t kSlotAllCached = 0x80000000u;  // flag in slot_chunks[].y: every chunk of the LOD has a cacheable vertex window

struct TriArgs {
    const uint32_t* cmds;        // view's draw list from the frame kernel, 5 words per command
    const uint32_t* src_slot;    // (mesh, lod) slot of each command: its LOD row gives the indexOffset push constant
    const uint4* lod_tab;        // per slot: index_len, first_index, vertex_offset, 0
    const uint2* slot_chunks;    // per slot: first chunk, number of chunks | kSlotAllCached (built on the host from the index buffer)
    const uint4* chunks;         // tri_begin (within the LOD), tri_count, v_min, v_count (0: gather path)
    const uint32_t* count;       // how many commands (device)
    const float* model;          // mat4 column-major, slot = firstInstance - draw_index_base
    const float* positions;      // ConsolidatedMeshBuffers.position_buffer, vec3 scalar layout
    const uint32_t* indices;     // ConsolidatedMeshBuffers.index_buffer
    uint32_t* out_indices;       // CulledIndexBuffer (uvec3 out_index_buffer[], generate_work.comp:40-42)
    uint32_t* refined;           // [count] refined indexCount of every command
    uint32_t* ticket;            // [0], [1] command tickets of the two cull kernels, [2] tile ticket of the compaction
    uint32_t* error_flag;
    uint64_t n_vertices, n_indices, out_capacity;
    uint32_t cap_cmds, draw_index_base, n_slots, n_chunks;
    uint32_t mixed_kinds;        // both kinds of LOD exist: each tri_cull kernel skips the commands of the other
    Mat4x2 pv;  // read straight out of the constant bank (kernel parameter): no loads
};

// GLSL mat4 * vec4: ((c0 * x + c1 * y) + c2 * z) + c3 * w per row, every product and sum rounded.
__device__ __forceinline__ float4 glsl_mul(const Mat4& a, float4 v) {
    const float* m = a.m;
    float4 r;
    r.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[0], v.x), __fmul_rn(m[4], v.y)), __fmul_rn(m[8], v.z)), __fmul_rn(m[12], v.w));
    r.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[1], v.x), __fmul_rn(m[5], v.y)), __fmul_rn(m[9], v.z)), __fmul_rn(m[13], v.w));
    r.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[2], v.x), __fmul_rn(m[6], v.y)), __fmul_rn(m[10], v.z)), __fmul_rn(m[14], v.w));
    r.w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[3], v.x), __fmul_rn(m[7], v.y)), __fmul_rn(m[11], v.z)), __fmul_rn(m[15], v.w));
    return r;
}
// the same with v.w == 1.0: c3 * 1.0 is c3 itself, bit for bit, so the four products are skipped
__device__ __forceinline__ float4 glsl_mul_w1(const Mat4& a, float x, float y, float z) {
    const float* m = a.m;
    float4 r;
    r.x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[0], x), __fmul_rn(m[4], y)), __fmul_rn(m[8], z)), m[12]);
    r.y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[1], x), __fmul_rn(m[5], y)), __fmul_rn(m[9], z)), m[13]);
    r.z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[2], x), __fmul_rn(m[6], y)), __fmul_rn(m[10], z)), m[14]);
    r.w = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[3], x), __fmul_rn(m[7], y)), __fmul_rn(m[11], z)), m[15]);
    return r;
}

// generate_work.comp:139-155.  true = culled.  v[k] = the k-th vertex position.
template <typename PV>
__device__ __forceinline__ bool tri_culled(const Mat4& model, const PV& pv, const float (&v)[3][3]) {
    const float4 c0 = glsl_mul(pv, glsl_mul_w1(model, v[0][0], v[0][1], v[0][2]));
    const float4 c1 = glsl_mul(pv, glsl_mul_w1(model, v[1][0], v[1][1], v[1][2]));
    const float4 c2 = glsl_mul(pv, glsl_mul_w1(model, v[2][0], v[2][1], v[2][2]));
    // determinant(mat3(c0.xyw, c1.xyw, c2.xyw)) > 0: cofactors along the first row of columns
    const float t0 = __fsub_rn(__fmul_rn(c1.y, c2.w), __fmul_rn(c2.y, c1.w));
    const float t1 = __fsub_rn(__fmul_rn(c0.y, c2.w), __fmul_rn(c2.y, c0.w));
    const float t2 = __fsub_rn(__fmul_rn(c0.y, c1.w), __fmul_rn(c1.y, c0.w));
    const float det = __fadd_rn(__fsub_rn(__fmul_rn(c0.x, t0), __fmul_rn(c1.x, t1)), __fmul_rn(c2.x, t2));
    if (det > 0.0f) return true;
    const float x0 = __fdiv_rn(c0.x, c0.w), y0 = __fdiv_rn(c0.y, c0.w);
    const float x1 = __fdiv_rn(c1.x, c1.w), y1 = __fdiv_rn(c1.y, c1.w);
    const float x2 = __fdiv_rn(c2.x, c2.w), y2 = __fdiv_rn(c2.y, c2.w);
    return (x0 < -1.0f && x1 < -1.0f && x2 < -1.0f) || (x0 > 1.0f && x1 > 1.0f && x2 > 1.0f) ||
           (y0 < -1.0f && y1 < -1.0f && y2 < -1.0f) || (y0 > 1.0f && y1 > 1.0f && y2 > 1.0f);
    // the area test of :159-167 is overwritten by `cull = false` (:166): nothing to compute
}

// ---- two triangles per thread: packed products, scalar sums ------------------------------
// Every product of the two transforms and of the determinant is an FMUL2 over the pair
// (triangle j = 0, 1 of the step in the two halves); every sum that consumes a packed product
// is a scalar FADD, so there is no packed mul + add pair for ptxas to contract into FFMA2
// (it does that even under --fmad=false; tests/test_abi.py checks the SASS).  The one packed
// add is sum + matrix entry.  Each half goes through exactly the operations of glsl_mul /
// glsl_mul_w1 above, in the same order.
__device__ __forceinline__ float2 sum3(float2 a, float2 b, float2 c) {
    return make_float2(__fadd_rn(__fadd_rn(a.x, b.x), c.x), __fadd_rn(__fadd_rn(a.y, b.y), c.y));
}
__device__ __forceinline__ float2 sub2s(float2 a, float2 b) { return make_float2(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y)); }
// row r of model * vec4(v, 1)
__device__ __forceinline__ float2 row_w1(const Mat4x2& a, int r, float2 x, float2 y, float2 z) {
    return add2_rn(sum3(mul2_rn(a.m[r], x), mul2_rn(a.m[4 + r], y), mul2_rn(a.m[8 + r], z)), a.m[12 + r]);
}
// row r of pv * v
__device__ __forceinline__ float2 row_full(const Mat4x2& a, int r, const float2 (&v)[4]) {
    const float2 s = sum3(mul2_rn(a.m[r], v[0]), mul2_rn(a.m[4 + r], v[1]), mul2_rn(a.m[8 + r], v[2]));
    const float2 w = mul2_rn(a.m[12 + r], v[3]);
    return make_float2(__fadd_rn(s.x, w.x), __fadd_rn(s.y, w.y));
}
__device__ __forceinline__ bool ndc_culled(float c0x, float c0y, float c0w, float c1x, float c1y, float c1w, float c2x, float c2y, float c2w) {
    const float x0 = __fdiv_rn(c0x, c0w), y0 = __fdiv_rn(c0y, c0w);
    const float x1 = __fdiv_rn(c1x, c1w), y1 = __fdiv_rn(c1y, c1w);
    const float x2 = __fdiv_rn(c2x, c2w), y2 = __fdiv_rn(c2y, c2w);
    return (x0 < -1.0f && x1 < -1.0f && x2 < -1.0f) || (x0 > 1.0f && x1 > 1.0f && x2 > 1.0f) ||
           (y0 < -1.0f && y1 < -1.0f && y2 < -1.0f) || (y0 > 1.0f && y1 > 1.0f && y2 > 1.0f);
}
// clip = pv * (model * vec4(v, 1)) for TWO vertices at once (one per half); only the x, y, w rows of pv are needed.
template <typename PV>
__device__ __forceinline__ void clip2(const Mat4x2& model, const PV& pv, float2 x, float2 y, float2 z, float2& cx, float2& cy, float2& cw) {
    const float2 w[4] = {row_w1(model, 0, x, y, z), row_w1(model, 1, x, y, z), row_w1(model, 2, x, y, z), row_w1(model, 3, x, y, z)};
    cx = row_full(pv, 0, w);
    cy = row_full(pv, 1, w);
    cw = row_full(pv, 3, w);
}
// generate_work.comp:139-155 on the clip-space corners of two triangles (one per half): cx[k] = clip x of corner k.
__device__ __forceinline__ void cull_from_clip(const float2 (&cx)[3], const float2 (&cy)[3], const float2 (&cw)[3], bool& culled0, bool& culled1) {
    // determinant(mat3(c0.xyw, c1.xyw, c2.xyw)) > 0: cofactors along the first row of columns
    const float2 t0 = sub2s(mul2_rn(cy[1], cw[2]), mul2_rn(cy[2], cw[1]));
    const float2 t1 = sub2s(mul2_rn(cy[0], cw[2]), mul2_rn(cy[2], cw[0]));
    const float2 t2 = sub2s(mul2_rn(cy[0], cw[1]), mul2_rn(cy[1], cw[0]));
    const float2 d0 = mul2_rn(cx[0], t0), d1 = mul2_rn(cx[1], t1), d2 = mul2_rn(cx[2], t2);
    const float det0 = __fadd_rn(__fsub_rn(d0.x, d1.x), d2.x), det1 = __fadd_rn(__fsub_rn(d0.y, d1.y), d2.y);
    // The four NDC tests each need ALL three vertices beyond the same side.  A vertex with w > 0,
    // |x| <= w and |y| <= w has x / w and y / w inside [-1, 1] after rounding too (division rounds
    // monotonically and +-1 are representable; inf / inf = NaN compares false as well), so one such
    // vertex decides "not culled" without the six divisions -- the common case for a mesh on screen.
    bool in0 = false, in1 = false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        in0 = in0 || (cw[k].x > 0.0f && fabsf(cx[k].x) <= cw[k].x && fabsf(cy[k].x) <= cw[k].x);
        in1 = in1 || (cw[k].y > 0.0f && fabsf(cx[k].y) <= cw[k].y && fabsf(cy[k].y) <= cw[k].y);
    }
    culled0 = det0 > 0.0f;
    culled1 = det1 > 0.0f;
    if (!culled0 && !in0) culled0 = ndc_culled(cx[0].x, cy[0].x, cw[0].x, cx[1].x, cy[1].x, cw[1].x, cx[2].x, cy[2].x, cw[2].x);
    if (!culled1 && !in1) culled1 = ndc_culled(cx[0].y, cy[0].y, cw[0].y, cx[1].y, cy[1].y, cw[1].y, cx[2].y, cy[2].y, cw[2].y);
    // the area test of :159-167 is overwritten by `cull = false` (:166): nothing to compute
}
// generate_work.comp:139-155 for triangles 0 and 1 of the step; v[j][k] = vertex k of triangle j.
template <typename PV>
__device__ __forceinline__ void tri_culled2(const Mat4x2& model, const PV& pv, const float (&v)[2][3][3], bool& culled0, bool& culled1) {
    float2 cx[3], cy[3], cw[3];  // clip-space x, y, w of vertex k, one triangle per half
#pragma unroll
    for (int k = 0; k < 3; ++k)
        clip2(model, pv, make_float2(v[0][k][0], v[1][k][0]), make_float2(v[0][k][1], v[1][k][1]), make_float2(v[0][k][2], v[1][k][2]), cx[k], cy[k], cw[k]);
    cull_from_clip(cx, cy, cw, culled0, culled1);
}

// One CTA per command at a time (ticket order).  The LOD behind the command comes as CHUNKS of its triangle list
// (built on the host from the index buffer, rcull.cu): a chunk whose triangles reference a window of at most
// kTriVertCap vertices has that window transformed ONCE into shared memory (clip-space x, y, w), and its triangles --
// 256 per step, two per thread -- pick their corners from there: an indexed mesh shares every vertex among ~6
// triangles, so this is a third to a sixth of the transforms of the reference's shader, which transforms three
// vertices per triangle (generate_work.comp:132-137).  Chunks with a wide window (random index triples) gather and
// transform per triangle as the shader does.  Either way each vertex goes through exactly the operations of
// glsl_mul / glsl_mul_w1, so the result does not depend on which path ran.  Kept triples land in ascending triangle order at
// out_index_buffer[firstIndex / 3 + rank] through ballot compaction with a running base (the reference's order
// inside a mesh is atomic-arrival order; ascending is one of its outcomes).
// Two instantiations split the commands between them by the kind of their LOD: kWindow = true takes the LODs whose
// chunks all have a cacheable vertex window (24 KB of shared memory per CTA), kWindow = false the others with NO shared
// memory -- the gathers of a mesh with random index triples live on a large L1 (a mesh's vertices fit it), which a
// shared-memory carveout for eight CTAs would take away (measured: 0.78 -> 1.5 ms on the random-triple scene).
template <bool kWindow>
__global__ void __launch_bounds__(kTriThreads) tri_cull(const __grid_constant__ TriArgs a) {
    __shared__ uint32_t s_cmd, s_warp[2][kTriItems][kTriThreads / 32];
    __shared__ float s_cx[kWindow ? kTriVertCap : 1], s_cy[kWindow ? kTriVertCap : 1], s_cw[kWindow ? kTriVertCap : 1];
    uint32_t* const ticket = a.ticket + (kWindow ? 0 : 1);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    uint32_t count = *a.count;
    if (count > a.cap_cmds) count = a.cap_cmds;
    for (;;) {
        if (tid == 0) {  // the next command of this kernel's kind
            uint32_t c = atomicAdd(ticket, 1u);
            while (a.mixed_kinds && c < count) {
                const uint32_t sl = a.src_slot[c];
                if (((a.slot_chunks[sl < a.n_slots ? sl : 0u].y & kSlotAllCached) != 0u) == kWindow) break;
                c = atomicAdd(ticket, 1u);
            }
            s_cmd = c;
        }
        __syncthreads();
        const uint32_t c = s_cmd;
        if (c >= count) break;
        const uint32_t* cmd = a.cmds + (size_t)c * 5;
        const uint32_t n_tri = cmd[0] / 3u, out_first = cmd[2] / 3u, first_instance = cmd[4];
        const int32_t vertex_offset = (int32_t)cmd[3];
        uint32_t slot = a.src_slot[c];
        if (slot >= a.n_slots) slot = 0;  // cannot happen: the frame kernel wrote it
        const uint64_t src = a.lod_tab[slot].y / 3u;
        uint2 sc = a.slot_chunks[slot];
        sc.y &= ~kSlotAllCached;
        Mat4x2 model;  // the command's model matrix in registers, every entry twice: one load per command, not per product
        {
            const float4* mp = reinterpret_cast<const float4*>(a.model + (size_t)(first_instance - a.draw_index_base) * 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 col = __ldg(mp + i);
                model.m[4 * i] = make_float2(col.x, col.x); model.m[4 * i + 1] = make_float2(col.y, col.y);
                model.m[4 * i + 2] = make_float2(col.z, col.z); model.m[4 * i + 3] = make_float2(col.w, col.w);
            }
        }
        __syncthreads();  // everyone has read s_cmd
        // Range checks once per command, not per triangle: the command's index range must lie inside the
        // index buffer, and an index is in range iff it is below v_limit (32-bit compare).
        bool bad = 3ull * (src + n_tri) > a.n_indices || vertex_offset < 0 || (uint64_t)vertex_offset >= a.n_vertices;
        if (bad && tid == 0) atomicOr(a.error_flag, 2u);
        const uint64_t room = a.n_vertices - (uint64_t)(vertex_offset < 0 ? 0 : vertex_offset);
        const uint32_t v_limit = room > 0xffffffffull ? 0xffffffffu : (uint32_t)room;
        const uint32_t* __restrict__ ib = a.indices + 3ull * src;
        const float* __restrict__ vb = a.positions + 3ll * vertex_offset;
        uint32_t running = 0, step = 0;
        for (uint32_t ci = sc.x; ci < sc.x + sc.y && !bad; ++ci) {
            const uint4 ch = a.chunks[ci];  // tri_begin, tri_count, v_min, v_count
            const uint32_t t_end = min(ch.x + ch.y, n_tri);
            const bool cached = kWindow && ch.w != 0u && ch.z + ch.w <= v_limit;
            if (cached) {
                // the chunk's vertex window through model and pv, two vertices per thread and trip
                for (uint32_t v = 2 * tid; v < ch.w; v += 2 * kTriThreads) {
                    const uint32_t v1 = v + 1 < ch.w ? v + 1 : v;
                    const float* q0 = vb + 3ull * (ch.z + v);
                    const float* q1 = vb + 3ull * (ch.z + v1);
                    float2 cx, cy, cw;
                    clip2(model, a.pv, make_float2(__ldg(q0), __ldg(q1)), make_float2(__ldg(q0 + 1), __ldg(q1 + 1)), make_float2(__ldg(q0 + 2), __ldg(q1 + 2)), cx, cy, cw);
                    s_cx[v] = cx.x; s_cy[v] = cy.x; s_cw[v] = cw.x;
                    s_cx[v1] = cx.y; s_cy[v1] = cy.y; s_cw[v1] = cw.y;
                }
                __syncthreads();
            }
            for (uint32_t t0 = ch.x; t0 < t_end; t0 += kTriThreads * kTriItems, ++step) {
                bool keep[kTriItems];
                uint32_t ix[kTriItems][3], b[kTriItems];
                // all loads of the step first (index triples, then the corners), then the math
#pragma unroll
                for (int j = 0; j < kTriItems; ++j) {  // item-major: triangle t0 + j * 128 + tid (coalesced index loads)
                    const uint32_t t = t0 + j * kTriThreads + tid;
                    const bool act = t < t_end;
#pragma unroll
                    for (int k = 0; k < 3; ++k) ix[j][k] = act ? __ldg(ib + 3u * t + k) : 0u;
                    keep[j] = act;
                }
                bool culled0, culled1;
                if (cached) {
                    float2 cx[3], cy[3], cw[3];
#pragma unroll
                    for (int j = 0; j < kTriItems; ++j) {
                        // the chunk table was built from this index buffer: an index outside the window means the buffers changed under us
                        const bool ok = ix[j][0] - ch.z < ch.w && ix[j][1] - ch.z < ch.w && ix[j][2] - ch.z < ch.w;
                        if (keep[j] && !ok) atomicOr(a.error_flag, 2u);
                        keep[j] = keep[j] && ok;
                    }
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        const uint32_t i0 = keep[0] ? ix[0][k] - ch.z : 0u, i1 = keep[1] ? ix[1][k] - ch.z : 0u;
                        cx[k] = make_float2(s_cx[i0], s_cx[i1]);
                        cy[k] = make_float2(s_cy[i0], s_cy[i1]);
                        cw[k] = make_float2(s_cw[i0], s_cw[i1]);
                    }
                    cull_from_clip(cx, cy, cw, culled0, culled1);
                } else {
                    float v[kTriItems][3][3];
#pragma unroll
                    for (int j = 0; j < kTriItems; ++j) {
                        const bool ok = ix[j][0] < v_limit && ix[j][1] < v_limit && ix[j][2] < v_limit;
                        if (keep[j] && !ok) atomicOr(a.error_flag, 2u);
                        keep[j] = keep[j] && ok;
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            const float* q = vb + 3ull * (keep[j] ? ix[j][k] : 0u);  // a safe address for inactive lanes
                            v[j][k][0] = __ldg(q);
                            v[j][k][1] = __ldg(q + 1);
                            v[j][k][2] = __ldg(q + 2);
                        }
                    }
                    static_assert(kTriItems == 2, "tri_culled2 works on the two triangles of a step");
                    tri_culled2(model, a.pv, v, culled0, culled1);
                }
                keep[0] = keep[0] && !culled0;
                keep[1] = keep[1] && !culled1;
                // per-warp kept counts, double-buffered by step parity: one barrier per step is enough
                uint32_t (*sw)[kTriThreads / 32] = s_warp[step & 1u];
#pragma unroll
                for (int j = 0; j < kTriItems; ++j) {
                    b[j] = __ballot_sync(kFullMask, keep[j]);
                    if (lane == 0) sw[j][warp] = __popc(b[j]);
                }
                __syncthreads();  // also: every corner of this step has been read from the vertex window
                uint32_t base = running;
#pragma unroll
                for (int j = 0; j < kTriItems; ++j) {
                    uint32_t before = 0, total = 0;
#pragma unroll
                    for (int w = 0; w < kTriThreads / 32; ++w) {
                        const uint32_t x = sw[j][w];
                        before += (uint32_t)w < warp ? x : 0u;
                        total += x;
                    }
                    if (keep[j]) {
                        const uint64_t dst = 3ull * ((uint64_t)out_first + base + before + __popc(b[j] & ((1u << lane) - 1u)));
                        if (dst + 2 < a.out_capacity) {
                            a.out_indices[dst] = ix[j][0];
                            a.out_indices[dst + 1] = ix[j][1];
                            a.out_indices[dst + 2] = ix[j][2];
                        } else {
                            atomicOr(a.error_flag, 4u);  // CulledIndexBuffer too small
                        }
                    }
                    base += total;
                }
                running = base;
            }
        }
        if (tid == 0) a.refined[c] = 3u * running;
    }
}

// compact_draw_stream.comp:34-63 on the refined commands: keep indexCount > 0, stable, dense
// from slot 0, count written by the last tile.  Single pass, decoupled look-back over tiles of
// 1024 commands; a descriptor is four u64 {flag:2 | kept commands:32, aggregate index sum,
// inclusive index sum, pad}, cleared by the host beforehand.
struct CompactArgs {
    const uint32_t* cmds;      // 5 words per command (input list)
    const uint32_t* refined;   // refined indexCount per command
    const uint32_t* count;     // input count (device)
    uint32_t* out_cmds;        // compacted list
    uint32_t* out_count;
    uint64_t* kept_indices;  // sum of refined indexCount
    uint64_t* status;        // [4 * n_tiles] zeroed
    uint32_t* ticket;
    uint32_t cap_cmds;
};

__global__ void __launch_bounds__(kCompactThreads) compact_draw_stream(const __grid_constant__ CompactArgs a) {
    __shared__ uint32_t s_tile, s_warp[kCompactThreads / 32], s_base;
    __shared__ uint64_t s_sum[kCompactThreads / 32];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    uint32_t count = *a.count;
    if (count > a.cap_cmds) count = a.cap_cmds;
    const uint32_t n_tiles = (count + kCompactTile - 1) / kCompactTile;
    if (n_tiles == 0) {
        if (blockIdx.x == 0 && tid == 0) {
            *a.out_count = 0;
            *a.kept_indices = 0;
        }
        return;
    }
    for (;;) {
        if (tid == 0) s_tile = atomicAdd(a.ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) break;
        // blocked arrangement: thread owns kCompactItems consecutive commands (stable order)
        const uint32_t first = tile * kCompactTile + tid * kCompactItems;
        uint32_t len[kCompactItems], mine = 0;
        uint64_t sum = 0;
#pragma unroll
        for (int j = 0; j < kCompactItems; ++j) {
            len[j] = first + j < count ? a.refined[first + j] : 0u;
            mine += len[j] > 0u ? 1u : 0u;
            sum += len[j];
        }
        const uint32_t incl = warp_incl_scan_u32(mine, lane);
        const uint64_t wsum = warp_sum_u64(sum);
        if (lane == 31) s_warp[warp] = incl;
        if (lane == 0) s_sum[warp] = wsum;
        __syncthreads();
        uint32_t before = 0, total = 0;
        uint64_t tsum = 0;
#pragma unroll
        for (int w = 0; w < kCompactThreads / 32; ++w) {
            before += (uint32_t)w < warp ? s_warp[w] : 0u;
            total += s_warp[w];
            tsum += s_sum[w];
        }
        // look-back (warp 0): value = kept commands (32 bits); the index sum rides in a second word
        if (warp == 0) {
            uint64_t* st = a.status + 4ull * tile;
            if (lane == 0) {
                st_relaxed(st + (tile == 0 ? 2 : 1), tsum);
                __threadfence();
                st_relaxed(st, ((tile == 0 ? 2ull : 1ull) << 32) | total);
            }
            uint32_t excl = 0;
            uint64_t excl_sum = 0;
            int p = (int)tile - 1 - (int)lane;
            bool done = tile == 0;
            while (!done) {
                uint64_t d = p >= 0 ? ld_relaxed(a.status + 4ull * p) : (2ull << 32);
                while (__any_sync(kFullMask, (d >> 32) == 0ull)) {
                    __nanosleep(100);
                    d = p >= 0 ? ld_relaxed(a.status + 4ull * p) : (2ull << 32);
                }
                __threadfence();
                const uint32_t incl_mask = __ballot_sync(kFullMask, (d >> 32) == 2ull);
                const uint32_t stop = incl_mask ? (uint32_t)__ffs(incl_mask) - 1u : 31u;
                if (lane <= stop) {
                    excl += (uint32_t)d;
                    // word 1 = aggregate index sum, word 2 = inclusive index sum (written before its flag)
                    excl_sum += p >= 0 ? ld_relaxed(a.status + 4ull * p + ((d >> 32) == 2ull ? 2 : 1)) : 0ull;
                }
                done = incl_mask != 0u;
                p -= 32;
            }
            excl = __reduce_add_sync(kFullMask, excl);
            excl_sum = warp_sum_u64(excl_sum);
            if (lane == 0) {
                s_base = excl;
                if (tile != 0) {
                    st_relaxed(st + 2, excl_sum + tsum);
                    __threadfence();
                    st_relaxed(st, (2ull << 32) | (excl + total));
                }
                if (tile == n_tiles - 1) {
                    *a.out_count = excl + total;
                    *a.kept_indices = excl_sum + tsum;
                }
            }
        }
        __syncthreads();
        uint32_t pos = s_base + before + incl - mine;
#pragma unroll
        for (int j = 0; j < kCompactItems; ++j) {
            if (len[j] > 0u) {
                const uint32_t* src = a.cmds + (size_t)(first + j) * 5;
                uint32_t* dst = a.out_cmds + (size_t)pos * 5;
                dst[0] = len[j];
                dst[1] = src[1];
                dst[2] = src[2];
                dst[3] = src[3];
                dst[4] = src[4];
                ++pos;
            }
        }
        __syncthreads();
    }
}

}  // namespace rcull
