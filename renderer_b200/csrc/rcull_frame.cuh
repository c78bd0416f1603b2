// rcull_frame.cuh -- the fused transform + AABB + multi-view cull + compaction
// kernel (SURVEY 8a rows A1-A7 in one launch).
//
// Shape: persistent CTAs (grid = SMs x resident CTAs), tiles of kTile entities
// handed out by an atomic ticket, SoA input tiles staged into shared memory by
// the TMA engine (cp.async.bulk + mbarrier, two stages: tile i+1 streams in
// while tile i is processed), one thread per entity for the arithmetic, warp
// ballot/popc compaction into per-warp visible lists in shared memory, and a
// single-pass decoupled-lookback scan over tiles that carries, per view, the
// survivor count and the running sum of index_len (firstIndex).
#pragma once
#include "rcull_device.cuh"

namespace rcull {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 4;                  // entities per thread per tile
constexpr int kTile = kThreads * kItems;   // 1024 entities
constexpr int kWarpSpan = 32 * kItems;     // entities owned by one warp
constexpr int kStages = 2;
constexpr int kMaxViews = 8;
constexpr int kMaxLods = 6;
constexpr uint32_t kStageBytes = kTile * (16 + 16 + 4);  // pos_scale + rot + mesh

// Tile status: 4 u64 per (view, tile): [0] = epoch:30 | flag:2 | count:32,
// [1] = index sum of the tile alone, [2] = inclusive index sum, [3] pad.
constexpr uint64_t kFlagAgg = 1, kFlagIncl = 2;
constexpr int kStatusWords = 4;

struct FrameArgs {
    const float4* pos_scale;
    const float4* rot;
    const uint32_t* mesh;
    const float4* mesh_aabb;  // [M][2]: (min.xyz, n_lods), (max.xyz, -)
    const uint4* lod_tab;     // [M*kMaxLods]: index_len, first_index, vertex_offset, 0
    float4* model;
    float* aabb;              // optional [n][6]
    uint8_t* flags[kMaxViews];  // optional [n]
    uint32_t* cmds[kMaxViews];  // 5 words per command
    uint32_t* visible[kMaxViews];
    uint32_t* counts;
    unsigned long long* totals;
    uint64_t* status;
    uint32_t* ticket;
    uint32_t* error_flag;
    uint32_t ticket_base;
    uint32_t epoch;
    uint32_t n;
    uint32_t n_tiles;
    uint32_t n_meshes;
    uint32_t draw_index_base;
    uint32_t cap_visible;
    float cam[3];
    float lod_d2_max;  // lod 1 iff |cam-pos|^2 > lod_d2_max  (== sqrt(d2) > lod_distance)
    float planes[kMaxViews][6][4];
    float absn[kMaxViews][6][3];
};

struct FrameSmem {
    // stage buffers first: 16-B aligned bulk-copy destinations
    float4 pos_scale[kStages][kTile];
    float4 rot[kStages][kTile];
    uint32_t mesh[kStages][kTile];          // overwritten in place by the (mesh, lod) slot
    float4 mat[kWarps][128];                // per-warp 2 KB matrix transpose buffer
    uint64_t wlen[kMaxViews][kWarps];
    uint64_t excl_len[kMaxViews];
    uint64_t bar[kStages];
    uint32_t wcnt[kMaxViews][kWarps];
    uint32_t excl_cnt[kMaxViews];
    uint32_t tile_of_stage[kStages];
    // followed by: uint16_t list[V][kTile]
};

template <int V>
constexpr size_t frame_smem_bytes() {
    return sizeof(FrameSmem) + sizeof(uint16_t) * V * kTile;
}

// Lookback for one view, run by one whole warp.  Returns the exclusive prefix
// (count, index sum) of `tile` and publishes this tile's words.
__device__ __forceinline__ void lookback(uint64_t* status_view, uint32_t tile, uint32_t epoch, uint32_t agg_cnt,
                                         uint64_t agg_len, uint32_t& excl_cnt, uint64_t& excl_len) {
    const uint32_t lane = lane_id();
    const uint64_t tag = (uint64_t)epoch << 34;
    uint64_t* mine = status_view + (size_t)tile * kStatusWords;
    uint32_t ecnt = 0;
    uint64_t elen = 0;
    if (tile != 0) {
        if (lane == 0) {
            st_relaxed(mine + 1, agg_len);
            st_release(mine, tag | (kFlagAgg << 32) | agg_cnt);
        }
        int pred = (int)tile - 1 - (int)lane;
        while (true) {
            uint64_t a = tag | (kFlagIncl << 32);  // virtual tile -1: inclusive prefix 0
            const uint64_t* w = status_view + (size_t)(pred < 0 ? 0 : pred) * kStatusWords;
            bool ok = true;
            do {  // spin until every real predecessor in the window has published for this frame
                if (pred >= 0) {
                    a = ld_acquire(w);
                    ok = (a >> 34) == (uint64_t)epoch && ((a >> 32) & 3) != 0;
                }
            } while (!__all_sync(kFullMask, ok));
            const uint32_t flag = (uint32_t)(a >> 32) & 3u;
            const uint32_t incl_mask = __ballot_sync(kFullMask, flag == kFlagIncl);
            const uint32_t first = incl_mask ? (uint32_t)__ffs(incl_mask) - 1u : 32u;
            const bool use = lane <= first;
            uint32_t c = use ? (uint32_t)a : 0u;
            uint64_t l = 0;
            if (use && pred >= 0) l = ld_relaxed(w + (flag == kFlagIncl ? 2 : 1));
            ecnt += __reduce_add_sync(kFullMask, c);
            elen += warp_sum_u64(l);
            if (incl_mask) break;
            pred -= 32;
        }
    }
    if (lane == 0) {
        st_relaxed(mine + 2, elen + agg_len);
        st_release(mine, tag | (kFlagIncl << 32) | (uint32_t)(ecnt + agg_cnt));
    }
    excl_cnt = ecnt;
    excl_len = elen;
}

template <int V>
__global__ void __launch_bounds__(kThreads, 2) transform_cull_compact(const __grid_constant__ FrameArgs args) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    FrameSmem& sm = *reinterpret_cast<FrameSmem*>(smem_raw);
    uint16_t* list = reinterpret_cast<uint16_t*>(smem_raw + sizeof(FrameSmem));  // [V][kTile]

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31u;
    const uint32_t warp = tid >> 5;
    const uint32_t lt = lanemask_lt();

    auto issue_tile = [&](uint32_t stage, uint32_t tile) {
        uint64_t* bar = &sm.bar[stage];
        mbar_expect_tx(bar, kStageBytes);
        bulk_g2s(sm.pos_scale[stage], args.pos_scale + (size_t)tile * kTile, kTile * 16, bar);
        bulk_g2s(sm.rot[stage], args.rot + (size_t)tile * kTile, kTile * 16, bar);
        bulk_g2s(sm.mesh[stage], args.mesh + (size_t)tile * kTile, kTile * 4, bar);
    };

    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        fence_mbar_init();
        const uint32_t t0 = atomicAdd(args.ticket, 1u) - args.ticket_base;
        sm.tile_of_stage[0] = t0;
        if (t0 < args.n_tiles) issue_tile(0, t0);
    }
    __syncthreads();

    for (uint32_t it = 0;; ++it) {
        const uint32_t stage = it & 1u;
        const uint32_t tile = sm.tile_of_stage[stage];
        if (tile >= args.n_tiles) break;
        if (tid == 0) {  // prefetch the next tile into the other stage
            const uint32_t t1 = atomicAdd(args.ticket, 1u) - args.ticket_base;
            sm.tile_of_stage[stage ^ 1u] = t1;
            if (t1 < args.n_tiles) {
                fence_proxy_async();
                issue_tile(stage ^ 1u, t1);
            }
        }
        mbar_wait(&sm.bar[stage], (it >> 1) & 1u);

        const uint32_t tile_base = tile * kTile;
        uint32_t wcnt[V];
        uint64_t wlen[V];
#pragma unroll
        for (int k = 0; k < V; ++k) {
            wcnt[k] = 0;
            wlen[k] = 0;
        }

        // ---------------- phase A: one entity per thread per item ----------------
#pragma unroll 1
        for (int item = 0; item < kItems; ++item) {
            const uint32_t e = warp * kWarpSpan + item * 32 + lane;  // slot in tile
            const uint32_t g = tile_base + e;                         // slot in frame
            const bool valid = g < args.n;
            float4 ps = sm.pos_scale[stage][e];
            float4 q = sm.rot[stage][e];
            uint32_t mid = sm.mesh[stage][e];
            if (!valid) {
                ps = make_float4(0.f, 0.f, 0.f, 1.f);
                q = make_float4(0.f, 0.f, 0.f, 1.f);
                mid = 0;
            }
            if (mid >= args.n_meshes) {
                atomicOr(args.error_flag, 1u);
                mid = 0;
            }
            const Mat34 m = model_matrix(ps, q);

            // model matrix out: transpose through the warp's smem buffer so that
            // every global store instruction covers 512 contiguous bytes.
            {
                float4* buf = sm.mat[warp];
                const uint32_t sw = (lane >> 1) & 3u;
                buf[lane * 4 + (0u ^ sw)] = make_float4(m.m00, m.m10, m.m20, 0.f);
                buf[lane * 4 + (1u ^ sw)] = make_float4(m.m01, m.m11, m.m21, 0.f);
                buf[lane * 4 + (2u ^ sw)] = make_float4(m.m02, m.m12, m.m22, 0.f);
                buf[lane * 4 + (3u ^ sw)] = make_float4(m.tx, m.ty, m.tz, 1.f);
                __syncwarp();
                const uint32_t g0 = tile_base + warp * kWarpSpan + item * 32;  // first entity of this pass
                float4* out = args.model + (size_t)g0 * 4;
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const uint32_t p = r * 32 + lane;
                    const uint32_t el = p >> 2;
                    const uint32_t col = (p & 3u) ^ ((el >> 1) & 3u);
                    const float4 v = buf[p];
                    if (g0 + el < args.n) __stcs(out + el * 4 + col, v);
                }
                __syncwarp();
            }

            // A2: world AABB
            const float4 mn = __ldg(args.mesh_aabb + 2 * mid);
            const float4 mx = __ldg(args.mesh_aabb + 2 * mid + 1);
            float lox, hix, loy, hiy, loz, hiz;
            row_extent(m.m00, m.m01, m.m02, m.tx, mn, mx, lox, hix);
            row_extent(m.m10, m.m11, m.m12, m.ty, mn, mx, loy, hiy);
            row_extent(m.m20, m.m21, m.m22, m.tz, mn, mx, loz, hiz);
            float minx, maxx, miny, maxy, minz, maxz;
            aabb_store_form(lox, hix, minx, maxx);
            aabb_store_form(loy, hiy, miny, maxy);
            aabb_store_form(loz, hiz, minz, maxz);
            if (args.aabb != nullptr && valid) {
                float* a = args.aabb + (size_t)g * 6;
                a[0] = minx; a[1] = miny; a[2] = minz;
                a[3] = maxx; a[4] = maxy; a[5] = maxz;
            }
            // A3 inputs: ncollide half_extents() / center() of the STORED box
            const float hx = __fmul_rn(__fsub_rn(maxx, minx), 0.5f), cx = __fmul_rn(__fadd_rn(minx, maxx), 0.5f);
            const float hy = __fmul_rn(__fsub_rn(maxy, miny), 0.5f), cy = __fmul_rn(__fadd_rn(miny, maxy), 0.5f);
            const float hz = __fmul_rn(__fsub_rn(maxz, minz), 0.5f), cz = __fmul_rn(__fadd_rn(minz, maxz), 0.5f);

            // A4: pick_lod, then the (mesh, lod) slot replaces the mesh id in smem
            const float dx = __fsub_rn(args.cam[0], ps.x), dy = __fsub_rn(args.cam[1], ps.y),
                        dz = __fsub_rn(args.cam[2], ps.z);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            const uint32_t n_lods = __float_as_uint(mn.w);
            const uint32_t lod = (d2 > args.lod_d2_max && n_lods > 1u) ? 1u : 0u;
            const uint32_t slot = mid * kMaxLods + lod;
            sm.mesh[stage][e] = slot;

            // A3 per view + ballot compaction into the warp's list segment
            bool any_vis = false;
            bool vis[V];
#pragma unroll
            for (int k = 0; k < V; ++k) {
                bool outside = false;
#pragma unroll
                for (int p = 0; p < 6; ++p) {
                    const float ee = __fadd_rn(
                        __fadd_rn(__fmul_rn(hx, args.absn[k][p][0]), __fmul_rn(hy, args.absn[k][p][1])),
                        __fmul_rn(hz, args.absn[k][p][2]));
                    const float s = __fadd_rn(
                        __fadd_rn(__fmul_rn(args.planes[k][p][0], cx), __fmul_rn(args.planes[k][p][2], cz)),
                        __fadd_rn(__fmul_rn(args.planes[k][p][1], cy), args.planes[k][p][3]));
                    outside = outside || (__fsub_rn(s, ee) > 0.0f);
                }
                vis[k] = valid && !outside;
                any_vis = any_vis || vis[k];
                if (args.flags[k] != nullptr && valid) args.flags[k][g] = outside ? 1 : 0;
            }
            uint32_t len = 0;
            if (any_vis) len = __ldg(args.lod_tab + slot).x;
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const uint32_t b = __ballot_sync(kFullMask, vis[k]);
                if (vis[k]) {
                    list[k * kTile + warp * kWarpSpan + wcnt[k] + __popc(b & lt)] = (uint16_t)e;
                    wlen[k] += len;
                }
                wcnt[k] += __popc(b);
            }
        }

        // warp aggregates -> smem
#pragma unroll
        for (int k = 0; k < V; ++k) {
            const uint64_t t = warp_sum_u64(wlen[k]);
            if (lane == 0) {
                sm.wcnt[k][warp] = wcnt[k];
                sm.wlen[k][warp] = t;
            }
        }
        __syncthreads();

        // ---------------- tile prefix: warp k looks back for view k ----------------
        for (int k = warp; k < V; k += kWarps) {
            uint32_t c = lane < kWarps ? sm.wcnt[k][lane] : 0u;
            uint64_t l = lane < kWarps ? sm.wlen[k][lane] : 0ull;
            c = __reduce_add_sync(kFullMask, c);
            l = warp_sum_u64(l);
            uint32_t ec;
            uint64_t el;
            lookback(args.status + (size_t)k * args.n_tiles * kStatusWords, tile, args.epoch, c, l, ec, el);
            if (lane == 0) {
                sm.excl_cnt[k] = ec;
                sm.excl_len[k] = el;
                if (tile == args.n_tiles - 1) {
                    args.counts[k] = ec + c;
                    args.totals[k] = el + l;
                }
            }
        }
        __syncthreads();

        // ---------------- phase B: emit commands from the warp's list ----------------
#pragma unroll
        for (int k = 0; k < V; ++k) {
            const uint32_t cw = wcnt[k];
            if (cw == 0) continue;
            uint32_t cnt_base = sm.excl_cnt[k];
            uint32_t len_base = (uint32_t)sm.excl_len[k];
            for (uint32_t w = 0; w < warp; ++w) {
                cnt_base += sm.wcnt[k][w];
                len_base += (uint32_t)sm.wlen[k][w];
            }
            uint32_t* cmds = args.cmds[k];
            uint32_t* visible = args.visible[k];
            for (uint32_t j0 = 0; j0 < cw; j0 += 32) {
                const uint32_t j = j0 + lane;
                const bool act = j < cw;
                uint32_t e = 0;
                uint4 L = make_uint4(0, 0, 0, 0);
                if (act) {
                    e = list[k * kTile + warp * kWarpSpan + j];
                    L = __ldg(args.lod_tab + sm.mesh[stage][e]);
                }
                const uint32_t incl = warp_incl_scan_u32(L.x, lane);
                const uint32_t first_index = len_base + incl - L.x;
                len_base += __shfl_sync(kFullMask, incl, 31);
                const uint32_t pos = cnt_base + j;
                if (act && pos < args.cap_visible) {
                    const uint32_t draw_index = args.draw_index_base + tile_base + e;
                    uint32_t* c = cmds + (size_t)pos * 5;
                    c[0] = L.x;          // indexCount (entity-level upper bound)
                    c[1] = 1u;           // instanceCount
                    c[2] = first_index;  // running sum of earlier survivors' index_len
                    c[3] = L.z;          // vertexOffset
                    c[4] = draw_index;   // firstInstance
                    visible[pos] = draw_index;
                }
            }
        }
        __syncthreads();  // stage buffers, list and aggregates are free again
    }
}

}  // namespace rcull
