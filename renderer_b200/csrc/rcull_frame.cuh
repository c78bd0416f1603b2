// rcull_frame.cuh -- the fused transform + AABB + multi-view cull + compaction
// kernel (SURVEY 8a rows A1-A7 in one launch).
//
// Shape: persistent, warp-specialised CTAs (grid = SMs x resident CTAs; one CTA of 15 + 1
// warps per SM in the default build, tile = 1920 entities).
//   * COMPUTE warps: one thread per entity, 4 entities per thread per tile.
//     Inputs come out of shared memory (staged there by TMA), model matrices go
//     back through a swizzled shared-memory box and a 2-D TMA tensor store, the
//     plane tests run planes-outer / entities-inner so each plane is fetched once
//     per 4 entities, survivors are ballot-compacted into per-warp lists.
//     Thread 0 is also the producer: it draws tile tickets and issues the TMA loads
//     (two stages).  The commands of a tile are emitted by the compute warps a few
//     tiles later, once its output offsets are known.
//   * 1 SCAN warp: for every finished tile, the decoupled look-back (count and index
//     sum per view in one 16-byte descriptor) as a non-blocking state machine, while
//     the compute warps are already working on the next tiles.
// Warps synchronise through mbarriers only; there is no CTA-wide barrier in the
// tile loop.
#pragma once
#include <cuda.h>  // CUtensorMap (type only; the encoder is resolved at run time)
#include <cuda/std/type_traits>

#include "rcull_aux.cuh"  // RecordArgs: the range-shard merge sends its records from this kernel
#include "rcull_device.cuh"

namespace rcull {

#ifndef RCULL_COMPUTE_WARPS
#define RCULL_COMPUTE_WARPS 15
#endif
#ifndef RCULL_VIEW_UNROLL
#define RCULL_VIEW_UNROLL 1
#endif
#ifndef RCULL_SCAN_IDLE_NS
#define RCULL_SCAN_IDLE_NS 400  // scan warp: pause between passes that made no progress
#endif
#ifndef RCULL_FRAME_REGS
#define RCULL_FRAME_REGS 112  // 512 x 112 registers leave 8 K of the SM's 64 K for a resident exchange CTA (128 x 64)
#endif
constexpr int kViewUnroll = RCULL_VIEW_UNROLL;  // 1: the view loop stays rolled
constexpr int kComputeWarps = RCULL_COMPUTE_WARPS;  // + 1 scan warp; 7 + 1: two CTAs fill an SM's register file at 112 regs, 15 + 1: one CTA does
constexpr int kCtasPerSm = kComputeWarps <= 7 ? 2 : 1;
// 228 KB per SM, 1 KB reserved per CTA; the one-CTA shape leaves a second reserved KB free so that a small
// shared-memory-free CTA of another stream (the merge's peer-store kernel: 128 threads x 64 registers)
// can be resident beside it and overlap with the frame
constexpr size_t kSmemPerCta = (228 * 1024 - (kCtasPerSm == 1 ? 2 : kCtasPerSm) * 1024) / kCtasPerSm;
constexpr int kThreads = (kComputeWarps + 1) * 32;  // + scan warp
#ifndef RCULL_ITEMS
#define RCULL_ITEMS 4
#endif
constexpr int kItems = RCULL_ITEMS;                 // entities per thread per tile
constexpr int kPairs = kItems / 2;                  // entity pairs whose plane tests share packed (f32x2) instructions
constexpr bool kOddItem = (kItems % 2) != 0;        // the last entity of an odd count runs the same arithmetic unpacked
constexpr int kTile = kComputeWarps * 32 * kItems;  // 1920 entities with 15 compute warps
constexpr int kWarpSpan = 32 * kItems;              // entities owned by one compute warp
constexpr int kStages = 2;
#ifndef RCULL_MAX_BUFS
#define RCULL_MAX_BUFS 4
#endif
constexpr int kMaxBufs = RCULL_MAX_BUFS;   // survivor bookkeeping ring (tiles whose commands are not emitted yet)
constexpr int kRing = 8;      // tile-id ring: bookkeeping ring + the tile being loaded + slack
constexpr int kMaxViews = 8;
constexpr int kMaxLods = 6;
constexpr uint32_t kStageBytes = kTile * (16 + 16 + 4);  // pos_scale + rot + mesh
constexpr int kMatBoxBytes = 32 * 64;                    // 32 model matrices = one TMA store box

#ifdef RCULL_SCAN_STATS  // diagnostic build only (profiles/scan_stats.py): where the scan and the waits spend their time
__device__ unsigned long long g_scan_stats[16];
#define RCULL_STAT(i, v) atomicAdd(&rcull::g_scan_stats[i], (unsigned long long)(v))
#define RCULL_CLOCK() clock64()
#else
#define RCULL_STAT(i, v) ((void)0)
#define RCULL_CLOCK() 0ll
#endif

// Tile descriptor: x = epoch:30 | flag:2 | count:32, y = index sum.
constexpr uint64_t kFlagAgg = 1, kFlagIncl = 2;

// One frustum plane as the kernel reads it: two 16-byte words.  A packed (f32x2) product takes the plane value as a
// SCALAR operand broadcast to both halves (SASS `FMUL2 R, Ra.F32x2, Rb.F32`), so nothing is stored twice.
struct PlaneC {
    float ax, ay, az, px;  // |p.x|, |p.y|, |p.z|, p.x
    float py, pz, pw, pad;
};

struct FrameArgs {
    const float4* pos_scale;
    const float4* rot;
    const uint32_t* mesh;
    const float4* mesh_aabb;  // [M][2]: (min.xyz, n_lods), (max.xyz, index_len of lod 0)
    const uint32_t* mesh_len1;  // [M]: index_len of lod 1 (pick_lod only ever returns lod 0 or 1)
    const uint4* lod_tab;     // [M*kMaxLods]: index_len, first_index, vertex_offset, 0
    float* aabb;              // optional [n][6]
    uint8_t* flags[kMaxViews];  // optional [n]
    uint32_t* cmds[kMaxViews];  // 5 words per command
    uint32_t* visible[kMaxViews];
    uint32_t* src_slot;  // optional, view 0 only: (mesh, lod) slot behind each command (work list of the triangle stage)
    uint32_t* counts;
    unsigned long long* totals;
    uint32_t max_views;
    ulonglong2* status;   // [views][n_tiles]
    uint32_t* ticket;
    uint32_t* error_flag;
    uint32_t ticket_base;
    uint32_t epoch;
    uint32_t n;
    uint32_t n_tiles;
    uint32_t n_meshes;
    uint32_t draw_index_base;
    uint32_t cap_visible;
    uint32_t per_mesh;   // RCULL_PER_MESH_INSTANCED: emit (slot key, draw_index) pairs for the bucketing passes
    uint32_t len_small;  // every index_len in the mesh table is < 2^25: 128 of them add up in 32 bits
    uint32_t planes_ftz_safe;  // every plane normal component is zero or >= 2^-63 in magnitude (packed-sum plane tests allowed)
    float cam[3];
    float lod_d2_max;  // lod 1 iff |cam-pos|^2 > lod_d2_max  (== sqrt(d2) > lod_distance)
    PlaneC planes[kMaxViews][6];
    RecordArgs rec;  // rec.world > 0: a range-shard merge is attached -- survivors leave as 8-byte records to every rank's window
};

struct FrameSmem {
    // TMA destinations / sources first (1024-B aligned for the 128-B swizzle)
    float4 mat[kComputeWarps][kMatBoxBytes / 16];  // per-warp model-matrix store box
    float4 pos_scale[kStages][kTile];
    float4 rot[kStages][kTile];
    uint32_t mesh[kStages][kTile];
    // Survivor bookkeeping is a ring of frame_bufs<V>() tiles (buffer = iteration % bufs): the
    // scan warp resolves the offsets of tile i while the compute warps run several tiles ahead.
    uint64_t full[kStages];
    uint64_t empty[kStages];
    uint64_t done[kMaxBufs];   // compute warps -> scan warp: lists and aggregates of a tile are written
    uint64_t bases[kMaxBufs];  // scan warp -> compute warps: output offsets of that tile are known
    uint64_t wlen[kMaxBufs][kMaxViews][kComputeWarps];   // per-warp index sums
    uint32_t wcnt[kMaxBufs][kMaxViews][kComputeWarps];   // per-warp survivor counts
    uint32_t adj_cnt[kMaxBufs][kMaxViews][kComputeWarps];  // output slot of combined-list entry i of view k: adj_cnt + i
    uint32_t adj_len[kMaxBufs][kMaxViews][kComputeWarps];  // firstIndex of it: adj_len + exclusive len scan up to i
    uint64_t agg_len[kMaxBufs][kMaxViews];  // tile aggregates kept between publish and resolve
    uint32_t agg_cnt[kMaxBufs][kMaxViews];
    uint32_t tile_ring[kRing];  // tile id of iteration it at [it % kRing]
    // followed by: PlaneC planes[V][6] (a shared-memory copy: LDS.128 instead of indexed LDC), then
    // uint8_t wlist[bufs][kComputeWarps][V * kWarpSpan] -- per-warp combined survivor list,
    // view-major, entry = lod << 7 | slot in the warp's span (the view follows from the entry's
    // position and the per-view counts) --, uint8_t mid8[bufs][kTile] (mesh id of every entity of the tile, kept
    // until its commands are emitted) and, when it fits, a copy of the mesh table (float4 aabb[M][2],
    // uint4 lod01[M][2]: 64 bytes per mesh)
};

// Per bookkeeping buffer: the survivor lists (V bytes per entity) and one byte per entity with its mesh id (kept until
// the tile's commands are emitted, so that emission needs no global load when the mesh table is small).
template <int V>
__host__ __device__ constexpr int frame_bufs() {  // what still lets kCtasPerSm CTAs share an SM's shared memory
    // 4 buffers when they leave room for a useful mesh table behind them, else 3, never fewer than 2
    constexpr size_t fixed = sizeof(FrameSmem) + V * 6 * sizeof(PlaneC) + 1024, per_buf = (V + 1) * kTile;
    constexpr size_t b = (kSmemPerCta - fixed) / per_buf;
    constexpr size_t want0 = (kSmemPerCta - fixed) >= 4 * per_buf + 12 * 1024 ? 4 : 3;
    constexpr size_t want = want0 > (size_t)kMaxBufs ? (size_t)kMaxBufs : want0;
    return b > want ? (int)want : (int)b;
}
template <int V>
__host__ __device__ constexpr size_t frame_mid8_off() {  // mesh-id bytes, behind the survivor lists
    return sizeof(FrameSmem) + V * 6 * sizeof(PlaneC) + frame_bufs<V>() * V * kTile;
}
template <int V>
__host__ __device__ constexpr size_t frame_lists_end() {  // offset of the first byte behind the per-buffer bookkeeping
    return (frame_mid8_off<V>() + frame_bufs<V>() * kTile + 15) / 16 * 16;
}
// Shared-memory copy of the mesh table, per mesh: float4 aabb[2] (+ n_lods, index_len of lod 0), and for lods 0 and 1
// (the only ones pick_lod returns, helpers.rs:3-11) the uint4 (index_len, first_index, vertex_offset, 0) rows that
// emission needs: 64 bytes.  Mesh ids must fit the byte kept per entity.
constexpr uint32_t kMeshSmemBytes = 64;
template <int V>
__host__ __device__ constexpr uint32_t frame_mesh_cap() {
    constexpr size_t left = kSmemPerCta - 1024 - frame_lists_end<V>();
    return left / kMeshSmemBytes > 256 ? 256u : (uint32_t)(left / kMeshSmemBytes);
}
template <int V>
constexpr size_t frame_smem_bytes() {
    return frame_lists_end<V>() + frame_mesh_cap<V>() * kMeshSmemBytes + 1024;  // + slack to align the base
}
static_assert(frame_bufs<1>() >= 2 && frame_bufs<kMaxViews>() >= 2, "at least two tiles of survivor bookkeeping");

__device__ __forceinline__ FrameSmem& frame_smem(unsigned char* raw) {
    const uint32_t a = smem_u32(raw);
    return *reinterpret_cast<FrameSmem*>(raw + ((1024u - (a & 1023u)) & 1023u));
}

// ---------------------------------------------------------------------------------
// scan warp: decoupled look-back over tiles, all views of one tile together
// ---------------------------------------------------------------------------------
__device__ __forceinline__ bool desc_ready(const ulonglong2& d, uint32_t epoch) {
    return (uint32_t)(d.x >> 34) == epoch && ((d.x >> 32) & 3ull) != 0ull;
}


// The scan warp never blocks on another CTA.  It is a small state machine over the
// CTA's tiles in claim order:
//   publish  -- as soon as the compute warps are done with a tile, reduce their
//               per-warp counts / index sums and publish the tile AGGREGATE; this only
//               depends on the CTA's own work, so aggregates appear without delay;
//   resolve  -- for the oldest unresolved tile, one look-back round per pass: all views
//               together, 32*U predecessors per view with every load in flight at
//               once (CTAs run in waves, so the last tile of a wave walks back over a
//               whole wave of aggregates).  A window that is not fully published yet
//               is simply retried on the next pass.
// (A two-level variant -- groups of 64 tiles with milestone descriptors -- was measured
// slower: every tile then needs two dependent rounds; see DESIGN.md.)
template <int V>
__device__ __forceinline__ void scan_warp(const FrameArgs& args, FrameSmem& sm, uint32_t lane) {
    constexpr int U = V <= 4 ? 4 : 2;
    constexpr uint32_t B = frame_bufs<V>();
    const uint32_t epoch = args.epoch, n_tiles = args.n_tiles;
    const uint64_t tag = (uint64_t)epoch << 34;
    const ulonglong2 virt = make_ulonglong2(tag | (kFlagIncl << 32), 0ull);  // tile -1: inclusive prefix 0
    ulonglong2* const status = args.status;
    uint32_t pub = 0, res = 0;  // tiles published / resolved so far (iteration numbers)
    bool ended = false, active = false;
    uint32_t acc_c[V], tile = 0;
    uint64_t acc_l[V];
    bool open[V];
    int base = 0;
    long long t_pub[kMaxBufs] = {}, t_act = 0;
    while (true) {
        bool progress = false;
        // ---- publish ----
        if (!ended && pub - res < B && mbar_test(&sm.done[pub % B], (pub / B) & 1u)) {
            const uint32_t buf = pub % B, t = sm.tile_ring[pub % kRing];
            progress = true;
            if (t >= n_tiles) {
                ended = true;  // end marker, passed on by the compute warps
            } else {
                uint32_t vstart_c = 0, vstart_l = 0;
#pragma unroll
                for (int k = 0; k < V; ++k) {  // lane w stands for compute warp w
                    const uint32_t c = lane < kComputeWarps ? sm.wcnt[buf][k][lane] : 0u;
                    const uint64_t l = lane < kComputeWarps ? sm.wlen[buf][k][lane] : 0ull;
                    uint32_t ic = c;
                    uint64_t il = l;
#pragma unroll
                    for (int o = 1; o < kComputeWarps; o <<= 1) {
                        const uint32_t tc = __shfl_up_sync(kFullMask, ic, o);
                        const uint64_t tl = __shfl_up_sync(kFullMask, il, o);
                        if (lane >= (uint32_t)o) {
                            ic += tc;
                            il += tl;
                        }
                    }
                    if (lane < kComputeWarps) {
                        // (prefix over lower warps) - (start of view k in the warp's combined list);
                        // the tile's exclusive prefix is added when it is resolved
                        sm.adj_cnt[buf][k][lane] = (ic - c) - vstart_c;
                        sm.adj_len[buf][k][lane] = (uint32_t)(il - l) - vstart_l;
                    }
                    if (lane == kComputeWarps - 1) {
                        sm.agg_cnt[buf][k] = ic;
                        sm.agg_len[buf][k] = il;
                        st_relaxed_v2(status + (size_t)k * n_tiles + t, tag | ((t == 0 ? kFlagIncl : kFlagAgg) << 32) | ic, il);
                    }
                    vstart_c += c;
                    vstart_l += (uint32_t)l;
                }
                __syncwarp();
                t_pub[pub % B] = RCULL_CLOCK();
                ++pub;
            }
        }
        // ---- resolve ----
        if (res < pub) {
            if (!active) {
                tile = sm.tile_ring[res % kRing];
                base = (int)tile - 1 - (int)lane;  // this lane's nearest predecessor in the current window
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    acc_c[k] = 0;
                    acc_l[k] = 0;
                    open[k] = tile != 0;
                }
                active = true;
                t_act = RCULL_CLOCK();
            }
            if (lane == 0) RCULL_STAT(2, 1);
            ulonglong2 d[V][U];
#pragma unroll
            for (int k = 0; k < V; ++k)
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int p = base - 32 * u;
                    d[k][u] = (open[k] && p >= 0) ? ld_relaxed_v2(status + (size_t)k * n_tiles + p) : virt;
                }
            // all views walk the same windows: a round only counts when every open view can use it
            bool ok = true;
#pragma unroll
            for (int k = 0; k < V; ++k)
#pragma unroll
                for (int u = 0; u < U; ++u) ok = ok && desc_ready(d[k][u], epoch);
            if (__all_sync(kFullMask, ok)) {
                progress = true;
                if (lane == 0) RCULL_STAT(3, 1);
                bool any_open = false;
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    if (!open[k]) continue;
                    bool stopped = false;
#pragma unroll
                    for (int u = 0; u < U; ++u) {
                        const uint32_t incl_mask = __ballot_sync(kFullMask, ((uint32_t)(d[k][u].x >> 32) & 3u) == kFlagIncl);
                        if (!stopped) {
                            const uint32_t stop = incl_mask ? (uint32_t)__ffs(incl_mask) - 1u : 31u;
                            if (lane <= stop) {
                                acc_c[k] += (uint32_t)d[k][u].x;
                                acc_l[k] += d[k][u].y;
                            }
                            stopped = incl_mask != 0u;
                        }
                    }
                    open[k] = !stopped;
                    any_open = any_open || !stopped;
                }
                base -= 32 * U;
                if (!any_open) {  // resolved: publish inclusive values, hand the offsets to the compute warps
                    const uint32_t buf = res % B;
#pragma unroll
                    for (int k = 0; k < V; ++k) {
                        const uint32_t ec = __reduce_add_sync(kFullMask, acc_c[k]);
                        const uint64_t el = warp_sum_u64(acc_l[k]);
                        if (lane < kComputeWarps) {
                            sm.adj_cnt[buf][k][lane] += ec;
                            sm.adj_len[buf][k][lane] += (uint32_t)el;  // firstIndex is u32 and wraps like the reference's
                        }
                        if (lane == 0) {
                            const uint32_t ic = ec + sm.agg_cnt[buf][k];
                            const uint64_t il = el + sm.agg_len[buf][k];
                            if (tile != 0) st_relaxed_v2(status + (size_t)k * n_tiles + tile, tag | (kFlagIncl << 32) | ic, il);
                            if (tile == n_tiles - 1) {
                                args.counts[k] = ic;
                                args.totals[k] = il;
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm.bases[buf]);
                    if (lane == 0) {
                        [[maybe_unused]] const long long now = RCULL_CLOCK();
                        RCULL_STAT(0, 1);
                        RCULL_STAT(1, now - t_pub[res % B]);
                        RCULL_STAT(9, now - t_act);
                    }
                    ++res;
                    active = false;
                }
            }
        } else if (ended) {
            break;
        }
        if (!progress) __nanosleep(RCULL_SCAN_IDLE_NS);
    }
}

// Compute-warp emission of its own survivors of a pending tile: one pass over the
// warp's combined (all views) list, stable within every view.
// sm_lod01 != nullptr: the mesh table is in shared memory (mesh ids fit a byte): the survivor's mesh id comes from
// the tile's mid8 bytes and its LOD row from the shared copy -- no global load between a resolved tile and its
// commands (the two dependent L2 round trips of the general path were 13 % of the compute warps' time).
template <int V>
__device__ __forceinline__ void emit_pending(const FrameArgs& args, const FrameSmem& sm, const uint8_t* wlist, const uint8_t* mid8,
                                             const uint4* sm_lod01, uint32_t buf, uint32_t tile_base, uint32_t warp, uint32_t lane) {
    uint32_t vend[V];  // end of view k in the combined list
    uint32_t total = 0;
#pragma unroll
    for (int k = 0; k < V; ++k) {
        total += sm.wcnt[buf][k][warp];
        vend[k] = total;
    }
    uint32_t carry = 0;
    for (uint32_t i0 = 0; i0 < total; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool act = i < total;
        uint32_t k = 0, e = 0, slot = 0;
        uint4 L = make_uint4(0, 0, 0, 0);
        if (act) {
#pragma unroll
            for (int kk = 0; kk < V - 1; ++kk) k += i >= vend[kk] ? 1u : 0u;
            const uint32_t entry = wlist[i];  // lod << 7 | slot in the warp's span
            e = warp * kWarpSpan + (entry & 127u);
            if (sm_lod01 != nullptr) {
                const uint32_t mid = mid8[e];  // already clamped by the compute pass
                slot = mid * kMaxLods + (entry >> 7);
                L = sm_lod01[2 * mid + (entry >> 7)];
            } else {
                uint32_t mid = __ldg(args.mesh + tile_base + e);  // survivors only: the column is still in L2
                if (mid >= args.n_meshes) mid = 0;                // same clamp as the compute pass (error already flagged)
                slot = mid * kMaxLods + (entry >> 7);
                L = __ldg(args.lod_tab + slot);
            }
        }
        const uint32_t incl = warp_incl_scan_u32(L.x, lane);
        const uint32_t excl = carry + incl - L.x;
        carry += __shfl_sync(kFullMask, incl, 31);
        if (args.rec.world != 0u) {  // merge attached: the record is the whole message; the receivers rebuild the command
            if (act) {
                const uint32_t pos = sm.adj_cnt[buf][k][warp] + i;
                record_store(args.rec, k, pos, args.draw_index_base + tile_base + e, slot);
                record_checkpoint(args.rec, k, pos, sm.adj_len[buf][k][warp] + excl);  // every 256th survivor: its local firstIndex
            }
            continue;
        }
        if (act) {
            const uint32_t pos = sm.adj_cnt[buf][k][warp] + i;
            if (pos < args.cap_visible) {
                const uint32_t draw_index = args.draw_index_base + tile_base + e;
                args.visible[k][pos] = draw_index;
                if (args.per_mesh) {  // per-mesh mode: the survivor's (mesh, lod) slot is the sort key
                    args.cmds[k][pos] = slot;
                    continue;
                }
                uint32_t* c = args.cmds[k] + (size_t)pos * 5;
                c[0] = L.x;                               // indexCount (entity-level upper bound)
                c[1] = 1u;                                // instanceCount
                c[2] = sm.adj_len[buf][k][warp] + excl;   // firstIndex: running sum of earlier survivors' index_len
                c[3] = L.z;                               // vertexOffset
                c[4] = draw_index;                        // firstInstance
                if (k == 0 && args.src_slot != nullptr) args.src_slot[pos] = slot;  // names the LOD: indexOffset of generate_work.comp's push constants = its first_index
            }
        }
    }
}

// ---------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------
// kMat: the optional AABB / CoarseCulled outputs may exist (checked per pointer) and index
// lengths may need 64-bit warp sums (args.len_small == 0); the hot variant (kMat = false)
// carries none of that code.
template <int V, bool kMat>
__global__ void __maxnreg__(RCULL_FRAME_REGS)
    transform_cull_compact(const __grid_constant__ FrameArgs args, const __grid_constant__ CUtensorMap model_map) {
    extern __shared__ unsigned char smem_raw[];
    FrameSmem& sm = frame_smem(smem_raw);

    const uint32_t tid = threadIdx.x;
    const uint32_t lane = tid & 31u;
    const uint32_t warp = tid >> 5;

    // the CTA's first tile ticket: requested before the barrier set-up so that its round trip overlaps it
    uint32_t first_t = 0;
    if (tid == 0) first_t = atomicAdd(args.ticket, 1u) - args.ticket_base;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&sm.full[s], 1);
            mbar_init(&sm.empty[s], kComputeWarps);
        }
#pragma unroll
        for (int b = 0; b < kMaxBufs; ++b) {
            mbar_init(&sm.done[b], kComputeWarps);
            mbar_init(&sm.bases[b], 1);
        }
        fence_mbar_init();
    }
    // Compute warp 0, lane 0 is also the producer: it draws tile tickets and issues the TMA
    // loads one tile ahead, as soon as the other stage has been read by every compute warp.
    // Tickets are drawn one tile EARLIER than they are needed (the atomic's round trip through L2
    // would otherwise sit in warp 0's timeline every tile, and every other warp ends up paced by
    // the warp that feeds them).  EVERY tile is a ticket, the first one of a CTA included: a tile is
    // only ever owned by a CTA that is running, so the look-back never waits on a CTA the hardware has
    // not scheduled yet (fewer SMs than the grid was sized for: MPS limits, green contexts).
    const bool producer = tid == 0;
    bool more = true;      // producer only: the last ticket used was a real tile
    uint32_t next_t = 0;   // producer only: the ticket drawn for the tile after the one being loaded
    auto draw_ticket = [&]() { return atomicAdd(args.ticket, 1u) - args.ticket_base; };
    auto load_tile = [&](uint32_t for_it, uint32_t t) {
        const uint32_t stage = for_it & 1u;
        sm.tile_ring[for_it % kRing] = t;
        uint64_t* bar = &sm.full[stage];
        if (t < args.n_tiles) {
            fence_proxy_async();
            mbar_expect_tx(bar, kStageBytes);
            bulk_g2s(sm.pos_scale[stage], args.pos_scale + (size_t)t * kTile, kTile * 16, bar);
            bulk_g2s(sm.rot[stage], args.rot + (size_t)t * kTile, kTile * 16, bar);
            bulk_g2s(sm.mesh[stage], args.mesh + (size_t)t * kTile, kTile * 4, bar);
        } else {
            mbar_arrive(bar);  // end marker: every warp reads tile_ring and leaves
            more = false;
        }
    };
    if (producer) {  // in flight while the CTA copies its tables below
        load_tile(0, first_t);
        if (more) next_t = draw_ticket();
    }
    float4* const sm_planes = reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(&sm) + sizeof(FrameSmem));
    for (uint32_t i = tid; i < V * 6 * 2; i += kThreads) sm_planes[i] = reinterpret_cast<const float4*>(&args.planes[0][0])[i];
    // Mesh table: a shared-memory copy when it fits (the L1 left beside ~220 KB of shared memory
    // is a few KB: table lines kept falling out of it), else the global one through generic loads.
    const float4* mesh_tab = args.mesh_aabb;
    const uint32_t* len1_tab = args.mesh_len1;
    const uint4* sm_lod01 = nullptr;  // set: emission reads mesh ids and LOD rows from shared memory
    if (args.n_meshes <= frame_mesh_cap<V>()) {
        float4* sm_mesh = reinterpret_cast<float4*>(reinterpret_cast<unsigned char*>(&sm) + frame_lists_end<V>());
        uint4* sm_lod = reinterpret_cast<uint4*>(sm_mesh + 2 * frame_mesh_cap<V>());
        for (uint32_t i = tid; i < 2 * args.n_meshes; i += kThreads) {
            sm_mesh[i] = args.mesh_aabb[i];
            sm_lod[i] = args.lod_tab[(i >> 1) * kMaxLods + (i & 1u)];  // lods 0 and 1 of mesh i / 2
        }
        mesh_tab = sm_mesh;
        sm_lod01 = sm_lod;
    }
    if (args.rec.world != 0u) record_wait_window_free(args.rec, tid);  // the region these records go to has been expanded by every rank
    __syncthreads();

    if (warp == kComputeWarps) {
        scan_warp<V>(args, sm, lane);
        return;
    }

    // ================================= compute warps =================================
    uint8_t* const wlist_base = reinterpret_cast<unsigned char*>(&sm) + sizeof(FrameSmem) + V * 6 * sizeof(PlaneC) + warp * (V * kWarpSpan);
    uint8_t* const mid8_base = reinterpret_cast<unsigned char*>(&sm) + frame_mid8_off<V>();  // [bufs][kTile]
    const uint32_t lt = (1u << lane) - 1u;
    // swizzled (128-B pattern) byte offsets of this lane's four matrix columns in the warp's store box
    unsigned char* const box = reinterpret_cast<unsigned char*>(sm.mat[warp]);
    const uint32_t mrow = lane >> 1, mhalf = (lane & 1u) * 4u;
    const uint32_t moff0 = mrow * 128u + (((mhalf + 0u) ^ (mrow & 7u)) << 4);
    const uint32_t moff1 = mrow * 128u + (((mhalf + 1u) ^ (mrow & 7u)) << 4);
    const uint32_t moff2 = mrow * 128u + (((mhalf + 2u) ^ (mrow & 7u)) << 4);
    const uint32_t moff3 = mrow * 128u + (((mhalf + 3u) ^ (mrow & 7u)) << 4);
    const uint32_t e0 = warp * kWarpSpan + lane;  // slot in tile of item 0; item j adds 32*j


    uint32_t it = 0;
    constexpr uint32_t B = frame_bufs<V>();
    [[maybe_unused]] const long long k0 = RCULL_CLOCK();
    uint32_t emitted = 0;  // iterations [emitted, it) of this CTA still have their commands pending
    auto emit_next = [&]() {
        const uint32_t b = emitted % B;
        emit_pending<V>(args, sm, wlist_base + b * (V * kTile), mid8_base + b * kTile, sm_lod01, b, sm.tile_ring[emitted % kRing] * kTile, warp, lane);
        ++emitted;
    };
    for (;; ++it) {
        const uint32_t stage = it & 1u;  // input stage of this tile
        const uint32_t buf = it % B;     // bookkeeping buffer of this tile
        if (producer && more) {
            if (it >= 1) mbar_wait(&sm.empty[stage ^ 1u], ((it - 1) >> 1) & 1u);
            const uint32_t t = next_t;
            if (t < args.n_tiles) next_t = draw_ticket();  // for tile it + 2: lands while this tile is computed
            load_tile(it + 1, t);
        }
        [[maybe_unused]] const long long f0 = RCULL_CLOCK();
        mbar_wait(&sm.full[stage], (it >> 1) & 1u);
        if (tid == 0) RCULL_STAT(6, RCULL_CLOCK() - f0);
        const uint32_t tile = sm.tile_ring[it % kRing];
        if (tile >= args.n_tiles) break;
        const uint32_t tile_base = tile * kTile;
        uint8_t* const wlist = wlist_base + buf * (V * kTile);

        float4 ps[kItems], q[kItems];
        uint32_t mid[kItems];
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            ps[j] = sm.pos_scale[stage][e0 + 32 * j];
            q[j] = sm.rot[stage][e0 + 32 * j];
            mid[j] = sm.mesh[stage][e0 + 32 * j];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.empty[stage]);  // inputs are in registers: the stage can be refilled

        float cx[kItems], cy[kItems], cz[kItems], hx[kItems], hy[kItems], hz[kItems];
        uint32_t len[kItems], entry[kItems];
        bool valid[kItems];
        // Input hygiene off the hot path: only the frame's last tile has padding slots (they
        // become identity entities that are never visible), and a mesh id outside the table
        // is an error the whole tile reports once.
        if (tile_base + kTile > args.n) {  // uniform
#pragma unroll
            for (int j = 0; j < kItems; ++j) {
                if (tile_base + e0 + 32 * j >= args.n) {
                    ps[j] = make_float4(0.f, 0.f, 0.f, 1.f);
                    q[j] = make_float4(0.f, 0.f, 0.f, 1.f);
                    mid[j] = 0;
                }
            }
        }
        {
            uint32_t worst = mid[0];
#pragma unroll
            for (int j = 1; j < kItems; ++j) worst = max(worst, mid[j]);
            if (worst >= args.n_meshes) {
                atomicOr(args.error_flag, 1u);
#pragma unroll
                for (int j = 0; j < kItems; ++j) mid[j] = mid[j] >= args.n_meshes ? 0u : mid[j];
            }
        }
#pragma unroll
        for (int j = 0; j < kItems; ++j) {
            const uint32_t g = tile_base + e0 + 32 * j;  // slot in frame
            valid[j] = g < args.n;
            const Mat34 m = model_matrix(ps[j], q[j]);

            // A2: world AABB
            const float4 mn = mesh_tab[2 * mid[j]];
            const float4 mx = mesh_tab[2 * mid[j] + 1];
            float lox, hix, loy, hiy, loz, hiz;
            row_extent(m.m00, m.m01, m.m02, m.tx, mn, mx, lox, hix);
            row_extent(m.m10, m.m11, m.m12, m.ty, mn, mx, loy, hiy);
            row_extent(m.m20, m.m21, m.m22, m.tz, mn, mx, loz, hiz);
            float minx, maxx, miny, maxy, minz, maxz;
            aabb_store_form(lox, hix, minx, maxx);
            aabb_store_form(loy, hiy, miny, maxy);
            aabb_store_form(loz, hiz, minz, maxz);
            if (kMat && args.aabb != nullptr && valid[j]) {
                float* a = args.aabb + (size_t)g * 6;
                a[0] = minx; a[1] = miny; a[2] = minz;
                a[3] = maxx; a[4] = maxy; a[5] = maxz;
            }
            // A3 inputs: ncollide half_extents() / center() of the STORED box
            hx[j] = __fmul_rn(__fsub_rn(maxx, minx), 0.5f); cx[j] = __fmul_rn(__fadd_rn(minx, maxx), 0.5f);
            hy[j] = __fmul_rn(__fsub_rn(maxy, miny), 0.5f); cy[j] = __fmul_rn(__fadd_rn(miny, maxy), 0.5f);
            hz[j] = __fmul_rn(__fsub_rn(maxz, minz), 0.5f); cz[j] = __fmul_rn(__fadd_rn(minz, maxz), 0.5f);

            // A4: pick_lod -> (mesh, lod) slot
            const float dx = __fsub_rn(args.cam[0], ps[j].x), dy = __fsub_rn(args.cam[1], ps[j].y),
                        dz = __fsub_rn(args.cam[2], ps[j].z);
            const float d2 = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
            const uint32_t n_lods = __float_as_uint(mn.w);
            const bool lod1 = d2 > args.lod_d2_max && n_lods > 1u;
            len[j] = lod1 ? (sm_lod01 != nullptr ? sm_lod01[2 * mid[j] + 1].x : len1_tab[mid[j]]) : __float_as_uint(mx.w);  // index_len of the chosen lod
            entry[j] = (lod1 ? 128u : 0u) + lane + 32u * j;            // survivor-list entry: lod << 7 | slot in span

            // model matrix out: swizzled box in smem -> one 2 KB TMA tensor store per warp and item.  LAST in the
            // item: the wait for the previous item's store to have read the box then has the whole A2 / A4 arithmetic of
            // this item in front of it.
            if (lane == 0) bulk_wait_read0();
            __syncwarp();
            *reinterpret_cast<float4*>(box + moff0) = make_float4(m.m00, m.m10, m.m20, 0.f);
            *reinterpret_cast<float4*>(box + moff1) = make_float4(m.m01, m.m11, m.m21, 0.f);
            *reinterpret_cast<float4*>(box + moff2) = make_float4(m.m02, m.m12, m.m22, 0.f);
            *reinterpret_cast<float4*>(box + moff3) = make_float4(m.tx, m.ty, m.tz, 1.f);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                tma_store_2d(&model_map, box, 0, (int)((tile_base + warp * kWarpSpan + 32 * j) >> 1));
                bulk_commit();
            }
        }

        float2 c2x[kPairs], c2y[kPairs], c2z[kPairs], h2x[kPairs], h2y[kPairs], h2z[kPairs];
#pragma unroll
        for (int i = 0; i < kPairs; ++i) {
            c2x[i] = make_float2(cx[2 * i], cx[2 * i + 1]); h2x[i] = make_float2(hx[2 * i], hx[2 * i + 1]);
            c2y[i] = make_float2(cy[2 * i], cy[2 * i + 1]); h2y[i] = make_float2(hy[2 * i], hy[2 * i + 1]);
            c2z[i] = make_float2(cz[2 * i], cz[2 * i + 1]); h2z[i] = make_float2(hz[2 * i], hz[2 * i + 1]);
        }

        // Emit the commands of earlier tiles whose output offsets the scan warp has resolved
        // by now; only the tile that owns this iteration's bookkeeping buffer is waited for.
        while (emitted < it) {
            const uint32_t b = emitted % B, parity = (emitted / B) & 1u;
            if (emitted + B <= it) {
                [[maybe_unused]] const long long w0 = RCULL_CLOCK();
                mbar_wait(&sm.bases[b], parity);
                if (tid == 0) {
                    RCULL_STAT(5, RCULL_CLOCK() - w0);
                    RCULL_STAT(8, 1);
                }
            } else {
                uint32_t ready = lane == 0 ? (uint32_t)mbar_test(&sm.bases[b], parity) : 0u;
                ready = __shfl_sync(kFullMask, ready, 0);
                if (!ready) break;
            }
            emit_next();
        }
        __syncwarp();
        const uint32_t wlist_s = smem_u32(wlist);
        if (sm_lod01 != nullptr) {  // this tile's mesh ids, for its emission a few tiles from now (the buffer's previous tile is emitted)
            const uint32_t m8 = smem_u32(mid8_base + buf * kTile + e0);
#pragma unroll
            for (int j = 0; j < kItems; ++j) sts_u8(m8 + 32 * j, mid[j]);
        }

        // A3 per view: planes outer, the thread's 4 entities inner
        // (the view loop stays rolled: unrolled, the kernel outgrows the instruction cache).
        // Two bodies, chosen per warp and tile:
        //   packed  -- products FMUL2.FTZ, every sum FADD2 (13 instructions per plane and entity pair).  Exact as
        //              long as no factor is a nonzero number below 2^-63 (mul2_ftz, rcull_device.cuh);
        //   scalar  -- products FMUL2, sums scalar FADD (17 instructions): any input.
        // The packed body needs a second use of nothing and no FFMA2: the .ftz mismatch is what keeps ptxas from
        // contracting the pairs (tests/test_abi.py checks the SASS for FFMA / FFMA2).
        bool packed = V >= 2 && args.planes_ftz_safe != 0u;  // one view is HBM-bound either way: it keeps the plain body
        if (V >= 2) {
            float tiny = fminf(fminf(fabsf(cx[0]), fabsf(cy[0])), fabsf(cz[0]));
            tiny = fminf(tiny, fminf(fminf(fabsf(hx[0]), fabsf(hy[0])), fabsf(hz[0])));
#pragma unroll
            for (int j = 1; j < kItems; ++j) {
                tiny = fminf(tiny, fminf(fminf(fabsf(cx[j]), fabsf(cy[j])), fabsf(cz[j])));
                tiny = fminf(tiny, fminf(fminf(fabsf(hx[j]), fabsf(hy[j])), fabsf(hz[j])));
            }
            if (__any_sync(kFullMask, tiny < kFtzSafe)) {  // rare: an exact zero (fine) or a tiny nonzero factor (not fine)
                bool bad = false;
#pragma unroll
                for (int j = 0; j < kItems; ++j) {
                    const float v6[6] = {cx[j], cy[j], cz[j], hx[j], hy[j], hz[j]};
#pragma unroll
                    for (int a = 0; a < 6; ++a) bad = bad || (v6[a] != 0.f && fabsf(v6[a]) < kFtzSafe);
                }
                if (__any_sync(kFullMask, bad)) packed = false;
            }
        }
        uint32_t woff = 0;  // start of view k in the warp's combined list
        const float4* vp = sm_planes;  // planes of view k: 6 x 2 float4
        auto view_loop = [&](auto packed_c) {
            constexpr bool kPacked = decltype(packed_c)::value;
#pragma unroll kViewUnroll
            for (int k = 0; k < V; ++k, vp += 12) {
                bool outside[kItems];
#pragma unroll
                for (int j = 0; j < kItems; ++j) outside[j] = !valid[j];  // padding slots are never visible
#pragma unroll
                for (int p = 0; p < 6; ++p) {
                    const float4* pp = vp + p * 2;
                    const float4 q0 = pp[0], q1 = pp[1];  // (ax ay az px) (py pz pw -)
                    const float pw = q1.z;
                    const float2 ax = make_float2(q0.x, q0.x), ay = make_float2(q0.y, q0.y), az = make_float2(q0.z, q0.z);
                    const float2 px = make_float2(q0.w, q0.w), py = make_float2(q1.x, q1.x), pz = make_float2(q1.y, q1.y);
                    if constexpr (kOddItem) {  // the unpaired entity: the same operations, one at a time (scalar IEEE products and sums)
                        constexpr int o = kItems - 1;
                        const float e = __fadd_rn(__fadd_rn(__fmul_rn(hx[o], ax.x), __fmul_rn(hy[o], ay.x)), __fmul_rn(hz[o], az.x));
                        const float sa = __fadd_rn(__fmul_rn(cx[o], px.x), __fmul_rn(cz[o], pz.x));
                        const float sb = __fadd_rn(__fmul_rn(cy[o], py.x), pw);
                        outside[o] = outside[o] || (__fadd_rn(sa, sb) > e);
                    }
#pragma unroll
                    for (int i = 0; i < kPairs; ++i) {
                        if constexpr (kPacked) {
                            const float2 pw2 = make_float2(pw, pw);
                            const float2 e = add2_rn(add2_rn(mul2_ftz(h2x[i], ax), mul2_ftz(h2y[i], ay)), mul2_ftz(h2z[i], az));
                            const float2 a = add2_rn(mul2_ftz(c2x[i], px), mul2_ftz(c2z[i], pz));
                            const float2 b = add2_rn(mul2_ftz(c2y[i], py), pw2);
                            const float2 s = add2_rn(a, b);
                            outside[2 * i] = outside[2 * i] || (s.x > e.x);
                            outside[2 * i + 1] = outside[2 * i + 1] || (s.y > e.y);
                        } else {
                            // Products of two entities per FMUL2; every sum that consumes a packed product
                            // is a scalar FADD, so ptxas has no packed mul+add pair to contract into FFMA2
                            // (it does that even under --fmad=false; tests/test_abi.py checks the SASS).
                            const float2 m1 = mul2_rn(h2x[i], ax), m2 = mul2_rn(h2y[i], ay), m3 = mul2_rn(h2z[i], az);
                            const float e0 = __fadd_rn(__fadd_rn(m1.x, m2.x), m3.x);
                            const float e1 = __fadd_rn(__fadd_rn(m1.y, m2.y), m3.y);
                            const float2 p1 = mul2_rn(c2x[i], px), p3 = mul2_rn(c2z[i], pz), p2 = mul2_rn(c2y[i], py);
                            const float2 a = make_float2(__fadd_rn(p1.x, p3.x), __fadd_rn(p1.y, p3.y));
                            const float2 b = make_float2(__fadd_rn(p2.x, pw), __fadd_rn(p2.y, pw));
                            const float2 s = add2_rn(a, b);
                            // (s - e > 0.0) == (s > e): the rounded difference of two floats keeps its sign,
                            // and both forms are false for NaN and for inf - inf
                            outside[2 * i] = outside[2 * i] || (s.x > e0);
                            outside[2 * i + 1] = outside[2 * i + 1] || (s.y > e1);
                        }
                    }
                }
                // ballot compaction, branch-free
                uint32_t wc = 0, lo = 0, hi = 0;
#pragma unroll
                for (int j = 0; j < kItems; ++j) {
                    const bool vis = !outside[j];
                    if (kMat && args.flags[k] != nullptr && valid[j]) args.flags[k][tile_base + e0 + 32 * j] = outside[j] ? 1 : 0;
                    const uint32_t b = __ballot_sync(kFullMask, vis);
                    if (vis) sts_u8(wlist_s + woff + wc + __popc(b & lt), entry[j]);  // compiles to a predicated STS
                    if (!kMat || args.len_small) {  // compile-time in the hot variant
                        lo += vis ? len[j] : 0u;
                    } else {
                        lo += vis ? (len[j] & 0xffffu) : 0u;
                        hi += vis ? (len[j] >> 16) : 0u;
                    }
                    wc += __popc(b);
                }
                uint64_t wl;
                if (!kMat || args.len_small) {  // compile-time in the hot variant
                    wl = __reduce_add_sync(kFullMask, lo);  // 128 values below 2^25
                } else {
                    lo = __reduce_add_sync(kFullMask, lo);  // 128 x 16-bit pieces: no overflow
                    hi = __reduce_add_sync(kFullMask, hi);
                    wl = (uint64_t)lo + ((uint64_t)hi << 16);
                }
                if (lane == 0) {
                    sm.wcnt[buf][k][warp] = wc;
                    sm.wlen[buf][k][warp] = wl;
                }
                woff += wc;
            }
        };
        if (V >= 2 && packed)
            view_loop(cuda::std::true_type{});
        else
            view_loop(cuda::std::false_type{});
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.done[buf]);
    }
    // the last tiles of this CTA are still pending
    while (emitted < it) {
        mbar_wait(&sm.bases[emitted % B], (emitted / B) & 1u);
        emit_next();
    }
    // pass the end marker on to the scan warp (after the waits above, so that this barrier
    // never runs two phases ahead of its reader)
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.done[it % B]);
    if (lane == 0) bulk_wait_read0();
    if (args.rec.world != 0u) {
        // Every record of this CTA is visible system-wide before the CTA counts itself: the barrier orders all compute
        // threads' stores before thread 0's system-scope fence (fences are cumulative), so ONE fence per CTA does it.
        // The last CTA of the launch sends the frame's (count, index total) words and raises the tag in every window.
        asm volatile("bar.sync 1, %0;" ::"r"(kComputeWarps * 32) : "memory");  // the compute warps only: the scan warp has left
        if (tid == 0) {
            __threadfence_system();
            record_finish(args.rec, args.counts, args.totals, V);
        }
    }
    if (tid == 0) {
        RCULL_STAT(7, RCULL_CLOCK() - k0);
        RCULL_STAT(10, 1);
    }
}

}  // namespace rcull
