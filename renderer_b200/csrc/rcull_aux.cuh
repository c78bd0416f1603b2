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

// ---- range-shard merge over peer memory (NVLink P2P) ----------------------
// Every rank holds a MERGED allocation with the same layout:
//   cmds of view k at  k * cap * 20 bytes,  visible of view k at  V * cap * 20 + k * cap * 4.
// After an all_gather of the per-rank (count, index_total) words, ONE kernel per rank
// computes its slot from the gathered words and stores its rebased list into that slot
// of every rank's allocation -- its own and, through peer-mapped pointers, the others'.
// The stores are the transfer: no staging copy, no per-slice collective, no host round trip.
constexpr int kMaxRanks = 8;

struct MergeArgs {
    unsigned char* peer[kMaxRanks];         // merged allocation of every rank (peer-mapped)
    const uint32_t* src_cmds[8];            // this rank's lists, per view
    const uint32_t* src_visible[8];
    const unsigned long long* meta_all;     // [world][2 * max_views]: counts then index totals (u64)
    uint32_t* merged_counts;                // [max_views] outputs in the local allocation's header
    unsigned long long* merged_totals;
    uint32_t world, rank, n_views, max_views;
    uint32_t cap;                           // merged list capacity (commands per view)
    // fused exchange (merge_exchange): no collective library on the path
    const unsigned long long* meta_mine;    // this rank's (counts, index totals) of the frame being merged
    size_t sync_off;                        // byte offset of the MergeSync block in every merged allocation
    unsigned long long epoch;               // exchange number, the same on every rank
    uint32_t* cta_done;                     // local counter: CTAs of this launch that have stored their share
    unsigned char* mc;                      // NVLS multicast mapping of every rank's merged allocation (one store, the switch replicates), or null
    __host__ __device__ size_t n_views_cap_bytes() const { return (size_t)max_views * cap * 20; }
};

__global__ void __launch_bounds__(256) pack_meta(const uint32_t* __restrict__ counts,
                                                 const unsigned long long* __restrict__ totals, uint32_t max_views,
                                                 unsigned long long* __restrict__ meta_local) {
    const uint32_t i = threadIdx.x;
    if (i < max_views) {
        meta_local[i] = counts[i];
        meta_local[max_views + i] = totals[i];
    }
}

// Flags other GPUs write into this GPU's memory over NVLink: system-scope release / acquire.
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Stores through an NVLS multicast address: one store leaves the GPU, the NVSwitch writes it into every
// member's memory (egress per rank = its own lists once, not once per peer).
__device__ __forceinline__ void multimem_st_v4(void* mc, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void multimem_st_u32(void* mc, uint32_t v) {
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

// Per-rank mailbox at the end of every merged allocation; slot [r] is written by rank r only.
struct MergeSync {
    // two slots, by exchange parity: a rank can be one exchange ahead of a peer's late CTAs, never two
    // (its exchange e + 2 needs the peer's words of e + 1, which the peer sends after finishing e)
    unsigned long long meta[2][kMaxRanks][16];  // counts[8] then index totals[8] of rank r
    unsigned long long meta_tag[2][kMaxRanks];  // epoch of the exchange meta[.][r] belongs to
    unsigned long long done_tag[kMaxRanks];  // rank r has stored all its slices of exchange `epoch` here
};

// Copies n_words 32-bit words of one of this rank's lists to byte offset dst_off of EVERY
// rank's merged allocation.  Stores are 16 bytes wide wherever the destination is 16-byte
// aligned (NVLink moves full 16-byte-multiple packets: word-wide stores reached 0.3 TB/s, a third
// of the link); the up-to-3 words before the first aligned address and after the last go out as
// single words.  kCmds: word w of the 5-word command stream is firstIndex iff w % 5 == 2.
template <bool kCmds>
__device__ __forceinline__ void store_to_peers(const MergeArgs& a, const uint32_t* __restrict__ src, size_t n_words, size_t dst_off,
                                               uint32_t add, size_t t0, size_t stride) {
    const size_t head = ((16u - (uint32_t)(dst_off & 15u)) & 15u) / 4u;  // words until the destination is 16-byte aligned
    const size_t h = head < n_words ? head : n_words;
    const size_t n_vec = (n_words - h) / 4;
    // four vectors per thread in flight: the loop is bound by the latency of the local loads, not by issue
    constexpr int kUnroll = 4;
    for (size_t v0 = t0; v0 < n_vec; v0 += kUnroll * stride) {
        uint4 x[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t v = v0 + u * stride;
            if (v < n_vec) {
                const size_t w = h + 4 * v;
                x[u] = make_uint4(src[w], src[w + 1], src[w + 2], src[w + 3]);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const size_t v = v0 + u * stride;
            if (v < n_vec) {
                const size_t w = h + 4 * v;
                if (kCmds) {
                    const uint32_t m = (uint32_t)(w % 5);  // firstIndex sits where (w + j) % 5 == 2
                    if (m == 2) x[u].x += add;
                    if (m == 1) x[u].y += add;
                    if (m == 0) x[u].z += add;
                    if (m == 4) x[u].w += add;
                }
                if (a.mc != nullptr) {
                    multimem_st_v4(a.mc + dst_off + 4 * w, x[u]);  // NVSwitch replicates the store into every rank's allocation
                } else {
                    for (uint32_t p = 0; p < a.world; ++p) *reinterpret_cast<uint4*>(a.peer[p] + dst_off + 4 * w) = x[u];
                }
            }
        }
    }
    // head and tail words (at most 6 in total)
    const size_t tail0 = h + 4 * n_vec;
    for (size_t i = t0; i < h + (n_words - tail0); i += stride) {
        const size_t w = i < h ? i : tail0 + (i - h);
        uint32_t x = src[w];
        if (kCmds && w % 5 == 2) x += add;
        if (a.mc != nullptr) {
            multimem_st_u32(a.mc + dst_off + 4 * w, x);
        } else {
            for (uint32_t p = 0; p < a.world; ++p) *reinterpret_cast<uint32_t*>(a.peer[p] + dst_off + 4 * w) = x;
        }
    }
}

__global__ void __launch_bounds__(256) scatter_rebased(const __grid_constant__ MergeArgs a) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (uint32_t k = 0; k < a.n_views; ++k) {
        unsigned long long cnt_prefix = 0, len_prefix = 0, cnt_all = 0, len_all = 0;
        for (uint32_t r = 0; r < a.world; ++r) {
            const unsigned long long c = a.meta_all[(size_t)r * 2 * a.max_views + k];
            const unsigned long long l = a.meta_all[(size_t)r * 2 * a.max_views + a.max_views + k];
            if (r < a.rank) {
                cnt_prefix += c;
                len_prefix += l;
            }
            cnt_all += c;
            len_all += l;
        }
        if (t0 == 0) {
            a.merged_counts[k] = (uint32_t)cnt_all;
            a.merged_totals[k] = len_all;
        }
        unsigned long long count = a.meta_all[(size_t)a.rank * 2 * a.max_views + k];
        if (cnt_prefix >= a.cap) continue;
        if (cnt_prefix + count > a.cap) count = a.cap - cnt_prefix;  // clamp writes, like cap_visible
        const size_t words = (size_t)count * 5;
        const uint32_t add = (uint32_t)len_prefix;
        const uint32_t* __restrict__ sc = a.src_cmds[k];
        const uint32_t* __restrict__ sv = a.src_visible[k];
        const size_t cmd_off = ((size_t)k * a.cap + cnt_prefix) * 20;
        const size_t vis_off = (size_t)a.n_views_cap_bytes() + ((size_t)k * a.cap + cnt_prefix) * 4;
        store_to_peers<true>(a, sc, words, cmd_off, add, t0, stride);
        store_to_peers<false>(a, sv, (size_t)count, vis_off, 0u, t0, stride);
    }
}

// The whole exchange step in ONE kernel, no collective library call on the path:
//   1. CTA 0 stores this rank's (count, index total) words into every rank's mailbox and raises its tag;
//   2. every CTA waits until all ranks' words of this epoch have arrived in the LOCAL mailbox;
//   3. exclusive prefixes over ranks -> this rank's slot; the rebased lists are stored into that
//      slot of every rank's merged allocation (16-byte peer stores over NVLink);
//   4. the last CTA to finish raises this rank's done tag in every mailbox.
// 128 threads x 64 registers (the 8192 registers the frame kernel's CTA leaves free) and no shared memory: such a CTA is resident beside
// the frame kernel's CTA, so the exchange of frame i runs under frame i + 1.
__global__ void __maxnreg__(64) merge_exchange(const __grid_constant__ MergeArgs a) {
    const uint32_t tid = threadIdx.x;
    if (blockIdx.x == 0) {
        if (tid < 16) {
            const uint32_t k = tid & 7u;
            const unsigned long long v = k < a.max_views ? a.meta_mine[(tid >> 3) * a.max_views + k] : 0ull;
            for (uint32_t p = 0; p < a.world; ++p) reinterpret_cast<MergeSync*>(a.peer[p] + a.sync_off)->meta[a.epoch & 1][a.rank][tid] = v;
        }
        __threadfence_system();
        __syncthreads();
        if (tid < a.world) st_release_sys(&reinterpret_cast<MergeSync*>(a.peer[tid] + a.sync_off)->meta_tag[a.epoch & 1][a.rank], a.epoch);
    }
    const MergeSync* mine = reinterpret_cast<const MergeSync*>(a.peer[a.rank] + a.sync_off);
    if (tid < a.world) {
        while (ld_acquire_sys(&mine->meta_tag[a.epoch & 1][tid]) != a.epoch) __nanosleep(200);
    }
    __syncthreads();
    const unsigned long long (*meta)[16] = mine->meta[a.epoch & 1];  // every rank's words, now in local memory

    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t t0 = (size_t)blockIdx.x * blockDim.x + tid;
    for (uint32_t k = 0; k < a.n_views; ++k) {
        unsigned long long cnt_prefix = 0, len_prefix = 0, cnt_all = 0, len_all = 0;
        for (uint32_t r = 0; r < a.world; ++r) {
            const unsigned long long c = ld_relaxed_sys(&meta[r][k]), l = ld_relaxed_sys(&meta[r][8 + k]);
            if (r < a.rank) {
                cnt_prefix += c;
                len_prefix += l;
            }
            cnt_all += c;
            len_all += l;
        }
        if (t0 == 0) {
            a.merged_counts[k] = (uint32_t)cnt_all;
            a.merged_totals[k] = len_all;
        }
        unsigned long long count = ld_relaxed_sys(&meta[a.rank][k]);
        if (cnt_prefix >= a.cap) continue;
        if (cnt_prefix + count > a.cap) count = a.cap - cnt_prefix;  // clamp writes, like cap_visible
        const size_t cmd_off = ((size_t)k * a.cap + cnt_prefix) * 20;
        const size_t vis_off = (size_t)a.n_views_cap_bytes() + ((size_t)k * a.cap + cnt_prefix) * 4;
        store_to_peers<true>(a, a.src_cmds[k], (size_t)count * 5, cmd_off, (uint32_t)len_prefix, t0, stride);
        store_to_peers<false>(a, a.src_visible[k], (size_t)count, vis_off, 0u, t0, stride);
    }
    __threadfence_system();  // this thread's peer stores are visible system-wide before the CTA counts itself
    __syncthreads();
    if (tid == 0) {
        const uint32_t done = atomicAdd(a.cta_done, 1u);
        if (done == gridDim.x - 1) {
            *a.cta_done = 0;
            __threadfence_system();
            for (uint32_t p = 0; p < a.world; ++p)
                st_release_sys(&reinterpret_cast<MergeSync*>(a.peer[p] + a.sync_off)->done_tag[a.rank], a.epoch);
        }
    }
}

// Returns when every rank's slices of exchange `epoch` have landed in this rank's allocation.
__global__ void merge_wait(const MergeSync* mine, uint32_t world, unsigned long long epoch) {
    if (threadIdx.x < world) {
        while (ld_acquire_sys(&mine->done_tag[threadIdx.x]) < epoch) __nanosleep(200);
    }
}

// ---- per-mesh instanced emission (BASELINE config 4; compact_draw_stream.comp:41 keep-rule on
// instanceCount) ----------------------------------------------------------------------
// The frame kernel leaves, per view, the survivors in draw_index order as (slot key,
// draw_index) pairs.  A stable counting sort buckets them by slot:
//   bucket_hist     one warp per tile of `tile_len` pairs: per-slot counts -> hist[tile][slot]
//   bucket_colscan  per slot, exclusive scan down the tiles (in place) + slot totals
//   bucket_cmds     exclusive scan over the slot totals -> bucket starts, one command per
//                   NON-EMPTY (mesh, lod) slot in slot order
//   bucket_scatter  same tiling as bucket_hist: rank of each pair inside its slot =
//                   bucket start + tile offset + arrivals before it in the tile (stable)
// Slot counters live in shared memory, kBucketChunk slots per warp at a time; tables with more
// slots take one sweep of the tile per chunk.  The survivor count is read on the device.
constexpr uint32_t kBucketChunk = 2048;
constexpr uint32_t kBucketWarps = 4;

struct BucketArgs {
    const uint32_t* keys[8];    // per view: slot key per survivor, draw order
    const uint32_t* draw_ix[8]; // per view: draw_index per survivor, draw order
    uint32_t* visible[8];       // out: draw indices grouped by slot
    uint32_t* cmds[8];          // out: 5 words per command
    uint32_t* counts;           // [view] in: survivors; out (bucket_cmds): commands
    unsigned long long* totals; // [view] out (bucket_cmds): instances
    uint32_t* hist;             // [view][rows_max][n_slots]
    uint32_t* slot_total;       // [view][n_slots]
    uint32_t* slot_base;        // [view][n_slots]
    const uint4* lod_tab;       // index_len, first_index, vertex_offset
    uint32_t n_slots, cap;      // cap: pairs the lists can hold
    uint32_t rows_max;          // histogram rows available per view
    uint32_t rows_target;       // tiles wanted: about two per resident warp
};

// The view this CTA works on, its survivor count and the tile length all sweeps agree on:
// as many tiles as there are warps to run them, never more rows than the histogram holds.
struct BucketView {
    uint32_t view, n_surv, tile_len, n_tiles;
    uint32_t* hist;
};
__device__ __forceinline__ BucketView bucket_view(const BucketArgs& a, bool after_cmds) {
    BucketView b;
    b.view = blockIdx.y;
    const uint32_t n = after_cmds ? (uint32_t)a.totals[b.view] : a.counts[b.view];
    b.n_surv = n < a.cap ? n : a.cap;
    const uint32_t rows = a.rows_target < a.rows_max ? a.rows_target : a.rows_max;
    const uint32_t len = ((b.n_surv + rows - 1) / rows + 31u) & ~31u;
    b.tile_len = len < 128u ? 128u : len;
    b.n_tiles = (b.n_surv + b.tile_len - 1) / b.tile_len;
    b.hist = a.hist + (size_t)b.view * a.rows_max * a.n_slots;
    return b;
}

template <bool kScatter>
__device__ __forceinline__ void bucket_sweep(const BucketArgs& a) {
    __shared__ uint32_t s_cnt[kBucketWarps][kBucketChunk];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t* cnt = s_cnt[warp];
    const BucketView bv = bucket_view(a, kScatter);
    const uint32_t* __restrict__ keys = a.keys[bv.view];
    const uint32_t* __restrict__ draw_ix = a.draw_ix[bv.view];
    uint32_t* __restrict__ visible = a.visible[bv.view];
    const uint32_t* slot_base = a.slot_base + (size_t)bv.view * a.n_slots;
    const uint32_t lt = (1u << lane) - 1u;
    for (uint32_t tile = blockIdx.x * kBucketWarps + warp; tile < bv.n_tiles; tile += gridDim.x * kBucketWarps) {
        const uint32_t lo = tile * bv.tile_len;
        const uint32_t hi = min(lo + bv.tile_len, bv.n_surv);
        uint32_t* row = bv.hist + (size_t)tile * a.n_slots;
        for (uint32_t c0 = 0; c0 < a.n_slots; c0 += kBucketChunk) {
            const uint32_t nc = min(kBucketChunk, a.n_slots - c0);
            for (uint32_t j = lane; j < nc; j += 32)
                cnt[j] = kScatter ? row[c0 + j] + slot_base[c0 + j] : 0u;
            __syncwarp();
            // four steps of 32 pairs at a time: their loads are issued together
            for (uint32_t i0 = lo; i0 < hi; i0 += 128) {
                uint32_t key[4], dix[4];
#pragma unroll
                for (uint32_t u = 0; u < 4; ++u) {
                    const uint32_t i = i0 + u * 32 + lane;
                    key[u] = i < hi ? __ldg(keys + i) - c0 : 0xffffffffu;
                    if (kScatter) dix[u] = i < hi ? __ldg(draw_ix + i) : 0u;
                }
#pragma unroll
                for (uint32_t u = 0; u < 4; ++u) {
                    const bool in = key[u] < nc;
                    const uint32_t peers = __match_any_sync(kFullMask, key[u]);
                    const uint32_t leader = __ffs(peers) - 1;
                    uint32_t old = 0;
                    if (in && lane == leader) {
                        old = cnt[key[u]];
                        cnt[key[u]] = old + __popc(peers);
                    }
                    if (kScatter) {
                        old = __shfl_sync(kFullMask, old, leader);
                        if (in) visible[old + __popc(peers & lt)] = dix[u];
                    }
                    __syncwarp();
                }
            }
            if (!kScatter)
                for (uint32_t j = lane; j < nc; j += 32) row[c0 + j] = cnt[j];
            __syncwarp();
        }
    }
}

__global__ void __launch_bounds__(kBucketWarps * 32) bucket_hist(const __grid_constant__ BucketArgs a) {
    bucket_sweep<false>(a);
}
__global__ void __launch_bounds__(kBucketWarps * 32) bucket_scatter(const __grid_constant__ BucketArgs a) {
    bucket_sweep<true>(a);
}

// CTA = 32 slots (x) by 32 runs of tiles (y): every thread sums its run of tiles for one slot,
// the runs are scanned through shared memory, and a second walk writes the exclusive offsets.
__global__ void __launch_bounds__(1024) bucket_colscan(const __grid_constant__ BucketArgs a) {
    __shared__ uint32_t s_run[32][33];
    const uint32_t x = threadIdx.x & 31u, y = threadIdx.x >> 5;
    const uint32_t slot = blockIdx.x * 32 + x;
    const BucketView bv = bucket_view(a, false);
    const uint32_t n_tiles = bv.n_tiles;
    const uint32_t per_run = (n_tiles + 31) / 32;
    const uint32_t t0 = min(y * per_run, n_tiles), t1 = min(t0 + per_run, n_tiles);
    uint32_t sum = 0;
    if (slot < a.n_slots)
        for (uint32_t t = t0; t < t1; ++t) sum += bv.hist[(size_t)t * a.n_slots + slot];
    s_run[y][x] = sum;
    __syncthreads();
    uint32_t before = 0;
    for (uint32_t r = 0; r < y; ++r) before += s_run[r][x];
    if (slot < a.n_slots) {
        for (uint32_t t = t0; t < t1; ++t) {
            uint32_t* h = bv.hist + (size_t)t * a.n_slots + slot;
            const uint32_t c = *h;
            *h = before;
            before += c;
        }
        if (y == 31) a.slot_total[(size_t)bv.view * a.n_slots + slot] = before;
    }
}

__global__ void __launch_bounds__(1024) bucket_cmds(const __grid_constant__ BucketArgs a) {
    __shared__ uint32_t warp_cmds[32], warp_inst[32];
    __shared__ uint32_t run_cmds, run_inst;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const BucketView bv = bucket_view(a, false);
    const uint32_t* slot_total = a.slot_total + (size_t)bv.view * a.n_slots;
    uint32_t* slot_base = a.slot_base + (size_t)bv.view * a.n_slots;
    if (tid == 0) run_cmds = run_inst = 0;
    __syncthreads();
    for (uint32_t s0 = 0; s0 < a.n_slots; s0 += blockDim.x) {
        const uint32_t s = s0 + tid;
        const uint32_t cnt = s < a.n_slots ? slot_total[s] : 0u;
        const uint32_t b = __ballot_sync(kFullMask, cnt > 0);
        uint32_t incl = cnt;  // inclusive scan of the instance counts across the warp
#pragma unroll
        for (uint32_t d = 1; d < 32; d <<= 1) {
            const uint32_t up = __shfl_up_sync(kFullMask, incl, d);
            if (lane >= d) incl += up;
        }
        if (lane == 31) {
            warp_cmds[warp] = __popc(b);
            warp_inst[warp] = incl;
        }
        __syncthreads();
        uint32_t cmds_before = run_cmds, inst_before = run_inst;
        for (uint32_t w = 0; w < warp; ++w) {
            cmds_before += warp_cmds[w];
            inst_before += warp_inst[w];
        }
        const uint32_t first = inst_before + incl - cnt;
        if (s < a.n_slots) slot_base[s] = first;
        if (cnt > 0) {
            const uint4 L = __ldg(a.lod_tab + s);
            uint32_t* c = a.cmds[bv.view] + (size_t)(cmds_before + __popc(b & ((1u << lane) - 1u))) * 5;
            c[0] = L.x;    // indexCount
            c[1] = cnt;    // instanceCount
            c[2] = L.y;    // firstIndex: the LOD's own first_index
            c[3] = L.z;    // vertexOffset
            c[4] = first;  // firstInstance: bucket start in the bucketed visible list
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t tc = run_cmds, ti = run_inst;
            for (uint32_t w = 0; w < (blockDim.x >> 5); ++w) {
                tc += warp_cmds[w];
                ti += warp_inst[w];
            }
            run_cmds = tc;
            run_inst = ti;
        }
        __syncthreads();
    }
    if (tid == 0) {
        a.counts[bv.view] = run_cmds;
        a.totals[bv.view] = bv.n_surv;
    }
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

// ---- acceleration-structure instance records ------------------------------
// build_acceleration_structures' CPU loop (acceleration_strucures.rs:419-445) writes one
// VkAccelerationStructureInstanceKHR per entity: the upper 3 rows of the model matrix row-major,
// instance_custom_index = draw_index (24 bits) | mask 0xFF, SBT offset 0 | flags
// TRIANGLE_FACING_CULL_DISABLE (= 1), and the BLAS device address of the entity's mesh.
// One thread per entity: 68 bytes in (model + mesh id), 64 bytes out.
__global__ void __launch_bounds__(256) as_instances(const float4* __restrict__ model, const uint32_t* __restrict__ mesh,
                                                    const unsigned long long* __restrict__ blas, uint32_t n_meshes, uint32_t n,
                                                    uint32_t draw_index_base, uint4* __restrict__ out) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 c0 = model[4 * (size_t)i], c1 = model[4 * (size_t)i + 1], c2 = model[4 * (size_t)i + 2], c3 = model[4 * (size_t)i + 3];
        uint32_t m = mesh[i];
        if (m >= n_meshes) m = 0;  // the frame kernel has already flagged the error
        const unsigned long long addr = blas[m];
        uint4* o = out + 4 * (size_t)i;
        o[0] = make_uint4(__float_as_uint(c0.x), __float_as_uint(c1.x), __float_as_uint(c2.x), __float_as_uint(c3.x));
        o[1] = make_uint4(__float_as_uint(c0.y), __float_as_uint(c1.y), __float_as_uint(c2.y), __float_as_uint(c3.y));
        o[2] = make_uint4(__float_as_uint(c0.z), __float_as_uint(c1.z), __float_as_uint(c2.z), __float_as_uint(c3.z));
        o[3] = make_uint4(((draw_index_base + i) & 0xffffffu) | 0xff000000u, 1u << 24, (uint32_t)addr, (uint32_t)(addr >> 32));
    }
}

// ---- FREEZE_CULLING frames (cull_pass_bypass, cull_pipeline.rs:331-421) ---------------------
// With culling frozen the reference keeps the last draw lists and still runs
// model_matrix_calculation + model_matrices_upload: model matrices only, 32 bytes in, 64 out.
__global__ void __launch_bounds__(256) model_only(const float4* __restrict__ pos_scale, const float4* __restrict__ rot, uint32_t n,
                                                  float4* __restrict__ model) {
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const Mat34 m = model_matrix(pos_scale[i], rot[i]);
        float4* o = model + 4 * (size_t)i;
        o[0] = make_float4(m.m00, m.m10, m.m20, 0.f);
        o[1] = make_float4(m.m01, m.m11, m.m21, 0.f);
        o[2] = make_float4(m.m02, m.m12, m.m22, 0.f);
        o[3] = make_float4(m.tx, m.ty, m.tz, 1.f);
    }
}

}  // namespace rcull
