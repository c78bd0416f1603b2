// rcull_aux.cuh -- small streaming kernels around the frame kernel: the
// range-shard merge step (firstIndex rebasing into the merged draw list) and
// the counter-based synthetic scene generator used by the harness.
#pragma once
#include <cuda.h>  // CUtensorMap (type only)

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

// ---- range-shard merge over peer memory (NVLink peer / NVLS multicast stores) --------
// What crosses NVLink is the information content of a survivor, not its command: an 8-byte RECORD
// {draw_index, slot = mesh * 6 + lod}.  The frame kernel itself sends it -- at emission time, one store per
// peer (or ONE multimem.st when the window has an NVLS multicast mapping) into the region of every rank's
// WINDOW that belongs to the sending rank -- so the transfer is spread over the whole frame, tile by tile, and
// there is no exchange kernel.  When a rank's frame kernel has sent everything it stores its per-view
// (count, index total) words and a tag into every window.  Every rank then EXPANDS the records of all ranks
// into the merged draw lists locally (merge_expand): indexCount / vertexOffset come from the mesh table,
// firstIndex from a running sum over the merged order -- the rank prefix of SURVEY 8e falls out of the same scan.
// Window layout (identical on every rank): one SLOT per (parity p = frame epoch & 1, source rank r) at
// (p * world + r) * slot_stride, holding everything that rank has to say about that frame in one contiguous piece --
//   records [max_views][cap_r] x 8 B | checkpoints [max_views][cap_r / 256 + 1] x 4 B | meta: 24 x 8 B
// (meta = counts[8], index totals[8], send bounds[8]) -- so that the copy-engine transport moves a slot with ONE peer
// copy; then the MergeSync mailbox (tags) at sync_off.
constexpr int kMaxRanks = 8;
constexpr int kExpandTile = 256;  // records per expand tile (one warp) = the sender's checkpoint spacing

struct MergeSync {
    // A slot's meta words: counts[8] then index totals[8] of the frame whose epoch is frame_tag[parity][rank], then
    // send bounds[8]: b + 1 = only the first b records of that view were DELIVERED into this window, the rest stay in
    // the source rank's own window and are read from there (copy-engine transport with rcull_merge_set_send_bound);
    // 0 = everything was delivered.  The frame kernel writes words 0..15 only.
    unsigned long long frame_tag[2][kMaxRanks];  // every record of that frame from that rank has landed in this window
    unsigned long long expand_tag[kMaxRanks];    // rank r has finished reading ITS window for every epoch <= this
};

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
__device__ __forceinline__ uint32_t ld_relaxed_gpu_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// A store through an NVLS multicast address: one store leaves the GPU, the NVSwitch writes it into every
// member's memory (egress per rank = its records once, not once per peer).
__device__ __forceinline__ void multimem_st_v2(void* mc, uint32_t a, uint32_t b) {
    asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc), "f"(__uint_as_float(a)), "f"(__uint_as_float(b)) : "memory");
}

// What the frame kernel needs to send records (part of FrameArgs; world == 0: no merge attached).
struct RecordArgs {
    unsigned char* peer[kMaxRanks];  // the n_dst windows this kernel stores into: every rank's (own included, peer-mapped), or --
                                     // copy-engine transport -- only its own (rcull_merge_send then copies the region to the peers)
    unsigned char* own;              // this rank's window (mailbox with the peers' expand tags)
    unsigned char* mc;               // NVLS multicast mapping of all windows, or null
    size_t base;                     // byte offset of this frame's records of view 0 from this rank
    size_t view_stride;              // cap_r * 8
    size_t sync_off;                 // MergeSync inside the window
    size_t meta_base;                // byte offset of this frame's meta words from this rank
    unsigned long long epoch;        // number of this merged frame (the same on every rank), from 1
    uint32_t* cta_done;              // local counter: CTAs of this launch whose records are out
    size_t ckpt_base;                // byte offset of this frame's checkpoints of view 0 from this rank
    size_t ckpt_view_stride;         // bytes of checkpoints per view
    uint32_t world, rank, cap_r, n_dst;
};

__device__ __forceinline__ void record_store(const RecordArgs& r, uint32_t view, uint32_t pos, uint32_t draw_index, uint32_t slot) {
    if (pos >= r.cap_r) return;  // clamp writes, like cap_visible; the count still says how many there were
    const size_t off = r.base + (size_t)view * r.view_stride + (size_t)pos * 8;
    if (r.mc != nullptr) {
        multimem_st_v2(r.mc + off, draw_index, slot);
    } else {
        for (uint32_t p = 0; p < r.n_dst; ++p) *reinterpret_cast<uint2*>(r.peer[p] + off) = make_uint2(draw_index, slot);
    }
}
// Every kExpandTile-th survivor of a view also leaves its LOCAL firstIndex behind (4 bytes per 256 survivors): the
// receivers restart their running sums there, so no expand tile depends on another.
__device__ __forceinline__ void record_checkpoint(const RecordArgs& r, uint32_t view, uint32_t pos, uint32_t first_index_local) {
    if (pos >= r.cap_r || (pos % kExpandTile) != 0u) return;
    const size_t off = r.ckpt_base + (size_t)view * r.ckpt_view_stride + (size_t)(pos / kExpandTile) * 4;
    for (uint32_t p = 0; p < r.n_dst; ++p) *reinterpret_cast<uint32_t*>(r.peer[p] + off) = first_index_local;
}
// Back-pressure: the records of epoch e reuse the region of epoch e - 2, which every rank must have expanded.
__device__ __forceinline__ void record_wait_window_free(const RecordArgs& r, uint32_t tid) {
    if (tid < r.world) {
        const MergeSync* mine = reinterpret_cast<const MergeSync*>(r.own + r.sync_off);
        while (ld_acquire_sys(&mine->expand_tag[tid]) + 2 < r.epoch) __nanosleep(200);
    }
}
// Called by ONE thread of a CTA after every record-storing thread of the CTA has fenced (system scope) and the CTA
// has synchronised: the last CTA of the launch sends the frame's counts / index totals and raises the tag everywhere.
__device__ __forceinline__ void record_finish(const RecordArgs& r, const uint32_t* counts, const unsigned long long* totals, uint32_t n_views) {
    const uint32_t done = atomicAdd(r.cta_done, 1u);
    if (done != gridDim.x - 1) return;
    *r.cta_done = 0;  // the next launch on this stream starts from zero
    __threadfence();
    const uint32_t par = (uint32_t)(r.epoch & 1ull);
    for (uint32_t k = 0; k < 8; ++k) {
        const unsigned long long c = k < n_views ? ld_relaxed_gpu_u32(counts + k) : 0ull;
        const unsigned long long t = k < n_views ? ld_relaxed(reinterpret_cast<const uint64_t*>(totals + k)) : 0ull;
        for (uint32_t p = 0; p < r.n_dst; ++p) {
            unsigned long long* mt = reinterpret_cast<unsigned long long*>(r.peer[p] + r.meta_base);
            mt[k] = c;
            mt[8 + k] = t;
        }
    }
    __threadfence_system();
    for (uint32_t p = 0; p < r.n_dst; ++p)
        st_release_sys(&reinterpret_cast<MergeSync*>(r.peer[p] + r.sync_off)->frame_tag[par][r.rank], r.epoch);
}

// An empty frame (n == 0) of a context with a merge attached: zero counts and the tag, nothing else.
__global__ void record_empty_frame(const __grid_constant__ RecordArgs r, const uint32_t* counts, const unsigned long long* totals) {
    record_wait_window_free(r, threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) record_finish(r, counts, totals, 0);
}

// ---- merge_expand: every rank's records -> the merged draw lists, on the receiving GPU -------------------------
// A tile is one CHUNK of kExpandTile (256) consecutive records of one (view, source rank), and it belongs to ONE WARP.
// Its commands start at merged position  (survivors of earlier ranks) + chunk * 256  and its first firstIndex is
// (index totals of earlier ranks) + (the sender's checkpoint of that chunk): both known without looking at any other
// tile, so the kernel is a plain streaming pass -- no look-back, no ticket, no barrier, no dependency between warps;
// latency is hidden by the number of independent warps (32 per SM).
// Two launch shapes of the same body:
//   merge_expand      CTAs of 8 warps, four per SM over the whole chip: the expand runs between frames;
//   merge_expand_big  ONE CTA of 32 warps and 200 KB of shared memory per SM: a grid of R such CTAs occupies exactly R
//                     SMs while the frame kernel (one persistent CTA per SM, grid cut to the other SMs) runs on the
//                     rest -- a deterministic split of the chip, whichever kernel is placed first.  (Light expand CTAs
//                     SHARING the frame kernel's SMs were measured and rejected: the frame kernel is latency-bound and
//                     its look-back couples every CTA to the slowest SM.)
// A warp stages its 256 commands in shared memory and writes them as 16-byte vectors: the 20-byte stride of a
// command would otherwise cost five scattered 4-byte stores per survivor.
constexpr int kExpandItems = kExpandTile / 32;           // 8 records per lane
constexpr int kExpandStageWords = kExpandTile * 5 + 12;  // + the alignment shift, rounded to 16 bytes

struct ExpandArgs {
    const unsigned char* window;       // this rank's window
    unsigned char* peer[kMaxRanks];    // every rank's window (for the expand tag)
    size_t sync_off, view_stride, ckpt_view_stride;
    size_t slot_stride, ckpt_in_slot, meta_in_slot;  // a (parity, rank) slot: records, then checkpoints, then meta words
    unsigned long long epoch;
    uint32_t world, rank, n_views, max_views, cap_r, cap;
    uint32_t* cmds[8];                 // merged lists (5 words per command), or slot keys in pair mode
    uint32_t* visible[8];              // merged visible indices, or draw indices in pair mode
    uint32_t* counts;                  // [max_views] merged survivor counts
    unsigned long long* totals;        // [max_views] merged index totals
    const uint4* lod_tab;              // index_len, first_index, vertex_offset per (mesh, lod) slot
    uint32_t n_slots;
    uint32_t pairs;                    // per-mesh mode: write (slot key, draw_index) pairs in merged draw order for the bucketing passes
    uint32_t* cta_done;                // CTAs of this launch that are through
};

template <int kWarps>
__device__ __forceinline__ void merge_expand_body(const ExpandArgs& a) {
    extern __shared__ uint4 s_stage4[];  // per warp: kExpandStageWords words
    __shared__ uint32_t s_cnt[8][kMaxRanks];      // records (view, rank) stored
    __shared__ uint32_t s_pos[8][kMaxRanks + 1];  // merged position of the rank's first survivor (exclusive prefix of s_cnt, clamped to cap)
    __shared__ uint32_t s_first[8][kMaxRanks];    // firstIndex the rank's first survivor starts from (exclusive prefix of the index totals)
    __shared__ uint32_t s_tiles[8 * kMaxRanks + 1];  // exclusive prefix of the tile counts over (view, rank)
    __shared__ unsigned long long s_meta[3][8][kMaxRanks];  // counts, index totals, send bounds: one mailbox word per thread
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t par = (uint32_t)(a.epoch & 1ull);
    const MergeSync* mine = reinterpret_cast<const MergeSync*>(a.window + a.sync_off);
    if (tid < a.world) {
        while (ld_acquire_sys(&mine->frame_tag[par][tid]) != a.epoch) __nanosleep(200);
    }
    __syncthreads();
    if (tid < 3 * 8 * kMaxRanks) {
        const uint32_t what = tid / (8 * kMaxRanks), k = (tid / kMaxRanks) % 8, r = tid % kMaxRanks;
        s_meta[what][k][r] = (k < a.n_views && r < a.world)
                                 ? ld_relaxed_sys(reinterpret_cast<const unsigned long long*>(a.window + (size_t)(par * a.world + r) * a.slot_stride + a.meta_in_slot) + what * 8 + k)
                                 : 0ull;
    }
    __syncthreads();
    if (tid < a.n_views) {
        unsigned long long run = 0, all = 0, len_all = 0;
        for (uint32_t r = 0; r < a.world; ++r) {
            const unsigned long long c = s_meta[0][tid][r];
            const uint32_t pos = (uint32_t)(run < a.cap ? run : a.cap);
            s_pos[tid][r] = pos;
            const uint32_t stored = (uint32_t)(c < a.cap_r ? c : a.cap_r);  // what the rank could send
            s_cnt[tid][r] = stored < a.cap - pos ? stored : a.cap - pos;    // what the merged list can take
            s_first[tid][r] = (uint32_t)len_all;                            // u32: firstIndex wraps like the reference's
            run += stored;
            all += c;
            len_all += s_meta[1][tid][r];
        }
        s_pos[tid][a.world] = (uint32_t)(run < a.cap ? run : a.cap);
        if (blockIdx.x == 0) {
            a.counts[tid] = (uint32_t)all;  // the true count, even when the lists are clamped
            a.totals[tid] = len_all;
        }
    }
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
        for (uint32_t k = 0; k < a.n_views; ++k)
            for (uint32_t r = 0; r < a.world; ++r) {
                s_tiles[k * a.world + r] = t;
                t += (s_cnt[k][r] + kExpandTile - 1) / kExpandTile;
            }
        s_tiles[a.n_views * a.world] = t;
    }
    __syncthreads();
    const uint32_t n_pairs = a.n_views * a.world, n_tiles = s_tiles[n_pairs];
    uint32_t* const stage = reinterpret_cast<uint32_t*>(s_stage4) + warp * kExpandStageWords;
    uint32_t kr = 0;  // (view, rank) of the current tile; a warp's tiles only move forward
    for (uint32_t t = blockIdx.x * kWarps + warp; t < n_tiles; t += gridDim.x * kWarps) {
        while (kr + 1 < n_pairs && t >= s_tiles[kr + 1]) ++kr;
        const uint32_t k = kr / a.world, r = kr % a.world, c = t - s_tiles[kr];
        const uint32_t n_kr = s_cnt[k][r], first_local = c * kExpandTile;
        const uint32_t n_here = n_kr - first_local < (uint32_t)kExpandTile ? n_kr - first_local : (uint32_t)kExpandTile;
        // delivered into this window, or -- past the sender's bound -- still in the sender's own window (same offset there)
        const uint32_t bound1 = (uint32_t)s_meta[2][k][r];
        const unsigned char* src = (bound1 == 0u || r == a.rank || first_local + n_here <= bound1 - 1u) ? a.window : a.peer[r];
        const size_t slot = (size_t)(par * a.world + r) * a.slot_stride;
        const uint2* recs = reinterpret_cast<const uint2*>(src + slot + (size_t)k * a.view_stride) + first_local;
        const uint32_t dst0 = s_pos[k][r] + first_local;  // merged position of the tile's first survivor
        uint2 rec[kExpandItems];  // lane-striped: coalesced 8-byte loads
#pragma unroll
        for (int q = 0; q < kExpandItems; ++q) rec[q] = lane + 32 * q < n_here ? __ldcg(recs + lane + 32 * q) : make_uint2(0u, 0xffffffffu);
        if (a.pairs) {  // merged (slot key, draw_index) pairs: the input of the bucketing passes
#pragma unroll
            for (int q = 0; q < kExpandItems; ++q)
                if (lane + 32 * q < n_here) {
                    a.cmds[k][dst0 + lane + 32 * q] = rec[q].y;
                    a.visible[k][dst0 + lane + 32 * q] = rec[q].x;
                }
            continue;
        }
        const uint32_t ckpt = __ldcg(reinterpret_cast<const uint32_t*>(a.window + slot + a.ckpt_in_slot + (size_t)k * a.ckpt_view_stride) + c);  // checkpoints are always delivered whole
        uint32_t len[kExpandItems], voff[kExpandItems];
#pragma unroll
        for (int q = 0; q < kExpandItems; ++q) {
            const bool ok = lane + 32 * q < n_here;
            const uint4 L = ok ? __ldg(a.lod_tab + (rec[q].y < a.n_slots ? rec[q].y : 0u)) : make_uint4(0, 0, 0, 0);
            len[q] = L.x;
            voff[q] = L.z;
        }
        // running sum of index_len inside the chunk, from the sender's checkpoint
        uint32_t carry = s_first[k][r] + ckpt;
        const size_t g0 = (size_t)dst0 * 5u;
        const uint32_t sh = (uint32_t)(g0 & 3u);  // word address of the tile's first command word in the list, and its phase in a 16-byte unit
#pragma unroll
        for (int q = 0; q < kExpandItems; ++q) {
            const uint32_t incl = warp_incl_scan_u32(len[q], lane);
            const uint32_t l = lane + 32 * q;
            if (l < n_here) {
                uint32_t* cw = stage + sh + l * 5;  // stride of 5 words across the lanes: no bank conflict
                cw[0] = len[q];                      // indexCount
                cw[1] = 1u;                          // instanceCount
                cw[2] = carry + incl - len[q];       // firstIndex: running sum over the MERGED order
                cw[3] = voff[q];                     // vertexOffset
                cw[4] = rec[q].x;                    // firstInstance = global draw_index
                a.visible[k][dst0 + l] = rec[q].x;
            }
            carry += __shfl_sync(kFullMask, incl, 31);
        }
        __syncwarp();
        {   // staged words [sh, sh + n_words) go to list words [g0, g0 + n_words): word i of the staging area and word
            // g0 - sh + i of the list have the same phase, so whole 16-byte units move as vectors
            uint32_t* const dst = a.cmds[k] + (g0 - sh);
            const uint32_t end = sh + n_here * 5u;
            const uint32_t v0 = sh ? 1u : 0u, v1 = end / 4u;
            for (uint32_t v = v0 + lane; v < v1; v += 32) reinterpret_cast<uint4*>(dst)[v] = reinterpret_cast<const uint4*>(stage)[v];
            if (sh && lane >= sh && lane < 4u && lane < end) dst[lane] = stage[lane];         // head: up to the first boundary
            if (v1 >= v0 && lane < end - 4u * v1 && 4u * v1 >= sh) dst[4u * v1 + lane] = stage[4u * v1 + lane];  // tail
        }
        __syncwarp();  // the staging area is free again
    }
    // this rank is done reading its window for this epoch: tell everyone (they may reuse the region two frames on)
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const uint32_t done = atomicAdd(a.cta_done, 1u);
        if (done == gridDim.x - 1) {
            *a.cta_done = 0;
            __threadfence_system();
            for (uint32_t p = 0; p < a.world; ++p)
                st_release_sys(&reinterpret_cast<MergeSync*>(a.peer[p] + a.sync_off)->expand_tag[a.rank], a.epoch);
        }
    }
}

__global__ void __launch_bounds__(256, 4) merge_expand(const __grid_constant__ ExpandArgs a) { merge_expand_body<8>(a); }
__global__ void __launch_bounds__(1024, 1) merge_expand_big(const __grid_constant__ ExpandArgs a) { merge_expand_body<32>(a); }
constexpr size_t kExpandSmem = 8 * (size_t)kExpandStageWords * 4;
constexpr size_t kExpandBigSmem = 32 * (size_t)kExpandStageWords * 4 + 36 * 1024;  // + ballast: no second CTA, of any kernel, fits beside it

// ---- per-mesh instanced emission (BASELINE config 4; compact_draw_stream.comp:41 keep-rule on
// instanceCount) ----------------------------------------------------------------------
// The frame kernel leaves, per view, the survivors in draw_index order as (slot key,
// draw_index) pairs.  A stable counting sort buckets them by slot:
//   bucket_hist     one warp per tile of `tile_len` pairs: per-slot counts -> hist[tile][slot]
//   bucket_colscan  per slot, exclusive scan down the tiles (in place) + slot totals; the LAST CTA of a view to finish
//                   goes on with bucket_cmds: exclusive scan over the slot totals -> bucket starts, one command per
//                   NON-EMPTY (mesh, lod) slot in slot order (one launch less on a path that is launch-bound)
//   bucket_scatter  same tiling as bucket_hist: rank of each pair inside its slot =
//                   bucket start + tile offset + arrivals before it in the tile (stable)
// Slot counters live in shared memory, kBucketChunk slots per warp at a time; tables with more
// slots take one sweep of the tile per chunk.  The survivor count is read on the device.
// The three launches are programmatic dependent launches: a kernel's CTAs are placed while its predecessor drains and
// wait in pdl_wait() for the predecessor's memory to be visible, instead of paying a full launch latency each.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
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
    uint32_t* colscan_done;     // [view] CTAs of bucket_colscan that are through (the last one builds the commands)
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
    pdl_wait();
    bucket_sweep<false>(a);
}
__global__ void __launch_bounds__(kBucketWarps * 32) bucket_scatter(const __grid_constant__ BucketArgs a) {
    pdl_wait();
    bucket_sweep<true>(a);
}

// CTA = 32 slots (x) by 32 runs of tiles (y): every thread sums its run of tiles for one slot,
// the runs are scanned through shared memory, and a second walk writes the exclusive offsets.
__device__ __forceinline__ void bucket_cmds_body(const BucketArgs& a, const BucketView& bv);

__global__ void __launch_bounds__(1024) bucket_colscan(const __grid_constant__ BucketArgs a) {
    __shared__ uint32_t s_run[32][33];
    __shared__ uint32_t s_last;
    pdl_wait();
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
    // the last CTA of this view to get here has every slot total in front of it: it builds the commands
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t done = atomicAdd(a.colscan_done + bv.view, 1u);
        s_last = done == gridDim.x - 1 ? 1u : 0u;
        if (s_last) a.colscan_done[bv.view] = 0;  // ready for the next frame
    }
    __syncthreads();
    if (s_last == 0u) return;
    __threadfence();
    bucket_cmds_body(a, bv);
}

__device__ __forceinline__ void bucket_cmds_body(const BucketArgs& a, const BucketView& bv) {
    __shared__ uint32_t warp_cmds[32], warp_inst[32];
    __shared__ uint32_t run_cmds, run_inst;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t* slot_total = a.slot_total + (size_t)bv.view * a.n_slots;
    uint32_t* slot_base = a.slot_base + (size_t)bv.view * a.n_slots;
    if (tid == 0) run_cmds = run_inst = 0;
    __syncthreads();
    for (uint32_t s0 = 0; s0 < a.n_slots; s0 += blockDim.x) {
        const uint32_t s = s0 + tid;
        const uint32_t cnt = s < a.n_slots ? __ldcg(slot_total + s) : 0u;  // written by other CTAs of this launch: L2, not this SM's L1
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

// ---- 64-byte records out through a swizzled shared-memory box and a TMA tensor store -----------------------------
// One thread produces one 64-byte record (a model matrix, an instance record); written directly that is four 16-byte
// stores per thread at a 64-byte stride -- quarter-line stores that the frame kernel measured 37 % slower than this:
// the warp writes its 32 records into a 2 KB box in the TMA 128-byte swizzle pattern (conflict-free STS.128) and one
// lane issues ONE 2-D tensor store (SASS UTMASTG) for the 32 records: full-line writes, no per-thread store instructions.
// The tensor map views the buffer as [records / 2 rows][32 floats], boxes of 16 rows (rcull.cu make_model_map).
struct RecordBox {
    unsigned char* box;  // this warp's 2 KB, 1024-byte aligned
    uint32_t off[4];     // swizzled byte offsets of this lane's four 16-byte quarters
    __device__ __forceinline__ RecordBox(unsigned char* warp_box, uint32_t lane) : box(warp_box) {
        const uint32_t row = lane >> 1, half = (lane & 1u) * 4u;
#pragma unroll
        for (uint32_t q = 0; q < 4; ++q) off[q] = row * 128u + (((half + q) ^ (row & 7u)) << 4);
    }
    // all 32 lanes call it, with the index of the warp's FIRST record (a multiple of 32)
    __device__ __forceinline__ void store(const void* tmap, uint32_t first_record, uint32_t lane, uint4 q0, uint4 q1, uint4 q2, uint4 q3) {
        if (lane == 0) bulk_wait_read0();  // the previous store has finished reading the box
        __syncwarp();
        *reinterpret_cast<uint4*>(box + off[0]) = q0;
        *reinterpret_cast<uint4*>(box + off[1]) = q1;
        *reinterpret_cast<uint4*>(box + off[2]) = q2;
        *reinterpret_cast<uint4*>(box + off[3]) = q3;
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
            tma_store_2d(tmap, box, 0, (int)(first_record >> 1));
            bulk_commit();
        }
    }
    __device__ __forceinline__ void drain(uint32_t lane) {
        if (lane == 0) bulk_wait_read0();
    }
};
constexpr int kRecordWarps = 8;  // warps per CTA of the two record kernels below

// ---- acceleration-structure instance records ------------------------------
// build_acceleration_structures' CPU loop (acceleration_strucures.rs:419-445) writes one
// VkAccelerationStructureInstanceKHR per entity: the upper 3 rows of the model matrix row-major,
// instance_custom_index = draw_index (24 bits) | mask 0xFF, SBT offset 0 | flags
// TRIANGLE_FACING_CULL_DISABLE (= 1), and the BLAS device address of the entity's mesh.
// One thread per entity: 68 bytes in (model + mesh id), 64 bytes out.
__global__ void __launch_bounds__(kRecordWarps * 32) as_instances(const float4* __restrict__ model, const uint32_t* __restrict__ mesh,
                                                                  const unsigned long long* __restrict__ blas, uint32_t n_meshes, uint32_t n,
                                                                  uint32_t draw_index_base, const __grid_constant__ CUtensorMap out_map) {
    __shared__ __align__(1024) unsigned char s_box[kRecordWarps][2048];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    RecordBox rb(s_box[warp], lane);
    const uint32_t stride = gridDim.x * kRecordWarps * 32;
    for (uint32_t i0 = (blockIdx.x * kRecordWarps + warp) * 32; i0 < n; i0 += stride) {
        const uint32_t i = i0 + lane < n ? i0 + lane : n - 1;  // lanes past the end repeat the last entity into padding slots
        const float4 c0 = model[4 * (size_t)i], c1 = model[4 * (size_t)i + 1], c2 = model[4 * (size_t)i + 2], c3 = model[4 * (size_t)i + 3];
        uint32_t m = mesh[i];
        if (m >= n_meshes) m = 0;  // the frame kernel has already flagged the error
        const unsigned long long addr = blas[m];
        rb.store(&out_map, i0, lane,
                 make_uint4(__float_as_uint(c0.x), __float_as_uint(c1.x), __float_as_uint(c2.x), __float_as_uint(c3.x)),
                 make_uint4(__float_as_uint(c0.y), __float_as_uint(c1.y), __float_as_uint(c2.y), __float_as_uint(c3.y)),
                 make_uint4(__float_as_uint(c0.z), __float_as_uint(c1.z), __float_as_uint(c2.z), __float_as_uint(c3.z)),
                 make_uint4(((draw_index_base + i) & 0xffffffu) | 0xff000000u, 1u << 24, (uint32_t)addr, (uint32_t)(addr >> 32)));
    }
    rb.drain(lane);
}

// ---- FREEZE_CULLING frames (cull_pass_bypass, cull_pipeline.rs:331-421) ---------------------
// With culling frozen the reference keeps the last draw lists and still runs
// model_matrix_calculation + model_matrices_upload: model matrices only, 32 bytes in, 64 out.
__global__ void __launch_bounds__(kRecordWarps * 32) model_only(const float4* __restrict__ pos_scale, const float4* __restrict__ rot, uint32_t n,
                                                                const __grid_constant__ CUtensorMap model_map) {
    __shared__ __align__(1024) unsigned char s_box[kRecordWarps][2048];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    RecordBox rb(s_box[warp], lane);
    const uint32_t stride = gridDim.x * kRecordWarps * 32;
    for (uint32_t i0 = (blockIdx.x * kRecordWarps + warp) * 32; i0 < n; i0 += stride) {
        const uint32_t i = i0 + lane;
        // padding slots of the last group become identity entities, as in the frame kernel
        const float4 ps = i < n ? pos_scale[i] : make_float4(0.f, 0.f, 0.f, 1.f), q = i < n ? rot[i] : make_float4(0.f, 0.f, 0.f, 1.f);
        const Mat34 m = model_matrix(ps, q);
        rb.store(&model_map, i0, lane, make_uint4(__float_as_uint(m.m00), __float_as_uint(m.m10), __float_as_uint(m.m20), 0u),
                 make_uint4(__float_as_uint(m.m01), __float_as_uint(m.m11), __float_as_uint(m.m21), 0u),
                 make_uint4(__float_as_uint(m.m02), __float_as_uint(m.m12), __float_as_uint(m.m22), 0u),
                 make_uint4(__float_as_uint(m.tx), __float_as_uint(m.ty), __float_as_uint(m.tz), __float_as_uint(1.f)));
    }
    rb.drain(lane);
}

}  // namespace rcull
