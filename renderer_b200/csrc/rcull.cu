// rcull.cu -- C ABI of librcull.so (see include/rcull.h).  Host side: context,
// device buffers, launch configuration.  There is no CPU fallback anywhere in
// this library: every compute entry point launches sm_100a kernels or fails.
#include "rcull.h"

#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>  // header-only NVTX 3: ranges cost nothing unless a tool (nsys / ncu --nvtx) is attached

#include <cmath>
#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>


#include "rcull_aux.cuh"
#include "rcull_frame.cuh"
#include "rcull_tri.cuh"

using namespace rcull;

static_assert(sizeof(rcull_mesh) == 80, "rcull_mesh layout");
static_assert(sizeof(rcull_draw_cmd) == 20, "VkDrawIndexedIndirectCommand stride");
static_assert(sizeof(rcull_config) == 56 && sizeof(rcull_view) == 96, "rcull_config / rcull_view layout");
static_assert(RCULL_MAX_VIEWS == kMaxViews && RCULL_MAX_LODS == kMaxLods, "header/kernel constants");
static_assert(frame_smem_bytes<1>() <= kSmemPerCta && frame_smem_bytes<5>() <= kSmemPerCta && frame_smem_bytes<8>() <= kSmemPerCta,
              "frame kernel must fit kCtasPerSm CTAs per SM");

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess)                                                                     \
            return fail(RCULL_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, \
                        __LINE__);                                                                 \
    } while (0)

// One NVTX range per entry point that enqueues device work, named like the scope!() of the system it
// replaces (src/ecs.rs:56,59,142,145; cull_pipeline.rs:104,107,439) -- domain "rcull".
struct Scope {
    static nvtxDomainHandle_t domain() {
        static nvtxDomainHandle_t d = nvtxDomainCreateA("rcull");
        return d;
    }
    explicit Scope(const char* name) {
        nvtxEventAttributes_t a = {};
        a.version = NVTX_VERSION;
        a.size = NVTX_EVENT_ATTRIB_STRUCT_SIZE;
        a.messageType = NVTX_MESSAGE_TYPE_ASCII;
        a.message.ascii = name;
        nvtxDomainRangePushEx(domain(), &a);
    }
    ~Scope() { nvtxDomainRangePop(domain()); }
    Scope(const Scope&) = delete;
    Scope& operator=(const Scope&) = delete;
};

struct OutputSet {
    float4* model = nullptr;
    float* aabb = nullptr;
    uint8_t* flags[kMaxViews] = {};
    uint32_t* cmds[kMaxViews] = {};
    uint32_t* visible[kMaxViews] = {};
    uint32_t* keys[kMaxViews] = {};     // per-mesh mode: slot key of every survivor, draw_index order
    uint32_t* vis_tmp[kMaxViews] = {};  // per-mesh mode: draw_index of every survivor before bucketing
    uint32_t* counts = nullptr;
    unsigned long long* totals = nullptr;
    uint32_t* src_slot = nullptr;  // triangle stage: (mesh, lod) slot of every camera-view command
    CUtensorMap model_map;  // model buffer as [cap/2 rows][32 floats], 16-row boxes, 128-B swizzle
};

}  // namespace

struct rcull_merge;
struct rcull_ctx {
    rcull_config cfg{};
    int sm_count = 0;
    uint64_t cap = 0;      // entity capacity padded to whole tiles
    uint64_t cap_vis = 0;  // per-view draw-list capacity
    float4* pos_scale = nullptr;
    float4* rot = nullptr;
    uint32_t* mesh = nullptr;
    float4* mesh_aabb = nullptr;
    uint4* lod_tab = nullptr;
    uint32_t* mesh_len1 = nullptr;
    uint32_t n_meshes = 0;
    uint32_t len_small = 0;
    uint32_t* bucket_hist = nullptr;  // per-mesh mode scratch: [tiles][slots] counts, see rcull_aux.cuh
    uint32_t* slot_total = nullptr;
    uint32_t* slot_base = nullptr;
    uint32_t* bucket_done = nullptr;  // [kMaxViews] bucket_colscan CTAs through, per view (zero between launches)
    size_t bucket_hist_words = 0;
    uint64_t cap_cmds = 0;  // command capacity per view: cap_vis, or max_meshes * 6 in per-mesh mode
    int reserved_sms = 0;  // SMs the frame kernel's grid leaves to a concurrent expand kernel (rcull_merge_set_overlap)
    struct rcull_merge* merge = nullptr;  // attached range-shard merge: frames send their survivors as records (rcull_merge_create)
    OutputSet out[3];
    ulonglong2* status = nullptr;
    size_t status_bytes = 0;
    uint32_t* ticket = nullptr;  // [0] = tile ticket, [1] = error flag
    uint32_t ticket_base = 0;
    uint32_t epoch = 0;
    uint64_t launches = 0;
    float lod_d2_max = 0.f;
    int blocks_per_sm[kMaxViews + 1][2] = {};
    cudaStream_t own_stream = nullptr;
    uint32_t* h_counts = nullptr;  // pinned
    unsigned long long* h_totals = nullptr;
    // stage 2 (rcull_tri.cuh); single-buffered like CullPassData (cull_pipeline.rs:74-84)
    float* positions = nullptr;
    uint32_t* indices = nullptr;
    uint64_t n_vertices = 0, n_indices = 0;
    // chunk table of the triangle stage (rcull_tri.cuh): built on the host from the index buffer and the mesh table
    std::vector<rcull_mesh> h_meshes;
    std::vector<uint32_t> h_indices;
    bool chunks_dirty = true;
    uint2* slot_chunks = nullptr;
    uint4* chunks = nullptr;
    uint32_t n_chunks = 0;
    bool any_window_slot = false, any_gather_slot = false;  // which of the two tri_cull kernels has anything to do
    uint32_t* tri_cmds = nullptr;
    uint32_t* tri_count = nullptr;
    uint64_t* tri_kept = nullptr;
    uint32_t* tri_refined = nullptr;
    uint32_t* tri_indices = nullptr;
    uint64_t* tri_status = nullptr;  // [4 * tiles] + 2 ticket words at the end
    size_t tri_status_bytes = 0;
    unsigned long long* blas = nullptr;  // F4: BLAS address per mesh
    uint32_t n_blas = 0;                 // entries rcull_set_blas_table filled
    // CullPassData is single-buffered in the reference (cull_pipeline.rs:70-84), only ModelData is
    // DoubleBuffered: a frozen frame on set j keeps the LISTS of the last unfrozen frame, whichever set holds them
    uint32_t lists_ix = 0;               // output set written by the last rcull_frame
    bool set_frozen[3] = {false, false, false};  // the last call on set j was rcull_frame_frozen
    uint32_t* h_error = nullptr;         // pinned: device error word read back by rcull_frame_host
    uint4* instances = nullptr;          // F4: VkAccelerationStructureInstanceKHR [cap]
    CUtensorMap instances_map;           // the instance buffer as [cap/2 rows][32 floats], like OutputSet.model_map
    std::vector<void*> allocs;
};

namespace {

template <typename T>
int dev_alloc(rcull_ctx* c, T** p, size_t count) {
    void* q = nullptr;
    CU(cudaMalloc(&q, count * sizeof(T) + 256));
    c->allocs.push_back(q);
    *p = static_cast<T*>(q);
    return 0;
}

// Output set whose draw lists belong to set j: its own, or -- after a frozen frame on j -- the set the
// last unfrozen frame wrote (cull_pass_bypass keeps CullPassData as it is, cull_pipeline.rs:331-421).
uint32_t lists_of(const rcull_ctx* c, uint32_t j) { return c->set_frozen[j] ? c->lists_ix : j; }

// Largest x with sqrtf(x) <= d, so that (sqrtf(d2) > d) == (d2 > x) exactly
// (helpers.rs:4-6; sqrtf is correctly rounded and monotone on host and device).
float lod_threshold(float d) {
    if (std::isnan(d)) return INFINITY;
    if (d < 0.f) return -1.f;
    uint32_t lo = 0, hi = 0x7f800000u;  // bit patterns of +0 .. +inf, monotone in value
    while (lo < hi) {
        uint32_t mid = lo + (hi - lo + 1) / 2;
        float x;
        memcpy(&x, &mid, 4);
        if (sqrtf(x) <= d) lo = mid; else hi = mid - 1;
    }
    float x;
    memcpy(&x, &lo, 4);
    return x;
}

// One kernel of the bucketing chain as a programmatic dependent launch: its CTAs may be placed while the kernel in front
// of it on the stream drains (they wait in pdl_wait()).  Falls back to an ordinary launch where the attribute is refused.
template <typename K>
cudaError_t launch_dependent(K kernel, dim3 grid, unsigned block, cudaStream_t s, const BucketArgs& b) {
    static bool pdl = getenv("RCULL_NO_PDL") == nullptr;
    if (pdl) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(block);
        cfg.dynamicSmemBytes = 0;
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (cudaLaunchKernelEx(&cfg, kernel, b) == cudaSuccess) return cudaSuccess;
        (void)cudaGetLastError();  // not supported here (stream kind, driver): ordinary launches from now on
        pdl = false;
    }
    kernel<<<grid, block, 0, s>>>(b);
    return cudaGetLastError();
}

// bucket_hist -> bucket_colscan (+ commands, by its last CTA) -> bucket_scatter (rcull_aux.cuh)
cudaError_t launch_bucketing(const BucketArgs& b, dim3 sweep_grid, uint32_t n_slots, uint32_t n_views, cudaStream_t s) {
    cudaError_t e = launch_dependent(bucket_hist, sweep_grid, kBucketWarps * 32, s, b);
    if (e != cudaSuccess) return e;
    e = launch_dependent(bucket_colscan, dim3((n_slots + 31) / 32, n_views), 1024, s, b);
    if (e != cudaSuccess) return e;
    return launch_dependent(bucket_scatter, sweep_grid, kBucketWarps * 32, s, b);
}

template <int V, bool M>
int configure(rcull_ctx* c) {
    auto k = transform_cull_compact<V, M>;
    CU(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)frame_smem_bytes<V>()));
    CU(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    int b = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, k, kThreads, frame_smem_bytes<V>()));
    if (b < 1) return fail(RCULL_ERR_CUDA, "transform_cull_compact<%d,%d> does not fit on an SM", V, (int)M);
    c->blocks_per_sm[V][M] = b;
    return 0;
}

template <int V>
cudaError_t launch(bool mat, const FrameArgs& a, const CUtensorMap& map, int grid, cudaStream_t s) {
    if (mat)
        transform_cull_compact<V, true><<<grid, kThreads, frame_smem_bytes<V>(), s>>>(a, map);
    else
        transform_cull_compact<V, false><<<grid, kThreads, frame_smem_bytes<V>(), s>>>(a, map);
    return cudaGetLastError();
}

// cuTensorMapEncodeTiled is a driver entry point; librcull links cudart statically
// and resolves it at run time instead of linking libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int make_model_map(CUtensorMap* map, float4* model, uint64_t cap) {
    static EncodeTiledFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (qres != cudaDriverEntryPointSuccess || !fn)
            return fail(RCULL_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
        encode = reinterpret_cast<EncodeTiledFn>(fn);
    }
    const cuuint64_t dims[2] = {32, cap / 2};      // 32 floats = 2 matrices per 128-B row
    const cuuint64_t strides[1] = {128};           // bytes between rows
    const cuuint32_t box[2] = {32, 16};            // 16 rows = 32 matrices = 2 KB
    const cuuint32_t elem[2] = {1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, model, dims, strides, box, elem,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(RCULL_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

}  // namespace

extern "C" {

const char* rcull_last_error(void) { return g_err; }
uint32_t rcull_version(void) { return (0u << 16) | 5u; }  // 0.2: triangle stage, rcull_config.capacity_tri_indices; 0.3: slot counts; 0.4: rcull_context_stream; 0.5: rcull_frame_noalloc, merge transports (rcull_merge_set_transport / send / set_send_bound)

void rcull_destroy(rcull_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->cfg.device);
    cudaDeviceSynchronize();
    for (void* p : c->allocs) cudaFree(p);
    if (c->positions) cudaFree(c->positions);
    if (c->indices) cudaFree(c->indices);
    if (c->slot_chunks) cudaFree(c->slot_chunks);
    if (c->chunks) cudaFree(c->chunks);
    if (c->h_counts) cudaFreeHost(c->h_counts);
    if (c->h_totals) cudaFreeHost(c->h_totals);
    if (c->h_error) cudaFreeHost(c->h_error);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

int rcull_create(const rcull_config* cfg, rcull_ctx** out) {
    if (!cfg || !out) return fail(RCULL_ERR_INVALID, "rcull_create: null argument");
    *out = nullptr;
    if (cfg->max_views < 1 || cfg->max_views > RCULL_MAX_VIEWS)
        return fail(RCULL_ERR_INVALID, "max_views must be 1..%d", RCULL_MAX_VIEWS);
    if (cfg->n_buffers < 1 || cfg->n_buffers > 3) return fail(RCULL_ERR_INVALID, "n_buffers must be 1..3");
    if (cfg->capacity_entities == 0 || cfg->capacity_entities > 0xfffff000ull)
        return fail(RCULL_ERR_INVALID, "capacity_entities must be 1..2^32-4096");
    if (cfg->max_meshes == 0 || (uint64_t)cfg->max_meshes * kMaxLods > 0xffffffffull)
        return fail(RCULL_ERR_INVALID, "max_meshes out of range");
    if (cfg->output_mode > RCULL_PER_MESH_INSTANCED) return fail(RCULL_ERR_INVALID, "bad output_mode");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0)
        return fail(RCULL_ERR_NO_DEVICE, "no CUDA device: librcull has no CPU path");
    if (cfg->device < 0 || cfg->device >= n_dev) return fail(RCULL_ERR_INVALID, "device %d of %d", cfg->device, n_dev);
    CU(cudaSetDevice(cfg->device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return fail(RCULL_ERR_NO_DEVICE, "device %d is sm_%d%d; librcull is built for sm_100a only", cfg->device,
                    prop.major, prop.minor);

    rcull_ctx* c = new rcull_ctx();
    c->cfg = *cfg;
    c->sm_count = prop.multiProcessorCount;
    c->cap = (cfg->capacity_entities + kTile - 1) / kTile * kTile;
    c->cap_vis = cfg->capacity_visible ? cfg->capacity_visible : cfg->capacity_entities;
    c->lod_d2_max = lod_threshold(cfg->lod_distance);
    c->cap_cmds = cfg->output_mode == RCULL_PER_MESH_INSTANCED ? (uint64_t)cfg->max_meshes * kMaxLods : c->cap_vis;
    int rc = 0;
#define TRY(x)                \
    if ((rc = (x)) != 0) {    \
        rcull_destroy(c);     \
        return rc;            \
    }
    TRY(dev_alloc(c, &c->pos_scale, c->cap));
    TRY(dev_alloc(c, &c->rot, c->cap));
    TRY(dev_alloc(c, &c->mesh, c->cap));
    TRY(dev_alloc(c, &c->mesh_aabb, (size_t)cfg->max_meshes * 2));
    TRY(dev_alloc(c, &c->lod_tab, (size_t)cfg->max_meshes * kMaxLods));
    TRY(dev_alloc(c, &c->mesh_len1, (size_t)cfg->max_meshes));
    // padding slots of the last tile are staged by TMA like real ones: keep them defined
    if (cudaMemset(c->pos_scale, 0, c->cap * 16) != cudaSuccess || cudaMemset(c->rot, 0, c->cap * 16) != cudaSuccess ||
        cudaMemset(c->mesh, 0, c->cap * 4) != cudaSuccess) {
        rcull_destroy(c);
        return fail(RCULL_ERR_CUDA, "cudaMemset of input columns failed");
    }
    for (uint32_t b = 0; b < cfg->n_buffers; ++b) {
        OutputSet& o = c->out[b];
        TRY(dev_alloc(c, &o.model, c->cap * 4));
        TRY(make_model_map(&o.model_map, o.model, c->cap));
        if (cfg->materialize_aabb) TRY(dev_alloc(c, &o.aabb, c->cap * 6));
        for (uint32_t v = 0; v < cfg->max_views; ++v) {
            if (cfg->materialize_flags) TRY(dev_alloc(c, &o.flags[v], c->cap));
            TRY(dev_alloc(c, &o.cmds[v], c->cap_cmds * 5));
            TRY(dev_alloc(c, &o.visible[v], c->cap_vis));
            if (cfg->output_mode == RCULL_PER_MESH_INSTANCED) {
                TRY(dev_alloc(c, &o.keys[v], c->cap_vis));
                TRY(dev_alloc(c, &o.vis_tmp[v], c->cap_vis));
            }
        }
        if (cfg->capacity_tri_indices) TRY(dev_alloc(c, &o.src_slot, c->cap_vis));
        TRY(dev_alloc(c, &o.counts, kMaxViews));
        TRY(dev_alloc(c, &o.totals, kMaxViews));
        cudaMemset(o.counts, 0, kMaxViews * 4);
        cudaMemset(o.totals, 0, kMaxViews * 8);
    }
    const size_t max_tiles = c->cap / kTile;
    c->status_bytes = max_tiles * cfg->max_views * sizeof(ulonglong2);
    TRY(dev_alloc(c, &c->status, max_tiles * cfg->max_views));
    TRY(dev_alloc(c, &c->ticket, 2));
    if (cfg->output_mode == RCULL_PER_MESH_INSTANCED) {
        if (c->cap_vis > 0xffffffffull) {
            rcull_destroy(c);
            return fail(RCULL_ERR_INVALID, "per-mesh mode buckets at most 2^32-1 survivors per view");
        }
        // histogram scratch: one row of slot counts per tile of pairs, capped at 128 MB over all
        // views (the kernels lengthen the tiles to fit, rcull_aux.cuh bucket_view)
        const size_t slots_max = (size_t)cfg->max_meshes * kMaxLods;
        const size_t rows = (c->cap_vis + 1023) / 1024;
        c->bucket_hist_words = rows * slots_max < (16u << 20) ? rows * slots_max : (16u << 20);
        c->bucket_hist_words *= cfg->max_views;  // every view has its own rows
        if (c->bucket_hist_words > (32u << 20)) c->bucket_hist_words = 32u << 20;
        if (c->bucket_hist_words < slots_max * cfg->max_views) c->bucket_hist_words = slots_max * cfg->max_views;
        TRY(dev_alloc(c, &c->bucket_hist, c->bucket_hist_words));
        TRY(dev_alloc(c, &c->slot_total, slots_max * cfg->max_views));
        TRY(dev_alloc(c, &c->slot_base, slots_max * cfg->max_views));
        TRY(dev_alloc(c, &c->bucket_done, (size_t)kMaxViews));
        cudaMemset(c->bucket_done, 0, kMaxViews * 4);
    }
    if (cfg->capacity_tri_indices) {
        if (cfg->output_mode != RCULL_PER_ENTITY) {
            rcull_destroy(c);
            return fail(RCULL_ERR_INVALID, "the per-triangle stage works on per-entity commands");
        }
        const size_t tiles = (c->cap_cmds + kCompactTile - 1) / kCompactTile;
        c->tri_status_bytes = (4 * tiles + 2) * 8;
        TRY(dev_alloc(c, &c->tri_cmds, c->cap_cmds * 5));
        TRY(dev_alloc(c, &c->tri_refined, c->cap_cmds));
        TRY(dev_alloc(c, &c->tri_count, 2));
        TRY(dev_alloc(c, &c->tri_kept, 1));
        TRY(dev_alloc(c, &c->tri_indices, cfg->capacity_tri_indices));
        TRY(dev_alloc(c, &c->tri_status, 4 * tiles + 2));
        cudaMemset(c->tri_count, 0, 8);
        cudaMemset(c->tri_kept, 0, 8);
    }
    if (cudaMemset(c->status, 0, c->status_bytes) != cudaSuccess || cudaMemset(c->ticket, 0, 8) != cudaSuccess) {
        rcull_destroy(c);
        return fail(RCULL_ERR_CUDA, "cudaMemset of scan state failed");
    }
    if (cudaMallocHost((void**)&c->h_counts, kMaxViews * 4) != cudaSuccess ||
        cudaMallocHost((void**)&c->h_totals, kMaxViews * 8) != cudaSuccess ||
        cudaMallocHost((void**)&c->h_error, 4) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        rcull_destroy(c);
        return fail(RCULL_ERR_CUDA, "host staging / stream creation failed");
    }
    TRY((configure<1, false>(c)));
    TRY((configure<1, true>(c)));
    TRY((configure<2, false>(c)));
    TRY((configure<2, true>(c)));
    TRY((configure<3, false>(c)));
    TRY((configure<3, true>(c)));
    TRY((configure<4, false>(c)));
    TRY((configure<4, true>(c)));
    TRY((configure<5, false>(c)));
    TRY((configure<5, true>(c)));
    TRY((configure<6, false>(c)));
    TRY((configure<6, true>(c)));
    TRY((configure<7, false>(c)));
    TRY((configure<7, true>(c)));
    TRY((configure<8, false>(c)));
    TRY((configure<8, true>(c)));
#undef TRY
    if (cudaError_t e = cudaDeviceSynchronize(); e != cudaSuccess) {
        rcull_destroy(c);
        return fail(RCULL_ERR_CUDA, "rcull_create: %s", cudaGetErrorString(e));
    }
    *out = c;
    return RCULL_OK;
}

int rcull_set_mesh_table(rcull_ctx* c, const rcull_mesh* meshes, uint32_t n) {
    if (!c || !meshes) return fail(RCULL_ERR_INVALID, "rcull_set_mesh_table: null argument");
    if (n == 0 || n > c->cfg.max_meshes) return fail(RCULL_ERR_INVALID, "mesh count %u exceeds max_meshes %u", n, c->cfg.max_meshes);
    std::vector<float4> aabb(2 * (size_t)n);
    std::vector<uint4> lods((size_t)n * kMaxLods, make_uint4(0, 0, 0, 0));
    std::vector<uint32_t> len1(n);
    uint32_t max_len = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const rcull_mesh& m = meshes[i];
        // pick_lod .expect("empty index buffer LODs") (helpers.rs:7-9)
        if (m.n_lods < 1 || m.n_lods > kMaxLods) return fail(RCULL_ERR_INVALID, "mesh %u: n_lods %u not in 1..%d", i, m.n_lods, kMaxLods);
        float nl;
        memcpy(&nl, &m.n_lods, 4);
        aabb[2 * i] = make_float4(m.aabb_min[0], m.aabb_min[1], m.aabb_min[2], nl);
        float l0;  // index_len of lod 0 rides in the spare word; lod 1's (the only other one pick_lod returns) in len1
        memcpy(&l0, &m.lod[0].index_len, 4);
        aabb[2 * i + 1] = make_float4(m.aabb_max[0], m.aabb_max[1], m.aabb_max[2], l0);
        len1[i] = m.n_lods > 1 ? m.lod[1].index_len : m.lod[0].index_len;
        for (uint32_t l = 0; l < m.n_lods; ++l) {
            if (m.lod[l].index_len > max_len) max_len = m.lod[l].index_len;
            lods[(size_t)i * kMaxLods + l] = make_uint4(m.lod[l].index_len, m.lod[l].first_index, (uint32_t)m.vertex_offset, 0);
        }
    }
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaMemcpy(c->mesh_aabb, aabb.data(), aabb.size() * sizeof(float4), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->lod_tab, lods.data(), lods.size() * sizeof(uint4), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->mesh_len1, len1.data(), len1.size() * 4, cudaMemcpyHostToDevice));
    c->n_meshes = n;
    c->len_small = max_len < (1u << 25) ? 1u : 0u;
    if (c->cfg.capacity_tri_indices) {
        c->h_meshes.assign(meshes, meshes + n);
        c->chunks_dirty = true;
    }
    return RCULL_OK;
}

int rcull_upload_entities(rcull_ctx* c, uint64_t first, uint64_t n, const float* pos_scale, const float* rot,
                          const uint32_t* mesh, void* stream) {
    Scope scope("rcull::upload_entities");
    if (!c || (n && (!pos_scale || !rot || !mesh))) return fail(RCULL_ERR_INVALID, "rcull_upload_entities: null argument");
    if (first + n > c->cfg.capacity_entities) return fail(RCULL_ERR_INVALID, "entities [%llu, %llu) exceed capacity %llu", (unsigned long long)first, (unsigned long long)(first + n), (unsigned long long)c->cfg.capacity_entities);
    if (n == 0) return RCULL_OK;
    CU(cudaSetDevice(c->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    CU(cudaMemcpyAsync(c->pos_scale + first, pos_scale, n * 16, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(c->rot + first, rot, n * 16, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(c->mesh + first, mesh, n * 4, cudaMemcpyHostToDevice, s));
    return RCULL_OK;
}

int rcull_context_stream(rcull_ctx* c, void** stream) {
    if (!c || !stream) return fail(RCULL_ERR_INVALID, "rcull_context_stream: null argument");
    *stream = c->own_stream;
    return RCULL_OK;
}

int rcull_entity_ptrs(rcull_ctx* c, float** pos_scale, float** rot, uint32_t** mesh) {
    if (!c) return fail(RCULL_ERR_INVALID, "rcull_entity_ptrs: null context");
    if (pos_scale) *pos_scale = reinterpret_cast<float*>(c->pos_scale);
    if (rot) *rot = reinterpret_cast<float*>(c->rot);
    if (mesh) *mesh = c->mesh;
    return RCULL_OK;
}

int rcull_pack_components(uint64_t n, const float* position, const float* rotation, const float* scale,
                          float* pos_scale, float* rot) {
    if (n && (!position || !rotation || !scale || !pos_scale || !rot))
        return fail(RCULL_ERR_INVALID, "rcull_pack_components: null argument");
    for (uint64_t i = 0; i < n; ++i) {
        pos_scale[4 * i + 0] = position[3 * i + 0];
        pos_scale[4 * i + 1] = position[3 * i + 1];
        pos_scale[4 * i + 2] = position[3 * i + 2];
        pos_scale[4 * i + 3] = scale[i];
    }
    memcpy(rot, rotation, n * 16);
    return RCULL_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------
// frame
// ---------------------------------------------------------------------------------
namespace {
int merge_frame_issued(rcull_ctx* c, uint32_t n_views, const uint32_t* counts, const unsigned long long* totals, RecordArgs* out, bool empty,
                       cudaStream_t s);
int merge_frame_launched(rcull_ctx* c, cudaStream_t s);
}
extern "C" {

int rcull_frame(rcull_ctx* c, uint64_t n, const rcull_view* views, uint32_t n_views, const float cam_pos[3],
                uint32_t draw_index_base, uint32_t buffer_ix, void* stream) {
    Scope scope("rcull::frame (model_matrix_calculation + aabb_calculation + coarse_culling + cull_pass)");
    if (!c || !views || !cam_pos) return fail(RCULL_ERR_INVALID, "rcull_frame: null argument");
    if (n > c->cfg.capacity_entities) return fail(RCULL_ERR_INVALID, "n %llu exceeds capacity_entities %llu", (unsigned long long)n, (unsigned long long)c->cfg.capacity_entities);
    if (n_views < 1 || n_views > c->cfg.max_views) return fail(RCULL_ERR_INVALID, "n_views %u not in 1..%u", n_views, c->cfg.max_views);
    if (buffer_ix >= c->cfg.n_buffers) return fail(RCULL_ERR_INVALID, "buffer_ix %u of %u", buffer_ix, c->cfg.n_buffers);
    if (c->n_meshes == 0) return fail(RCULL_ERR_INVALID, "rcull_set_mesh_table has not been called");
    const bool per_mesh = c->cfg.output_mode == RCULL_PER_MESH_INSTANCED;
    CU(cudaSetDevice(c->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    OutputSet& o = c->out[buffer_ix];
    c->lists_ix = buffer_ix;
    c->set_frozen[buffer_ix] = false;
    if (n == 0) {  // empty query: cull_pass still clears the buffers (cull_pipeline.rs:498-511)
        CU(cudaMemsetAsync(o.counts, 0, kMaxViews * 4, s));
        CU(cudaMemsetAsync(o.totals, 0, kMaxViews * 8, s));
        if (c->merge) return merge_frame_issued(c, n_views, o.counts, o.totals, nullptr, true, s);
        if (per_mesh) CU(cudaMemsetAsync(c->slot_total, 0, (size_t)c->n_meshes * kMaxLods * c->cfg.max_views * 4, s));  // rcull_get_slot_counts
        return RCULL_OK;
    }
    const uint32_t n_tiles = (uint32_t)((n + kTile - 1) / kTile);
    if (++c->epoch >= (1u << 30)) {  // epoch tag would wrap: start over from clean status words
        CU(cudaMemsetAsync(c->status, 0, c->status_bytes, s));
        c->epoch = 1;
    }
    FrameArgs a;
    memset(&a, 0, sizeof a);
    a.pos_scale = c->pos_scale;
    a.rot = c->rot;
    a.mesh = c->mesh;
    a.mesh_aabb = c->mesh_aabb;
    a.lod_tab = c->lod_tab;
    a.mesh_len1 = c->mesh_len1;
    a.aabb = o.aabb;
    for (uint32_t v = 0; v < n_views; ++v) {
        a.flags[v] = o.flags[v];
        a.cmds[v] = per_mesh ? o.keys[v] : o.cmds[v];
        a.visible[v] = per_mesh ? o.vis_tmp[v] : o.visible[v];
        for (int p = 0; p < 6; ++p) {
            const float* pl = views[v].planes[p];
            PlaneC& pc = a.planes[v][p];
            pc.ax = fabsf(pl[0]);
            pc.ay = fabsf(pl[1]);
            pc.az = fabsf(pl[2]);
            pc.px = pl[0];
            pc.py = pl[1];
            pc.pz = pl[2];
            pc.pw = pl[3];
        }
    }
    // packed-sum plane tests multiply with .ftz (rcull_frame.cuh): allowed while no plane factor is a tiny nonzero number
    a.planes_ftz_safe = 1u;
    for (uint32_t v = 0; v < n_views; ++v)
        for (int p = 0; p < 6; ++p)
            for (int x = 0; x < 3; ++x) {
                const float f = fabsf(views[v].planes[p][x]);
                if (!(f == 0.f || f >= kFtzSafe)) a.planes_ftz_safe = 0u;  // tiny, or NaN
            }
    if (getenv("RCULL_NO_PACKED_SUMS")) a.planes_ftz_safe = 0u;  // A/B switch: the scalar-sum body for every tile
    a.src_slot = o.src_slot;
    a.counts = o.counts;
    a.totals = o.totals;
    if (c->merge) {
        int mrc = merge_frame_issued(c, n_views, o.counts, o.totals, &a.rec, false, s);
        if (mrc) return mrc;
    }
    a.max_views = c->cfg.max_views;
    a.status = c->status;
    a.ticket = c->ticket;
    a.error_flag = c->ticket + 1;
    a.ticket_base = c->ticket_base;
    a.epoch = c->epoch;
    a.n = (uint32_t)n;
    a.n_tiles = n_tiles;
    a.n_meshes = c->n_meshes;
    a.len_small = c->len_small;
    a.per_mesh = per_mesh ? 1u : 0u;
    a.draw_index_base = draw_index_base;
    a.cap_visible = (uint32_t)(c->cap_vis > 0xffffffffull ? 0xffffffffull : c->cap_vis);
    a.cam[0] = cam_pos[0];
    a.cam[1] = cam_pos[1];
    a.cam[2] = cam_pos[2];
    a.lod_d2_max = c->lod_d2_max;
    const bool mat = c->cfg.materialize_aabb || c->cfg.materialize_flags || !c->len_small;  // the general variant
    int grid = (c->sm_count - c->reserved_sms) * c->blocks_per_sm[n_views][mat];
    if ((uint32_t)grid > n_tiles) grid = (int)n_tiles;
    cudaError_t e = cudaSuccess;
    switch (n_views) {
        case 1: e = launch<1>(mat, a, o.model_map, grid, s); break;
        case 2: e = launch<2>(mat, a, o.model_map, grid, s); break;
        case 3: e = launch<3>(mat, a, o.model_map, grid, s); break;
        case 4: e = launch<4>(mat, a, o.model_map, grid, s); break;
        case 5: e = launch<5>(mat, a, o.model_map, grid, s); break;
        case 6: e = launch<6>(mat, a, o.model_map, grid, s); break;
        case 7: e = launch<7>(mat, a, o.model_map, grid, s); break;
        case 8: e = launch<8>(mat, a, o.model_map, grid, s); break;
    }
    if (e != cudaSuccess) return fail(RCULL_ERR_CUDA, "transform_cull_compact launch failed: %s", cudaGetErrorString(e));
    c->ticket_base += n_tiles + (uint32_t)grid;  // every tile is a ticket, and every CTA draws one ticket past the end
    c->launches += 1;
    if (c->merge) {
        int mrc = merge_frame_launched(c, s);
        if (mrc) return mrc;
    }
    if (per_mesh && !c->merge) {  // with a merge attached the pairs leave as records and are bucketed after the expand (merged order)
        // Stable bucketing by (mesh, lod) slot: counting sort over tiles of the survivor pairs
        // (rcull_aux.cuh), then one command per non-empty slot.
        const uint32_t n_slots = c->n_meshes * kMaxLods;
        const uint64_t n_pairs = n < c->cap_vis ? n : c->cap_vis;
        BucketArgs b;
        memset(&b, 0, sizeof b);
        for (uint32_t v = 0; v < n_views; ++v) {
            b.keys[v] = o.keys[v];
            b.draw_ix[v] = o.vis_tmp[v];
            b.visible[v] = o.visible[v];
            b.cmds[v] = o.cmds[v];
        }
        b.counts = o.counts;
        b.totals = o.totals;
        b.hist = c->bucket_hist;
        b.slot_total = c->slot_total;
        b.slot_base = c->slot_base;
        b.lod_tab = c->lod_tab;
        b.n_slots = n_slots;
        b.cap = (uint32_t)n_pairs;
        const uint64_t rows_max = c->bucket_hist_words / ((uint64_t)n_slots * n_views);
        b.rows_max = (uint32_t)(rows_max < 0xffffffffull ? rows_max : 0xffffffffull);
        uint32_t sweep_ctas = (uint32_t)c->sm_count * 7 / n_views;  // 32 KB of counters per CTA: seven CTAs per SM
        if (sweep_ctas == 0) sweep_ctas = 1;
        b.rows_target = sweep_ctas * kBucketWarps;
        const dim3 sweep_grid(sweep_ctas, n_views);
        b.colscan_done = c->bucket_done;
        CU(launch_bucketing(b, sweep_grid, n_slots, n_views, s));
        c->launches += 3;
    }
    return RCULL_OK;
}

static int check_device_error(rcull_ctx* c) {
    uint32_t flag = 0;
    CU(cudaMemcpy(&flag, c->ticket + 1, 4, cudaMemcpyDeviceToHost));
    if (flag) {
        cudaMemset(c->ticket + 1, 0, 4);
        if (flag & 1u)
            return fail(RCULL_ERR_INVALID, "mesh id out of range of the mesh table (Vertex/Index buffer not consolidated, cull_pipeline.rs:543)");
        if (flag & 2u) return fail(RCULL_ERR_INVALID, "triangle stage: index or vertex outside the consolidated buffers");
        return fail(RCULL_ERR_INVALID, "triangle stage: culled index buffer too small (capacity_tri_indices)");
    }
    return RCULL_OK;
}

static int frame_host_impl(rcull_ctx* c, uint64_t n, const float* pos_scale, const float* rot, const uint32_t* mesh,
                           const rcull_view* views, uint32_t n_views, const float cam_pos[3], uint32_t draw_index_base,
                           uint32_t buffer_ix, uint32_t* counts_out, uint64_t* index_totals_out,
                           rcull_draw_cmd* const* cmds_out, uint32_t* const* visible_out, uint64_t capacity) {
    if (!c) return fail(RCULL_ERR_INVALID, "rcull_frame_host: null context");
    int rc = rcull_upload_entities(c, 0, n, pos_scale, rot, mesh, c->own_stream);
    if (rc) return rc;
    rc = rcull_frame(c, n, views, n_views, cam_pos, draw_index_base, buffer_ix, c->own_stream);
    if (rc) return rc;
    OutputSet& o = c->out[buffer_ix];
    CU(cudaMemcpyAsync(c->h_counts, o.counts, kMaxViews * 4, cudaMemcpyDeviceToHost, c->own_stream));
    CU(cudaMemcpyAsync(c->h_totals, o.totals, kMaxViews * 8, cudaMemcpyDeviceToHost, c->own_stream));
    CU(cudaMemcpyAsync(c->h_error, c->ticket + 1, 4, cudaMemcpyDeviceToHost, c->own_stream));
    CU(cudaStreamSynchronize(c->own_stream));
    if (*c->h_error) {  // the reference panics here ("Vertex buffer not consolidated", cull_pipeline.rs:543)
        CU(cudaMemsetAsync(c->ticket + 1, 0, 4, c->own_stream));
        CU(cudaStreamSynchronize(c->own_stream));
        return fail(RCULL_ERR_INVALID, "mesh id out of range of the mesh table (Vertex/Index buffer not consolidated, cull_pipeline.rs:543)");
    }
    for (uint32_t v = 0; v < n_views; ++v) {
        if (counts_out) counts_out[v] = c->h_counts[v];
        if (index_totals_out) index_totals_out[v] = c->h_totals[v];
    }
    if (cmds_out || visible_out) {  // the lists themselves: sizes are known now
        const bool per_mesh = c->cfg.output_mode == RCULL_PER_MESH_INSTANCED;
        for (uint32_t v = 0; v < n_views; ++v) {
            // per-mesh mode: counts = commands, index totals = instances in the bucketed visible list
            uint64_t n_cmd = c->h_counts[v], n_vis = per_mesh ? c->h_totals[v] : c->h_counts[v];
            if (n_cmd > c->cap_cmds) n_cmd = c->cap_cmds;
            if (n_vis > c->cap_vis) n_vis = c->cap_vis;
            if (n_cmd > capacity || n_vis > capacity)
                return fail(RCULL_ERR_INVALID, "view %u: %llu commands / %llu visible entries exceed the host capacity %llu", v,
                            (unsigned long long)n_cmd, (unsigned long long)n_vis, (unsigned long long)capacity);
            if (cmds_out && cmds_out[v] && n_cmd) CU(cudaMemcpyAsync(cmds_out[v], o.cmds[v], n_cmd * 20, cudaMemcpyDeviceToHost, c->own_stream));
            if (visible_out && visible_out[v] && n_vis) CU(cudaMemcpyAsync(visible_out[v], o.visible[v], n_vis * 4, cudaMemcpyDeviceToHost, c->own_stream));
        }
        CU(cudaStreamSynchronize(c->own_stream));
    }
    return RCULL_OK;
}

int rcull_frame_host(rcull_ctx* c, uint64_t n, const float* pos_scale, const float* rot, const uint32_t* mesh,
                     const rcull_view* views, uint32_t n_views, const float cam_pos[3], uint32_t draw_index_base,
                     uint32_t buffer_ix, uint32_t* counts_out, uint64_t* index_totals_out) {
    Scope scope("rcull::frame_host (upload + frame + counts)");
    return frame_host_impl(c, n, pos_scale, rot, mesh, views, n_views, cam_pos, draw_index_base, buffer_ix, counts_out,
                           index_totals_out, nullptr, nullptr, 0);
}

int rcull_frame_host_lists(rcull_ctx* c, uint64_t n, const float* pos_scale, const float* rot, const uint32_t* mesh,
                           const rcull_view* views, uint32_t n_views, const float cam_pos[3], uint32_t draw_index_base,
                           uint32_t buffer_ix, uint32_t* counts_out, uint64_t* index_totals_out,
                           rcull_draw_cmd* const* cmds_out, uint32_t* const* visible_out, uint64_t capacity) {
    Scope scope("rcull::frame_host_lists (upload + frame + counts + draw lists back)");
    if (!cmds_out && !visible_out) return fail(RCULL_ERR_INVALID, "rcull_frame_host_lists: no list destination");
    return frame_host_impl(c, n, pos_scale, rot, mesh, views, n_views, cam_pos, draw_index_base, buffer_ix, counts_out,
                           index_totals_out, cmds_out, visible_out, capacity);
}

int rcull_get_outputs(rcull_ctx* c, uint32_t buffer_ix, rcull_outputs* out) {
    if (!c || !out) return fail(RCULL_ERR_INVALID, "rcull_get_outputs: null argument");
    if (buffer_ix >= c->cfg.n_buffers) return fail(RCULL_ERR_INVALID, "buffer_ix %u of %u", buffer_ix, c->cfg.n_buffers);
    const OutputSet& o = c->out[buffer_ix];
    const OutputSet& l = c->out[lists_of(c, buffer_ix)];  // frozen set: the lists of the last unfrozen frame
    memset(out, 0, sizeof *out);
    out->model = reinterpret_cast<float*>(o.model);
    out->aabb = o.aabb;
    for (int v = 0; v < kMaxViews; ++v) {
        out->flags[v] = o.flags[v];
        out->cmds[v] = reinterpret_cast<rcull_draw_cmd*>(l.cmds[v]);
        out->visible[v] = l.visible[v];
    }
    out->counts = l.counts;
    out->index_totals = reinterpret_cast<uint64_t*>(l.totals);
    return RCULL_OK;
}

#define DL_PROLOGUE(what)                                                                         \
    if (!c || !dst) return fail(RCULL_ERR_INVALID, what ": null argument");                       \
    if (buffer_ix >= c->cfg.n_buffers) return fail(RCULL_ERR_INVALID, what ": buffer_ix %u of %u", buffer_ix, c->cfg.n_buffers); \
    CU(cudaSetDevice(c->cfg.device));                                                             \
    CU(cudaDeviceSynchronize());

int rcull_download_counts(rcull_ctx* c, uint32_t buffer_ix, uint32_t* dst, uint64_t* index_totals) {
    DL_PROLOGUE("rcull_download_counts");
    int rc = check_device_error(c);
    if (rc) return rc;
    const OutputSet& l = c->out[lists_of(c, buffer_ix)];
    CU(cudaMemcpy(dst, l.counts, c->cfg.max_views * 4, cudaMemcpyDeviceToHost));
    if (index_totals) CU(cudaMemcpy(index_totals, l.totals, c->cfg.max_views * 8, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}
int rcull_download_model(rcull_ctx* c, uint32_t buffer_ix, uint64_t first, uint64_t n, float* dst) {
    DL_PROLOGUE("rcull_download_model");
    if (first + n > c->cap) return fail(RCULL_ERR_INVALID, "model range out of bounds");
    CU(cudaMemcpy(dst, c->out[buffer_ix].model + first * 4, n * 64, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}
int rcull_download_aabb(rcull_ctx* c, uint32_t buffer_ix, uint64_t first, uint64_t n, float* dst) {
    DL_PROLOGUE("rcull_download_aabb");
    if (!c->out[buffer_ix].aabb) return fail(RCULL_ERR_INVALID, "context created without materialize_aabb");
    if (first + n > c->cap) return fail(RCULL_ERR_INVALID, "aabb range out of bounds");
    CU(cudaMemcpy(dst, c->out[buffer_ix].aabb + first * 6, n * 24, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}
int rcull_download_flags(rcull_ctx* c, uint32_t buffer_ix, uint32_t view, uint64_t first, uint64_t n, uint8_t* dst) {
    DL_PROLOGUE("rcull_download_flags");
    if (view >= c->cfg.max_views || !c->out[buffer_ix].flags[view]) return fail(RCULL_ERR_INVALID, "flags not materialised for view %u", view);
    if (first + n > c->cap) return fail(RCULL_ERR_INVALID, "flags range out of bounds");
    CU(cudaMemcpy(dst, c->out[buffer_ix].flags[view] + first, n, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}
int rcull_download_cmds(rcull_ctx* c, uint32_t buffer_ix, uint32_t view, uint64_t first, uint64_t n, rcull_draw_cmd* dst) {
    DL_PROLOGUE("rcull_download_cmds");
    if (view >= c->cfg.max_views) return fail(RCULL_ERR_INVALID, "view %u of %u", view, c->cfg.max_views);
    if (first + n > c->cap_cmds) return fail(RCULL_ERR_INVALID, "command range out of bounds");
    CU(cudaMemcpy(dst, c->out[lists_of(c, buffer_ix)].cmds[view] + first * 5, n * 20, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}
int rcull_download_visible(rcull_ctx* c, uint32_t buffer_ix, uint32_t view, uint64_t first, uint64_t n, uint32_t* dst) {
    DL_PROLOGUE("rcull_download_visible");
    if (view >= c->cfg.max_views) return fail(RCULL_ERR_INVALID, "view %u of %u", view, c->cfg.max_views);
    if (first + n > c->cap_vis) return fail(RCULL_ERR_INVALID, "visible range out of bounds");
    CU(cudaMemcpy(dst, c->out[lists_of(c, buffer_ix)].visible[view] + first, n * 4, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}

int rcull_rebase_into(rcull_ctx* c, uint32_t buffer_ix, uint32_t view, uint64_t index_prefix, rcull_draw_cmd* dst_cmds,
                      uint32_t* dst_visible, void* stream) {
    Scope scope("rcull::rebase_into");
    if (!c || !dst_cmds || !dst_visible) return fail(RCULL_ERR_INVALID, "rcull_rebase_into: null argument");
    if (c->merge) return fail(RCULL_ERR_INVALID, "rcull_rebase_into: a merge is attached -- the frame kernel sends records, the rank's own lists are not written");
    if (buffer_ix >= c->cfg.n_buffers || view >= c->cfg.max_views) return fail(RCULL_ERR_INVALID, "rcull_rebase_into: bad buffer/view");
    CU(cudaSetDevice(c->cfg.device));
    const OutputSet& o = c->out[lists_of(c, buffer_ix)];
    rebase_into<<<c->sm_count * 4, 256, 0, (cudaStream_t)stream>>>(
        o.cmds[view], o.visible[view], o.counts + view, (uint32_t)(c->cap_vis > 0xffffffffull ? 0xffffffffull : c->cap_vis),
        (uint32_t)index_prefix, reinterpret_cast<uint32_t*>(dst_cmds), dst_visible);
    CU(cudaGetLastError());
    c->launches += 1;
    return RCULL_OK;
}

int rcull_synth_entities(rcull_ctx* c, uint64_t seed, uint64_t first_global, uint64_t n, float half_box,
                         uint32_t n_meshes, void* stream) {
    Scope scope("rcull::synth_entities");
    if (!c) return fail(RCULL_ERR_INVALID, "rcull_synth_entities: null context");
    if (n > c->cfg.capacity_entities) return fail(RCULL_ERR_INVALID, "n exceeds capacity_entities");
    if (n == 0) return RCULL_OK;
    CU(cudaSetDevice(c->cfg.device));
    synth_entities<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, first_global, (uint32_t)n, half_box,
                                                                             n_meshes, c->pos_scale, c->rot, c->mesh);
    CU(cudaGetLastError());
    c->launches += 1;
    return RCULL_OK;
}

uint64_t rcull_launch_count(rcull_ctx* c) { return c ? c->launches : 0; }

int rcull_frame_frozen(rcull_ctx* c, uint64_t n, uint32_t buffer_ix, void* stream) {
    Scope scope("rcull::frame_frozen (cull_pass_bypass)");
    if (!c) return fail(RCULL_ERR_INVALID, "rcull_frame_frozen: null context");
    if (n > c->cfg.capacity_entities) return fail(RCULL_ERR_INVALID, "n = %llu exceeds capacity_entities", (unsigned long long)n);
    if (buffer_ix >= c->cfg.n_buffers) return fail(RCULL_ERR_INVALID, "buffer_ix %u of %u", buffer_ix, c->cfg.n_buffers);
    c->set_frozen[buffer_ix] = buffer_ix != c->lists_ix;  // the lists of this set are now those of set lists_ix
    if (n == 0) return RCULL_OK;
    CU(cudaSetDevice(c->cfg.device));
    model_only<<<c->sm_count * 8, kRecordWarps * 32, 0, (cudaStream_t)stream>>>(c->pos_scale, c->rot, (uint32_t)n, c->out[buffer_ix].model_map);
    CU(cudaGetLastError());
    c->launches += 1;
    return RCULL_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------
// rcull_frame_noalloc: the frame kernel without a context and without an allocation
// (no counterpart in the reference)
// ---------------------------------------------------------------------------------
// The scratch of the frame kernel in a file-scope device array: every GPU that loads librcull holds these
// bytes until the process exits.  In HOST code this name is only a stand-in for the symbol -- it is not a
// device address; the address comes from cudaGetSymbolAddress, once per device (noalloc_device below).
static __device__ __align__(256) unsigned char g_noalloc_scratch[RCULL_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr uint32_t kNoallocMaxTiles = (RCULL_NOALLOC_MAX_ENTITIES + kTile - 1) / kTile;
// layout of g_noalloc_scratch (byte offsets; every part 256-byte aligned)
constexpr size_t kNaTicket = 0;                                                          // tile ticket word
constexpr size_t kNaMeshAabb = 256;                                                      // float4 [meshes][2]
constexpr size_t kNaLodTab = kNaMeshAabb + RCULL_NOALLOC_MAX_MESHES * 2 * sizeof(float4);  // uint4 [meshes][kMaxLods]
constexpr size_t kNaLen1 = kNaLodTab + RCULL_NOALLOC_MAX_MESHES * kMaxLods * sizeof(uint4);  // uint32 [meshes]
constexpr size_t kNaStatus = kNaLen1 + RCULL_NOALLOC_MAX_MESHES * sizeof(uint32_t);      // ulonglong2 [views][tiles]
constexpr size_t kNaEnd = kNaStatus + (size_t)kMaxViews * kNoallocMaxTiles * sizeof(ulonglong2);
static_assert(RCULL_NOALLOC_TILE == kTile, "rcull.h states the tile the no-alloc inputs are padded to");
static_assert(kNaLodTab % 256 == 0 && kNaLen1 % 256 == 0 && kNaStatus % 256 == 0, "scratch parts stay aligned");
static_assert(kNaEnd <= RCULL_NOALLOC_SCRATCH_BYTES, "scratch layout exceeds the stated maximum");

struct NoallocDevice {
    bool ready = false;
    unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
    int sm_count = 0;
    int blocks_per_sm[kMaxViews + 1][2] = {};
};
constexpr int kNoallocMaxDevices = 64;
NoallocDevice g_noalloc_dev[kNoallocMaxDevices];
std::mutex g_noalloc_mu;

template <int V, bool M>
int noalloc_configure(NoallocDevice& d) {
    auto k = transform_cull_compact<V, M>;
    CU(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)frame_smem_bytes<V>()));
    CU(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    int b = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, k, kThreads, frame_smem_bytes<V>()));
    if (b < 1) return fail(RCULL_ERR_CUDA, "transform_cull_compact<%d,%d> does not fit on an SM", V, (int)M);
    d.blocks_per_sm[V][M] = b;
    return 0;
}

// Set-up, once per device and with that device current: the device address of the scratch array and the
// launch attributes of the frame kernel.  Nothing here allocates device memory.
int noalloc_device(int device, NoallocDevice** out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    NoallocDevice& d = g_noalloc_dev[device];
    if (!d.ready) {
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, device));
        if (prop.major != 10)
            return fail(RCULL_ERR_NO_DEVICE, "device %d is sm_%d%d; librcull is built for sm_100a only", device, prop.major, prop.minor);
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(RCULL_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d failed: %s", device, cudaGetErrorString(e));
        d.scratch = static_cast<unsigned char*>(p);
        d.sm_count = prop.multiProcessorCount;
        int rc = 0;
#define NA_TRY(x) \
    if ((rc = (x)) != 0) return rc;
        NA_TRY((noalloc_configure<1, false>(d)));
        NA_TRY((noalloc_configure<1, true>(d)));
        NA_TRY((noalloc_configure<2, false>(d)));
        NA_TRY((noalloc_configure<2, true>(d)));
        NA_TRY((noalloc_configure<3, false>(d)));
        NA_TRY((noalloc_configure<3, true>(d)));
        NA_TRY((noalloc_configure<4, false>(d)));
        NA_TRY((noalloc_configure<4, true>(d)));
        NA_TRY((noalloc_configure<5, false>(d)));
        NA_TRY((noalloc_configure<5, true>(d)));
        NA_TRY((noalloc_configure<6, false>(d)));
        NA_TRY((noalloc_configure<6, true>(d)));
        NA_TRY((noalloc_configure<7, false>(d)));
        NA_TRY((noalloc_configure<7, true>(d)));
        NA_TRY((noalloc_configure<8, false>(d)));
        NA_TRY((noalloc_configure<8, true>(d)));
#undef NA_TRY
        d.ready = true;
    }
    *out = &d;
    return 0;
}

// The caller's model buffer as [ceil(n / 2) rows][32 floats]: rows past the end are clipped by the TMA
// store, so the buffer needs no padding to whole tiles (make_model_map sizes a context's own buffer).
int noalloc_model_map(CUtensorMap* map, float* model, uint64_t n) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || !fn) return fail(RCULL_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
    const cuuint64_t dims[2] = {32, (n + 1) / 2};
    const cuuint64_t strides[1] = {128};
    const cuuint32_t box[2] = {32, 16};
    const cuuint32_t elem[2] = {1, 1};
    CUresult r = reinterpret_cast<EncodeTiledFn>(fn)(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, model, dims, strides, box, elem,
                                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                                     CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(RCULL_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

extern "C" {

int rcull_frame_noalloc(int32_t device, uint64_t n, const rcull_noalloc_io* io, const rcull_mesh* meshes, uint32_t n_meshes,
                        const rcull_view* views, uint32_t n_views, const float cam_pos[3], float lod_distance,
                        uint32_t draw_index_base, void* stream) {
    Scope scope("rcull::frame_noalloc");
    if (!io || !meshes || !views || !cam_pos) return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: null argument");
    if (n > RCULL_NOALLOC_MAX_ENTITIES)
        return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: n = %llu exceeds RCULL_NOALLOC_MAX_ENTITIES = %u (the static scratch is sized for it; use rcull_create + rcull_frame)",
                    (unsigned long long)n, RCULL_NOALLOC_MAX_ENTITIES);
    if (n_meshes < 1 || n_meshes > RCULL_NOALLOC_MAX_MESHES)
        return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: n_meshes = %u not in 1..RCULL_NOALLOC_MAX_MESHES = %u", n_meshes, RCULL_NOALLOC_MAX_MESHES);
    if (n_views < 1 || n_views > RCULL_MAX_VIEWS) return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: n_views %u not in 1..%d", n_views, RCULL_MAX_VIEWS);
    if (!io->counts || !io->index_totals || !io->error) return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: counts / index_totals / error are required");
    if (n && (!io->pos_scale || !io->rot || !io->mesh || !io->model)) return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: null entity column or model buffer");
    if (!aligned16(io->pos_scale) || !aligned16(io->rot) || !aligned16(io->mesh) || !aligned16(io->model))
        return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: pos_scale, rot, mesh and model must be 16-byte aligned");
    if (n && io->capacity_visible == 0) return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: capacity_visible is 0");
    for (uint32_t v = 0; n && v < n_views; ++v)
        if (!io->cmds[v] || !io->visible[v]) return fail(RCULL_ERR_INVALID, "rcull_frame_noalloc: view %u has no cmds / visible buffer", v);
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) return fail(RCULL_ERR_NO_DEVICE, "no CUDA device: librcull has no CPU path");
    if (device < 0 || device >= n_dev || device >= kNoallocMaxDevices) return fail(RCULL_ERR_INVALID, "device %d of %d", device, n_dev);
    CU(cudaSetDevice(device));
    NoallocDevice* d = nullptr;
    int rc = noalloc_device(device, &d);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;

    // mesh table in the kernel's form (rcull_set_mesh_table), built on the stack
    float4 aabb[2 * RCULL_NOALLOC_MAX_MESHES];
    uint4 lods[kMaxLods * RCULL_NOALLOC_MAX_MESHES];
    uint32_t len1[RCULL_NOALLOC_MAX_MESHES];
    uint32_t max_len = 0;
    for (uint32_t i = 0; i < n_meshes; ++i) {
        const rcull_mesh& m = meshes[i];
        if (m.n_lods < 1 || m.n_lods > kMaxLods) return fail(RCULL_ERR_INVALID, "mesh %u: n_lods %u not in 1..%d", i, m.n_lods, kMaxLods);
        float nl, l0;
        memcpy(&nl, &m.n_lods, 4);
        memcpy(&l0, &m.lod[0].index_len, 4);
        aabb[2 * i] = make_float4(m.aabb_min[0], m.aabb_min[1], m.aabb_min[2], nl);
        aabb[2 * i + 1] = make_float4(m.aabb_max[0], m.aabb_max[1], m.aabb_max[2], l0);
        len1[i] = m.n_lods > 1 ? m.lod[1].index_len : m.lod[0].index_len;
        for (uint32_t l = 0; l < (uint32_t)kMaxLods; ++l) {
            if (l < m.n_lods) {
                if (m.lod[l].index_len > max_len) max_len = m.lod[l].index_len;
                lods[(size_t)i * kMaxLods + l] = make_uint4(m.lod[l].index_len, m.lod[l].first_index, (uint32_t)m.vertex_offset, 0);
            } else {
                lods[(size_t)i * kMaxLods + l] = make_uint4(0, 0, 0, 0);
            }
        }
    }
    const uint32_t len_small = max_len < (1u << 25) ? 1u : 0u;

    // what must start at zero, on the caller's stream: outputs, then the scan state this frame will use
    CU(cudaMemsetAsync(io->counts, 0, kMaxViews * 4, s));
    CU(cudaMemsetAsync(io->index_totals, 0, kMaxViews * 8, s));
    CU(cudaMemsetAsync(io->error, 0, 4, s));
    if (n == 0) return RCULL_OK;  // empty query: cull_pass still clears the buffers (cull_pipeline.rs:498-511)
    const uint32_t n_tiles = (uint32_t)((n + kTile - 1) / kTile);
    CU(cudaMemsetAsync(d->scratch + kNaTicket, 0, 256, s));
    CU(cudaMemsetAsync(d->scratch + kNaStatus, 0, (size_t)n_views * n_tiles * sizeof(ulonglong2), s));
    // pageable (stack) sources: the runtime stages them before these calls return
    CU(cudaMemcpyAsync(d->scratch + kNaMeshAabb, aabb, (size_t)n_meshes * 2 * sizeof(float4), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d->scratch + kNaLodTab, lods, (size_t)n_meshes * kMaxLods * sizeof(uint4), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d->scratch + kNaLen1, len1, (size_t)n_meshes * 4, cudaMemcpyHostToDevice, s));

    CUtensorMap map;
    rc = noalloc_model_map(&map, io->model, n);
    if (rc) return rc;
    FrameArgs a;
    memset(&a, 0, sizeof a);
    a.pos_scale = reinterpret_cast<const float4*>(io->pos_scale);
    a.rot = reinterpret_cast<const float4*>(io->rot);
    a.mesh = io->mesh;
    a.mesh_aabb = reinterpret_cast<const float4*>(d->scratch + kNaMeshAabb);
    a.lod_tab = reinterpret_cast<const uint4*>(d->scratch + kNaLodTab);
    a.mesh_len1 = reinterpret_cast<const uint32_t*>(d->scratch + kNaLen1);
    a.planes_ftz_safe = 1u;
    for (uint32_t v = 0; v < n_views; ++v) {
        a.cmds[v] = reinterpret_cast<uint32_t*>(io->cmds[v]);
        a.visible[v] = io->visible[v];
        for (int p = 0; p < 6; ++p) {
            const float* pl = views[v].planes[p];
            PlaneC& pc = a.planes[v][p];
            pc.ax = fabsf(pl[0]);
            pc.ay = fabsf(pl[1]);
            pc.az = fabsf(pl[2]);
            pc.px = pl[0];
            pc.py = pl[1];
            pc.pz = pl[2];
            pc.pw = pl[3];
            for (int x = 0; x < 3; ++x) {
                const float f = fabsf(pl[x]);
                if (!(f == 0.f || f >= kFtzSafe)) a.planes_ftz_safe = 0u;  // tiny, or NaN: the scalar-sum body (as rcull_frame)
            }
        }
    }
    a.counts = io->counts;
    a.totals = reinterpret_cast<unsigned long long*>(io->index_totals);
    a.max_views = n_views;
    a.status = reinterpret_cast<ulonglong2*>(d->scratch + kNaStatus);
    a.ticket = reinterpret_cast<uint32_t*>(d->scratch + kNaTicket);
    a.error_flag = io->error;
    a.ticket_base = 0;  // ticket and descriptors were cleared above: first epoch, first ticket
    a.epoch = 1;
    a.n = (uint32_t)n;
    a.n_tiles = n_tiles;
    a.n_meshes = n_meshes;
    a.len_small = len_small;
    a.draw_index_base = draw_index_base;
    a.cap_visible = (uint32_t)(io->capacity_visible > 0xffffffffull ? 0xffffffffull : io->capacity_visible);
    a.cam[0] = cam_pos[0];
    a.cam[1] = cam_pos[1];
    a.cam[2] = cam_pos[2];
    a.lod_d2_max = lod_threshold(lod_distance);
    const bool mat = !len_small;
    int grid = d->sm_count * d->blocks_per_sm[n_views][mat];
    if ((uint32_t)grid > n_tiles) grid = (int)n_tiles;
    cudaError_t e = cudaSuccess;
    switch (n_views) {
        case 1: e = launch<1>(mat, a, map, grid, s); break;
        case 2: e = launch<2>(mat, a, map, grid, s); break;
        case 3: e = launch<3>(mat, a, map, grid, s); break;
        case 4: e = launch<4>(mat, a, map, grid, s); break;
        case 5: e = launch<5>(mat, a, map, grid, s); break;
        case 6: e = launch<6>(mat, a, map, grid, s); break;
        case 7: e = launch<7>(mat, a, map, grid, s); break;
        case 8: e = launch<8>(mat, a, map, grid, s); break;
    }
    if (e != cudaSuccess) return fail(RCULL_ERR_CUDA, "transform_cull_compact launch failed: %s", cudaGetErrorString(e));
    return RCULL_OK;
}

int rcull_get_slot_counts(rcull_ctx* c, uint32_t view, uint32_t** counts, uint32_t* n_slots) {
    if (!c || !counts || !n_slots) return fail(RCULL_ERR_INVALID, "rcull_get_slot_counts: null argument");
    if (c->cfg.output_mode != RCULL_PER_MESH_INSTANCED) return fail(RCULL_ERR_INVALID, "slot counts exist in per-mesh instanced mode only");
    if (view >= c->cfg.max_views) return fail(RCULL_ERR_INVALID, "view %u of %u", view, c->cfg.max_views);
    if (c->n_meshes == 0) return fail(RCULL_ERR_INVALID, "rcull_set_mesh_table has not been called");
    *n_slots = c->n_meshes * kMaxLods;
    *counts = c->slot_total + (size_t)view * *n_slots;  // rows are packed at the table's slot count (BucketArgs.n_slots)
    return RCULL_OK;
}

int rcull_download_slot_counts(rcull_ctx* c, uint32_t view, uint32_t* counts, uint32_t n_slots) {
    uint32_t* src = nullptr;
    uint32_t have = 0;
    int rc = rcull_get_slot_counts(c, view, &src, &have);
    if (rc) return rc;
    if (!counts || n_slots > have) return fail(RCULL_ERR_INVALID, "rcull_download_slot_counts: %u slots asked, %u exist", n_slots, have);
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(counts, src, (size_t)n_slots * 4, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}

int rcull_set_blas_table(rcull_ctx* c, const uint64_t* blas, uint32_t n) {
    if (!c || !blas) return fail(RCULL_ERR_INVALID, "rcull_set_blas_table: null argument");
    if (n == 0 || n > c->cfg.max_meshes) return fail(RCULL_ERR_INVALID, "blas table of %u entries, max_meshes %u", n, c->cfg.max_meshes);
    CU(cudaSetDevice(c->cfg.device));
    if (!c->blas) {
        int rc = dev_alloc(c, &c->blas, c->cfg.max_meshes);
        if (rc) return rc;
        rc = dev_alloc(c, &c->instances, c->cap * 4);
        if (rc) return rc;
        rc = make_model_map(&c->instances_map, reinterpret_cast<float4*>(c->instances), c->cap);  // same shape as a model buffer: 64-byte records
        if (rc) return rc;
    }
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(c->blas, blas, (size_t)n * 8, cudaMemcpyHostToDevice));
    c->n_blas = n;
    return RCULL_OK;
}

int rcull_as_instances(rcull_ctx* c, uint64_t n, uint32_t draw_index_base, uint32_t buffer_ix, void* stream) {
    Scope scope("rcull::as_instances (build_acceleration_structures instance loop)");
    if (!c) return fail(RCULL_ERR_INVALID, "rcull_as_instances: null context");
    if (!c->blas) return fail(RCULL_ERR_INVALID, "rcull_set_blas_table has not been called");
    // a mesh id past the BLAS table would pack an address nobody set (the reference indexes a Vec: it would panic)
    if (c->n_blas < c->n_meshes) return fail(RCULL_ERR_INVALID, "BLAS table has %u entries, the mesh table %u", c->n_blas, c->n_meshes);
    if (n > c->cfg.capacity_entities) return fail(RCULL_ERR_INVALID, "n exceeds capacity_entities");
    if (buffer_ix >= c->cfg.n_buffers) return fail(RCULL_ERR_INVALID, "buffer_ix %u of %u", buffer_ix, c->cfg.n_buffers);
    if (n == 0) return RCULL_OK;
    CU(cudaSetDevice(c->cfg.device));
    as_instances<<<c->sm_count * 8, kRecordWarps * 32, 0, (cudaStream_t)stream>>>(c->out[buffer_ix].model, c->mesh, c->blas, c->n_meshes, (uint32_t)n,
                                                                                  draw_index_base, c->instances_map);
    CU(cudaGetLastError());
    c->launches += 1;
    return RCULL_OK;
}

int rcull_get_as_instances(rcull_ctx* c, void** instances) {
    if (!c || !instances) return fail(RCULL_ERR_INVALID, "rcull_get_as_instances: null argument");
    if (!c->instances) return fail(RCULL_ERR_INVALID, "rcull_set_blas_table has not been called");
    *instances = c->instances;
    return RCULL_OK;
}

int rcull_download_as_instances(rcull_ctx* c, uint64_t first, uint64_t n, void* dst) {
    if (!c || !dst) return fail(RCULL_ERR_INVALID, "rcull_download_as_instances: null argument");
    if (!c->instances) return fail(RCULL_ERR_INVALID, "rcull_set_blas_table has not been called");
    if (first + n > c->cap) return fail(RCULL_ERR_INVALID, "instance range out of bounds");
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(dst, c->instances + first * 4, n * 64, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}

int rcull_set_geometry(rcull_ctx* c, const float* positions, uint64_t n_vertices, const uint32_t* indices, uint64_t n_indices) {
    if (!c || !positions || !indices) return fail(RCULL_ERR_INVALID, "rcull_set_geometry: null argument");
    if (!c->cfg.capacity_tri_indices) return fail(RCULL_ERR_INVALID, "context created without capacity_tri_indices");
    if (n_vertices == 0 || n_indices == 0 || n_indices > 0xffffffffull) return fail(RCULL_ERR_INVALID, "rcull_set_geometry: empty or oversized buffers");
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaDeviceSynchronize());
    if (c->positions) cudaFree(c->positions);
    if (c->indices) cudaFree(c->indices);
    c->positions = nullptr;
    c->indices = nullptr;
    CU(cudaMalloc((void**)&c->positions, n_vertices * 12 + 256));
    CU(cudaMalloc((void**)&c->indices, n_indices * 4 + 256));
    CU(cudaMemcpy(c->positions, positions, n_vertices * 12, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->indices, indices, n_indices * 4, cudaMemcpyHostToDevice));
    c->n_vertices = n_vertices;
    c->n_indices = n_indices;
    c->h_indices.assign(indices, indices + n_indices);
    c->chunks_dirty = true;
    return RCULL_OK;
}

// Chunk table of the triangle stage: every LOD's triangle list cut into runs of 256-triangle steps whose indices stay
// inside a window of at most kTriVertCap vertices (those runs get their window transformed once into shared memory);
// a step with a wider window -- random index triples -- becomes a chunk of its own on the gather path (v_count = 0).
static int build_tri_chunks(rcull_ctx* c) {
    const uint32_t n_slots = c->n_meshes * kMaxLods;
    std::vector<uint2> sc(n_slots, make_uint2(0, 0));
    std::vector<uint4> ch;
    bool any_window = false, any_gather = false;
    const uint32_t kStep = kTriThreads * kTriItems;
    for (uint32_t m = 0; m < c->n_meshes && m < c->h_meshes.size(); ++m) {
        const rcull_mesh& mesh = c->h_meshes[m];
        for (uint32_t l = 0; l < mesh.n_lods && l < (uint32_t)kMaxLods; ++l) {
            const uint64_t first = mesh.lod[l].first_index / 3u, n_tri = mesh.lod[l].index_len / 3u;
            sc[m * kMaxLods + l].x = (uint32_t)ch.size();
            if (3 * (first + n_tri) > c->h_indices.size()) continue;  // the kernel flags the command
            bool open = false, wide = false;  // wide: a run of steps that each reference more than kTriVertCap vertices (gather path)
            uint32_t c_begin = 0, c_count = 0, c_lo = 0, c_hi = 0;
            auto flush = [&]() {
                if (!open) return;
                const uint64_t window = (uint64_t)c_hi - c_lo + 1;
                // worth it when the window holds fewer vertices than the triangles have corners (3 per triangle), with some
                // slack for tiny LODs so that they do not drag a second kernel launch in
                const bool cached = !wide && window <= (uint64_t)kTriVertCap && (2 * window <= 5ull * c_count || window <= 512);
                ch.push_back(make_uint4(c_begin, c_count, cached ? c_lo : 0u, cached ? (uint32_t)window : 0u));
                open = false;
            };
            for (uint64_t s0 = 0; s0 < n_tri; s0 += kStep) {
                const uint64_t e = s0 + kStep < n_tri ? s0 + kStep : n_tri;
                uint32_t lo = 0xffffffffu, hi = 0;
                for (uint64_t i = 3 * (first + s0); i < 3 * (first + e); ++i) {
                    const uint32_t x = c->h_indices[i];
                    lo = x < lo ? x : lo;
                    hi = x > hi ? x : hi;
                }
                const bool step_wide = (uint64_t)hi - lo + 1 > (uint64_t)kTriVertCap;
                if (open && wide == step_wide) {
                    const uint32_t nlo = lo < c_lo ? lo : c_lo, nhi = hi > c_hi ? hi : c_hi;
                    if (wide || (uint64_t)nhi - nlo + 1 <= (uint64_t)kTriVertCap) {  // wide runs just grow: one chunk for the whole run
                        c_lo = nlo;
                        c_hi = nhi;
                        c_count += (uint32_t)(e - s0);
                        continue;
                    }
                }
                flush();
                open = true;
                wide = step_wide;
                c_begin = (uint32_t)s0;
                c_count = (uint32_t)(e - s0);
                c_lo = lo;
                c_hi = hi;
            }
            flush();
            sc[m * kMaxLods + l].y = (uint32_t)ch.size() - sc[m * kMaxLods + l].x;
            bool all_cached = sc[m * kMaxLods + l].y > 0;
            for (uint32_t i = sc[m * kMaxLods + l].x; i < ch.size(); ++i) all_cached = all_cached && ch[i].w != 0u;
            if (all_cached) sc[m * kMaxLods + l].y |= kSlotAllCached;
            if (l < 2) (all_cached ? any_window : any_gather) = true;  // pick_lod only ever returns lod 0 or 1 (helpers.rs:3-11)
        }
    }
    CU(cudaDeviceSynchronize());
    if (c->slot_chunks) cudaFree(c->slot_chunks);
    if (c->chunks) cudaFree(c->chunks);
    c->slot_chunks = nullptr;
    c->chunks = nullptr;
    CU(cudaMalloc((void**)&c->slot_chunks, sc.size() * sizeof(uint2) + 256));
    CU(cudaMalloc((void**)&c->chunks, ch.size() * sizeof(uint4) + 256));
    CU(cudaMemcpy(c->slot_chunks, sc.data(), sc.size() * sizeof(uint2), cudaMemcpyHostToDevice));
    if (!ch.empty()) CU(cudaMemcpy(c->chunks, ch.data(), ch.size() * sizeof(uint4), cudaMemcpyHostToDevice));
    c->n_chunks = (uint32_t)ch.size();
    c->any_window_slot = any_window;
    c->any_gather_slot = any_gather;
    c->chunks_dirty = false;
    return RCULL_OK;
}

int rcull_tri_cull(rcull_ctx* c, const float pv[16], uint32_t draw_index_base, uint32_t buffer_ix, void* stream) {
    Scope scope("rcull::tri_cull (generate_work + compact_draw_stream)");
    if (!c || !pv) return fail(RCULL_ERR_INVALID, "rcull_tri_cull: null argument");
    if (!c->cfg.capacity_tri_indices) return fail(RCULL_ERR_INVALID, "context created without capacity_tri_indices");
    if (!c->positions) return fail(RCULL_ERR_INVALID, "rcull_set_geometry has not been called");
    if (buffer_ix >= c->cfg.n_buffers) return fail(RCULL_ERR_INVALID, "buffer_ix %u of %u", buffer_ix, c->cfg.n_buffers);
    if (c->merge) return fail(RCULL_ERR_INVALID, "rcull_tri_cull: a merge is attached -- the rank's own lists are not written");
    CU(cudaSetDevice(c->cfg.device));
    if (c->chunks_dirty) {
        int rc = build_tri_chunks(c);
        if (rc) return rc;
    }
    cudaStream_t s = (cudaStream_t)stream;
    OutputSet& o = c->out[lists_of(c, buffer_ix)];  // a frozen set draws from the lists of the last unfrozen frame
    CU(cudaMemsetAsync(c->tri_status, 0, c->tri_status_bytes, s));  // descriptors + the two tickets behind them
    const size_t tiles = (c->cap_cmds + kCompactTile - 1) / kCompactTile;
    uint32_t* tickets = reinterpret_cast<uint32_t*>(c->tri_status + 4 * tiles);
    TriArgs t;
    memset(&t, 0, sizeof t);
    t.cmds = o.cmds[0];
    t.src_slot = o.src_slot;
    t.lod_tab = c->lod_tab;
    t.slot_chunks = c->slot_chunks;
    t.chunks = c->chunks;
    t.n_slots = c->n_meshes * kMaxLods;
    t.n_chunks = c->n_chunks;
    t.mixed_kinds = c->any_window_slot && c->any_gather_slot ? 1u : 0u;
    t.count = o.counts;
    t.model = reinterpret_cast<const float*>(c->out[buffer_ix].model);
    t.positions = c->positions;
    t.indices = c->indices;
    t.out_indices = c->tri_indices;
    t.refined = c->tri_refined;
    t.ticket = tickets;
    t.error_flag = c->ticket + 1;
    t.n_vertices = c->n_vertices;
    t.n_indices = c->n_indices;
    t.out_capacity = c->cfg.capacity_tri_indices;
    t.cap_cmds = (uint32_t)(c->cap_cmds > 0xffffffffull ? 0xffffffffull : c->cap_cmds);
    t.draw_index_base = draw_index_base;
    for (int i = 0; i < 16; ++i) t.pv.m[i] = make_float2(pv[i], pv[i]);
    // the LODs with a cacheable vertex window, then the others (no shared memory: a large L1 for their gathers);
    // the two kernels touch disjoint commands and may overlap
    if (c->any_window_slot || !c->any_gather_slot) {
        tri_cull<true><<<c->sm_count * 8, kTriThreads, 0, s>>>(t);
        c->launches += 1;
    }
    if (c->any_gather_slot) {
        tri_cull<false><<<c->sm_count * 8, kTriThreads, 0, s>>>(t);
        c->launches += 1;
    }
    CU(cudaGetLastError());
    CompactArgs k;
    memset(&k, 0, sizeof k);
    k.cmds = o.cmds[0];
    k.refined = c->tri_refined;
    k.count = o.counts;
    k.out_cmds = c->tri_cmds;
    k.out_count = c->tri_count;
    k.kept_indices = c->tri_kept;
    k.status = c->tri_status;
    k.ticket = tickets + 2;
    k.cap_cmds = t.cap_cmds;
    compact_draw_stream<<<c->sm_count * 2, kCompactThreads, 0, s>>>(k);
    CU(cudaGetLastError());
    c->launches += 1;
    return RCULL_OK;
}

int rcull_get_tri_outputs(rcull_ctx* c, rcull_tri_outputs* out) {
    if (!c || !out) return fail(RCULL_ERR_INVALID, "rcull_get_tri_outputs: null argument");
    if (!c->cfg.capacity_tri_indices) return fail(RCULL_ERR_INVALID, "context created without capacity_tri_indices");
    out->cmds = reinterpret_cast<rcull_draw_cmd*>(c->tri_cmds);
    out->count = c->tri_count;
    out->indices = c->tri_indices;
    out->kept_indices = c->tri_kept;
    return RCULL_OK;
}

int rcull_download_tri(rcull_ctx* c, uint32_t* count, uint64_t* kept_indices, rcull_draw_cmd* cmds, uint64_t n_cmds,
                       uint32_t* indices, uint64_t first_index, uint64_t n_indices) {
    if (!c || !count) return fail(RCULL_ERR_INVALID, "rcull_download_tri: null argument");
    if (!c->cfg.capacity_tri_indices) return fail(RCULL_ERR_INVALID, "context created without capacity_tri_indices");
    CU(cudaSetDevice(c->cfg.device));
    CU(cudaDeviceSynchronize());
    int rc = check_device_error(c);
    if (rc) return rc;
    CU(cudaMemcpy(count, c->tri_count, 4, cudaMemcpyDeviceToHost));
    if (kept_indices) CU(cudaMemcpy(kept_indices, c->tri_kept, 8, cudaMemcpyDeviceToHost));
    if (cmds) {
        if (n_cmds > c->cap_cmds) return fail(RCULL_ERR_INVALID, "command range out of bounds");
        CU(cudaMemcpy(cmds, c->tri_cmds, n_cmds * 20, cudaMemcpyDeviceToHost));
    }
    if (indices) {
        if (first_index + n_indices > c->cfg.capacity_tri_indices) return fail(RCULL_ERR_INVALID, "index range out of bounds");
        CU(cudaMemcpy(indices, c->tri_indices + first_index, n_indices * 4, cudaMemcpyDeviceToHost));
    }
    return RCULL_OK;
}

#ifdef RCULL_SCAN_STATS  // diagnostic build only, not part of include/rcull.h
int rcull_debug_stats(uint64_t out[16], int reset) {
    if (cudaMemcpyFromSymbol(out, rcull::g_scan_stats, 16 * 8) != cudaSuccess) return -1;
    if (reset) {
        uint64_t z[16] = {0};
        if (cudaMemcpyToSymbol(rcull::g_scan_stats, z, 16 * 8) != cudaSuccess) return -1;
    }
    return 0;
}
#endif

}  // extern "C"

// ---------------------------------------------------------------------------------
// range-shard merge over peer memory (records + expand, rcull_aux.cuh)
// ---------------------------------------------------------------------------------
struct rcull_merge {
    rcull_ctx* ctx = nullptr;
    uint32_t world = 0, rank = 0;
    uint64_t cap = 0;    // merged list capacity per view
    uint64_t cap_r = 0;  // records one rank can send per view
    size_t window_bytes = 0, sync_off = 0;
    size_t slot_stride = 0, ckpt_in_slot = 0, meta_in_slot = 0;  // window layout (rcull_aux.cuh)
    unsigned char* window = nullptr;  // peer-visible: the record regions of every source rank + the mailbox
    bool owns_window = true;          // false: caller's allocation (symmetric memory), rcull_merge_create_on
    unsigned char* peer[kMaxRanks] = {};
    bool opened[kMaxRanks] = {};
    unsigned char* mc = nullptr;      // NVLS multicast mapping of all windows, or null
    // local outputs: the merged draw lists
    uint32_t* cmds[kMaxViews] = {};
    uint32_t* visible[kMaxViews] = {};
    uint32_t* keys[kMaxViews] = {};     // per-mesh mode: merged slot keys / draw indices before bucketing
    uint32_t* vis_tmp[kMaxViews] = {};
    uint32_t* counts = nullptr;
    unsigned long long* totals = nullptr;
    uint32_t* ticket = nullptr;    // [0] expand tile ticket, [1] expand CTAs done, [2] frame CTAs done
    unsigned long long frame_epoch = 0, expand_epoch = 0;  // merged frames issued / expanded
    uint32_t views_of[2] = {0, 0};  // n_views of the frame with that epoch parity
    int expand_sms = 0;             // > 0: the expand runs on that many SMs of its own beside the next frame (rcull_merge_set_overlap)
    // transport 1 (rcull_merge_set_transport): the frame kernel keeps its records in its OWN window and rcull_merge_send
    // moves them to the peers with the copy engines -- no SM stores over NVLink, no store traffic beside the frame's own
    int transport = 0;
    unsigned long long sent_epoch = 0;          // merged frames whose records rcull_merge_send has queued
    cudaEvent_t frame_done[2] = {nullptr, nullptr};  // [parity] the frame of that parity has finished on its stream
    cudaEvent_t sent[2] = {nullptr, nullptr};        // [parity] the copies of that parity's region have finished
    bool sent_valid[2] = {false, false};
    cudaStream_t ce_stream[kMaxRanks] = {};  // one stream per peer: the peer copies of a send run on several copy engines at once
    cudaEvent_t ce_fork = nullptr, ce_join[kMaxRanks] = {};
    int ce_streams = 0;  // 0 (default): every copy on the caller's stream, one after the other; 1: one stream per peer (RCULL_CE_STREAMS=1)
    uint32_t send_bound[kMaxViews] = {};  // records of a view that rcull_merge_send copies (rcull_merge_set_send_bound); 0 = the whole capacity
    std::vector<void*> allocs;
};

static_assert(RCULL_MAX_RANKS == kMaxRanks && RCULL_IPC_HANDLE_BYTES == sizeof(cudaIpcMemHandle_t), "merge ABI constants");

namespace {

// window = [2][world] slots { records [views][cap_r] x 8 B | checkpoints [views][cap_r / 256 + 1] x 4 B | meta 24 x 8 B } | mailbox
constexpr size_t kMetaBytes = 24 * 8;
size_t merge_ckpt_view_stride(uint64_t cap_r) { return ((size_t)(cap_r / kExpandTile + 1) * 4 + 15) / 16 * 16; }
size_t merge_window_layout(uint32_t world, uint32_t max_views, uint64_t cap_r, size_t* slot_stride, size_t* ckpt_in_slot, size_t* meta_in_slot,
                           size_t* sync_off) {
    const size_t ck = (size_t)max_views * cap_r * 8;  // cap_r is even: 16-byte aligned
    const size_t mt = ck + (size_t)max_views * merge_ckpt_view_stride(cap_r);
    const size_t st = (mt + kMetaBytes + 255) / 256 * 256;
    const size_t so = (size_t)2 * world * st;
    if (slot_stride) *slot_stride = st;
    if (ckpt_in_slot) *ckpt_in_slot = ck;
    if (meta_in_slot) *meta_in_slot = mt;
    if (sync_off) *sync_off = so;
    return so + sizeof(MergeSync);
}

// Called by rcull_frame on a context with a merge attached: numbers the frame, fills the kernel's RecordArgs (or, for
// an empty frame, sends the zero counts and the tag right away).
int merge_frame_issued(rcull_ctx* c, uint32_t n_views, const uint32_t* counts, const unsigned long long* totals, RecordArgs* out,
                       bool empty, cudaStream_t s) {
    rcull_merge* m = c->merge;
    for (uint32_t p = 0; p < m->world; ++p)
        if (!m->peer[p]) return fail(RCULL_ERR_INVALID, "merge: peer %u has not been mapped (rcull_merge_open_peer / rcull_merge_set_peer_ptr)", p);
    if (m->frame_epoch - m->expand_epoch >= 2)
        return fail(RCULL_ERR_INVALID, "merge: two merged frames are in flight -- call rcull_merge_expand before the next rcull_frame");
    RecordArgs r;
    memset(&r, 0, sizeof r);
    if (m->transport == 1) {  // copy engines: the kernel writes into its own window only
        if (m->frame_epoch != m->sent_epoch)
            return fail(RCULL_ERR_INVALID, "merge: the records of the last frame have not been sent -- call rcull_merge_send before the next rcull_frame");
        r.peer[0] = m->window;
        r.n_dst = 1;
        // this frame overwrites the region the copies of two frames ago read from
        if (m->sent_valid[(m->frame_epoch + 1) & 1ull]) CU(cudaStreamWaitEvent(s, m->sent[(m->frame_epoch + 1) & 1ull], 0));
    } else {
        for (uint32_t p = 0; p < m->world; ++p) r.peer[p] = m->peer[p];
        r.n_dst = m->world;
        r.mc = m->mc;
    }
    r.own = m->window;
    r.epoch = ++m->frame_epoch;
    r.base = (size_t)((r.epoch & 1ull) * m->world + m->rank) * m->slot_stride;
    r.view_stride = (size_t)m->cap_r * 8;
    r.ckpt_view_stride = merge_ckpt_view_stride(m->cap_r);
    r.ckpt_base = r.base + m->ckpt_in_slot;
    r.meta_base = r.base + m->meta_in_slot;
    r.sync_off = m->sync_off;
    r.cta_done = m->ticket + 2;
    r.world = m->world;
    r.rank = m->rank;
    r.cap_r = (uint32_t)m->cap_r;
    m->views_of[r.epoch & 1ull] = n_views;
    if (empty) {
        record_empty_frame<<<1, 32, 0, s>>>(r, counts, totals);
        CU(cudaGetLastError());
        c->launches += 1;
        if (m->transport == 1) CU(cudaEventRecord(m->frame_done[r.epoch & 1ull], s));
        return RCULL_OK;
    }
    *out = r;
    return RCULL_OK;
}

// After the frame kernel of a merged frame has been launched on s (copy-engine transport): the point rcull_merge_send waits for.
int merge_frame_launched(rcull_ctx* c, cudaStream_t s) {
    rcull_merge* m = c->merge;
    if (m->transport == 1) CU(cudaEventRecord(m->frame_done[m->frame_epoch & 1ull], s));
    return RCULL_OK;
}

template <typename T>
int merge_alloc(rcull_merge* m, T** p, size_t count) {
    void* q = nullptr;
    CU(cudaMalloc(&q, count * sizeof(T) + 256));
    m->allocs.push_back(q);
    *p = static_cast<T*>(q);
    return 0;
}

int merge_create(rcull_ctx* c, uint32_t world, uint32_t rank, uint64_t capacity, uint64_t capacity_per_rank, void* external_window,
                 rcull_merge** out) {
    if (!c || !out) return fail(RCULL_ERR_INVALID, "rcull_merge_create: null argument");
    *out = nullptr;
    if (world < 1 || world > kMaxRanks || rank >= world) return fail(RCULL_ERR_INVALID, "rank %u of world %u (max %d)", rank, world, kMaxRanks);
    if (capacity == 0 || capacity > 0xffffffffull) return fail(RCULL_ERR_INVALID, "merged capacity must be 1..2^32-1 commands");
    if (capacity_per_rank == 0) capacity_per_rank = capacity;
    if (capacity_per_rank > capacity) capacity_per_rank = capacity;
    capacity_per_rank = (capacity_per_rank + 1) & ~1ull;  // record regions are copied in 16-byte units
    if (c->merge) return fail(RCULL_ERR_INVALID, "the context already has a merge attached");
    if (c->cfg.capacity_tri_indices) return fail(RCULL_ERR_INVALID, "the per-triangle stage works on a rank's own lists: create the context without capacity_tri_indices");
    CU(cudaSetDevice(c->cfg.device));
    rcull_merge* m = new rcull_merge();
    m->ctx = c;
    m->world = world;
    m->rank = rank;
    m->cap = capacity;
    m->cap_r = capacity_per_rank;
    const uint32_t mv = c->cfg.max_views;
    const bool per_mesh = c->cfg.output_mode == RCULL_PER_MESH_INSTANCED;
    m->window_bytes = merge_window_layout(world, mv, m->cap_r, &m->slot_stride, &m->ckpt_in_slot, &m->meta_in_slot, &m->sync_off);
    int rc = 0;
#define MTRY(x)                      \
    if ((rc = (x)) != 0) {           \
        rcull_merge_destroy(m);      \
        return rc;                   \
    }
    if (external_window) {
        m->window = static_cast<unsigned char*>(external_window);
        m->owns_window = false;
    } else if (cudaMalloc((void**)&m->window, m->window_bytes) != cudaSuccess) {
        rcull_merge_destroy(m);
        return fail(RCULL_ERR_CUDA, "rcull_merge_create: cudaMalloc of a %zu-byte window failed", m->window_bytes);
    }
    const uint64_t cmd_cap = per_mesh ? (uint64_t)c->cfg.max_meshes * kMaxLods : capacity;
    for (uint32_t v = 0; v < mv; ++v) {
        MTRY(merge_alloc(m, &m->cmds[v], cmd_cap * 5));
        MTRY(merge_alloc(m, &m->visible[v], capacity));
        if (per_mesh) {
            MTRY(merge_alloc(m, &m->keys[v], capacity));
            MTRY(merge_alloc(m, &m->vis_tmp[v], capacity));
        }
    }
    MTRY(merge_alloc(m, &m->counts, kMaxViews));
    MTRY(merge_alloc(m, &m->totals, kMaxViews));
    MTRY(merge_alloc(m, &m->ticket, 4));
#undef MTRY
    cudaMemset(m->window + m->sync_off, 0, sizeof(MergeSync));
    for (uint32_t sl = 0; sl < 2 * world; ++sl) cudaMemset(m->window + (size_t)sl * m->slot_stride + m->meta_in_slot, 0, kMetaBytes);
    cudaMemset(m->counts, 0, kMaxViews * 4);
    cudaMemset(m->totals, 0, kMaxViews * 8);
    cudaMemset(m->ticket, 0, 16);
    if (cudaDeviceSynchronize() != cudaSuccess) {
        rcull_merge_destroy(m);
        return fail(RCULL_ERR_CUDA, "rcull_merge_create: clearing the window failed");
    }
    cudaFuncSetAttribute(merge_expand, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kExpandSmem);
    cudaFuncSetAttribute(merge_expand_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kExpandBigSmem);
    m->peer[rank] = m->window;
    c->merge = m;  // frames issued from now on send records
    *out = m;
    return RCULL_OK;
}

}  // namespace

extern "C" {

void rcull_merge_destroy(rcull_merge* m) {
    if (!m) return;
    cudaSetDevice(m->ctx->cfg.device);
    cudaDeviceSynchronize();
    if (m->ctx->merge == m) {
        m->ctx->merge = nullptr;
        m->ctx->reserved_sms = 0;
    }
    for (uint32_t p = 0; p < m->world; ++p)
        if (m->opened[p]) cudaIpcCloseMemHandle(m->peer[p]);
    for (int i = 0; i < 2; ++i) {
        if (m->frame_done[i]) cudaEventDestroy(m->frame_done[i]);
        if (m->sent[i]) cudaEventDestroy(m->sent[i]);
    }
    for (int i = 0; i < kMaxRanks; ++i) {
        if (m->ce_stream[i]) cudaStreamDestroy(m->ce_stream[i]);
        if (m->ce_join[i]) cudaEventDestroy(m->ce_join[i]);
    }
    if (m->ce_fork) cudaEventDestroy(m->ce_fork);
    if (m->owns_window && m->window) cudaFree(m->window);
    for (void* p : m->allocs) cudaFree(p);
    delete m;
}

int rcull_merge_window_bytes(rcull_ctx* c, uint32_t world, uint64_t capacity_per_rank, uint64_t* bytes) {
    if (!c || !bytes || world < 1 || world > kMaxRanks || capacity_per_rank == 0) return fail(RCULL_ERR_INVALID, "rcull_merge_window_bytes: bad argument");
    *bytes = merge_window_layout(world, c->cfg.max_views, (capacity_per_rank + 1) & ~1ull, nullptr, nullptr, nullptr, nullptr);
    return RCULL_OK;
}

int rcull_merge_create(rcull_ctx* c, uint32_t world, uint32_t rank, uint64_t capacity, uint64_t capacity_per_rank, rcull_merge** out) {
    return merge_create(c, world, rank, capacity, capacity_per_rank, nullptr, out);
}

int rcull_merge_create_on(rcull_ctx* c, uint32_t world, uint32_t rank, uint64_t capacity, uint64_t capacity_per_rank, void* window,
                          rcull_merge** out) {
    if (!window) return fail(RCULL_ERR_INVALID, "rcull_merge_create_on: null window");
    return merge_create(c, world, rank, capacity, capacity_per_rank, window, out);
}

int rcull_merge_set_multicast_ptr(rcull_merge* m, void* multicast_window) {
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_set_multicast_ptr: null argument");
    m->mc = static_cast<unsigned char*>(multicast_window);
    return RCULL_OK;
}

int rcull_merge_ipc_handle(rcull_merge* m, void* handle_out) {
    if (!m || !handle_out) return fail(RCULL_ERR_INVALID, "rcull_merge_ipc_handle: null argument");
    CU(cudaSetDevice(m->ctx->cfg.device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, m->window));
    memcpy(handle_out, &h, sizeof h);
    return RCULL_OK;
}

int rcull_merge_open_peer(rcull_merge* m, uint32_t peer, const void* handle) {
    if (!m || !handle) return fail(RCULL_ERR_INVALID, "rcull_merge_open_peer: null argument");
    if (peer >= m->world || peer == m->rank) return fail(RCULL_ERR_INVALID, "peer %u is not another rank of world %u", peer, m->world);
    CU(cudaSetDevice(m->ctx->cfg.device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    void* p = nullptr;
    CU(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    m->peer[peer] = static_cast<unsigned char*>(p);
    m->opened[peer] = true;
    return RCULL_OK;
}

int rcull_merge_set_peer_ptr(rcull_merge* m, uint32_t peer, void* window) {
    if (!m || !window) return fail(RCULL_ERR_INVALID, "rcull_merge_set_peer_ptr: null argument");
    if (peer >= m->world) return fail(RCULL_ERR_INVALID, "peer %u of world %u", peer, m->world);
    m->peer[peer] = static_cast<unsigned char*>(window);
    return RCULL_OK;
}

int rcull_merge_window(rcull_merge* m, void** window) {
    if (!m || !window) return fail(RCULL_ERR_INVALID, "rcull_merge_window: null argument");
    *window = m->window;
    return RCULL_OK;
}

int rcull_merge_set_overlap(rcull_merge* m, uint32_t expand_sms) {
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_set_overlap: null argument");
    if ((int)expand_sms >= m->ctx->sm_count / 2) return fail(RCULL_ERR_INVALID, "rcull_merge_set_overlap: %u of %d SMs", expand_sms, m->ctx->sm_count);
    CU(cudaSetDevice(m->ctx->cfg.device));
    m->expand_sms = (int)expand_sms;
    m->ctx->reserved_sms = (int)expand_sms;  // the frame kernel's persistent grid leaves them free
    return RCULL_OK;
}

int rcull_merge_set_transport(rcull_merge* m, uint32_t transport) {
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_set_transport: null argument");
    if (transport > 1) return fail(RCULL_ERR_INVALID, "rcull_merge_set_transport: 0 = peer stores from the frame kernel, 1 = copy engines");
    if (m->frame_epoch != m->expand_epoch || m->frame_epoch != (m->transport == 1 ? m->sent_epoch : m->frame_epoch))
        return fail(RCULL_ERR_INVALID, "rcull_merge_set_transport: a merged frame is in flight");
    CU(cudaSetDevice(m->ctx->cfg.device));
    if (transport == 1) {
        for (int i = 0; i < 2; ++i) {
            if (!m->frame_done[i]) CU(cudaEventCreateWithFlags(&m->frame_done[i], cudaEventDisableTiming));
            if (!m->sent[i]) CU(cudaEventCreateWithFlags(&m->sent[i], cudaEventDisableTiming));
        }
        const char* e = getenv("RCULL_CE_STREAMS");  // A/B switch: 1 = one stream per peer instead of serial copies on the caller's stream
        m->ce_streams = (e && atoi(e) != 0) ? 1 : 0;
        if (m->ce_streams) {
            if (!m->ce_fork) CU(cudaEventCreateWithFlags(&m->ce_fork, cudaEventDisableTiming));
            for (uint32_t i = 1; i < m->world; ++i) {
                if (!m->ce_stream[i]) CU(cudaStreamCreateWithFlags(&m->ce_stream[i], cudaStreamNonBlocking));
                if (!m->ce_join[i]) CU(cudaEventCreateWithFlags(&m->ce_join[i], cudaEventDisableTiming));
            }
        }
    }
    m->transport = (int)transport;
    m->sent_epoch = m->frame_epoch;
    return RCULL_OK;
}

int rcull_merge_set_send_bound(rcull_merge* m, const uint32_t* records_per_view, uint32_t n_views) {
    if (!m || (n_views && !records_per_view)) return fail(RCULL_ERR_INVALID, "rcull_merge_set_send_bound: null argument");
    if (m->transport != 1) return fail(RCULL_ERR_INVALID, "rcull_merge_set_send_bound: only the copy-engine transport has a send (rcull_merge_set_transport(m, 1))");
    if (n_views > m->ctx->cfg.max_views) return fail(RCULL_ERR_INVALID, "rcull_merge_set_send_bound: %u views of %u", n_views, m->ctx->cfg.max_views);
    if (m->frame_epoch != m->expand_epoch || m->frame_epoch != m->sent_epoch) return fail(RCULL_ERR_INVALID, "rcull_merge_set_send_bound: a merged frame is in flight");
    CU(cudaSetDevice(m->ctx->cfg.device));
    CU(cudaDeviceSynchronize());
    unsigned long long words[2][8] = {};
    for (uint32_t v = 0; v < (uint32_t)kMaxViews; ++v) {
        uint64_t b = v < n_views ? ((uint64_t)records_per_view[v] + kExpandTile - 1) / kExpandTile * kExpandTile : 0;  // whole expand tiles
        if (v >= n_views || b >= m->cap_r) b = 0;  // 0: the whole capacity is copied, nothing is left behind
        m->send_bound[v] = (uint32_t)b;
        words[0][v] = words[1][v] = b ? b + 1 : 0;
    }
    // the bound words ride behind this rank's counts in its own mailbox slot (both parities); every send copies them along
    for (uint32_t par = 0; par < 2; ++par)
        CU(cudaMemcpy(m->window + (size_t)(par * m->world + m->rank) * m->slot_stride + m->meta_in_slot + 16 * 8, words[par], 8 * 8, cudaMemcpyHostToDevice));
    return RCULL_OK;
}

int rcull_merge_send(rcull_merge* m, void* stream) {
    Scope scope("rcull::merge_send (a rank's records -> every peer's window, copy engines)");
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_send: null argument");
    if (m->transport != 1) return fail(RCULL_ERR_INVALID, "rcull_merge_send: the merge sends from the frame kernel (rcull_merge_set_transport(m, 1) selects the copy engines)");
    if (m->sent_epoch >= m->frame_epoch) return fail(RCULL_ERR_INVALID, "rcull_merge_send: no merged frame is waiting to be sent");
    for (uint32_t p = 0; p < m->world; ++p)
        if (!m->peer[p]) return fail(RCULL_ERR_INVALID, "merge: peer %u has not been mapped", p);
    rcull_ctx* c = m->ctx;
    CU(cudaSetDevice(c->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned long long epoch = ++m->sent_epoch;
    const uint32_t par = (uint32_t)(epoch & 1ull);
    // The slot of (parity, this rank): records | checkpoints | meta in one piece.  A peer copy costs ~13 us before its
    // first byte moves (profiles/p2p_copy_bench.py), so the NUMBER of copies is what matters: one for the slot, one for
    // the tag.  With send bounds the records go view by view (what the bound covers; the rest stays here and the
    // receivers read it over NVLink themselves) and the checkpoints + meta follow as one copy.
    const size_t slot_off = (size_t)(par * m->world + m->rank) * m->slot_stride;
    const size_t slot_bytes = m->meta_in_slot + kMetaBytes;
    bool bounded = false;
    for (uint32_t v = 0; v < c->cfg.max_views; ++v) bounded = bounded || m->send_bound[v] != 0;
    auto copy_slot = [&](unsigned char* dst, cudaStream_t t) -> cudaError_t {
        if (!bounded) return cudaMemcpyAsync(dst + slot_off, m->window + slot_off, slot_bytes, cudaMemcpyDeviceToDevice, t);
        const uint32_t nv = m->views_of[par];
        for (uint32_t v = 0; v < nv; ++v) {
            const size_t off = slot_off + (size_t)v * m->cap_r * 8, bytes = (size_t)(m->send_bound[v] ? m->send_bound[v] : m->cap_r) * 8;
            cudaError_t e = cudaMemcpyAsync(dst + off, m->window + off, bytes, cudaMemcpyDeviceToDevice, t);
            if (e != cudaSuccess) return e;
        }
        return cudaMemcpyAsync(dst + slot_off + m->ckpt_in_slot, m->window + slot_off + m->ckpt_in_slot, slot_bytes - m->ckpt_in_slot,
                               cudaMemcpyDeviceToDevice, t);
    };
    const size_t tag_off = m->sync_off + offsetof(MergeSync, frame_tag) + ((size_t)par * kMaxRanks + m->rank) * 8;
    CU(cudaStreamWaitEvent(s, m->frame_done[par], 0));  // the frame kernel has written its slot and its own tag
    // Copies on one stream run in order: for every peer the tag goes last, behind everything it announces.  Peers are
    // visited starting behind this rank so that the ranks do not all write into the same GPU at once.
    if (m->ce_streams) {  // one stream per peer, forked from and joined back into the caller's (RCULL_CE_STREAMS=1; measured no faster, and heavier on the host)
        CU(cudaEventRecord(m->ce_fork, s));
        for (uint32_t i = 1; i < m->world; ++i) {
            unsigned char* dst = m->peer[(m->rank + i) % m->world];
            cudaStream_t t = m->ce_stream[i];
            CU(cudaStreamWaitEvent(t, m->ce_fork, 0));
            CU(copy_slot(dst, t));
            CU(cudaMemcpyAsync(dst + tag_off, m->window + tag_off, 8, cudaMemcpyDeviceToDevice, t));
            CU(cudaEventRecord(m->ce_join[i], t));
            CU(cudaStreamWaitEvent(s, m->ce_join[i], 0));
        }
    } else {
        for (uint32_t i = 1; i < m->world; ++i) {
            unsigned char* dst = m->peer[(m->rank + i) % m->world];
            CU(copy_slot(dst, s));
            CU(cudaMemcpyAsync(dst + tag_off, m->window + tag_off, 8, cudaMemcpyDeviceToDevice, s));
        }
    }
    CU(cudaEventRecord(m->sent[par], s));
    m->sent_valid[par] = true;
    return RCULL_OK;
}

int rcull_merge_expand(rcull_merge* m, void* stream) {
    Scope scope("rcull::merge_expand (records of every rank -> merged draw lists)");
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_expand: null argument");
    rcull_ctx* c = m->ctx;
    if (m->expand_epoch >= m->frame_epoch) return fail(RCULL_ERR_INVALID, "rcull_merge_expand: no merged frame is waiting (call rcull_frame first)");
    for (uint32_t p = 0; p < m->world; ++p)
        if (!m->peer[p]) return fail(RCULL_ERR_INVALID, "merge: peer %u has not been mapped", p);
    CU(cudaSetDevice(c->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned long long epoch = ++m->expand_epoch;
    const uint32_t n_views = m->views_of[epoch & 1ull];
    const bool per_mesh = c->cfg.output_mode == RCULL_PER_MESH_INSTANCED;
    ExpandArgs a;
    memset(&a, 0, sizeof a);
    a.window = m->window;
    for (uint32_t p = 0; p < m->world; ++p) a.peer[p] = m->peer[p];
    a.sync_off = m->sync_off;
    a.view_stride = (size_t)m->cap_r * 8;
    a.slot_stride = m->slot_stride;
    a.ckpt_in_slot = m->ckpt_in_slot;
    a.meta_in_slot = m->meta_in_slot;
    a.ckpt_view_stride = merge_ckpt_view_stride(m->cap_r);
    a.epoch = epoch;
    a.world = m->world;
    a.rank = m->rank;
    a.n_views = n_views;
    a.max_views = c->cfg.max_views;
    a.cap_r = (uint32_t)m->cap_r;
    a.cap = (uint32_t)m->cap;
    for (uint32_t v = 0; v < n_views; ++v) {
        a.cmds[v] = per_mesh ? m->keys[v] : m->cmds[v];
        a.visible[v] = per_mesh ? m->vis_tmp[v] : m->visible[v];
    }
    a.counts = m->counts;
    a.totals = m->totals;
    a.lod_tab = c->lod_tab;
    a.n_slots = c->n_meshes * kMaxLods;
    a.pairs = per_mesh ? 1u : 0u;
    a.cta_done = m->ticket + 1;
    if (m->expand_sms > 0) {
        // R CTAs that each take a whole SM, on the caller's second stream: beside the next frame, whose grid leaves R SMs
        merge_expand_big<<<m->expand_sms, 1024, kExpandBigSmem, s>>>(a);
    } else {
        merge_expand<<<c->sm_count * 4, 256, kExpandSmem, s>>>(a);
    }
    CU(cudaGetLastError());
    c->launches += 1;
    if (per_mesh && n_views) {
        // the merged pairs are in draw order (rank order = ascending ranges): the single-GPU bucketing passes finish the job
        const uint32_t n_slots = c->n_meshes * kMaxLods;
        BucketArgs b;
        memset(&b, 0, sizeof b);
        for (uint32_t v = 0; v < n_views; ++v) {
            b.keys[v] = m->keys[v];
            b.draw_ix[v] = m->vis_tmp[v];
            b.visible[v] = m->visible[v];
            b.cmds[v] = m->cmds[v];
        }
        b.counts = m->counts;
        b.totals = m->totals;
        b.hist = c->bucket_hist;
        b.slot_total = c->slot_total;
        b.slot_base = c->slot_base;
        b.lod_tab = c->lod_tab;
        b.n_slots = n_slots;
        b.cap = (uint32_t)m->cap;
        const uint64_t rows_max = c->bucket_hist_words / ((uint64_t)n_slots * n_views);
        b.rows_max = (uint32_t)(rows_max < 0xffffffffull ? rows_max : 0xffffffffull);
        uint32_t sweep_ctas = (uint32_t)c->sm_count * 7 / n_views;
        if (sweep_ctas == 0) sweep_ctas = 1;
        b.rows_target = sweep_ctas * kBucketWarps;
        const dim3 sweep_grid(sweep_ctas, n_views);
        b.colscan_done = c->bucket_done;
        CU(launch_bucketing(b, sweep_grid, n_slots, n_views, s));
        c->launches += 3;
    }
    return RCULL_OK;
}

int rcull_merge_lists(rcull_merge* m, uint32_t view, rcull_draw_cmd** cmds, uint32_t** visible, uint32_t** counts, uint64_t** index_totals) {
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_lists: null argument");
    const uint32_t mv = m->ctx->cfg.max_views;
    if (view >= mv) return fail(RCULL_ERR_INVALID, "view %u of %u", view, mv);
    if (cmds) *cmds = reinterpret_cast<rcull_draw_cmd*>(m->cmds[view]);
    if (visible) *visible = m->visible[view];
    if (counts) *counts = m->counts;
    if (index_totals) *index_totals = reinterpret_cast<uint64_t*>(m->totals);
    return RCULL_OK;
}

int rcull_merge_download(rcull_merge* m, uint32_t view, uint32_t* count, uint64_t* index_total, rcull_draw_cmd* cmds, uint32_t* visible,
                         uint64_t n) {
    if (!m) return fail(RCULL_ERR_INVALID, "rcull_merge_download: null argument");
    const uint32_t mv = m->ctx->cfg.max_views;
    if (view >= mv) return fail(RCULL_ERR_INVALID, "view %u of %u", view, mv);
    CU(cudaSetDevice(m->ctx->cfg.device));
    CU(cudaDeviceSynchronize());
    uint32_t c[kMaxViews];
    unsigned long long t[kMaxViews];
    CU(cudaMemcpy(c, m->counts, sizeof c, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(t, m->totals, sizeof t, cudaMemcpyDeviceToHost));
    if (count) *count = c[view];
    if (index_total) *index_total = t[view];
    const bool per_mesh = m->ctx->cfg.output_mode == RCULL_PER_MESH_INSTANCED;
    const uint64_t have_cmds = per_mesh ? (uint64_t)m->ctx->cfg.max_meshes * kMaxLods : m->cap;
    if (cmds && n) CU(cudaMemcpy(cmds, m->cmds[view], (n < have_cmds ? n : have_cmds) * 20, cudaMemcpyDeviceToHost));
    if (visible && n) CU(cudaMemcpy(visible, m->visible[view], (n < m->cap ? n : m->cap) * 4, cudaMemcpyDeviceToHost));
    return RCULL_OK;
}

}  // extern "C"
