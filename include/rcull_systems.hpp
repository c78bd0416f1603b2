// rcull_systems.hpp -- C++17 host side above the C ABI (include/rcull.h): the reference's ECS
// component / resource / system interface for the transform -> cull -> compact stage, restated for
// a host without rustc.  Header-only; nothing here computes on entities -- every system ends in
// calls into librcull.so, and there is no CPU fallback (without a CUDA device RcullRes panics).
//
// What mirrors what (paths relative to the reference tree):
//   Position / Rotation / Scale / ModelMatrix / AABB / Light   src/ecs/components.rs:5-27
//   GltfMesh (as the row of the mesh table)                    src/renderer.rs:117-126
//   Camera + project_camera                                    src/ecs/camera_controller.rs:10-24, src/ecs.rs:66-91
//   shadow_mapping_mvp_calculation (per-light pv)              src/renderer/systems/shadow_mapping.rs:314-342
//   ImageIndex / DoubleBuffered::current                       src/renderer/helpers/double_buffered.rs:17-19
//   RuntimeConfiguration.freeze_culling                        src/renderer/systems/cull_pipeline.rs:331-421
//   gpu_transform_cull = model_matrix_calculation + aabb_calculation + coarse_culling +
//       model_matrices_upload + cull_pass' emit loop           src/ecs.rs:52-64,138-181; cull_pipeline.rs:99-120,534-577;
//                                                              src/renderer.rs:2266-2288
//   gpu_cull_triangles = cull_pass' generate_work dispatches + compact_draw_stream
//                                                              cull_pipeline.rs:534-605
//   build_acceleration_structures' instance loop               src/renderer/systems/acceleration_strucures.rs:419-445
//
// Error behaviour: the reference unwraps / expects and is built with panic = "abort"
// (Cargo.toml:133,138).  rcull::ecs::expect() does the same -- message to stderr, std::abort() --
// unless the host installs a panic hook (tests install one that throws).
//
// Query iteration order: the caller hands the systems the component columns of the entities
// `With<Position, GltfMeshBaseColorTexture, GltfMesh>, Without<Deleting>` in query order;
// slot i is draw_index i (assign_draw_index, src/ecs.rs:117-136).
#ifndef RCULL_SYSTEMS_HPP
#define RCULL_SYSTEMS_HPP

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "rcull.h"

namespace rcull {
namespace ecs {

// ---- components (src/ecs/components.rs:5-27) ------------------------------------------------
struct Position { float x, y, z; };      // na::Point3<f32>
struct Rotation { float i, j, k, w; };   // na::UnitQuaternion<f32>, nalgebra storage order
struct Scale { float s; };               // uniform
struct ModelMatrix { float m[16]; };     // glm::Mat4, column-major
struct AABB { float mins[3], maxs[3]; }; // ncollide3d AABB<f32>
struct CoarseCulled { uint8_t culled; }; // cull_pipeline.rs:64-65 (1 = outside the frustum)
struct DrawIndex { uint32_t ix; };       // src/ecs.rs:117-136
struct GltfMesh { uint32_t table_row; }; // row of MeshTable below
struct Light { Position position; Rotation rotation; };  // Query<(&Position, &Rotation), With<Light>>
static_assert(sizeof(Position) == 12 && sizeof(Rotation) == 16 && sizeof(Scale) == 4, "component layout");
static_assert(sizeof(ModelMatrix) == 64 && sizeof(AABB) == 24 && sizeof(rcull_draw_cmd) == 20, "component layout");

// ---- resources ------------------------------------------------------------------------------
// Camera (src/ecs/camera_controller.rs:10-24): default position (0, 1, 2), identity rotation.
struct Camera {
    float position[3] = {0.f, 1.f, 2.f};
    float rotation[4] = {0.f, 0.f, 0.f, 1.f};  // i, j, k, w
    float projection[16] = {};
    float view[16] = {};
    float pv[16] = {};              // CameraMatrices.pv (generate_work.comp:33)
    rcull_view frustum_planes = {}; // left, right, bottom, top, near, far
};
struct ImageIndex { uint32_t ix = 0; };
struct RuntimeConfiguration { bool freeze_culling = false; };
// What GltfMesh + ConsolidatedMeshBuffers give cull_pass (cull_pipeline.rs:540-548).
struct MeshTable { std::vector<rcull_mesh> rows; };

// ---- panics ---------------------------------------------------------------------------------
using PanicHook = void (*)(const char* message);
inline PanicHook& panic_hook() {
    static PanicHook hook = nullptr;
    return hook;
}
[[noreturn]] inline void panic(const std::string& message) {
    if (panic_hook()) panic_hook()(message.c_str());  // a hook that returns still aborts
    std::fprintf(stderr, "rcull panic: %s\n", message.c_str());
    std::abort();
}
// `.expect(what)` on an rcull_status.
inline void expect(int status, const char* what) {
    if (status != RCULL_OK) panic(std::string(what) + ": " + rcull_last_error() + " (status " + std::to_string(status) + ")");
}

// ---- the context resource (ResMut<RcullRes>: exclusive access, one GPU) -----------------------
class RcullRes {
  public:
    explicit RcullRes(const rcull_config& cfg) : cfg_(cfg) {
        expect(rcull_create(&cfg_, &ctx_), "rcull_create");
        expect(rcull_context_stream(ctx_, &stream_), "rcull_context_stream");
    }
    ~RcullRes() { destroy(); }
    RcullRes(const RcullRes&) = delete;
    RcullRes& operator=(const RcullRes&) = delete;
    // explicit .destroy(device) convention (cull_pipeline.rs:276-286); idempotent
    void destroy() {
        if (ctx_) rcull_destroy(ctx_);
        ctx_ = nullptr;
    }
    rcull_ctx* ctx() const { return ctx_; }
    // Every system enqueues on this one stream (the one rcull_frame_host uses): frame i + 1's upload and
    // frame kernel are ordered after frame i's triangle stage / instance packing without a host sync.
    void* stream() const { return stream_; }
    const rcull_config& config() const { return cfg_; }

    // consolidate_mesh_buffers changed its offset maps (consolidate_mesh_buffers.rs:99-162)
    void set_mesh_table(const MeshTable& t) {
        expect(rcull_set_mesh_table(ctx_, t.rows.data(), static_cast<uint32_t>(t.rows.size())), "rcull_set_mesh_table");
    }
    void set_geometry(const std::vector<float>& positions_xyz, const std::vector<uint32_t>& indices) {
        expect(rcull_set_geometry(ctx_, positions_xyz.data(), positions_xyz.size() / 3, indices.data(), indices.size()),
               "rcull_set_geometry");
    }
    void set_blas_table(const std::vector<uint64_t>& addresses) {
        expect(rcull_set_blas_table(ctx_, addresses.data(), static_cast<uint32_t>(addresses.size())), "rcull_set_blas_table");
    }

    // Results of the last gpu_transform_cull (IndirectCommandsCount and the index sums per view).
    uint64_t n_entities = 0;
    uint32_t n_views = 0;
    uint32_t counts[RCULL_MAX_VIEWS] = {};
    uint64_t index_totals[RCULL_MAX_VIEWS] = {};

    // staging columns reused across frames (the Vec<f32> fields of the Rust resource)
    std::vector<float> pos_scale, rot;
    std::vector<uint32_t> mesh;

  private:
    rcull_config cfg_;
    rcull_ctx* ctx_ = nullptr;
    void* stream_ = nullptr;
};

// ---- systems ----------------------------------------------------------------------------------
// project_camera (src/ecs.rs:66-91): projection, view, pv and the six frustum planes.
inline void project_camera(Camera& camera, float aspect) {
    rcull_project_camera(aspect, camera.position, camera.rotation, camera.projection, camera.view, camera.pv,
                         &camera.frustum_planes);
}

// The frustum of one light's projection * view (shadow_mapping.rs:324-329) under the plane formula of
// src/ecs.rs:82-90 -- BASELINE configs[2]'s shadow views; the reference itself does not cull per light.
inline rcull_view light_frustum(const Light& light) {
    const float pos[3] = {light.position.x, light.position.y, light.position.z};
    const float rot[4] = {light.rotation.i, light.rotation.j, light.rotation.k, light.rotation.w};
    float proj[16], view[16], pv[16];
    rcull_view out;
    rcull_light_view(pos, rot, proj, view, pv, &out);
    return out;
}

// One call per frame in place of model_matrix_calculation, aabb_calculation, coarse_culling,
// model_matrices_upload and cull_pass' emit loop.  Columns are the query's components in
// iteration order; views = camera + one per light.  Outputs stay on the device
// (rcull_get_outputs); counts / index totals land in `rc`.
// With runtime_config.freeze_culling the reference runs cull_pass_bypass (cull_pipeline.rs:331-421):
// the entities keep moving and their model matrices follow, the draw lists, counts and culled
// indices stay those of the last unfrozen frame.
inline void gpu_transform_cull(RcullRes& rc, ImageIndex image_index, const Camera& camera, const std::vector<Light>& lights,
                               const RuntimeConfiguration& runtime_config, const Position* position, const Rotation* rotation,
                               const Scale* scale, const GltfMesh* gltf_mesh, uint64_t n_entities, uint32_t draw_index_base = 0) {
    rc.pos_scale.resize(4 * n_entities);
    rc.rot.resize(4 * n_entities);
    rc.mesh.resize(n_entities);
    expect(rcull_pack_components(n_entities, reinterpret_cast<const float*>(position), reinterpret_cast<const float*>(rotation),
                                 reinterpret_cast<const float*>(scale), rc.pos_scale.data(), rc.rot.data()),
           "rcull_pack_components");
    for (uint64_t e = 0; e < n_entities; ++e) rc.mesh[e] = gltf_mesh[e].table_row;

    if (runtime_config.freeze_culling) {
        expect(rcull_upload_entities(rc.ctx(), 0, n_entities, rc.pos_scale.data(), rc.rot.data(), rc.mesh.data(), rc.stream()),
               "rcull_upload_entities");
        expect(rcull_frame_frozen(rc.ctx(), n_entities, image_index.ix, rc.stream()), "rcull_frame_frozen");
        rc.n_entities = n_entities;
        return;  // rc.counts / rc.index_totals / rc.n_views keep describing the lists that are drawn
    }

    std::vector<rcull_view> views;
    views.reserve(1 + lights.size());
    views.push_back(camera.frustum_planes);
    for (const Light& l : lights) views.push_back(light_frustum(l));
    if (views.size() > rc.config().max_views) panic("gpu_transform_cull: more views than RcullRes was created for");

    expect(rcull_frame_host(rc.ctx(), n_entities, rc.pos_scale.data(), rc.rot.data(), rc.mesh.data(), views.data(),
                            static_cast<uint32_t>(views.size()), camera.position, draw_index_base, image_index.ix, rc.counts,
                            rc.index_totals),
           "rcull_frame_host");
    rc.n_entities = n_entities;
    rc.n_views = static_cast<uint32_t>(views.size());
}

// The second half of cull_pass: per-triangle cull of the camera view's survivors + compact_draw_stream.
// Frozen frames keep the triangle-stage outputs of the last unfrozen frame (cull_pass_bypass copies the
// buffers onto themselves), so there is nothing to launch.
inline void gpu_cull_triangles(RcullRes& rc, ImageIndex image_index, const Camera& camera, const RuntimeConfiguration& runtime_config,
                               uint32_t draw_index_base = 0) {
    if (runtime_config.freeze_culling) return;
    expect(rcull_tri_cull(rc.ctx(), camera.pv, draw_index_base, image_index.ix, rc.stream()), "rcull_tri_cull");
}

// build_acceleration_structures' per-entity instance loop (acceleration_strucures.rs:419-445).
inline void gpu_as_instances(RcullRes& rc, ImageIndex image_index, uint32_t draw_index_base = 0) {
    expect(rcull_as_instances(rc.ctx(), rc.n_entities, draw_index_base, image_index.ix, rc.stream()), "rcull_as_instances");
}

// ---- range shards over several GPUs of one process ---------------------------------
// The same stage with the query's entities split into contiguous ranges, one per GPU: every GPU culls its range, the
// frame kernels send their survivors to every GPU as 8-byte records, every GPU expands them into the merged draw
// lists -- byte-identical on every GPU to what one GPU would have produced (rcull_mgpu_*, rcull_merge_* in rcull.h).
class MultiGpuRes {
  public:
    MultiGpuRes(const rcull_config& cfg, uint32_t n_gpus, uint64_t merged_capacity, uint64_t capacity_per_rank = 0, bool multicast = true)
        : n_gpus_(n_gpus) {
        if (rcull_mgpu_create(&cfg, n_gpus, nullptr, merged_capacity, capacity_per_rank, multicast ? 1u : 0u, &g_) != RCULL_OK)
            panic(std::string("rcull_mgpu_create: ") + rcull_mgpu_last_error());
    }
    ~MultiGpuRes() { destroy(); }
    MultiGpuRes(const MultiGpuRes&) = delete;
    MultiGpuRes& operator=(const MultiGpuRes&) = delete;
    void destroy() {
        if (g_) rcull_mgpu_destroy(g_);
        g_ = nullptr;
    }
    uint32_t n_gpus() const { return n_gpus_; }
    rcull_mgpu* handle() const { return g_; }
    rcull_ctx* ctx(uint32_t rank) const {
        rcull_ctx* c = nullptr;
        expect(rcull_mgpu_rank(g_, rank, &c, nullptr, nullptr), "rcull_mgpu_rank");
        return c;
    }
    rcull_merge* merge(uint32_t rank) const {
        rcull_merge* m = nullptr;
        expect(rcull_mgpu_rank(g_, rank, nullptr, &m, nullptr), "rcull_mgpu_rank");
        return m;
    }
    void* stream(uint32_t rank) const {
        void* s = nullptr;
        expect(rcull_mgpu_rank(g_, rank, nullptr, nullptr, &s), "rcull_mgpu_rank");
        return s;
    }
    void set_mesh_table(const MeshTable& t) {
        for (uint32_t r = 0; r < n_gpus_; ++r)
            expect(rcull_set_mesh_table(ctx(r), t.rows.data(), static_cast<uint32_t>(t.rows.size())), "rcull_set_mesh_table");
    }
    std::vector<float> pos_scale, rot;  // staging columns (all ranks, entity order)
    std::vector<uint32_t> mesh;

  private:
    uint32_t n_gpus_ = 0;
    rcull_mgpu* g_ = nullptr;
};

// Contiguous range of rank r: [n * r / world, n * (r + 1) / world).
inline uint64_t shard_begin(uint64_t n, uint32_t world, uint32_t rank) { return n * rank / world; }

// gpu_transform_cull over range shards: rank r gets entities [shard_begin(r), shard_begin(r + 1)) with
// draw_index_base = shard_begin(r); asynchronous -- mg.synchronize-style waits are the caller's (download_merged does).
inline void gpu_transform_cull_sharded(MultiGpuRes& mg, const Camera& camera, const std::vector<Light>& lights, const Position* position,
                                       const Rotation* rotation, const Scale* scale, const GltfMesh* gltf_mesh, uint64_t n_entities) {
    mg.pos_scale.resize(4 * n_entities);
    mg.rot.resize(4 * n_entities);
    mg.mesh.resize(n_entities);
    expect(rcull_pack_components(n_entities, reinterpret_cast<const float*>(position), reinterpret_cast<const float*>(rotation),
                                 reinterpret_cast<const float*>(scale), mg.pos_scale.data(), mg.rot.data()),
           "rcull_pack_components");
    for (uint64_t e = 0; e < n_entities; ++e) mg.mesh[e] = gltf_mesh[e].table_row;
    std::vector<rcull_view> views;
    views.push_back(camera.frustum_planes);
    for (const Light& l : lights) views.push_back(light_frustum(l));
    std::vector<uint64_t> n(mg.n_gpus());
    std::vector<uint32_t> base(mg.n_gpus());
    for (uint32_t r = 0; r < mg.n_gpus(); ++r) {
        const uint64_t lo = shard_begin(n_entities, mg.n_gpus(), r), hi = shard_begin(n_entities, mg.n_gpus(), r + 1);
        n[r] = hi - lo;
        base[r] = static_cast<uint32_t>(lo);
        expect(rcull_upload_entities(mg.ctx(r), 0, n[r], mg.pos_scale.data() + 4 * lo, mg.rot.data() + 4 * lo, mg.mesh.data() + lo, mg.stream(r)),
               "rcull_upload_entities");
    }
    if (rcull_mgpu_frame(mg.handle(), n.data(), base.data(), views.data(), static_cast<uint32_t>(views.size()), camera.position) != RCULL_OK)
        panic(std::string("rcull_mgpu_frame: ") + rcull_mgpu_last_error());
}

// The merged draw list of one view as one rank holds it (blocking).
struct MergedList {
    uint32_t count = 0;
    uint64_t index_total = 0;
    std::vector<rcull_draw_cmd> commands;
    std::vector<DrawIndex> visible;
};
inline MergedList download_merged(MultiGpuRes& mg, uint32_t rank, uint32_t view) {
    if (rcull_mgpu_synchronize(mg.handle()) != RCULL_OK) panic(std::string("rcull_mgpu_synchronize: ") + rcull_mgpu_last_error());
    MergedList out;
    expect(rcull_merge_download(mg.merge(rank), view, &out.count, &out.index_total, nullptr, nullptr, 0), "rcull_merge_download");
    out.commands.resize(out.count);
    out.visible.resize(out.count);
    if (out.count)
        expect(rcull_merge_download(mg.merge(rank), view, &out.count, &out.index_total, out.commands.data(), &out.visible[0].ix, out.count),
               "rcull_merge_download");
    return out;
}

// ---- read-back (tests, debug_aabb-style consumers; blocking) -----------------------------------
inline std::vector<ModelMatrix> download_model(RcullRes& rc, ImageIndex ix) {
    std::vector<ModelMatrix> out(rc.n_entities);
    if (!out.empty()) expect(rcull_download_model(rc.ctx(), ix.ix, 0, out.size(), out[0].m), "rcull_download_model");
    return out;
}
inline std::vector<AABB> download_aabb(RcullRes& rc, ImageIndex ix) {
    std::vector<AABB> out(rc.n_entities);
    if (!out.empty()) expect(rcull_download_aabb(rc.ctx(), ix.ix, 0, out.size(), out[0].mins), "rcull_download_aabb");
    return out;
}
inline std::vector<CoarseCulled> download_coarse_culled(RcullRes& rc, ImageIndex ix, uint32_t view) {
    std::vector<CoarseCulled> out(rc.n_entities);
    if (!out.empty()) expect(rcull_download_flags(rc.ctx(), ix.ix, view, 0, out.size(), &out[0].culled), "rcull_download_flags");
    return out;
}
inline std::vector<rcull_draw_cmd> download_commands(RcullRes& rc, ImageIndex ix, uint32_t view) {
    std::vector<rcull_draw_cmd> out(rc.counts[view]);
    if (!out.empty()) expect(rcull_download_cmds(rc.ctx(), ix.ix, view, 0, out.size(), out.data()), "rcull_download_cmds");
    return out;
}
inline std::vector<DrawIndex> download_visible(RcullRes& rc, ImageIndex ix, uint32_t view) {
    // per-entity mode: one entry per command; per-mesh instanced mode: one per instance (index_totals counts instances there)
    const uint64_t n = rc.config().output_mode == RCULL_PER_MESH_INSTANCED ? rc.index_totals[view] : rc.counts[view];
    std::vector<DrawIndex> out(n);
    if (!out.empty()) expect(rcull_download_visible(rc.ctx(), ix.ix, view, 0, out.size(), &out[0].ix), "rcull_download_visible");
    return out;
}


// After gpu_cull_triangles: IndirectCommandsCount, the compacted commands and the sum of the refined indexCount
// (CullPassData.culled_commands_count_buffer / culled_commands_buffer, cull_pipeline.rs:74-84).
struct TriangleStage {
    uint32_t count = 0;
    uint64_t kept_indices = 0;
    std::vector<rcull_draw_cmd> commands;
};
inline TriangleStage download_triangle_stage(RcullRes& rc) {
    TriangleStage t;
    expect(rcull_download_tri(rc.ctx(), &t.count, &t.kept_indices, nullptr, 0, nullptr, 0, 0), "rcull_download_tri");
    t.commands.resize(t.count);
    if (t.count) expect(rcull_download_tri(rc.ctx(), &t.count, nullptr, t.commands.data(), t.count, nullptr, 0, 0), "rcull_download_tri");
    return t;
}
// CullPassData.culled_index_buffer [first, first + n): a command's triples start at its firstIndex.
inline std::vector<uint32_t> download_culled_indices(RcullRes& rc, uint64_t first, uint64_t n) {
    std::vector<uint32_t> out(n);
    uint32_t count = 0;
    if (n) expect(rcull_download_tri(rc.ctx(), &count, nullptr, nullptr, 0, out.data(), first, n), "rcull_download_tri");
    return out;
}
struct AsInstance { unsigned char bytes[64]; };  // VkAccelerationStructureInstanceKHR
inline std::vector<AsInstance> download_as_instances(RcullRes& rc) {
    std::vector<AsInstance> out(rc.n_entities);
    if (!out.empty()) expect(rcull_download_as_instances(rc.ctx(), 0, out.size(), out.data()), "rcull_download_as_instances");
    return out;
}

}  // namespace ecs
}  // namespace rcull
#endif
