// systems_driver.cpp -- test harness for include/rcull_systems.hpp: loads a scene file written by
// tests/test_zz_cpp_systems.py, runs the C++ systems (project_camera, gpu_transform_cull and, on request,
// a frozen frame and the acceleration-structure records) and writes what they produced to a result
// file.  The CHECK against the oracle happens in the Python test; nothing here links oracle/.
//
//   systems_driver planes <scene> <result>    host-only: camera + light frusta (no CUDA device needed)
//   systems_driver frame  <scene> <result>    the full stage through librcull.so (panics without a GPU)
//   systems_driver stage2 <scene> <result>    camera view only, then gpu_cull_triangles + gpu_as_instances
//   systems_driver mgpu<N> <scene> <result>   N = 1..8 GPUs of this process: range shards, record merge, every rank's merged lists
//
// Scene file (little endian): u64 n, u32 n_meshes, u32 n_lights, u32 draw_index_base, f32 aspect,
// f32 lod_distance, Camera position[3] rotation[4], Light[n_lights] (position[3], rotation[4]),
// rcull_mesh[n_meshes], Position[n], Rotation[n], Scale[n], GltfMesh[n].
// Result file: u32 n_views, u32 counts[8], u64 index_totals[8], rcull_view[n_views]; then for `frame`:
// ModelMatrix[n], AABB[n], per view: CoarseCulled[n], rcull_draw_cmd[count], DrawIndex[count];
// ModelMatrix[n] after a frozen frame over entities whose Scale was doubled, u32 counts[8] after it.
// `mgpu<N>` result: u32 n_views, u32 n_gpus, the transport string (u32 length + bytes), then per rank and view: u32 count,
// u64 index_total, rcull_draw_cmd[count], DrawIndex[count] (two frames are run; the second one's lists are written).
// `stage2` scenes continue with: u64 n_vertices, u64 n_indices, u64 capacity_tri_indices, f32 positions[3 * n_vertices],
// u32 indices[n_indices], u64 blas[n_meshes]; its result: u32 n_views (= 1), u32 counts[8], u64 index_totals[8],
// f32 pv[16], u32 tri_count, u64 kept_indices, rcull_draw_cmd[tri_count], u32 culled_indices[index_totals[0]],
// AsInstance[n].
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <vector>

#include "rcull_systems.hpp"

using namespace rcull::ecs;

namespace {

template <class T>
void get(std::ifstream& f, T* dst, size_t count = 1) {
    f.read(reinterpret_cast<char*>(dst), static_cast<std::streamsize>(sizeof(T) * count));
    if (!f) panic("scene file is truncated");
}
template <class T>
void put(std::ofstream& f, const T* src, size_t count = 1) {
    f.write(reinterpret_cast<const char*>(src), static_cast<std::streamsize>(sizeof(T) * count));
}

struct Scene {
    uint64_t n = 0;
    uint32_t n_meshes = 0, n_lights = 0, draw_index_base = 0;
    float aspect = 1.f, lod_distance = 10.f;
    Camera camera;
    std::vector<Light> lights;
    MeshTable meshes;
    std::vector<float> positions_xyz;  // stage2: consolidated position / index buffers, BLAS addresses
    std::vector<uint32_t> indices;
    std::vector<uint64_t> blas;
    uint64_t capacity_tri_indices = 0;
    std::vector<Position> position;
    std::vector<Rotation> rotation;
    std::vector<Scale> scale;
    std::vector<GltfMesh> gltf_mesh;
};

Scene load(const char* path, bool stage2) {
    std::ifstream f(path, std::ios::binary);
    if (!f) panic(std::string("cannot open ") + path);
    Scene s;
    get(f, &s.n);
    get(f, &s.n_meshes);
    get(f, &s.n_lights);
    get(f, &s.draw_index_base);
    get(f, &s.aspect);
    get(f, &s.lod_distance);
    get(f, s.camera.position, 3);
    get(f, s.camera.rotation, 4);
    s.lights.resize(s.n_lights);
    s.meshes.rows.resize(s.n_meshes);
    s.position.resize(s.n);
    s.rotation.resize(s.n);
    s.scale.resize(s.n);
    s.gltf_mesh.resize(s.n);
    if (s.n_lights) get(f, s.lights.data(), s.n_lights);
    if (s.n_meshes) get(f, s.meshes.rows.data(), s.n_meshes);
    if (s.n) {
        get(f, s.position.data(), s.n);
        get(f, s.rotation.data(), s.n);
        get(f, s.scale.data(), s.n);
        get(f, s.gltf_mesh.data(), s.n);
    }
    if (stage2) {
        uint64_t n_vertices = 0, n_indices = 0;
        get(f, &n_vertices);
        get(f, &n_indices);
        get(f, &s.capacity_tri_indices);
        s.positions_xyz.resize(3 * n_vertices);
        s.indices.resize(n_indices);
        s.blas.resize(s.n_meshes);
        if (n_vertices) get(f, s.positions_xyz.data(), s.positions_xyz.size());
        if (n_indices) get(f, s.indices.data(), s.indices.size());
        if (s.n_meshes) get(f, s.blas.data(), s.blas.size());
    }
    return s;
}

}  // namespace

int main(int argc, char** argv) {
    const bool mgpu = argc == 4 && !std::strncmp(argv[1], "mgpu", 4) && argv[1][4] >= '1' && argv[1][4] <= '8' && !argv[1][5];
    if (argc != 4 || (!mgpu && std::strcmp(argv[1], "planes") && std::strcmp(argv[1], "frame") && std::strcmp(argv[1], "stage2"))) {
        std::fprintf(stderr, "usage: %s planes|frame|stage2|mgpu<N> <scene> <result>\n", argv[0]);
        return 2;
    }
    const bool stage2 = !std::strcmp(argv[1], "stage2");
    const bool full = !std::strcmp(argv[1], "frame");
    Scene s = load(argv[2], stage2);
    project_camera(s.camera, s.aspect);  // the system that runs before the cull stage (src/ecs.rs:66-91)
    const uint32_t n_views = 1 + s.n_lights;

    std::ofstream out(argv[3], std::ios::binary);
    if (!out) panic(std::string("cannot create ") + argv[3]);

    if (mgpu) {
        const uint32_t n_gpus = static_cast<uint32_t>(argv[1][4] - '0');
        rcull_config cfg;
        std::memset(&cfg, 0, sizeof cfg);
        cfg.max_views = n_views;
        cfg.capacity_entities = s.n ? s.n : 1;  // any rank could own everything
        cfg.max_meshes = s.n_meshes;
        cfg.output_mode = RCULL_PER_ENTITY;
        cfg.n_buffers = 1;
        cfg.lod_distance = s.lod_distance;
        MultiGpuRes mg(cfg, n_gpus, s.n ? s.n : 1);  // panics without n_gpus CUDA devices
        mg.set_mesh_table(s.meshes);
        for (int frame = 0; frame < 2; ++frame)  // twice: both record regions, tags reused
            gpu_transform_cull_sharded(mg, s.camera, s.lights, s.position.data(), s.rotation.data(), s.scale.data(), s.gltf_mesh.data(), s.n);
        put(out, &n_views);
        put(out, &n_gpus);
        const std::string transport = rcull_mgpu_transport(mg.handle());
        const uint32_t tl = static_cast<uint32_t>(transport.size());
        put(out, &tl);
        put(out, transport.data(), tl);
        for (uint32_t r = 0; r < n_gpus; ++r)
            for (uint32_t v = 0; v < n_views; ++v) {
                const MergedList m = download_merged(mg, r, v);
                put(out, &m.count);
                put(out, &m.index_total);
                put(out, m.commands.data(), m.commands.size());
                put(out, m.visible.data(), m.visible.size());
            }
        mg.destroy();
        out.flush();
        return out.good() ? 0 : 1;
    }

    if (stage2) {
        rcull_config cfg;
        std::memset(&cfg, 0, sizeof cfg);
        cfg.max_views = 1;  // the reference culls triangles against the camera only
        cfg.capacity_entities = s.n ? s.n : 1;
        cfg.max_meshes = s.n_meshes;
        cfg.output_mode = RCULL_PER_ENTITY;
        cfg.n_buffers = 1;
        cfg.lod_distance = s.lod_distance;
        cfg.capacity_tri_indices = s.capacity_tri_indices;
        RcullRes rc(cfg);
        rc.set_mesh_table(s.meshes);
        rc.set_geometry(s.positions_xyz, s.indices);
        rc.set_blas_table(s.blas);
        const ImageIndex image_index{0};
        gpu_transform_cull(rc, image_index, s.camera, {}, RuntimeConfiguration{}, s.position.data(), s.rotation.data(),
                           s.scale.data(), s.gltf_mesh.data(), s.n, s.draw_index_base);
        gpu_cull_triangles(rc, image_index, s.camera, RuntimeConfiguration{}, s.draw_index_base);
        const TriangleStage t = download_triangle_stage(rc);
        const std::vector<uint32_t> culled = download_culled_indices(rc, 0, rc.index_totals[0]);
        gpu_as_instances(rc, image_index, s.draw_index_base);
        const std::vector<AsInstance> instances = download_as_instances(rc);
        put(out, &rc.n_views);
        put(out, rc.counts, RCULL_MAX_VIEWS);
        put(out, rc.index_totals, RCULL_MAX_VIEWS);
        put(out, s.camera.pv, 16);
        put(out, &t.count);
        put(out, &t.kept_indices);
        put(out, t.commands.data(), t.commands.size());
        put(out, culled.data(), culled.size());
        put(out, instances.data(), instances.size());
        rc.destroy();
        out.flush();
        return out.good() ? 0 : 1;
    }

    if (!full) {
        const uint32_t zero32[RCULL_MAX_VIEWS] = {};
        const uint64_t zero64[RCULL_MAX_VIEWS] = {};
        put(out, &n_views);
        put(out, zero32, RCULL_MAX_VIEWS);
        put(out, zero64, RCULL_MAX_VIEWS);
        put(out, &s.camera.frustum_planes);
        for (const Light& l : s.lights) {
            const rcull_view v = light_frustum(l);
            put(out, &v);
        }
        return out.good() ? 0 : 1;
    }

    rcull_config cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.device = 0;
    cfg.max_views = n_views;
    cfg.capacity_entities = s.n ? s.n : 1;
    cfg.max_meshes = s.n_meshes;
    cfg.output_mode = RCULL_PER_ENTITY;
    cfg.n_buffers = 2;  // DoubleBuffered: this run uses image index 1
    cfg.materialize_aabb = 1;
    cfg.materialize_flags = 1;
    cfg.lod_distance = s.lod_distance;
    RcullRes rc(cfg);  // panics (aborts) without a CUDA device: there is no CPU path
    rc.set_mesh_table(s.meshes);

    const ImageIndex image_index{1};
    gpu_transform_cull(rc, image_index, s.camera, s.lights, RuntimeConfiguration{}, s.position.data(), s.rotation.data(),
                       s.scale.data(), s.gltf_mesh.data(), s.n, s.draw_index_base);

    put(out, &rc.n_views);
    put(out, rc.counts, RCULL_MAX_VIEWS);
    put(out, rc.index_totals, RCULL_MAX_VIEWS);
    put(out, &s.camera.frustum_planes);
    for (const Light& l : s.lights) {
        const rcull_view v = light_frustum(l);
        put(out, &v);
    }
    const std::vector<ModelMatrix> model = download_model(rc, image_index);
    const std::vector<AABB> aabb = download_aabb(rc, image_index);
    put(out, model.data(), model.size());
    put(out, aabb.data(), aabb.size());
    for (uint32_t v = 0; v < rc.n_views; ++v) {
        const std::vector<CoarseCulled> culled = download_coarse_culled(rc, image_index, v);
        const std::vector<rcull_draw_cmd> cmds = download_commands(rc, image_index, v);
        const std::vector<DrawIndex> visible = download_visible(rc, image_index, v);
        put(out, culled.data(), culled.size());
        put(out, cmds.data(), cmds.size());
        put(out, visible.data(), visible.size());
    }

    // freeze_culling: the entities keep moving (here: every Scale doubles) but only the model
    // matrices follow; counts and lists stay those of the frame above (cull_pipeline.rs:331-421).
    for (Scale& sc : s.scale) sc.s *= 2.f;
    RuntimeConfiguration frozen_config;
    frozen_config.freeze_culling = true;
    gpu_transform_cull(rc, image_index, s.camera, s.lights, frozen_config, s.position.data(), s.rotation.data(), s.scale.data(),
                       s.gltf_mesh.data(), s.n, s.draw_index_base);
    gpu_cull_triangles(rc, image_index, s.camera, frozen_config, s.draw_index_base);  // nothing to launch
    const std::vector<ModelMatrix> frozen = download_model(rc, image_index);
    put(out, frozen.data(), frozen.size());
    uint32_t counts_after[RCULL_MAX_VIEWS] = {};
    uint64_t totals_after[RCULL_MAX_VIEWS] = {};
    expect(rcull_download_counts(rc.ctx(), image_index.ix, counts_after, totals_after), "rcull_download_counts");
    put(out, counts_after, RCULL_MAX_VIEWS);

    rc.destroy();
    out.flush();
    return out.good() ? 0 : 1;
}
