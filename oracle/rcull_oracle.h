/*
 * rcull_oracle.h -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * Scalar C restatement of farnoy/renderer's per-frame transform -> AABB ->
 * frustum cull -> draw-command emission path.  Only tests/, the smoke check in
 * __graft_entry__.py and bench.py's cpu_baseline / --impl reference legs may
 * load this library; renderer_b200 itself never does.
 *
 * PARITY UNPINNED: the reference (Rust, bevy_ecs 0.6) holds no test, golden
 * vector or fixture for this path, and neither rustc nor the crates that hold
 * the arithmetic are on this box:
 *   nalgebra 0.29.0, nalgebra-glm 0.15.0, ncollide3d 0.32.0, simba 0.6.0
 *   (Cargo.lock:1703,1719,1751,2653).
 * The operation order below restates those crates' published algorithms
 *  at the reference's own call sites, cited per function.
 *
 * Build: gcc -O2 -std=c11 -ffp-contract=off -fno-fast-math (see Makefile).
 */
#ifndef RCULL_ORACLE_H
#define RCULL_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORACLE_MAX_LODS 6

/* GltfMesh fields the path reads (src/renderer.rs:117-126) plus the offsets
 * consolidate_mesh_buffers assigns (consolidate_mesh_buffers.rs:100-159). */
typedef struct {
    float aabb_min[3];
    float aabb_max[3];
    uint32_t n_lods;
    struct { uint32_t first_index, index_len; } lod[ORACLE_MAX_LODS];
    int32_t vertex_offset;
} oracle_mesh;

/* VkDrawIndexedIndirectCommand (src/shaders/generate_work.comp:9-15). */
typedef struct {
    uint32_t indexCount, instanceCount, firstIndex;
    int32_t vertexOffset;
    uint32_t firstInstance;
} oracle_cmd;

/* A1: src/ecs.rs:61-62.  m is column-major mat4. */
void oracle_model_matrix(const float pos[3], const float quat_ijkw[4], float scale, float m[16]);
/* A2: src/ecs.rs:147-179.  Writes ncollide AABB {mins, maxs}. */
void oracle_aabb(const float m[16], const float mesh_min[3], const float mesh_max[3],
                 float mins[3], float maxs[3]);
/* A3: cull_pipeline.rs:108-118.  Returns CoarseCulled (1 = outside). */
int oracle_coarse_culled(const float mins[3], const float maxs[3], const float planes[6][4]);
/* A4: helpers.rs:3-11.  Returns the LOD index. */
uint32_t oracle_pick_lod(const float cam[3], const float pos[3], uint32_t n_lods, float lod_distance);
/* A5: planes from a column-major projection*view (src/ecs.rs:82-90). */
void oracle_frustum_planes(const float pv[16], float planes[6][4]);
/* A5: camera projection/view (src/ecs.rs:69-80); NOT bit-pinned (tanf, look_at). */
void oracle_project_camera(float aspect, const float pos[3], const float rot_ijkw[4],
                           float proj[16], float view[16], float pv[16], float planes[6][4]);
/* A5: light pv (shadow_mapping.rs:321-329) and the same plane formula. */
void oracle_light_view(const float pos[3], const float rot_ijkw[4],
                       float proj[16], float view[16], float pv[16], float planes[6][4]);
/* UnitQuaternion::look_at_lh(dir, up) as used for the lights (main.rs:369-381). */
void oracle_quat_look_at_lh(const float dir[3], const float up[3], float quat_ijkw[4]);
/* column-major 4x4 product, nalgebra gemv order. */
void oracle_mat4_mul(const float a[16], const float b[16], float out[16]);

/*
 * Whole frame, reference-shaped: three separate passes over AoS components
 * (A1, A2, A3 per view) and a serial emit loop (A4, A6) per view.
 * Inputs are the reference's components: Position (3 floats), Rotation
 * (i,j,k,w), Scale, mesh id per entity in draw_index order (A0).
 * Outputs (any may be NULL): model [16n], aabb [6n] (mins,maxs), flags
 * [n_views][n] (1 = culled), cmds/visible [n_views][n] dense from slot 0,
 * counts [n_views], index_totals [n_views].
 * n_threads > 1 splits the three passes by contiguous range (the emit stays
 * serial, as in the reference, cull_pipeline.rs:536).
 * Returns 0, or -1 on a mesh id out of range.
 */
int oracle_frame(const float *pos, const float *quat, const float *scale, const uint32_t *mesh_id,
                 uint64_t n, const oracle_mesh *meshes, uint32_t n_meshes,
                 const float *planes /* [n_views][6][4] */, uint32_t n_views,
                 const float cam_pos[3], float lod_distance, uint32_t draw_index_base,
                 float *model, float *aabb, uint8_t *flags,
                 oracle_cmd *cmds, uint32_t *visible, uint32_t *counts, uint64_t *index_totals,
                 int n_threads);

/* The same frame with a fused schedule (CPU baseline (ii)): one pass per entity, per-thread
 * counts, an exclusive prefix over the threads, a second pass that writes each thread's slice of the
 * lists.  Same arguments, bit-identical outputs; n_views <= 8. */
int oracle_frame_fused(const float *pos, const float *quat, const float *scale, const uint32_t *mesh_id,
                       uint64_t n, const oracle_mesh *meshes, uint32_t n_meshes,
                       const float *planes /* [n_views][6][4] */, uint32_t n_views,
                       const float cam_pos[3], float lod_distance, uint32_t draw_index_base,
                       float *model, float *aabb, uint8_t *flags,
                       oracle_cmd *cmds, uint32_t *visible, uint32_t *counts, uint64_t *index_totals,
                       int n_threads);

/* Per-mesh instanced emission (BASELINE config 4): one command per (mesh, lod)
 * slot with instanceCount = survivors, firstInstance = bucket start in the
 * bucketed visible list (stable by draw_index), firstIndex = the lod's own
 * first_index; empty slots dropped (compact_draw_stream.comp:41 predicate on
 * instanceCount).  flags as written by oracle_frame for one view. */
int oracle_emit_per_mesh(const float *pos, const uint32_t *mesh_id, const uint8_t *flags, uint64_t n,
                         const oracle_mesh *meshes, uint32_t n_meshes, const float cam_pos[3],
                         float lod_distance, uint32_t draw_index_base,
                         oracle_cmd *cmds, uint32_t *visible, uint32_t *count, uint64_t *n_instances);

/*
 * A8 + A7: per-triangle cull of one view's survivors and compaction of the
 * commands that kept a triangle (src/shaders/generate_work.comp:60-201,
 * src/shaders/compact_draw_stream.comp:34-63; dispatch loop cull_pipeline.rs:534-605).
 * For command c (as written by oracle_frame: indexCount = index_len, firstIndex =
 * indexOffsetInOutput, firstInstance = gltfIndex): triangle t reads
 * index_buffer[src_first_index[c] / 3 + t] (uvec3), vertices
 * vertex_buffer[vertexOffset + ix] (vec3), clip = pv * (model[gltfIndex] * vec4(v, 1))
 * (GLSL mat * vec = ((col0 * x + col1 * y) + col2 * z) + col3 * w, every product and
 * sum rounded, no contraction); culled iff determinant(mat3(c0.xyw, c1.xyw, c2.xyw)) > 0
 * (cofactors along the first column of GLSL's column-major mat3, glm order) or all
 * three ndc.x < -1 / > 1 / ndc.y < -1 / > 1 (ndc = xyz / w, IEEE division).  The area
 * test of :159-167 is computed and then disabled in the reference (:166) -- no effect.
 * Kept triples go to out_indices[(firstIndex + 3 * rank) ...] in ascending triangle
 * order (the reference's order inside a mesh is atomic-arrival order; this is one of
 * its legal outcomes), indexCount becomes 3 * kept.  Commands with indexCount > 0 are
 * packed to the front of out_cmds in order; *out_count = how many.
 * model is indexed by firstInstance - draw_index_base.  Returns 0, -1 on bad input.
 */
int oracle_tri_cull(const oracle_cmd *cmds, const uint32_t *src_first_index, uint32_t count,
                    const float *model, uint32_t draw_index_base, const float pv[16],
                    const float *positions, uint64_t n_vertices, const uint32_t *indices, uint64_t n_indices,
                    uint32_t *out_indices, uint64_t out_capacity,
                    oracle_cmd *out_cmds, uint32_t *out_count, uint64_t *kept_indices, int n_threads);

/* F4: VkAccelerationStructureInstanceKHR records (acceleration_strucures.rs:419-449): rows 0..2
 * of the column-major model matrix transposed into 12 row-major floats, custom index = draw_index
 * (24 bits) | mask 0xFF << 24, SBT offset 0 | TRIANGLE_FACING_CULL_DISABLE (1) << 24, BLAS address
 * of the entity's mesh.  out: 64 bytes per entity. */
void oracle_as_instances(const float *model, const uint32_t *mesh_id, const uint64_t *blas, uint64_t n,
                         uint32_t draw_index_base, void *out);

/* Deterministic scene generator twin (see renderer_b200/scene.py): entity i of
 * the global scene, counter-based so shards regenerate their slice. */
void oracle_synth_entities(uint64_t seed, uint64_t first, uint64_t n, float half_box, uint32_t n_meshes,
                           float *pos, float *quat, float *scale, uint32_t *mesh_id);

#ifdef __cplusplus
}
#endif
#endif
