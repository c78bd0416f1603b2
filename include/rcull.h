/*
 * rcull.h -- C ABI of librcull.so: B200 (sm_100a) transform -> AABB -> frustum
 * cull -> prefix-sum compaction for farnoy/renderer's per-frame hot path.
 *
 * The reference has no FFI for this stage; its "operator API" is the set of
 * bevy systems below.  A replacement system makes ONE rcull_frame call per
 * frame; each entry point cites what it stands in for (paths relative to the
 * reference tree):
 *   model_matrix_calculation   src/ecs.rs:52-64
 *   aabb_calculation           src/ecs.rs:138-181
 *   coarse_culling             src/renderer/systems/cull_pipeline.rs:99-120
 *   model_matrices_upload      src/renderer.rs:2266-2288
 *   cull_pass (emit loop)      src/renderer/systems/cull_pipeline.rs:534-577
 *   compact_draw_stream.comp   src/shaders/compact_draw_stream.comp:34-63
 *   pick_lod                   src/renderer/helpers.rs:3-11
 *
 * Conventions: every call returns 0 on success or a negative rcull_status and
 * records a message for rcull_last_error() (thread-local).  The reference
 * panics (panic = "abort", Cargo.toml:133); a Rust shim would .expect() on the
 * code.  A context owns all device memory it hands out; one context = one GPU
 * = externally serialised calls (a system holding ResMut<Ctx>).  There is no
 * CPU fallback: without a CUDA device rcull_create fails.
 */
#ifndef RCULL_H
#define RCULL_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RCULL_MAX_VIEWS 8
#define RCULL_MAX_LODS 6 /* scene_loader.rs:739-753: lod 0 + up to 5 simplified */

typedef enum {
    RCULL_OK = 0,
    RCULL_ERR_INVALID = -1,  /* bad argument / capacity exceeded */
    RCULL_ERR_CUDA = -2,     /* a CUDA runtime call failed */
    RCULL_ERR_NO_DEVICE = -3 /* no sm_100 device visible */
} rcull_status;

typedef enum {
    RCULL_PER_ENTITY = 0,        /* one command per visible entity (cull_pass) */
    RCULL_PER_MESH_INSTANCED = 1 /* one command per non-empty (mesh, lod) slot */
} rcull_output_mode;

typedef struct rcull_ctx rcull_ctx;

/* Replaces the fixed capacities of generate_work.comp:26,37 (4096 / 2400). */
typedef struct {
    int32_t device;             /* CUDA ordinal */
    uint32_t max_views;         /* 1..RCULL_MAX_VIEWS: camera + shadow views */
    uint64_t capacity_entities; /* per frame call */
    uint64_t capacity_visible;  /* per view draw-list capacity; 0 = capacity_entities */
    uint32_t max_meshes;
    uint32_t output_mode;       /* rcull_output_mode */
    uint32_t n_buffers;         /* 1..3 output sets; DoubleBuffered (double_buffered.rs:6-39) */
    uint32_t materialize_aabb;  /* write the AABB component (debug_aabb consumer) */
    uint32_t materialize_flags; /* write CoarseCulled per view */
    float lod_distance;         /* helpers.rs:6 uses 10.0 */
    uint64_t capacity_tri_indices; /* CulledIndexBuffer capacity in indices (generate_work.comp:40-42 has
                                      20 M triples); 0 = no per-triangle stage */
} rcull_config;

/* GltfMesh (src/renderer.rs:117-126) + consolidate_mesh_buffers offsets
 * (consolidate_mesh_buffers.rs:100-159).  80 bytes. */
typedef struct {
    float aabb_min[3];
    float aabb_max[3];
    uint32_t n_lods; /* >= 1 */
    struct { uint32_t first_index, index_len; } lod[RCULL_MAX_LODS];
    int32_t vertex_offset;
} rcull_mesh;

/* Camera.frustum_planes (camera_controller.rs:15-16): left,right,bottom,top,near,far. */
typedef struct {
    float planes[6][4];
} rcull_view;

/* VkDrawIndexedIndirectCommand, stride 20 (generate_work.comp:9-15). */
typedef struct {
    uint32_t indexCount;
    uint32_t instanceCount;
    uint32_t firstIndex;
    int32_t vertexOffset;
    uint32_t firstInstance;
} rcull_draw_cmd;

/* Device pointers of one output set.  model is shared by all views. */
typedef struct {
    float *model;                              /* mat4 column-major [n]   (ModelData.model_buffer) */
    float *aabb;                               /* mins,maxs [n][6] or NULL */
    uint8_t *flags[RCULL_MAX_VIEWS];           /* CoarseCulled [n] per view or NULL */
    rcull_draw_cmd *cmds[RCULL_MAX_VIEWS];     /* IndirectCommandsBuffer, dense from 0 */
    uint32_t *visible[RCULL_MAX_VIEWS];        /* draw_index of each survivor */
    uint32_t *counts;                          /* [max_views]  IndirectCommandsCount */
    uint64_t *index_totals;                    /* [max_views]  sum of indexCount */
} rcull_outputs;

const char *rcull_last_error(void);
/* Library/ABI version: major<<16 | minor. */
uint32_t rcull_version(void);

/* Allocates every device buffer once (replaces StaticBuffer::new / VMA,
 * cull_pipeline.rs:186-243, renderer.rs:1225-1244). */
int rcull_create(const rcull_config *cfg, rcull_ctx **out);
/* Mirrors the explicit .destroy(device) convention (cull_pipeline.rs:276-286). */
void rcull_destroy(rcull_ctx *ctx);

/* Mesh table = what GltfMesh + ConsolidatedMeshBuffers give cull_pass
 * (cull_pipeline.rs:540-548).  Host pointer; copied synchronously. */
int rcull_set_mesh_table(rcull_ctx *ctx, const rcull_mesh *meshes, uint32_t n);

/* SoA component columns: pos_scale = (Position.xyz, Scale), rot = Rotation
 * (i,j,k,w), mesh = index into the mesh table; entity slot i IS draw_index i
 * (assign_draw_index, src/ecs.rs:117-136).  Host pointers; async on stream
 * (0 = default stream) when the host memory is pinned.
 * Domain: finite components whose model matrix is finite (zero, negative,
 * denormal and huge scales, non-unit quaternions are all fine and bit-exact).
 * A NaN / inf component only changes that entity's own matrix, box and
 * visibility, and not necessarily the way the reference's generic 4x4
 * products would (inf * 0 terms, src/ecs.rs:60-62). */
int rcull_upload_entities(rcull_ctx *ctx, uint64_t first, uint64_t n, const float *pos_scale,
                          const float *rot, const uint32_t *mesh, void *stream);
/* The context's own stream (created non-blocking): rcull_frame_host uploads and launches on it.  A
 * host that mixes rcull_frame_host with the asynchronous entry points (rcull_tri_cull,
 * rcull_as_instances, rcull_frame_frozen, rcull_upload_entities) passes this stream to them, so that
 * all per-frame work of a context is ordered on ONE stream -- frame i + 1's upload must not overtake
 * frame i's triangle stage (the render graph orders cull_pass after the uploads the same way,
 * src/renderer.rs:2323-2391). */
int rcull_context_stream(rcull_ctx *ctx, void **stream);
/* Device pointers of the input columns, for callers that fill them in place. */
int rcull_entity_ptrs(rcull_ctx *ctx, float **pos_scale, float **rot, uint32_t **mesh);
/* AoS bevy components -> the SoA columns above (host side, no device work):
 * Position [n][3], Rotation [n][4], Scale [n]. */
int rcull_pack_components(uint64_t n, const float *position, const float *rotation, const float *scale,
                          float *pos_scale, float *rot);

/* One frame: model_matrix_calculation + aabb_calculation + coarse_culling per
 * view + model_matrices_upload + cull_pass emit/prefix-sum + compact_draw_stream,
 * fused.  views[0] is the camera; cam_pos feeds pick_lod.  draw_index_base is
 * added to firstInstance / visible indices (range sharding); buffer_ix plays
 * ImageIndex.  Asynchronous on stream. */
int rcull_frame(rcull_ctx *ctx, uint64_t n, const rcull_view *views, uint32_t n_views,
                const float cam_pos[3], uint32_t draw_index_base, uint32_t buffer_ix, void *stream);

/* A frame with RuntimeConfiguration.freeze_culling on (cull_pass_bypass, cull_pipeline.rs:331-421):
 * the reference keeps the draw lists, count and culled index buffer of the last unfrozen frame
 * (it copies the buffers onto themselves) while model_matrix_calculation and
 * model_matrices_upload still run.  Writes the model matrices of entities [0, n) into output set
 * buffer_ix and nothing else.  CullPassData is single-buffered in the reference and only ModelData
 * is DoubleBuffered (cull_pipeline.rs:70-84): after this call rcull_get_outputs(buffer_ix), the
 * list / count downloads and rcull_tri_cull on buffer_ix resolve to the LISTS of the output set the
 * last rcull_frame wrote (whatever its buffer_ix was) and to the MODEL matrices of buffer_ix. */
int rcull_frame_frozen(rcull_ctx *ctx, uint64_t n, uint32_t buffer_ix, void *stream);

/* Host-buffer convenience used by the end-to-end path: upload the columns,
 * run the frame, read back counts and index totals; blocks until done. */
int rcull_frame_host(rcull_ctx *ctx, uint64_t n, const float *pos_scale, const float *rot,
                     const uint32_t *mesh, const rcull_view *views, uint32_t n_views,
                     const float cam_pos[3], uint32_t draw_index_base, uint32_t buffer_ix,
                     uint32_t *counts_out, uint64_t *index_totals_out);

/* rcull_frame_host + the draw lists back on the host: after the frame, cmds_out[k][0 .. counts[k]) and
 * visible_out[k][0 .. counts[k]) of every view k < n_views (per-mesh mode: index_totals[k] visible
 * entries).  Each destination holds `capacity` entries (host pointers, pinned for full PCIe speed; a
 * NULL array or NULL entry skips that list).  The reference's consumers are GPU draws
 * (cmd_draw_indexed_indirect_count, depth_pass.rs:183-192), so lists normally stay resident: this is the
 * "results back" variant of the end-to-end call.  Blocks until the copies are done. */
int rcull_frame_host_lists(rcull_ctx *ctx, uint64_t n, const float *pos_scale, const float *rot,
                           const uint32_t *mesh, const rcull_view *views, uint32_t n_views,
                           const float cam_pos[3], uint32_t draw_index_base, uint32_t buffer_ix,
                           uint32_t *counts_out, uint64_t *index_totals_out,
                           rcull_draw_cmd *const *cmds_out, uint32_t *const *visible_out, uint64_t capacity);

int rcull_get_outputs(rcull_ctx *ctx, uint32_t buffer_ix, rcull_outputs *out);

/* Blocking device->host copies for tests and the harness. */
int rcull_download_counts(rcull_ctx *ctx, uint32_t buffer_ix, uint32_t *counts, uint64_t *index_totals);
int rcull_download_model(rcull_ctx *ctx, uint32_t buffer_ix, uint64_t first, uint64_t n, float *model);
int rcull_download_aabb(rcull_ctx *ctx, uint32_t buffer_ix, uint64_t first, uint64_t n, float *aabb);
int rcull_download_flags(rcull_ctx *ctx, uint32_t buffer_ix, uint32_t view, uint64_t first, uint64_t n, uint8_t *flags);
int rcull_download_cmds(rcull_ctx *ctx, uint32_t buffer_ix, uint32_t view, uint64_t first, uint64_t n, rcull_draw_cmd *cmds);
int rcull_download_visible(rcull_ctx *ctx, uint32_t buffer_ix, uint32_t view, uint64_t first, uint64_t n, uint32_t *visible);

/* RCULL_PER_MESH_INSTANCED only: the per-slot survivor counts of one view as the last rcull_frame
 * left them -- slot = mesh * RCULL_MAX_LODS + lod, n_slots = meshes in the table * RCULL_MAX_LODS,
 * every slot's instanceCount, empty slots included.  This is what a range-shard merge of the
 * per-mesh lists all_gathers (renderer_b200/shard.py plan_merge_per_mesh).  The scratch
 * is shared by the output sets: read it before the next rcull_frame. */
int rcull_get_slot_counts(rcull_ctx *ctx, uint32_t view, uint32_t **counts, uint32_t *n_slots);
/* Blocking copy of counts[0 .. n_slots) for tests and host-side merges. */
int rcull_download_slot_counts(rcull_ctx *ctx, uint32_t view, uint32_t *counts, uint32_t n_slots);

/* ---- Stage 2: per-triangle cull + command compaction -------------
 * Consolidated position / index buffers (ConsolidatedMeshBuffers.position_buffer / index_buffer,
 * consolidate_mesh_buffers.rs:26-35; vec3 and u32, scalar layout).  Host pointers, copied
 * synchronously.  rcull_mesh.lod[].first_index and .vertex_offset index into them. */
int rcull_set_geometry(rcull_ctx *ctx, const float *positions, uint64_t n_vertices,
                       const uint32_t *indices, uint64_t n_indices);
/* One call replaces cull_pass' dispatch loop over generate_work.comp (cull_pipeline.rs:534-577,
 * one vkCmdDispatch per visible entity) and the compact_draw_stream dispatch (:589-605): every
 * triangle of every survivor of the CAMERA view (view 0 of the last rcull_frame on buffer_ix;
 * the reference culls triangles against the camera only) is tested with
 * pv * (model * vec4(v, 1)) -- back-face determinant and the four NDC side tests of
 * generate_work.comp:139-155 -- kept index triples are written to the culled index buffer at
 * the command's firstIndex in ascending triangle order, indexCount becomes 3 * kept, and the
 * commands with indexCount > 0 are packed to the front of the triangle-stage command list.
 * pv = projection * view, column-major (CameraMatrices.pv, generate_work.comp:33).
 * draw_index_base must be the value given to rcull_frame.  Asynchronous on stream. */
int rcull_tri_cull(rcull_ctx *ctx, const float pv[16], uint32_t draw_index_base, uint32_t buffer_ix, void *stream);
typedef struct {
    rcull_draw_cmd *cmds;   /* compacted IndirectCommandsBuffer after stage 2 */
    uint32_t *count;        /* IndirectCommandsCount */
    uint32_t *indices;      /* CulledIndexBuffer: triples at [firstIndex, firstIndex + indexCount) */
    uint64_t *kept_indices; /* sum of the refined indexCount */
} rcull_tri_outputs;
int rcull_get_tri_outputs(rcull_ctx *ctx, rcull_tri_outputs *out);
/* Blocking copies for tests: count and kept_indices always; cmds[0..n_cmds) and
 * indices[first_index .. first_index + n_indices) when the pointers are not NULL. */
int rcull_download_tri(rcull_ctx *ctx, uint32_t *count, uint64_t *kept_indices, rcull_draw_cmd *cmds, uint64_t n_cmds,
                       uint32_t *indices, uint64_t first_index, uint64_t n_indices);

/* ---- Acceleration-structure instance records ---------------------------
 * The data-parallel half of build_acceleration_structures (acceleration_strucures.rs:419-445):
 * one 64-byte VkAccelerationStructureInstanceKHR per entity -- rows 0..2 of the model matrix
 * row-major, instance_custom_index = draw_index | mask 0xFF << 24, SBT offset 0 | flags
 * TRIANGLE_FACING_CULL_DISABLE << 24, BLAS device address of the entity's mesh.  The BLAS/TLAS
 * builds themselves are Vulkan ray-tracing calls and stay with the caller.
 * blas_addresses: one u64 per mesh of the mesh table (host pointer, copied synchronously). */
int rcull_set_blas_table(rcull_ctx *ctx, const uint64_t *blas_addresses, uint32_t n);
/* Packs records [0, n) from the model buffer the last rcull_frame wrote into output set
 * buffer_ix; asynchronous on stream. */
int rcull_as_instances(rcull_ctx *ctx, uint64_t n, uint32_t draw_index_base, uint32_t buffer_ix, void *stream);
/* Device pointer of the instance buffer (capacity_entities records of 64 bytes). */
int rcull_get_as_instances(rcull_ctx *ctx, void **instances);
int rcull_download_as_instances(rcull_ctx *ctx, uint64_t first, uint64_t n, void *dst);

/* Range-shard merge step: writes this rank's draw list of one view
 * into dst_cmds/dst_visible (device pointers into the MERGED lists, already
 * offset to this rank's slot), adding index_prefix to every firstIndex.  The
 * collective that moves the slices is the caller's (NCCL allgatherv). */
int rcull_rebase_into(rcull_ctx *ctx, uint32_t buffer_ix, uint32_t view, uint64_t index_prefix,
                      rcull_draw_cmd *dst_cmds, uint32_t *dst_visible, void *stream);

/* ---- Range-shard merge over peer memory (NVLink peer / NVLS multicast stores) ------
 * The entity set is sharded by contiguous range, one rank per GPU; the merged draw lists (every
 * rank's survivors, ascending global draw_index, byte-identical to the single-GPU lists) land on
 * EVERY rank.  What crosses NVLink is an 8-byte record per survivor {draw_index, slot = mesh * 6 +
 * lod}, a third of the 24 bytes of a command + visible index; the frame kernel sends it itself, at
 * emission time, with one store per peer (or one multimem.st through an NVLS multicast mapping)
 * into the part of every rank's WINDOW that belongs to the sender -- the transfer is spread over
 * the whole frame and there is no exchange kernel and no collective call.  rcull_merge_expand then
 * rebuilds the commands on the receiving GPU: indexCount / vertexOffset from the mesh table,
 * firstIndex as the running sum over the merged order (cull_pipeline.rs:558 over all ranks),
 * firstInstance = draw_index (generate_work.comp:63-66).  In RCULL_PER_MESH_INSTANCED mode the
 * expand merges the ranks' (slot, draw_index) pairs in draw order and runs the bucketing passes on
 * them: one command per non-empty slot with instanceCount summed over ranks.
 *
 * Protocol: after rcull_merge_create every rcull_frame on the context is a MERGED frame (the rank's
 * own per-rank lists are not written; every rank must issue the same frames in the same order).
 * Each merged frame is followed -- at once, or one frame later to hide the skew between ranks --
 * by one rcull_merge_expand; at most two merged frames may be in flight.  The expand kernel waits
 * for every rank's records of that frame; a frame kernel waits until every rank has expanded the
 * frame before the previous one (the record regions are double-buffered).  A single process that
 * drives several ranks on ONE GPU must therefore complete all ranks' frames before any expand. */
typedef struct rcull_merge rcull_merge;
#define RCULL_MAX_RANKS 8
#define RCULL_IPC_HANDLE_BYTES 64
/* capacity: commands one merged list can hold (per view); capacity_per_rank: survivors of one
 * view one rank can send (0 = capacity).  The window -- record regions of every source rank,
 * double-buffered, plus a mailbox -- is allocated by the library (cudaMalloc, exportable with
 * rcull_merge_ipc_handle) or supplied by the caller (rcull_merge_create_on: e.g. symmetric
 * memory that every rank has mapped, rcull_merge_window_bytes bytes). */
int rcull_merge_window_bytes(rcull_ctx *ctx, uint32_t world, uint64_t capacity_per_rank, uint64_t *bytes);
int rcull_merge_create(rcull_ctx *ctx, uint32_t world, uint32_t rank, uint64_t capacity,
                       uint64_t capacity_per_rank, rcull_merge **out);
int rcull_merge_create_on(rcull_ctx *ctx, uint32_t world, uint32_t rank, uint64_t capacity,
                          uint64_t capacity_per_rank, void *window, rcull_merge **out);
void rcull_merge_destroy(rcull_merge *m);
/* cudaIpcMemHandle_t of this rank's window, to be sent to the other processes. */
int rcull_merge_ipc_handle(rcull_merge *m, void *handle_out);
/* Maps a peer process' window (cudaIpcOpenMemHandle). */
int rcull_merge_open_peer(rcull_merge *m, uint32_t peer, const void *handle);
/* Same-process peers (single-process multi-GPU with peer access enabled, tests) or windows mapped
 * by the caller: the peer's window address itself. */
int rcull_merge_set_peer_ptr(rcull_merge *m, uint32_t peer, void *window);
int rcull_merge_window(rcull_merge *m, void **window);
/* NVLS multicast mapping of all ranks' windows: records then leave the GPU ONCE and the NVSwitch
 * replicates them (multimem.st), instead of one store per peer. */
int rcull_merge_set_multicast_ptr(rcull_merge *m, void *multicast_window);
/* expand_sms > 0: split the chip between the frame kernel and the expand.  The frame kernel's
 * persistent grid leaves expand_sms SMs free and rcull_merge_expand launches exactly that many CTAs,
 * each of which takes a whole SM; issued on a SECOND stream after frame i (event), the expand of
 * frame i then runs beside frame i + 1 instead of between frames.  0 (default): the expand is an
 * ordinary full-chip kernel on the stream it is given. */
int rcull_merge_set_overlap(rcull_merge *m, uint32_t expand_sms);
/* How a rank's records reach the peers.  0 (default): the frame kernel stores every record into every rank's window as
 * it emits (peer stores, or one multimem.st through the multicast mapping).  1: the COPY ENGINES move them -- the frame
 * kernel keeps its records, checkpoints, counts and tag in its own window (no store leaves the GPU from an SM, so the
 * frame runs at its single-GPU speed) and rcull_merge_send(m, stream) queues, on a stream of the caller's choice, one
 * peer copy of the rank's region per peer and the tag behind it.  The library orders the streams itself: the send waits
 * for the frame it belongs to (event), the frame two epochs later waits for the send that read the region it
 * overwrites.  Per merged frame: rcull_frame, rcull_merge_send, rcull_merge_expand (the expand may be queued behind
 * the next frame; the records then travel while that frame runs).  No kernel is involved in the transfer, so nothing
 * a spinning expand could be waiting for can be held up behind another kernel.  Callable only between merged frames. */
int rcull_merge_set_transport(rcull_merge *m, uint32_t transport);
int rcull_merge_send(rcull_merge *m, void *stream);
/* Copy-engine transport: how many records of each view a send copies (a host cannot know a frame's survivor counts when
 * it queues the copies; without a bound the whole capacity_per_rank of every view travels).  records_per_view[k] is an
 * ESTIMATE, rounded up to whole expand tiles of 256 -- e.g. the last known count plus a margin.  It is never a
 * correctness matter: records past the bound stay in the sender's window and every receiver's expand reads them from
 * there over NVLink.  An estimate of 0, or one at or above capacity_per_rank, leaves that view unbounded (its whole
 * capacity is copied); n_views = 0 removes the bounds.  With bounds a send takes one peer copy per view instead of one
 * per slot, and a peer copy has a fixed cost of ~13 us: worth it only when the capacities are far above the survivor
 * counts.  Callable only between merged frames. */
int rcull_merge_set_send_bound(rcull_merge *m, const uint32_t *records_per_view, uint32_t n_views);
/* Expands the oldest merged frame that has not been expanded yet into this rank's merged lists;
 * asynchronous on stream.  After it, rcull_merge_lists names the results. */
int rcull_merge_expand(rcull_merge *m, void *stream);
/* Device pointers of the merged lists of one view and of the merged counts / index totals
 * (per-mesh mode: commands per non-empty slot, bucketed visible list, counts = commands,
 * index totals = instances). */
int rcull_merge_lists(rcull_merge *m, uint32_t view, rcull_draw_cmd **cmds, uint32_t **visible,
                      uint32_t **counts, uint64_t **index_totals);

/* Blocking copy for tests and hosts without CUDA code of their own: the merged count and index total of
 * one view, and (pointers not NULL) the first n commands / visible entries (per-mesh mode: n commands and
 * n visible entries as far as each list goes). */
int rcull_merge_download(rcull_merge *m, uint32_t view, uint32_t *count, uint64_t *index_total,
                         rcull_draw_cmd *cmds, uint32_t *visible, uint64_t n);

/* ---- Single-process multi-GPU bootstrap (rcull_mgpu.cpp) ---------------------------------------
 * The reference's caller is ONE process stepping one schedule per frame (src/main.rs:907-910).
 * rcull_mgpu_create stands up n_gpus ranks for such a host without Python, torch or NCCL: one context
 * (cfg with device = devices[r], or r when devices is NULL) and one merge per GPU, every window
 * mapped into every device (peer access), and -- use_multicast != 0 and the driver offers it -- the
 * windows allocated with the virtual-memory API and bound to an NVLS multicast object
 * (cuMulticastCreate), whose mapping becomes rcull_merge_set_multicast_ptr.  Falls back to plain peer
 * pointers; rcull_mgpu_transport says which path is live.  Multi-PROCESS hosts exchange
 * rcull_merge_ipc_handle blobs over their own transport instead (renderer_b200/shard.py does it with
 * torch.distributed). */
typedef struct rcull_mgpu rcull_mgpu;
int rcull_mgpu_create(const rcull_config *cfg, uint32_t n_gpus, const int32_t *devices, uint64_t capacity,
                      uint64_t capacity_per_rank, uint32_t use_multicast, rcull_mgpu **out);
void rcull_mgpu_destroy(rcull_mgpu *g);
/* The context, merge and stream of one rank (upload its shard with rcull_upload_entities, read its
 * merged lists with rcull_merge_lists). */
int rcull_mgpu_rank(rcull_mgpu *g, uint32_t rank, rcull_ctx **ctx, rcull_merge **merge, void **stream);
const char *rcull_mgpu_transport(rcull_mgpu *g);
const char *rcull_mgpu_last_error(void);
/* One merged frame on every rank: rank r culls entities [0, n[r]) of its shard with
 * draw_index_base[r], then every rank expands; asynchronous (each rank's own stream).  Every rank's
 * frame is enqueued before any rank's expand. */
int rcull_mgpu_frame(rcull_mgpu *g, const uint64_t *n, const uint32_t *draw_index_base, const rcull_view *views,
                     uint32_t n_views, const float cam_pos[3]);
int rcull_mgpu_synchronize(rcull_mgpu *g);

/* rcull_frame for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that
 * forbid cudaMalloc.  It needs no rcull_ctx: the caller passes every input and output as device memory,
 * and the kernel's scratch (tile tickets, the look-back descriptors of the prefix scan, the device form of
 * the mesh table) lives in a file-scope `static __device__` array of RCULL_NOALLOC_SCRATCH_BYTES that every
 * GPU which loads librcull holds until the process exits.  The call issues no cudaMalloc / cudaFree and no
 * allocating library routine; it is asynchronous on `stream` (per-entity output mode, no optional
 * AABB / flag outputs, same kernel and bit-identical results as rcull_frame).
 *   limits   n <= RCULL_NOALLOC_MAX_ENTITIES, n_meshes <= RCULL_NOALLOC_MAX_MESHES, n_views <= RCULL_MAX_VIEWS;
 *            anything larger is rejected with RCULL_ERR_INVALID and a message naming the limit.
 *   sharing  every call on a device uses the same array and re-initialises the part it needs on `stream`:
 *            calls that can overlap on the GPU (different streams without an ordering between them) race.
 *   sizes    pos_scale / rot / mesh are read in whole tiles: they must be READABLE up to
 *            RCULL_NOALLOC_PADDED(n) entities (contents past n are ignored), 16-byte aligned.
 *            model holds n rounded up to 2 matrices, 16-byte aligned.  cmds[k] / visible[k] hold
 *            capacity_visible entries for every k < n_views (survivors past it are counted, not stored,
 *            as in rcull_frame).  counts / index_totals hold RCULL_MAX_VIEWS entries (the first n_views
 *            are written).  error is one word: 0, or nonzero when a mesh id was outside the table. */
#define RCULL_NOALLOC_TILE 1920u
#define RCULL_NOALLOC_MAX_ENTITIES (1u << 20)
#define RCULL_NOALLOC_MAX_MESHES 256u
#define RCULL_NOALLOC_SCRATCH_BYTES (128u * 1024u)
#define RCULL_NOALLOC_PADDED(n) (((uint64_t)(n) + RCULL_NOALLOC_TILE - 1) / RCULL_NOALLOC_TILE * RCULL_NOALLOC_TILE)
typedef struct {
    const float *pos_scale;                /* in:  [RCULL_NOALLOC_PADDED(n)][4] */
    const float *rot;                      /* in:  [RCULL_NOALLOC_PADDED(n)][4] */
    const uint32_t *mesh;                  /* in:  [RCULL_NOALLOC_PADDED(n)] */
    float *model;                          /* out: mat4 column-major [n rounded up to 2] */
    rcull_draw_cmd *cmds[RCULL_MAX_VIEWS]; /* out: [capacity_visible] per view */
    uint32_t *visible[RCULL_MAX_VIEWS];    /* out: [capacity_visible] per view */
    uint32_t *counts;                      /* out: [RCULL_MAX_VIEWS] */
    uint64_t *index_totals;                /* out: [RCULL_MAX_VIEWS] */
    uint32_t *error;                       /* out: [1] */
    uint64_t capacity_visible;
} rcull_noalloc_io;
int rcull_frame_noalloc(int32_t device, uint64_t n, const rcull_noalloc_io *io, const rcull_mesh *meshes,
                        uint32_t n_meshes, const rcull_view *views, uint32_t n_views, const float cam_pos[3],
                        float lod_distance, uint32_t draw_index_base, void *stream);

/* Harness utility: fills the input columns on the device with the counter-based
 * synthetic scene (entity i of the global scene = first_global + slot), bit-
 * identical to renderer_b200/scene.py; avoids shipping 128 M entities from host. */
int rcull_synth_entities(rcull_ctx *ctx, uint64_t seed, uint64_t first_global, uint64_t n,
                         float half_box, uint32_t n_meshes, void *stream);

/* Number of kernel launches issued by this context so far (bench accounting). */
uint64_t rcull_launch_count(rcull_ctx *ctx);

/* Host-side A5 helpers (src/ecs.rs:66-91, shadow_mapping.rs:321-329):
 * column-major matrices; planes in rcull_view order. */
void rcull_frustum_planes(const float pv[16], rcull_view *out);
void rcull_project_camera(float aspect, const float pos[3], const float rot_ijkw[4],
                          float proj[16], float view[16], float pv[16], rcull_view *out);
void rcull_light_view(const float pos[3], const float rot_ijkw[4],
                      float proj[16], float view[16], float pv[16], rcull_view *out);

#ifdef __cplusplus
}
#endif
#endif
