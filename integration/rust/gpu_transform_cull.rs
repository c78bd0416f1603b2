// gpu_transform_cull.rs -- the bevy system that replaces model_matrix_calculation (src/ecs.rs:52-64),
// aabb_calculation (:138-181), coarse_culling (cull_pipeline.rs:99-120), model_matrices_upload
// (renderer.rs:2266-2288) and cull_pass' emit loop + dispatches (cull_pipeline.rs:534-605) with calls into
// librcull.so.  Registered where those systems are (src/main.rs:780-839).  Source only: never compiled on
// the box that wrote it (no rustc there); include/rcull_systems.hpp is the same system in C++, which is
// what the tests drive.
use bevy_ecs::prelude::*;

use crate::{
    ecs::{
        components::{Deleting, Light, Position, Rotation, Scale},
        resources::Camera,
        systems::RuntimeConfiguration,
    },
    renderer::{
        rcull_ffi::*, GltfMesh, GltfMeshBaseColorTexture, ImageIndex,
    },
};

/// ResMut => exclusive access: one context = one GPU = externally serialised calls.
pub(crate) struct RcullRes {
    pub(crate) ctx: *mut RcullCtx,
    /// the context's own stream: every call of a frame is enqueued on it (rcull_context_stream)
    stream: *mut std::ffi::c_void,
    ps: Vec<f32>,
    rot: Vec<f32>,
    mesh: Vec<u32>,
    pub(crate) counts: [u32; RCULL_MAX_VIEWS],
    pub(crate) index_totals: [u64; RCULL_MAX_VIEWS],
}
unsafe impl Send for RcullRes {}
unsafe impl Sync for RcullRes {}

impl RcullRes {
    pub(crate) fn new(cfg: &RcullConfig) -> RcullRes {
        let mut ctx = std::ptr::null_mut();
        check(unsafe { rcull_create(cfg, &mut ctx) }); // aborts without a CUDA device: there is no CPU path
        let mut stream = std::ptr::null_mut();
        check(unsafe { rcull_context_stream(ctx, &mut stream) });
        RcullRes { ctx, stream, ps: vec![], rot: vec![], mesh: vec![], counts: [0; 8], index_totals: [0; 8] }
    }
    /// consolidate_mesh_buffers changed its offset maps (consolidate_mesh_buffers.rs:99-162)
    pub(crate) fn set_mesh_table(&mut self, rows: &[RcullMesh]) {
        check(unsafe { rcull_set_mesh_table(self.ctx, rows.as_ptr(), rows.len() as u32) });
    }
    /// explicit destroy, like the renderer's `.destroy(device)` convention (cull_pipeline.rs:276-286)
    pub(crate) fn destroy(&mut self) {
        unsafe { rcull_destroy(self.ctx) };
        self.ctx = std::ptr::null_mut();
    }
}

/// GltfMesh vertex-buffer handle (the key of ConsolidatedMeshBuffers.vertex_offsets, cull_pipeline.rs:540-543)
/// -> row of the table given to rcull_set_mesh_table.
pub(crate) struct MeshTableIds(pub(crate) hashbrown::HashMap<u64, u32>);

pub(crate) fn gpu_transform_cull(
    mut rc: ResMut<RcullRes>,
    image_index: Res<ImageIndex>,
    camera: Res<Camera>,
    runtime_config: Res<RuntimeConfiguration>,
    mesh_ids: Res<MeshTableIds>,
    lights: Query<(&Position, &Rotation), With<Light>>,
    query: Query<(&Position, &Rotation, &Scale, &GltfMesh), (With<GltfMeshBaseColorTexture>, Without<Deleting>)>,
) {
    let rc = &mut *rc;
    rc.ps.clear();
    rc.rot.clear();
    rc.mesh.clear();
    for (pos, rot, scale, mesh) in query.iter() {
        // iteration order == assign_draw_index order (src/ecs.rs:117-136): slot i is draw_index i
        rc.ps.extend_from_slice(&[pos.0.x, pos.0.y, pos.0.z, scale.0]);
        rc.rot.extend_from_slice(rot.0.as_vector().as_slice()); // [i, j, k, w]
        rc.mesh.push(mesh_ids.0[&mesh.vertex_buffer.handle.as_raw()]);
    }
    let n = rc.mesh.len() as u64;

    if runtime_config.freeze_culling {
        // cull_pass_bypass (cull_pipeline.rs:331-421): lists, count and culled indices of the last unfrozen frame stay
        check(unsafe { rcull_upload_entities(rc.ctx, 0, n, rc.ps.as_ptr(), rc.rot.as_ptr(), rc.mesh.as_ptr(), rc.stream) });
        check(unsafe { rcull_frame_frozen(rc.ctx, n, image_index.0, rc.stream) });
        return;
    }

    let mut views = vec![RcullView { planes: camera.frustum_planes.map(|p| [p.x, p.y, p.z, p.w]) }];
    for (pos, rot) in lights.iter() {
        // shadow views: the plane formula of src/ecs.rs:82-90 on each light's projection * view (shadow_mapping.rs:324-329)
        let (p, q) = ([pos.0.x, pos.0.y, pos.0.z], rot.0.as_vector());
        let mut v = RcullView { planes: [[0.0; 4]; 6] };
        let (mut proj, mut view, mut pv) = ([0f32; 16], [0f32; 16], [0f32; 16]);
        unsafe { rcull_light_view(p.as_ptr(), q.as_slice().as_ptr(), proj.as_mut_ptr(), view.as_mut_ptr(), pv.as_mut_ptr(), &mut v) };
        views.push(v);
    }
    let cam = [camera.position.x, camera.position.y, camera.position.z];
    check(unsafe {
        rcull_frame_host(rc.ctx, n, rc.ps.as_ptr(), rc.rot.as_ptr(), rc.mesh.as_ptr(), views.as_ptr(), views.len() as u32,
                         cam.as_ptr(), 0, image_index.0, rc.counts.as_mut_ptr(), rc.index_totals.as_mut_ptr())
    });

    // second half of cull_pass: per-triangle cull of the camera view's survivors + compact_draw_stream
    let pv = camera.projection * camera.view; // what camera_matrices_upload writes as `pv` (generate_work.comp:33); column-major
    let pv: [f32; 16] = pv.as_slice().try_into().unwrap();
    check(unsafe { rcull_tri_cull(rc.ctx, pv.as_ptr(), 0, image_index.0, rc.stream) });

    // Outputs stay on the device: rcull_get_outputs(..).model / cmds[k] / counts and rcull_get_tri_outputs(..)
    // .{cmds, count, indices} are what depth_only_pass / render_frame bind in place of ModelData.model_buffer and
    // CullPassData.{culled_commands_buffer, culled_commands_count_buffer, culled_index_buffer}
    // (cmd_draw_indexed_indirect_count(cmds, count, max, stride 20), depth_pass.rs:183-192).  Sharing those
    // allocations with Vulkan (external memory) is outside this repository.
}
