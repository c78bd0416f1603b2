/*
 * rcull_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 * See rcull_oracle.h for the scope statement and the "parity unpinned" note.
 *
 * Every function restates one reference call site; the arithmetic of the
 * un-vendored crates (nalgebra 0.29.0 / nalgebra-glm 0.15.0 / ncollide3d
 * 0.32.0) is written out generically (full 4x4 products, 8 corners, 6 planes)
 * so that the CUDA path's algebraic shortcuts are checked against the long form.
 * Compile with -ffp-contract=off: Rust/LLVM never contracts a*b+c into an FMA.
 */
#include "rcull_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ---- nalgebra building blocks ------------------------------------------------ */

/* Matrix4 * Vector4 as nalgebra's gemv (blas.rs): y = (1*A[:,0])*x0, then
 * y = (1*A[:,k])*xk + 1*y for k = 1..3.  a is column-major. */
static void gemv4(const float a[16], const float x[4], float y[4]) {
    for (int i = 0; i < 4; i++) y[i] = (1.0f * a[0 * 4 + i]) * x[0];
    for (int k = 1; k < 4; k++)
        for (int i = 0; i < 4; i++) y[i] = (1.0f * a[k * 4 + i]) * x[k] + 1.0f * y[i];
}

/* Matrix4 * Matrix4: one gemv per column of the right operand. */
void oracle_mat4_mul(const float a[16], const float b[16], float out[16]) {
    float tmp[16];
    for (int j = 0; j < 4; j++) gemv4(a, b + 4 * j, tmp + 4 * j);
    memcpy(out, tmp, sizeof tmp);
}

static void mat4_identity(float m[16]) {
    memset(m, 0, 16 * sizeof(float));
    m[0] = m[5] = m[10] = m[15] = 1.0f;
}

/* UnitQuaternion::to_rotation_matrix -> to_homogeneous (nalgebra
 * geometry/quaternion.rs), storage order [i, j, k, w]. */
static void quat_to_homogeneous(const float q[4], float m[16]) {
    float i = q[0], j = q[1], k = q[2], w = q[3];
    float ww = w * w, ii = i * i, jj = j * j, kk = k * k;
    float ij = i * j * 2.0f, wk = w * k * 2.0f, wj = w * j * 2.0f;
    float ik = i * k * 2.0f, jk = j * k * 2.0f, wi = w * i * 2.0f;
    mat4_identity(m);
    /* row r, column c lives at m[c*4 + r] */
    m[0 * 4 + 0] = ww + ii - jj - kk; m[1 * 4 + 0] = ij - wk;           m[2 * 4 + 0] = wj + ik;
    m[0 * 4 + 1] = wk + ij;           m[1 * 4 + 1] = ww - ii + jj - kk; m[2 * 4 + 1] = jk - wi;
    m[0 * 4 + 2] = ik - wj;           m[1 * 4 + 2] = wi + jk;           m[2 * 4 + 2] = ww - ii - jj + kk;
}

/* dot for statically sized vectors (nalgebra base/blas.rs dotx unrolls):
 * dim 3: (a + b) + c;  dim 4: (a + c) + (b + d). */
static float dot3(const float x[3], const float y[3]) {
    float a = x[0] * y[0], b = x[1] * y[1], c = x[2] * y[2];
    return a + b + c;
}
static float dot4(const float x[4], const float y[4]) {
    float a = x[0] * y[0], b = x[1] * y[1], c = x[2] * y[2], d = x[3] * y[3];
    a += c;
    b += d;
    return a + b;
}

/* ---- A1: model_matrix_calculation (src/ecs.rs:61-62) -------------------------- */
void oracle_model_matrix(const float pos[3], const float quat[4], float scale, float m[16]) {
    float t[16], r[16], s[16], tr[16];
    mat4_identity(t); /* glm::translation(&pos.coords) */
    t[12] = pos[0]; t[13] = pos[1]; t[14] = pos[2];
    quat_to_homogeneous(quat, r); /* rot.to_homogeneous() */
    mat4_identity(s); /* glm::scaling(&Vec3::repeat(scale)) */
    s[0] = s[5] = s[10] = scale;
    oracle_mat4_mul(t, r, tr);
    oracle_mat4_mul(tr, s, m);
}

/* ---- A2: aabb_calculation (src/ecs.rs:147-179) -------------------------------- */
void oracle_aabb(const float m[16], const float mn[3], const float mx[3], float mins[3], float maxs[3]) {
    /* corner order exactly as listed at ecs.rs:151-159 */
    const float corners[8][3] = {
        {mn[0], mn[1], mn[2]}, {mx[0], mn[1], mn[2]}, {mn[0], mn[1], mx[2]}, {mx[0], mn[1], mx[2]},
        {mn[0], mx[1], mn[2]}, {mx[0], mx[1], mn[2]}, {mn[0], mx[1], mx[2]}, {mx[0], mx[1], mx[2]},
    };
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    float hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX}; /* f32::MIN */
    for (int c = 0; c < 8; c++) {
        float v[4] = {corners[c][0], corners[c][1], corners[c][2], 1.0f}; /* to_homogeneous */
        float p[4];
        gemv4(m, v, p);
        for (int a = 0; a < 3; a++) {
            float e = p[a] / p[3]; /* vertex.xyz() / vertex.w */
            lo[a] = fminf(lo[a], e); /* f32::min */
            hi[a] = fmaxf(hi[a], e);
        }
    }
    for (int a = 0; a < 3; a++) {
        float c = (hi[a] + lo[a]) / 2.0f;
        float h = (hi[a] - lo[a]) / 2.0f;
        /* ncollide3d AABB::from_half_extents: new(center - half, center + half) */
        mins[a] = c - h;
        maxs[a] = c + h;
    }
}

/* ---- A3: coarse_culling (cull_pipeline.rs:108-118) ---------------------------- */
int oracle_coarse_culled(const float mins[3], const float maxs[3], const float planes[6][4]) {
    float half[3], center[4];
    for (int a = 0; a < 3; a++) {
        half[a] = (maxs[a] - mins[a]) * 0.5f;   /* AABB::half_extents() */
        center[a] = (mins[a] + maxs[a]) * 0.5f; /* AABB::center() = na::center(mins, maxs) */
    }
    center[3] = 1.0f; /* to_homogeneous */
    int outside = 0;
    for (int p = 0; p < 6; p++) {
        float ab[3] = {fabsf(planes[p][0]), fabsf(planes[p][1]), fabsf(planes[p][2])};
        float e = dot3(half, ab);
        float s = dot4(planes[p], center);
        if (s - e > 0.0f) {
            outside = 1;
            break;
        }
    }
    return outside;
}

/* ---- A4: pick_lod (helpers.rs:3-11) ------------------------------------------- */
uint32_t oracle_pick_lod(const float cam[3], const float pos[3], uint32_t n_lods, float lod_distance) {
    float d[3] = {cam[0] - pos[0], cam[1] - pos[1], cam[2] - pos[2]};
    float n2 = 0.0f;
    n2 += dot3(d, d); /* norm_squared(): res = 0; res += col.dotc(col) */
    float dist = sqrtf(n2);
    return (dist > lod_distance && n_lods > 1) ? 1u : 0u;
}

/* ---- A5: per-view matrices and planes ----------------------------------------- */

/* planes = -(row3 +/- row_i) for i = 0,1,2 (src/ecs.rs:82-90); order
 * left, right, bottom, top, near, far (camera_controller.rs:15). */
void oracle_frustum_planes(const float pv[16], float planes[6][4]) {
    for (int i = 0; i < 3; i++)
        for (int c = 0; c < 4; c++) {
            float r3 = pv[c * 4 + 3], ri = pv[c * 4 + i];
            planes[2 * i + 0][c] = -(r3 + ri);
            planes[2 * i + 1][c] = -(r3 - ri);
        }
}

/* glm::perspective_lh_zo (nalgebra-glm ext/matrix_clip_space.rs). */
static void perspective_lh_zo(float aspect, float fovy, float near, float far, float m[16]) {
    memset(m, 0, 16 * sizeof(float));
    float tan_half = tanf(fovy / 2.0f);
    m[0 * 4 + 0] = 1.0f / (aspect * tan_half);
    m[1 * 4 + 1] = 1.0f / tan_half;
    m[2 * 4 + 2] = far / (far - near);
    m[3 * 4 + 2] = -(far * near) / (far - near);
    m[2 * 4 + 3] = 1.0f;
}

static void normalize3(float v[3]) {
    float n = sqrtf(dot3(v, v));
    v[0] /= n; v[1] /= n; v[2] /= n;
}
static void cross3(const float a[3], const float b[3], float o[3]) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
/* UnitQuaternion * Vector3 (nalgebra 0.29 geometry/quaternion_ops.rs, what
 * transform_vector and `rot * v` evaluate): t = (q.ijk x v) * 2;
 * result = ((t * w) + (q.ijk x t)) + v -- the cross-product form, not the matrix. */
static void quat_rotate(const float q[4], const float v[3], float o[3]) {
    float t[3], c[3];
    cross3(q, v, t);
    for (int a = 0; a < 3; a++) t[a] = t[a] * 2.0f;
    cross3(q, t, c);
    for (int a = 0; a < 3; a++) o[a] = (t[a] * q[3] + c[a]) + v[a];
}

/* project_camera (src/ecs.rs:66-91): near 0.1, far 100, fov 70 deg, view =
 * look_at_lh(position, position + rot*z, rot*y): the direction the look-at normalises is
 * (position + dir) - position, as the reference builds it (ecs.rs:76-80), and
 * Matrix4::look_at_lh goes through Rotation3::face_towards(dir, up).inverse() with the
 * translation rot * (-eye) (nalgebra geometry/isometry_construction.rs).  tanf is not
 * bit-pinned; planes are INPUTS to both the oracle frame and the GPU. */
void oracle_project_camera(float aspect, const float pos[3], const float rot[4],
                           float proj[16], float view[16], float pv[16], float planes[6][4]) {
    const float fwd[3] = {0, 0, 1}, upv[3] = {0, 1, 0};
    float fovy = 70.0f * 0.017453292519943295f;
    perspective_lh_zo(aspect, fovy, 0.1f, 100.0f, proj);
    float f[3], up[3], s[3], u[3];
    float dir[3];
    quat_rotate(rot, fwd, dir);
    quat_rotate(rot, upv, up);
    for (int a = 0; a < 3; a++) f[a] = (pos[a] + dir[a]) - pos[a]; /* target - eye */
    normalize3(f);
    cross3(up, f, s);
    normalize3(s);
    cross3(f, s, u);
    mat4_identity(view);
    for (int c = 0; c < 3; c++) {
        view[c * 4 + 0] = s[c];
        view[c * 4 + 1] = u[c];
        view[c * 4 + 2] = f[c];
    }
    view[12] = -dot3(s, pos);
    view[13] = -dot3(u, pos);
    view[14] = -dot3(f, pos);
    oracle_mat4_mul(proj, view, pv);
    oracle_frustum_planes(pv, planes);
}

/* UnitQuaternion::look_at_lh(dir, up): the rotation taking world space into a
 * frame whose +z is dir (left-handed view).  Built from the basis rows. */
void oracle_quat_look_at_lh(const float dir[3], const float up[3], float q[4]) {
    float z[3] = {dir[0], dir[1], dir[2]}, x[3], y[3];
    normalize3(z);
    cross3(up, z, x);
    normalize3(x);
    cross3(z, x, y);
    /* rotation matrix rows = x, y, z  (world -> view) */
    float m00 = x[0], m01 = x[1], m02 = x[2];
    float m10 = y[0], m11 = y[1], m12 = y[2];
    float m20 = z[0], m21 = z[1], m22 = z[2];
    float tr = m00 + m11 + m22, s;
    if (tr > 0.0f) {
        s = sqrtf(tr + 1.0f) * 2.0f;
        q[3] = 0.25f * s; q[0] = (m21 - m12) / s; q[1] = (m02 - m20) / s; q[2] = (m10 - m01) / s;
    } else if (m00 > m11 && m00 > m22) {
        s = sqrtf(1.0f + m00 - m11 - m22) * 2.0f;
        q[3] = (m21 - m12) / s; q[0] = 0.25f * s; q[1] = (m01 + m10) / s; q[2] = (m02 + m20) / s;
    } else if (m11 > m22) {
        s = sqrtf(1.0f + m11 - m00 - m22) * 2.0f;
        q[3] = (m02 - m20) / s; q[0] = (m01 + m10) / s; q[1] = 0.25f * s; q[2] = (m12 + m21) / s;
    } else {
        s = sqrtf(1.0f + m22 - m00 - m11) * 2.0f;
        q[3] = (m10 - m01) / s; q[0] = (m02 + m20) / s; q[1] = (m12 + m21) / s; q[2] = 0.25f * s;
    }
    float n = sqrtf(dot4(q, q));
    for (int a = 0; a < 4; a++) q[a] /= n;
}

/* shadow_mapping_mvp_calculation (shadow_mapping.rs:321-329):
 * projection = perspective_lh_zo(1, 70 deg, 10, 400);
 * view = translation(rot * (-pos)) * rot.to_homogeneous().  The reference derives
 * no planes for lights; BASELINE config 3 applies ecs.rs:82-90 to this pv. */
void oracle_light_view(const float pos[3], const float rot[4],
                       float proj[16], float view[16], float pv[16], float planes[6][4]) {
    float fovy = 70.0f * 0.017453292519943295f;
    perspective_lh_zo(1.0f, fovy, 10.0f, 400.0f, proj);
    float neg[3] = {-pos[0], -pos[1], -pos[2]}, t[3], tm[16], rm[16];
    quat_rotate(rot, neg, t);
    mat4_identity(tm);
    tm[12] = t[0]; tm[13] = t[1]; tm[14] = t[2];
    quat_to_homogeneous(rot, rm);
    oracle_mat4_mul(tm, rm, view);
    oracle_mat4_mul(proj, view, pv);
    oracle_frustum_planes(pv, planes);
}

/* ---- whole frame, reference-shaped -------------------------------------------- */

typedef struct {
    const float *pos, *quat, *scale;
    const uint32_t *mesh_id;
    const oracle_mesh *meshes;
    const float *planes;
    uint32_t n_views;
    uint64_t n, lo, hi;
    float *model, *aabb;
    uint8_t *flags;
    int pass;
} frame_job;

static void frame_pass(frame_job *j) {
    if (j->pass == 0) { /* model_matrix_calculation: Query<(&Position,&Rotation,&Scale,&mut ModelMatrix)> */
        for (uint64_t i = j->lo; i < j->hi; i++)
            oracle_model_matrix(j->pos + 3 * i, j->quat + 4 * i, j->scale[i], j->model + 16 * i);
    } else if (j->pass == 1) { /* aabb_calculation: Query<(&ModelMatrix,&GltfMesh,&mut AABB)> */
        for (uint64_t i = j->lo; i < j->hi; i++) {
            const oracle_mesh *m = j->meshes + j->mesh_id[i];
            oracle_aabb(j->model + 16 * i, m->aabb_min, m->aabb_max, j->aabb + 6 * i, j->aabb + 6 * i + 3);
        }
    } else { /* coarse_culling, once per view: Query<(&AABB,&mut CoarseCulled)> */
        for (uint32_t v = 0; v < j->n_views; v++) {
            const float(*pl)[4] = (const float(*)[4])(j->planes + 24 * v);
            for (uint64_t i = j->lo; i < j->hi; i++)
                j->flags[(uint64_t)v * j->n + i] = (uint8_t)oracle_coarse_culled(j->aabb + 6 * i, j->aabb + 6 * i + 3, pl);
        }
    }
}
static void *frame_thread(void *p) {
    frame_pass((frame_job *)p);
    return NULL;
}

static void run_pass(frame_job *base, int pass, int n_threads) {
    if (n_threads <= 1 || base->n < 1024) {
        frame_job j = *base;
        j.pass = pass; j.lo = 0; j.hi = base->n;
        frame_pass(&j);
        return;
    }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * n_threads);
    frame_job *jobs = (frame_job *)malloc(sizeof(frame_job) * n_threads);
    for (int t = 0; t < n_threads; t++) {
        jobs[t] = *base;
        jobs[t].pass = pass;
        jobs[t].lo = base->n * (uint64_t)t / n_threads;
        jobs[t].hi = base->n * (uint64_t)(t + 1) / n_threads;
        pthread_create(&th[t], NULL, frame_thread, &jobs[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

int oracle_frame(const float *pos, const float *quat, const float *scale, const uint32_t *mesh_id,
                 uint64_t n, const oracle_mesh *meshes, uint32_t n_meshes,
                 const float *planes, uint32_t n_views, const float cam_pos[3], float lod_distance,
                 uint32_t draw_index_base, float *model, float *aabb, uint8_t *flags,
                 oracle_cmd *cmds, uint32_t *visible, uint32_t *counts, uint64_t *index_totals,
                 int n_threads) {
    for (uint64_t i = 0; i < n; i++)
        if (mesh_id[i] >= n_meshes) return -1;
    float *own_model = NULL, *own_aabb = NULL;
    uint8_t *own_flags = NULL;
    if (!model) model = own_model = (float *)malloc(sizeof(float) * 16 * (n ? n : 1));
    if (!aabb) aabb = own_aabb = (float *)malloc(sizeof(float) * 6 * (n ? n : 1));
    if (!flags) flags = own_flags = (uint8_t *)malloc((size_t)(n ? n : 1) * (n_views ? n_views : 1));
    frame_job base = {pos, quat, scale, mesh_id, meshes, planes, n_views, n, 0, n, model, aabb, flags, 0};
    run_pass(&base, 0, n_threads);
    run_pass(&base, 1, n_threads);
    run_pass(&base, 2, n_threads);

    /* cull_pass emit loop (cull_pipeline.rs:534-577) + generate_work.comp:61-67
     * header write + compact_draw_stream.comp:41 keep-rule, in draw_index order. */
    for (uint32_t v = 0; v < n_views; v++) {
        uint32_t index_offset_in_output = 0, count = 0;
        uint64_t total = 0;
        const uint8_t *f = flags + (uint64_t)v * n;
        for (uint64_t i = 0; i < n; i++) {
            if (f[i]) continue;
            const oracle_mesh *m = meshes + mesh_id[i];
            uint32_t lod = oracle_pick_lod(cam_pos, pos + 3 * i, m->n_lods, lod_distance);
            uint32_t index_len = m->lod[lod].index_len;
            if (cmds) {
                oracle_cmd *c = cmds + (uint64_t)v * n + count;
                c->indexCount = index_len; /* entity-level upper bound */
                c->instanceCount = 1;
                c->firstIndex = index_offset_in_output;
                c->vertexOffset = m->vertex_offset;
                c->firstInstance = draw_index_base + (uint32_t)i;
            }
            if (visible) visible[(uint64_t)v * n + count] = draw_index_base + (uint32_t)i;
            index_offset_in_output += index_len; /* u32, wraps like release Rust */
            total += index_len;
            count++;
        }
        if (counts) counts[v] = count;
        if (index_totals) index_totals[v] = total;
    }
    free(own_model);
    free(own_aabb);
    free(own_flags);
    return 0;
}

/* ---- fused CPU variant (CPU baseline (ii)) ----------------------------
 * The same per-entity functions as oracle_frame, scheduled the way a CPU implementation
 * that cared about memory traffic would: ONE pass per entity (A1 -> A2 -> A3 for every
 * view, matrix and box kept in registers / L1 unless the caller wants them), a per-thread
 * count of survivors and index lengths, an exclusive prefix over the threads, and a second
 * pass in which every thread writes its own slice of the draw lists.  Outputs are
 * bit-identical to oracle_frame's (tests/test_oracle.py); only the schedule differs.  It is
 * the stronger of the two CPU baselines bench.py reports, not a second definition. */
typedef struct {
    const float *pos, *quat, *scale;
    const uint32_t *mesh_id;
    const oracle_mesh *meshes;
    const float *planes;
    const float *cam_pos;
    float lod_distance;
    uint32_t n_views, draw_index_base;
    uint64_t n, lo, hi;
    float *model, *aabb;
    uint8_t *flags;      /* optional output [n_views][n] */
    uint8_t *culled_in;  /* [n]: bit v = culled in view v (n_views <= 8) */
    oracle_cmd *cmds;
    uint32_t *visible;
    uint32_t count[8];      /* phase A out: survivors of this range per view */
    uint64_t len[8];        /*              their index lengths */
    uint32_t count_base[8]; /* phase B in: exclusive prefixes over the earlier ranges */
    uint64_t len_base[8];
} fused_job;

static void *fused_phase_a(void *arg) {
    fused_job *j = (fused_job *)arg;
    for (uint32_t v = 0; v < j->n_views; v++) { j->count[v] = 0; j->len[v] = 0; }
    for (uint64_t i = j->lo; i < j->hi; i++) {
        float m_local[16], box_local[6];
        float *m = j->model ? j->model + 16 * i : m_local;
        float *box = j->aabb ? j->aabb + 6 * i : box_local;
        const oracle_mesh *mesh = j->meshes + j->mesh_id[i];
        oracle_model_matrix(j->pos + 3 * i, j->quat + 4 * i, j->scale[i], m);
        oracle_aabb(m, mesh->aabb_min, mesh->aabb_max, box, box + 3);
        uint8_t mask = 0;
        for (uint32_t v = 0; v < j->n_views; v++) {
            const int culled = oracle_coarse_culled(box, box + 3, (const float(*)[4])(j->planes + 24 * v));
            if (culled) mask |= (uint8_t)(1u << v);
            if (j->flags) j->flags[(uint64_t)v * j->n + i] = (uint8_t)culled;
        }
        j->culled_in[i] = mask;
        if (mask != (uint8_t)((1u << j->n_views) - 1u)) { /* visible somewhere: pick_lod once */
            const uint32_t lod = oracle_pick_lod(j->cam_pos, j->pos + 3 * i, mesh->n_lods, j->lod_distance);
            const uint32_t index_len = mesh->lod[lod].index_len;
            for (uint32_t v = 0; v < j->n_views; v++)
                if (!(mask & (1u << v))) { j->count[v]++; j->len[v] += index_len; }
        }
    }
    return NULL;
}

static void *fused_phase_b(void *arg) {
    fused_job *j = (fused_job *)arg;
    uint32_t at[8], first_index[8];
    for (uint32_t v = 0; v < j->n_views; v++) { at[v] = j->count_base[v]; first_index[v] = (uint32_t)j->len_base[v]; }
    const uint8_t all = (uint8_t)((1u << j->n_views) - 1u);
    for (uint64_t i = j->lo; i < j->hi; i++) {
        const uint8_t mask = j->culled_in[i];
        if (mask == all) continue;
        const oracle_mesh *mesh = j->meshes + j->mesh_id[i];
        const uint32_t lod = oracle_pick_lod(j->cam_pos, j->pos + 3 * i, mesh->n_lods, j->lod_distance);
        const uint32_t index_len = mesh->lod[lod].index_len;
        for (uint32_t v = 0; v < j->n_views; v++) {
            if (mask & (1u << v)) continue;
            if (j->cmds) {
                oracle_cmd *c = j->cmds + (uint64_t)v * j->n + at[v];
                c->indexCount = index_len;
                c->instanceCount = 1;
                c->firstIndex = first_index[v];
                c->vertexOffset = mesh->vertex_offset;
                c->firstInstance = j->draw_index_base + (uint32_t)i;
            }
            if (j->visible) j->visible[(uint64_t)v * j->n + at[v]] = j->draw_index_base + (uint32_t)i;
            first_index[v] += index_len; /* u32, wraps like oracle_frame's */
            at[v]++;
        }
    }
    return NULL;
}

static void fused_run(fused_job *jobs, int n_threads, void *(*fn)(void *)) {
    if (n_threads == 1) { fn(&jobs[0]); return; }
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * n_threads);
    for (int t = 0; t < n_threads; t++) pthread_create(&th[t], NULL, fn, &jobs[t]);
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    free(th);
}

int oracle_frame_fused(const float *pos, const float *quat, const float *scale, const uint32_t *mesh_id,
                       uint64_t n, const oracle_mesh *meshes, uint32_t n_meshes,
                       const float *planes, uint32_t n_views, const float cam_pos[3], float lod_distance,
                       uint32_t draw_index_base, float *model, float *aabb, uint8_t *flags,
                       oracle_cmd *cmds, uint32_t *visible, uint32_t *counts, uint64_t *index_totals,
                       int n_threads) {
    if (n_views < 1 || n_views > 8) return -1;
    for (uint64_t i = 0; i < n; i++)
        if (mesh_id[i] >= n_meshes) return -1;
    if (n_threads < 1 || n < 1024) n_threads = 1;
    uint8_t *mask = (uint8_t *)malloc(n ? n : 1);
    fused_job *jobs = (fused_job *)calloc((size_t)n_threads, sizeof(fused_job));
    for (int t = 0; t < n_threads; t++) {
        fused_job *j = &jobs[t];
        j->pos = pos; j->quat = quat; j->scale = scale; j->mesh_id = mesh_id; j->meshes = meshes;
        j->planes = planes; j->cam_pos = cam_pos; j->lod_distance = lod_distance;
        j->n_views = n_views; j->draw_index_base = draw_index_base; j->n = n;
        j->lo = n * (uint64_t)t / (uint64_t)n_threads;
        j->hi = n * (uint64_t)(t + 1) / (uint64_t)n_threads;
        j->model = model; j->aabb = aabb; j->flags = flags; j->culled_in = mask; j->cmds = cmds; j->visible = visible;
    }
    fused_run(jobs, n_threads, fused_phase_a);
    for (uint32_t v = 0; v < n_views; v++) {
        uint32_t c = 0;
        uint64_t l = 0;
        for (int t = 0; t < n_threads; t++) {
            jobs[t].count_base[v] = c;
            jobs[t].len_base[v] = l;
            c += jobs[t].count[v];
            l += jobs[t].len[v];
        }
        if (counts) counts[v] = c;
        if (index_totals) index_totals[v] = l;
    }
    if (cmds || visible) fused_run(jobs, n_threads, fused_phase_b);
    free(jobs);
    free(mask);
    return 0;
}

/* ---- per-mesh instanced emission (BASELINE config 4) -------------------------- */
int oracle_emit_per_mesh(const float *pos, const uint32_t *mesh_id, const uint8_t *flags, uint64_t n,
                         const oracle_mesh *meshes, uint32_t n_meshes, const float cam_pos[3],
                         float lod_distance, uint32_t draw_index_base,
                         oracle_cmd *cmds, uint32_t *visible, uint32_t *count, uint64_t *n_instances) {
    uint32_t n_slots = n_meshes * ORACLE_MAX_LODS;
    uint32_t *bucket = (uint32_t *)calloc(n_slots + 1, sizeof(uint32_t));
    uint32_t *cursor = (uint32_t *)calloc(n_slots + 1, sizeof(uint32_t));
    for (uint64_t i = 0; i < n; i++) {
        if (flags[i]) continue;
        if (mesh_id[i] >= n_meshes) { free(bucket); free(cursor); return -1; }
        const oracle_mesh *m = meshes + mesh_id[i];
        uint32_t lod = oracle_pick_lod(cam_pos, pos + 3 * i, m->n_lods, lod_distance);
        bucket[mesh_id[i] * ORACLE_MAX_LODS + lod]++;
    }
    uint32_t run = 0, kept = 0;
    for (uint32_t s = 0; s < n_slots; s++) {
        cursor[s] = run;
        if (bucket[s] > 0) { /* instanceCount > 0 keep-rule */
            const oracle_mesh *m = meshes + s / ORACLE_MAX_LODS;
            oracle_cmd *c = cmds + kept++;
            c->indexCount = m->lod[s % ORACLE_MAX_LODS].index_len;
            c->instanceCount = bucket[s];
            c->firstIndex = m->lod[s % ORACLE_MAX_LODS].first_index;
            c->vertexOffset = m->vertex_offset;
            c->firstInstance = run;
        }
        run += bucket[s];
    }
    for (uint64_t i = 0; i < n; i++) {
        if (flags[i]) continue;
        const oracle_mesh *m = meshes + mesh_id[i];
        uint32_t lod = oracle_pick_lod(cam_pos, pos + 3 * i, m->n_lods, lod_distance);
        visible[cursor[mesh_id[i] * ORACLE_MAX_LODS + lod]++] = draw_index_base + (uint32_t)i;
    }
    *count = kept;
    *n_instances = run;
    free(bucket);
    free(cursor);
    return 0;
}

/* ---- synthetic scene generator twin ------------------------------------------- */
static uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint64_t chan(uint64_t seed, uint64_t i, uint64_t c) {
    return mix64(seed + (i * 16 + c + 1) * 0x9E3779B97F4A7C15ull);
}
static float unit24(uint64_t h) { return (float)(h >> 40) * 5.9604644775390625e-08f; /* 2^-24 */ }

void oracle_synth_entities(uint64_t seed, uint64_t first, uint64_t n, float half_box, uint32_t n_meshes,
                           float *pos, float *quat, float *scale, uint32_t *mesh_id) {
    for (uint64_t k = 0; k < n; k++) {
        uint64_t i = first + k;
        for (int a = 0; a < 3; a++) pos[3 * k + a] = (unit24(chan(seed, i, a)) * 2.0f - 1.0f) * half_box;
        scale[k] = unit24(chan(seed, i, 3)) * 3.75f + 0.25f;
        float c[4];
        for (int a = 0; a < 4; a++) c[a] = unit24(chan(seed, i, 4 + a)) * 2.0f - 1.0f;
        float n2 = ((c[0] * c[0] + c[1] * c[1]) + c[2] * c[2]) + c[3] * c[3];
        if (n2 < 1e-4f) {
            c[0] = c[1] = c[2] = 0.0f;
            c[3] = 1.0f;
        } else {
            float nn = sqrtf(n2);
            for (int a = 0; a < 4; a++) c[a] = c[a] / nn;
        }
        /* UnitQuaternion::new_normalize (scene_loader.rs:590): q / sqrt(dot4(q,q)) */
        float nn = sqrtf(dot4(c, c));
        for (int a = 0; a < 4; a++) quat[4 * k + a] = c[a] / nn;
        mesh_id[k] = (uint32_t)(((chan(seed, i, 8) >> 32) * (uint64_t)n_meshes) >> 32);
    }
}

/* ------------------------------------------------------------------------------------
 * A8: per-triangle cull (src/shaders/generate_work.comp:60-201) + A7 on the refined
 * commands (src/shaders/compact_draw_stream.comp:34-63).
 * ---------------------------------------------------------------------------------- */

/* GLSL `mat4 * vec4`: linear combination of the columns, left to right
 * (generate_work.comp:139-141 applies it twice: pv * (model * vec4(v, 1.0))). */
static void glsl_mat4_mul_vec4(const float m[16], const float v[4], float out[4]) {
    for (int r = 0; r < 4; r++) {
        float a = m[0 + r] * v[0];
        a = a + m[4 + r] * v[1];
        a = a + m[8 + r] * v[2];
        a = a + m[12 + r] * v[3];
        out[r] = a;
    }
}

/* generate_work.comp:139-155 for one triangle; returns 1 when culled. */
static int tri_culled(const float model[16], const float pv[16], const float *p0, const float *p1, const float *p2) {
    const float *p[3] = {p0, p1, p2};
    float c[3][4];
    for (int k = 0; k < 3; k++) {
        const float v[4] = {p[k][0], p[k][1], p[k][2], 1.0f};
        float w[4];
        glsl_mat4_mul_vec4(model, v, w);
        glsl_mat4_mul_vec4(pv, w, c[k]);
    }
    /* determinant(mat3(c0.xyw, c1.xyw, c2.xyw)) > 0 (:143): columns a, b, d */
    const float ax = c[0][0], ay = c[0][1], az = c[0][3];
    const float bx = c[1][0], by = c[1][1], bz = c[1][3];
    const float dx = c[2][0], dy = c[2][1], dz = c[2][3];
    const float t0 = by * dz - dy * bz;
    const float t1 = ay * dz - dy * az;
    const float t2 = ay * bz - by * az;
    const float det = (ax * t0 - bx * t1) + dx * t2;
    if (det > 0.0f) return 1;
    /* ndc = xyz / w (:145-147); only x and y are tested (:152-155) */
    const float x0 = c[0][0] / c[0][3], y0 = c[0][1] / c[0][3];
    const float x1 = c[1][0] / c[1][3], y1 = c[1][1] / c[1][3];
    const float x2 = c[2][0] / c[2][3], y2 = c[2][1] / c[2][3];
    if (x0 < -1.0f && x1 < -1.0f && x2 < -1.0f) return 1;
    if (x0 > 1.0f && x1 > 1.0f && x2 > 1.0f) return 1;
    if (y0 < -1.0f && y1 < -1.0f && y2 < -1.0f) return 1;
    if (y0 > 1.0f && y1 > 1.0f && y2 > 1.0f) return 1;
    return 0; /* the area test (:159-167) is overwritten by `cull = false` at :166 */
}

typedef struct {
    const oracle_cmd *cmds;
    const uint32_t *src_first_index;
    uint32_t lo, hi;
    const float *model;
    uint32_t draw_index_base;
    const float *pv;
    const float *positions;
    uint64_t n_vertices;
    const uint32_t *indices;
    uint64_t n_indices;
    uint32_t *out_indices;
    uint64_t out_capacity;
    uint32_t *kept; /* [count]: refined indexCount */
    int err;
} tri_job;

static void *tri_thread(void *arg) {
    tri_job *j = (tri_job *)arg;
    for (uint32_t c = j->lo; c < j->hi; c++) {
        const oracle_cmd *cmd = &j->cmds[c];
        const uint32_t n_tri = cmd->indexCount / 3;           /* :73 indexCount / 3 */
        const uint64_t src = j->src_first_index[c] / 3;       /* :74 indexOffset / 3 */
        const float *model = j->model + 16 * (uint64_t)(cmd->firstInstance - j->draw_index_base);
        uint32_t kept = 0;
        for (uint32_t t = 0; t < n_tri; t++) {
            const uint64_t at = 3 * (src + t);
            if (at + 2 >= j->n_indices) { j->err = 1; break; }
            const uint32_t *ix = j->indices + at;
            uint64_t v[3];
            int bad = 0;
            for (int k = 0; k < 3; k++) {
                v[k] = (uint64_t)((int64_t)cmd->vertexOffset + (int64_t)ix[k]);
                if (v[k] >= j->n_vertices) bad = 1;
            }
            if (bad) { j->err = 1; break; }
            if (tri_culled(model, j->pv, j->positions + 3 * v[0], j->positions + 3 * v[1], j->positions + 3 * v[2])) continue;
            const uint64_t dst = ((uint64_t)cmd->firstIndex + 3ull * kept) / 3 * 3; /* :190-196: (ix + indexOffsetInOutput) / 3 */
            if (dst + 2 < j->out_capacity) {
                j->out_indices[dst] = ix[0];
                j->out_indices[dst + 1] = ix[1];
                j->out_indices[dst + 2] = ix[2];
            }
            kept++;
        }
        j->kept[c] = 3 * kept;
    }
    return NULL;
}

int oracle_tri_cull(const oracle_cmd *cmds, const uint32_t *src_first_index, uint32_t count,
                    const float *model, uint32_t draw_index_base, const float pv[16],
                    const float *positions, uint64_t n_vertices, const uint32_t *indices, uint64_t n_indices,
                    uint32_t *out_indices, uint64_t out_capacity,
                    oracle_cmd *out_cmds, uint32_t *out_count, uint64_t *kept_indices, int n_threads) {
    uint32_t *kept = (uint32_t *)calloc(count ? count : 1, sizeof(uint32_t));
    if (!kept) return -1;
    if (n_threads < 1) n_threads = 1;
    if (count < 64) n_threads = 1;
    pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * n_threads);
    tri_job *jobs = (tri_job *)malloc(sizeof(tri_job) * n_threads);
    for (int t = 0; t < n_threads; t++) {
        tri_job j = {cmds, src_first_index, (uint32_t)((uint64_t)count * t / n_threads),
                     (uint32_t)((uint64_t)count * (t + 1) / n_threads), model, draw_index_base, pv, positions,
                     n_vertices, indices, n_indices, out_indices, out_capacity, kept, 0};
        jobs[t] = j;
        if (n_threads == 1) tri_thread(&jobs[t]);
        else pthread_create(&th[t], NULL, tri_thread, &jobs[t]);
    }
    int err = 0;
    for (int t = 0; t < n_threads; t++) {
        if (n_threads > 1) pthread_join(th[t], NULL);
        err |= jobs[t].err;
    }
    /* compact_draw_stream.comp:40-62: keep indexCount > 0, pack to the front, in order */
    uint32_t n_out = 0;
    uint64_t total = 0;
    for (uint32_t c = 0; c < count; c++) {
        if (kept[c] == 0) continue;
        if (out_cmds) {
            out_cmds[n_out] = cmds[c];
            out_cmds[n_out].indexCount = kept[c];
        }
        total += kept[c];
        n_out++;
    }
    if (out_count) *out_count = n_out;
    if (kept_indices) *kept_indices = total;
    free(kept); free(th); free(jobs);
    return err ? -1 : 0;
}

/* F4: acceleration_strucures.rs:419-449. */
void oracle_as_instances(const float *model, const uint32_t *mesh_id, const uint64_t *blas, uint64_t n,
                         uint32_t draw_index_base, void *out) {
    unsigned char *o = (unsigned char *)out;
    for (uint64_t i = 0; i < n; i++, o += 64) {
        const float *m = model + 16 * i;
        float rows[12];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 4; c++) rows[4 * r + c] = m[4 * c + r]; /* model_matrix.rows(0, 3).transpose().as_slice() (:427) */
        const uint32_t custom = (((uint32_t)(draw_index_base + i)) & 0xffffffu) | (0xffu << 24); /* Packed24_8::new(draw_index, 0xFF) */
        const uint32_t sbt = 0u | (1u << 24); /* Packed24_8::new(0, TRIANGLE_FACING_CULL_DISABLE) */
        const uint64_t addr = blas[mesh_id[i]];
        memcpy(o, rows, 48);
        memcpy(o + 48, &custom, 4);
        memcpy(o + 52, &sbt, 4);
        memcpy(o + 56, &addr, 8);
    }
}
