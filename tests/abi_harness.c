/* C harness for include/xpbd_b200.h (Next round 8): the header compiles as plain C11 with -Wall -Werror,
 * every record has the size / field offsets the Rust `#[repr(C)]` mirrors (rust/src/ffi.rs) and the ctypes / numpy mirrors
 * assume, and — when a CUDA device is present — the boundary is driven the way a maintainer's FFI would drive it:
 * create -> add_hull -> upload -> step -> download, with the reference's default cube (src/world.rs:13,16-20).
 *
 *   gcc -std=c11 -Wall -Werror -Iinclude tests/abi_harness.c -o harness -Lconstraint_solver_b200/lib -lxpbd_b200
 *   ./harness layout      # sizes / offsets only (no GPU)
 *   ./harness run         # + the device round trip; prints the cube's state after 10 frames as hex doubles
 */
#include <stddef.h>
#include <stdio.h>
#include <string.h>

#include "xpbd_b200.h"

#define CHECK(cond)                                                   \
    do {                                                              \
        if (!(cond)) {                                                \
            fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond); \
            return 1;                                                 \
        }                                                             \
    } while (0)

static int layout(void) {
    /* xpbd_rigid = struct Rigid without color: 38 doubles in the reference's field order (src/rigid.rs:9-47) */
    CHECK(sizeof(xpbd_rigid) == 304);
    CHECK(offsetof(xpbd_rigid, inverse_mass) == 0 && offsetof(xpbd_rigid, inverse_inertia) == 8);
    CHECK(offsetof(xpbd_rigid, external_force) == 80 && offsetof(xpbd_rigid, internal_force) == 104);
    CHECK(offsetof(xpbd_rigid, external_torque) == 128 && offsetof(xpbd_rigid, internal_torque) == 152);
    CHECK(offsetof(xpbd_rigid, velocity) == 176 && offsetof(xpbd_rigid, angular_velocity) == 200);
    CHECK(offsetof(xpbd_rigid, center_of_mass) == 224 && offsetof(xpbd_rigid, position) == 248);
    CHECK(offsetof(xpbd_rigid, rotation) == 272);
    CHECK(sizeof(xpbd_frame) == 56 && offsetof(xpbd_frame, rotation) == 24);
    CHECK(sizeof(xpbd_constraint) == 64 && offsetof(xpbd_constraint, rigid_b) == 4);
    CHECK(offsetof(xpbd_constraint, contact0) == 8 && offsetof(xpbd_constraint, contact1) == 32 && offsetof(xpbd_constraint, distance) == 56);
    CHECK(sizeof(xpbd_rigid_metrics) == 112 && offsetof(xpbd_rigid_metrics, inertia_tensor) == 40);
    CHECK(sizeof(xpbd_joint) == 64 && offsetof(xpbd_joint, anchor_a) == 8 && offsetof(xpbd_joint, anchor_b) == 32 && offsetof(xpbd_joint, distance) == 56);
    CHECK(sizeof(xpbd_polytope) == 80 && offsetof(xpbd_polytope, n_vertices) == 8 && offsetof(xpbd_polytope, edges) == 16);
    CHECK(offsetof(xpbd_polytope, n_edges) == 24 && offsetof(xpbd_polytope, face_offsets) == 32 && offsetof(xpbd_polytope, face_indices) == 40);
    CHECK(offsetof(xpbd_polytope, n_faces) == 48 && offsetof(xpbd_polytope, centroid) == 56);
    CHECK(sizeof(xpbd_batch_desc) == 48 && offsetof(xpbd_batch_desc, n_worlds) == 8 && offsetof(xpbd_batch_desc, bodies_per_world) == 16);
    CHECK(offsetof(xpbd_batch_desc, joints_per_world) == 20 && offsetof(xpbd_batch_desc, body_body_contacts) == 24);
    CHECK(offsetof(xpbd_batch_desc, compliance) == 32 && offsetof(xpbd_batch_desc, friction) == 40);
    CHECK(sizeof(xpbd_step_stats) == 64 && offsetof(xpbd_step_stats, elapsed_ms) == 24 && offsetof(xpbd_step_stats, kernel_launches) == 28);
    CHECK(offsetof(xpbd_step_stats, stage_ms) == 32 && offsetof(xpbd_step_stats, candidate_pairs) == 48);
    CHECK(offsetof(xpbd_step_stats, sub_batches) == 56 && offsetof(xpbd_step_stats, nan_worlds) == 60);
    CHECK(XPBD_OK == 0 && XPBD_ERR_INVALID == -1 && XPBD_ERR_CUDA == -2 && XPBD_ERR_CAPACITY == -3 && XPBD_ERR_SINGULAR == -4 && XPBD_ERR_STATE == -5);
    CHECK(XPBD_MAX_MANIFOLD == 4);
    printf("layout ok\n");
    return 0;
}

/* Polytope::new_cube (src/geometry.rs:113-149) */
static const double cube_vertices[8 * 3] = {0, 0, 0, 1, 0, 0, 0, 1, 0, 1, 1, 0, 0, 0, 1, 1, 0, 1, 0, 1, 1, 1, 1, 1};
static const int32_t cube_edges[12 * 2] = {0, 1, 1, 3, 3, 2, 2, 0, 4, 5, 5, 7, 7, 6, 6, 4, 0, 4, 1, 5, 3, 7, 2, 6};
static const int32_t cube_face_offsets[7] = {0, 4, 8, 12, 16, 20, 24};
static const int32_t cube_face_indices[24] = {0, 2, 3, 1, 4, 5, 7, 6, 4, 0, 1, 5, 5, 1, 3, 7, 7, 3, 2, 6, 6, 2, 0, 4};

static int run(void) {
    xpbd_polytope cube;
    memset(&cube, 0, sizeof cube);
    cube.vertices = cube_vertices; cube.n_vertices = 8;
    cube.edges = cube_edges; cube.n_edges = 12;
    cube.face_offsets = cube_face_offsets; cube.face_indices = cube_face_indices; cube.n_faces = 6;
    cube.centroid[0] = cube.centroid[1] = cube.centroid[2] = 0.5;

    /* World::new body a (src/world.rs:13,16-20): host-side f64 setup math through the ABI */
    xpbd_rigid_metrics m;
    CHECK(xpbd_polytope_rigid_metrics(&cube, 0.1, &m) == XPBD_OK);
    CHECK(m.mass > 0.0999999 && m.mass < 0.1000001 && m.volume > 0.999999 && m.volume < 1.000001);
    xpbd_rigid a;
    CHECK(xpbd_rigid_new(&m, &a) == XPBD_OK);
    a.position[2] = 4.0; a.velocity[1] = 2.5; a.angular_velocity[0] = -4.0; a.angular_velocity[1] = 1.0; a.external_force[2] = -2.0;

    int32_t ndev = 0;
    if (xpbd_device_count(&ndev) != XPBD_OK || ndev == 0) {
        /* no CPU fallback: every compute entry point must fail loudly */
        xpbd_batch_desc d0 = {0, 4, 1, 0, 0, 1e-6, 1.0};
        xpbd_batch* none = NULL;
        CHECK(xpbd_batch_create(&d0, &none) == XPBD_ERR_CUDA && none == NULL);
        CHECK(strstr(xpbd_last_error(), "no CPU fallback") != NULL);
        printf("no device: compute entry points refuse (%s)\n", xpbd_last_error());
        return 0;
    }

    enum { N = 3 };
    xpbd_batch_desc desc = {0, N, 1, 0, 0, 1e-6, 1.0};
    xpbd_batch* batch = NULL;
    CHECK(xpbd_batch_create(&desc, &batch) == XPBD_OK && batch != NULL);
    int32_t hull_id = -1;
    CHECK(xpbd_batch_add_hull(batch, &cube, &hull_id) == XPBD_OK && hull_id == 0);
    xpbd_rigid worlds[N];
    for (int i = 0; i < N; ++i) worlds[i] = a;
    CHECK(xpbd_batch_upload(batch, 0, N, worlds) == XPBD_OK);
    CHECK(xpbd_batch_step(batch, 1.0 / 60.0, 25, 10) == XPBD_OK);   /* World::integrate x 10 */
    CHECK(xpbd_batch_sync(batch) == XPBD_OK);
    xpbd_step_stats st;
    CHECK(xpbd_batch_last_stats(batch, &st) == XPBD_OK);
    CHECK(st.body_substeps == (uint64_t)N * 25 * 10 && st.kernel_launches >= 1 && st.nan_worlds == 0);
    xpbd_rigid out[N];
    CHECK(xpbd_batch_download(batch, 0, N, out) == XPBD_OK);
    /* identical worlds give identical results; free fall closed form to fp32 accuracy: the
     * fp32 and f64 runs of this scene are 2e-3 m apart by frame 10 */
    CHECK(memcmp(&out[0], &out[1], sizeof out[0]) == 0 && memcmp(&out[0], &out[2], sizeof out[0]) == 0);
    const double dts = 1.0 / 60.0 / 25.0, n = 250.0;
    const double z = 4.0 - 20.0 * dts * dts * n * (n + 1.0) / 2.0, y = 2.5 * n * dts;
    CHECK(out[0].position[2] > z - 5e-3 && out[0].position[2] < z + 5e-3);
    CHECK(out[0].position[1] > y - 5e-3 && out[0].position[1] < y + 5e-3);
    /* the same frame through the records-in-host-memory call: in place, constants untouched */
    xpbd_rigid host[N];
    for (int i = 0; i < N; ++i) host[i] = a;
    for (int f = 0; f < 10; ++f) {
        CHECK(xpbd_batch_step_host_rigid(batch, 1.0 / 60.0, 25, 1, host, host) == XPBD_OK);
        CHECK(xpbd_batch_sync(batch) == XPBD_OK);
    }
    CHECK(memcmp(host[0].position, out[0].position, sizeof out[0].position) == 0);
    CHECK(memcmp(host[0].rotation, out[0].rotation, sizeof out[0].rotation) == 0);
    CHECK(memcmp(&host[0].inverse_mass, &a.inverse_mass, 22 * sizeof(double)) == 0);
    CHECK(xpbd_batch_destroy(batch) == XPBD_OK);
    /* the single-call drop-in: solver::step(&mut rigid, &polytope, dt, 25) ten times */
    xpbd_rigid one = a;
    for (int f = 0; f < 10; ++f) CHECK(xpbd_solver_step(&one, &cube, 1.0 / 60.0, 25) == XPBD_OK);
    CHECK(memcmp(one.position, out[0].position, sizeof one.position) == 0);
    printf("run ok: position %a %a %a rotation %a %a %a %a\n", out[0].position[0], out[0].position[1], out[0].position[2],
           out[0].rotation[0], out[0].rotation[1], out[0].rotation[2], out[0].rotation[3]);
    return 0;
}

int main(int argc, char** argv) {
    if (layout()) return 1;
    if (argc > 1 && strcmp(argv[1], "run") == 0) return run();
    return 0;
}
