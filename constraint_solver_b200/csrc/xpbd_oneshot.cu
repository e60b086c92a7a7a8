// Single-call drop-ins for the reference's own entry points: solver::step (src/solver.rs:3-17) and
// World::integrate (src/world.rs:34-43).  Each builds a throw-away batch (upload, ONE fused launch, download), so
// the arithmetic is exactly the batched kernel's.  Meant for bringing a caller over one call at a time; a caller
// that steps many bodies should hold an xpbd_batch instead (the launch + PCIe latency dominates here).
#include "xpbd_internal.hpp"

using namespace xpbd;

namespace {

int step_bodies_once(xpbd_rigid* const* rigids, int n, const xpbd_polytope* hull, double dt, int32_t substeps) {
    xpbd_batch_desc desc{};
    desc.device = 0;
    if (cudaGetDevice(&desc.device) != cudaSuccess) desc.device = 0;
    desc.n_worlds = n;
    desc.bodies_per_world = 1;
    desc.joints_per_world = 0;
    desc.body_body_contacts = 0;
    desc.compliance = 1e-6;  // src/solver.rs:20
    desc.friction = 1.0;     // src/collision.rs:29
    xpbd_batch* b = nullptr;
    int rc = xpbd_batch_create(&desc, &b);
    if (rc != XPBD_OK) return rc;
    std::vector<xpbd_rigid> host((size_t)n);
    for (int i = 0; i < n; ++i) host[(size_t)i] = *rigids[i];
    rc = xpbd_batch_add_hull(b, hull, nullptr);
    if (rc == XPBD_OK) rc = xpbd_batch_upload(b, 0, n, host.data());
    if (rc == XPBD_OK) rc = xpbd_batch_step(b, dt, substeps, 1);
    if (rc == XPBD_OK) rc = xpbd_batch_download(b, 0, n, host.data());
    if (rc == XPBD_OK) {
        // solver::step only mutates velocity, angular_velocity, position and rotation (src/solver.rs:3-17): write back
        // those 13 values (at fp32 precision) and leave the caller's f64 constants — inverse mass / inertia, forces,
        // torques, centre of mass — untouched, so that a caller mixing this drop-in with its own f64 code keeps them.
        for (int i = 0; i < n; ++i) {
            const xpbd_rigid& h = host[(size_t)i];
            xpbd_rigid& r = *rigids[i];
            for (int k = 0; k < 3; ++k) { r.velocity[k] = h.velocity[k]; r.angular_velocity[k] = h.angular_velocity[k]; r.position[k] = h.position[k]; }
            for (int k = 0; k < 4; ++k) r.rotation[k] = h.rotation[k];
        }
    }
    xpbd_batch_destroy(b);
    return rc;
}

}  // namespace

extern "C" {

int xpbd_solver_step(xpbd_rigid* rigid, const xpbd_polytope* hull, double dt, int32_t substep_count) {
    if (!rigid || !hull) return fail(XPBD_ERR_INVALID, "xpbd_solver_step: null argument");
    xpbd_rigid* one[1] = {rigid};
    return step_bodies_once(one, 1, hull, dt, substep_count);
}

// World::integrate steps `a` and `b` independently, BOTH against p1, with 25 substeps; p2 is ignored
// (src/world.rs:41-42 — a quirk of the reference that is kept).
int xpbd_world_integrate(xpbd_rigid* a, xpbd_rigid* b, double dt, const xpbd_polytope* p1, const xpbd_polytope* p2) {
    (void)p2;
    if (!a || !b || !p1) return fail(XPBD_ERR_INVALID, "xpbd_world_integrate: null argument");
    xpbd_rigid* two[2] = {a, b};
    return step_bodies_once(two, 2, p1, dt, 25);
}

}  // extern "C"
