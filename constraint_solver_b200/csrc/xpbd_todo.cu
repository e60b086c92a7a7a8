// Entry points not implemented yet: they fail loudly (no CPU fallback, no silent success).
#include "xpbd_internal.hpp"
using namespace xpbd;
namespace xpbd {
int step_world(xpbd_batch*, float, int, int) { return fail(XPBD_ERR_STATE, "multi-body step not implemented yet"); }
void release_world_scratch(xpbd_batch*) {}
}
#define TODO(name) return fail(XPBD_ERR_STATE, name ": not implemented yet")
extern "C" {
int xpbd_quickhull(const double*, int32_t, double*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, double*) { TODO("xpbd_quickhull"); }
int xpbd_solver_step(xpbd_rigid*, const xpbd_polytope*, double, int32_t) { TODO("xpbd_solver_step"); }
int xpbd_world_integrate(xpbd_rigid*, xpbd_rigid*, double, const xpbd_polytope*, const xpbd_polytope*) { TODO("xpbd_world_integrate"); }
int xpbd_rigid_integrate_batch(xpbd_rigid*, int64_t, double) { TODO("xpbd_rigid_integrate_batch"); }
int xpbd_rigid_derive_batch(xpbd_rigid*, const double*, int64_t, double) { TODO("xpbd_rigid_derive_batch"); }
int xpbd_rigid_apply_impulse_batch(xpbd_rigid*, const double*, int64_t) { TODO("xpbd_rigid_apply_impulse_batch"); }
int xpbd_rigid_frame_batch(const xpbd_rigid*, int64_t, xpbd_frame*) { TODO("xpbd_rigid_frame_batch"); }
int xpbd_ground_batch(const xpbd_rigid*, const xpbd_frame*, int64_t, const xpbd_polytope*, xpbd_constraint*, int32_t*) { TODO("xpbd_ground_batch"); }
int xpbd_solve_batch(xpbd_rigid*, int64_t, const xpbd_constraint*, const int32_t*, int32_t, double, double) { TODO("xpbd_solve_batch"); }
int xpbd_face_axes_separation_batch(const xpbd_frame*, const xpbd_frame*, int64_t, const xpbd_polytope*, const xpbd_polytope*, double*, int32_t*) { TODO("xpbd_face_axes_separation_batch"); }
int xpbd_edge_axes_separation_batch(const xpbd_frame*, const xpbd_frame*, int64_t, const xpbd_polytope*, const xpbd_polytope*, double*, int32_t*) { TODO("xpbd_edge_axes_separation_batch"); }
int xpbd_sat_manifold_batch(const xpbd_frame*, const xpbd_frame*, int64_t, const xpbd_polytope*, const xpbd_polytope*, xpbd_constraint*, int32_t*, int32_t*) { TODO("xpbd_sat_manifold_batch"); }
int xpbd_fp32_peak_microbench(int32_t, int32_t, double*) { TODO("xpbd_fp32_peak_microbench"); }
}
