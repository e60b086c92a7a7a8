// Entry points not implemented yet: they fail loudly (no CPU fallback, no silent success).
#include "xpbd_internal.hpp"
using namespace xpbd;
#define TODO(name) return fail(XPBD_ERR_STATE, name ": not implemented yet")
extern "C" {
int xpbd_quickhull(const double*, int32_t, double*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, int32_t*, double*) { TODO("xpbd_quickhull"); }
int xpbd_solver_step(xpbd_rigid*, const xpbd_polytope*, double, int32_t) { TODO("xpbd_solver_step"); }
int xpbd_world_integrate(xpbd_rigid*, xpbd_rigid*, double, const xpbd_polytope*, const xpbd_polytope*) { TODO("xpbd_world_integrate"); }
}
