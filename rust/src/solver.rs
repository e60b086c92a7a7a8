//! Drop-in for src/solver.rs: `step` = every substep of a frame fused in ONE CUDA launch; `solve` = one
//! Gauss-Seidel sweep in `Vec` order.
use crate::{constraint::Constraint, ffi, geometry::Polytope, rigid::Rigid};

pub fn step(rigid: &mut Rigid, polytope: &Polytope, dt: f64, substep_count: usize) {
    let mut r = rigid.to_ffi();
    let hull = polytope.to_ffi(); // borrowed view: vertices / edges / CSR faces / centroid (see worlds.rs)
    ffi::check(unsafe { ffi::xpbd_solver_step(&mut r, &hull.view, dt, substep_count as i32) });
    rigid.update_from_ffi(&r);
}

pub fn solve(rigid: &mut Rigid, constraints: Vec<Constraint>, dt: f64) {
    let mut r = rigid.to_ffi();
    let list: Vec<ffi::xpbd_constraint> = constraints.iter().map(|c| c.to_ffi()).collect();
    let count = [list.len() as i32];
    let stride = list.len().max(1) as i32;
    let dummy = [ffi::xpbd_constraint { rigid: 0, rigid_b: -1, contact0: [0.0; 3], contact1: [0.0; 3], distance: 0.0 }];
    let ptr = if list.is_empty() { dummy.as_ptr() } else { list.as_ptr() };
    ffi::check(unsafe { ffi::xpbd_solve_batch(&mut r, 1, ptr, count.as_ptr(), stride, dt, 1e-6) });
    rigid.update_from_ffi(&r);
}
