//! Drop-in for src/world.rs.  `World` stays `Copy` with public `a` and `b`, so the app's rewind buffer
//! (`states: Vec<(World, DebugLines)>`, src/app.rs:48) and `world.a.frame()` (src/app.rs:228) keep working.
//! `integrate` forwards to the CUDA library and keeps the reference's quirk of stepping BOTH bodies against `p1`
//! with 25 substeps (src/world.rs:41-42).
use cgmath::{Deg, Euler, Quaternion, Vector3};

use crate::{debug::DebugLines, ffi, geometry::Polytope, rigid::Rigid};

#[derive(Debug, Clone, Copy)]
pub struct World {
    pub a: Rigid,
    pub b: Rigid,
}

/// Initial conditions of one body of the default scene (values of src/world.rs:16-28).
struct Launch {
    density: f64,
    position: [f64; 3],
    velocity: [f64; 3],
    angular_velocity: [f64; 3],
    euler_deg: Option<[f64; 3]>,
}

const GRAVITY_FORCE_Z: f64 = -2.0; // a force, not an acceleration: a = F / m = -20 m/s^2 for the 0.1 kg cube

const BODY_A: Launch = Launch { density: 0.1, position: [0.0, 0.0, 4.0], velocity: [0.0, 2.5, 0.0], angular_velocity: [-4.0, 1.0, 0.0], euler_deg: None };
const BODY_B: Launch = Launch { density: 5.0, position: [4.0, 0.0, 4.0], velocity: [0.0, 0.0, 7.0], angular_velocity: [-5.0, 5.0, 0.0], euler_deg: Some([10.0, 15.0, 5.0]) };

fn launch(hull: &Polytope, l: &Launch) -> Rigid {
    let mut r = Rigid::new(hull.rigid_metrics(l.density));
    r.position = Vector3::from(l.position);
    r.velocity = Vector3::from(l.velocity);
    r.angular_velocity = Vector3::from(l.angular_velocity);
    r.external_force = Vector3::new(0.0, 0.0, GRAVITY_FORCE_Z);
    if let Some([x, y, z]) = l.euler_deg {
        r.rotation = Quaternion::from(Euler::new(Deg(x), Deg(y), Deg(z)));
    }
    r
}

impl World {
    pub fn new(p1: &Polytope, p2: &Polytope) -> World {
        World { a: launch(p1, &BODY_A), b: launch(p2, &BODY_B) }
    }

    /// One frame: `xpbd_world_integrate` uploads both bodies, runs all 25 substeps of both in one launch and
    /// downloads the result.  `p2` and `debug` are accepted and ignored, as in the reference.
    #[allow(unused_variables)]
    pub fn integrate(&mut self, dt: f64, p1: &Polytope, p2: &Polytope, debug: &mut DebugLines) {
        let mut bodies = [self.a.to_ffi(), self.b.to_ffi()];
        let (hull1, hull2) = (p1.to_ffi(), p2.to_ffi());
        let (first, second) = bodies.split_at_mut(1);
        ffi::check(unsafe { ffi::xpbd_world_integrate(&mut first[0], &mut second[0], dt, &hull1.view, &hull2.view) });
        self.a.update_from_ffi(&bodies[0]);
        self.b.update_from_ffi(&bodies[1]);
    }
}
