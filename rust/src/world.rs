//! Drop-in for src/world.rs: `World` stays `Copy` with public `a` and `b` (the app's rewind buffer and
//! `world.a.frame()` keep working, src/app.rs:48,228).  `integrate` keeps the reference's quirk of stepping both
//! bodies against `p1` (src/world.rs:41-42).
use cgmath::{Deg, Euler};

use crate::{debug, ffi, geometry, rigid};

#[derive(Debug, Clone, Copy)]
pub struct World {
    pub a: rigid::Rigid,
    pub b: rigid::Rigid,
}

impl World {
    pub fn new(p1: &geometry::Polytope, p2: &geometry::Polytope) -> World {
        let mut a = rigid::Rigid::new(p1.rigid_metrics(0.1));
        let mut b = rigid::Rigid::new(p2.rigid_metrics(5.0));
        a.position.z = 4.0;
        a.velocity.y = 2.5;
        a.angular_velocity.x = -4.0;
        a.angular_velocity.y = 1.0;
        a.external_force.z = -2.0;
        b.position.x = 4.0;
        b.position.z = 4.0;
        b.velocity.z = 7.0;
        b.angular_velocity.x = -5.0;
        b.angular_velocity.y = 5.0;
        b.external_force.z = -2.0;
        b.rotation = Euler::new(Deg(10.0), Deg(15.0), Deg(5.0)).into();
        World { a, b }
    }

    #[allow(unused)]
    pub fn integrate(&mut self, dt: f64, p1: &geometry::Polytope, p2: &geometry::Polytope, debug: &mut debug::DebugLines) {
        let (mut a, mut b) = (self.a.to_ffi(), self.b.to_ffi());
        let (h1, h2) = (p1.to_ffi(), p2.to_ffi());
        ffi::check(unsafe { ffi::xpbd_world_integrate(&mut a, &mut b, dt, &h1.view, &h2.view) });
        self.a.update_from_ffi(&a);
        self.b.update_from_ffi(&b);
    }
}
