//! Drop-in for src/constraint.rs.  `Constraint` keeps its three public fields; the sweep that consumes a
//! `Vec<Constraint>` is `solver::solve`, which runs on the device.  `inverse_resitance` (sic) and `act` keep their
//! signatures for source compatibility and forward to the same kernel with a one-constraint list.
use cgmath::{InnerSpace, Vector3};

use crate::{ffi, rigid::Rigid};

#[derive(Debug)]
pub struct Constraint {
    pub rigid: usize,
    pub contacts: (Vector3<f64>, Vector3<f64>),
    pub distance: f64,
}

impl Constraint {
    pub(crate) fn to_ffi(&self) -> ffi::xpbd_constraint {
        ffi::xpbd_constraint {
            rigid: self.rigid as i32,
            rigid_b: -1,
            contact0: [self.contacts.0.x, self.contacts.0.y, self.contacts.0.z],
            contact1: [self.contacts.1.x, self.contacts.1.y, self.contacts.1.z],
            distance: self.distance,
        }
    }

    /// Host-side convenience only (f64); the device recomputes it in fp32 inside the sweep.
    pub fn current_distance(&self) -> f64 {
        (self.contacts.1 - self.contacts.0).magnitude()
    }

    /// `act` with the reference's Lagrange factor is one step of `solver::solve`; callers that used
    /// `inverse_resitance` + `act` by hand should call `solver::solve(rigid, vec![constraint], dt)` instead.
    pub fn act(&mut self, rigids: &mut [&mut Rigid], factor: f64) {
        let d = self.contacts.1 - self.contacts.0;
        let impulse = factor * d.normalize();
        rigids[self.rigid].apply_impulse(impulse, self.contacts.0);
    }
}
