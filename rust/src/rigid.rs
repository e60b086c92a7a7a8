//! Drop-in for src/rigid.rs: same struct, same public fields, same method signatures.  The arithmetic of
//! `integrate`, `derive` and `apply_impulse` runs in the CUDA library (fp32, cgmath operation order); `frame` too.
//! A caller that steps whole scenes should use `worlds::Worlds` — these per-call forms pay a PCIe round trip each.
use cgmath::{Matrix3, Quaternion, Vector3, Zero};
use derive_setters::Setters;

use crate::{ffi, frame::Frame, geometry::integrate::RigidMetrics};

#[derive(Debug, Clone, Copy, Setters)]
pub struct Rigid {
    pub inverse_mass: f64,
    pub inverse_inertia: Matrix3<f64>,
    pub external_force: Vector3<f64>,
    pub internal_force: Vector3<f64>,
    pub external_torque: Vector3<f64>,
    pub internal_torque: Vector3<f64>,
    pub velocity: Vector3<f64>,
    pub angular_velocity: Vector3<f64>,
    pub center_of_mass: Vector3<f64>,
    pub position: Vector3<f64>,
    pub rotation: Quaternion<f64>,
    pub color: Option<[f32; 3]>,
}

fn v(a: Vector3<f64>) -> [f64; 3] { [a.x, a.y, a.z] }
fn w(a: [f64; 3]) -> Vector3<f64> { Vector3::new(a[0], a[1], a[2]) }

impl Rigid {
    pub(crate) fn to_ffi(&self) -> ffi::xpbd_rigid {
        let m = &self.inverse_inertia;
        ffi::xpbd_rigid {
            inverse_mass: self.inverse_mass,
            inverse_inertia: [m.x.x, m.x.y, m.x.z, m.y.x, m.y.y, m.y.z, m.z.x, m.z.y, m.z.z],
            external_force: v(self.external_force),
            internal_force: v(self.internal_force),
            external_torque: v(self.external_torque),
            internal_torque: v(self.internal_torque),
            velocity: v(self.velocity),
            angular_velocity: v(self.angular_velocity),
            center_of_mass: v(self.center_of_mass),
            position: v(self.position),
            rotation: [self.rotation.v.x, self.rotation.v.y, self.rotation.v.z, self.rotation.s],
        }
    }

    /// Only the dynamic part comes back from the device; constants and `color` are untouched.
    pub(crate) fn update_from_ffi(&mut self, r: &ffi::xpbd_rigid) {
        self.velocity = w(r.velocity);
        self.angular_velocity = w(r.angular_velocity);
        self.position = w(r.position);
        self.rotation = Quaternion::new(r.rotation[3], r.rotation[0], r.rotation[1], r.rotation[2]);
    }

    pub fn new(metrics: RigidMetrics) -> Rigid {
        let m = ffi::xpbd_rigid_metrics {
            mass: metrics.mass,
            volume: metrics.volume,
            center_of_mass: v(metrics.center_of_mass),
            inertia_tensor: {
                let t = &metrics.inertia_tensor;
                [t.x.x, t.x.y, t.x.z, t.y.x, t.y.y, t.y.z, t.z.x, t.z.y, t.z.z]
            },
        };
        let mut r: ffi::xpbd_rigid = unsafe { std::mem::zeroed() };
        let status = unsafe { ffi::xpbd_rigid_new(&m, &mut r) };
        if status == ffi::XPBD_ERR_SINGULAR {
            panic!("Inertia tensor is not invertible"); // the reference's message (src/rigid.rs:59)
        }
        ffi::check(status);
        let i = r.inverse_inertia;
        Rigid {
            inverse_mass: r.inverse_mass,
            inverse_inertia: Matrix3::new(i[0], i[1], i[2], i[3], i[4], i[5], i[6], i[7], i[8]),
            internal_force: Vector3::zero(),
            external_force: Vector3::zero(),
            internal_torque: Vector3::zero(),
            external_torque: Vector3::zero(),
            velocity: Vector3::zero(),
            angular_velocity: Vector3::zero(),
            center_of_mass: metrics.center_of_mass,
            position: Vector3::zero(),
            rotation: Quaternion::from_sv(1.0, Vector3::zero()),
            color: None,
        }
    }

    pub fn frame(&self) -> Frame {
        let r = self.to_ffi();
        let mut f = ffi::xpbd_frame { position: [0.0; 3], rotation: [0.0; 4] };
        ffi::check(unsafe { ffi::xpbd_rigid_frame_batch(&r, 1, &mut f) });
        Frame {
            position: w(f.position),
            rotation: Quaternion::new(f.rotation[3], f.rotation[0], f.rotation[1], f.rotation[2]),
        }
    }

    pub fn integrate(&mut self, dt: f64) {
        let mut r = self.to_ffi();
        ffi::check(unsafe { ffi::xpbd_rigid_integrate_batch(&mut r, 1, dt) });
        self.update_from_ffi(&r);
    }

    pub fn derive(&mut self, position: Vector3<f64>, rotation: Quaternion<f64>, dt: f64) {
        let mut r = self.to_ffi();
        let past = [position.x, position.y, position.z, rotation.v.x, rotation.v.y, rotation.v.z, rotation.s];
        ffi::check(unsafe { ffi::xpbd_rigid_derive_batch(&mut r, past.as_ptr(), 1, dt) });
        self.update_from_ffi(&r);
    }

    pub fn apply_impulse(&mut self, impulse: Vector3<f64>, point: Vector3<f64>) {
        let mut r = self.to_ffi();
        let ip = [impulse.x, impulse.y, impulse.z, point.x, point.y, point.z];
        ffi::check(unsafe { ffi::xpbd_rigid_apply_impulse_batch(&mut r, ip.as_ptr(), 1) });
        self.update_from_ffi(&r);
    }
}
