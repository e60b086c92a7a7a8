//! Raw bindings of include/xpbd_b200.h.  `#[repr(C)]` mirrors of the records; field order = the reference's own
//! structs (src/rigid.rs:7-50, src/frame.rs:7-10, src/constraint.rs:6-10, src/geometry/integrate.rs:18-24).
#![allow(non_camel_case_types, dead_code)]
use std::os::raw::{c_char, c_int, c_void};

pub const XPBD_OK: c_int = 0;
pub const XPBD_ERR_SINGULAR: c_int = -4;

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct xpbd_rigid {
    pub inverse_mass: f64,
    pub inverse_inertia: [f64; 9], // cgmath Matrix3, column-major
    pub external_force: [f64; 3],
    pub internal_force: [f64; 3],
    pub external_torque: [f64; 3],
    pub internal_torque: [f64; 3],
    pub velocity: [f64; 3],
    pub angular_velocity: [f64; 3],
    pub center_of_mass: [f64; 3],
    pub position: [f64; 3],
    pub rotation: [f64; 4], // v.x v.y v.z s
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct xpbd_frame {
    pub position: [f64; 3],
    pub rotation: [f64; 4],
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct xpbd_constraint {
    pub rigid: i32,
    pub rigid_b: i32,
    pub contact0: [f64; 3],
    pub contact1: [f64; 3],
    pub distance: f64,
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct xpbd_rigid_metrics {
    pub mass: f64,
    pub volume: f64,
    pub center_of_mass: [f64; 3],
    pub inertia_tensor: [f64; 9],
}

#[repr(C)]
pub struct xpbd_polytope {
    pub vertices: *const f64,
    pub n_vertices: i32,
    pub edges: *const i32,
    pub n_edges: i32,
    pub face_offsets: *const i32,
    pub face_indices: *const i32,
    pub n_faces: i32,
    pub centroid: [f64; 3],
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct xpbd_joint {
    pub body_a: i32,
    pub body_b: i32,
    pub anchor_a: [f64; 3],
    pub anchor_b: [f64; 3],
    pub distance: f64,
}

#[repr(C)]
#[derive(Debug, Clone, Copy)]
pub struct xpbd_batch_desc {
    pub device: i32,
    pub n_worlds: i64,
    pub bodies_per_world: i32,
    pub joints_per_world: i32,
    pub body_body_contacts: i32,
    pub compliance: f64,
    pub friction: f64,
}

#[repr(C)]
#[derive(Debug, Clone, Copy, Default)]
pub struct xpbd_step_stats {
    pub body_substeps: u64,
    pub constraints: u64,
    pub overflow_worlds: u64,
    pub elapsed_ms: f32,
    pub kernel_launches: i32,
    pub stage_ms: [f32; 4],
    pub candidate_pairs: u64,
    pub sub_batches: i32,
    pub nan_worlds: i32,
}

#[repr(C)]
pub struct xpbd_batch {
    _private: [u8; 0],
}

pub const XPBD_MAX_MANIFOLD: usize = 4;
pub const XPBD_STATIC_MAX_BODIES: i64 = 4096;
pub const XPBD_STATIC_MAX_CONSTRAINTS: i64 = 32768;

// Every entry point of include/xpbd_b200.h, in the header's order (tests/test_abi.py diffs names and argument counts).
extern "C" {
    pub fn xpbd_last_error() -> *const c_char;
    pub fn xpbd_version() -> *const c_char;
    pub fn xpbd_device_count(count: *mut i32) -> c_int;
    pub fn xpbd_polytope_rigid_metrics(hull: *const xpbd_polytope, density: f64, out: *mut xpbd_rigid_metrics) -> c_int;
    pub fn xpbd_rigid_new(metrics: *const xpbd_rigid_metrics, out: *mut xpbd_rigid) -> c_int;
    pub fn xpbd_polytope_planes(hull: *const xpbd_polytope, out: *mut f64) -> c_int;
    pub fn xpbd_quickhull(points: *const f64, n_points: i32, vertices: *mut f64, n_vertices: *mut i32, edges: *mut i32,
                          n_edges: *mut i32, face_offsets: *mut i32, face_indices: *mut i32, n_faces: *mut i32,
                          n_face_indices: *mut i32, centroid: *mut f64) -> c_int;
    pub fn xpbd_batch_create(desc: *const xpbd_batch_desc, out: *mut *mut xpbd_batch) -> c_int;
    pub fn xpbd_batch_destroy(batch: *mut xpbd_batch) -> c_int;
    pub fn xpbd_batch_add_hull(batch: *mut xpbd_batch, hull: *const xpbd_polytope, hull_id: *mut i32) -> c_int;
    pub fn xpbd_batch_set_hulls(batch: *mut xpbd_batch, first: i64, n: i64, ids: *const i32) -> c_int;
    pub fn xpbd_batch_set_joints(batch: *mut xpbd_batch, first: i64, n: i64, joints: *const xpbd_joint) -> c_int;
    pub fn xpbd_batch_set_stream(batch: *mut xpbd_batch, cuda_stream: *mut c_void) -> c_int;
    pub fn xpbd_batch_set_sub_batches(batch: *mut xpbd_batch, count: i32) -> c_int;
    pub fn xpbd_batch_step_host_f32(batch: *mut xpbd_batch, dt: f64, substep_count: i32, n_frames: i32, input: *const f32, output: *mut f32) -> c_int;
    pub fn xpbd_batch_step_host_rigid(batch: *mut xpbd_batch, dt: f64, substep_count: i32, n_frames: i32, input: *const xpbd_rigid, output: *mut xpbd_rigid) -> c_int;
    pub fn xpbd_batch_upload(batch: *mut xpbd_batch, first: i64, n: i64, rigids: *const xpbd_rigid) -> c_int;
    pub fn xpbd_batch_download(batch: *mut xpbd_batch, first: i64, n: i64, rigids: *mut xpbd_rigid) -> c_int;
    pub fn xpbd_batch_upload_broadcast(batch: *mut xpbd_batch, world_bodies: *const xpbd_rigid) -> c_int;
    pub fn xpbd_batch_download_dynamic_f32(batch: *mut xpbd_batch, first: i64, n: i64, out: *mut f32) -> c_int;
    pub fn xpbd_batch_upload_dynamic_f32(batch: *mut xpbd_batch, first: i64, n: i64, input: *const f32) -> c_int;
    pub fn xpbd_batch_step(batch: *mut xpbd_batch, dt: f64, substep_count: i32, n_frames: i32) -> c_int;
    pub fn xpbd_batch_sync(batch: *mut xpbd_batch) -> c_int;
    pub fn xpbd_batch_last_stats(batch: *mut xpbd_batch, out: *mut xpbd_step_stats) -> c_int;
    pub fn xpbd_solver_step(rigid: *mut xpbd_rigid, hull: *const xpbd_polytope, dt: f64, substep_count: i32) -> c_int;
    pub fn xpbd_world_integrate(a: *mut xpbd_rigid, b: *mut xpbd_rigid, dt: f64, p1: *const xpbd_polytope, p2: *const xpbd_polytope) -> c_int;
    pub fn xpbd_rigid_integrate_batch(rigids: *mut xpbd_rigid, n: i64, dt: f64) -> c_int;
    pub fn xpbd_rigid_derive_batch(rigids: *mut xpbd_rigid, past: *const f64, n: i64, dt: f64) -> c_int;
    pub fn xpbd_rigid_apply_impulse_batch(rigids: *mut xpbd_rigid, impulse_point: *const f64, n: i64) -> c_int;
    pub fn xpbd_rigid_frame_batch(rigids: *const xpbd_rigid, n: i64, out: *mut xpbd_frame) -> c_int;
    pub fn xpbd_ground_batch(rigids: *const xpbd_rigid, past: *const xpbd_frame, n: i64, hull: *const xpbd_polytope,
                             out: *mut xpbd_constraint, counts: *mut i32) -> c_int;
    pub fn xpbd_solve_batch(rigids: *mut xpbd_rigid, n: i64, constraints: *const xpbd_constraint, counts: *const i32,
                            stride: i32, dt: f64, compliance: f64) -> c_int;
    /// No counterpart in the reference: `xpbd_solve_batch` with its device working copies in a static device array
    /// (no cudaMalloc); at most 4096 bodies and 32768 constraint slots per call, calls must not overlap on the GPU.
    pub fn xpbd_solve_batch_static(rigids: *mut xpbd_rigid, n: i64, constraints: *const xpbd_constraint, counts: *const i32,
                                   stride: i32, dt: f64, compliance: f64) -> c_int;
    pub fn xpbd_face_axes_separation_batch(frames_a: *const xpbd_frame, frames_b: *const xpbd_frame, n: i64, hull_a: *const xpbd_polytope,
                                           hull_b: *const xpbd_polytope, distance: *mut f64, face: *mut i32) -> c_int;
    pub fn xpbd_edge_axes_separation_batch(frames_a: *const xpbd_frame, frames_b: *const xpbd_frame, n: i64, hull_a: *const xpbd_polytope,
                                           hull_b: *const xpbd_polytope, distance: *mut f64, edges: *mut i32) -> c_int;
    pub fn xpbd_sat_manifold_batch(frames_a: *const xpbd_frame, frames_b: *const xpbd_frame, n: i64, hull_a: *const xpbd_polytope,
                                   hull_b: *const xpbd_polytope, out: *mut xpbd_constraint, counts: *mut i32, feature: *mut i32) -> c_int;
    pub fn xpbd_sat_feature_batch(frames_a: *const xpbd_frame, frames_b: *const xpbd_frame, n: i64, hull_a: *const xpbd_polytope,
                                  hull_b: *const xpbd_polytope, schedule: i32, cull: i32, feature: *mut i32, index: *mut i32) -> c_int;
    pub fn xpbd_fp32_peak_microbench(device: i32, use_fma: i32, tflops: *mut f64) -> c_int;
}

/// Panics with the library's error text: the reference's own failure mode is a panic (src/rigid.rs:59).
pub fn check(status: c_int) {
    if status != XPBD_OK {
        let msg = unsafe { std::ffi::CStr::from_ptr(xpbd_last_error()) }.to_string_lossy().into_owned();
        panic!("xpbd_b200: {} (status {})", msg, status);
    }
}
