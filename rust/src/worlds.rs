//! The batched form the reference does not have: N independent worlds resident on one B200 as fp32
//! structure-of-arrays, stepped by the CUDA kernels; one `Worlds` per GPU, sharded by world index.
use crate::{ffi, geometry::Polytope, rigid::Rigid};

/// Owns the flattened arrays an `xpbd_polytope` view borrows.
pub struct PolytopeFfi {
    _vertices: Vec<f64>,
    _edges: Vec<i32>,
    _face_offsets: Vec<i32>,
    _face_indices: Vec<i32>,
    pub view: ffi::xpbd_polytope,
}

impl Polytope {
    pub fn to_ffi(&self) -> PolytopeFfi {
        let vertices: Vec<f64> = self.vertices.iter().flat_map(|v| [v.x, v.y, v.z]).collect();
        let edges: Vec<i32> = self.edges.iter().flat_map(|&(a, b)| [a as i32, b as i32]).collect();
        let mut face_offsets = vec![0i32];
        let mut face_indices = Vec::new();
        for f in &self.faces {
            face_indices.extend(f.iter().map(|&i| i as i32));
            face_offsets.push(face_indices.len() as i32);
        }
        let view = ffi::xpbd_polytope {
            vertices: vertices.as_ptr(),
            n_vertices: self.vertices.len() as i32,
            edges: edges.as_ptr(),
            n_edges: self.edges.len() as i32,
            face_offsets: face_offsets.as_ptr(),
            face_indices: face_indices.as_ptr(),
            n_faces: self.faces.len() as i32,
            centroid: [self.centroid.x, self.centroid.y, self.centroid.z],
        };
        PolytopeFfi { _vertices: vertices, _edges: edges, _face_offsets: face_offsets, _face_indices: face_indices, view }
    }
}

pub struct Worlds {
    handle: *mut ffi::xpbd_batch,
    pub n_worlds: usize,
    pub bodies_per_world: usize,
}

impl Worlds {
    /// `device` = CUDA ordinal; with G GPUs, rank r owns worlds [r * n / G, (r + 1) * n / G) and builds its own
    /// `Worlds` of that size — there is no collective on the step path.
    pub fn new(device: i32, n_worlds: usize, bodies_per_world: usize, joints_per_world: usize, body_body_contacts: bool) -> Worlds {
        let desc = ffi::xpbd_batch_desc {
            device,
            n_worlds: n_worlds as i64,
            bodies_per_world: bodies_per_world as i32,
            joints_per_world: joints_per_world as i32,
            body_body_contacts: body_body_contacts as i32,
            compliance: 1e-6, // src/solver.rs:20
            friction: 1.0,    // src/collision.rs:29
        };
        let mut handle = std::ptr::null_mut();
        ffi::check(unsafe { ffi::xpbd_batch_create(&desc, &mut handle) });
        Worlds { handle, n_worlds, bodies_per_world }
    }

    pub fn add_hull(&mut self, polytope: &Polytope) -> i32 {
        let h = polytope.to_ffi();
        let mut id = -1;
        ffi::check(unsafe { ffi::xpbd_batch_add_hull(self.handle, &h.view, &mut id) });
        id
    }

    pub fn upload(&mut self, first_world: usize, rigids: &[Rigid]) {
        let recs: Vec<ffi::xpbd_rigid> = rigids.iter().map(|r| r.to_ffi()).collect();
        let n = recs.len() / self.bodies_per_world;
        ffi::check(unsafe { ffi::xpbd_batch_upload(self.handle, first_world as i64, n as i64, recs.as_ptr()) });
    }

    pub fn download(&mut self, first_world: usize, rigids: &mut [Rigid]) {
        let mut recs: Vec<ffi::xpbd_rigid> = rigids.iter().map(|r| r.to_ffi()).collect();
        let n = recs.len() / self.bodies_per_world;
        ffi::check(unsafe { ffi::xpbd_batch_download(self.handle, first_world as i64, n as i64, recs.as_mut_ptr()) });
        for (r, rec) in rigids.iter_mut().zip(&recs) {
            r.update_from_ffi(rec);
        }
    }

    /// Hull id of every body of worlds [first_world, ...): `ids.len()` = n_worlds x bodies_per_world (mixed-hull scenes, C4).
    pub fn set_hulls(&mut self, first_world: usize, ids: &[i32]) {
        let n = ids.len() / self.bodies_per_world;
        ffi::check(unsafe { ffi::xpbd_batch_set_hulls(self.handle, first_world as i64, n as i64, ids.as_ptr()) });
    }

    /// Distance / attachment joints (`Constraint.distance`, src/constraint.rs:9) of worlds [first_world, ...):
    /// `joints.len()` = n_worlds x joints_per_world; `body_a < 0` marks an unused slot, `body_b < 0` a world-space anchor.
    pub fn set_joints(&mut self, first_world: usize, joints: &[ffi::xpbd_joint], joints_per_world: usize) {
        let n = joints.len() / joints_per_world;
        ffi::check(unsafe { ffi::xpbd_batch_set_joints(self.handle, first_world as i64, n as i64, joints.as_ptr()) });
    }

    /// World ranges the multi-body pipeline steps concurrently (0 = automatic); results do not depend on it.
    pub fn set_sub_batches(&mut self, count: i32) {
        ffi::check(unsafe { ffi::xpbd_batch_set_sub_batches(self.handle, count) });
    }

    /// The rewind buffer of the app (`states: Vec<(World, DebugLines)>`, src/app.rs:48,146-148) without the 304-byte
    /// records: 13 packed fp32 values per body (position, rotation, velocity, angular velocity).
    pub fn download_dynamic(&mut self, first_world: usize, n_worlds: usize) -> Vec<f32> {
        let mut out = vec![0f32; n_worlds * self.bodies_per_world * 13];
        ffi::check(unsafe { ffi::xpbd_batch_download_dynamic_f32(self.handle, first_world as i64, n_worlds as i64, out.as_mut_ptr()) });
        out
    }

    pub fn upload_dynamic(&mut self, first_world: usize, snapshot: &[f32]) {
        let n = snapshot.len() / (self.bodies_per_world * 13);
        ffi::check(unsafe { ffi::xpbd_batch_upload_dynamic_f32(self.handle, first_world as i64, n as i64, snapshot.as_ptr()) });
    }

    /// `World::integrate` for callers whose state lives in host memory as the reference's own records: uploads
    /// `rigids` (all worlds), steps, and writes the new velocity / angular velocity / position / rotation back; every
    /// other field keeps its f64 value (xpbd_batch_step_host_rigid; the copies are pipelined per world range).
    pub fn integrate_host(&mut self, rigids: &mut [Rigid], dt: f64, substep_count: usize, n_frames: usize) {
        assert_eq!(rigids.len(), self.n_worlds * self.bodies_per_world);
        let mut recs: Vec<ffi::xpbd_rigid> = rigids.iter().map(|r| r.to_ffi()).collect();
        let p = recs.as_mut_ptr();
        ffi::check(unsafe { ffi::xpbd_batch_step_host_rigid(self.handle, dt, substep_count as i32, n_frames as i32, p, p) });
        ffi::check(unsafe { ffi::xpbd_batch_sync(self.handle) });
        for (r, rec) in rigids.iter_mut().zip(&recs) {
            r.update_from_ffi(rec);
        }
    }

    /// `World::integrate` for every world: `n_frames` frames of `substep_count` substeps.
    pub fn integrate(&mut self, dt: f64, substep_count: usize, n_frames: usize) {
        ffi::check(unsafe { ffi::xpbd_batch_step(self.handle, dt, substep_count as i32, n_frames as i32) });
        ffi::check(unsafe { ffi::xpbd_batch_sync(self.handle) });
    }

    pub fn stats(&mut self) -> ffi::xpbd_step_stats {
        let mut s = ffi::xpbd_step_stats::default();
        ffi::check(unsafe { ffi::xpbd_batch_last_stats(self.handle, &mut s) });
        s
    }
}

impl Drop for Worlds {
    fn drop(&mut self) {
        unsafe { ffi::xpbd_batch_destroy(self.handle) };
    }
}
