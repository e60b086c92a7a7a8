//! Drop-in for src/collision.rs: `ground` and the SAT queries keep their names and signatures; each forwards to the
//! CUDA kernel that restates it (fp32, the reference's operation order and tie rules).  `sat` is finished here the way
//! its comments plan it (src/collision.rs:42-118): both face queries, the edge query, the feature of minimal
//! penetration and the clipped manifold; it keeps the reference's return sense (`true` = separated) and additionally
//! returns the contact constraints through `sat_manifold`.
use cgmath::Vector3;

use crate::{constraint::Constraint, debug, ffi, frame::Frame, geometry::Polytope, rigid::Rigid};

fn frame_to_ffi(f: &Frame) -> ffi::xpbd_frame {
    ffi::xpbd_frame {
        position: [f.position.x, f.position.y, f.position.z],
        rotation: [f.rotation.v.x, f.rotation.v.y, f.rotation.v.z, f.rotation.s],
    }
}

fn constraint_from_ffi(c: &ffi::xpbd_constraint) -> Constraint {
    Constraint {
        rigid: c.rigid as usize,
        contacts: (Vector3::from(c.contact0), Vector3::from(c.contact1)),
        distance: c.distance,
    }
}

/// collision::ground (src/collision.rs:13-35): constraints in hull-vertex order (= the Gauss-Seidel order).
#[allow(dead_code)]
pub fn ground(rigid: &Rigid, past: Frame, polytope: &Polytope) -> Vec<Constraint> {
    let r = rigid.to_ffi();
    let p = frame_to_ffi(&past);
    let hull = polytope.to_ffi();
    let nv = polytope.vertices.len();
    let blank = ffi::xpbd_constraint { rigid: 0, rigid_b: -1, contact0: [0.0; 3], contact1: [0.0; 3], distance: 0.0 };
    let mut out = vec![blank; nv];
    let mut count = 0i32;
    ffi::check(unsafe { ffi::xpbd_ground_batch(&r, &p, 1, &hull.view, out.as_mut_ptr(), &mut count) });
    out[..count as usize].iter().map(constraint_from_ffi).collect()
}

/// collision::face_axes_separation (src/collision.rs:123-149): (max separation, face index of polytopes.0).
pub fn face_axes_separation(frames: (Frame, Frame), polytopes: (&Polytope, &Polytope)) -> (f64, usize) {
    let (fa, fb) = (frame_to_ffi(&frames.0), frame_to_ffi(&frames.1));
    let (ha, hb) = (polytopes.0.to_ffi(), polytopes.1.to_ffi());
    let (mut distance, mut face) = (0.0f64, 0i32);
    ffi::check(unsafe { ffi::xpbd_face_axes_separation_batch(&fa, &fb, 1, &ha.view, &hb.view, &mut distance, &mut face) });
    (distance, face as usize) // -1 (no face: usize::MAX in the reference) wraps to usize::MAX as well
}

/// collision::edge_axes_separation (src/collision.rs:151-197): (max separation, (edge of 0, edge of 1)).
pub fn edge_axes_separation(
    frames: (Frame, Frame),
    polytopes: (&Polytope, &Polytope),
    _debug: &mut debug::DebugLines,
) -> (f64, (usize, usize)) {
    let (fa, fb) = (frame_to_ffi(&frames.0), frame_to_ffi(&frames.1));
    let (ha, hb) = (polytopes.0.to_ffi(), polytopes.1.to_ffi());
    let mut distance = 0.0f64;
    let mut edges = [0i32; 2];
    ffi::check(unsafe { ffi::xpbd_edge_axes_separation_batch(&fa, &fb, 1, &ha.view, &hb.view, &mut distance, edges.as_mut_ptr()) });
    (distance, (edges[0] as usize, edges[1] as usize))
}

/// The finished narrowphase: contact constraints between the two bodies (`rigid` = 0 / 1: which body of the pair
/// carries `contacts.0`), empty when they are separated.
pub fn sat_manifold(frames: (Frame, Frame), polytopes: (&Polytope, &Polytope)) -> Vec<Constraint> {
    let (fa, fb) = (frame_to_ffi(&frames.0), frame_to_ffi(&frames.1));
    let (ha, hb) = (polytopes.0.to_ffi(), polytopes.1.to_ffi());
    let blank = ffi::xpbd_constraint { rigid: 0, rigid_b: -1, contact0: [0.0; 3], contact1: [0.0; 3], distance: 0.0 };
    let mut out = [blank; ffi::XPBD_MAX_MANIFOLD];
    let (mut count, mut feature) = (0i32, 0i32);
    ffi::check(unsafe {
        ffi::xpbd_sat_manifold_batch(&fa, &fb, 1, &ha.view, &hb.view, out.as_mut_ptr(), &mut count, &mut feature)
    });
    out[..count as usize].iter().map(constraint_from_ffi).collect()
}

/// collision::sat (src/collision.rs:37-121): `true` when the bodies are SEPARATED (the reference's sense).
pub fn sat(frames: (Frame, Frame), polytopes: (&Polytope, &Polytope), _debug: &mut debug::DebugLines) -> bool {
    sat_manifold(frames, polytopes).is_empty()
}
