/* xpbd_b200.h — C ABI of the B200-native batched XPBD substep path.
 *
 * This is the drop-in boundary for jim-ec/constraint_solver's step path.  The reference has no
 * FFI of its own (it is an in-process Rust module API); each entry point
 * below names the Rust item whose work it takes over.  The Rust-side shim that binds these
 * symbols is in rust/ and INTEGRATION.md.  Plain pointers and sizes only: no torch, no C++ types.
 *
 * Conventions
 *  - Every call returns XPBD_OK (0) or a negative xpbd_status; xpbd_last_error() gives the text
 *    of the calling thread's most recent failure.  Nothing throws or aborts across the boundary
 *    (the reference panics instead: src/rigid.rs:59, src/geometry.rs:264,280).
 *  - A batch handle is bound to one CUDA device and is NOT thread-safe: one handle per host
 *    thread / per GPU (worlds are sharded across GPUs by world index, no collective on the
 *    step path).
 *  - Host-side records are f64 in the reference's own field order; device state is fp32
 *    structure-of-arrays.  Upload rounds f64 -> f32 (round-to-nearest-even), download widens.
 *  - There is no CPU fallback: without a CUDA device every compute entry point fails with
 *    XPBD_ERR_CUDA.
 */
#ifndef XPBD_B200_H
#define XPBD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum xpbd_status {
    XPBD_OK = 0,
    XPBD_ERR_INVALID = -1,   /* bad argument (null, out of range, malformed hull ...) */
    XPBD_ERR_CUDA = -2,      /* CUDA runtime error or no device */
    XPBD_ERR_CAPACITY = -3,  /* a fixed capacity (hull size, contacts per world, shared memory) exceeded */
    XPBD_ERR_SINGULAR = -4,  /* inertia tensor not invertible (src/rigid.rs:59 panics here) */
    XPBD_ERR_STATE = -5      /* call not valid in the handle's current state */
} xpbd_status;

/* Mirror of `struct Rigid` (src/rigid.rs:7-50) without `color` (host-only, src/app.rs:228):
 * 38 doubles, same field order, #[repr(C)]-compatible. */
typedef struct xpbd_rigid {
    double inverse_mass;
    double inverse_inertia[9]; /* cgmath Matrix3, column-major: x.x x.y x.z y.x ... */
    double external_force[3];
    double internal_force[3];
    double external_torque[3];
    double internal_torque[3];
    double velocity[3];
    double angular_velocity[3];
    double center_of_mass[3];
    double position[3];
    double rotation[4];        /* cgmath Quaternion field order: v.x v.y v.z s */
} xpbd_rigid;

/* Mirror of `struct Frame` (src/frame.rs:7-10). */
typedef struct xpbd_frame {
    double position[3];
    double rotation[4];        /* v.x v.y v.z s */
} xpbd_frame;

/* Mirror of `struct Constraint` (src/constraint.rs:6-10), extended to two bodies
 * (SURVEY.md §8 row F2): rigid_b < 0 means "world" (the reference's single-body form). */
typedef struct xpbd_constraint {
    int32_t rigid;
    int32_t rigid_b;
    double contact0[3];
    double contact1[3];
    double distance;
} xpbd_constraint;

/* Mirror of `struct RigidMetrics` (src/geometry/integrate.rs:18-24). */
typedef struct xpbd_rigid_metrics {
    double mass;
    double volume;
    double center_of_mass[3];
    double inertia_tensor[9];  /* column-major */
} xpbd_rigid_metrics;

/* Borrowed view of a `Polytope` (src/geometry.rs:81-93). Faces are CSR: face f uses
 * face_indices[face_offsets[f] .. face_offsets[f+1]). */
typedef struct xpbd_polytope {
    const double* vertices;     /* n_vertices x 3 */
    int32_t n_vertices;
    const int32_t* edges;       /* n_edges x 2 */
    int32_t n_edges;
    const int32_t* face_offsets; /* n_faces + 1 */
    const int32_t* face_indices;
    int32_t n_faces;
    double centroid[3];
} xpbd_polytope;

/* A joint: two-body distance constraint between body-local anchor points
 * (Constraint.distance, src/constraint.rs:9). body_b < 0 anchors to a
 * fixed world-space point given in anchor_b. */
typedef struct xpbd_joint {
    int32_t body_a;
    int32_t body_b;
    double anchor_a[3];
    double anchor_b[3];
    double distance;
} xpbd_joint;

typedef struct xpbd_batch xpbd_batch; /* opaque */

typedef struct xpbd_batch_desc {
    int32_t device;             /* CUDA ordinal */
    int64_t n_worlds;
    int32_t bodies_per_world;   /* 1 = the reference's live path (body vs ground only) */
    int32_t joints_per_world;   /* 0 for none */
    int32_t body_body_contacts; /* 0: ground only (src/solver.rs:12); 1: SAT narrowphase between bodies */
    double compliance;          /* src/solver.rs:20 hard-codes 1e-6 */
    double friction;            /* src/collision.rs:29 hard-codes 1.0 */
} xpbd_batch_desc;

typedef struct xpbd_step_stats {
    uint64_t body_substeps;     /* worlds x bodies x substeps x frames of the last step call */
    uint64_t constraints;       /* constraints applied in the last step call (all worlds) */
    uint64_t overflow_worlds;   /* worlds that hit the per-world contact capacity (0 in a valid run) */
    float elapsed_ms;           /* device time of the last step call (CUDA events on the batch stream) */
    int32_t kernel_launches;    /* kernels launched by the last step call */
    /* Multi-body staged pipeline only (0 otherwise): device time of each stage's kernel in the LAST substep of the
     * call (prepare, sat, manifold, solve) and the number of candidate body pairs that substep's SAT stage processed. */
    float stage_ms[4];
    uint64_t candidate_pairs;
    int32_t sub_batches;        /* world ranges the last step call ran concurrently (see xpbd_batch_set_sub_batches) */
    int32_t nan_worlds;         /* worlds whose state holds a NaN / Inf after the last step call (the reference panics or
                                 * silently propagates; here the world is isolated — neighbours are unaffected — and counted) */
} xpbd_step_stats;

/* ---- library / device ---------------------------------------------------------------- */
const char* xpbd_last_error(void);
const char* xpbd_version(void);
int xpbd_device_count(int32_t* count);

/* ---- setup-time host math (f64, follows the reference exactly; no GPU needed) -------- */
/* Polytope::rigid_metrics (src/geometry.rs:291) -> integrate::rigid_metrics (src/geometry/integrate.rs:26-75) */
int xpbd_polytope_rigid_metrics(const xpbd_polytope* hull, double density, xpbd_rigid_metrics* out);
/* Rigid::new (src/rigid.rs:53-71); XPBD_ERR_SINGULAR where the reference panics */
int xpbd_rigid_new(const xpbd_rigid_metrics* metrics, xpbd_rigid* out);
/* Polytope::plane for every face (src/geometry.rs:256-271); out = n_faces x 4 (normal, displacement) */
int xpbd_polytope_planes(const xpbd_polytope* hull, double* out);
/* QuickHull: convex hull of a point cloud -> Polytope arrays (README.md:20).
 * Caller provides capacities; counts are written back.  Coplanar triangles are merged into polygons. */
int xpbd_quickhull(const double* points, int32_t n_points, double* vertices, int32_t* n_vertices,
                   int32_t* edges, int32_t* n_edges, int32_t* face_offsets, int32_t* face_indices,
                   int32_t* n_faces, int32_t* n_face_indices, double centroid[3]);

/* ---- batch lifetime ---------------------------------------------------------------------- */
int xpbd_batch_create(const xpbd_batch_desc* desc, xpbd_batch** out);
int xpbd_batch_destroy(xpbd_batch* batch);
/* Registers a hull in the batch's device hull table (validates: >= 4 vertices, every face >= 3
 * vertices, indices in range).  hull ids are dense from 0. */
int xpbd_batch_add_hull(xpbd_batch* batch, const xpbd_polytope* hull, int32_t* hull_id);
/* Hull id for every body of worlds [first_world, first_world + n_worlds): n_worlds x bodies_per_world.
 * ids == NULL assigns hull 0 everywhere. */
int xpbd_batch_set_hulls(xpbd_batch* batch, int64_t first_world, int64_t n_worlds, const int32_t* ids);
/* Joints of worlds [first_world, ...): n_worlds x joints_per_world records. */
int xpbd_batch_set_joints(xpbd_batch* batch, int64_t first_world, int64_t n_worlds, const xpbd_joint* joints);
/* Use the caller's CUDA stream (a cudaStream_t passed as void*) for every later launch/copy. */
int xpbd_batch_set_stream(xpbd_batch* batch, void* cuda_stream);
/* Multi-body staged pipeline: number of contiguous world ranges ("sub-batches") a step call runs concurrently, each
 * as its own chain of launches on an internal stream forked from / joined into the batch stream.  Worlds are
 * independent (src/world.rs:41-42), so results do not depend on it; kernels of different stages then share the SMs.
 * 0 = automatic (one range per 2^18 bodies, at most 4), 1 = one chain of launches on the batch stream (what
 * xpbd_step_stats.stage_ms needs to time each kernel alone), up to 8. */
int xpbd_batch_set_sub_batches(xpbd_batch* batch, int32_t count);

/* ---- state: Rigid AoS f64 <-> fp32 SoA in HBM ------------------------------------------ */
/* rigids: n_worlds x bodies_per_world records, world-major. */
int xpbd_batch_upload(xpbd_batch* batch, int64_t first_world, int64_t n_worlds, const xpbd_rigid* rigids);
int xpbd_batch_download(xpbd_batch* batch, int64_t first_world, int64_t n_worlds, xpbd_rigid* rigids);
/* Replicates one world's bodies (bodies_per_world records) into every world on the device. */
int xpbd_batch_upload_broadcast(xpbd_batch* batch, const xpbd_rigid* world_bodies);
/* Dynamic state only (position, rotation, velocity, angular velocity) as packed fp32, 13 floats per
 * body: px py pz qx qy qz qs vx vy vz wx wy wz.  The cheap path for per-frame snapshots
 * (the reference's rewind buffer, src/app.rs:48,204-215). */
int xpbd_batch_download_dynamic_f32(xpbd_batch* batch, int64_t first_world, int64_t n_worlds, float* out);
int xpbd_batch_upload_dynamic_f32(xpbd_batch* batch, int64_t first_world, int64_t n_worlds, const float* in);

/* ---- the hot path ---------------------------------------------------------------------- */
/* solver::step (src/solver.rs:3-17) for every body of every world, `n_frames` times in a row
 * (World::integrate, src/world.rs:34-43, is one frame with substep_count 25).
 * Asynchronous on the batch stream; all substeps of all frames run fused in one launch. */
int xpbd_batch_step(xpbd_batch* batch, double dt, int32_t substep_count, int32_t n_frames);
/* The same for callers whose state lives in HOST memory (the reference's `World` values, src/app.rs:204-215): uploads
 * the packed dynamic state `in` (13 floats per body, the layout of xpbd_batch_upload_dynamic_f32; NULL = keep the device
 * state), steps, and downloads the result into `out` (NULL = none).  Asynchronous: `out` is valid after xpbd_batch_sync.
 * When the staged pipeline runs several world ranges, each range's copies sit on that range's stream, so the host
 * copies of one range run under the kernels of the others.  `in` / `out` should be page-locked; pageable memory works but
 * serialises the copies. */
int xpbd_batch_step_host_f32(xpbd_batch* batch, double dt, int32_t substep_count, int32_t n_frames, const float* in,
                             float* out);
/* The same for callers holding the reference's own records (`Vec<Rigid>` / `World`, src/rigid.rs:7-50, src/app.rs:204-215:
 * 304 bytes per body): uploads `in` (all 38 fields, f64 -> f32; NULL = keep the device state), steps, and writes the records
 * into `out` (NULL = none; may equal `in`): the 13 values solver::step mutates at fp32 precision, every other field exactly
 * as it was uploaded through this call (the caller's f64 bits are not rounded; without an earlier `in` they are the device's
 * fp32 values).  Asynchronous like xpbd_batch_step_host_f32; copies are pipelined per world range. */
int xpbd_batch_step_host_rigid(xpbd_batch* batch, double dt, int32_t substep_count, int32_t n_frames, const xpbd_rigid* in,
                               xpbd_rigid* out);
int xpbd_batch_sync(xpbd_batch* batch);
/* Valid after xpbd_batch_sync. */
int xpbd_batch_last_stats(xpbd_batch* batch, xpbd_step_stats* out);

/* ---- single-call drop-ins (upload, one launch, download) ----------------------------- */
/* solver::step(&mut Rigid, &Polytope, dt, substep_count) (src/solver.rs:3) */
int xpbd_solver_step(xpbd_rigid* rigid, const xpbd_polytope* hull, double dt, int32_t substep_count);
/* World::integrate (src/world.rs:34-43): steps a and b, both against p1, 25 substeps. */
int xpbd_world_integrate(xpbd_rigid* a, xpbd_rigid* b, double dt, const xpbd_polytope* p1, const xpbd_polytope* p2);

/* ---- batched building blocks (each a CUDA kernel; for parity tests and reuse) -------- */
/* Rigid::integrate (src/rigid.rs:82-99) on n bodies in place. */
int xpbd_rigid_integrate_batch(xpbd_rigid* rigids, int64_t n, double dt);
/* Rigid::derive (src/rigid.rs:101-109): past = n x 7 doubles (position xyz, rotation v.x v.y v.z s). */
int xpbd_rigid_derive_batch(xpbd_rigid* rigids, const double* past, int64_t n, double dt);
/* Rigid::apply_impulse (src/rigid.rs:113-123): impulse_point = n x 6 doubles (impulse xyz, point xyz). */
int xpbd_rigid_apply_impulse_batch(xpbd_rigid* rigids, const double* impulse_point, int64_t n);
/* Rigid::frame (src/rigid.rs:75-80). */
int xpbd_rigid_frame_batch(const xpbd_rigid* rigids, int64_t n, xpbd_frame* out);
/* collision::ground (src/collision.rs:13-35): for body i, up to hull->n_vertices constraints are
 * written at out[i * n_vertices ...]; counts[i] = how many. */
int xpbd_ground_batch(const xpbd_rigid* rigids, const xpbd_frame* past, int64_t n, const xpbd_polytope* hull,
                      xpbd_constraint* out, int32_t* counts);
/* solver::solve (src/solver.rs:19-27): body i sweeps constraints[i * stride .. + counts[i]). */
int xpbd_solve_batch(xpbd_rigid* rigids, int64_t n, const xpbd_constraint* constraints, const int32_t* counts,
                     int32_t stride, double dt, double compliance);
/* xpbd_solve_batch that allocates nothing.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, same kernel and bit-identical results; the device-side working copies of rigids,
 * constraints and counts live in a static device array of 3.2 MiB inside the library (held by every GPU that loads the
 * library until the process exits) instead of cudaMalloc'ed buffers.  No cudaMalloc / cudaFree / allocating library
 * routine is called.  Limits: n <= XPBD_STATIC_MAX_BODIES and n * stride <= XPBD_STATIC_MAX_CONSTRAINTS, otherwise
 * XPBD_ERR_CAPACITY (xpbd_last_error says which).  All calls on one device share the array, which keeps the last call's
 * contents (every byte the kernel reads is overwritten by the call first): calls that overlap on the GPU — other host
 * threads, other streams — race; serialise them.  Runs on the current device's default stream and returns when done. */
#define XPBD_STATIC_MAX_BODIES 4096
#define XPBD_STATIC_MAX_CONSTRAINTS 32768
int xpbd_solve_batch_static(xpbd_rigid* rigids, int64_t n, const xpbd_constraint* constraints, const int32_t* counts,
                            int32_t stride, double dt, double compliance);
/* collision::face_axes_separation (src/collision.rs:123-149) on n frame pairs. */
int xpbd_face_axes_separation_batch(const xpbd_frame* frames_a, const xpbd_frame* frames_b, int64_t n,
                                    const xpbd_polytope* hull_a, const xpbd_polytope* hull_b, double* distance,
                                    int32_t* face);
/* collision::edge_axes_separation (src/collision.rs:151-197) on n frame pairs; edges: n x 2. */
int xpbd_edge_axes_separation_batch(const xpbd_frame* frames_a, const xpbd_frame* frames_b, int64_t n,
                                    const xpbd_polytope* hull_a, const xpbd_polytope* hull_b, double* distance,
                                    int32_t* edges);
/* The finished `sat` (src/collision.rs:37-121 + its commented plan): contact
 * manifold for n body pairs.  out: n x XPBD_MAX_MANIFOLD constraints (rigid = incident body 0/1,
 * rigid_b = the other), counts[i] points, feature[i] = 0 none, 1 face of A, 2 face of B, 3 edge-edge. */
#define XPBD_MAX_MANIFOLD 4
int xpbd_sat_manifold_batch(const xpbd_frame* frames_a, const xpbd_frame* frames_b, int64_t n,
                            const xpbd_polytope* hull_a, const xpbd_polytope* hull_b, xpbd_constraint* out,
                            int32_t* counts, int32_t* feature);

/* The SAT stage of the multi-body step (the three queries + feature choice, without the manifold) on n synthetic
 * candidate pairs, through the step's own kernels: schedule 0 = warp per candidate (any hulls), 1 = two candidates per
 * warp, 2 = thread per candidate (1, 2: cube-shaped hulls, 8 vertices / 12 edges / 6 faces); cull 0 = the reference's
 * full edge-pair scan (src/collision.rs:151-197), 1 = with the witness cull.  feature[i] as above, index[i] = face
 * index or flat edge pair index ia * n_edges_b + jb (-1 when separated).  Every (schedule, cull) combination must
 * return the same values: this is the A/B harness for that claim (tests/test_sat_fuzz.py). */
int xpbd_sat_feature_batch(const xpbd_frame* frames_a, const xpbd_frame* frames_b, int64_t n,
                           const xpbd_polytope* hull_a, const xpbd_polytope* hull_b, int32_t schedule, int32_t cull,
                           int32_t* feature, int32_t* index);

/* FP32 issue-rate microbenchmark (non-FMA FADD/FMUL chain) for the roofline denominator.
 * Returns achieved TFLOP/s on `device`. */
int xpbd_fp32_peak_microbench(int32_t device, int32_t use_fma, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* XPBD_B200_H */
