// ORACLE — TEST INFRASTRUCTURE ONLY (see xo_math.hpp header). PARITY UNPINNED.
// Flat C entry points over the templated restatement so that tests/ and bench.py's
// cpu_baseline / --impl reference legs can drive it through ctypes.
//
// Body record layout ("rigid38"): 38 scalars in the field order of src/rigid.rs:7-50
//   [0] inverse_mass  [1..9] inverse_inertia (column-major)  [10..12] external_force
//   [13..15] internal_force [16..18] external_torque [19..21] internal_torque
//   [22..24] velocity [25..27] angular_velocity [28..30] center_of_mass [31..33] position
//   [34..37] rotation as cgmath stores it: v.x v.y v.z s
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "xo_physics.hpp"
#include "xo_world.hpp"

using namespace xo;

namespace {

template <class S>
Rigid<S> load_rigid(const S* r) {
    Rigid<S> b;
    b.inverse_mass = r[0];
    b.inverse_inertia.x = {r[1], r[2], r[3]};
    b.inverse_inertia.y = {r[4], r[5], r[6]};
    b.inverse_inertia.z = {r[7], r[8], r[9]};
    b.external_force = {r[10], r[11], r[12]};
    b.internal_force = {r[13], r[14], r[15]};
    b.external_torque = {r[16], r[17], r[18]};
    b.internal_torque = {r[19], r[20], r[21]};
    b.velocity = {r[22], r[23], r[24]};
    b.angular_velocity = {r[25], r[26], r[27]};
    b.center_of_mass = {r[28], r[29], r[30]};
    b.position = {r[31], r[32], r[33]};
    b.rotation = Quat<S>(r[37], r[34], r[35], r[36]);
    return b;
}
template <class S>
void store_rigid(const Rigid<S>& b, S* r) {
    r[0] = b.inverse_mass;
    for (int c = 0; c < 3; ++c) for (int k = 0; k < 3; ++k) r[1 + 3 * c + k] = b.inverse_inertia[c][k];
    auto put = [&](int o, Vec3<S> v) { r[o] = v.x; r[o + 1] = v.y; r[o + 2] = v.z; };
    put(10, b.external_force); put(13, b.internal_force); put(16, b.external_torque); put(19, b.internal_torque);
    put(22, b.velocity); put(25, b.angular_velocity); put(28, b.center_of_mass); put(31, b.position);
    r[34] = b.rotation.v.x; r[35] = b.rotation.v.y; r[36] = b.rotation.v.z; r[37] = b.rotation.s;
}

template <class S>
Polytope<S> load_hull(const S* verts, int n_verts, const int32_t* edges, int n_edges, const int32_t* face_offsets,
                      const int32_t* face_indices, int n_faces, const S* centroid) {
    Polytope<S> p;
    for (int i = 0; i < n_verts; ++i) p.vertices.push_back({verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]});
    for (int i = 0; i < n_edges; ++i) p.edges.push_back({edges[2 * i], edges[2 * i + 1]});
    for (int f = 0; f < n_faces; ++f)
        p.faces.emplace_back(face_indices + face_offsets[f], face_indices + face_offsets[f + 1]);
    if (centroid) p.centroid = {centroid[0], centroid[1], centroid[2]};
    return p;
}

template <class S>
Frame<S> load_frame(const S* f) {  // px py pz qx qy qz qs
    Frame<S> fr;
    fr.position = {f[0], f[1], f[2]};
    fr.rotation = Quat<S>(f[6], f[3], f[4], f[5]);
    return fr;
}

// Step n independent bodies (each its own world, as world.rs:41-42 steps a and b independently)
// for `frames` calls of solver::step. Static partition over `threads` host threads.
template <class S>
uint64_t step_bodies(S* rigids, int64_t n, const Polytope<S>& hull, S dt, int substeps, int frames, int threads) {
    if (threads < 1) threads = 1;
    std::vector<uint64_t> applied((size_t)threads, 0);
    auto work = [&](int t) {
        int64_t lo = n * t / threads, hi = n * (t + 1) / threads;
        uint64_t a = 0;
        for (int64_t i = lo; i < hi; ++i) {
            Rigid<S> b = load_rigid(rigids + 38 * i);
            for (int f = 0; f < frames; ++f) a += step(b, hull, dt, (size_t)substeps);
            store_rigid(b, rigids + 38 * i);
        }
        applied[(size_t)t] = a;
    };
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& th : pool) th.join();
    }
    uint64_t total = 0;
    for (uint64_t a : applied) total += a;
    return total;
}

}  // namespace

#define XO_HULL_ARGS(S) const S *verts, int n_verts, const int32_t *edges, int n_edges, const int32_t *face_offsets, \
                        const int32_t *face_indices, int n_faces, const S *centroid
#define XO_HULL_PASS verts, n_verts, edges, n_edges, face_offsets, face_indices, n_faces, centroid

extern "C" {

// solver::step on n independent bodies (solver.rs:3-17). Returns constraints applied.
uint64_t xo_step_bodies_f32(float* rigids, int64_t n, const float* verts, int n_verts, float dt, int substeps,
                            int frames, int threads) {
    Polytope<float> hull = load_hull<float>(verts, n_verts, nullptr, 0, nullptr, nullptr, 0, nullptr);
    return step_bodies<float>(rigids, n, hull, dt, substeps, frames, threads);
}
uint64_t xo_step_bodies_f64(double* rigids, int64_t n, const double* verts, int n_verts, double dt, int substeps,
                            int frames, int threads) {
    Polytope<double> hull = load_hull<double>(verts, n_verts, nullptr, 0, nullptr, nullptr, 0, nullptr);
    return step_bodies<double>(rigids, n, hull, dt, substeps, frames, threads);
}

// Rigid::integrate / derive / apply_impulse / frame, one body (rigid.rs:75-123)
void xo_rigid_integrate_f32(float* r, float dt) { auto b = load_rigid(r); b.integrate(dt); store_rigid(b, r); }
void xo_rigid_integrate_f64(double* r, double dt) { auto b = load_rigid(r); b.integrate(dt); store_rigid(b, r); }
void xo_rigid_derive_f32(float* r, const float* past_pos, const float* past_rot, float dt) {
    auto b = load_rigid(r);
    b.derive({past_pos[0], past_pos[1], past_pos[2]}, Quat<float>(past_rot[3], past_rot[0], past_rot[1], past_rot[2]), dt);
    store_rigid(b, r);
}
void xo_rigid_derive_f64(double* r, const double* past_pos, const double* past_rot, double dt) {
    auto b = load_rigid(r);
    b.derive({past_pos[0], past_pos[1], past_pos[2]}, Quat<double>(past_rot[3], past_rot[0], past_rot[1], past_rot[2]), dt);
    store_rigid(b, r);
}
void xo_rigid_apply_impulse_f32(float* r, const float* impulse, const float* point) {
    auto b = load_rigid(r);
    b.apply_impulse({impulse[0], impulse[1], impulse[2]}, {point[0], point[1], point[2]});
    store_rigid(b, r);
}
void xo_rigid_apply_impulse_f64(double* r, const double* impulse, const double* point) {
    auto b = load_rigid(r);
    b.apply_impulse({impulse[0], impulse[1], impulse[2]}, {point[0], point[1], point[2]});
    store_rigid(b, r);
}
void xo_rigid_frame_f32(const float* r, float* out7) {
    auto f = load_rigid(r).frame();
    out7[0] = f.position.x; out7[1] = f.position.y; out7[2] = f.position.z;
    out7[3] = f.rotation.v.x; out7[4] = f.rotation.v.y; out7[5] = f.rotation.v.z; out7[6] = f.rotation.s;
}
void xo_rigid_frame_f64(const double* r, double* out7) {
    auto f = load_rigid(r).frame();
    out7[0] = f.position.x; out7[1] = f.position.y; out7[2] = f.position.z;
    out7[3] = f.rotation.v.x; out7[4] = f.rotation.v.y; out7[5] = f.rotation.v.z; out7[6] = f.rotation.s;
}

// collision::ground (collision.rs:13-35). out: up to n_verts records of 7 scalars
// (contact0 xyz, contact1 xyz, distance). Returns the number of constraints.
int xo_ground_f32(const float* r, const float* past7, const float* verts, int n_verts, float* out) {
    Polytope<float> hull = load_hull<float>(verts, n_verts, nullptr, 0, nullptr, nullptr, 0, nullptr);
    auto cs = ground(load_rigid(r), load_frame(past7), hull);
    for (size_t i = 0; i < cs.size(); ++i) {
        float* o = out + 7 * i;
        o[0] = cs[i].contact0.x; o[1] = cs[i].contact0.y; o[2] = cs[i].contact0.z;
        o[3] = cs[i].contact1.x; o[4] = cs[i].contact1.y; o[5] = cs[i].contact1.z; o[6] = cs[i].distance;
    }
    return (int)cs.size();
}
int xo_ground_f64(const double* r, const double* past7, const double* verts, int n_verts, double* out) {
    Polytope<double> hull = load_hull<double>(verts, n_verts, nullptr, 0, nullptr, nullptr, 0, nullptr);
    auto cs = ground(load_rigid(r), load_frame(past7), hull);
    for (size_t i = 0; i < cs.size(); ++i) {
        double* o = out + 7 * i;
        o[0] = cs[i].contact0.x; o[1] = cs[i].contact0.y; o[2] = cs[i].contact0.z;
        o[3] = cs[i].contact1.x; o[4] = cs[i].contact1.y; o[5] = cs[i].contact1.z; o[6] = cs[i].distance;
    }
    return (int)cs.size();
}

// solver::solve (solver.rs:19-27) with an explicit constraint list (7 scalars each).
void xo_solve_f32(float* r, const float* cons, int n, float dt) {
    auto b = load_rigid(r);
    std::vector<Constraint<float>> cs((size_t)n);
    for (int i = 0; i < n; ++i) {
        const float* c = cons + 7 * i;
        cs[i].contact0 = {c[0], c[1], c[2]}; cs[i].contact1 = {c[3], c[4], c[5]}; cs[i].distance = c[6];
    }
    solve(b, cs, dt);
    store_rigid(b, r);
}
void xo_solve_f64(double* r, const double* cons, int n, double dt) {
    auto b = load_rigid(r);
    std::vector<Constraint<double>> cs((size_t)n);
    for (int i = 0; i < n; ++i) {
        const double* c = cons + 7 * i;
        cs[i].contact0 = {c[0], c[1], c[2]}; cs[i].contact1 = {c[3], c[4], c[5]}; cs[i].distance = c[6];
    }
    solve(b, cs, dt);
    store_rigid(b, r);
}

// Polytope::plane for every face (geometry.rs:256-271). out: n_faces x 4 (normal, displacement)
int xo_hull_planes_f64(XO_HULL_ARGS(double), double* out) {
    try {
        auto p = load_hull<double>(XO_HULL_PASS);
        for (int f = 0; f < n_faces; ++f) {
            auto pl = p.plane((size_t)f);
            out[4 * f] = pl.normal.x; out[4 * f + 1] = pl.normal.y; out[4 * f + 2] = pl.normal.z; out[4 * f + 3] = pl.displacement;
        }
    } catch (...) { return -1; }
    return 0;
}
int xo_hull_planes_f32(XO_HULL_ARGS(float), float* out) {
    try {
        auto p = load_hull<float>(XO_HULL_PASS);
        for (int f = 0; f < n_faces; ++f) {
            auto pl = p.plane((size_t)f);
            out[4 * f] = pl.normal.x; out[4 * f + 1] = pl.normal.y; out[4 * f + 2] = pl.normal.z; out[4 * f + 3] = pl.displacement;
        }
    } catch (...) { return -1; }
    return 0;
}

// integrate::rigid_metrics (integrate.rs:26-75). out[14]: mass, volume, com(3), inertia (9, column-major)
int xo_rigid_metrics_f64(XO_HULL_ARGS(double), double density, double* out) {
    try {
        auto p = load_hull<double>(XO_HULL_PASS);
        auto m = rigid_metrics(p, density);
        out[0] = m.mass; out[1] = m.volume;
        out[2] = m.center_of_mass.x; out[3] = m.center_of_mass.y; out[4] = m.center_of_mass.z;
        for (int c = 0; c < 3; ++c) for (int k = 0; k < 3; ++k) out[5 + 3 * c + k] = m.inertia_tensor[c][k];
    } catch (...) { return -1; }
    return 0;
}
// Rigid::new (rigid.rs:53-71): metrics -> rigid38. Returns -1 where the reference panics.
int xo_rigid_new_f64(const double* metrics14, double* rigid38) {
    RigidMetrics<double> m;
    m.mass = metrics14[0]; m.volume = metrics14[1];
    m.center_of_mass = {metrics14[2], metrics14[3], metrics14[4]};
    for (int c = 0; c < 3; ++c) for (int k = 0; k < 3; ++k) m.inertia_tensor[c][k] = metrics14[5 + 3 * c + k];
    try { store_rigid(Rigid<double>::from_metrics(m), rigid38); } catch (...) { return -1; }
    return 0;
}

// World::new (world.rs:12-31) with the app's hulls (app.rs:29-30): cube and 0.5 * tetrahedron.
// out: 2 x rigid38 (a then b).
void xo_world_new_f64(double* out76) {
    auto p1 = Polytope<double>::new_cube();
    auto p2 = Polytope<double>::new_tetrahedron().scaled(0.5);
    auto w = World<double>::create(p1, p2);
    store_rigid(w.a, out76);
    store_rigid(w.b, out76 + 38);
}
// World::integrate (world.rs:34-43), `frames` calls. Both bodies against the cube.
uint64_t xo_world_integrate_f64(double* ab76, double dt, int frames) {
    auto p1 = Polytope<double>::new_cube();
    World<double> w; w.a = load_rigid(ab76); w.b = load_rigid(ab76 + 38);
    uint64_t n = 0;
    for (int f = 0; f < frames; ++f) n += w.integrate(dt, p1, p1);
    store_rigid(w.a, ab76); store_rigid(w.b, ab76 + 38);
    return n;
}
uint64_t xo_world_integrate_f32(float* ab76, float dt, int frames) {
    auto p1 = Polytope<float>::new_cube();
    World<float> w; w.a = load_rigid(ab76); w.b = load_rigid(ab76 + 38);
    uint64_t n = 0;
    for (int f = 0; f < frames; ++f) n += w.integrate(dt, p1, p1);
    store_rigid(w.a, ab76); store_rigid(w.b, ab76 + 38);
    return n;
}

// Built-in hulls (geometry.rs:97-230). kind: 0 tetrahedron, 1 cube, 2 icosahedron.
// Any out pointer may be NULL; counts returned through n[3] = {verts, edges, faces}; face_indices total in n[3].
void xo_builtin_hull_f64(int kind, double* verts, int32_t* edges, int32_t* face_offsets, int32_t* face_indices,
                         double* centroid, int32_t* n) {
    Polytope<double> p = kind == 0 ? Polytope<double>::new_tetrahedron()
                        : kind == 1 ? Polytope<double>::new_cube() : Polytope<double>::new_icosahedron();
    int total = 0;
    for (auto& f : p.faces) total += (int)f.size();
    if (n) { n[0] = (int)p.vertices.size(); n[1] = (int)p.edges.size(); n[2] = (int)p.faces.size(); n[3] = total; }
    if (verts) for (size_t i = 0; i < p.vertices.size(); ++i) { verts[3 * i] = p.vertices[i].x; verts[3 * i + 1] = p.vertices[i].y; verts[3 * i + 2] = p.vertices[i].z; }
    if (edges) for (size_t i = 0; i < p.edges.size(); ++i) { edges[2 * i] = p.edges[i].first; edges[2 * i + 1] = p.edges[i].second; }
    if (face_offsets && face_indices) {
        int o = 0;
        for (size_t f = 0; f < p.faces.size(); ++f) {
            face_offsets[f] = o;
            for (int v : p.faces[f]) face_indices[o++] = v;
        }
        face_offsets[p.faces.size()] = o;
    }
    if (centroid) { centroid[0] = p.centroid.x; centroid[1] = p.centroid.y; centroid[2] = p.centroid.z; }
}

// collision::face_axes_separation (collision.rs:123-149). frames: 7 scalars each. face index -> *face.
float xo_face_axes_separation_f32(const float* fa, const float* fb, XO_HULL_ARGS(float), const float* verts_b,
                                  int n_verts_b, int32_t* face) {
    auto pa = load_hull<float>(XO_HULL_PASS);
    auto pb = load_hull<float>(verts_b, n_verts_b, nullptr, 0, nullptr, nullptr, 0, nullptr);
    auto q = face_axes_separation(load_frame(fa), load_frame(fb), pa, pb);
    *face = (int32_t)q.face;
    return q.distance;
}
double xo_face_axes_separation_f64(const double* fa, const double* fb, XO_HULL_ARGS(double), const double* verts_b,
                                   int n_verts_b, int32_t* face) {
    auto pa = load_hull<double>(XO_HULL_PASS);
    auto pb = load_hull<double>(verts_b, n_verts_b, nullptr, 0, nullptr, nullptr, 0, nullptr);
    auto q = face_axes_separation(load_frame(fa), load_frame(fb), pa, pb);
    *face = (int32_t)q.face;
    return q.distance;
}

// collision::edge_axes_separation (collision.rs:151-197). Hull A: verts/edges/centroid; hull B: verts/edges.
float xo_edge_axes_separation_f32(const float* fa, const float* fb, const float* verts_a, int n_verts_a,
                                  const int32_t* edges_a, int n_edges_a, const float* centroid_a, const float* verts_b,
                                  int n_verts_b, const int32_t* edges_b, int n_edges_b, int32_t* edge_pair) {
    auto pa = load_hull<float>(verts_a, n_verts_a, edges_a, n_edges_a, nullptr, nullptr, 0, centroid_a);
    auto pb = load_hull<float>(verts_b, n_verts_b, edges_b, n_edges_b, nullptr, nullptr, 0, nullptr);
    auto q = edge_axes_separation(load_frame(fa), load_frame(fb), pa, pb);
    edge_pair[0] = (int32_t)q.edge0; edge_pair[1] = (int32_t)q.edge1;
    return q.distance;
}
double xo_edge_axes_separation_f64(const double* fa, const double* fb, const double* verts_a, int n_verts_a,
                                   const int32_t* edges_a, int n_edges_a, const double* centroid_a,
                                   const double* verts_b, int n_verts_b, const int32_t* edges_b, int n_edges_b,
                                   int32_t* edge_pair) {
    auto pa = load_hull<double>(verts_a, n_verts_a, edges_a, n_edges_a, nullptr, nullptr, 0, centroid_a);
    auto pb = load_hull<double>(verts_b, n_verts_b, edges_b, n_edges_b, nullptr, nullptr, 0, nullptr);
    auto q = edge_axes_separation(load_frame(fa), load_frame(fb), pa, pb);
    edge_pair[0] = (int32_t)q.edge0; edge_pair[1] = (int32_t)q.edge1;
    return q.distance;
}

int xo_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"

#include "xo_world_capi.inc"
