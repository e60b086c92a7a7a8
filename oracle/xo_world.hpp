// ORACLE — TEST INFRASTRUCTURE ONLY (see xo_math.hpp header). PARITY UNPINNED.
//
// Multi-body worlds.  The reference has NO algorithm here: its `sat` is
// unfinished dead code (src/collision.rs:37-121), `Constraint` indexes one body, and `World` steps
// its two bodies independently.  This file is the builder's specification of the finished path,
// built from the reference's own pieces and conventions:
//   * face_axes_separation / edge_axes_separation exactly as written (src/collision.rs:123-197),
//     wired the way the comments in `sat` plan it (A faces, B faces, edges, early-out when >= 0,
//     feature of minimal penetration), with Gregorius' relative/absolute bias towards face contacts;
//   * reference plane = Frame * Polytope::plane, incident face = most anti-parallel normal with the
//     reference's strict `<` scan (src/collision.rs:66-88);
//   * Sutherland-Hodgman clipping of the incident face against the side planes of the reference face,
//     points below the reference plane kept, reduced to <= 4 (README.md:18-19);
//   * `Constraint` generalised to two bodies (src/constraint.rs:7,25,34 were shaped for it):
//     lambda = c / (w_a + w_b + compliance/dt^2), +impulse on a, -impulse on b, with the reference's
//     inverse_resitance and apply_impulse per body.
// Per substep: save past, integrate all bodies, generate ALL constraints from the post-integrate poses
// (ground body-major in hull-vertex order, then body pairs i<j lexicographic, then joints), ONE
// Gauss-Seidel sweep in that order, derive all, then one velocity pass over the contacts that were active.
//   * Ground constraints are the reference's, unchanged: stale contact points and the position-level
//     friction offset (src/collision.rs:13-35, src/solver.rs:12-13).  With one body, no joints and
//     body_body = false the step reduces to solver::step bit for bit.
//   * Body-body contacts and joints are evaluated FRESH at sweep time from body-local anchors (Mueller et
//     al. 2020, README.md:14): measured on the 10-cube stack, stale two-body contacts cannot carry a stack
//     (each of the n stale points re-applies the full correction; the stack explodes within 10 frames).
//   * The velocity pass (Mueller 2020 "solveVelocities", restitution 0) removes the relative normal velocity
//     and `friction` x the relative tangential velocity at every active contact.  Without it the
//     penetration-recovery velocity of a compressed stack is kept by derive() and the stack bounces apart.
#pragma once
#include <algorithm>
#include <array>

#include "xo_physics.hpp"

namespace xo {

constexpr int kMaxManifold = 4;
constexpr int kMaxClip = 16;

template <class S>
struct Tuning {
    // Gregorius (GDC 2013/2015) feature-flip tolerances; part of this specification.
    static constexpr S rel_edge() { return S(0.90); }
    static constexpr S rel_face() { return S(0.98); }
    static constexpr S abs_tol() { return S(0.0025); }
};

enum ConstraintKind : int { kGround = 0, kContact = 1, kJoint = 2 };

// kGround : a, b = contact0, contact1 in WORLD space (stale, as the reference), distance = 0.
// kContact: a = point in body rigid_a's OBJECT space (the incident body), b = point on the reference plane in
//           body rigid_b's object space, normal = world-space unit direction a must move along.
// kJoint  : a, b = anchors in object space (b in world space when rigid_b < 0), distance = rest length.
template <class S>
struct Constraint2 {
    int rigid_a = 0;
    int rigid_b = -1;  // < 0: world
    int kind = kGround;
    Vec3<S> a{}, b{}, normal{};
    S distance{};
};

template <class S>
struct Joint {
    int body_a = -1, body_b = -1;  // body_a < 0: inactive slot; body_b < 0: anchor_b is a world point
    Vec3<S> anchor_a{}, anchor_b{};
    S distance{};
};

template <class S>
struct WorldParams {
    S compliance = S(1e-6);  // src/solver.rs:20
    S friction = S(1.0);     // src/collision.rs:29
    bool body_body = true;
};

// Hull as the world step sees it: polytope + cached local planes + bounding radius about the centroid.
template <class S>
struct Hull {
    Polytope<S> poly;
    std::vector<Plane<S>> planes;  // Polytope::plane(i), local space
    S radius{};
    static Hull make(const Polytope<double>& p64) {
        Hull h;
        h.poly = p64.template cast<S>();
        for (size_t f = 0; f < h.poly.faces.size(); ++f) h.planes.push_back(h.poly.plane(f));
        double r = 0;
        for (auto& v : p64.vertices) r = std::max(r, magnitude(v - p64.centroid));
        h.radius = S(r);  // computed in f64 from the registered hull, then rounded (same as the device table)
        return h;
    }
};

// Phases of a substep, for the flop bookkeeping of the Counted<T> instantiation (xo_count.hpp).
enum Phase : int { kPhIntegrate = 0, kPhGround, kPhBroadphase, kPhFaceQueries, kPhEdgeQuery, kPhManifold, kPhSweep, kPhDerive,
                   kPhVelocity, kPhCount };
template <class S>
struct StepCounters {
    uint64_t constraints = 0, pairs_tested = 0, manifolds = 0, active_contacts = 0;
    uint64_t flops[kPhCount][4] = {};  // add, mul, div, sqrt per phase (all zero unless S = Counted<T>)
};
// Adds the flops performed during its lifetime to one phase of the counters (compiles to nothing for plain scalars).
template <class S>
struct PhaseScope {
    StepCounters<S>* c;
    int phase;
    FlopCounters t0;
    PhaseScope(StepCounters<S>* c_, int phase_) : c(c_), phase(phase_), t0(FlopProbe<S>::now()) {}
    ~PhaseScope() {
        if (!c) return;
        const FlopCounters t1 = FlopProbe<S>::now();
        c->flops[phase][0] += t1.add - t0.add; c->flops[phase][1] += t1.mul - t0.mul;
        c->flops[phase][2] += t1.div - t0.div; c->flops[phase][3] += t1.sqrt - t0.sqrt;
    }
};

// ---- narrowphase ---------------------------------------------------------------------------
enum Feature : int { kNone = 0, kFaceA = 1, kFaceB = 2, kEdgeEdge = 3 };

template <class S>
struct Manifold {
    int feature = kNone;
    int count = 0;
    int incident = 0;  // 0: body A carries contact0, 1: body B
    std::array<Vec3<S>, kMaxManifold> point{};   // on the incident body (contact0)
    std::array<Vec3<S>, kMaxManifold> target{};  // on the other body, before friction
};

// Clip `poly` (n points) against the half-space dot(s, p - o) <= 0. Returns new count.
template <class S>
int clip_polygon(const Vec3<S>* in, int n, Vec3<S> o, Vec3<S> s, Vec3<S>* out) {
    int m = 0;
    for (int k = 0; k < n; ++k) {
        Vec3<S> p1 = in[k], p2 = in[(k + 1) % n];
        S d1 = dot(s, p1 - o), d2 = dot(s, p2 - o);
        bool in1 = d1 <= S(0), in2 = d2 <= S(0);
        if (in1 != in2) {
            S t = d1 / (d1 - d2);
            if (m < kMaxClip) out[m++] = p1 + (p2 - p1) * t;
        }
        if (in2) {
            if (m < kMaxClip) out[m++] = p2;
        }
    }
    return m;
}

// Face contact: reference body R (face r), incident body I.
template <class S>
void face_manifold(Frame<S> fr, Frame<S> fi, const Hull<S>& hr, const Hull<S>& hi, int r, Manifold<S>& out) {
    Plane<S> ref_plane = fr * hr.planes[(size_t)r];  // src/collision.rs:66
    // incident face: least dot, first strict minimum (src/collision.rs:74-85)
    S least = std::numeric_limits<S>::max();
    int inc = -1;
    for (size_t k = 0; k < hi.planes.size(); ++k) {
        Plane<S> pk = fi * hi.planes[k];
        S d = dot(pk.normal, ref_plane.normal);
        if (d < least) { least = d; inc = (int)k; }
    }
    if (inc < 0) return;
    Vec3<S> a[kMaxClip], b[kMaxClip];
    int n = 0;
    for (int v : hi.poly.faces[(size_t)inc])
        if (n < kMaxClip) a[n++] = fi * hi.poly.vertices[(size_t)v];
    // side planes of the reference face
    const std::vector<int>& rf = hr.poly.faces[(size_t)r];
    int nr = (int)rf.size();
    Vec3<S>*src = a, *dst = b;
    for (int m = 0; m < nr && n > 0; ++m) {
        Vec3<S> w0 = fr * hr.poly.vertices[(size_t)rf[m]];
        Vec3<S> w1 = fr * hr.poly.vertices[(size_t)rf[(m + 1) % nr]];
        Vec3<S> w2 = fr * hr.poly.vertices[(size_t)rf[(m + 2) % nr]];
        Vec3<S> s = cross(w1 - w0, ref_plane.normal);
        if (dot(s, w2 - w0) > S(0)) s = -s;  // outward: the rest of the (convex) face is on the inner side
        n = clip_polygon(src, n, w0, s, dst);
        std::swap(src, dst);
    }
    // keep points strictly below the reference plane
    Vec3<S> pts[kMaxClip];
    S depth[kMaxClip];
    int c = 0;
    for (int k = 0; k < n; ++k) {
        S d = ref_plane.distance(src[k]);
        if (d < S(0)) { pts[c] = src[k]; depth[c] = d; ++c; }
    }
    int keep[kMaxManifold];
    int nk = 0;
    if (c <= kMaxManifold) {
        for (int k = 0; k < c; ++k) keep[nk++] = k;
    } else {
        int i0 = 0;
        for (int k = 1; k < c; ++k) if (depth[k] < depth[i0]) i0 = k;
        int i1 = -1; S far = S(-1);
        for (int k = 0; k < c; ++k) {
            if (k == i0) continue;
            S d2 = magnitude2(pts[k] - pts[i0]);
            if (d2 > far) { far = d2; i1 = k; }
        }
        int i2 = -1, i3 = -1; S amax = S(0), amin = S(0);
        Vec3<S> e01 = pts[i1] - pts[i0];
        for (int k = 0; k < c; ++k) {
            if (k == i0 || k == i1) continue;
            S area = dot(cross(e01, pts[k] - pts[i0]), ref_plane.normal);
            if (area > amax) { amax = area; i2 = k; }
            if (area < amin) { amin = area; i3 = k; }
        }
        keep[nk++] = i0; keep[nk++] = i1;
        if (i2 >= 0) keep[nk++] = i2;
        if (i3 >= 0) keep[nk++] = i3;
        std::sort(keep, keep + nk);  // polygon order
    }
    for (int k = 0; k < nk; ++k) {
        Vec3<S> p = pts[keep[k]];
        out.point[(size_t)out.count] = p;
        out.target[(size_t)out.count] = p - depth[keep[k]] * ref_plane.normal;  // Plane::project
        ++out.count;
    }
}

template <class S>
inline S clamp01(S x) { return x < S(0) ? S(0) : (x > S(1) ? S(1) : x); }

// Closest points of two segments (Ericson, Real-Time Collision Detection 5.1.9), non-degenerate edges.
template <class S>
void edge_manifold(Vec3<S> p1, Vec3<S> q1, Vec3<S> p2, Vec3<S> q2, Manifold<S>& out) {
    Vec3<S> d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
    S a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
    S c = dot(d1, r), b = dot(d1, d2);
    S denom = a * e - b * b;
    S s = denom != S(0) ? clamp01((b * f - c * e) / denom) : S(0);
    S t = (b * s + f) / e;
    if (t < S(0)) { t = S(0); s = clamp01(-c / a); }
    else if (t > S(1)) { t = S(1); s = clamp01((b - c) / a); }
    out.point[0] = p1 + d1 * s;
    out.target[0] = p2 + d2 * t;
    out.count = 1;
}

// The finished `sat`: contact manifold between body A (frames.0, hulls.0) and body B.
template <class S>
Manifold<S> sat_manifold(Frame<S> fa, Frame<S> fb, const Hull<S>& ha, const Hull<S>& hb, StepCounters<S>* ctr = nullptr) {
    Manifold<S> m;
    FaceQuery<S> qa, qb;
    EdgeQuery<S> qe;
    {
        PhaseScope<S> ph(ctr, kPhFaceQueries);
        qa = face_axes_separation(fa, fb, ha.poly, hb.poly);
        if (qa.distance >= S(0)) return m;
        qb = face_axes_separation(fb, fa, hb.poly, ha.poly);
        if (qb.distance >= S(0)) return m;
    }
    {
        PhaseScope<S> ph(ctr, kPhEdgeQuery);
        qe = edge_axes_separation(fa, fb, ha.poly, hb.poly);
        if (qe.distance >= S(0)) return m;
    }
    PhaseScope<S> ph(ctr, kPhManifold);
    S face_max = qa.distance > qb.distance ? qa.distance : qb.distance;
    if (qe.edge0 >= 0 && qe.distance > Tuning<S>::rel_edge() * face_max + Tuning<S>::abs_tol()) {
        m.feature = kEdgeEdge;
        m.incident = 0;
        auto ea = ha.poly.edges[(size_t)qe.edge0];
        auto eb = hb.poly.edges[(size_t)qe.edge1];
        edge_manifold(fa * ha.poly.vertices[(size_t)ea.first], fa * ha.poly.vertices[(size_t)ea.second],
                      fb * hb.poly.vertices[(size_t)eb.first], fb * hb.poly.vertices[(size_t)eb.second], m);
    } else if (qb.distance > Tuning<S>::rel_face() * qa.distance + Tuning<S>::abs_tol()) {
        m.feature = kFaceB;
        m.incident = 0;  // B is the reference, A's face is clipped: contact0 on A
        face_manifold(fb, fa, hb, ha, (int)qb.face, m);
    } else {
        m.feature = kFaceA;
        m.incident = 1;  // A is the reference: contact0 on B
        face_manifold(fa, fb, ha, hb, (int)qa.face, m);
    }
    return m;
}

// One manifold point -> fresh contact constraint.  Dropped when point and target coincide (no direction).
template <class S>
bool contact_constraint(int inc, int oth, Frame<S> f_inc, Frame<S> f_oth, Vec3<S> point, Vec3<S> target,
                        Constraint2<S>& out) {
    Vec3<S> correction = target - point;
    if (magnitude2(correction) == S(0)) return false;
    out.rigid_a = inc; out.rigid_b = oth; out.kind = kContact;
    out.a = f_inc.inverse() * point;
    out.b = f_oth.inverse() * target;
    out.normal = normalize(correction);
    out.distance = S(0);
    return true;
}

// ---- the world step ------------------------------------------------------------------------
template <class S>
struct WorldState {
    std::vector<Rigid<S>> bodies;
    std::vector<int> hull_of;  // per body
    std::vector<Joint<S>> joints;
};

template <class S>
void generate_constraints(const WorldState<S>& w, const std::vector<Hull<S>>& hulls, const std::vector<Frame<S>>& past,
                          const WorldParams<S>& prm, std::vector<Constraint2<S>>& out, StepCounters<S>* ctr) {
    const int B = (int)w.bodies.size();
    std::vector<Frame<S>> frames((size_t)B);
    {
        PhaseScope<S> ph(ctr, kPhIntegrate);
        for (int b = 0; b < B; ++b) frames[(size_t)b] = w.bodies[(size_t)b].frame();
    }
    // 1. ground, body-major, hull vertex order (src/collision.rs:13-35 per body)
    for (int b = 0; b < B; ++b) {
        PhaseScope<S> ph(ctr, kPhGround);
        const Rigid<S>& rigid = w.bodies[(size_t)b];
        for (const Vec3<S>& vertex : hulls[(size_t)w.hull_of[(size_t)b]].poly.vertices) {
            Vec3<S> position = rigid.frame() * vertex;
            if (position.z >= S(0)) continue;
            Vec3<S> target{position.x, position.y, S(0)};
            Vec3<S> correction = target - position;
            Vec3<S> delta = rigid.frame().delta(past[(size_t)b], position);
            Vec3<S> delta_t = delta - project_on(delta, correction);
            Constraint2<S> c;
            c.rigid_a = b; c.rigid_b = -1; c.kind = kGround;
            c.a = position;
            c.b = target - prm.friction * delta_t;
            c.distance = S(0);
            out.push_back(c);
        }
    }
    // 2. body pairs
    if (prm.body_body) {
        for (int i = 0; i < B; ++i) {
            for (int j = i + 1; j < B; ++j) {
                const Hull<S>& hi = hulls[(size_t)w.hull_of[(size_t)i]];
                const Hull<S>& hj = hulls[(size_t)w.hull_of[(size_t)j]];
                {
                    PhaseScope<S> ph(ctr, kPhBroadphase);
                    Vec3<S> ci = frames[(size_t)i] * hi.poly.centroid, cj = frames[(size_t)j] * hj.poly.centroid;
                    S rsum = hi.radius + hj.radius;
                    if (magnitude2(cj - ci) > rsum * rsum) continue;  // bounding spheres
                }
                if (ctr) ctr->pairs_tested++;
                Manifold<S> m = sat_manifold(frames[(size_t)i], frames[(size_t)j], hi, hj, ctr);
                if (m.count && ctr) ctr->manifolds++;
                int inc = m.incident == 0 ? i : j, oth = m.incident == 0 ? j : i;
                PhaseScope<S> ph(ctr, kPhManifold);
                for (int k = 0; k < m.count; ++k) {
                    Constraint2<S> c;
                    if (contact_constraint(inc, oth, frames[(size_t)inc], frames[(size_t)oth], m.point[(size_t)k],
                                           m.target[(size_t)k], c))
                        out.push_back(c);
                }
            }
        }
    }
    // 3. joints
    for (const Joint<S>& j : w.joints) {
        if (j.body_a < 0) continue;
        Constraint2<S> c;
        c.rigid_a = j.body_a; c.rigid_b = j.body_b; c.kind = kJoint;
        c.a = j.anchor_a; c.b = j.anchor_b;
        c.distance = j.distance;
        out.push_back(c);
    }
}

template <class S>
S inverse_resistance_at(const Rigid<S>& r, Vec3<S> contact, Vec3<S> direction) {  // src/constraint.rs:25-32
    Vec3<S> angular_impulse = conjugate(r.rotation) * cross(contact - (r.position + r.center_of_mass), direction);
    return r.inverse_mass + dot(r.inverse_inertia * angular_impulse, angular_impulse);
}

// One Gauss-Seidel sweep. active[k] = 1 for every contact that was penetrating when it was solved.
template <class S>
void sweep(WorldState<S>& w, const std::vector<Constraint2<S>>& cs, S compliance, std::vector<char>& active) {
    active.assign(cs.size(), 0);
    for (size_t k = 0; k < cs.size(); ++k) {
        const Constraint2<S>& c = cs[k];
        Rigid<S>& A = w.bodies[(size_t)c.rigid_a];
        if (c.kind == kGround) {
            // solver::solve body (src/solver.rs:22-26) with the reference's Constraint methods
            Vec3<S> diff = c.b - c.a;
            S dist = magnitude(diff);
            Vec3<S> dir = diff * (S(1) / dist);
            S lambda = (dist - c.distance) / (inverse_resistance_at(A, c.a, dir) + compliance);
            A.apply_impulse(lambda * dir, c.a);
        } else if (c.kind == kContact) {
            Rigid<S>& Bb = w.bodies[(size_t)c.rigid_b];
            Vec3<S> pa = A.frame() * c.a;
            Vec3<S> pb = Bb.frame() * c.b;
            S depth = dot(pa - pb, c.normal);
            if (!(depth < S(0))) continue;
            active[k] = 1;
            S resistance = inverse_resistance_at(A, pa, c.normal) + inverse_resistance_at(Bb, pb, c.normal);
            S lambda = -depth / (resistance + compliance);
            Vec3<S> impulse = lambda * c.normal;
            A.apply_impulse(impulse, pa);
            Bb.apply_impulse(-impulse, pb);
        } else {
            Vec3<S> pa = A.frame() * c.a;
            Vec3<S> pb = c.rigid_b >= 0 ? w.bodies[(size_t)c.rigid_b].frame() * c.b : c.b;
            Vec3<S> diff = pb - pa;
            S dist = magnitude(diff);
            if (dist == S(0)) continue;  // no direction
            Vec3<S> dir = diff * (S(1) / dist);
            S resistance = inverse_resistance_at(A, pa, dir);
            if (c.rigid_b >= 0) resistance = resistance + inverse_resistance_at(w.bodies[(size_t)c.rigid_b], pb, dir);
            S lambda = (dist - c.distance) / (resistance + compliance);
            Vec3<S> impulse = lambda * dir;
            A.apply_impulse(impulse, pa);
            if (c.rigid_b >= 0) w.bodies[(size_t)c.rigid_b].apply_impulse(-impulse, pb);
        }
    }
}

// Velocity pass over the active contacts (after derive): relative velocity at the contact loses its normal
// part and `friction` x its tangential part.  Body-space inverse inertia is applied to world-space vectors
// directly, as Rigid::integrate does (src/rigid.rs:88).
template <class S>
void solve_velocities(WorldState<S>& w, const std::vector<Constraint2<S>>& cs, const std::vector<char>& active, S friction) {
    for (size_t k = 0; k < cs.size(); ++k) {
        if (!active[k]) continue;
        const Constraint2<S>& c = cs[k];
        Rigid<S>& A = w.bodies[(size_t)c.rigid_a];
        Rigid<S>& Bb = w.bodies[(size_t)c.rigid_b];
        Vec3<S> pa = A.frame() * c.a, pb = Bb.frame() * c.b;
        Vec3<S> ra = pa - (A.position + A.center_of_mass), rb = pb - (Bb.position + Bb.center_of_mass);
        Vec3<S> vrel = (A.velocity + cross(A.angular_velocity, ra)) - (Bb.velocity + cross(Bb.angular_velocity, rb));
        S vn = dot(vrel, c.normal);
        Vec3<S> vnormal = c.normal * vn;
        Vec3<S> vt = vrel - vnormal;
        Vec3<S> dv = -(vnormal + vt * friction);
        S mag = magnitude(dv);
        if (!(mag > S(0))) continue;
        Vec3<S> dir = dv * (S(1) / mag);
        S resistance = inverse_resistance_at(A, pa, dir) + inverse_resistance_at(Bb, pb, dir);
        Vec3<S> p = dir * (mag / resistance);
        A.velocity = A.velocity + p * A.inverse_mass;
        A.angular_velocity = A.angular_velocity + A.inverse_inertia * cross(ra, p);
        Bb.velocity = Bb.velocity - p * Bb.inverse_mass;
        Bb.angular_velocity = Bb.angular_velocity - Bb.inverse_inertia * cross(rb, p);
    }
}

template <class S>
void step_world(WorldState<S>& w, const std::vector<Hull<S>>& hulls, const WorldParams<S>& prm, S dt,
                size_t substep_count, StepCounters<S>* ctr = nullptr) {
    dt = dt / S(substep_count);
    const S compliance = prm.compliance / (dt * dt);
    const size_t B = w.bodies.size();
    std::vector<Vec3<S>> past_pos(B);
    std::vector<Quat<S>> past_rot(B);
    std::vector<Frame<S>> past(B);
    std::vector<Constraint2<S>> cs;
    std::vector<char> active;
    for (size_t k = 0; k < substep_count; ++k) {
        {
            PhaseScope<S> ph(ctr, kPhIntegrate);
            for (size_t b = 0; b < B; ++b) {
                past_pos[b] = w.bodies[b].position;
                past_rot[b] = w.bodies[b].rotation;
                past[b] = w.bodies[b].frame();
                w.bodies[b].integrate(dt);
            }
        }
        cs.clear();
        generate_constraints(w, hulls, past, prm, cs, ctr);
        { PhaseScope<S> ph(ctr, kPhSweep); sweep(w, cs, compliance, active); }
        { PhaseScope<S> ph(ctr, kPhDerive); for (size_t b = 0; b < B; ++b) w.bodies[b].derive(past_pos[b], past_rot[b], dt); }
        { PhaseScope<S> ph(ctr, kPhVelocity); solve_velocities(w, cs, active, prm.friction); }
        if (ctr) {
            ctr->constraints += cs.size();
            for (char a : active) ctr->active_contacts += (uint64_t)a;
        }
    }
}

}  // namespace xo
