// ORACLE — TEST INFRASTRUCTURE ONLY (see xo_math.hpp header). PARITY UNPINNED.
// Restates src/rigid.rs, src/constraint.rs, src/collision.rs, src/solver.rs, src/world.rs of
// the reference, operation for operation, keeping its quirks (stale contact points during the
// sweep, (I^-1 r) x p, body-space inverse inertia on world-space torque, b stepped with p1).
#pragma once
#include <limits>
#include <vector>

#include "xo_geometry.hpp"

namespace xo {

// ---- src/rigid.rs:7-124 ----------------------------------------------------------------
template <class S>
struct Rigid {
    S inverse_mass{};
    Mat3<S> inverse_inertia{};
    Vec3<S> external_force{}, internal_force{}, external_torque{}, internal_torque{};
    Vec3<S> velocity{}, angular_velocity{}, center_of_mass{}, position{};
    Quat<S> rotation{S(1), S(0), S(0), S(0)};

    // rigid.rs:53-71 — throws where the reference panics ("Inertia tensor is not invertible").
    static Rigid from_metrics(const RigidMetrics<S>& m) {
        Rigid r;
        r.inverse_mass = S(1) / m.mass;
        if (!invert(m.inertia_tensor, &r.inverse_inertia)) throw std::runtime_error("Inertia tensor is not invertible");
        r.center_of_mass = m.center_of_mass;
        return r;
    }
    // rigid.rs:75-80
    Frame<S> frame() const {
        Frame<S> f;
        f.position = position + center_of_mass + rotation * -center_of_mass;
        f.rotation = rotation;
        return f;
    }
    // rigid.rs:82-99
    void integrate(S dt) {
        Vec3<S> force = external_force + rotation * internal_force;
        velocity += dt * force * inverse_mass;
        position += dt * velocity;
        Vec3<S> torque = external_torque + rotation * internal_torque;
        angular_velocity += dt * inverse_inertia * torque;  // (dt * M) * torque
        Quat<S> delta_rotation =
            dt * S(0.5) * Quat<S>(S(0), angular_velocity.x, angular_velocity.y, angular_velocity.z) * rotation;
        rotation = normalize(rotation + delta_rotation);
    }
    // rigid.rs:101-109
    void derive(Vec3<S> past_position, Quat<S> past_rotation, S dt) {
        velocity = (position - past_position) / dt;
        Quat<S> delta = rotation * conjugate(past_rotation);
        if (delta.s < S(0)) delta = -delta;
        angular_velocity = S(2) * delta.v / dt;
    }
    // rigid.rs:113-123
    void apply_impulse(Vec3<S> impulse, Vec3<S> point) {
        position += impulse * inverse_mass;
        rotation = rotation +
                   S(0.5) * Quat<S>::from_sv(S(0), cross(inverse_inertia * (point - (position + center_of_mass)), impulse)) *
                       rotation;
        rotation = normalize(rotation);
    }
};

// ---- src/constraint.rs:6-37 ------------------------------------------------------------
template <class S>
struct Constraint {
    int rigid = 0;
    Vec3<S> contact0{}, contact1{};
    S distance{};

    Vec3<S> difference() const { return contact1 - contact0; }
    Vec3<S> direction() const { return normalize(difference()); }
    S current_distance() const { return magnitude(difference()); }
    // (sic) constraint.rs:25-32
    S inverse_resitance(const Rigid<S>* const* rigids) const {
        const Rigid<S>& r = *rigids[rigid];
        Vec3<S> angular_impulse = conjugate(r.rotation) * cross(contact0 - (r.position + r.center_of_mass), direction());
        return r.inverse_mass + dot(r.inverse_inertia * angular_impulse, angular_impulse);
    }
    void act(Rigid<S>* const* rigids, S factor) const {  // constraint.rs:34-37
        Vec3<S> impulse = factor * direction();
        rigids[rigid]->apply_impulse(impulse, contact0);
    }
};

// ---- src/collision.rs:13-35 ------------------------------------------------------------
template <class S>
std::vector<Constraint<S>> ground(const Rigid<S>& rigid, Frame<S> past, const Polytope<S>& polytope) {
    std::vector<Constraint<S>> constraints;
    for (const Vec3<S>& vertex : polytope.vertices) {
        Vec3<S> position = rigid.frame() * vertex;
        if (position.z >= S(0)) continue;
        Vec3<S> target_position{position.x, position.y, S(0)};
        Vec3<S> correction = target_position - position;
        Vec3<S> delta_position = rigid.frame().delta(past, position);
        Vec3<S> delta_tangential_position = delta_position - project_on(delta_position, correction);
        Constraint<S> c;
        c.rigid = 0;
        c.contact0 = position;
        c.contact1 = target_position - S(1.0) * delta_tangential_position;
        c.distance = S(0);
        constraints.push_back(c);
    }
    return constraints;
}

// ---- src/collision.rs:123-149 (dead code in the reference) -----------------------------
template <class S>
struct FaceQuery { S distance; long face; };
template <class S>
FaceQuery<S> face_axes_separation(Frame<S> f0, Frame<S> f1, const Polytope<S>& p0, const Polytope<S>& p1) {
    S max_distance = std::numeric_limits<S>::lowest();  // f64::MIN
    long face_index = -1;                               // usize::MAX
    for (size_t i = 0; i < p0.faces.size(); ++i) {
        Plane<S> plane = p0.plane(i);
        Vec3<S> neg_normal = -plane.normal;
        // max_by(total_cmp) over B's vertices mapped world -> A-local: last maximum wins.
        bool have = false;
        Vec3<S> support{};
        S support_key{};
        for (const Vec3<S>& v : p1.vertices) {
            Vec3<S> p = f0.inverse() * (f1 * v);
            S key = dot(p, neg_normal);
            if (!have || total_cmp(support_key, key) <= 0) { support = p; support_key = key; have = true; }
        }
        if (!have) throw std::runtime_error("face_axes_separation: empty polytope");
        S distance = plane.distance(support);
        if (distance > max_distance) { max_distance = distance; face_index = (long)i; }
    }
    return {max_distance, face_index};
}

// ---- src/collision.rs:151-197 (dead code in the reference) -----------------------------
template <class S>
struct EdgeQuery { S distance; long edge0, edge1; };
template <class S>
EdgeQuery<S> edge_axes_separation(Frame<S> f0, Frame<S> f1, const Polytope<S>& p0, const Polytope<S>& p1) {
    S max_distance = std::numeric_limits<S>::lowest();
    long e0 = -1, e1 = -1;
    // itertools::cartesian_product: A-major order
    for (size_t ia = 0; ia < p0.edges.size(); ++ia) {
        for (size_t jb = 0; jb < p1.edges.size(); ++jb) {
            auto i = p0.edges[ia];
            auto j = p1.edges[jb];
            Vec3<S> foot = f0 * p0.vertices[i.first];
            Vec3<S> edge_a = f0 * p0.vertices[i.second] - foot;
            Vec3<S> edge_b = f1 * p1.vertices[j.second] - f1 * p1.vertices[j.first];
            Vec3<S> axis = normalize(cross(edge_a, edge_b));  // parallel edges -> NaN axis, no guard
            if (dot(axis, foot - f0 * p0.centroid) < S(0)) axis = -axis;
            if (dot(p0.support(f0, axis), axis) > dot(foot, axis)) continue;
            Plane<S> plane = Plane<S>::from_point_normal(foot, axis);
            S distance = plane.distance(p1.support(f1, -axis));
            if (distance > max_distance) { max_distance = distance; e0 = (long)ia; e1 = (long)jb; }
        }
    }
    return {max_distance, e0, e1};
}

// ---- src/collision.rs:37-121 `sat` as written (dead, unfinished): returns true when SEPARATED
// on A's face axes, else picks reference/incident planes and returns false.
template <class S>
struct SatAsWritten { bool separated; Plane<S> ref_plane; Plane<S> incident_plane; long ref_face; };
template <class S>
SatAsWritten<S> sat_as_written(Frame<S> f0, Frame<S> f1, const Polytope<S>& p0, const Polytope<S>& p1) {
    SatAsWritten<S> out{};
    FaceQuery<S> a = face_axes_separation(f0, f1, p0, p1);
    out.ref_face = a.face;
    if (a.distance >= S(0)) { out.separated = true; return out; }
    out.ref_plane = f0 * p0.plane((size_t)a.face);
    S least_dot = std::numeric_limits<S>::max();
    for (size_t k = 0; k < p1.faces.size(); ++k) {
        Plane<S> plane = f1 * p1.plane(k);
        S d = dot(plane.normal, out.ref_plane.normal);
        if (d < least_dot) { out.incident_plane = plane; least_dot = d; }
    }
    out.separated = false;
    return out;
}

// ---- src/solver.rs:19-27 ---------------------------------------------------------------
template <class S>
size_t solve(Rigid<S>& rigid, const std::vector<Constraint<S>>& constraints, S dt) {
    S compliance = S(1e-6) / (dt * dt);
    for (const Constraint<S>& constraint : constraints) {
        const Rigid<S>* ro[1] = {&rigid};
        Rigid<S>* rw[1] = {&rigid};
        S difference = constraint.current_distance() - constraint.distance;
        S lagrange_factor = difference / (constraint.inverse_resitance(ro) + compliance);
        constraint.act(rw, lagrange_factor);
    }
    return constraints.size();
}

// ---- src/solver.rs:3-17 ----------------------------------------------------------------
// Returns the number of constraints applied (instrumentation; not in the reference).
template <class S>
size_t step(Rigid<S>& rigid, const Polytope<S>& polytope, S dt, size_t substep_count) {
    dt = dt / S(substep_count);
    size_t applied = 0;
    for (size_t k = 0; k < substep_count; ++k) {
        Vec3<S> past_position = rigid.position;
        Quat<S> past_rotation = rigid.rotation;
        Frame<S> past_frame = rigid.frame();
        rigid.integrate(dt);
        std::vector<Constraint<S>> constraints = ground(rigid, past_frame, polytope);
        applied += solve(rigid, constraints, dt);
        rigid.derive(past_position, past_rotation, dt);
    }
    return applied;
}

// ---- src/world.rs:5-43 -----------------------------------------------------------------
template <class S>
struct World {
    Rigid<S> a, b;
    // world.rs:12-31. Setup arithmetic is done in the scalar type S.
    static World create(const Polytope<S>& p1, const Polytope<S>& p2) {
        World w;
        w.a = Rigid<S>::from_metrics(rigid_metrics(p1, S(0.1)));
        w.b = Rigid<S>::from_metrics(rigid_metrics(p2, S(5.0)));
        w.a.position.z = S(4.0);
        w.a.velocity.y = S(2.5);
        w.a.angular_velocity.x = S(-4.0);
        w.a.angular_velocity.y = S(1.0);
        w.a.external_force.z = S(-2.0);
        w.b.position.x = S(4.0);
        w.b.position.z = S(4.0);
        w.b.velocity.z = S(7.0);
        w.b.angular_velocity.x = S(-5.0);
        w.b.angular_velocity.y = S(5.0);
        w.b.external_force.z = S(-2.0);
        w.b.rotation = quat_from_euler_rad(deg_to_rad(S(10.0)), deg_to_rad(S(15.0)), deg_to_rad(S(5.0)));
        return w;
    }
    // world.rs:34-43 — both bodies are stepped against p1 (reference quirk); p2 is ignored.
    size_t integrate(S dt, const Polytope<S>& p1, const Polytope<S>& /*p2*/) {
        size_t n = step(a, p1, dt, 25);
        n += step(b, p1, dt, 25);
        return n;
    }
};

}  // namespace xo
