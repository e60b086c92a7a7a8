// ORACLE — TEST INFRASTRUCTURE ONLY (see xo_math.hpp header). PARITY UNPINNED.
// Restates src/geometry.rs, src/frame.rs and src/geometry/integrate.rs of the reference.
#pragma once
#include <stdexcept>
#include <utility>
#include <vector>

#include "xo_math.hpp"

namespace xo {

// ---- src/geometry.rs:8-69 Plane --------------------------------------------------------
template <class S>
struct Plane {
    Vec3<S> normal{};
    S displacement{};

    // geometry.rs:16-24
    static Plane from_points(Vec3<S> p0, Vec3<S> p1, Vec3<S> p2) {
        Plane pl;
        pl.normal = normalize(cross(p1 - p0, p2 - p0));
        pl.displacement = dot(pl.normal, p0);
        return pl;
    }
    // geometry.rs:27-36
    static Plane from_point_normal(Vec3<S> point, Vec3<S> normal) {
        S displacement = magnitude(project_on(point, normal));
        if (dot(point, normal) < S(0)) displacement *= S(-1);
        Plane pl; pl.normal = normal; pl.displacement = displacement;
        return pl;
    }
    S distance(Vec3<S> point) const { return dot(normal, point) - displacement; }   // :39-41
    Vec3<S> project(Vec3<S> point) const { return point - distance(point) * normal; }  // :45-47
    S constant() const { return -displacement; }                                       // :50-52
    Plane flip() const { Plane p; p.normal = -normal; p.displacement = -displacement; return p; }  // :55-60
    Vec3<S> support() const { return displacement * normal; }                          // :62-64
    bool facing(Vec3<S> p) const { return dot(normal, p - support()) >= S(0); }       // :66-68
};

// ---- src/frame.rs:7-76 Frame -----------------------------------------------------------
template <class S>
struct Frame {
    Vec3<S> position{};
    Quat<S> rotation{S(1), S(0), S(0), S(0)};

    Frame inverse() const {  // frame.rs:30-37
        Frame f;
        f.rotation = conjugate(rotation);
        f.position = f.rotation * -position;
        return f;
    }
    Vec3<S> operator*(Vec3<S> rhs) const { return rotation * rhs + position; }  // frame.rs:47-53
    // frame.rs:40-44
    Vec3<S> delta(Frame past, Vec3<S> global) const {
        Vec3<S> local = inverse() * global;
        Vec3<S> past_global = past * local;
        return global - past_global;
    }
    Plane<S> operator*(Plane<S> rhs) const {  // frame.rs:55-64
        Vec3<S> support = rhs.displacement * rhs.normal;
        support = (*this) * support;
        Vec3<S> normal = rotation * rhs.normal;
        return Plane<S>::from_point_normal(support, normal);
    }
    Frame operator*(Frame rhs) const {  // frame.rs:67-76
        Frame f;
        f.position = position + rotation * rhs.position;
        f.rotation = rotation * rhs.rotation;
        return f;
    }
};

// ---- src/geometry/integrate.rs:18-24 ---------------------------------------------------
template <class S>
struct RigidMetrics {
    S mass{}, volume{};
    Vec3<S> center_of_mass{};
    Mat3<S> inertia_tensor{};
};

// ---- src/geometry.rs:81-307 Polytope ---------------------------------------------------
template <class S>
struct Polytope {
    std::vector<Vec3<S>> vertices;
    std::vector<std::pair<int, int>> edges;
    std::vector<std::vector<int>> faces;
    Vec3<S> centroid{};

    static Polytope new_tetrahedron() {  // geometry.rs:97-109
        Polytope p;
        p.centroid = {S(0.25), S(0.25), S(0.25)};
        p.vertices = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
        p.edges = {{0, 1}, {0, 2}, {0, 3}, {1, 2}, {1, 3}, {2, 3}};
        p.faces = {{0, 3, 2}, {3, 0, 1}, {2, 1, 0}, {1, 2, 3}};
        return p;
    }
    static Polytope new_cube() {  // geometry.rs:113-149
        Polytope p;
        p.centroid = {S(0.5), S(0.5), S(0.5)};
        p.vertices = {{0, 0, 0}, {1, 0, 0}, {0, 1, 0}, {1, 1, 0}, {0, 0, 1}, {1, 0, 1}, {0, 1, 1}, {1, 1, 1}};
        p.edges = {{0, 1}, {1, 3}, {3, 2}, {2, 0}, {4, 5}, {5, 7}, {7, 6}, {6, 4}, {0, 4}, {1, 5}, {3, 7}, {2, 6}};
        p.faces = {{0, 2, 3, 1}, {4, 5, 7, 6}, {4, 0, 1, 5}, {5, 1, 3, 7}, {7, 3, 2, 6}, {6, 2, 0, 4}};
        return p;
    }
    static Polytope new_icosahedron() {  // geometry.rs:153-230
        S phi = (S(1) + std::sqrt(S(5))) / S(2);
        S mag = std::sqrt(phi * phi + S(1));
        S a = phi / mag, b = S(1) / mag;
        Polytope p;
        p.centroid = {0, 0, 0};
        p.vertices = {{a, b, 0}, {a, -b, 0}, {-a, b, 0}, {-a, -b, 0}, {0, a, b}, {0, a, -b},
                      {0, -a, b}, {0, -a, -b}, {b, 0, a}, {-b, 0, a}, {b, 0, -a}, {-b, 0, -a}};
        p.edges = {{8, 9}, {8, 0}, {8, 1}, {1, 0}, {9, 2}, {9, 3}, {2, 3}, {2, 5}, {2, 11}, {5, 11},
                   {3, 7}, {3, 11}, {7, 11}, {0, 5}, {1, 7}, {4, 5}, {4, 0}, {4, 8}, {4, 9}, {4, 2},
                   {6, 9}, {6, 8}, {6, 1}, {6, 7}, {6, 3}, {10, 1}, {10, 0}, {10, 5}, {10, 11}, {10, 7}};
        p.faces = {{0, 5, 4}, {2, 4, 5}, {1, 6, 7}, {3, 7, 6}, {1, 0, 8}, {0, 1, 10}, {2, 3, 9},
                   {3, 2, 11}, {4, 9, 8}, {6, 8, 9}, {5, 10, 11}, {7, 11, 10}, {0, 4, 8}, {0, 10, 5},
                   {2, 9, 4}, {2, 5, 11}, {1, 8, 6}, {1, 7, 10}, {3, 6, 9}, {3, 11, 7}};
        return p;
    }

    // geometry.rs:262-271 — plane of face i, pointing away from the centroid.
    Plane<S> plane(size_t i) const {
        const std::vector<int>& face = faces.at(i);
        if (face.size() < 3) throw std::out_of_range("face with fewer than 3 vertices");
        Plane<S> pl = Plane<S>::from_points(vertices[face[0]], vertices[face[1]], vertices[face[2]]);
        return !pl.facing(centroid) ? pl : pl.flip();
    }
    // geometry.rs:274-281 — world-space support vertex; LAST maximum under total_cmp
    // (Rust Iterator::max_by keeps the later element on Ordering::Equal).
    Vec3<S> support(Frame<S> frame, Vec3<S> direction) const {
        if (vertices.empty()) throw std::runtime_error("support of empty polytope");
        Vec3<S> best = frame * vertices[0];
        S best_key = dot(best, direction);
        for (size_t i = 1; i < vertices.size(); ++i) {
            Vec3<S> w = frame * vertices[i];
            S key = dot(w, direction);
            if (total_cmp(best_key, key) <= 0) { best = w; best_key = key; }
        }
        return best;
    }
    // geometry.rs:283-289
    Vec3<S> minkowski_support(Frame<S> f0, Frame<S> f1, Vec3<S> direction) const {
        return support(f0, direction) - support(f1, -direction);
    }
    // geometry.rs:296-307  `f64 * Polytope`
    Polytope scaled(S k) const {
        Polytope p = *this;
        for (auto& v : p.vertices) v = k * v;
        p.centroid = k * centroid;
        return p;
    }
    template <class T>
    Polytope<T> cast() const {
        Polytope<T> p;
        for (auto& v : vertices) p.vertices.push_back({T(v.x), T(v.y), T(v.z)});
        p.edges = edges; p.faces = faces;
        p.centroid = {T(centroid.x), T(centroid.y), T(centroid.z)};
        return p;
    }
};

// ---- src/geometry/integrate.rs (Mirtich 1996) ------------------------------------------
namespace mirtich {
template <class S> struct Face { Vec3<S> normal; S displacement; const std::vector<int>* vertices; };
template <class S> struct ProjectionIntegrals { S e{}, a{}, b{}, aa{}, ab{}, bb{}, aaa{}, aab{}, abb{}, bbb{}; };
template <class S> struct FaceIntegrals { S a{}, b{}, c{}, aa{}, bb{}, cc{}, aaa{}, bbb{}, ccc{}, aab{}, bbc{}, cca{}; };
template <class S> struct VolumeIntegrals { S t0{}; Vec3<S> t1{}, t2{}, tp{}; };
template <class S> inline S sq(S x) { return x * x; }        // powi(2)
template <class S> inline S cb(S x) { return x * x * x; }    // powi(3)

// integrate.rs:223-288
template <class S>
ProjectionIntegrals<S> projection_integrals(const std::vector<Vec3<S>>& verts, const Face<S>& f, int alpha, int beta) {
    ProjectionIntegrals<S> I;
    const std::vector<int>& fv = *f.vertices;
    size_t n = fv.size();
    for (size_t i = 0; i < n; ++i) {
        S a0 = verts[fv[i]][alpha], b0 = verts[fv[i]][beta];
        S a1 = verts[fv[(i + 1) % n]][alpha], b1 = verts[fv[(i + 1) % n]][beta];
        S da = a1 - a0, db = b1 - b0;
        S a0_2 = a0 * a0, a0_3 = a0_2 * a0, a0_4 = a0_3 * a0;
        S b0_2 = b0 * b0, b0_3 = b0_2 * b0, b0_4 = b0_3 * b0;
        S a1_2 = a1 * a1, a1_3 = a1_2 * a1;
        S b1_2 = b1 * b1, b1_3 = b1_2 * b1;
        S c_1 = a1 + a0;
        S c_a = a1 * c_1 + a0_2;
        S c_aa = a1 * c_a + a0_3;
        S c_aaa = a1 * c_aa + a0_4;
        S c_b = b1 * (b1 + b0) + b0_2;
        S c_bb = b1 * c_b + b0_3;
        S c_bbb = b1 * c_bb + b0_4;
        S c_ab = S(3) * a1_2 + S(2) * a1 * a0 + a0_2;
        S k_ab = a1_2 + S(2) * a1 * a0 + S(3) * a0_2;
        S c_aab = a0 * c_ab + S(4) * a1_3;
        S k_aab = a1 * k_ab + S(4) * a0_3;
        S c_abb = S(4) * b1_3 + S(3) * b1_2 * b0 + S(2) * b1 * b0_2 + b0_3;
        S k_abb = b1_3 + S(2) * b1_2 * b0 + S(3) * b1 * b0_2 + S(4) * b0_3;
        I.e += db * c_1;
        I.a += db * c_a;
        I.aa += db * c_aa;
        I.aaa += db * c_aaa;
        I.b += da * c_b;
        I.bb += da * c_bb;
        I.bbb += da * c_bbb;
        I.ab += db * (b1 * c_ab + b0 * k_ab);
        I.aab += db * (b1 * c_aab + b0 * k_aab);
        I.abb += da * (a1 * c_abb + a0 * k_abb);
    }
    I.e /= S(2); I.a /= S(6); I.aa /= S(12); I.aaa /= S(20);
    I.b /= S(-6); I.bb /= S(-12); I.bbb /= S(-20);
    I.ab /= S(24); I.aab /= S(60); I.abb /= S(-60);
    return I;
}

// integrate.rs:171-220
template <class S>
FaceIntegrals<S> face_integrals(const std::vector<Vec3<S>>& verts, const Face<S>& f, int alpha, int beta, int gamma) {
    ProjectionIntegrals<S> p = projection_integrals(verts, f, alpha, beta);
    S w = f.displacement;
    Vec3<S> n = f.normal;
    S k1 = S(1) / n[gamma], k2 = k1 * k1, k3 = k2 * k1, k4 = k3 * k1;
    S na = n[alpha], nb = n[beta];
    FaceIntegrals<S> F;
    F.a = k1 * p.a;
    F.b = k1 * p.b;
    F.c = -k2 * (na * p.a + nb * p.b + w * p.e);
    F.aa = k1 * p.aa;
    F.bb = k1 * p.bb;
    F.cc = k3 * (sq(na) * p.aa + S(2) * na * nb * p.ab + sq(nb) * p.bb + w * (S(2) * (na * p.a + nb * p.b) + w * p.e));
    F.aaa = k1 * p.aaa;
    F.bbb = k1 * p.bbb;
    F.ccc = -k4 * (cb(na) * p.aaa + S(3) * sq(na) * nb * p.aab + S(3) * na * sq(nb) * p.abb + cb(nb) * p.bbb +
                   S(3) * w * (sq(na) * p.aa + S(2) * na * nb * p.ab + sq(nb) * p.bb) +
                   w * w * (S(3) * (na * p.a + nb * p.b) + w * p.e));
    F.aab = k1 * p.aab;
    F.bbc = -k2 * (na * p.abb + nb * p.bbb + w * p.bb);
    F.cca = k3 * (sq(na) * p.aaa + S(2) * na * nb * p.aab + sq(nb) * p.abb + w * (S(2) * (na * p.aa + nb * p.ab) + w * p.a));
    return F;
}

// integrate.rs:126-169
template <class S>
VolumeIntegrals<S> volume_integrals(const std::vector<Vec3<S>>& verts, const std::vector<Face<S>>& faces) {
    VolumeIntegrals<S> t;
    for (const Face<S>& face : faces) {
        Vec3<S> n{std::fabs(face.normal.x), std::fabs(face.normal.y), std::fabs(face.normal.z)};
        int gamma = (n.x > n.y && n.x > n.z) ? 0 : (n.y > n.z ? 1 : 2);
        int alpha = (gamma + 1) % 3, beta = (alpha + 1) % 3;
        FaceIntegrals<S> f = face_integrals(verts, face, alpha, beta, gamma);
        t.t0 += face.normal.x * (alpha == 0 ? f.a : (beta == 0 ? f.b : f.c));
        t.t1[alpha] += face.normal[alpha] * f.aa;
        t.t1[beta] += face.normal[beta] * f.bb;
        t.t1[gamma] += face.normal[gamma] * f.cc;
        t.t2[alpha] += face.normal[alpha] * f.aaa;
        t.t2[beta] += face.normal[beta] * f.bbb;
        t.t2[gamma] += face.normal[gamma] * f.ccc;
        t.tp[alpha] += face.normal[alpha] * f.aab;
        t.tp[beta] += face.normal[beta] * f.bbc;
        t.tp[gamma] += face.normal[gamma] * f.cca;
    }
    t.t1 /= S(2); t.t2 /= S(3); t.tp /= S(2);
    return t;
}
}  // namespace mirtich

// integrate.rs:26-75
template <class S>
RigidMetrics<S> rigid_metrics(const Polytope<S>& polytope, S density) {
    std::vector<mirtich::Face<S>> faces;
    for (size_t i = 0; i < polytope.faces.size(); ++i) {
        Plane<S> pl = polytope.plane(i);
        faces.push_back({pl.normal, pl.constant(), &polytope.faces[i]});
    }
    mirtich::VolumeIntegrals<S> I = mirtich::volume_integrals(polytope.vertices, faces);
    S m = density * I.t0;
    Vec3<S> r = I.t1 / I.t0;
    Mat3<S> j;
    j.x.x = density * (I.t2.y + I.t2.z);
    j.y.y = density * (I.t2.z + I.t2.x);
    j.z.z = density * (I.t2.x + I.t2.y);
    j.x.y = -density * I.tp.x;
    j.y.z = -density * I.tp.y;
    j.z.x = -density * I.tp.z;
    j.y.x = j.x.y; j.z.y = j.y.z; j.x.z = j.z.x;
    j.x.x -= m * (r.y * r.y + r.z * r.z);
    j.y.y -= m * (r.z * r.z + r.x * r.x);
    j.z.z -= m * (r.x * r.x + r.y * r.y);
    j.x.y += m * r.x * r.y;
    j.y.z += m * r.y * r.z;
    j.z.x += m * r.z * r.x;
    j.y.x = j.x.y; j.z.y = j.y.z; j.x.z = j.z.x;
    RigidMetrics<S> out;
    out.mass = m; out.volume = I.t0; out.center_of_mass = r; out.inertia_tensor = j;
    return out;
}

}  // namespace xo
