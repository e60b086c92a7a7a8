// Host-side f64 setup math of the product library: face planes (src/geometry.rs:14-69,256-271),
// Mirtich mass properties (src/geometry/integrate.rs) and Matrix3 inversion for Rigid::new
// (src/rigid.rs:53-71).  Runs once per hull / body at setup time; never on the step path.
// Expression order follows the reference (and cgmath 0.18.0) so that the f64 results agree with it.
#pragma once
#include <array>
#include <cmath>
#include <vector>

#include "../../include/xpbd_b200.h"

namespace xpbd_host {

using d3 = std::array<double, 3>;

inline d3 sub(const d3& a, const d3& b) { return {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; }
inline double dot(const d3& a, const d3& b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
inline d3 cross(const d3& a, const d3& b) {
    return {(a[1] * b[2]) - (a[2] * b[1]), (a[2] * b[0]) - (a[0] * b[2]), (a[0] * b[1]) - (a[1] * b[0])};
}
inline d3 scale(const d3& a, double s) { return {a[0] * s, a[1] * s, a[2] * s}; }
inline d3 unit(const d3& a) { return scale(a, 1.0 / std::sqrt(dot(a, a))); }

struct PlaneD {
    d3 n;
    double d;
};

inline d3 hull_vertex(const xpbd_polytope& h, int i) { return {h.vertices[3 * i], h.vertices[3 * i + 1], h.vertices[3 * i + 2]}; }

// Polytope::plane (src/geometry.rs:262-271): from the first three face vertices, flipped to point
// away from the centroid.
inline PlaneD face_plane(const xpbd_polytope& h, int f) {
    const int32_t* idx = h.face_indices + h.face_offsets[f];
    d3 p0 = hull_vertex(h, idx[0]), p1 = hull_vertex(h, idx[1]), p2 = hull_vertex(h, idx[2]);
    PlaneD pl;
    pl.n = unit(cross(sub(p1, p0), sub(p2, p0)));
    pl.d = dot(pl.n, p0);
    d3 c = {h.centroid[0], h.centroid[1], h.centroid[2]};
    d3 support = {pl.d * pl.n[0], pl.d * pl.n[1], pl.d * pl.n[2]};
    bool facing = dot(pl.n, sub(c, support)) >= 0.0;
    if (facing) {
        pl.n = {-pl.n[0], -pl.n[1], -pl.n[2]};
        pl.d = -pl.d;
    }
    return pl;
}

// Structural validation the reference leaves to panics (src/geometry.rs:264,280).
inline const char* validate_hull(const xpbd_polytope* h) {
    if (!h || !h->vertices || !h->face_offsets || !h->face_indices) return "null hull arrays";
    if (h->n_vertices < 4) return "hull needs at least 4 vertices";
    if (h->n_faces < 4) return "hull needs at least 4 faces";
    if (h->n_edges < 0 || (h->n_edges > 0 && !h->edges)) return "bad edge array";
    if (h->face_offsets[0] != 0) return "face_offsets[0] must be 0";
    for (int f = 0; f < h->n_faces; ++f) {
        int b = h->face_offsets[f], e = h->face_offsets[f + 1];
        if (e - b < 3) return "face with fewer than 3 vertices";
        for (int k = b; k < e; ++k)
            if (h->face_indices[k] < 0 || h->face_indices[k] >= h->n_vertices) return "face index out of range";
    }
    for (int e = 0; e < 2 * h->n_edges; ++e)
        if (h->edges[e] < 0 || h->edges[e] >= h->n_vertices) return "edge index out of range";
    for (int i = 0; i < 3 * h->n_vertices; ++i)
        if (!std::isfinite(h->vertices[i])) return "non-finite vertex";
    return nullptr;
}

// ---- Mirtich 1996 (src/geometry/integrate.rs:126-288) ------------------------------------
struct Proj { double e = 0, a = 0, b = 0, aa = 0, ab = 0, bb = 0, aaa = 0, aab = 0, abb = 0, bbb = 0; };

inline Proj projection_integrals(const xpbd_polytope& h, int f, int A, int B) {
    Proj P;
    const int32_t* idx = h.face_indices + h.face_offsets[f];
    const int n = h.face_offsets[f + 1] - h.face_offsets[f];
    for (int i = 0; i < n; ++i) {
        const double a0 = h.vertices[3 * idx[i] + A], b0 = h.vertices[3 * idx[i] + B];
        const double a1 = h.vertices[3 * idx[(i + 1) % n] + A], b1 = h.vertices[3 * idx[(i + 1) % n] + B];
        const double da = a1 - a0, db = b1 - b0;
        const double a0_2 = a0 * a0, a0_3 = a0_2 * a0, a0_4 = a0_3 * a0;
        const double b0_2 = b0 * b0, b0_3 = b0_2 * b0, b0_4 = b0_3 * b0;
        const double a1_2 = a1 * a1, a1_3 = a1_2 * a1, b1_2 = b1 * b1, b1_3 = b1_2 * b1;
        const double c1 = a1 + a0;
        const double ca = a1 * c1 + a0_2, caa = a1 * ca + a0_3, caaa = a1 * caa + a0_4;
        const double cb = b1 * (b1 + b0) + b0_2, cbb = b1 * cb + b0_3, cbbb = b1 * cbb + b0_4;
        const double cab = 3.0 * a1_2 + 2.0 * a1 * a0 + a0_2, kab = a1_2 + 2.0 * a1 * a0 + 3.0 * a0_2;
        const double caab = a0 * cab + 4.0 * a1_3, kaab = a1 * kab + 4.0 * a0_3;
        const double cabb = 4.0 * b1_3 + 3.0 * b1_2 * b0 + 2.0 * b1 * b0_2 + b0_3;
        const double kabb = b1_3 + 2.0 * b1_2 * b0 + 3.0 * b1 * b0_2 + 4.0 * b0_3;
        P.e += db * c1; P.a += db * ca; P.aa += db * caa; P.aaa += db * caaa;
        P.b += da * cb; P.bb += da * cbb; P.bbb += da * cbbb;
        P.ab += db * (b1 * cab + b0 * kab);
        P.aab += db * (b1 * caab + b0 * kaab);
        P.abb += da * (a1 * cabb + a0 * kabb);
    }
    P.e /= 2.0; P.a /= 6.0; P.aa /= 12.0; P.aaa /= 20.0;
    P.b /= -6.0; P.bb /= -12.0; P.bbb /= -20.0;
    P.ab /= 24.0; P.aab /= 60.0; P.abb /= -60.0;
    return P;
}

inline void rigid_metrics(const xpbd_polytope& h, double density, xpbd_rigid_metrics* out) {
    double t0 = 0;
    d3 t1{0, 0, 0}, t2{0, 0, 0}, tp{0, 0, 0};
    for (int f = 0; f < h.n_faces; ++f) {
        const PlaneD pl = face_plane(h, f);
        const d3 n = pl.n;
        const double w = -pl.d;  // Plane::constant
        const double ax = std::fabs(n[0]), ay = std::fabs(n[1]), az = std::fabs(n[2]);
        const int C = (ax > ay && ax > az) ? 0 : (ay > az ? 1 : 2);
        const int A = (C + 1) % 3, B = (A + 1) % 3;
        const Proj p = projection_integrals(h, f, A, B);
        const double k1 = 1.0 / n[C], k2 = k1 * k1, k3 = k2 * k1, k4 = k3 * k1;
        const double na = n[A], nb = n[B];
        const double na2 = na * na, nb2 = nb * nb, na3 = na * na * na, nb3 = nb * nb * nb;
        const double Fa = k1 * p.a, Fb = k1 * p.b;
        const double Fc = -k2 * (na * p.a + nb * p.b + w * p.e);
        const double Faa = k1 * p.aa, Fbb = k1 * p.bb;
        const double Fcc = k3 * (na2 * p.aa + 2.0 * na * nb * p.ab + nb2 * p.bb + w * (2.0 * (na * p.a + nb * p.b) + w * p.e));
        const double Faaa = k1 * p.aaa, Fbbb = k1 * p.bbb;
        const double Fccc = -k4 * (na3 * p.aaa + 3.0 * na2 * nb * p.aab + 3.0 * na * nb2 * p.abb + nb3 * p.bbb +
                                   3.0 * w * (na2 * p.aa + 2.0 * na * nb * p.ab + nb2 * p.bb) +
                                   w * w * (3.0 * (na * p.a + nb * p.b) + w * p.e));
        const double Faab = k1 * p.aab;
        const double Fbbc = -k2 * (na * p.abb + nb * p.bbb + w * p.bb);
        const double Fcca = k3 * (na2 * p.aaa + 2.0 * na * nb * p.aab + nb2 * p.abb + w * (2.0 * (na * p.aa + nb * p.ab) + w * p.a));
        t0 += n[0] * (A == 0 ? Fa : (B == 0 ? Fb : Fc));
        t1[A] += n[A] * Faa; t1[B] += n[B] * Fbb; t1[C] += n[C] * Fcc;
        t2[A] += n[A] * Faaa; t2[B] += n[B] * Fbbb; t2[C] += n[C] * Fccc;
        tp[A] += n[A] * Faab; tp[B] += n[B] * Fbbc; tp[C] += n[C] * Fcca;
    }
    for (int i = 0; i < 3; ++i) { t1[i] /= 2.0; t2[i] /= 3.0; tp[i] /= 2.0; }
    const double m = density * t0;
    const d3 r = {t1[0] / t0, t1[1] / t0, t1[2] / t0};
    // j[col][row], cgmath naming j.<col>.<row>
    double j[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    j[0][0] = density * (t2[1] + t2[2]);
    j[1][1] = density * (t2[2] + t2[0]);
    j[2][2] = density * (t2[0] + t2[1]);
    j[0][1] = -density * tp[0];
    j[1][2] = -density * tp[1];
    j[2][0] = -density * tp[2];
    j[0][0] -= m * (r[1] * r[1] + r[2] * r[2]);
    j[1][1] -= m * (r[2] * r[2] + r[0] * r[0]);
    j[2][2] -= m * (r[0] * r[0] + r[1] * r[1]);
    j[0][1] += m * r[0] * r[1];
    j[1][2] += m * r[1] * r[2];
    j[2][0] += m * r[2] * r[0];
    j[1][0] = j[0][1]; j[2][1] = j[1][2]; j[0][2] = j[2][0];
    out->mass = m;
    out->volume = t0;
    for (int i = 0; i < 3; ++i) out->center_of_mass[i] = r[i];
    for (int c = 0; c < 3; ++c) for (int k = 0; k < 3; ++k) out->inertia_tensor[3 * c + k] = j[c][k];
}

// Matrix3::invert (cgmath 0.18): columns = cross products / det, then transpose. m column-major.
inline bool invert3(const double m[9], double out[9]) {
    const d3 c0{m[0], m[1], m[2]}, c1{m[3], m[4], m[5]}, c2{m[6], m[7], m[8]};
    const double det = c0[0] * (c1[1] * c2[2] - c2[1] * c1[2]) - c1[0] * (c0[1] * c2[2] - c2[1] * c0[2]) +
                       c2[0] * (c0[1] * c1[2] - c1[1] * c0[2]);
    if (det == 0.0 || !std::isfinite(det)) return false;
    const d3 r0 = cross(c1, c2), r1 = cross(c2, c0), r2 = cross(c0, c1);
    // rows of the inverse are r0/det, r1/det, r2/det; store column-major
    for (int k = 0; k < 3; ++k) {
        out[3 * k + 0] = r0[k] / det;
        out[3 * k + 1] = r1[k] / det;
        out[3 * k + 2] = r2[k] / det;
    }
    return true;
}

}  // namespace xpbd_host
