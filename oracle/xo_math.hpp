// ORACLE — TEST INFRASTRUCTURE ONLY. Nothing under oracle/ is shipped or measured as the
// product; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build or call it.
//
// PARITY UNPINNED: the reference (jim-ec/constraint_solver, Rust) has no tests or golden
// vectors, cannot be compiled here (no cargo/rustc), and its arithmetic lives in the
// un-vendored crate cgmath 0.18.0 (Cargo.lock:372-373).  This header restates cgmath 0.18.0's
// published, non-SIMD formulas from memory; the evaluation order of every expression below
// is part of the oracle's own specification.
//
// Scalar-generic: S = double follows the reference (which is f64 end to end);
// S = float is the parity oracle for the fp32 CUDA path.
// Build with -O2 -ffp-contract=off, never -ffast-math.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>

#include "xo_count.hpp"

namespace xo {

// ---- cgmath::Vector3 -------------------------------------------------------------------
template <class S>
struct Vec3 {
    S x{}, y{}, z{};
    Vec3() = default;
    Vec3(S x_, S y_, S z_) : x(x_), y(y_), z(z_) {}
    S& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
    const S& operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

template <class S> inline Vec3<S> operator+(Vec3<S> a, Vec3<S> b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class S> inline Vec3<S> operator-(Vec3<S> a, Vec3<S> b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class S> inline Vec3<S> operator-(Vec3<S> a) { return {-a.x, -a.y, -a.z}; }
// Vector3 * scalar and scalar * Vector3 (cgmath: component-wise, scalar on the written side)
template <class S> inline Vec3<S> operator*(Vec3<S> a, S s) { return {a.x * s, a.y * s, a.z * s}; }
template <class S> inline Vec3<S> operator*(S s, Vec3<S> a) { return {s * a.x, s * a.y, s * a.z}; }
// Vector3 / scalar: one division per component (NOT multiply by reciprocal)
template <class S> inline Vec3<S> operator/(Vec3<S> a, S s) { return {a.x / s, a.y / s, a.z / s}; }
template <class S> inline Vec3<S>& operator+=(Vec3<S>& a, Vec3<S> b) { a = a + b; return a; }
template <class S> inline Vec3<S>& operator/=(Vec3<S>& a, S s) { a = a / s; return a; }

// InnerSpace::dot = mul_element_wise(..).sum(), sum is a left fold: (x*x' + y*y') + z*z'
template <class S> inline S dot(Vec3<S> a, Vec3<S> b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
template <class S> inline Vec3<S> cross(Vec3<S> a, Vec3<S> b) {
    return {(a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)};
}
template <class S> inline S magnitude2(Vec3<S> a) { return dot(a, a); }
template <class S> inline S magnitude(Vec3<S> a) { return sqrt_s(magnitude2(a)); }
// normalize = normalize_to(1) = self * (1 / magnitude)
template <class S> inline Vec3<S> normalize(Vec3<S> a) { return a * (S(1) / magnitude(a)); }
// project_on(other) = other * (self.dot(other) / other.magnitude2())
template <class S> inline Vec3<S> project_on(Vec3<S> a, Vec3<S> other) {
    return other * (dot(a, other) / magnitude2(other));
}

// ---- cgmath::Quaternion (field order v then s) -----------------------------------------
template <class S>
struct Quat {
    Vec3<S> v{};
    S s{};
    Quat() = default;
    // Quaternion::new(w, xi, yj, zk)
    Quat(S w, S xi, S yj, S zk) : v(xi, yj, zk), s(w) {}
    static Quat from_sv(S s, Vec3<S> v) { Quat q; q.s = s; q.v = v; return q; }
};

template <class S> inline Quat<S> operator+(Quat<S> a, Quat<S> b) { return Quat<S>::from_sv(a.s + b.s, a.v + b.v); }
template <class S> inline Quat<S> operator-(Quat<S> a) { return Quat<S>::from_sv(-a.s, -a.v); }
template <class S> inline Quat<S> operator*(S k, Quat<S> q) { return Quat<S>::from_sv(k * q.s, k * q.v); }
template <class S> inline Quat<S> operator*(Quat<S> q, S k) { return Quat<S>::from_sv(q.s * k, q.v * k); }
template <class S> inline Quat<S> conjugate(Quat<S> q) { return Quat<S>::from_sv(q.s, -q.v); }
// Hamilton product, each component evaluated left to right
template <class S> inline Quat<S> operator*(Quat<S> l, Quat<S> r) {
    return Quat<S>(
        l.s * r.s - l.v.x * r.v.x - l.v.y * r.v.y - l.v.z * r.v.z,
        l.s * r.v.x + l.v.x * r.s + l.v.y * r.v.z - l.v.z * r.v.y,
        l.s * r.v.y + l.v.y * r.s + l.v.z * r.v.x - l.v.x * r.v.z,
        l.s * r.v.z + l.v.z * r.s + l.v.x * r.v.y - l.v.y * r.v.x);
}
// Quaternion * Vector3: tmp = q.v x v + v*q.s ; (q.v x tmp) * 2 + v
template <class S> inline Vec3<S> operator*(Quat<S> q, Vec3<S> r) {
    Vec3<S> tmp = cross(q.v, r) + (r * q.s);
    return (cross(q.v, tmp) * S(2)) + r;
}
// InnerSpace for Quaternion: dot = s*s' + v.dot(v')
template <class S> inline S dot(Quat<S> a, Quat<S> b) { return a.s * b.s + dot(a.v, b.v); }
template <class S> inline S magnitude(Quat<S> q) { return sqrt_s(dot(q, q)); }
template <class S> inline Quat<S> normalize(Quat<S> q) { return q * (S(1) / magnitude(q)); }

// Quaternion::from(Euler{x,y,z}) with angles in radians (cgmath 0.18 quaternion.rs)
template <class S> inline Quat<S> quat_from_euler_rad(S ex, S ey, S ez) {
    S half = S(0.5);
    S sx = std::sin(ex * half), cx = std::cos(ex * half);
    S sy = std::sin(ey * half), cy = std::cos(ey * half);
    S sz = std::sin(ez * half), cz = std::cos(ez * half);
    return Quat<S>(-sx * sy * sz + cx * cy * cz,
                   sx * cy * cz + sy * sz * cx,
                   -sx * sz * cy + sy * cx * cz,
                   sx * sy * cz + sz * cx * cy);
}
// Rad::from(Deg): deg * (PI / 180)
template <class S> inline S deg_to_rad(S d) { return d * S(3.14159265358979323846264338327950288 / 180.0); }

// ---- cgmath::Matrix3 (column-major: columns x, y, z) -----------------------------------
template <class S>
struct Mat3 {
    Vec3<S> x{}, y{}, z{};  // columns
    Vec3<S>& operator[](int c) { return c == 0 ? x : (c == 1 ? y : z); }
    const Vec3<S>& operator[](int c) const { return c == 0 ? x : (c == 1 ? y : z); }
    Vec3<S> row(int r) const { return {x[r], y[r], z[r]}; }
};
template <class S> inline Mat3<S> operator*(S k, Mat3<S> m) { Mat3<S> o; o.x = k * m.x; o.y = k * m.y; o.z = k * m.z; return o; }
// Matrix3 * Vector3 = (row(0).dot(v), row(1).dot(v), row(2).dot(v))
template <class S> inline Vec3<S> operator*(const Mat3<S>& m, Vec3<S> v) {
    return {dot(m.row(0), v), dot(m.row(1), v), dot(m.row(2), v)};
}
template <class S> inline S determinant(const Mat3<S>& m) {
    return m[0][0] * (m[1][1] * m[2][2] - m[2][1] * m[1][2]) -
           m[1][0] * (m[0][1] * m[2][2] - m[2][1] * m[0][2]) +
           m[2][0] * (m[0][1] * m[1][2] - m[1][1] * m[0][2]);
}
template <class S> inline Mat3<S> transpose(const Mat3<S>& m) {
    Mat3<S> t; t.x = m.row(0); t.y = m.row(1); t.z = m.row(2); return t;
}
// SquareMatrix::invert: cross-product / determinant form, then transpose. false when det == 0.
template <class S> inline bool invert(const Mat3<S>& m, Mat3<S>* out) {
    S det = determinant(m);
    if (det == S(0)) return false;
    Mat3<S> c;
    c.x = cross(m[1], m[2]) / det;
    c.y = cross(m[2], m[0]) / det;
    c.z = cross(m[0], m[1]) / det;
    *out = transpose(c);
    return true;
}

// ---- f32/f64::total_cmp (IEEE 754 totalOrder): -NaN < -inf < .. < -0 < +0 < .. < +inf < +NaN
inline int total_cmp(double a, double b) {
    int64_t l, r;
    std::memcpy(&l, &a, 8); std::memcpy(&r, &b, 8);
    l ^= (int64_t)((uint64_t)(l >> 63) >> 1);
    r ^= (int64_t)((uint64_t)(r >> 63) >> 1);
    return l < r ? -1 : (l > r ? 1 : 0);
}
inline int total_cmp(float a, float b) {
    int32_t l, r;
    std::memcpy(&l, &a, 4); std::memcpy(&r, &b, 4);
    l ^= (int32_t)((uint32_t)(l >> 31) >> 1);
    r ^= (int32_t)((uint32_t)(r >> 31) >> 1);
    return l < r ? -1 : (l > r ? 1 : 0);
}

}  // namespace xo
