// ORACLE — TEST INFRASTRUCTURE ONLY (see xo_math.hpp header).
//
// Counted<T>: a scalar that computes exactly like T and counts every add/sub, mul, div and sqrt it performs
// (add = mul = div = sqrt = 1 flop).  Instantiating the scalar-generic restatement with
// S = Counted<float> runs the very same arithmetic and control flow as the S = float parity oracle and yields the
// ALGORITHMIC flop count of the specification as written — the reference's face / edge queries brute force
// (src/collision.rs:123-197) included — per phase of the substep.  bench.py divides by the measured time to get the
// FP32 fraction of the step; what the CUDA kernels EXECUTE (fewer: common subexpressions hoisted, edge pairs culled)
// is reported next to it from ncu.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>
#include <type_traits>

namespace xo {

struct FlopCounters {
    uint64_t add = 0, mul = 0, div = 0, sqrt = 0;
    uint64_t total() const { return add + mul + div + sqrt; }
};
inline thread_local FlopCounters t_flops;

template <class T>
struct Counted {
    T v{};
    Counted() = default;
    template <class U, class = std::enable_if_t<std::is_arithmetic<U>::value>>
    Counted(U x) : v(static_cast<T>(x)) {}
};
template <class T> inline Counted<T> operator+(Counted<T> a, Counted<T> b) { ++t_flops.add; return Counted<T>(a.v + b.v); }
template <class T> inline Counted<T> operator-(Counted<T> a, Counted<T> b) { ++t_flops.add; return Counted<T>(a.v - b.v); }
template <class T> inline Counted<T> operator*(Counted<T> a, Counted<T> b) { ++t_flops.mul; return Counted<T>(a.v * b.v); }
template <class T> inline Counted<T> operator/(Counted<T> a, Counted<T> b) { ++t_flops.div; return Counted<T>(a.v / b.v); }
template <class T> inline Counted<T> operator-(Counted<T> a) { return Counted<T>(-a.v); }  // sign flip: not a flop
template <class T> inline Counted<T>& operator*=(Counted<T>& a, Counted<T> b) { a = a * b; return a; }
template <class T> inline Counted<T>& operator+=(Counted<T>& a, Counted<T> b) { a = a + b; return a; }
template <class T> inline Counted<T>& operator-=(Counted<T>& a, Counted<T> b) { a = a - b; return a; }
template <class T> inline Counted<T>& operator/=(Counted<T>& a, Counted<T> b) { a = a / b; return a; }
template <class T> inline bool operator<(Counted<T> a, Counted<T> b) { return a.v < b.v; }
template <class T> inline bool operator>(Counted<T> a, Counted<T> b) { return a.v > b.v; }
template <class T> inline bool operator<=(Counted<T> a, Counted<T> b) { return a.v <= b.v; }
template <class T> inline bool operator>=(Counted<T> a, Counted<T> b) { return a.v >= b.v; }
template <class T> inline bool operator==(Counted<T> a, Counted<T> b) { return a.v == b.v; }
template <class T> inline bool operator!=(Counted<T> a, Counted<T> b) { return a.v != b.v; }

// scalar functions the restatement calls unqualified (xo_math.hpp): plain scalars forward to <cmath>
inline float sqrt_s(float x) { return std::sqrt(x); }
inline double sqrt_s(double x) { return std::sqrt(x); }
template <class T> inline Counted<T> sqrt_s(Counted<T> x) { ++t_flops.sqrt; return Counted<T>(std::sqrt(x.v)); }
inline int total_cmp(float a, float b);
inline int total_cmp(double a, double b);
template <class T> inline int total_cmp(Counted<T> a, Counted<T> b) { return total_cmp(a.v, b.v); }

// flop counters "now" for a scalar type: zero for plain scalars (the phase bookkeeping compiles away)
template <class S> struct FlopProbe { static FlopCounters now() { return FlopCounters(); } };
template <class T> struct FlopProbe<Counted<T>> { static FlopCounters now() { return t_flops; } };

}  // namespace xo

namespace std {
template <class T>
struct numeric_limits<xo::Counted<T>> {
    static constexpr bool is_specialized = true;
    static xo::Counted<T> max() { return xo::Counted<T>(numeric_limits<T>::max()); }
    static xo::Counted<T> lowest() { return xo::Counted<T>(numeric_limits<T>::lowest()); }
    static xo::Counted<T> min() { return xo::Counted<T>(numeric_limits<T>::min()); }
};
}  // namespace std
