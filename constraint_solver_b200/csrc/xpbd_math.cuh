// fp32 device math for the XPBD kernels.
//
// Every operation is an explicit round-to-nearest intrinsic (__fadd_rn/__fmul_rn/__fdiv_rn/
// __fsqrt_rn): nvcc never contracts these into FMA and never substitutes approximate div/sqrt,
// whatever flags the translation unit is built with.  That is what makes the kernels
// bit-reproducible against a plain fp32 CPU evaluation of the same expression trees, which the
// 1e-4 parity bar needs (the scene is chaotic: a 1-ulp difference grows to centimetres within
// ~80 frames).  Expression order follows cgmath 0.18.0 (the reference's math crate):
// dot = (x*x' + y*y') + z*z', normalize = v * (1/|v|), q*v = (q.v x (q.v x v + v*s))*2 + v, ...
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdio>

namespace xpbd {

#define XD __device__ __forceinline__

// Index checks of the debug build `make variant NAME=bounds EXTRA=-DXPBD_BOUNDS` (compute-sanitizer is not available on the
// GPU pool): a failed check prints its location and traps, which surfaces as a CUDA error at the next synchronisation.
// scripts/gpu_bounds.sh runs the parity tests on that library.  Compiled out of the product build.
#ifdef XPBD_BOUNDS
#define XPBD_CHECK(cond) do { if (!(cond)) { printf("XPBD_CHECK failed: %s (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, (int)blockIdx.x, (int)threadIdx.x); __trap(); } } while (0)
#else
#define XPBD_CHECK(cond) do { } while (0)
#endif

XD float fadd(float a, float b) { return __fadd_rn(a, b); }
XD float fsub(float a, float b) { return __fsub_rn(a, b); }
XD float fmul(float a, float b) { return __fmul_rn(a, b); }
XD float fdiv(float a, float b) { return __fdiv_rn(a, b); }
XD float fsqrt(float a) { return __fsqrt_rn(a); }
// 1/x, correctly rounded: the same bits as fdiv(1.0f, x) (both are the RN rounding of the same real number), fewer instructions
XD float frcp(float a) { return __frcp_rn(a); }

struct V3 {
    float x, y, z;
};
struct Q4 {  // cgmath Quaternion: vector part v = (x, y, z), scalar part s
    float x, y, z, s;
};
struct M3 {  // column-major: c0, c1, c2 are columns
    V3 c0, c1, c2;
};
struct Fr {  // Frame (src/frame.rs:7-10)
    V3 p;
    Q4 q;
};

XD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
XD V3 operator+(V3 a, V3 b) { return v3(fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
XD V3 operator-(V3 a, V3 b) { return v3(fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)); }
XD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
XD V3 operator*(V3 a, float s) { return v3(fmul(a.x, s), fmul(a.y, s), fmul(a.z, s)); }
XD V3 operator*(float s, V3 a) { return v3(fmul(s, a.x), fmul(s, a.y), fmul(s, a.z)); }
XD V3 operator/(V3 a, float s) { return v3(fdiv(a.x, s), fdiv(a.y, s), fdiv(a.z, s)); }
XD float dot(V3 a, V3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }
XD V3 cross(V3 a, V3 b) {
    return v3(fsub(fmul(a.y, b.z), fmul(a.z, b.y)), fsub(fmul(a.z, b.x), fmul(a.x, b.z)),
              fsub(fmul(a.x, b.y), fmul(a.y, b.x)));
}
XD float magnitude2(V3 a) { return dot(a, a); }
XD float magnitude(V3 a) { return fsqrt(dot(a, a)); }
XD V3 normalize(V3 a) { return a * frcp(magnitude(a)); }
XD V3 project_on(V3 a, V3 other) { return other * fdiv(dot(a, other), magnitude2(other)); }

XD Q4 q4(float s, float x, float y, float z) { Q4 q; q.x = x; q.y = y; q.z = z; q.s = s; return q; }
XD V3 qv(Q4 q) { return v3(q.x, q.y, q.z); }
XD Q4 operator+(Q4 a, Q4 b) { return q4(fadd(a.s, b.s), fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
XD Q4 operator-(Q4 a) { return q4(-a.s, -a.x, -a.y, -a.z); }
XD Q4 operator*(float k, Q4 q) { return q4(fmul(k, q.s), fmul(k, q.x), fmul(k, q.y), fmul(k, q.z)); }
XD Q4 operator*(Q4 q, float k) { return q4(fmul(q.s, k), fmul(q.x, k), fmul(q.y, k), fmul(q.z, k)); }
XD Q4 conjugate(Q4 q) { return q4(q.s, -q.x, -q.y, -q.z); }
// Hamilton product, every component evaluated left to right.
XD Q4 operator*(Q4 l, Q4 r) {
    return q4(fsub(fsub(fsub(fmul(l.s, r.s), fmul(l.x, r.x)), fmul(l.y, r.y)), fmul(l.z, r.z)),
              fsub(fadd(fadd(fmul(l.s, r.x), fmul(l.x, r.s)), fmul(l.y, r.z)), fmul(l.z, r.y)),
              fsub(fadd(fadd(fmul(l.s, r.y), fmul(l.y, r.s)), fmul(l.z, r.x)), fmul(l.x, r.z)),
              fsub(fadd(fadd(fmul(l.s, r.z), fmul(l.z, r.s)), fmul(l.x, r.y)), fmul(l.y, r.x)));
}
// Quaternion * Vector3
XD V3 operator*(Q4 q, V3 r) {
    V3 v = qv(q);
    V3 tmp = cross(v, r) + (r * q.s);
    return (cross(v, tmp) * 2.0f) + r;
}
XD float qdot(Q4 a, Q4 b) { return fadd(fmul(a.s, b.s), dot(qv(a), qv(b))); }
XD Q4 qnormalize(Q4 q) { return q * frcp(fsqrt(qdot(q, q))); }

XD M3 operator*(float k, M3 m) { M3 o; o.c0 = k * m.c0; o.c1 = k * m.c1; o.c2 = k * m.c2; return o; }
// Matrix3 * Vector3: rows dotted with v.
XD V3 operator*(const M3& m, V3 v) {
    return v3(dot(v3(m.c0.x, m.c1.x, m.c2.x), v), dot(v3(m.c0.y, m.c1.y, m.c2.y), v),
              dot(v3(m.c0.z, m.c1.z, m.c2.z), v));
}

// Frame ops (src/frame.rs:30-53)
XD V3 operator*(const Fr& f, V3 v) { return f.q * v + f.p; }
XD Fr frame_inverse(const Fr& f) {
    Fr o;
    o.q = conjugate(f.q);
    o.p = o.q * (-f.p);
    return o;
}

// f32::total_cmp key: signed-integer order of this key == IEEE totalOrder of the floats.
XD int32_t total_key(float f) {
    int32_t b = __float_as_int(f);
    return b ^ (int32_t)((uint32_t)(b >> 31) >> 1);
}

}  // namespace xpbd
