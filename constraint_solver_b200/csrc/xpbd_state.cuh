// Device-resident body state: fp32 structure-of-arrays, indexed by global body index
// g = world * bodies_per_world + body.
//
//   dynamic (13 floats, read + written by every step call):
//     dynA[g] = (position.x, position.y, position.z, velocity.x)
//     dynB[g] = (velocity.y, velocity.z, angular_velocity.x, angular_velocity.y)
//     dynC[g] = (rotation.v.x, rotation.v.y, rotation.v.z, rotation.s)
//     dynD[g] =  angular_velocity.z
//   constant (25 floats, read by every step call):
//     k0[g] = (inverse_mass, center_of_mass.xyz)
//     k1[g] = (I^-1 column-major elements 0..3)   k2[g] = (elements 4..7)
//     k3[g] = (element 8, external_force.xyz)
//     k4[g] = (internal_force.xyz, external_torque.x)
//     k5[g] = (external_torque.y, external_torque.z, internal_torque.x, internal_torque.y)
//     k6[g] =  internal_torque.z
// A warp reading consecutive g issues fully coalesced 16-byte loads per array.
// Algorithmic traffic per body per step call: 38 floats read + 13 written = 204 B.
#pragma once
#include "xpbd_math.cuh"

namespace xpbd {

struct BodyArrays {
    float4* dynA; float4* dynB; float4* dynC; float* dynD;
    float4* k0; float4* k1; float4* k2; float4* k3; float4* k4; float4* k5; float* k6;
};

struct BodyDyn {   // Rigid's mutable part
    V3 position; Q4 rotation; V3 velocity; V3 angular_velocity;
};
struct BodyConst { // Rigid's constant part
    float inverse_mass; M3 inverse_inertia; V3 center_of_mass;
    V3 external_force, internal_force, external_torque, internal_torque;
};

XD void load_dyn(const BodyArrays& s, int64_t g, BodyDyn& d) {
    float4 a = s.dynA[g], b = s.dynB[g], c = s.dynC[g];
    float w = s.dynD[g];
    d.position = v3(a.x, a.y, a.z);
    d.velocity = v3(a.w, b.x, b.y);
    d.angular_velocity = v3(b.z, b.w, w);
    d.rotation = q4(c.w, c.x, c.y, c.z);
}
XD void store_dyn(const BodyArrays& s, int64_t g, const BodyDyn& d) {
    s.dynA[g] = make_float4(d.position.x, d.position.y, d.position.z, d.velocity.x);
    s.dynB[g] = make_float4(d.velocity.y, d.velocity.z, d.angular_velocity.x, d.angular_velocity.y);
    s.dynC[g] = make_float4(d.rotation.x, d.rotation.y, d.rotation.z, d.rotation.s);
    s.dynD[g] = d.angular_velocity.z;
}
// true when any of the 13 dynamic values is NaN or +-Inf (x * 0 is 0 for finite x and NaN otherwise)
XD bool dyn_nonfinite(const BodyDyn& d) {
    float s = fmul(d.position.x, 0.0f);
    s = fadd(s, fmul(d.position.y, 0.0f)); s = fadd(s, fmul(d.position.z, 0.0f));
    s = fadd(s, fmul(d.rotation.x, 0.0f)); s = fadd(s, fmul(d.rotation.y, 0.0f)); s = fadd(s, fmul(d.rotation.z, 0.0f)); s = fadd(s, fmul(d.rotation.s, 0.0f));
    s = fadd(s, fmul(d.velocity.x, 0.0f)); s = fadd(s, fmul(d.velocity.y, 0.0f)); s = fadd(s, fmul(d.velocity.z, 0.0f));
    s = fadd(s, fmul(d.angular_velocity.x, 0.0f)); s = fadd(s, fmul(d.angular_velocity.y, 0.0f)); s = fadd(s, fmul(d.angular_velocity.z, 0.0f));
    return s != s;
}
XD void load_const(const BodyArrays& s, int64_t g, BodyConst& k) {
    float4 a = __ldg(&s.k0[g]), b = __ldg(&s.k1[g]), c = __ldg(&s.k2[g]), d = __ldg(&s.k3[g]);
    float4 e = __ldg(&s.k4[g]), f = __ldg(&s.k5[g]);
    float t = __ldg(&s.k6[g]);
    k.inverse_mass = a.x;
    k.center_of_mass = v3(a.y, a.z, a.w);
    k.inverse_inertia.c0 = v3(b.x, b.y, b.z);
    k.inverse_inertia.c1 = v3(b.w, c.x, c.y);
    k.inverse_inertia.c2 = v3(c.z, c.w, d.x);
    k.external_force = v3(d.y, d.z, d.w);
    k.internal_force = v3(e.x, e.y, e.z);
    k.external_torque = v3(e.w, f.x, f.y);
    k.internal_torque = v3(f.z, f.w, t);
}

}  // namespace xpbd
