// Device versions of Rigid (src/rigid.rs) and Constraint (src/constraint.rs) arithmetic.
// Same expression trees as the reference; redundant recomputation removed only where the
// recomputed value is bit-identical (common subexpressions).
#pragma once
#include "xpbd_state.cuh"

namespace xpbd {

// Rigid::frame (src/rigid.rs:75-80): position + center_of_mass + rotation * -center_of_mass
XD Fr rigid_frame(V3 position, Q4 rotation, V3 com) {
    Fr f;
    f.p = (position + com) + rotation * (-com);
    f.q = rotation;
    return f;
}

// Per-step-call constants hoisted out of the substep loop (each is the same value every substep).
struct StepConsts {
    float dt;          // dt / substep_count (src/solver.rs:4)
    float half_dt;     // dt * 0.5 (src/rigid.rs:89-90)
    float compliance;  // compliance / (dt*dt) (src/solver.rs:20)
};
XD StepConsts make_step_consts(float frame_dt, int substeps, float compliance) {
    StepConsts c;
    c.dt = fdiv(frame_dt, (float)substeps);
    c.half_dt = fmul(c.dt, 0.5f);
    c.compliance = fdiv(compliance, fmul(c.dt, c.dt));
    return c;
}

// Rigid::integrate (src/rigid.rs:82-99). dtI = dt * inverse_inertia (src/rigid.rs:88 evaluates
// `dt * self.inverse_inertia * torque` left to right, i.e. scaled matrix first).
XD void rigid_integrate(BodyDyn& d, const BodyConst& k, const M3& dtI, const StepConsts& sc) {
    V3 force = k.external_force + d.rotation * k.internal_force;
    d.velocity = d.velocity + (sc.dt * force) * k.inverse_mass;
    d.position = d.position + sc.dt * d.velocity;
    V3 torque = k.external_torque + d.rotation * k.internal_torque;
    d.angular_velocity = d.angular_velocity + dtI * torque;
    Q4 w = q4(0.0f, d.angular_velocity.x, d.angular_velocity.y, d.angular_velocity.z);
    Q4 delta_rotation = (sc.half_dt * w) * d.rotation;
    d.rotation = qnormalize(d.rotation + delta_rotation);
}

// Rigid::derive (src/rigid.rs:101-109)
XD void rigid_derive(BodyDyn& d, V3 past_position, Q4 past_rotation, float dt) {
    d.velocity = (d.position - past_position) / dt;
    Q4 delta = d.rotation * conjugate(past_rotation);
    if (delta.s < 0.0f) delta = -delta;
    d.angular_velocity = (2.0f * qv(delta)) / dt;
}

// Rigid::apply_impulse (src/rigid.rs:113-123): position first, then rotation with the UPDATED
// position; (I^-1 r) x p with body-space I^-1 (reference quirks kept).
XD void rigid_apply_impulse(V3& position, Q4& rotation, float inverse_mass, const M3& inverse_inertia, V3 com,
                            V3 impulse, V3 point) {
    position = position + impulse * inverse_mass;
    V3 arm = inverse_inertia * (point - (position + com));
    Q4 dq = (0.5f * q4(0.0f, cross(arm, impulse).x, cross(arm, impulse).y, cross(arm, impulse).z)) * rotation;
    rotation = qnormalize(rotation + dq);
}

// Constraint::inverse_resitance (sic, src/constraint.rs:25-32) for one body.
XD float inverse_resistance(V3 position, Q4 rotation, float inverse_mass, const M3& inverse_inertia, V3 com,
                            V3 contact, V3 direction) {
    V3 angular_impulse = conjugate(rotation) * cross(contact - (position + com), direction);
    return fadd(inverse_mass, dot(inverse_inertia * angular_impulse, angular_impulse));
}

}  // namespace xpbd
