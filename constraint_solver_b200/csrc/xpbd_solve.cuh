// Per-constraint pieces of the multi-body Gauss-Seidel sweep and of the velocity pass, written once and used by
// both multi-body paths (bodies in shared memory: xpbd_world.cu; bodies in HBM/L2: xpbd_staged.cu) through a
// body accessor `Acc` with
//     V3 pos(b), Q4 rot(b), V3 com(b), float invm(b), M3 inertia(b), V3 vel(b), V3 ang(b), V3 ppos(b), Q4 prot(b)
//     void set_pose(b, V3, Q4), void set_vel(b, V3, V3)
//
// A LANE PAIR works on one world: side 0 owns the constraint's body a, side 1 its body b.  Frame, lever arm,
// generalized inverse mass (Constraint::inverse_resitance, src/constraint.rs:25-32) and impulse application
// (Rigid::apply_impulse, src/rigid.rs:113-123) of the two bodies are independent; the contact point (3 shuffles)
// and the inverse mass (1 shuffle) are exchanged inside the pair.  Values the serial specification computes once
// (depth, lambda, direction) are computed redundantly by both lanes from identical inputs, so every result is
// bit-identical to the serial evaluation order of the specification.
#pragma once
#include "xpbd_rigid.cuh"

namespace xpbd {

// exchange with the other lane of the pair (pm = the pair's two-lane mask)
XD float swap1(float v, unsigned pm) { return __shfl_xor_sync(pm, v, 1); }
XD V3 swap3(V3 v, unsigned pm) { return v3(swap1(v.x, pm), swap1(v.y, pm), swap1(v.z, pm)); }

template <class Acc>
XD Fr acc_frame(const Acc& A, int b) { return rigid_frame(A.pos(b), A.rot(b), A.com(b)); }
template <class Acc>
XD float acc_inverse_resistance(const Acc& A, int b, V3 contact, V3 direction) {
    return inverse_resistance(A.pos(b), A.rot(b), A.invm(b), A.inertia(b), A.com(b), contact, direction);
}
template <class Acc>
XD void acc_apply_impulse(Acc& A, int b, V3 impulse, V3 point) {
    V3 position = A.pos(b);
    Q4 rotation = A.rot(b);
    rigid_apply_impulse(position, rotation, A.invm(b), A.inertia(b), A.com(b), impulse, point);
    A.set_pose(b, position, rotation);
}

// Ground constraint (the reference's single-body form, src/solver.rs:22-26): stale world-space contact point c0,
// unit direction and |contact1 - contact0| were fixed at generation time.  Side 0 only.
template <class Acc>
XD void solve_ground(Acc& A, int a, V3 c0, V3 dir, float dist, const StepConsts& sc) {
    const float w = acc_inverse_resistance(A, a, c0, dir);
    const float lambda = fdiv(fsub(dist, 0.0f), fadd(w, sc.compliance));
    acc_apply_impulse(A, a, lambda * dir, c0);
}

// Body-body contact, evaluated fresh from the body-local anchors.  `me` = this side's body, `anchor` = its local
// point.  Returns true when the contact was active (depth < 0) — identical on both lanes.
template <class Acc>
XD bool solve_contact(Acc& A, int side, unsigned pm, int me, V3 anchor, V3 n, const StepConsts& sc) {
    const V3 p_me = acc_frame(A, me) * anchor;
    const V3 p_ot = swap3(p_me, pm);
    const float depth = side ? dot(p_ot - p_me, n) : dot(p_me - p_ot, n);  // dot(pa - pb, n)
    if (!(depth < 0.0f)) return false;
    const float w_me = acc_inverse_resistance(A, me, p_me, n);
    const float resistance = fadd(w_me, swap1(w_me, pm));  // w_a + w_b (IEEE addition commutes)
    const float lambda = fdiv(-depth, fadd(resistance, sc.compliance));
    const V3 impulse = lambda * n;
    acc_apply_impulse(A, me, side ? -impulse : impulse, p_me);
    return true;
}

// Joint (two-body distance constraint between body-local anchors; me < 0 on side 1: world-space anchor).
template <class Acc>
XD void solve_joint(Acc& A, int side, unsigned pm, int me, bool two_body, V3 anchor, float rest, const StepConsts& sc) {
    const V3 p_me = me >= 0 ? acc_frame(A, me) * anchor : anchor;
    const V3 p_ot = swap3(p_me, pm);
    const V3 diff = side ? p_me - p_ot : p_ot - p_me;  // pb - pa
    const float dist = magnitude(diff);
    if (dist == 0.0f) return;
    const V3 dir = diff * frcp(dist);
    const float w_me = me >= 0 ? acc_inverse_resistance(A, me, p_me, dir) : 0.0f;
    const float w_ot = swap1(w_me, pm);
    const float w_a = side ? w_ot : w_me, w_b = side ? w_me : w_ot;
    const float resistance = two_body ? fadd(w_a, w_b) : w_a;
    const float lambda = fdiv(fsub(dist, rest), fadd(resistance, sc.compliance));
    const V3 impulse = lambda * dir;
    if (me >= 0) acc_apply_impulse(A, me, side ? -impulse : impulse, p_me);
}

// Rigid::derive (src/rigid.rs:101-109) for body b.
template <class Acc>
XD void derive_body(Acc& A, int b, float dt) {
    BodyDyn d;
    d.position = A.pos(b); d.rotation = A.rot(b);
    rigid_derive(d, A.ppos(b), A.prot(b), dt);
    A.set_vel(b, d.velocity, d.angular_velocity);
}

// Velocity pass for one contact that was active in the sweep (Mueller et al. 2020 "solveVelocities",
// restitution 0): removes the relative normal velocity and friction x the relative tangential velocity.
template <class Acc>
XD void velocity_contact(Acc& A, int side, unsigned pm, int me, V3 anchor, V3 n, float friction) {
    const V3 p_me = acc_frame(A, me) * anchor;
    const V3 r_me = p_me - (A.pos(me) + A.com(me));
    V3 v_me = A.vel(me), w_me = A.ang(me);
    const V3 u_me = v_me + cross(w_me, r_me);
    const V3 u_ot = swap3(u_me, pm);
    const V3 vrel = side ? u_ot - u_me : u_me - u_ot;  // (va + wa x ra) - (vb + wb x rb)
    const float vn = dot(vrel, n);
    const V3 vnormal = n * vn;
    const V3 vt = vrel - vnormal;
    const V3 dv = -(vnormal + vt * friction);
    const float mag = magnitude(dv);
    if (!(mag > 0.0f)) return;
    const V3 dir = dv * frcp(mag);
    const float res_me = acc_inverse_resistance(A, me, p_me, dir);
    const float resistance = fadd(res_me, swap1(res_me, pm));
    const V3 p = dir * fdiv(mag, resistance);
    const V3 dlin = p * A.invm(me);
    const V3 dang = A.inertia(me) * cross(r_me, p);
    if (side) { v_me = v_me - dlin; w_me = w_me - dang; }
    else      { v_me = v_me + dlin; w_me = w_me + dang; }
    A.set_vel(me, v_me, w_me);
}

}  // namespace xpbd
