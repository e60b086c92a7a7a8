// Multi-body world step: every substep of every frame of W worlds
// per block in ONE launch, the whole world resident in shared memory.
//
// Mapping (BASELINE.json north_star: "one warp or block per independent world"):
//   * block = W warps = W worlds; hull table staged into shared memory by one TMA bulk copy;
//   * phase 1, warp w <-> world w, 32 lanes cooperate on CONTACT GENERATION: lanes over bodies (integrate,
//     frames), over (body, vertex) (world-space vertices + ground test, ballot-compacted in the reference's
//     body-major / vertex order), over body pairs (bounding spheres), over faces / edge pairs (SAT queries with
//     shuffle arg-max), over candidate pairs (clipping into manifolds);
//   * phase 2, lane l of warp 0 <-> world l: the Gauss-Seidel sweep, derive and the velocity pass are strictly
//     sequential per world (same constraint order as the CPU oracle), so worlds are transposed onto lanes: one
//     warp instruction advances up to W worlds instead of one.
// HBM is touched once per step call: 38 floats read + 13 written per body (xpbd_state.cuh).
// The arithmetic is the fp32 restatement of the specification in DESIGN.md ("multi-body world step"),
// expression for expression, so the result is bit-identical to the CPU oracle's S = float instantiation.
#include "xpbd_internal.hpp"
#include "xpbd_narrow.cuh"
#include "xpbd_rigid.cuh"
#include "xpbd_tma.cuh"
#include "xpbd_world.cuh"

namespace xpbd {

struct WorldStepParams {
    BodyArrays bodies;
    int64_t n_worlds;
    const int32_t* hull_ids;  // per body or nullptr
    const uint32_t* hull_blob;
    uint32_t hull_blob_bytes;
    const DeviceJoint* joints;
    WorldLayout L;
    int W;  // worlds per block
    float frame_dt;
    int substeps, n_frames;
    float compliance, friction;
    int body_body;
    unsigned long long* counters;  // [0] constraints, [1] overflow worlds, [2] tma fallbacks
};

XD Fr body_frame(const float* R, const WorldLayout& L, int b) {
    return rigid_frame(sld3(R + L.pos, L.B, b), sldq(R + L.rot, L.B, b), sld3(R + L.com, L.B, b));
}

XD void write_constraint_head(float* c, int kind, int ra, int rb) {
    reinterpret_cast<int*>(c)[0] = kind;
    reinterpret_cast<int*>(c)[1] = ra | ((rb + 1) << 16);
}

// ---- phase 1a: per body — save past, Rigid::integrate, frames (lanes over bodies) -------------------
XD void phase_bodies(float* R, const WorldLayout& L, const HullTable& T, const StepConsts& sc, int lane) {
    const int B = L.B;
    for (int b = lane; b < B; b += 32) {
        BodyDyn d;
        BodyConst k;
        d.position = sld3(R + L.pos, B, b); d.rotation = sldq(R + L.rot, B, b);
        d.velocity = sld3(R + L.vel, B, b); d.angular_velocity = sld3(R + L.ang, B, b);
        k.inverse_mass = R[L.invm + b]; k.inverse_inertia = sldm(R + L.inertia, B, b);
        k.center_of_mass = sld3(R + L.com, B, b);
        k.external_force = sld3(R + L.fext, B, b); k.internal_force = sld3(R + L.fint, B, b);
        k.external_torque = sld3(R + L.text, B, b); k.internal_torque = sld3(R + L.tint, B, b);
        const Fr past = rigid_frame(d.position, d.rotation, k.center_of_mass);
        sst3(R + L.ppos, B, b, d.position); sstq(R + L.prot, B, b, d.rotation); sst3(R + L.pfpos, B, b, past.p);
        rigid_integrate(d, k, sc.dt * k.inverse_inertia, sc);
        sst3(R + L.pos, B, b, d.position); sstq(R + L.rot, B, b, d.rotation);
        sst3(R + L.vel, B, b, d.velocity); sst3(R + L.ang, B, b, d.angular_velocity);
        const Fr frame = rigid_frame(d.position, d.rotation, k.center_of_mass);
        const Fr finv = frame_inverse(frame);
        sst3(R + L.fpos, B, b, frame.p); sst3(R + L.finvp, B, b, finv.p);
        const HullDesc& hd = T.desc(reinterpret_cast<const int*>(R + L.hull)[b]);
        sst3(R + L.cen, B, b, frame * T.centroid(hd));
    }
}

// ---- phase 1b: world-space vertices + collision::ground (lanes over (body, vertex)) ------------------
// Appends ground constraints in body-major, hull-vertex order (src/collision.rs:16). Returns the new count.
XD int phase_vertices_ground(float* R, const WorldLayout& L, const HullTable& T, float friction, int lane, int n_cons,
                             bool& overflow) {
    const int B = L.B, maxV = L.maxV;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    for (int t0 = 0; t0 < B * maxV; t0 += 32) {
        const int t = t0 + lane;
        const int b = t / maxV, v = t - b * maxV;
        bool below = false;
        V3 p = v3(0.f, 0.f, 0.f);
        Fr frame, past;
        if (b < B) {
            const HullDesc& hd = T.desc(hull[b]);
            if (v < hd.n_verts) {
                frame.p = sld3(R + L.fpos, B, b); frame.q = sldq(R + L.rot, B, b);
                p = frame * T.vertex(hd, v);
                st3(R + L.verts + 3 * maxV * b, v, p);
                below = !(p.z >= 0.0f);
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, below);
        if (below) {
            const int idx = n_cons + __popc(m & ((1u << lane) - 1u));
            if (idx < L.capC) {
                past.p = sld3(R + L.pfpos, B, b); past.q = sldq(R + L.prot, B, b);
                Fr finv; finv.q = conjugate(frame.q); finv.p = sld3(R + L.finvp, B, b);
                const V3 target = v3(p.x, p.y, 0.0f);
                const V3 correction = target - p;
                const V3 delta = p - (past * (finv * p));
                const V3 delta_t = delta - project_on(delta, correction);
                const V3 c1 = target - friction * delta_t;
                const V3 diff = c1 - p;
                const float dist = magnitude(diff);
                const V3 dir = diff * fdiv(1.0f, dist);
                float* c = R + L.cons + 12 * idx;
                write_constraint_head(c, kGround, b, -1);
                c[2] = p.x; c[3] = p.y; c[4] = p.z; c[5] = dir.x; c[6] = dir.y; c[7] = dir.z; c[11] = dist;
            } else {
                overflow = true;
            }
        }
        n_cons += __popc(m);
    }
    return n_cons;
}

// ---- phase 1c: broadphase (bounding spheres about the hull centroids), pairs i<j in lexicographic order ---
XD int phase_broadphase(float* R, const WorldLayout& L, const HullTable& T, int lane, bool& overflow) {
    const int B = L.B;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    int* pairs = reinterpret_cast<int*>(R + L.pairs);
    int n = 0;
    for (int i = 0; i + 1 < B; ++i) {
        const V3 ci = sld3(R + L.cen, B, i);
        const float ri = T.desc(hull[i]).radius;
        for (int j0 = i + 1; j0 < B; j0 += 32) {
            const int j = j0 + lane;
            bool cand = false;
            if (j < B) {
                const V3 d = sld3(R + L.cen, B, j) - ci;
                const float rsum = fadd(ri, T.desc(hull[j]).radius);
                cand = !(magnitude2(d) > fmul(rsum, rsum));
            }
            const unsigned m = __ballot_sync(0xffffffffu, cand);
            if (cand) {
                const int idx = n + __popc(m & ((1u << lane) - 1u));
                if (idx < L.capP) pairs[3 * idx] = i | (j << 16);
                else overflow = true;
            }
            n += __popc(m);
        }
    }
    return n < L.capP ? n : L.capP;
}

// ---- phase 1d: SAT queries per candidate pair (warp-cooperative) -----------------------------------
XD void phase_sat(float* R, const WorldLayout& L, const HullTable& T, int n_pairs, int lane) {
    const int B = L.B;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    int* pairs = reinterpret_cast<int*>(R + L.pairs);
    for (int p = 0; p < n_pairs; ++p) {
        const int ij = pairs[3 * p], i = ij & 0xffff, j = ij >> 16;
        const HullDesc& hA = T.desc(hull[i]);
        const HullDesc& hB = T.desc(hull[j]);
        const float* WA = R + L.verts + 3 * L.maxV * i;
        const float* WB = R + L.verts + 3 * L.maxV * j;
        Fr finvA, finvB;
        finvA.q = conjugate(sldq(R + L.rot, B, i)); finvA.p = sld3(R + L.finvp, B, i);
        finvB.q = conjugate(sldq(R + L.rot, B, j)); finvB.p = sld3(R + L.finvp, B, j);
        float da, db = 0.f, de = 0.f;
        int fa, fb = -1, ep = -1;
        PairFeature pf; pf.feature = kNone; pf.index = -1;
        face_query(T, hA, finvA, WB, hB.n_verts, R + L.scratchL, lane, da, fa);
        if (!(da >= 0.0f)) {
            face_query(T, hB, finvB, WA, hA.n_verts, R + L.scratchL, lane, db, fb);
            if (!(db >= 0.0f)) {
                edge_query(T, hA, hB, WA, WB, sld3(R + L.cen, B, i), lane, de, ep);
                if (!(de >= 0.0f)) pf = choose_feature(da, fa, db, fb, de, ep);
            }
        }
        if (lane == 0) { pairs[3 * p + 1] = pf.feature; pairs[3 * p + 2] = pf.index; }
    }
    __syncwarp();
}

// ---- phase 1e: manifolds (one lane per candidate pair) -> fresh contact constraints, in pair order -----
XD int phase_manifolds(float* R, const WorldLayout& L, const HullTable& T, int n_pairs, int lane, int n_cons, bool& overflow) {
    const int B = L.B;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    const int* pairs = reinterpret_cast<const int*>(R + L.pairs);
    for (int p0 = 0; p0 < n_pairs; p0 += 32) {
        const int p = p0 + lane;
        int count = 0;
        int inc = 0, oth = 0;
        V3 ca[kMaxManifold], cb[kMaxManifold], cn[kMaxManifold];
        if (p < n_pairs && pairs[3 * p + 1] != kNone) {
            const int ij = pairs[3 * p], i = ij & 0xffff, j = ij >> 16;
            PairFeature pf; pf.feature = pairs[3 * p + 1]; pf.index = pairs[3 * p + 2];
            Fr FA, FB;
            FA.p = sld3(R + L.fpos, B, i); FA.q = sldq(R + L.rot, B, i);
            FB.p = sld3(R + L.fpos, B, j); FB.q = sldq(R + L.rot, B, j);
            Manifold m;
            int incident;
            pair_manifold(T, T.desc(hull[i]), T.desc(hull[j]), FA, FB, R + L.verts + 3 * L.maxV * i,
                          R + L.verts + 3 * L.maxV * j, pf, m, incident);
            inc = incident == 0 ? i : j;
            oth = incident == 0 ? j : i;
            Fr finvI, finvO;
            finvI.q = conjugate(sldq(R + L.rot, B, inc)); finvI.p = sld3(R + L.finvp, B, inc);
            finvO.q = conjugate(sldq(R + L.rot, B, oth)); finvO.p = sld3(R + L.finvp, B, oth);
            for (int k = 0; k < m.count; ++k) {
                const V3 correction = m.target[k] - m.point[k];
                if (magnitude2(correction) == 0.0f) continue;  // no direction: dropped
                ca[count] = finvI * m.point[k];
                cb[count] = finvO * m.target[k];
                cn[count] = normalize(correction);
                ++count;
            }
        }
        // exclusive scan of `count` over the lanes: constraints land in pair order
        int incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        const int base = n_cons + incl - count;
        for (int k = 0; k < count; ++k) {
            const int idx = base + k;
            if (idx < L.capC) {
                float* c = R + L.cons + 12 * idx;
                write_constraint_head(c, kContact, inc, oth);
                c[2] = ca[k].x; c[3] = ca[k].y; c[4] = ca[k].z;
                c[5] = cb[k].x; c[6] = cb[k].y; c[7] = cb[k].z;
                c[8] = cn[k].x; c[9] = cn[k].y; c[10] = cn[k].z;
                c[11] = 0.0f;
            } else {
                overflow = true;
            }
        }
        n_cons += __shfl_sync(0xffffffffu, incl, 31);
    }
    return n_cons;
}

// ---- phase 1f: joints (lanes over joints, index order) ---------------------------------------------------
XD int phase_joints(float* R, const WorldLayout& L, const DeviceJoint* joints, int lane, int n_cons, bool& overflow) {
    for (int j0 = 0; j0 < L.J; j0 += 32) {
        const int j = j0 + lane;
        DeviceJoint q;
        q.body_a = -1;
        if (j < L.J) q = joints[j];
        const bool act = q.body_a >= 0;
        const unsigned m = __ballot_sync(0xffffffffu, act);
        if (act) {
            const int idx = n_cons + __popc(m & ((1u << lane) - 1u));
            if (idx < L.capC) {
                float* c = R + L.cons + 12 * idx;
                write_constraint_head(c, kJoint, q.body_a, q.body_b);
                c[2] = q.ax; c[3] = q.ay; c[4] = q.az; c[5] = q.bx; c[6] = q.by; c[7] = q.bz; c[11] = q.distance;
            } else {
                overflow = true;
            }
        }
        n_cons += __popc(m);
    }
    return n_cons;
}

// ---- phase 2: one lane per world — Gauss-Seidel sweep, derive, velocity pass ----------------------------
XD float body_inverse_resistance(const float* R, const WorldLayout& L, int b, V3 contact, V3 direction) {
    return inverse_resistance(sld3(R + L.pos, L.B, b), sldq(R + L.rot, L.B, b), R[L.invm + b], sldm(R + L.inertia, L.B, b),
                              sld3(R + L.com, L.B, b), contact, direction);
}
XD void body_apply_impulse(float* R, const WorldLayout& L, int b, V3 impulse, V3 point) {
    V3 position = sld3(R + L.pos, L.B, b);
    Q4 rotation = sldq(R + L.rot, L.B, b);
    rigid_apply_impulse(position, rotation, R[L.invm + b], sldm(R + L.inertia, L.B, b), sld3(R + L.com, L.B, b), impulse, point);
    sst3(R + L.pos, L.B, b, position);
    sstq(R + L.rot, L.B, b, rotation);
}

XD void solve_world(float* R, const WorldLayout& L, int n_cons, const StepConsts& sc, float friction) {
    const int B = L.B;
    // --- sweep, in list order
    for (int k = 0; k < n_cons; ++k) {
        float* c = R + L.cons + 12 * k;
        const int kind = reinterpret_cast<const int*>(c)[0];
        const int ab = reinterpret_cast<const int*>(c)[1];
        const int a = ab & 0xffff, b = (ab >> 16) - 1;
        if (kind == kGround) {
            const V3 c0 = v3(c[2], c[3], c[4]), dir = v3(c[5], c[6], c[7]);
            const float dist = c[11];
            const float w = body_inverse_resistance(R, L, a, c0, dir);
            const float lambda = fdiv(fsub(dist, 0.0f), fadd(w, sc.compliance));
            body_apply_impulse(R, L, a, lambda * dir, c0);
        } else if (kind == kContact) {
            const V3 n = v3(c[8], c[9], c[10]);
            const V3 pa = body_frame(R, L, a) * v3(c[2], c[3], c[4]);
            const V3 pb = body_frame(R, L, b) * v3(c[5], c[6], c[7]);
            const float depth = dot(pa - pb, n);
            if (!(depth < 0.0f)) continue;
            reinterpret_cast<int*>(c)[0] = kContact | 0x100;  // active
            const float resistance = fadd(body_inverse_resistance(R, L, a, pa, n), body_inverse_resistance(R, L, b, pb, n));
            const float lambda = fdiv(-depth, fadd(resistance, sc.compliance));
            const V3 impulse = lambda * n;
            body_apply_impulse(R, L, a, impulse, pa);
            body_apply_impulse(R, L, b, -impulse, pb);
        } else {
            const V3 pa = body_frame(R, L, a) * v3(c[2], c[3], c[4]);
            const V3 anchor_b = v3(c[5], c[6], c[7]);
            const V3 pb = b >= 0 ? body_frame(R, L, b) * anchor_b : anchor_b;
            const V3 diff = pb - pa;
            const float dist = magnitude(diff);
            if (dist == 0.0f) continue;
            const V3 dir = diff * fdiv(1.0f, dist);
            float resistance = body_inverse_resistance(R, L, a, pa, dir);
            if (b >= 0) resistance = fadd(resistance, body_inverse_resistance(R, L, b, pb, dir));
            const float lambda = fdiv(fsub(dist, c[11]), fadd(resistance, sc.compliance));
            const V3 impulse = lambda * dir;
            body_apply_impulse(R, L, a, impulse, pa);
            if (b >= 0) body_apply_impulse(R, L, b, -impulse, pb);
        }
    }
    // --- Rigid::derive for every body (src/rigid.rs:101-109)
    for (int b = 0; b < B; ++b) {
        BodyDyn d;
        d.position = sld3(R + L.pos, B, b); d.rotation = sldq(R + L.rot, B, b);
        rigid_derive(d, sld3(R + L.ppos, B, b), sldq(R + L.prot, B, b), sc.dt);
        sst3(R + L.vel, B, b, d.velocity); sst3(R + L.ang, B, b, d.angular_velocity);
    }
    // --- velocity pass over the contacts that were active in the sweep
    for (int k = 0; k < n_cons; ++k) {
        const float* c = R + L.cons + 12 * k;
        if (reinterpret_cast<const int*>(c)[0] != (kContact | 0x100)) continue;
        const int ab = reinterpret_cast<const int*>(c)[1];
        const int a = ab & 0xffff, b = (ab >> 16) - 1;
        const V3 n = v3(c[8], c[9], c[10]);
        const V3 pa = body_frame(R, L, a) * v3(c[2], c[3], c[4]);
        const V3 pb = body_frame(R, L, b) * v3(c[5], c[6], c[7]);
        const V3 ra = pa - (sld3(R + L.pos, B, a) + sld3(R + L.com, B, a));
        const V3 rb = pb - (sld3(R + L.pos, B, b) + sld3(R + L.com, B, b));
        V3 va = sld3(R + L.vel, B, a), wa = sld3(R + L.ang, B, a);
        V3 vb = sld3(R + L.vel, B, b), wb = sld3(R + L.ang, B, b);
        const V3 vrel = (va + cross(wa, ra)) - (vb + cross(wb, rb));
        const float vn = dot(vrel, n);
        const V3 vnormal = n * vn;
        const V3 vt = vrel - vnormal;
        const V3 dv = -(vnormal + vt * friction);
        const float mag = magnitude(dv);
        if (!(mag > 0.0f)) continue;
        const V3 dir = dv * fdiv(1.0f, mag);
        const float resistance = fadd(body_inverse_resistance(R, L, a, pa, dir), body_inverse_resistance(R, L, b, pb, dir));
        const V3 p = dir * fdiv(mag, resistance);
        va = va + p * R[L.invm + a];
        wa = wa + sldm(R + L.inertia, B, a) * cross(ra, p);
        vb = vb - p * R[L.invm + b];
        wb = wb - sldm(R + L.inertia, B, b) * cross(rb, p);
        sst3(R + L.vel, B, a, va); sst3(R + L.ang, B, a, wa);
        sst3(R + L.vel, B, b, vb); sst3(R + L.ang, B, b, wb);
    }
}

// ---- the kernel ------------------------------------------------------------------------------------------
__global__ void step_world_kernel(WorldStepParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const WorldLayout& L = P.L;
    const HullTable T{smem};
    float* regions = reinterpret_cast<float*>(smem) + P.hull_blob_bytes / 4;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t world = (int64_t)blockIdx.x * P.W + warp;
    const bool valid = world < P.n_worlds;
    float* R = regions + (size_t)warp * L.stride;
    const int B = L.B;
    int* counters = reinterpret_cast<int*>(R + L.counters);

    // load the world: coalesced float4 reads of its B consecutive bodies
    if (valid) {
        for (int b = lane; b < B; b += 32) {
            const int64_t g = world * B + b;
            BodyDyn d; BodyConst k;
            load_dyn(P.bodies, g, d);
            load_const(P.bodies, g, k);
            sst3(R + L.pos, B, b, d.position); sstq(R + L.rot, B, b, d.rotation);
            sst3(R + L.vel, B, b, d.velocity); sst3(R + L.ang, B, b, d.angular_velocity);
            R[L.invm + b] = k.inverse_mass;
            float* I = R + L.inertia;
            I[b] = k.inverse_inertia.c0.x; I[B + b] = k.inverse_inertia.c0.y; I[2 * B + b] = k.inverse_inertia.c0.z;
            I[3 * B + b] = k.inverse_inertia.c1.x; I[4 * B + b] = k.inverse_inertia.c1.y; I[5 * B + b] = k.inverse_inertia.c1.z;
            I[6 * B + b] = k.inverse_inertia.c2.x; I[7 * B + b] = k.inverse_inertia.c2.y; I[8 * B + b] = k.inverse_inertia.c2.z;
            sst3(R + L.com, B, b, k.center_of_mass);
            sst3(R + L.fext, B, b, k.external_force); sst3(R + L.fint, B, b, k.internal_force);
            sst3(R + L.text, B, b, k.external_torque); sst3(R + L.tint, B, b, k.internal_torque);
            reinterpret_cast<int*>(R + L.hull)[b] = P.hull_ids ? P.hull_ids[g] : 0;
        }
        if (lane == 0) { counters[0] = 0; counters[1] = 0; counters[2] = 0; }
    }
    __syncwarp();

    const StepConsts sc = make_step_consts(P.frame_dt, P.substeps, P.compliance);
    const int total = P.n_frames * P.substeps;
    unsigned long long applied = 0;
    for (int s = 0; s < total; ++s) {
        if (valid) {
            bool overflow = false;
            phase_bodies(R, L, T, sc, lane);
            __syncwarp();
            int n_cons = phase_vertices_ground(R, L, T, P.friction, lane, 0, overflow);
            __syncwarp();
            if (P.body_body) {
                const int n_pairs = phase_broadphase(R, L, T, lane, overflow);
                __syncwarp();
                phase_sat(R, L, T, n_pairs, lane);
                n_cons = phase_manifolds(R, L, T, n_pairs, lane, n_cons, overflow);
            }
            if (L.J > 0) n_cons = phase_joints(R, L, P.joints + world * L.J, lane, n_cons, overflow);
            overflow = __any_sync(0xffffffffu, overflow);
            if (lane == 0) {
                counters[0] = n_cons < L.capC ? n_cons : L.capC;
                if (overflow) counters[2] = 1;
            }
        }
        __syncthreads();
        if (warp == 0 && lane < P.W && (int64_t)blockIdx.x * P.W + lane < P.n_worlds) {
            float* Rl = regions + (size_t)lane * L.stride;
            const int n_cons = reinterpret_cast<const int*>(Rl + L.counters)[0];
            solve_world(Rl, L, n_cons, sc, P.friction);
            applied += (unsigned long long)n_cons;
        }
        __syncthreads();
    }

    if (valid) {
        for (int b = lane; b < B; b += 32) {
            BodyDyn d;
            d.position = sld3(R + L.pos, B, b); d.rotation = sldq(R + L.rot, B, b);
            d.velocity = sld3(R + L.vel, B, b); d.angular_velocity = sld3(R + L.ang, B, b);
            store_dyn(P.bodies, world * B + b, d);
        }
        if (lane == 0 && counters[2]) atomicAdd(P.counters + 1, 1ull);
    }
    if (applied) atomicAdd(P.counters, applied);
}

struct WorldScratch {
    int W = 0;
    WorldLayout L{};
    size_t smem_bytes = 0;
};

void release_world_scratch(xpbd_batch* b) {
    delete static_cast<WorldScratch*>(b->world_scratch);
    b->world_scratch = nullptr;
}

int step_world(xpbd_batch* b, float frame_dt, int substeps, int n_frames) {
    const int B = b->desc.bodies_per_world, J = b->desc.joints_per_world;
    const DeviceHulls& H = b->dev_hulls;
    if (B > 1024) return fail(XPBD_ERR_CAPACITY, "bodies_per_world > 1024");
    if (H.max_verts > 64 || H.max_face_verts > kMaxClip / 2)
        return fail(XPBD_ERR_CAPACITY, "multi-body path supports hulls with <= 64 vertices and <= %d vertices per face", kMaxClip / 2);
    if (b->desc.body_body_contacts) {
        for (const HostHull& h : b->hulls)
            if (h.edges.empty()) return fail(XPBD_ERR_INVALID, "body-body contacts need hull edges (edge query)");
    }
    // capacities: every body can rest on the ground with a full face, every body touches ~4 neighbours
    const int maxV = H.max_verts;
    const int capC = 8 * B + J + 8;
    const int capP = B * (B - 1) / 2 < 4 * B ? B * (B - 1) / 2 + 1 : 4 * B;
    const WorldLayout L = make_world_layout(B, maxV, capC, capP, J);
    int dev = 0, smem_max = 0;
    XPBD_CUDA_TRY(cudaGetDevice(&dev));
    XPBD_CUDA_TRY(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const size_t region = (size_t)L.stride * 4;
    const size_t budget = (size_t)smem_max - H.bytes - 64;
    if (region > budget)
        return fail(XPBD_ERR_CAPACITY, "one world needs %zu bytes of shared memory, %zu available", region, budget);
    // W worlds per block: enough blocks per SM to overlap the serial solve phase of one block with the
    // contact generation of the others (target ~4 resident blocks), W <= 32 (one solve lane per world).
    int W = (int)((budget / 4) / region);
    if (W < 1) W = 1;
    if (W > 8) W = 8;
    if ((int64_t)W > b->desc.n_worlds) W = (int)b->desc.n_worlds;
    const size_t smem_bytes = H.bytes + (size_t)W * region;
    XPBD_CUDA_TRY(cudaFuncSetAttribute(step_world_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));

    WorldStepParams P;
    P.bodies = b->arrays;
    P.n_worlds = b->desc.n_worlds;
    P.hull_ids = b->d_hull_ids;
    P.hull_blob = H.blob;
    P.hull_blob_bytes = H.bytes;
    P.joints = b->d_joints;
    P.L = L;
    P.W = W;
    P.frame_dt = frame_dt;
    P.substeps = substeps;
    P.n_frames = n_frames;
    P.compliance = (float)b->desc.compliance;
    P.friction = (float)b->desc.friction;
    P.body_body = b->desc.body_body_contacts;
    P.counters = b->d_counters;
    const int64_t grid = (P.n_worlds + W - 1) / W;
    step_world_kernel<<<(unsigned)grid, 32 * W, smem_bytes, b->stream>>>(P);
    XPBD_CUDA_TRY(cudaGetLastError());
    b->last.kernel_launches = 1;
    return XPBD_OK;
}

}  // namespace xpbd
