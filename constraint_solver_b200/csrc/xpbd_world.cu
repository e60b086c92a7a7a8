// Multi-body world step: every substep of every frame of NW worlds
// per block in ONE launch, the worlds resident in shared memory for the whole call.
//
// Mapping (BASELINE.json north_star: "one warp or block per independent world"):
//   * block = NW resident worlds, NWARP warps; hull table staged into shared memory by one TMA bulk copy;
//   * phase 1 (contact generation), one warp per world at a time, 32 lanes cooperating: lanes over bodies
//     (integrate, frames), over (body, vertex) (world-space vertices + ground test, ballot-compacted in the
//     reference's body-major / vertex order), over body pairs (bounding spheres), over faces / edge pairs (SAT
//     queries with shuffle arg-max; the edge query compacts its survivors through a warp-private queue), over
//     candidate pairs (clipping into manifolds);
//   * phase 2 (Gauss-Seidel sweep, derive, velocity pass) is strictly sequential per world (same constraint
//     order as the CPU oracle), so worlds are transposed onto lanes: a LANE PAIR per world, side 0 working on
//     the constraint's body a and side 1 on body b (frame, lever arm, generalized inverse mass and impulse
//     application of the two bodies are independent; three shuffles exchange the contact point, one the
//     inverse mass).  16 worlds fill one solver warp.
// HBM is touched once per step call: 38 floats read + 13 written per body (xpbd_state.cuh).
// The arithmetic is the fp32 restatement of the specification in DESIGN.md ("multi-body world step"),
// expression for expression, so the result is bit-identical to the CPU oracle's S = float instantiation.
#include <cstdlib>

#include "xpbd_internal.hpp"
#include "xpbd_narrow.cuh"
#include "xpbd_rigid.cuh"
#include "xpbd_solve.cuh"
#include "xpbd_tma.cuh"
#include "xpbd_world.cuh"

namespace xpbd {

constexpr int kWorldThreads = 512;  // upper bound on the block size (-> at most 128 registers per thread)

struct WorldStepParams {
    BodyArrays bodies;
    int64_t n_worlds;
    const int32_t* hull_ids;  // per body or nullptr
    const uint32_t* hull_blob;
    uint32_t hull_blob_bytes;
    const DeviceJoint* joints;
    WorldLayout L;
    int NW;  // resident worlds per block (multiple of 16 unless the batch is smaller)
    float frame_dt;
    int substeps, n_frames;
    float compliance, friction;
    int body_body;
    unsigned long long* counters;  // [0] constraints, [1] overflow worlds, [2] tma fallbacks
    unsigned long long* phase_cycles;  // diagnostics (XPBD_PHASE_TIMERS=1), else nullptr: cycles summed over warps
};
enum PhaseTimer { kTBodies, kTVerts, kTBroad, kTSat, kTManifold, kTWait1, kTSolve, kTWait2, kTCount };

XD Fr body_frame(const float* R, const WorldLayout& L, int b) {
    return rigid_frame(sld3(R + L.pos, L.B, b), sldq(R + L.rot, L.B, b), sld3(R + L.com, L.B, b));
}

// ---- phase 1a: per body — save past, Rigid::integrate, frames (lanes over bodies) -------------------
XD void phase_bodies(float* R, float* S, const WorldLayout& L, const HullTable& T, const StepConsts& sc, int lane) {
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
        sst3(R + L.ppos, B, b, d.position); sstq(R + L.prot, B, b, d.rotation); sst3(S + L.s_pfpos, B, b, past.p);
        rigid_integrate(d, k, sc.dt * k.inverse_inertia, sc);
        sst3(R + L.pos, B, b, d.position); sstq(R + L.rot, B, b, d.rotation);
        sst3(R + L.vel, B, b, d.velocity); sst3(R + L.ang, B, b, d.angular_velocity);
        const Fr frame = rigid_frame(d.position, d.rotation, k.center_of_mass);
        const Fr finv = frame_inverse(frame);
        sst3(S + L.s_fpos, B, b, frame.p); sst3(S + L.s_finvp, B, b, finv.p);
        const HullDesc& hd = T.desc(reinterpret_cast<const int*>(R + L.hull)[b]);
        sst3(S + L.s_cen, B, b, frame * T.centroid(hd));
    }
}

// ---- phase 1b: world-space vertices + collision::ground (lanes over (body, vertex)) ------------------
// Appends ground constraints in body-major, hull-vertex order (src/collision.rs:16). Returns the new count.
XD int phase_vertices_ground(float* R, float* S, const WorldLayout& L, const HullTable& T, float friction, int lane,
                             int n_cons, bool& overflow) {
    const int B = L.B, maxV = L.maxV;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    int b = 0, v = lane;  // (body, vertex) of item t = b * maxV + v, advanced incrementally
    for (int t0 = 0; t0 < B * maxV; t0 += 32) {
        while (v >= maxV) { v -= maxV; ++b; }
        bool below = false;
        V3 p = v3(0.f, 0.f, 0.f);
        Fr frame, past;
        if (b < B) {
            const HullDesc& hd = T.desc(hull[b]);
            if (v < hd.n_verts) {
                frame.p = sld3(S + L.s_fpos, B, b); frame.q = sldq(R + L.rot, B, b);
                p = frame * T.vertex(hd, v);
                st3(S + L.s_verts + kVertWords * maxV * b, v, p);
                below = !(p.z >= 0.0f);
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, below);
        if (below) {
            const int idx = n_cons + __popc(m & ((1u << lane) - 1u));
            if (idx < L.capC) {
                past.p = sld3(S + L.s_pfpos, B, b); past.q = sldq(R + L.prot, B, b);
                Fr finv; finv.q = conjugate(frame.q); finv.p = sld3(S + L.s_finvp, B, b);
                const V3 target = v3(p.x, p.y, 0.0f);
                const V3 correction = target - p;
                const V3 delta = p - (past * (finv * p));
                const V3 delta_t = delta - project_on(delta, correction);
                const V3 c1 = target - friction * delta_t;
                const V3 diff = c1 - p;
                const float dist = magnitude(diff);
                const V3 dir = diff * frcp(dist);
                float* c = R + L.cons + kConsWords * idx;
                reinterpret_cast<int*>(c)[0] = cons_head(kGround, b, -1);
                c[1] = p.x; c[2] = p.y; c[3] = p.z; c[4] = dir.x; c[5] = dir.y; c[6] = dir.z; c[7] = dist;
            } else {
                overflow = true;
            }
        }
        n_cons += __popc(m);
        v += 32;
    }
    return n_cons;
}

// ---- phase 1c: broadphase (bounding spheres about the hull centroids), pairs i<j in lexicographic order ---
XD int phase_broadphase(float* R, float* S, const WorldLayout& L, const HullTable& T, int lane, bool& overflow) {
    const int B = L.B;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    int* pairs = reinterpret_cast<int*>(S + L.s_pairs);
    int n = 0;
    for (int i = 0; i + 1 < B; ++i) {
        const V3 ci = sld3(S + L.s_cen, B, i);
        const float ri = T.desc(hull[i]).radius;
        for (int j0 = i + 1; j0 < B; j0 += 32) {
            const int j = j0 + lane;
            bool cand = false;
            if (j < B) {
                const V3 d = sld3(S + L.s_cen, B, j) - ci;
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
XD void phase_sat(float* R, float* S, const WorldLayout& L, const HullTable& T, int n_pairs, int lane) {
    const int B = L.B;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    int* pairs = reinterpret_cast<int*>(S + L.s_pairs);
    for (int p = 0; p < n_pairs; ++p) {
        const int ij = pairs[3 * p], i = ij & 0xffff, j = ij >> 16;
        const HullDesc& hA = T.desc(hull[i]);
        const HullDesc& hB = T.desc(hull[j]);
        const float* WA = S + L.s_verts + kVertWords * L.maxV * i;
        const float* WB = S + L.s_verts + kVertWords * L.maxV * j;
        Fr finvA, finvB;
        finvA.q = conjugate(sldq(R + L.rot, B, i)); finvA.p = sld3(S + L.s_finvp, B, i);
        finvB.q = conjugate(sldq(R + L.rot, B, j)); finvB.p = sld3(S + L.s_finvp, B, j);
        float da, db = 0.f, de = 0.f;
        int fa, fb = -1, ep = -1;
        PairFeature pf; pf.feature = kNone; pf.index = -1;
        face_query(T, hA, finvA, WB, hB.n_verts, S + L.s_local, lane, da, fa);
        if (!(da >= 0.0f)) {
            face_query(T, hB, finvB, WA, hA.n_verts, S + L.s_local, lane, db, fb);
            if (!(db >= 0.0f)) {
                edge_query(T, hA, hB, WA, WB, sld3(S + L.s_cen, B, i), S + L.s_queue, lane, de, ep);
                if (!(de >= 0.0f)) pf = choose_feature(da, fa, db, fb, de, ep);
            }
        }
        if (lane == 0) { pairs[3 * p + 1] = pf.feature; pairs[3 * p + 2] = pf.index; }
    }
    __syncwarp();
}

// ---- phase 1e: manifolds (one lane per candidate pair) -> fresh contact constraints, in pair order -----
XD int phase_manifolds(float* R, float* S, const WorldLayout& L, const HullTable& T, int n_pairs, int lane, int n_cons,
                       bool& overflow) {
    const int B = L.B;
    const int* hull = reinterpret_cast<const int*>(R + L.hull);
    const int* pairs = reinterpret_cast<const int*>(S + L.s_pairs);
    for (int p0 = 0; p0 < n_pairs; p0 += 32) {
        const int p = p0 + lane;
        int count = 0;
        int inc = 0, oth = 0;
        V3 ca[kMaxManifold], cb[kMaxManifold], cn[kMaxManifold];
        if (p < n_pairs && pairs[3 * p + 1] != kNone) {
            const int ij = pairs[3 * p], i = ij & 0xffff, j = ij >> 16;
            PairFeature pf; pf.feature = pairs[3 * p + 1]; pf.index = pairs[3 * p + 2];
            Fr FA, FB;
            FA.p = sld3(S + L.s_fpos, B, i); FA.q = sldq(R + L.rot, B, i);
            FB.p = sld3(S + L.s_fpos, B, j); FB.q = sldq(R + L.rot, B, j);
            Manifold m;
            int incident;
            pair_manifold(T, T.desc(hull[i]), T.desc(hull[j]), FA, FB, S + L.s_verts + kVertWords * L.maxV * i,
                          S + L.s_verts + kVertWords * L.maxV * j, pf, m, incident);
            inc = incident == 0 ? i : j;
            oth = incident == 0 ? j : i;
            Fr finvI, finvO;
            finvI.q = conjugate(sldq(R + L.rot, B, inc)); finvI.p = sld3(S + L.s_finvp, B, inc);
            finvO.q = conjugate(sldq(R + L.rot, B, oth)); finvO.p = sld3(S + L.s_finvp, B, oth);
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
                float* c = R + L.cons + kConsWords * idx;
                reinterpret_cast<int*>(c)[0] = cons_head(kContact, inc, oth);
                c[1] = ca[k].x; c[2] = ca[k].y; c[3] = ca[k].z;
                c[4] = cb[k].x; c[5] = cb[k].y; c[6] = cb[k].z;
                c[7] = cn[k].x; c[8] = cn[k].y; c[9] = cn[k].z;
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
                float* c = R + L.cons + kConsWords * idx;
                reinterpret_cast<int*>(c)[0] = cons_head(kJoint, q.body_a, q.body_b);
                c[1] = q.ax; c[2] = q.ay; c[3] = q.az; c[4] = q.bx; c[5] = q.by; c[6] = q.bz; c[7] = q.distance;
            } else {
                overflow = true;
            }
        }
        n_cons += __popc(m);
    }
    return n_cons;
}

// ---- phase 2: a lane pair per world — Gauss-Seidel sweep, derive, velocity pass (xpbd_solve.cuh) -----------
struct SmemBodies {  // body accessor over a world's persistent shared-memory region
    float* R;
    const WorldLayout& L;
    XD V3 pos(int b) const { return sld3(R + L.pos, L.B, b); }
    XD Q4 rot(int b) const { return sldq(R + L.rot, L.B, b); }
    XD V3 com(int b) const { return sld3(R + L.com, L.B, b); }
    XD float invm(int b) const { return R[L.invm + b]; }
    XD M3 inertia(int b) const { return sldm(R + L.inertia, L.B, b); }
    XD V3 vel(int b) const { return sld3(R + L.vel, L.B, b); }
    XD V3 ang(int b) const { return sld3(R + L.ang, L.B, b); }
    XD V3 ppos(int b) const { return sld3(R + L.ppos, L.B, b); }
    XD Q4 prot(int b) const { return sldq(R + L.prot, L.B, b); }
    XD void set_pose(int b, V3 p, Q4 q) { sst3(R + L.pos, L.B, b, p); sstq(R + L.rot, L.B, b, q); }
    XD void set_vel(int b, V3 v, V3 w) { sst3(R + L.vel, L.B, b, v); sst3(R + L.ang, L.B, b, w); }
};

XD void solve_world(float* R, const WorldLayout& L, int n_cons, const StepConsts& sc, float friction, int side, unsigned pm) {
    SmemBodies A{R, L};
    // --- sweep, in list order
    for (int k = 0; k < n_cons; ++k) {
        __syncwarp(pm);  // the other lane's writes of the previous constraint are visible
        float* c = R + L.cons + kConsWords * k;
        const int head = reinterpret_cast<const int*>(c)[0];
        const int kind = head_kind(head), a = head_a(head), b = head_b(head);
        if (kind == kGround) {
            if (side == 0) solve_ground(A, a, v3(c[1], c[2], c[3]), v3(c[4], c[5], c[6]), c[7], sc);
        } else if (kind == kContact) {
            const bool active = solve_contact(A, side, pm, side ? b : a, side ? v3(c[4], c[5], c[6]) : v3(c[1], c[2], c[3]),
                                              v3(c[7], c[8], c[9]), sc);
            if (active && side == 0) reinterpret_cast<int*>(c)[0] = head | kActiveBit;
        } else {
            solve_joint(A, side, pm, side ? b : a, b >= 0, side ? v3(c[4], c[5], c[6]) : v3(c[1], c[2], c[3]), c[7], sc);
        }
    }
    __syncwarp(pm);
    // --- Rigid::derive for every body, bodies split between the two lanes
    for (int b = side; b < L.B; b += 2) derive_body(A, b, sc.dt);
    // --- velocity pass over the contacts that were active in the sweep
    for (int k = 0; k < n_cons; ++k) {
        __syncwarp(pm);
        const float* c = R + L.cons + kConsWords * k;
        const int head = reinterpret_cast<const int*>(c)[0];
        if ((head & (3 | kActiveBit)) != (kContact | kActiveBit)) continue;
        velocity_contact(A, side, pm, side ? head_b(head) : head_a(head),
                         side ? v3(c[4], c[5], c[6]) : v3(c[1], c[2], c[3]), v3(c[7], c[8], c[9]), friction);
    }
    __syncwarp(pm);
}

// ---- the kernel ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWorldThreads) step_world_kernel(WorldStepParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const WorldLayout& L = P.L;
    const HullTable T{smem};
    const int B = L.B, NW = P.NW;
    const int nwarps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // [hull blob][per-warp scratch, 16-byte aligned][per-world persistent regions]
    float* scratch = reinterpret_cast<float*>(smem) + P.hull_blob_bytes / 4;
    float* regions = scratch + (size_t)nwarps * L.sstride;
    float* S = scratch + (size_t)warp * L.sstride;
    const int64_t world0 = (int64_t)blockIdx.x * NW;
    const int n_here = (int)min((int64_t)NW, P.n_worlds - world0);

    // load the block's worlds: their n_here * B bodies are consecutive in the SoA arrays -> coalesced float4 reads
    for (int t = threadIdx.x; t < n_here * B; t += blockDim.x) {
        const int wl = t / B, b = t - wl * B;
        const int64_t g = world0 * B + t;
        float* R = regions + (size_t)wl * L.stride;
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
        if (b == 0) { reinterpret_cast<int*>(R + L.counters)[0] = 0; reinterpret_cast<int*>(R + L.counters)[1] = 0; }
    }
    __syncthreads();

    const StepConsts sc = make_step_consts(P.frame_dt, P.substeps, P.compliance);
    const int total = P.n_frames * P.substeps;
    // solver role: thread t < 2 * NW is side (t & 1) of world (t >> 1)
    const int swl = threadIdx.x >> 1, side = threadIdx.x & 1;
    const bool solver = swl < n_here;
    const unsigned pm = 3u << (lane & ~1);
    unsigned long long applied = 0;
    const bool timed = P.phase_cycles != nullptr;
    long long tacc[kTCount] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long t0 = 0;
#define TICK() do { if (timed) t0 = clock64(); } while (0)
#define TOCK(which) do { if (timed) { const long long t1 = clock64(); tacc[which] += t1 - t0; t0 = t1; } } while (0)
    for (int s = 0; s < total; ++s) {
        for (int wl = warp; wl < n_here; wl += nwarps) {
            float* R = regions + (size_t)wl * L.stride;
            int* counters = reinterpret_cast<int*>(R + L.counters);
            bool overflow = false;
            TICK();
            phase_bodies(R, S, L, T, sc, lane);
            __syncwarp();
            TOCK(kTBodies);
            int n_cons = phase_vertices_ground(R, S, L, T, P.friction, lane, 0, overflow);
            __syncwarp();
            TOCK(kTVerts);
            if (P.body_body) {
                const int n_pairs = phase_broadphase(R, S, L, T, lane, overflow);
                __syncwarp();
                TOCK(kTBroad);
                phase_sat(R, S, L, T, n_pairs, lane);
                TOCK(kTSat);
                n_cons = phase_manifolds(R, S, L, T, n_pairs, lane, n_cons, overflow);
                TOCK(kTManifold);
            }
            if (L.J > 0) n_cons = phase_joints(R, L, P.joints + (world0 + wl) * L.J, lane, n_cons, overflow);
            overflow = __any_sync(0xffffffffu, overflow);
            if (lane == 0) {
                counters[0] = n_cons < L.capC ? n_cons : L.capC;
                if (overflow) counters[1] = 1;
            }
            __syncwarp();
        }
        TICK();
        __syncthreads();
        TOCK(kTWait1);
        if (solver) {
            float* R = regions + (size_t)swl * L.stride;
            const int n_cons = reinterpret_cast<const int*>(R + L.counters)[0];
            solve_world(R, L, n_cons, sc, P.friction, side, pm);
            if (side == 0) applied += (unsigned long long)n_cons;
        }
        __syncwarp();
        TOCK(kTSolve);
        __syncthreads();
        TOCK(kTWait2);
    }
#undef TICK
#undef TOCK
    if (timed && lane == 0)
        for (int k = 0; k < kTCount; ++k) atomicAdd(P.phase_cycles + k, (unsigned long long)tacc[k]);

    for (int t = threadIdx.x; t < n_here * B; t += blockDim.x) {
        const int wl = t / B, b = t - wl * B;
        const float* R = regions + (size_t)wl * L.stride;
        BodyDyn d;
        d.position = sld3(R + L.pos, B, b); d.rotation = sldq(R + L.rot, B, b);
        d.velocity = sld3(R + L.vel, B, b); d.angular_velocity = sld3(R + L.ang, B, b);
        store_dyn(P.bodies, world0 * B + t, d);
        if (b == 0 && reinterpret_cast<const int*>(R + L.counters)[1]) atomicAdd(P.counters + 1, 1ull);
    }
    if (applied) atomicAdd(P.counters, applied);
}


// Tuning knobs (environment, read once): XPBD_WORLD_NW = resident worlds per block, XPBD_WORLD_WARPS = warps per block.
static int env_int(const char* name, int fallback) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : fallback;
}

int step_world_fused(xpbd_batch* b, float frame_dt, int substeps, int n_frames) {
    const int B = b->desc.bodies_per_world, J = b->desc.joints_per_world;
    const DeviceHulls& H = b->dev_hulls;
    if (B > 1024) return fail(XPBD_ERR_CAPACITY, "bodies_per_world > 1024");
    if (H.max_verts > 64 || H.max_face_verts > kMaxClip / 2)
        return fail(XPBD_ERR_CAPACITY, "multi-body path supports hulls with <= 64 vertices and <= %d vertices per face", kMaxClip / 2);
    if (b->desc.body_body_contacts) {
        for (const HostHull& h : b->hulls)
            if (h.edges.empty()) return fail(XPBD_ERR_INVALID, "body-body contacts need hull edges (edge query)");
    }
    // capacities: every body can rest on the ground with a full face and touches ~4 neighbours with <= 4 points
    const int maxV = H.max_verts;
    const int capC = 6 * B + J + 4;
    const int capP = B * (B - 1) / 2 < 4 * B ? B * (B - 1) / 2 + 1 : 4 * B;
    const WorldLayout L = make_world_layout(B, maxV, capC, capP, J);
    int dev = 0, smem_max = 0, n_sm = 0;
    XPBD_CUDA_TRY(cudaGetDevice(&dev));
    XPBD_CUDA_TRY(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    XPBD_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    const size_t region = (size_t)L.stride * 4, scratch = (size_t)L.sstride * 4;
    const size_t avail = (size_t)smem_max - H.bytes - 64;
    int nwarps = env_int("XPBD_WORLD_WARPS", 8);
    if (nwarps < 1) nwarps = 1;
    if (nwarps > kWorldThreads / 32) nwarps = kWorldThreads / 32;
    if (region + scratch > avail)
        return fail(XPBD_ERR_CAPACITY, "one world needs %zu bytes of shared memory, %zu available", region + scratch, avail);
    // NW resident worlds per block.  Default: 16 (one full solver warp of lane pairs) with two blocks per SM so that
    // one block's sweep overlaps the other's contact generation; fewer when the worlds are large.
    int NW = env_int("XPBD_WORLD_NW", 16);
    const int blocks_per_sm = env_int("XPBD_WORLD_BLOCKS_PER_SM", 2);
    while (NW > 1 && (NW * region + nwarps * scratch + H.bytes + 1024) * blocks_per_sm > (size_t)227 * 1024) NW >>= 1;
    while (NW > 1 && NW * region + nwarps * scratch > avail) NW >>= 1;
    if ((int64_t)NW > b->desc.n_worlds) NW = (int)b->desc.n_worlds;
    if (nwarps > NW) nwarps = NW;
    if (nwarps * 32 < 2 * NW) nwarps = (2 * NW + 31) / 32;  // enough threads for the lane-pair solver
    if (NW * region + nwarps * scratch > avail)
        return fail(XPBD_ERR_CAPACITY, "%d worlds of %zu bytes + %d scratch regions of %zu bytes exceed %zu bytes of shared memory",
                    NW, region, nwarps, scratch, avail);
    const size_t smem_bytes = H.bytes + (size_t)nwarps * scratch + (size_t)NW * region;
    XPBD_CUDA_TRY(cudaFuncSetAttribute(step_world_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));

    WorldStepParams P;
    P.bodies = b->arrays;
    P.n_worlds = b->desc.n_worlds;
    P.hull_ids = b->d_hull_ids;
    P.hull_blob = H.blob;
    P.hull_blob_bytes = H.bytes;
    P.joints = b->d_joints;
    P.L = L;
    P.NW = NW;
    P.frame_dt = frame_dt;
    P.substeps = substeps;
    P.n_frames = n_frames;
    P.compliance = (float)b->desc.compliance;
    P.friction = (float)b->desc.friction;
    P.body_body = b->desc.body_body_contacts;
    P.counters = b->d_counters;
    P.phase_cycles = nullptr;
    static const int timers = env_int("XPBD_PHASE_TIMERS", 0);
    unsigned long long* d_phase = nullptr;
    if (timers) {
        XPBD_CUDA_TRY(cudaMalloc(&d_phase, kTCount * sizeof(unsigned long long)));
        XPBD_CUDA_TRY(cudaMemsetAsync(d_phase, 0, kTCount * sizeof(unsigned long long), b->stream));
        P.phase_cycles = d_phase;
    }
    const int64_t grid = (P.n_worlds + NW - 1) / NW;
    step_world_kernel<<<(unsigned)grid, 32 * nwarps, smem_bytes, b->stream>>>(P);
    XPBD_CUDA_TRY(cudaGetLastError());
    if (timers) {  // diagnostics only: synchronous
        unsigned long long h[kTCount];
        XPBD_CUDA_TRY(cudaMemcpyAsync(h, d_phase, sizeof h, cudaMemcpyDeviceToHost, b->stream));
        XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
        cudaFree(d_phase);
        static const char* names[kTCount] = {"bodies", "verts+ground", "broadphase", "sat", "manifolds", "wait(gen->solve)", "solve", "wait(solve->gen)"};
        const double warps = (double)grid * nwarps, steps = (double)substeps * n_frames;
        fprintf(stderr, "[xpbd phase timers] NW=%d warps=%d grid=%lld smem=%zu: cycles per warp per substep:", NW, nwarps, (long long)grid, smem_bytes);
        for (int k = 0; k < kTCount; ++k) fprintf(stderr, " %s=%.0f", names[k], (double)h[k] / warps / steps);
        fprintf(stderr, "\n");
    }
    b->last.kernel_launches = 1;
    return XPBD_OK;
}

}  // namespace xpbd
