// SAT stage, THREAD per candidate pair, for batches in which every hull is cube-shaped (8 vertices, 12 edges,
// 6 faces: C3 / C5).  Included by xpbd_staged.cu after StagedParams.
//
// Same queries, same arithmetic per value as the warp-cooperative kernels (face_query_t / edge_query_culled_t in
// xpbd_narrow.cuh; the reference's face_axes_separation / edge_axes_separation, src/collision.rs:123-197) — only the
// schedule differs: one thread walks its candidate's 6 + 6 faces and 144 edge pairs serially, so the "first strict
// maximum" of the reference's scans is a plain `>` and there are no shuffles, votes, ballots, queues or warp syncs.
// The warp-per-two-candidates kernel spends 862 warp-instructions per candidate, i.e. 27.6 k thread slots, on 9.1 k
// useful flops (profiles/r01j_stage_sat.txt); here the instruction stream of a thread is little more than the flops.
//
// Data placement per thread:
//   registers     : the 16 world-space vertices of the two hulls (every support scan indexes them with compile-time
//                   constants after unrolling), B's 12 edge vectors during the cull;
//   shared memory : a per-thread copy of the 16 vertices as float4 [vertex][thread] (LDS.128, conflict-free for any
//                   vertex index) for the accesses whose vertex index is hull data (edge end points, witnesses), and
//                   the per-thread survivor masks of the cull, 12 bits per edge of A.
// No two threads share scratch, so the kernel has no barrier after the hull table is staged.
// (Measured and rejected: running the manifold stage as the epilogue of this thread, with the clip polygons in the same
// scratch column.  SAT + manifold as one kernel took 0.72 ms against 0.48 + 0.20 ms as two: the executed code grows from
// ~2 500 to ~5 000 instructions, "no instruction" stalls rise from 0.16 to 0.55 per issue and issue utilisation drops
// from 75 % to 61 % — instruction cache.)
#pragma once

namespace xpbd {

constexpr int kSatThreadBlock = 128;
#ifndef XPBD_SAT_THREAD_MIN_BLOCKS
#define XPBD_SAT_THREAD_MIN_BLOCKS 4
#endif
// float4 slots per thread: 0..15 the vertices, 16..18 the survivor masks (12 words), 19 = (centroid of A, edge-table
// offsets of the two hulls), 20 = the candidate's best edge pair so far (64-bit key)
constexpr int kSatThreadSlots = 21;
constexpr int kSatQueueCap = 768;  // survivor entries (16 bit) per warp: 24 per candidate; more -> the per-lane walk
__host__ __device__ inline size_t sat_thread_smem_bytes(uint32_t hull_blob_bytes) {
    return (size_t)hull_blob_bytes + (size_t)kSatThreadBlock * kSatThreadSlots * sizeof(float4) +
           (size_t)(kSatThreadBlock / 32) * kSatQueueCap * sizeof(uint16_t);
}

// sqrt for the witness heuristic only (never feeds a result): one MUFU instead of the correctly rounded sequence
XD float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// The thread's scratch: one COLUMN of float4 slots, [slot][thread]; a thread only ever touches its own column.
struct SatThreadScratch {
    float4* col;  // already offset by the thread index: slots 0..7 A's vertices, 8..15 B's, 16..18 survivor masks
    XD V3 vert(int v) const { const float4 x = col[v * kSatThreadBlock]; return v3(x.x, x.y, x.z); }
    XD void set_vert(int v, float4 x) const { col[v * kSatThreadBlock] = x; }
    XD uint32_t* mask(int ia) const { return reinterpret_cast<uint32_t*>(col + (16 + (ia >> 2)) * kSatThreadBlock) + (ia & 3); }
};
// face_axes_separation (src/collision.rs:123-149) of a 6-face hull against 8 vertices `WO` of the other hull; L = those
// vertices in the face hull's local space (kept by the caller for the witnesses).
XD void face_query_thread(const float4* planes, const Fr& Finv, const V3 (&WO)[8], V3 (&L)[8], float& out_d, int& out_face) {
#pragma unroll
    for (int v = 0; v < 8; ++v) L[v] = Finv * WO[v];  // frames.0.inverse() * (frames.1 * p)
    float best = -FLT_MAX;
    int bi = -1;
#pragma unroll 1
    for (int f = 0; f < 6; ++f) {
        const float4 pl = planes[f];
        const V3 n = v3(pl.x, pl.y, pl.z);
        float m = dot(n, L[0]);
#pragma unroll
        for (int v = 1; v < 8; ++v) m = fminf(m, dot(n, L[v]));
        const float distance = fsub(m, pl.w);  // plane.distance(support)
        if (distance > best) { best = distance; bi = f; }
    }
    out_d = best;
    out_face = bi;
}

// The cull's two witnesses (see witness_vertices in xpbd_narrow.cuh): B's deepest vertex below the plane and a second deep
// vertex far from the first.  Heuristic only — any two vertices of B are valid witnesses.
XD int2 witness_vertices_thread(float4 plane, const V3 (&L)[8]) {
    const V3 n = v3(plane.x, plane.y, plane.z);
    float dep[8];
    int v1 = 0;
#pragma unroll
    for (int v = 0; v < 8; ++v) dep[v] = dot(n, L[v]);
    float d1 = dep[0];
#pragma unroll
    for (int v = 1; v < 8; ++v) if (dep[v] < d1) { d1 = dep[v]; v1 = v; }
    V3 p1 = L[0];
#pragma unroll
    for (int v = 1; v < 8; ++v) if (v == v1) p1 = L[v];
    int v2 = 0;
    float k2 = FLT_MAX;
#pragma unroll
    for (int v = 0; v < 8; ++v) {
        const V3 r = L[v] - p1;
        const float k = dep[v] - 0.5f * sqrt_approx(r.x * r.x + r.y * r.y + r.z * r.z);
        if (k < k2) { k2 = k; v2 = v; }
    }
    return make_int2(v1, v2);
}

// What a lane keeps of its candidate between the phases of the kernel loop.
struct SatThreadCandidate {
    float da = 0.f, db = 0.f;
    int fa = -1, fb = -1;
    bool open = false;  // both face queries found penetration: the edge query decides
};

// Phase 1 (thread per candidate): face query A, witnesses, face query B, and stage 1 of the culled edge query — which of
// the 144 edge pairs can matter — into the lane's survivor masks.  A lane without an open candidate leaves zero masks.
template <bool kCull>
XD SatThreadCandidate sat_thread_phase1(const StagedParams& P, const HullTable& T, int4 d, bool valid, const SatThreadScratch& S) {
    SatThreadCandidate out;
#pragma unroll
    for (int k = 0; k < 3; ++k) S.col[(16 + k) * kSatThreadBlock] = make_float4(0.f, 0.f, 0.f, 0.f);  // masks
    S.col[20 * kSatThreadBlock] = make_float4(0.f, 0.f, 0.f, 0.f);                                      // best key = none
    if (!valid) return out;
    const int B = P.B, maxV = P.S.maxV;
    const int64_t world = d.w;
    const int64_t gi = world * B + (d.y & 0xffff), gj = world * B + (d.y >> 16);
    const HullDesc& hA = T.desc(d.z & 0xffff);
    const HullDesc& hB = T.desc(d.z >> 16);
    uint32_t rkey = 0;
    int2 wi = make_int2(0, 0);
    V3 CA;
    {
        V3 WA[8], WB[8], L[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) {
            const float4 x = P.S.verts[gi * maxV + v], y = P.S.verts[gj * maxV + v];
            WA[v] = v3(x.x, x.y, x.z); WB[v] = v3(y.x, y.y, y.z);
            S.set_vert(v, x); S.set_vert(8 + v, y);
            if (kCull) rkey = max(rkey, max(abs_coord_key(x), abs_coord_key(y)));
        }
        const float4 ra = P.bodies.dynC[gi], rb = P.bodies.dynC[gj];
        const float4 fa4 = P.S.finvp[gi], fb4 = P.S.finvp[gj], ca4 = P.S.cen[gi];
        CA = v3(ca4.x, ca4.y, ca4.z);
        S.col[19 * kSatThreadBlock] = make_float4(CA.x, CA.y, CA.z, __int_as_float(hA.edge_off | (hB.edge_off << 16)));
        Fr finvA, finvB;
        finvA.q = conjugate(q4(ra.w, ra.x, ra.y, ra.z)); finvA.p = v3(fa4.x, fa4.y, fa4.z);
        finvB.q = conjugate(q4(rb.w, rb.x, rb.y, rb.z)); finvB.p = v3(fb4.x, fb4.y, fb4.z);
        const float4* planesA = reinterpret_cast<const float4*>(T.words + hA.plane_off);
        const float4* planesB = reinterpret_cast<const float4*>(T.words + hB.plane_off);
        face_query_thread(planesA, finvA, WB, L, out.da, out.fa);
        if (out.da >= 0.0f) return out;
        if (kCull) wi = witness_vertices_thread(planesA[out.fa < 0 ? 0 : out.fa], L);
        face_query_thread(planesB, finvB, WA, L, out.db, out.fb);
        if (out.db >= 0.0f) return out;
    }
    out.open = true;
    const int2* EA = reinterpret_cast<const int2*>(T.words + hA.edge_off);
    const int2* EB = reinterpret_cast<const int2*>(T.words + hB.edge_off);
    if (!kCull) {
#pragma unroll
        for (int k = 0; k < 3; ++k) S.col[(16 + k) * kSatThreadBlock] = make_float4(__int_as_float(0xfff), __int_as_float(0xfff), __int_as_float(0xfff), __int_as_float(0xfff));
        return out;
    }
    const float da = out.da, db = out.db;
    const EdgeCull cull = make_edge_cull(da, db, __int_as_float((int)rkey), S.vert(8 + wi.x), S.vert(8 + wi.y));
    const float tm = fabsf(cull.T) + cull.M;
    const float K = tm * tm, Kf = cull.M * cull.M;
    // B's 12 edge vectors in registers, sorted by parallel class when B is box-like (HullDesc::box_classes; entry k
    // is hull edge perm[k], classes are k = 0-3, 4-7, 8-11), else in hull order
    const bool boxB = hB.box_classes != 0;
    const unsigned long long perm = boxB ? ((unsigned long long)hB.edge_perm_lo | ((unsigned long long)hB.edge_perm_hi << 32))
                                         : 0xBA9876543210ull;
    V3 eb[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) { const int2 e = EB[(int)(perm >> (4 * k)) & 15]; eb[k] = S.vert(8 + e.y) - S.vert(8 + e.x); }
    // Class-level cull (box-like B).  The world-space vectors of B's parallel edges are the same real vector up to the
    // fp32 rounding of their end points.  With u = 2^-24, rho = largest |object-space vertex| of B and R = largest
    // |world coordinate|, frame * vertex = (qv x (qv x v + v s)) 2 + v + p carries at most 49 u rho + u R of rounding
    // per component (worst case, every operation's error aligned), so two parallel edge vectors differ by at most
    // beta = 200 u rho + 4 u R per component (4 end points + the two subtractions), i.e. by sqrt(3) beta in length.
    // Hence c_p = edge_a x b_p differs from +-c_rep by at most e_c = |a| (sqrt(3) beta + rounding of the two cross
    // products), and with N' = -us_rep - e_c max|d_i| - tm e_c and G' = |f_rep| - e_c |dC| - M e_c the conditions
    //     N' > tm |c_rep|,  G' > M |c_rep|,  |c_rep| > 2 e_c
    // imply, for each of the four pairs, exactly the per-pair conditions below (us < 0, us^2 > K c2, f^2 > Kf c2,
    // c2 > 1e-24) in real arithmetic — the flip sign of every pair equals the representative's (up to the pair's
    // edge orientation, which cancels in s c).  One test then drops four pairs; a class that is not provably
    // droppable (the contact's own classes, nearly parallel classes whose cross product is rounding noise,
    // far-away coordinates) falls back to the per-pair tests, so nothing is decided differently — only sooner.
    // (Lengths through sqrt.approx, rounded up by 1e-5.)
    const float rhoB = 1.00001f * sqrt_approx(hB.cx * hB.cx + hB.cy * hB.cy + hB.cz * hB.cz) + hB.radius;
    const float beta = 1.25e-5f * rhoB + 3e-7f * __int_as_float((int)rkey);  // 200 u = 1.19e-5, 4 u = 2.4e-7
    const float e_unit = 1.7321f * beta + 8e-7f * hB.radius;                  // per unit length of edge_a
#pragma unroll 1
    for (int ia = 0; ia < 12; ++ia) {
        const int2 ea = EA[ia];
        const V3 foot = S.vert(ea.x);
        const V3 edge_a = S.vert(ea.y) - foot;
        // per edge of A: the three vectors every pair of this edge is tested against (the estimate c . (w - foot) is
        // taken as one dot product with the difference: fewer operations and less cancellation than c.w - c.foot)
        const V3 d1 = cull.w1 - foot, d2 = cull.w2 - foot, dC = foot - CA;
        const float e_c = 1.00001f * sqrt_approx(edge_a.x * edge_a.x + edge_a.y * edge_a.y + edge_a.z * edge_a.z) * e_unit;
        const float dmax = 1.00001f * sqrt_approx(fmaxf(d1.x * d1.x + d1.y * d1.y + d1.z * d1.z, d2.x * d2.x + d2.y * d2.y + d2.z * d2.z));
        const float h1 = e_c * dmax + tm * e_c;
        const float g0 = e_c * 1.00001f * sqrt_approx(dC.x * dC.x + dC.y * dC.y + dC.z * dC.z) + cull.M * e_c;
        const float c2min = 4.0f * e_c * e_c + 4e-24f;
        uint32_t bits = 0;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            bool class_culled = false;
            if (boxB) {
                const V3 c = cross(edge_a, eb[4 * j]);
                const float c2 = dot(c, c);
                const float f = dot(c, dC);
                const float u1 = dot(c, d1), u2 = dot(c, d2);
                const float us = f < 0.0f ? -fmaxf(u1, u2) : fminf(u1, u2);
                const float Np = -us - h1, Gp = fabsf(f) - g0;
                class_culled = Np > 0.0f && Np * Np > K * c2 && Gp > 0.0f && Gp * Gp > Kf * c2 && c2 > c2min;
            }
            if (!class_culled) {
#pragma unroll
                for (int k = 4 * j; k < 4 * j + 4; ++k) {
                    const V3 c = cross(edge_a, eb[k]);
                    const float c2 = dot(c, c);
                    const float f = dot(c, dC);
                    const float u1 = dot(c, d1), u2 = dot(c, d2);
                    const float us = f < 0.0f ? -fmaxf(u1, u2) : fminf(u1, u2);  // the smaller of the two bounds, flipped like the axis
                    const bool below = us < 0.0f && fmul(us, us) > fmul(K, c2) && fmul(f, f) > fmul(Kf, c2) && c2 > 1e-24f;
                    const bool keep = !below && c2 > 0.0f;
                    bits |= (keep ? 1u : 0u) << ((int)(perm >> (4 * k)) & 15);
                }
            }
        }
        *S.mask(ia) = bits;
    }
    return out;
}

// Phase 2, per-lane form (fallback when the warp's survivors do not fit the queue — no cull, or coordinates so large that
// the cull keeps most pairs): every lane walks ITS OWN survivors with the reference's arithmetic (src/collision.rs:166-193).
XD void sat_thread_walk_own(const HullTable& T, const SatThreadScratch& S, bool open, float& best, int& bi) {
    best = -FLT_MAX;
    bi = -1;
    if (!open) return;
    const float4 info = S.col[19 * kSatThreadBlock];
    const V3 CA = v3(info.x, info.y, info.z);
    const int offs = __float_as_int(info.w);
    const int2* EA = reinterpret_cast<const int2*>(T.words + (offs & 0xffff));
    const int2* EB = reinterpret_cast<const int2*>(T.words + ((unsigned)offs >> 16));
    V3 XA[8], XB[8];
#pragma unroll
    for (int v = 0; v < 8; ++v) { XA[v] = S.vert(v); XB[v] = S.vert(8 + v); }
    int ia = -1;
    uint32_t bits = 0;
    V3 foot = v3(0.f, 0.f, 0.f), edge_a = foot, fc = foot;
    for (;;) {
        if (bits == 0) {
            do {
                if (++ia >= 12) break;
                bits = *S.mask(ia);
            } while (bits == 0);
            if (ia >= 12) break;
            const int2 ea = EA[ia];
            foot = S.vert(ea.x);
            edge_a = S.vert(ea.y) - foot;
            fc = foot - CA;
        }
        const int jb = __ffs((int)bits) - 1;
        bits &= bits - 1;
        const int2 e = EB[jb];
        const V3 edge_b = S.vert(8 + e.y) - S.vert(8 + e.x);
        V3 axis = normalize(cross(edge_a, edge_b));  // parallel edges -> NaN axis, no guard (as the reference)
        if (dot(axis, fc) < 0.0f) axis = -axis;
        // Ignore if another point on `a` is further in the direction to `b`.
        float sa = dot(axis, XA[0]);
#pragma unroll
        for (int v = 1; v < 8; ++v) sa = fmaxf(sa, dot(axis, XA[v]));
        if (sa > dot(foot, axis)) continue;
        const float disp = plane_displacement(foot, axis);
        float sb = dot(axis, XB[0]);
#pragma unroll
        for (int v = 1; v < 8; ++v) sb = fminf(sb, dot(axis, XB[v]));
        const float distance = fsub(sb, disp);
        if (distance > best) { best = distance; bi = ia * 12 + jb; }  // serial scan: strict > keeps the first maximum
    }
}

// (distance, pair) -> 64-bit key whose unsigned order is "larger distance first, then lower pair index": the serial scan's
// first strict maximum.  -0 is folded into +0 (the scan's `>` cannot tell them apart either).  0 = no pair.
XD unsigned long long edge_key(float distance, int pair) {
    const uint32_t u = (uint32_t)__float_as_int(fadd(distance, 0.0f));
    const uint32_t k = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ((unsigned long long)k << 32) | (unsigned long long)(0xffffffffu - (uint32_t)pair);
}

// Phase 2, warp-cooperative form: the survivors of the warp's 32 candidates are compacted into one queue (owner lane,
// edge of A, edge of B) and handed out one per lane, 32 at a time, so that the reference's arithmetic runs on full warps
// whatever the spread of survivor counts between candidates (per-lane walks measured 20 of 32 lanes active before the
// support test of A and 6 of 32 after it).  A lane reads the OWNER's scratch column: consecutive queue entries belong to
// one or two owners, so the 32 lanes of an LDS.128 touch one or two float4 (broadcast).  Every value is computed by the
// same expressions as in the per-lane walk; the per-candidate maximum is a segmented shuffle reduction of edge_key over
// the lanes of one owner, merged into the owner's column by the first lane of the segment.
XD void sat_thread_exact_warp(const HullTable& T, const SatThreadScratch& S, uint16_t* queue, int lane, int total, int my_offset) {
    // every lane appends its own survivors, in scan order
    {
        int off = my_offset, ia = -1;
        uint32_t bits = 0;
        for (;;) {
            if (bits == 0) {
                do {
                    if (++ia >= 12) break;
                    bits = *S.mask(ia);
                } while (bits == 0);
                if (ia >= 12) break;
            }
            const int jb = __ffs((int)bits) - 1;
            bits &= bits - 1;
            queue[off++] = (uint16_t)((lane << 8) | (ia << 4) | jb);
        }
    }
    __syncwarp();
    const float4* warp_cols = S.col - lane;
    for (int g0 = 0; g0 < total; g0 += 32) {
        const int g = g0 + lane;
        const bool active = g < total;
        const int e = active ? (int)queue[g] : 0xff00;
        const int o = e >> 8, ia = (e >> 4) & 15, jb = e & 15;
        unsigned long long key = 0;
        if (active) {
            const float4* col = warp_cols + o;
            const float4 info = col[19 * kSatThreadBlock];
            const V3 CA = v3(info.x, info.y, info.z);
            const int offs = __float_as_int(info.w);
            const int2 ea = reinterpret_cast<const int2*>(T.words + (offs & 0xffff))[ia];
            const int2 eb = reinterpret_cast<const int2*>(T.words + ((unsigned)offs >> 16))[jb];
            auto vert = [&](int v) { const float4 x = col[v * kSatThreadBlock]; return v3(x.x, x.y, x.z); };
            const V3 foot = vert(ea.x);
            const V3 edge_a = vert(ea.y) - foot;
            const V3 edge_b = vert(8 + eb.y) - vert(8 + eb.x);
            V3 axis = normalize(cross(edge_a, edge_b));  // parallel edges -> NaN axis, no guard (as the reference)
            if (dot(axis, foot - CA) < 0.0f) axis = -axis;
            // Ignore if another point on `a` is further in the direction to `b`.
            float sa = dot(axis, vert(0));
#pragma unroll
            for (int v = 1; v < 8; ++v) sa = fmaxf(sa, dot(axis, vert(v)));
            if (!(sa > dot(foot, axis))) {
                const float disp = plane_displacement(foot, axis);
                float sb = dot(axis, vert(8));
#pragma unroll
                for (int v = 1; v < 8; ++v) sb = fminf(sb, dot(axis, vert(8 + v)));
                const float distance = fsub(sb, disp);
                if (distance > -FLT_MAX) key = edge_key(distance, ia * 12 + jb);  // what `distance > best` admits from best = f32::MIN
            }
        }
        // segmented maximum over the lanes of one owner (entries of an owner are contiguous)
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const unsigned long long other = __shfl_down_sync(0xffffffffu, key, off);
            const int other_o = __shfl_down_sync(0xffffffffu, o, off);
            if (lane + off < 32 && other_o == o && other > key) key = other;
        }
        const int prev_o = __shfl_up_sync(0xffffffffu, o, 1);
        if (active && key && (lane == 0 || prev_o != o)) {
            unsigned long long* bk = reinterpret_cast<unsigned long long*>(const_cast<float4*>(warp_cols + o) + 20 * kSatThreadBlock);
            if (key > *bk) *bk = key;
        }
        __syncwarp();
    }
}

template <bool kCull>
__global__ void __launch_bounds__(kSatThreadBlock, XPBD_SAT_THREAD_MIN_BLOCKS) stage_sat_thread_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const HullTable T{smem};
    SatThreadScratch S;
    S.col = reinterpret_cast<float4*>(smem + P.hull_blob_bytes / 4) + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint16_t* queue = reinterpret_cast<uint16_t*>(reinterpret_cast<float4*>(smem + P.hull_blob_bytes / 4) + kSatThreadSlots * kSatThreadBlock) +
                      warp * kSatQueueCap;
    const unsigned count = *P.S.work_count;
    // warp-uniform trip count: the phases below are entered by all 32 lanes together
    for (unsigned base = blockIdx.x * blockDim.x + warp * 32; base < count; base += gridDim.x * blockDim.x) {
        const unsigned it = base + lane;
        const bool valid = it < count;
        int4 d = make_int4(0, 0, 0, 0);
        if (valid) d = P.S.worklist[it];  // slot, i | j << 16, hull_i | hull_j << 16, world
        const SatThreadCandidate c = sat_thread_phase1<kCull>(P, T, d, valid, S);
        // survivors of this lane, and of the warp
        int n = 0;
        if (c.open) {
#pragma unroll
            for (int ia = 0; ia < 12; ++ia) n += __popc(*S.mask(ia));
        }
        int incl = n;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) { const int up = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += up; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        float de = -FLT_MAX;
        int ep = -1;
        if (total <= kSatQueueCap) {
            __syncwarp();
            sat_thread_exact_warp(T, S, queue, lane, total, incl - n);
            const unsigned long long key = *reinterpret_cast<const unsigned long long*>(S.col + 20 * kSatThreadBlock);
            if (key) {
                const uint32_t m = (uint32_t)(key >> 32);
                de = __int_as_float((int)((m & 0x80000000u) ? (m & 0x7fffffffu) : ~m));
                ep = (int)(0xffffffffu - (uint32_t)key);
            }
        } else {
            sat_thread_walk_own(T, S, c.open, de, ep);
        }
        if (valid) {
            PairFeature pf; pf.feature = kNone; pf.index = -1;
            if (c.open && !(de >= 0.0f)) pf = choose_feature(c.da, c.fa, c.db, c.fb, de, ep);
            P.S.pairs[d.x].y = pf.feature;
            P.S.pairs[d.x].z = pf.index;
        }
        __syncwarp();  // the columns are rewritten by the next round's phase 1
    }
}

}  // namespace xpbd
