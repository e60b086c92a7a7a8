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
constexpr int kSatThreadSlots = 19;  // float4 slots per thread: 0..15 the vertices, 16..18 the survivor masks (12 words)
__host__ __device__ inline size_t sat_thread_smem_bytes(uint32_t hull_blob_bytes) {
    return (size_t)hull_blob_bytes + (size_t)kSatThreadBlock * kSatThreadSlots * sizeof(float4);
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

template <bool kCull>
XD void sat_thread_candidate(const StagedParams& P, const HullTable& T, int4 d, const SatThreadScratch& S) {
    const int B = P.B, maxV = P.S.maxV;
    const int64_t world = d.w;
    const int64_t gi = world * B + (d.y & 0xffff), gj = world * B + (d.y >> 16);
    const HullDesc& hA = T.desc(d.z & 0xffff);
    const HullDesc& hB = T.desc(d.z >> 16);
    float da, db = 0.f, de = 0.f;
    int fa, fb = -1, ep = -1;
    PairFeature pf; pf.feature = kNone; pf.index = -1;
    uint32_t rkey = 0;
    int2 wi = make_int2(0, 0);
    V3 CA;
#ifdef XPBD_CULL_STATS
    unsigned st_class = 0, st_class_culled = 0, st_fb_ill = 0, st_fb_well = 0, st_kept_ill = 0, st_kept_well = 0, st_walk_a = 0, st_walk_b = 0;
#endif
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
        Fr finvA, finvB;
        finvA.q = conjugate(q4(ra.w, ra.x, ra.y, ra.z)); finvA.p = v3(fa4.x, fa4.y, fa4.z);
        finvB.q = conjugate(q4(rb.w, rb.x, rb.y, rb.z)); finvB.p = v3(fb4.x, fb4.y, fb4.z);
        const float4* planesA = reinterpret_cast<const float4*>(T.words + hA.plane_off);
        const float4* planesB = reinterpret_cast<const float4*>(T.words + hB.plane_off);
        face_query_thread(planesA, finvA, WB, L, da, fa);
        if (da >= 0.0f) goto done;
        if (kCull) wi = witness_vertices_thread(planesA[fa < 0 ? 0 : fa], L);
        face_query_thread(planesB, finvB, WA, L, db, fb);
        if (db >= 0.0f) goto done;
    }
    {
        const int2* EA = reinterpret_cast<const int2*>(T.words + hA.edge_off);
        const int2* EB = reinterpret_cast<const int2*>(T.words + hB.edge_off);
        // ---- stage 1 of the culled edge query (edge_query_culled_t): which of the 144 pairs can matter
        if (kCull) {
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
            // fp32 rounding of their end points: b_p = +-(b_rep + delta_p).  The spread beta = max |delta_p| (per component) is
            // MEASURED on the twelve vectors in registers (nearly equal numbers subtract exactly; typically a few 1e-7 — the
            // a-priori bound 200 u rho + 4 u R used before was ~50 x larger and sent half of the cullable classes of a resting
            // stack to the per-pair tests).  Then c_p = edge_a x b_p differs from +-c_rep by at most
            //     e_c = |a| (sqrt(3) beta + 4 sqrt(3) u |b|)        (the second term: fp32 rounding of the two cross products,
            //                                                        2 u |a| |b| per component each, |b| <= 2 radius)
            // in length, and with N' = -us_rep - e_c max|d_i| - tm e_c and G' = |f_rep| - e_c |dC| - M e_c the conditions
            //     N' > tm |c_rep|,  G' > M |c_rep|,  |c_rep| > 2 e_c
            // imply, for each of the four pairs, exactly the per-pair conditions below (us < 0, us^2 > K c2, f^2 > Kf c2,
            // c2 > 1e-24) in real arithmetic — the flip sign of every pair equals the representative's (up to the pair's
            // edge orientation, which cancels in s c).  One test then drops four pairs; a class that is not provably
            // droppable (the contact's own classes, nearly parallel classes whose cross product is rounding noise) falls back
            // to the per-pair tests, so nothing is decided differently — only sooner.
            // (Lengths through sqrt.approx, rounded up by 1e-5.)
            float beta = 0.0f;
            if (boxB) {
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const V3 r = eb[4 * j];
#pragma unroll
                    for (int k = 4 * j + 1; k < 4 * j + 4; ++k) {
                        const V3 e = eb[k];
                        const bool opposite = e.x * r.x + e.y * r.y + e.z * r.z < 0.0f;
                        const float dx = opposite ? e.x + r.x : e.x - r.x, dy = opposite ? e.y + r.y : e.y - r.y, dz = opposite ? e.z + r.z : e.z - r.z;
                        beta = fmaxf(beta, fmaxf(fabsf(dx), fmaxf(fabsf(dy), fabsf(dz))));
                    }
                }
                beta = beta != beta ? FLT_MAX : beta;  // NaN vertices: no class is droppable
            }
            const float e_unit = 1.7322f * beta + 9e-7f * hB.radius;  // per unit length of edge_a
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
#ifdef XPBD_CULL_STATS
                        ++st_class; st_class_culled += class_culled; if (!class_culled) { if (c2 > c2min) ++st_fb_well; else ++st_fb_ill; }
#endif
                    }
#ifdef XPBD_CULL_STATS
                    const uint32_t bits_before = bits;
                    bool ill = false;
                    if (boxB) { const V3 c = cross(edge_a, eb[4 * j]); ill = !(dot(c, c) > c2min); }
#endif
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
#ifdef XPBD_CULL_STATS
                    { const unsigned kept = __popc(bits ^ bits_before); if (ill) st_kept_ill += kept; else st_kept_well += kept; }
#endif
                }
                *S.mask(ia) = bits;
            }
        }
        // ---- stage 2: the reference's arithmetic (src/collision.rs:166-193) for the surviving pairs, in scan order
        V3 XA[8], XB[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) { XA[v] = S.vert(v); XB[v] = S.vert(8 + v); }
        float best = -FLT_MAX;
        int bi = -1;
        // Every lane walks ITS OWN survivors (edge of A by edge of A, set bit by set bit), so a warp iterates
        // max-over-lanes(survivor count) times; a loop over the edges of A with an inner loop over the bits would iterate
        // sum-over-edges(max-over-lanes(bits of that edge)) times (measured: 32.7 instead of ~24 per warp on C3).
        int ia = -1;
        uint32_t bits = 0;
        V3 foot = v3(0.f, 0.f, 0.f), edge_a = foot, fc = foot;
        for (;;) {
            if (bits == 0) {
                do {
                    if (++ia >= 12) break;
                    bits = kCull ? *S.mask(ia) : 0xfffu;
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
#ifdef XPBD_CULL_STATS
            ++st_walk_a;
#endif
            if (sa > dot(foot, axis)) continue;
#ifdef XPBD_CULL_STATS
            ++st_walk_b;
#endif
            const float disp = plane_displacement(foot, axis);
            float sb = dot(axis, XB[0]);
#pragma unroll
            for (int v = 1; v < 8; ++v) sb = fminf(sb, dot(axis, XB[v]));
            const float distance = fsub(sb, disp);
            if (distance > best) { best = distance; bi = ia * 12 + jb; }  // serial scan: strict > keeps the first maximum
        }
        de = best; ep = bi;
        if (!(de >= 0.0f)) pf = choose_feature(da, fa, db, fb, de, ep);
#ifdef XPBD_CULL_STATS
        atomicAdd(&g_cull_stats[0], 1ull); atomicAdd(&g_cull_stats[1], (unsigned long long)(st_kept_ill + st_kept_well)); atomicAdd(&g_cull_stats[2], 144ull);
        atomicAdd(&g_cull_stats[3 + 20], (unsigned long long)st_class); atomicAdd(&g_cull_stats[3 + 21], (unsigned long long)st_class_culled);
        atomicAdd(&g_cull_stats[3 + 22], (unsigned long long)st_fb_ill); atomicAdd(&g_cull_stats[3 + 23], (unsigned long long)st_fb_well);
        atomicAdd(&g_cull_stats[3 + 24], (unsigned long long)st_kept_ill); atomicAdd(&g_cull_stats[3 + 25], (unsigned long long)st_kept_well);
        atomicAdd(&g_cull_stats[3 + 26], (unsigned long long)st_walk_a); atomicAdd(&g_cull_stats[3 + 27], (unsigned long long)st_walk_b);
        atomicAdd(&g_cull_stats[3 + 28], (unsigned long long)(pf.feature == kEdgeEdge));
#endif
    }
done:
    P.S.pairs[d.x].y = pf.feature;
    P.S.pairs[d.x].z = pf.index;
}

template <bool kCull>
__global__ void __launch_bounds__(kSatThreadBlock, XPBD_SAT_THREAD_MIN_BLOCKS) stage_sat_thread_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const HullTable T{smem};
    SatThreadScratch S;
    S.col = reinterpret_cast<float4*>(smem + P.hull_blob_bytes / 4) + threadIdx.x;
    if (P.by_level) {
        // Candidates in (pair slot, world) order: the 32 candidates of a warp are the SAME pair slot of 32 consecutive worlds.
        // In a batch of similar worlds (C3: pair slot p is the contact between cubes p and p + 1 of every stack) they take the
        // same branches — the same classes culled, the same number of survivors to walk — where work-list order (the pairs of
        // three or four worlds side by side) mixes the shallow top contacts with the deep bottom ones in every warp.
        const unsigned nw = (unsigned)P.n_worlds, max_pairs = P.S.work_count[1];
        const unsigned long long total = (unsigned long long)max_pairs * nw;
        for (unsigned long long it = blockIdx.x * blockDim.x + threadIdx.x; it < total; it += gridDim.x * blockDim.x) {
            const unsigned p = (unsigned)(it / nw), w = (unsigned)(it - (unsigned long long)p * nw);
            XPBD_CHECK(w < nw && max_pairs <= (unsigned)P.S.capP);
            if (p >= (unsigned)P.S.n_pairs[w]) continue;
            const unsigned slot = w * (unsigned)P.S.capP + p;
            const int ij = P.S.pairs[slot].x;
            XPBD_CHECK((ij & 0xffff) < P.B && (ij >> 16) < P.B);
            int hulls = 0;
            if (P.hull_ids) hulls = P.hull_ids[(int64_t)w * P.B + (ij & 0xffff)] | (P.hull_ids[(int64_t)w * P.B + (ij >> 16)] << 16);
            sat_thread_candidate<kCull>(P, T, make_int4((int)slot, ij, hulls, (int)w), S);
        }
        return;
    }
    const unsigned count = *P.S.work_count;
    for (unsigned it = blockIdx.x * blockDim.x + threadIdx.x; it < count; it += gridDim.x * blockDim.x)
        sat_thread_candidate<kCull>(P, T, P.S.worklist[it], S);
}

}  // namespace xpbd
