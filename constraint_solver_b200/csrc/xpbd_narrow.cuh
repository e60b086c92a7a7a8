// Narrowphase between two convex hulls: the reference's SAT queries (src/collision.rs:123-197,
// dead code there) finished into a contact manifold.
//
//   face_query / edge_query : warp-cooperative (lanes over faces / edge pairs, shuffle arg-max that
//                             honours the reference's "first strict maximum wins" scan order)
//   choose_feature          : A-face / B-face / edge-edge with Gregorius' bias towards faces
//   face_manifold / edge_manifold : one lane per body pair (Sutherland-Hodgman clipping is sequential)
//
// Inputs are world-space vertices W[b][v] = Rigid::frame() * vertex, computed once per body per substep
// and shared by the ground test and every pair the body takes part in (the reference recomputes
// `frames.0 * vertex` inside every loop; the value is identical).
//
// Support functions: the reference selects the support VERTEX with max_by(total_cmp) and then takes a dot
// product with it (src/collision.rs:138-141,182,188).  max_v dot(v, -n) == -min_v dot(n, v) exactly in
// round-to-nearest arithmetic (negation commutes with every rounding), so only the extreme VALUE is
// needed and vertex ties cannot change it.
#pragma once
#include <float.h>

#include "xpbd_hull.cuh"

namespace xpbd {

constexpr int kMaxManifold = 4;
constexpr int kMaxClip = 16;
constexpr float kRelEdge = 0.90f, kRelFace = 0.98f, kAbsTol = 0.0025f;
enum Feature : int { kNone = 0, kFaceA = 1, kFaceB = 2, kEdgeEdge = 3 };

// World-space / local vertex arrays are float4 per vertex (x, y, z, 0), 16-byte aligned: one LDS.128 per vertex.
XD V3 ld3(const float* a, int i) {
    const float4 v = reinterpret_cast<const float4*>(a)[i];
    return v3(v.x, v.y, v.z);
}
XD void st3(float* a, int i, V3 v) { reinterpret_cast<float4*>(a)[i] = make_float4(v.x, v.y, v.z, 0.0f); }
constexpr int kVertWords = 4;      // words per vertex in those arrays
// warp-private scratch of edge_query: survivor queue (64 x (axis.xyz, pair index)) followed by the per-candidate
// edge table of the compile-time-sized instantiation (edge vectors of A, their feet, edge vectors of B)
constexpr int kEdgeTableMax = 36;
constexpr int kEdgeQueueWords = 4 * (64 + kEdgeTableMax);        // whole warp on one pair: 64-entry queue + table
constexpr int kEdgeQueueWordsHalf = 4 * (36 + kEdgeTableMax);    // 16 lanes on one pair: 32-entry queue (culled query: 144 pair indices) + table

// (value, index) arg-max across the warp: larger value wins, equal values -> lower index (the serial scan
// keeps the FIRST strict maximum), index < 0 = "no candidate".  All lanes receive the result.
// Two warp reductions (REDUX.SYNC) instead of a 5-round shuffle butterfly: the value is mapped to an unsigned key
// whose integer order is the float order (never NaN here: `best` is only ever assigned through `distance > best`;
// -0 is folded into +0 first, which `>` cannot tell apart either), then the lowest index among the lanes that hold
// the maximum is taken (an index of -1 is the largest unsigned value, so real candidates win).
XD void warp_argmax_first(float& d, int& i, unsigned mask = 0xffffffffu) {  // mask: the lane group that cooperates
    const uint32_t u = (uint32_t)__float_as_int(fadd(d, 0.0f));
    const uint32_t key = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    const uint32_t m = __reduce_max_sync(mask, key);
    const uint32_t idx = __reduce_min_sync(mask, key == m ? (uint32_t)i : 0xffffffffu);
    d = __int_as_float((int)((m & 0x80000000u) ? (m & 0x7fffffffu) : ~m));
    i = (int)idx;
}

// The same for two 16-lane groups that run in lockstep (gshift = 0 for lanes 0-15, 16 for lanes 16-31): four
// full-mask reductions in which each half contributes the neutral element to the other half's reduction.  (Reductions
// with a run-time partial mask compile to MATCH-based loops: ~3x the instructions.)
XD void warp_argmax_first_halves(float& d, int& i, int gshift) {
    const uint32_t u = (uint32_t)__float_as_int(fadd(d, 0.0f));
    const uint32_t key = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    const bool hi = gshift != 0;
    const uint32_t m_lo = __reduce_max_sync(0xffffffffu, hi ? 0u : key);
    const uint32_t m_hi = __reduce_max_sync(0xffffffffu, hi ? key : 0u);
    const uint32_t m = hi ? m_hi : m_lo;
    const uint32_t cand = key == m ? (uint32_t)i : 0xffffffffu;
    const uint32_t i_lo = __reduce_min_sync(0xffffffffu, hi ? 0xffffffffu : cand);
    const uint32_t i_hi = __reduce_min_sync(0xffffffffu, hi ? cand : 0xffffffffu);
    d = __int_as_float((int)((m & 0x80000000u) ? (m & 0x7fffffffu) : ~m));
    i = (int)(hi ? i_hi : i_lo);
}

// Support scans: extreme value of dot(axis, v) over the first n vertices of a float4 vertex array.  NV > 0: fully
// unrolled (no loop control, FMNMX3 pairs) for the compile-time-sized instantiations; NV = 0: rolled loop over n.
// (A warp-uniform switch to unrolled 4/8/12-vertex variants inside the generic path was removed again: in
// mixed-hull batches the extra code cost more in instruction-cache misses than the unrolling saved, 1.72 vs 1.77 ms
// on C4.)  min/max are exact, so the order in which they are folded cannot change the value.
template <int NV, bool kMax>
XD float scan_dots_t(const float* W, int n, V3 axis) {
    float s = dot(axis, ld3(W, 0));
    if (NV > 0) {
#pragma unroll
        for (int v = 1; v < NV; ++v) { const float d = dot(axis, ld3(W, v)); s = kMax ? fmaxf(s, d) : fminf(s, d); }
    } else {
        for (int v = 1; v < n; ++v) { const float d = dot(axis, ld3(W, v)); s = kMax ? fmaxf(s, d) : fminf(s, d); }
    }
    return s;
}
template <bool kMax>
XD float scan_dots(const float* W, int n, V3 axis) { return scan_dots_t<0, kMax>(W, n, axis); }

// collision::face_axes_separation (src/collision.rs:123-149) for hull A's faces against hull B's vertices.
// L: warp-private scratch for B's vertices in A-local space (float4 per vertex).
// Lanes are spread over (face, vertex chunk): with nF <= 16 faces, G = 32 / pow2(nF) lane groups each scan every
// G-th vertex and a shuffle-min butterfly across the groups completes the support value of every face
// (cube vs cube: 4 groups x 2 vertices instead of 6 active lanes x 8 vertices).
// NF / NB > 0: face / vertex counts known at compile time (the cube-cube instantiation), 0: taken from nF / nB.
// LG = lanes that cooperate on one body pair: the whole warp, or 16 when a warp works on two pairs at once.  Then
// `lane` is the lane's index inside its half, gshift the index of the half's first lane, and the two halves MUST call
// this function together (lockstep): every sync, shuffle and vote names the full warp, shuffles stay inside a half
// because their xor distance is < 16.
template <int NF, int NB, int LG = 32>
XD void face_query_t(const float4* planes, int nF_rt, const Fr& FinvA, const float* WB, int nB_rt, float* L, int lane,
                     float& out_d, int& out_face, int gshift = 0) {
    const int nF = NF ? NF : nF_rt, nB = NB ? NB : nB_rt;
    for (int v = lane; v < nB; v += LG) st3(L, v, FinvA * ld3(WB, v));  // frames.0.inverse() * (frames.1 * p)
    __syncwarp();
    float best = -FLT_MAX;  // f64::MIN
    int bi = -1;            // usize::MAX
    int nFp = 4;
    while (nFp < nF && nFp < LG) nFp <<= 1;
    int G = LG / nFp;                 // lane groups
    while (G > nB) G >>= 1;           // every group must own at least one vertex
    if (G > 1) {
        const int f = lane & (nFp - 1), g = lane / nFp;
        float m = 0.0f;
        float4 pl = make_float4(0.f, 0.f, 0.f, 0.f);
        if (f < nF && g < G) {
            pl = planes[f];
            const V3 n = v3(pl.x, pl.y, pl.z);
            m = dot(n, ld3(L, g));
            for (int v = g + G; v < nB; v += G) m = fminf(m, dot(n, ld3(L, v)));
        }
        for (int o = nFp; o < nFp * G; o <<= 1) m = fminf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (f < nF && g == 0) {
            const float distance = fsub(m, pl.w);  // plane.distance(support)
            if (distance > best) { best = distance; bi = f; }
        }
    } else {
        for (int f = lane; f < nF; f += LG) {
            const float4 pl = planes[f];
            const float m = scan_dots<false>(L, nB, v3(pl.x, pl.y, pl.z));
            const float distance = fsub(m, pl.w);
            if (distance > best) { best = distance; bi = f; }
        }
    }
    if (LG == 32) warp_argmax_first(best, bi); else warp_argmax_first_halves(best, bi, gshift);
    out_d = best;
    out_face = bi;
    __syncwarp();
}
XD void face_query(const HullTable& T, const HullDesc& hA, const Fr& FinvA, const float* WB, int nB, float* L, int lane,
                   float& out_d, int& out_face) {
    const float4* planes = reinterpret_cast<const float4*>(T.words + hA.plane_off);
    const int nF = hA.n_faces;
    if (nF == 6 && nB == 8) face_query_t<6, 8>(planes, nF, FinvA, WB, nB, L, lane, out_d, out_face);
    else face_query_t<0, 0>(planes, nF, FinvA, WB, nB, L, lane, out_d, out_face);
}

// Plane::from_point_normal (src/geometry.rs:27-36): displacement only.
XD float plane_displacement(V3 point, V3 normal) {
    float d = magnitude(project_on(point, normal));
    if (dot(point, normal) < 0.0f) d = fmul(d, -1.0f);
    return d;
}

// collision::edge_axes_separation (src/collision.rs:151-197).  Edge pairs in A-major order.
// Two passes over the Cartesian product so that the expensive second half only runs on the pairs that survive
// the reference's "no vertex of A further along the axis than the edge foot" test (~40 % for cubes):
//   pass 1 (lanes over edge pairs): axis, orientation flip, support scan of A; survivors are pushed, in order,
//          into the warp-private queue Q (ballot compaction);
//   pass 2 (whenever 32 survivors are queued, and once at the end): plane displacement, support scan of B,
//          distance; lane-local arg-max with the explicit "lower pair index wins ties" rule, which is what the
//          serial scan's strict `>` amounts to (src/collision.rs:190).
// Every value is computed by the same expression as before; only the schedule differs.
// NVA/NEA/NVB/NEB > 0: vertex / edge counts known at compile time (the cube-cube instantiation: the pair loop is
// unrolled, index arithmetic and support scans are straight-line code); 0: taken from the run-time arguments.
// LG / gshift: lanes cooperating on this pair (see face_query_t); gshift = index of the group's first lane.
// With LG = 16 two pairs share the warp and run this function in lockstep: the decision to drain the survivor queue
// is then taken for both halves together (warp vote), each half draining what it has, so that the two halves
// never serialise the expensive second half.
template <int NVA, int NEA, int NVB, int NEB, int LG = 32>
XD void edge_query_t(const int2* EA, const int2* EB, int nA_rt, int nEa_rt, int nB_rt, int nEb_rt, const float* WA,
                     const float* WB, V3 CA, float* Q, int lane, float& out_d, int& out_pair,
                     int gshift = 0) {
    float best = -FLT_MAX;
    int bi = -1;
    const int nA = NVA ? NVA : nA_rt, nB = NVB ? NVB : nB_rt;
    const int nEb = NEB ? NEB : nEb_rt, nPairs = (NEA ? NEA : nEa_rt) * nEb;
    float4* Q4v = reinterpret_cast<float4*>(Q);  // 2 LG entries
    // Small hulls: every edge vector is needed nEb (A) / nEa (B) times — compute each once per candidate.
    constexpr bool kTable = NEA > 0 && NEB > 0 && 2 * NEA + NEB <= kEdgeTableMax;
    float* EV = Q + 4 * 2 * LG;  // [0, NEA) edge_a | [NEA, 2 NEA) foot | [2 NEA, 2 NEA + NEB) edge_b
    if (kTable) {
        for (int t = lane; t < NEA + NEB; t += LG) {
            if (t < NEA) {
                const int2 ea = EA[t];
                const V3 foot = ld3(WA, ea.x);
                st3(EV, t, ld3(WA, ea.y) - foot);
                st3(EV, NEA + t, foot);
            } else {
                const int2 eb = EB[t - NEA];
                st3(EV, NEA + t, ld3(WB, eb.y) - ld3(WB, eb.x));
            }
        }
        __syncwarp();
    }
    auto second_half = [&](float4 q) {
        const V3 axis = v3(q.x, q.y, q.z);
        const int e = __float_as_int(q.w);  // ia << 16 | jb: same order as the flat pair index
        const V3 foot = kTable ? ld3(EV, NEA + (e >> 16)) : ld3(WA, EA[e >> 16].x);
        const float disp = plane_displacement(foot, axis);
        // plane.distance(B.support(-axis)) = min_v dot(axis, v) - displacement
        const float sb = NVB ? scan_dots_t<NVB, false>(WB, nB, axis) : scan_dots<false>(WB, nB, axis);
        const float distance = fsub(sb, disp);
        if (distance > best || (distance == best && e < bi)) { best = distance; bi = e; }
    };
    int qn = 0;  // uniform across the lane group
    const int step_a = LG / nEb, step_b = LG - step_a * nEb;  // advancing the flat pair index by LG
    const unsigned below_me = (1u << lane) - 1u;
    int ia = lane / nEb, jb = lane - ia * nEb;
#ifndef XPBD_DUAL_UNROLL
#define XPBD_DUAL_UNROLL 1
#endif
#ifndef XPBD_SINGLE_UNROLL
#define XPBD_SINGLE_UNROLL 8
#endif
    // Code size matters here: the fully unrolled 16-lane loop measured 1.59 ms against 1.29 ms rolled (instruction
    // cache), and in mixed-hull batches the rolled cube instantiation beats the unrolled one (1.81 vs 2.16 ms on C4).
    // (Folding the final drain into the loop as an extra round, for a single call site of the second half, was
    // slower: 1.35 vs 1.29 ms.)
    constexpr int kUnroll = LG == 32 ? (NEA > 0 ? XPBD_SINGLE_UNROLL : 1) : XPBD_DUAL_UNROLL;
#pragma unroll kUnroll
    for (int e0 = 0; e0 < nPairs; e0 += LG) {
        const int e = e0 + lane;
        bool pass = false;
        V3 axis = v3(0.f, 0.f, 0.f);
        if (e0 + LG <= nPairs || e < nPairs) {
            V3 foot, edge_a, edge_b;
            if (kTable) {
                foot = ld3(EV, NEA + ia); edge_a = ld3(EV, ia); edge_b = ld3(EV, 2 * NEA + jb);
            } else {
                const int2 ea = EA[ia], eb = EB[jb];
                foot = ld3(WA, ea.x);
                edge_a = ld3(WA, ea.y) - foot;
                edge_b = ld3(WB, eb.y) - ld3(WB, eb.x);
            }
            axis = normalize(cross(edge_a, edge_b));  // parallel edges -> NaN axis, no guard (as the reference)
            if (dot(axis, foot - CA) < 0.0f) axis = -axis;
            // Ignore if another point on `a` is further in the direction to `b`.
            const float sa = NVA ? scan_dots_t<NVA, true>(WA, nA, axis) : scan_dots<true>(WA, nA, axis);  // dot(v, axis) == dot(axis, v): IEEE multiplication commutes
            pass = !(sa > dot(foot, axis));
        }
        const unsigned m = LG == 32 ? __ballot_sync(0xffffffffu, pass) : ((__ballot_sync(0xffffffffu, pass) >> gshift) & ((1u << (LG & 31)) - 1u));
        if (pass) Q4v[qn + __popc(m & below_me)] = make_float4(axis.x, axis.y, axis.z, __int_as_float((ia << 16) | jb));
        ia += step_a; jb += step_b;
        if (jb >= nEb) { jb -= nEb; ++ia; }
        qn += __popc(m);
        __syncwarp();
        const bool drain = LG == 32 ? qn >= 32 : __any_sync(0xffffffffu, qn >= LG);
        if (drain) {
            const int take = qn < LG ? qn : LG;
            qn -= take;
            if (lane < take) second_half(Q4v[qn + lane]);
            __syncwarp();
        }
    }
    if (lane < qn) second_half(Q4v[lane]);
    __syncwarp();
    if (LG == 32) warp_argmax_first(best, bi); else warp_argmax_first_halves(best, bi, gshift);
    out_d = best;
    out_pair = bi < 0 ? -1 : (bi >> 16) * nEb + (bi & 0xffff);  // flat pair index ia * nEdgesB + jb
}
// ---- culled edge query ------------------------------------------------------------------------------------------------
// What the step needs from the edge query is only (a) whether some edge axis separates the hulls (distance >= 0) and
// (b) whether the best edge axis beats the faces in choose_feature (distance > kRelEdge * face_max + kAbsTol) — and, only
// then, which pair it is.  So with T = min(0, kRelEdge * face_max + kAbsTol) any edge pair whose distance is PROVABLY
// below T can be dropped without changing the outcome: if the true maximum is >= T its pair is never dropped and
// (distance, pair) come out bit-identical; if it is < T both tests fail for the true and for the culled maximum alike.
//
// The proof per pair is a witness: for ANY vertex w of B,
//     distance = min_v dot(axis, B_v) - displacement  <=  dot(axis, w) - displacement  ~=  dot(axis, w - foot),
// and the sign / size of dot(axis, w - foot) can be read off the UNNORMALISED cross product c = edge_a x edge_b:
//     dot(axis, w - foot) = s * dot(c, w - foot) / |c|,   s = sign of dot(c, foot - centroid_A)   (the reference's flip).
// Stage 1 (all pairs; no sqrt, no reciprocal, no support scan): cull when s * dot(c, w - foot) < -(|T| + M) |c|, compared
// in squared form, where M bounds every rounding difference between this estimate and the fp32 value the reference
// arithmetic produces.  A pair whose flip sign is not clear-cut (|dot(c, foot - centroid_A)| <= M |c|) is kept.  A pair with
// |c|^2 == 0 or NaN (parallel edges; NaN vertices) has a NaN distance in the reference arithmetic, which can never win: culled.
//
// The margin.  u = 2^-24, R = largest |coordinate| of the two hulls' world vertices (so |v| <= sqrt(3) R and
// |B_v - foot| <= 2 sqrt(3) R), c = the fp32 cross product (the cull and the reference form it by the same expression from
// the same fp32 edge vectors: identical bits), c^ = c / |c| exactly.  Worst case, every error aligned:
//   reference  axis = fl(normalize(c)) = c^ (1 + e), |e| <= 4.5 u per component (|c|^2 3u, sqrt, reciprocal, product);
//              its direction c' differs from c^ by <= 9 u per component, its length from 1 by <= 4.5 u;
//              sb   = min_v fl(axis . B_v)                        : 3 u sqrt(3) R = 5.2 u R   of rounding per dot
//              disp = +-fl|fl(axis fl(fl(foot . axis) / fl(axis . axis)))| : 12.1 u R (quotient) + 6.1 u R (length)
//              distance = fl(sb - disp) = c' . (B_v* - foot) + (|axis| - 1) c' . B_v* + the roundings above
//                       <= c^ . (w - foot) + (4.5 sqrt(3) + 5.2 + 18.2) u R + 9 u |w - foot|   <=  c^ . (w - foot) + 62.4 u R;
//   estimate   fl(c . fl(w - foot)) / |c| (thread kernel) is within 4 u |w - foot| <= 13.9 u R of c^ . (w - foot);
//              fl(c . w) - fl(c . foot) (warp kernels) within 10.4 u R; the squared comparison adds a few u RELATIVE to tm.
// Hence estimate < T - M implies distance < T as soon as M >= 76.3 u R = 4.6e-6 R.  The sign of the flip is the sign of
// c^ . (foot - centroid_A) whenever that exceeds ~8 u |foot - centroid_A| (reference) resp. 10.4 u R (estimate) — far below M.
// M = 1e-5 (1 + R): 2.2 x that worst-case bound for large R, more below (round 1 used 4e-5 (1 + R); on C3's resting stacks,
// whose contact depths are 1e-5 ... 3e-4, the tighter margin halves the pairs that reach stage 2: 16.2 -> 8.6 per candidate).
// tests/test_sat_fuzz.py sweeps penetrations and gaps ulp by ulp through 0 and through T at offsets 0 / 1e3 / 1e4 for every
// schedule with and without the cull.
// Stage 2 (survivors, through the warp-private queue): the reference's arithmetic for that pair, unchanged — axis,
// flip, support scan of A, plane displacement, support scan of B, "first strict maximum" tie rule.
// Two witnesses are used (witness_vertices: B's deepest vertex below A's best face and a second deep vertex far from it);
// any vertices would be valid, these sit in the contact region, where they cull most.
#ifdef XPBD_CULL_STATS
__device__ unsigned long long g_cull_stats[3 + 40];  // + histogram of pairs kept per query (capped at 39)  // edge queries, pairs kept by stage 1, pairs seen (debug builds: scripts/cull_stats.sh)
#endif
#ifndef XPBD_CULL_MARGIN
#define XPBD_CULL_MARGIN 1e-5f
#endif
struct EdgeCull {
    float T;    // min(0, kRelEdge * face_max + kAbsTol)
    float M;    // rounding margin, XPBD_CULL_MARGIN * (1 + largest |coordinate| of the two hulls' world vertices)
    V3 w1, w2;  // witnesses: two world-space vertices of B (the smaller of the two bounds is used)
};
XD EdgeCull make_edge_cull(float da, float db, float max_abs_coord, V3 witness1, V3 witness2) {
    EdgeCull c;
    const float face_max = da > db ? da : db;
    c.T = fminf(0.0f, fadd(fmul(kRelEdge, face_max), kAbsTol));
    c.M = XPBD_CULL_MARGIN * (1.0f + max_abs_coord);
    c.w1 = witness1;
    c.w2 = witness2;
    return c;
}
// Index of (approximately) the vertex with the smallest dot(n, L_v): the low bits of the ordered key carry the index.
// Only feeds the cull's witness, so neither ties nor the few mantissa bits given up matter.
// largest |x|, |y|, |z| of a vertex as an unsigned key (bit pattern of a non-negative float; NaN sorts above everything,
// which switches the cull off)
XD uint32_t abs_coord_key(float4 v) {
    const uint32_t a = (uint32_t)__float_as_int(fabsf(v.x)), b = (uint32_t)__float_as_int(fabsf(v.y)), c = (uint32_t)__float_as_int(fabsf(v.z));
    return max(a, max(b, c));
}

template <int LG>
XD int min_key_vertex(uint32_t key, int nB, int gshift) {
    if (LG == 32) {
        key = __reduce_min_sync(0xffffffffu, key);
    } else {
        const bool hi = gshift != 0;
        const uint32_t lo_min = __reduce_min_sync(0xffffffffu, hi ? 0xffffffffu : key);
        const uint32_t hi_min = __reduce_min_sync(0xffffffffu, hi ? key : 0xffffffffu);
        key = hi ? hi_min : lo_min;
    }
    const int v = (int)(key & 63u);
    return v < nB ? v : 0;
}
XD uint32_t ordered_key_with_index(float d, int v) {
    const uint32_t u = (uint32_t)__float_as_int(d);
    return (((u & 0x80000000u) ? ~u : (u | 0x80000000u)) & ~63u) | (uint32_t)v;
}
// Two witnesses for the cull, as vertex indices of B: the deepest vertex below the plane (L = B's vertices in the
// plane's space), and a second one that is deep too but far from the first (depth minus half the distance to the first):
// two points of the contact region that are not collinear with most edge feet, so that only axes close to the contact
// normal survive both bounds.  Heuristics only — any two vertices are valid witnesses.
template <int LG>
XD int2 witness_vertices(float4 plane, const float* L, int nB, int lane, int gshift) {
    const V3 n = v3(plane.x, plane.y, plane.z);
    uint32_t key = 0xffffffffu;
    for (int v = lane; v < nB; v += LG) { const uint32_t k = ordered_key_with_index(dot(n, ld3(L, v)), v); key = k < key ? k : key; }
    const int v1 = min_key_vertex<LG>(key, nB, gshift);
    const V3 p1 = ld3(L, v1);
    key = 0xffffffffu;
    for (int v = lane; v < nB; v += LG) {
        const V3 p = ld3(L, v);
        const uint32_t k = ordered_key_with_index(dot(n, p) - 0.5f * sqrtf(magnitude2(p - p1)), v);
        key = k < key ? k : key;
    }
    return make_int2(v1, min_key_vertex<LG>(key, nB, gshift));
}

template <int NVA, int NEA, int NVB, int NEB, int LG = 32>
XD void edge_query_culled_t(const int2* EA, const int2* EB, int nA_rt, int nEa_rt, int nB_rt, int nEb_rt, const float* WA,
                            const float* WB, V3 CA, float* Q, int lane, EdgeCull cull, float& out_d, int& out_pair,
                            int gshift = 0) {
    float best = -FLT_MAX;
    int bi = -1;
    const int nA = NVA ? NVA : nA_rt, nB = NVB ? NVB : nB_rt;
    const int nEb = NEB ? NEB : nEb_rt, nPairs = (NEA ? NEA : nEa_rt) * nEb;
    constexpr bool kTable = NEA > 0 && NEB > 0 && 2 * NEA + NEB <= kEdgeTableMax;
    // Survivor queue (pair indices ia << 16 | jb) in the words ahead of the edge table.  When the hull sizes are known
    // at compile time and every pair fits (cube pair: 144), the scan never has to drain it.
    constexpr int kRegionWords = LG == 32 ? kEdgeQueueWords : kEdgeQueueWordsHalf;
    constexpr bool kNoDrain = kTable && NEA * NEB <= kRegionWords - 4 * kEdgeTableMax;
    constexpr int kQueueCap = kNoDrain ? kRegionWords - 4 * kEdgeTableMax : 4 * 2 * LG;
    int* Qi = reinterpret_cast<int*>(Q);
    float* EV = Q + kQueueCap;  // [0, NEA) edge_a | [NEA, 2 NEA) foot | [2 NEA, 2 NEA + NEB) edge_b
    if (kTable) {
        for (int t = lane; t < NEA + NEB; t += LG) {
            if (t < NEA) {
                const int2 ea = EA[t];
                const V3 foot = ld3(WA, ea.x);
                st3(EV, t, ld3(WA, ea.y) - foot);
                st3(EV, NEA + t, foot);
            } else {
                const int2 eb = EB[t - NEA];
                st3(EV, NEA + t, ld3(WB, eb.y) - ld3(WB, eb.x));
            }
        }
        __syncwarp();
    }
    auto load_pair = [&](int ia, int jb, V3& foot, V3& edge_a, V3& edge_b) {
        if (kTable) {
            foot = ld3(EV, NEA + ia); edge_a = ld3(EV, ia); edge_b = ld3(EV, 2 * NEA + jb);
        } else {
            const int2 ea = EA[ia], eb = EB[jb];
            foot = ld3(WA, ea.x);
            edge_a = ld3(WA, ea.y) - foot;
            edge_b = ld3(WB, eb.y) - ld3(WB, eb.x);
        }
    };
    // stage 2: the reference's arithmetic for one surviving pair (src/collision.rs:166-193)
    auto exact_pair = [&](int e) {
        V3 foot, edge_a, edge_b;
        load_pair(e >> 16, e & 0xffff, foot, edge_a, edge_b);
        V3 axis = normalize(cross(edge_a, edge_b));
        if (dot(axis, foot - CA) < 0.0f) axis = -axis;
        const float sa = NVA ? scan_dots_t<NVA, true>(WA, nA, axis) : scan_dots<true>(WA, nA, axis);
        if (sa > dot(foot, axis)) return;  // another point of `a` is further in the direction to `b`
        const float disp = plane_displacement(foot, axis);
        const float sb = NVB ? scan_dots_t<NVB, false>(WB, nB, axis) : scan_dots<false>(WB, nB, axis);
        const float distance = fsub(sb, disp);
        if (distance > best || (distance == best && e < bi)) { best = distance; bi = e; }
    };
    auto drain_once = [&](int& qn) {
        __syncwarp();
        const int take = qn < LG ? qn : LG;
        qn -= take;
        if (lane < take) exact_pair(Qi[qn + lane]);
        __syncwarp();
    };
    const float tm = fabsf(cull.T) + cull.M;
    const float K = tm * tm, Kf = cull.M * cull.M;
    int qn = 0;  // uniform across the lane group
#ifdef XPBD_CULL_STATS
    int n_kept = 0;
#endif
    const int step_a = LG / nEb, step_b = LG - step_a * nEb;  // advancing the flat pair index by LG
    // a lane group's share of a full-warp ballot, and the lanes of the group below this one
    const unsigned grp_mask = LG == 32 ? 0xffffffffu : (((1u << (LG & 31)) - 1u) << gshift);
    const unsigned below_me = ((1u << lane) - 1u) << gshift;
    int ia = lane / nEb, jb = lane - ia * nEb;
#ifndef XPBD_CULL_UNROLL
#define XPBD_CULL_UNROLL 1
#endif
    constexpr int kCullUnroll = XPBD_CULL_UNROLL;
#pragma unroll kCullUnroll
    for (int e0 = 0; e0 < nPairs; e0 += LG) {
        bool keep = false;
        if ((NEA * NEB > 0 && (NEA * NEB) % LG == 0) || e0 + LG <= nPairs || e0 + lane < nPairs) {
            V3 foot, edge_a, edge_b;
            load_pair(ia, jb, foot, edge_a, edge_b);
            const V3 c = cross(edge_a, edge_b);
            const float c2 = dot(c, c);
            const float df = dot(c, foot);
            const float f = df - dot(c, CA);
            const float u1 = dot(c, cull.w1) - df, u2 = dot(c, cull.w2) - df;
            const float us = f < 0.0f ? -fmaxf(u1, u2) : fminf(u1, u2);  // the smaller of the two bounds, flipped like the axis
            // (c2 > 1e-24: below that |c|^2 loses precision and the normalised axis is no longer unit length — left to stage 2)
            const bool below = us < 0.0f && fmul(us, us) > fmul(K, c2) && fmul(f, f) > fmul(Kf, c2) && c2 > 1e-24f;
            keep = !below && c2 > 0.0f;
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (keep) Qi[qn + __popc(m & below_me)] = (ia << 16) | jb;
        ia += step_a; jb += step_b;
        if (jb >= nEb) { jb -= nEb; ++ia; }
        qn += __popc(m & grp_mask);
#ifdef XPBD_CULL_STATS
        n_kept += __popc(m & grp_mask);
#endif
        if (!kNoDrain) {
            // Survivors are few, so they stay queued until the scan is over and are then processed LG at a time (full lane
            // groups; two halves in lockstep stay in step) — only a queue about to overflow is drained early.
            const bool drain = LG == 32 ? qn > kQueueCap - LG : __any_sync(0xffffffffu, qn > kQueueCap - LG);
            if (drain) drain_once(qn);
        }
    }
    for (;;) {
        const bool more = LG == 32 ? qn > 0 : __any_sync(0xffffffffu, qn > 0);
        if (!more) break;
        drain_once(qn);
    }
#ifdef XPBD_CULL_STATS
    if (lane == 0) { atomicAdd(&g_cull_stats[0], 1ull); atomicAdd(&g_cull_stats[1], (unsigned long long)n_kept); atomicAdd(&g_cull_stats[2], (unsigned long long)nPairs); atomicAdd(&g_cull_stats[3 + (n_kept < 39 ? n_kept : 39)], 1ull); }
#endif
    if (LG == 32) warp_argmax_first(best, bi); else warp_argmax_first_halves(best, bi, gshift);
    out_d = best;
    out_pair = bi < 0 ? -1 : (bi >> 16) * nEb + (bi & 0xffff);  // flat pair index ia * nEdgesB + jb
}

XD void edge_query(const HullTable& T, const HullDesc& hA, const HullDesc& hB, const float* WA, const float* WB, V3 CA,
                   float* Q, int lane, float& out_d, int& out_pair) {
    const int2* EA = reinterpret_cast<const int2*>(T.words + hA.edge_off);
    const int2* EB = reinterpret_cast<const int2*>(T.words + hB.edge_off);
    const int nA = hA.n_verts, nEa = hA.n_edges, nB = hB.n_verts, nEb = hB.n_edges;
    if (nA == 8 && nEa == 12 && nB == 8 && nEb == 12)
        edge_query_t<8, 12, 8, 12>(EA, EB, nA, nEa, nB, nEb, WA, WB, CA, Q, lane, out_d, out_pair);
    else
        edge_query_t<0, 0, 0, 0>(EA, EB, nA, nEa, nB, nEb, WA, WB, CA, Q, lane, out_d, out_pair);
}

// Result of the three queries for one body pair, kept until the per-lane manifold pass.
struct PairFeature {
    int feature;  // Feature
    int index;    // face index (kFaceA: of A, kFaceB: of B) or flat edge pair index ia * nEdgesB + jb
};

XD PairFeature choose_feature(float da, int fa, float db, int fb, float de, int epair) {
    PairFeature r;
    const float face_max = da > db ? da : db;
    if (epair >= 0 && de > fadd(fmul(kRelEdge, face_max), kAbsTol)) {
        r.feature = kEdgeEdge; r.index = epair;
    } else if (db > fadd(fmul(kRelFace, da), kAbsTol)) {
        r.feature = kFaceB; r.index = fb;
    } else {
        r.feature = kFaceA; r.index = fa;
    }
    return r;
}

struct Manifold {
    int count;
    V3 point[kMaxManifold];   // on the incident body
    V3 target[kMaxManifold];  // on the reference body (projection on the reference plane)
};

// Scratch of face_manifold: two polygon buffers and one depth array of kMaxClip entries each.
// ClipLocal: thread-private arrays (local memory) — the fused kernel and the query kernels.
struct ClipLocal {
    V3 buf[2][kMaxClip];
    float dep[kMaxClip];
    XD V3 get(int w, int k) const { return buf[w][k]; }
    XD void set(int w, int k, V3 v) { buf[w][k] = v; }
    XD float depth(int k) const { return dep[k]; }
    XD void set_depth(int k, float d) { dep[k] = d; }
    XD bool overflowed() const { return false; }
    XD void reset() {}
};
// ClipShared: the same arrays in shared memory, word-interleaved across the block's threads (bank-conflict free),
// with room for CAP <= kMaxClip entries.  A convex polygon gains at most one vertex per clip plane, so
// CAP = 2 x the largest face never overflows geometrically; should rounding noise ever produce more, the entry is
// dropped, overflowed() reports it and the caller repeats the pair with ClipLocal (kMaxClip, the specification's cap).
template <int CAP, int STRIDE>  // entries per buffer, threads per block (compile-time: addresses fold into immediates)
struct ClipShared {
    float* base;  // already offset by the thread's index inside the block
    bool over = false;
    static constexpr int kWordsPerThread = 7 * CAP;
    XD static constexpr int word(int w, int k, int c) { return ((w * CAP + k) * 3 + c) * STRIDE; }
    XD V3 get(int w, int k) const {
        k = k < CAP ? k : CAP - 1;  // only reachable after an overflow (result discarded): stay inside the scratch
        const float* p = base + word(w, k, 0);
        return v3(p[0], p[STRIDE], p[2 * STRIDE]);
    }
    XD void set(int w, int k, V3 v) {
        if (k >= CAP) { over = true; return; }
        float* p = base + word(w, k, 0);
        p[0] = v.x; p[STRIDE] = v.y; p[2 * STRIDE] = v.z;
    }
    XD float depth(int k) const { return base[(6 * CAP + (k < CAP ? k : CAP - 1)) * STRIDE]; }
    XD void set_depth(int k, float d) { if (k < CAP) base[(6 * CAP + k) * STRIDE] = d; }
    XD bool overflowed() const { return over; }
    XD void reset() { over = false; }
};

// Clip polygon `src` (n points) against the half-space dot(s, p - o) <= 0 into polygon `dst`.
template <class Clip>
XD int clip_polygon(Clip& C, int src, int n, V3 o, V3 s, int dst) {
    int m = 0;
    V3 p1 = C.get(src, 0);
    float d1 = dot(s, p1 - o);
    const V3 first = p1;
    const float dfirst = d1;
    for (int k = 0; k < n; ++k) {
        const bool last = k + 1 == n;
        const V3 p2 = last ? first : C.get(src, k + 1);
        const float d2 = last ? dfirst : dot(s, p2 - o);
        const bool in1 = d1 <= 0.0f, in2 = d2 <= 0.0f;
        if (in1 != in2) {
            const float t = fdiv(d1, fsub(d1, d2));
            if (m < kMaxClip) C.set(dst, m++, p1 + (p2 - p1) * t);
        }
        if (in2) {
            if (m < kMaxClip) C.set(dst, m++, p2);
        }
        p1 = p2; d1 = d2;
    }
    return m;
}

// Face contact: reference body R (face r, frame FR, world verts WR), incident body I.  One lane.
// emit(point, target) is called once per kept contact, in polygon order: point on the incident body, target = its
// projection on the reference plane.
template <class Clip, class Emit>
XD void face_manifold(const HullTable& T, const HullDesc& hR, const HullDesc& hI, const Fr& FR, const Fr& FI,
                      const float* WR, const float* WI, int r, Clip& C, Emit& emit) {
    // ref_plane = frames.0 * polytopes.0.plane(face)  (src/collision.rs:66, src/frame.rs:55-64)
    const float4 pl = T.plane(hR, r);
    const V3 nl = v3(pl.x, pl.y, pl.z);
    const V3 support = FR * (pl.w * nl);
    const V3 n = FR.q * nl;
    const float disp = plane_displacement(support, n);
    // incident face: least dot, first strict minimum (src/collision.rs:74-85)
    float least = FLT_MAX;
    int inc = -1;
    for (int k = 0; k < hI.n_faces; ++k) {
        const float4 pk = T.plane(hI, k);
        const float d = dot(FI.q * v3(pk.x, pk.y, pk.z), n);
        if (d < least) { least = d; inc = k; }
    }
    if (inc < 0) return;
    int cnt = 0;
    for (int k = T.face_begin(hI, inc); k < T.face_begin(hI, inc + 1); ++k)
        if (cnt < kMaxClip) C.set(0, cnt++, ld3(WI, T.face_vertex(hI, k)));
    const int rb = T.face_begin(hR, r), nr = T.face_begin(hR, r + 1) - rb;
    int src = 0;
    if (nr <= 4) {
        // triangles and quads: the reference face's vertices are loaded ONCE, all in flight together (as written, every
        // side plane re-loads three of them through the face index list: four dependent global-load latencies in a row)
        V3 wv[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) wv[m] = ld3(WR, T.face_vertex(hR, rb + (m < nr ? m : 0)));
        auto pick = [&](int k) {  // wv[k] for a run-time k (registers cannot be indexed: three selects per component)
            const V3 lo = k & 1 ? wv[1] : wv[0], hi = k & 1 ? wv[3] : wv[2];
            return k & 2 ? hi : lo;
        };
#pragma unroll 1
        for (int m = 0; m < nr && cnt > 0; ++m) {  // rolled: ONE inlined copy of clip_polygon
            const int m1 = (m + 1 >= nr) ? m + 1 - nr : m + 1, m2 = (m + 2 >= nr) ? m + 2 - nr : m + 2;
            const V3 w0 = pick(m), w1 = pick(m1), w2 = pick(m2);
            V3 s = cross(w1 - w0, n);
            if (dot(s, w2 - w0) > 0.0f) s = -s;
            cnt = clip_polygon(C, src, cnt, w0, s, src ^ 1);
            src ^= 1;
        }
    } else {
        for (int m = 0; m < nr && cnt > 0; ++m) {
            const int m1 = (m + 1 >= nr) ? m + 1 - nr : m + 1, m2 = (m + 2 >= nr) ? m + 2 - nr : m + 2;
            const V3 w0 = ld3(WR, T.face_vertex(hR, rb + m));
            const V3 w1 = ld3(WR, T.face_vertex(hR, rb + m1));
            const V3 w2 = ld3(WR, T.face_vertex(hR, rb + m2));
            V3 s = cross(w1 - w0, n);
            if (dot(s, w2 - w0) > 0.0f) s = -s;
            cnt = clip_polygon(C, src, cnt, w0, s, src ^ 1);
            src ^= 1;
        }
    }
    // keep points strictly below the reference plane (compacted into the other buffer)
    const int dst = src ^ 1;
    int c = 0;
    for (int k = 0; k < cnt; ++k) {
        const V3 p = C.get(src, k);
        const float d = fsub(dot(n, p), disp);
        if (d < 0.0f) { C.set(dst, c, p); C.set_depth(c, d); ++c; }
    }
    if (C.overflowed()) return;
    // the (up to four) kept points, in polygon order; scalars + compare-exchange network instead of an indexed array
    // (same result as an insertion sort: the four indices are distinct)
    int k0 = 0, k1 = 1, k2 = 2, k3 = 3, nk = c;
    if (c > kMaxManifold) {
        int i0 = 0;
        float dmin = C.depth(0);
        for (int k = 1; k < c; ++k) { const float d = C.depth(k); if (d < dmin) { dmin = d; i0 = k; } }
        const V3 q0 = C.get(dst, i0);
        int i1 = -1;
        float far = -1.0f;
        for (int k = 0; k < c; ++k) {
            if (k == i0) continue;
            const float d2 = magnitude2(C.get(dst, k) - q0);
            if (d2 > far) { far = d2; i1 = k; }
        }
        int i2 = -1, i3 = -1;
        float amax = 0.0f, amin = 0.0f;
        const V3 e01 = C.get(dst, i1) - q0;
        for (int k = 0; k < c; ++k) {
            if (k == i0 || k == i1) continue;
            const float area = dot(cross(e01, C.get(dst, k) - q0), n);
            if (area > amax) { amax = area; i2 = k; }
            if (area < amin) { amin = area; i3 = k; }
        }
        constexpr int kAbsent = 1 << 20;  // sorts behind every real index
        k0 = i0; k1 = i1; k2 = i2 >= 0 ? i2 : kAbsent; k3 = i3 >= 0 ? i3 : kAbsent;
        nk = 2 + (i2 >= 0) + (i3 >= 0);
        int t;
#define XPBD_CSWAP(a, b) if (b < a) { t = a; a = b; b = t; }
        XPBD_CSWAP(k0, k1) XPBD_CSWAP(k2, k3) XPBD_CSWAP(k0, k2) XPBD_CSWAP(k1, k3) XPBD_CSWAP(k1, k2)
#undef XPBD_CSWAP
    }
#pragma unroll 1
    for (int k = 0; k < nk; ++k) {  // rolled: one copy of the caller's contact code
        const int kk = k == 0 ? k0 : (k == 1 ? k1 : (k == 2 ? k2 : k3));
        const V3 p = C.get(dst, kk);
        emit(p, p - C.depth(kk) * n);  // Plane::project
    }
}

XD float clamp01(float x) { return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x); }

// Closest points of the two edges (Ericson 5.1.9).  One lane.
template <class Emit>
XD void edge_manifold(V3 p1, V3 q1, V3 p2, V3 q2, Emit& emit) {
    const V3 d1 = q1 - p1, d2 = q2 - p2, r = p1 - p2;
    const float a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
    const float c = dot(d1, r), b = dot(d1, d2);
    const float denom = fsub(fmul(a, e), fmul(b, b));
    float s = denom != 0.0f ? clamp01(fdiv(fsub(fmul(b, f), fmul(c, e)), denom)) : 0.0f;
    float t = fdiv(fadd(fmul(b, s), f), e);
    if (t < 0.0f) { t = 0.0f; s = clamp01(fdiv(-c, a)); }
    else if (t > 1.0f) { t = 1.0f; s = clamp01(fdiv(fsub(b, c), a)); }
    emit(p1 + d1 * s, p2 + d2 * t);
}

// Which body carries the contact points of a feature: 0 = A, 1 = B (kFaceA: A is the reference, B's face is clipped).
XD int feature_incident(int feature) { return feature == kFaceA ? 1 : 0; }

// Manifold of one pair from its chosen feature: emit(point, target) per contact, point on body feature_incident(feature).
// The two face cases share ONE inlined copy of face_manifold (operands selected, not two call sites): the clipping code
// is the bulk of the manifold stage's instruction footprint.
template <class Clip, class Emit>
XD void pair_manifold_emit(const HullTable& T, const HullDesc& hA, const HullDesc& hB, const Fr& FA, const Fr& FB,
                           const float* WA, const float* WB, PairFeature pf, Clip& C, Emit& emit) {
    if (pf.feature == kEdgeEdge) {
        const int ia = pf.index / hB.n_edges, jb = pf.index - ia * hB.n_edges;
        const int2 ea = T.edge(hA, ia), eb = T.edge(hB, jb);
        edge_manifold(ld3(WA, ea.x), ld3(WA, ea.y), ld3(WB, eb.x), ld3(WB, eb.y), emit);
    } else if (pf.feature == kFaceA || pf.feature == kFaceB) {
        const bool refB = pf.feature == kFaceB;  // B is the reference, A's face is clipped: the points ride on A
        const HullDesc& hR = refB ? hB : hA;
        const HullDesc& hI = refB ? hA : hB;
        Fr FR, FI;
        FR.p = refB ? FB.p : FA.p; FR.q = refB ? FB.q : FA.q;
        FI.p = refB ? FA.p : FB.p; FI.q = refB ? FA.q : FB.q;
        face_manifold(T, hR, hI, FR, FI, refB ? WB : WA, refB ? WA : WB, pf.index, C, emit);
    }
}
// The specification's own scratch (thread-private arrays of kMaxClip entries): the fallback for a polygon that outgrew a
// smaller shared-memory scratch (face_manifold returns before its first emit then).  Never seen in practice and kept out
// of line, so that it costs no instruction-cache footprint on the paths that run.
template <class Emit>
__device__ __noinline__ void pair_manifold_local_emit(const HullTable& T, const HullDesc& hA, const HullDesc& hB, const Fr& FA,
                                                      const Fr& FB, const float* WA, const float* WB, PairFeature pf, Emit& emit) {
    ClipLocal C;
    pair_manifold_emit(T, hA, hB, FA, FB, WA, WB, pf, C, emit);
}

// The same into a Manifold record (the fused world kernel and the per-function query kernels).
struct ManifoldFill {
    Manifold& m;
    XD void operator()(V3 point, V3 target) {
        if (m.count < kMaxManifold) { m.point[m.count] = point; m.target[m.count] = target; ++m.count; }
    }
};
XD void pair_manifold(const HullTable& T, const HullDesc& hA, const HullDesc& hB, const Fr& FA, const Fr& FB,
                      const float* WA, const float* WB, PairFeature pf, Manifold& m, int& incident) {
    m.count = 0;
    incident = feature_incident(pf.feature);
    ClipLocal C;
    ManifoldFill fill{m};
    pair_manifold_emit(T, hA, hB, FA, FB, WA, WB, pf, C, fill);
}

}  // namespace xpbd
