// Multi-body world step, STAGED path: the same specification and the
// same device functions as the fused kernel in xpbd_world.cu, but one kernel per stage of the substep, with the
// per-world intermediates (world-space hull vertices, frames, candidate pairs, manifolds) in HBM/L2 instead of
// shared memory.  Why: the fused kernel can only keep ~32 worlds resident per SM (shared memory) at 128
// registers per thread, so its FP32 work runs at 16 warps per SM and ~43 % issue utilisation; here every stage
// gets its own register budget and the parallel axis that suits it:
//
//   prepare  : warp per world       Rigid::integrate, frames, world-space vertices, collision::ground
//                                   (ballot-compacted in the reference's order), bounding-sphere broadphase ->
//                                   candidate pair slots + a global work list
//   sat      : warp per candidate   face query A, face query B, edge query (shuffle arg-max), feature choice
//   manifold : thread per candidate Sutherland-Hodgman clipping -> <= 4 fresh contacts in the pair's slot
//   solve    : lane pair per world  Gauss-Seidel sweep in the specification's order (ground, pairs i<j, joints),
//                                   Rigid::derive, velocity pass (xpbd_solve.cuh)
//
// Worlds stay independent (no atomics on the state; the only atomic is the work-list append, whose order does not
// influence results because every candidate writes into its own (world, pair) slot).  Per substep the extra HBM/L2
// traffic is ~5 KB per world — about 1 % of the measured HBM bandwidth at the achieved rate.
#include <cstdlib>
#include <cstring>

#include "xpbd_host_math.hpp"
#include "xpbd_internal.hpp"
#include "xpbd_narrow.cuh"
#include "xpbd_rigid.cuh"
#include "xpbd_solve.cuh"
#include "xpbd_tma.cuh"
#include "xpbd_world.cuh"

namespace xpbd {

constexpr int kGroundWords = 8;    // head, contact0.xyz, direction.xyz, distance

struct StagedBuffers {
    float4* ppos = nullptr;   // per body: pose at the start of the substep
    float4* prot = nullptr;
    float4* fpos = nullptr;   // per body: post-integrate frame position, inverse frame position, world centroid
    float4* finvp = nullptr;
    float4* cen = nullptr;
    float4* verts = nullptr;  // per body x maxV: world-space hull vertices
    float* gcons = nullptr;   // per world x capG x kGroundWords
    int* n_ground = nullptr;  // per world
    int4* pairs = nullptr;    // per world x capP: (i | j << 16, feature, index, n_contacts)
    int* n_pairs = nullptr;   // per world
    float4* pcons = nullptr;  // per (world, pair slot) x kMaxManifold x 3 float4
    int4* worklist = nullptr;  // per candidate: (world * capP + pair slot, i | j << 16, hull_i | hull_j << 16, world)
    uint32_t* work_count = nullptr;  // [0] candidates of this substep, [1] largest pair count of a world, [2] stage_sat_kernel's cursor
    int* overflow = nullptr;         // per world
    int maxV = 0, capG = 0, capP = 0;
    size_t bytes = 0;
    void release() {
        for (void* p : {(void*)ppos, (void*)prot, (void*)fpos, (void*)finvp, (void*)cen, (void*)verts, (void*)gcons,
                        (void*)n_ground, (void*)pairs, (void*)n_pairs, (void*)pcons, (void*)worklist, (void*)work_count,
                        (void*)overflow})
            if (p) cudaFree(p);
        *this = StagedBuffers();
    }
};

struct StagedParams {
    BodyArrays bodies;
    int64_t n_worlds;
    int B, J;
    const int32_t* hull_ids;
    const uint32_t* hull_blob;
    uint32_t hull_blob_bytes;
    const DeviceJoint* joints;
    StagedBuffers S;
    StepConsts sc;
    float friction;
    int body_body;
    int pair_table_len;  // B(B-1)/2 when stage_prepare keeps the pair table in shared memory, else 0
    int compact_verts;   // hulls with different vertex counts: stage_prepare enumerates (body, vertex) through a prefix sum
    int by_level;        // all-cube batches: candidates enumerated in (pair slot, world) order by the SAT and manifold kernels; no work list
    unsigned long long* counters;  // [0] constraints, [1] overflow worlds, [2] tma fallbacks
    unsigned long long* cand_stat;  // non-null on the last substep of a step call: += candidates of this launch (statistics)
};

// ---- stage 1: prepare (warp per world) -----------------------------------------------------------------
// Per-world shared scratch: kPrepSlots float4 per body, slot-major ([slot * B + body]: lanes over bodies read consecutive
// float4, lanes on the same body broadcast).  What the vertex / ground / broadphase loops need of a body is then one or two
// LDS.128 each instead of a dozen scalar loads and a hull-descriptor lookup:
//   0 rot   rotation (v.x, v.y, v.z, s)                 3 prot  rotation at the start of the substep
//   1 fpos  frame position | hull vertex info           4 pfpos past frame position | first flat vertex index (compact walk)
//   2 cen   world centroid | bounding radius            5 finvp inverse frame position | hull id
// (hull vertex info = word offset of the hull's float4 vertex array | n_verts << 20; integers travel as float bit patterns)
constexpr int kPrepSlots = 6;
constexpr int kPrepWordsPerBody = 4 * kPrepSlots;
enum PrepSlot : int { kPsRot = 0, kPsFpos = 1, kPsCen = 2, kPsProt = 3, kPsPfpos = 4, kPsFinvp = 5 };

#ifndef XPBD_PREP_MIN_BLOCKS
#define XPBD_PREP_MIN_BLOCKS 4
#endif
// G = lanes per world: worlds of up to 16 bodies share a warp two at a time (half-warp each; ballots are masked and
// shifted to the half), larger worlds take the whole warp.  Per block, after the hull table and the per-world
// scratch, shared memory holds the table of body pairs (i | j << 16, i < j, lexicographic) when it fits
// (P.pair_table_len > 0): the broadphase then runs over the flattened pair index — ceil(B(B-1)/2 / G) ballot
// rounds per world instead of (B-1) x ceil(B / G).
template <int G>
__global__ void __launch_bounds__(256, XPBD_PREP_MIN_BLOCKS) stage_prepare_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const HullTable T{smem};
    constexpr int NG = 32 / G;
    constexpr unsigned kFull = 0xffffffffu;
    const int B = P.B, maxV = P.S.maxV;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const int grp = lane / G, gl = lane % G;
    const int gshift = grp * G;
    const unsigned gmask = G == 32 ? kFull : (((1u << (G & 31)) - 1u) << gshift);
    const unsigned below_me = (1u << gl) - 1u;
    float4* W = reinterpret_cast<float4*>(smem + P.hull_blob_bytes / 4) + (size_t)(warp * NG + grp) * kPrepSlots * B;
    float4* s_rot = W + kPsRot * B; float4* s_fpos = W + kPsFpos * B; float4* s_cen = W + kPsCen * B;
    float4* s_prot = W + kPsProt * B; float4* s_pfpos = W + kPsPfpos * B; float4* s_finvp = W + kPsFinvp * B;
    auto voff = [&](int b) { return __float_as_int(s_pfpos[b].w); };  // compact walk: vertices of the bodies before b
    uint32_t* pair_table = smem + P.hull_blob_bytes / 4 + (size_t)nwarps * NG * kPrepWordsPerBody * B;
    const int n_table = P.pair_table_len;
    if (n_table > 0) {
        for (int i = threadIdx.x; i + 1 < B; i += blockDim.x) {
            const int row = (i * (2 * B - i - 1)) / 2;  // pairs before row i
            for (int j = i + 1; j < B; ++j) pair_table[row + j - i - 1] = (uint32_t)i | ((uint32_t)j << 16);
        }
        __syncthreads();
    }
    const StepConsts sc = P.sc;

    for (int64_t w0 = ((int64_t)blockIdx.x * nwarps + warp) * NG; w0 < P.n_worlds; w0 += (int64_t)gridDim.x * nwarps * NG) {
        const int64_t world = w0 + grp;
        const bool valid = world < P.n_worlds;  // odd world count: the second half of the last warp idles
        const int64_t g0 = world * B;
        // --- bodies: save past, Rigid::integrate, frames (lanes over bodies)
        for (int b = gl; b < B && valid; b += G) {
            const int64_t g = g0 + b;
            BodyDyn d; BodyConst k;
            load_dyn(P.bodies, g, d);
            load_const(P.bodies, g, k);
            const Fr past = rigid_frame(d.position, d.rotation, k.center_of_mass);
            const float4 prot4 = make_float4(d.rotation.x, d.rotation.y, d.rotation.z, d.rotation.s);
            P.S.ppos[g] = make_float4(d.position.x, d.position.y, d.position.z, 0.f);
            P.S.prot[g] = prot4;
            s_prot[b] = prot4;
            s_pfpos[b] = make_float4(past.p.x, past.p.y, past.p.z, 0.f);
            rigid_integrate(d, k, sc.dt * k.inverse_inertia, sc);
            store_dyn(P.bodies, g, d);
            const Fr frame = rigid_frame(d.position, d.rotation, k.center_of_mass);
            const Fr finv = frame_inverse(frame);
            const int h = P.hull_ids ? P.hull_ids[g] : 0;
            const HullDesc& hd = T.desc(h);
            const V3 c = frame * T.centroid(hd);
            s_rot[b] = make_float4(d.rotation.x, d.rotation.y, d.rotation.z, d.rotation.s);
            s_fpos[b] = make_float4(frame.p.x, frame.p.y, frame.p.z, __int_as_float(hd.vert_off | (hd.n_verts << 20)));
            s_cen[b] = make_float4(c.x, c.y, c.z, hd.radius);
            s_finvp[b] = make_float4(finv.p.x, finv.p.y, finv.p.z, __int_as_float(h));
            P.S.fpos[g] = make_float4(frame.p.x, frame.p.y, frame.p.z, 0.f);
            P.S.finvp[g] = make_float4(finv.p.x, finv.p.y, finv.p.z, 0.f);
            P.S.cen[g] = make_float4(c.x, c.y, c.z, 0.f);
        }
        __syncwarp();
        // --- world-space vertices + collision::ground (lanes over (body, vertex)), body-major / vertex order
        int n_ground = 0;
        {
            float* gc = P.S.gcons + (size_t)world * P.S.capG * kGroundWords;
            // Lanes over (body, vertex) in body-major / vertex order.  Batches whose hulls all have maxV vertices walk the
            // B x maxV grid directly; mixed batches (C4: 4 ... 32 vertices per hull) walk a compact enumeration built from a
            // prefix sum of the vertex counts, so that no lane is spent on the padding (C4: 64 -> ~28 rounds per world).
            int b = 0, v = gl, total = B * maxV;
            if (P.compact_verts) {
                int run = 0;
                for (int b0 = 0; b0 < B; b0 += G) {
                    const int bb = b0 + gl;
                    const int nv = (valid && bb < B) ? (__float_as_int(s_fpos[bb].w) >> 20) : 0;
                    int incl = nv;
#pragma unroll
                    for (int o = 1; o < G; o <<= 1) { const int up = __shfl_up_sync(kFull, incl, o, G); if (gl >= o) incl += up; }
                    if (valid && bb < B) s_pfpos[bb].w = __int_as_float(run + incl - nv);
                    run += __shfl_sync(kFull, incl, G - 1, G);
                }
                total = (int)__reduce_max_sync(kFull, (unsigned)run);  // a warp-uniform trip count (two worlds per warp when G = 16)
                __syncwarp();
                if (!valid || gl >= run) { b = B; }  // nothing (more) for this lane
                else { b = 0; }
                v = gl;  // here: the lane's flat vertex index
            }
            for (int t0 = 0; t0 < total; t0 += G) {
                bool has = false;
                int vv = 0;
                float4 fp = make_float4(0.f, 0.f, 0.f, 0.f);
                if (P.compact_verts) {
                    if (valid && b < B) {
                        while (b + 1 < B && v >= voff(b + 1)) ++b;
                        vv = v - voff(b);
                        fp = s_fpos[b];
                        has = vv < (__float_as_int(fp.w) >> 20);  // false once the flat index has run past the last body
                    }
                } else {
                    while (v >= maxV) { v -= maxV; ++b; }
                    vv = v;
                    has = valid && b < B;  // every hull of the batch has maxV vertices
                    if (has) fp = s_fpos[b];
                }
                bool below = false;
                V3 p = v3(0.f, 0.f, 0.f);
                Fr frame;
                if (has) {
                    XPBD_CHECK(b >= 0 && b < B && vv >= 0 && vv < maxV && vv < (__float_as_int(fp.w) >> 20));
                    XPBD_CHECK(((__float_as_int(fp.w) & 0xfffff) + 4 * vv + 4) * 4 <= (int)P.hull_blob_bytes);
                    const float4 rq = s_rot[b];
                    const float4 lv = reinterpret_cast<const float4*>(smem + (__float_as_int(fp.w) & 0xfffff))[vv];
                    frame.p = v3(fp.x, fp.y, fp.z); frame.q = q4(rq.w, rq.x, rq.y, rq.z);
                    p = frame * v3(lv.x, lv.y, lv.z);
                    P.S.verts[(g0 + b) * maxV + vv] = make_float4(p.x, p.y, p.z, 0.f);
                    below = !(p.z >= 0.0f);
                }
                const unsigned m = (__ballot_sync(kFull, below) & gmask) >> gshift;
                if (below) {
                    const int idx = n_ground + __popc(m & below_me);
                    if (idx < P.S.capG) {
                        const float4 pf = s_pfpos[b], pq = s_prot[b], fi = s_finvp[b];
                        Fr past, finv;
                        past.p = v3(pf.x, pf.y, pf.z); past.q = q4(pq.w, pq.x, pq.y, pq.z);
                        finv.q = conjugate(frame.q); finv.p = v3(fi.x, fi.y, fi.z);
                        const V3 target = v3(p.x, p.y, 0.0f);
                        const V3 correction = target - p;
                        const V3 delta = p - (past * (finv * p));
                        const V3 delta_t = delta - project_on(delta, correction);
                        const V3 c1 = target - P.friction * delta_t;
                        const V3 diff = c1 - p;
                        const float dist = magnitude(diff);
                        const V3 dir = diff * frcp(dist);
                        float4* c = reinterpret_cast<float4*>(gc + (size_t)idx * kGroundWords);
                        c[0] = make_float4(__int_as_float(cons_head(kGround, b, -1)), p.x, p.y, p.z);
                        c[1] = make_float4(dir.x, dir.y, dir.z, dist);
                    }
                }
                n_ground += __popc(m);
                v += G;
            }
        }
        // --- broadphase: bounding spheres about the hull centroids, pairs i<j in lexicographic order
        int n_pairs = 0;
        bool overflow = n_ground > P.S.capG;  // uniform in the lane group: a ground contact found no slot
        if (P.body_body) {
            int4* pairs = P.S.pairs + (size_t)world * P.S.capP;
            auto spheres_touch = [&](float4 a, float4 c) {  // a = sphere of body i, c = sphere of body j (i < j)
                const V3 d = v3(c.x, c.y, c.z) - v3(a.x, a.y, a.z);
                const float rsum = fadd(a.w, c.w);
                return !(magnitude2(d) > fmul(rsum, rsum));
            };
            if (n_table > 0) {
                const int my_table = valid ? n_table : 0;
                for (int p0 = 0; p0 < n_table; p0 += G) {
                    const int p = p0 + gl;
                    bool cand = false;
                    uint32_t ij = 0;
                    if (p < my_table) {
                        ij = pair_table[p];
                        XPBD_CHECK((int)(ij & 0xffff) < B && (int)(ij >> 16) < B && (ij & 0xffff) < (ij >> 16));
                        cand = spheres_touch(s_cen[ij & 0xffff], s_cen[ij >> 16]);
                    }
                    const unsigned m = (__ballot_sync(kFull, cand) & gmask) >> gshift;
                    if (cand) {
                        const int idx = n_pairs + __popc(m & below_me);
                        if (idx < P.S.capP) pairs[idx] = make_int4((int)ij, kNone, -1, 0);
                    }
                    n_pairs += __popc(m);
                }
            } else {
                for (int i = 0; i + 1 < B; ++i) {
                    const float4 si = valid ? s_cen[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                    for (int j0 = i + 1; j0 < B; j0 += G) {
                        const int j = j0 + gl;
                        const bool cand = valid && j < B && spheres_touch(si, s_cen[j < B ? j : 0]);
                        const unsigned m = (__ballot_sync(kFull, cand) & gmask) >> gshift;
                        if (cand) {
                            const int idx = n_pairs + __popc(m & below_me);
                            if (idx < P.S.capP) pairs[idx] = make_int4(i | (j << 16), kNone, -1, 0);
                        }
                        n_pairs += __popc(m);
                    }
                }
            }
            overflow = overflow || n_pairs > P.S.capP;  // uniform in the lane group: a candidate found no slot
            if (overflow) n_pairs = P.S.capP;
            if (P.by_level) {
                // no work list: the SAT and manifold kernels enumerate (pair slot, world) up to the largest pair count of a
                // world, kept in work_count[1] (never reset: an upper bound is all they need, and an atomic is issued only when
                // a world raises it — none in steady state, where the work-list append is one atomic per world on ONE
                // address: 100 k serialised L2 atomics per C3 substep)
                XPBD_CHECK(n_pairs >= 0 && n_pairs <= P.S.capP);
                if (gl == 0 && (unsigned)n_pairs > *reinterpret_cast<volatile unsigned*>(P.S.work_count + 1))
                    atomicMax(P.S.work_count + 1, (unsigned)n_pairs);
            } else {
                // append this world's candidates to the global work list (order irrelevant: results land in the slots)
                unsigned base = 0;
                if (gl == 0 && n_pairs) base = atomicAdd(P.S.work_count, (unsigned)n_pairs);
                base = __shfl_sync(kFull, base, gshift);
                __syncwarp();  // the pair slots written above are read back by other lanes
                for (int p = gl; p < n_pairs; p += G) {
                    const int ij = pairs[p].x;
                    const int hi = __float_as_int(s_finvp[ij & 0xffff].w), hj = __float_as_int(s_finvp[ij >> 16].w);
                    P.S.worklist[base + p] = make_int4((int)(world * P.S.capP + p), ij, hi | (hj << 16), (int)world);
                }
            }
        }
        if (gl == 0 && valid) {
            P.S.n_ground[world] = n_ground < P.S.capG ? n_ground : P.S.capG;
            P.S.n_pairs[world] = n_pairs;
            if (overflow && atomicExch(&P.S.overflow[world], 1) == 0) atomicAdd(P.counters + 1, 1ull);
        }
        __syncwarp();
    }
}

// ---- stage 2: SAT queries (warp per candidate pair) -------------------------------------------------------
#ifndef XPBD_SAT_MIN_BLOCKS
#define XPBD_SAT_MIN_BLOCKS 4
#endif
// (A cp.async double-buffered software pipeline over the work list was measured here and was SLOWER, 1.73 ms vs
// 1.56 ms: the kernel is issue-slot bound with 32 resident warps per SM hiding the per-candidate load chain, and the
// pipeline's extra address arithmetic and spills cost more than the exposed latency it removed.  A lighter look-ahead
// — prefetch.global.L1 of the next candidate's seven input lines, one per lane, issued after the first face query —
// was slower too: 1.57 ms vs 1.41 ms.)
// per-warp shared scratch, in words: two sets of (WA, WB, L) + the queue/table region (two half-warp regions, which
// together also hold the whole-warp layout)
__host__ __device__ inline int sat_words_per_warp(int maxV) { return 6 * kVertWords * maxV + 2 * kEdgeQueueWordsHalf; }
static_assert(2 * kEdgeQueueWordsHalf >= kEdgeQueueWords, "queue region must hold the whole-warp layout");

XD bool is_cube_shape(const HullDesc& h) { return h.n_verts == 8 && h.n_edges == 12 && h.n_faces == 6; }

// One candidate on the whole warp (any hull sizes).
template <bool kCull>
XD void sat_single(const StagedParams& P, const HullTable& T, int4 d, float* WA, float* WB, float* Lc, float* Q, int lane) {
    const int B = P.B, maxV = P.S.maxV;
    const int64_t world = d.w;
    const int64_t gi = world * B + (d.y & 0xffff), gj = world * B + (d.y >> 16);
    const HullDesc& hA = T.desc(d.z & 0xffff);
    const HullDesc& hB = T.desc(d.z >> 16);
    __syncwarp();
    uint32_t rkey = 0;
    for (int v = lane; v < hA.n_verts; v += 32) { const float4 x = P.S.verts[gi * maxV + v]; reinterpret_cast<float4*>(WA)[v] = x; if (kCull) rkey = max(rkey, abs_coord_key(x)); }
    for (int v = lane; v < hB.n_verts; v += 32) { const float4 x = P.S.verts[gj * maxV + v]; reinterpret_cast<float4*>(WB)[v] = x; if (kCull) rkey = max(rkey, abs_coord_key(x)); }
    const float4 ra = P.bodies.dynC[gi], rb = P.bodies.dynC[gj];
    const float4 fa4 = P.S.finvp[gi], fb4 = P.S.finvp[gj], ca4 = P.S.cen[gi];
    Fr finvA, finvB;
    finvA.q = conjugate(q4(ra.w, ra.x, ra.y, ra.z)); finvA.p = v3(fa4.x, fa4.y, fa4.z);
    finvB.q = conjugate(q4(rb.w, rb.x, rb.y, rb.z)); finvB.p = v3(fb4.x, fb4.y, fb4.z);
    __syncwarp();
    float da, db = 0.f, de = 0.f;
    int fa, fb = -1, ep = -1;
    PairFeature pf; pf.feature = kNone; pf.index = -1;
    // Generic instantiations only: all-cube batches run stage_sat_cubes_kernel, and in mixed batches the
    // size-specialised code paths cost more in instruction-cache misses than they save (C4: 1.62 vs 1.77 ms).
    const float4* planesA = reinterpret_cast<const float4*>(T.words + hA.plane_off);
    const float4* planesB = reinterpret_cast<const float4*>(T.words + hB.plane_off);
    face_query_t<0, 0>(planesA, hA.n_faces, finvA, WB, hB.n_verts, Lc, lane, da, fa);
    if (!(da >= 0.0f)) {
        int2 wi = make_int2(0, 0);  // the cull's witnesses: deep vertices of B below A's best face (Lc still holds B in A's space)
        if (kCull) wi = witness_vertices<32>(planesA[fa < 0 ? 0 : fa], Lc, hB.n_verts, lane, 0);
        __syncwarp();
        face_query_t<0, 0>(planesB, hB.n_faces, finvB, WA, hA.n_verts, Lc, lane, db, fb);
        if (!(db >= 0.0f)) {
            const int2* EA = reinterpret_cast<const int2*>(T.words + hA.edge_off);
            const int2* EB = reinterpret_cast<const int2*>(T.words + hB.edge_off);
            if (kCull) {
                const float R = __int_as_float((int)__reduce_max_sync(0xffffffffu, rkey));
                edge_query_culled_t<0, 0, 0, 0>(EA, EB, hA.n_verts, hA.n_edges, hB.n_verts, hB.n_edges, WA, WB, v3(ca4.x, ca4.y, ca4.z), Q,
                                                lane, make_edge_cull(da, db, R, ld3(WB, wi.x), ld3(WB, wi.y)), de, ep);
            } else
            edge_query_t<0, 0, 0, 0>(EA, EB, hA.n_verts, hA.n_edges, hB.n_verts, hB.n_edges, WA, WB, v3(ca4.x, ca4.y, ca4.z), Q,
                                     lane, de, ep);
            if (!(de >= 0.0f)) pf = choose_feature(da, fa, db, fb, de, ep);
        }
    }
    if (lane == 0) { P.S.pairs[d.x].y = pf.feature; P.S.pairs[d.x].z = pf.index; }
}

// Two cube-cube candidates on one warp, 16 lanes each, in lockstep: the 144 edge pairs of a cube pair fill 4.5
// rounds of 32 lanes but exactly 9 rounds of 16, the face queries use 16 of 16 lanes instead of 24 of 32, and the
// per-candidate setup (descriptor, vertices, frames, edge table) is issued once for both.  Each half runs exactly
// the single-candidate arithmetic on its own scratch; the queries are entered by both halves together (a half whose
// candidate is already separated idles through them) so that the warp never serialises the two halves.
template <bool kCull>
XD void sat_dual_cubes(const StagedParams& P, const HullTable& T, int4 d, bool valid, float* W, float* Qregion, int lane) {
    const int B = P.B, maxV = P.S.maxV;
    const int half = lane >> 4, hl = lane & 15, gshift = half * 16;
    float* WA = W + half * 3 * kVertWords * maxV; float* WB = WA + kVertWords * maxV; float* Lc = WB + kVertWords * maxV;
    float* Q = Qregion + half * kEdgeQueueWordsHalf;
    const int64_t world = d.w;
    const int64_t gi = world * B + (d.y & 0xffff), gj = world * B + (d.y >> 16);
    const HullDesc& hA = T.desc(d.z & 0xffff);
    const HullDesc& hB = T.desc(d.z >> 16);
    __syncwarp();
    float4 ra, rb, fa4, fb4, ca4;
    ra = rb = fa4 = fb4 = ca4 = make_float4(0.f, 0.f, 0.f, 1.f);
    uint32_t rkey = 0;
    if (valid) {  // an odd work list leaves the last warp's second half without a candidate: it idles through
        const float4 x = hl < 8 ? P.S.verts[gi * maxV + hl] : P.S.verts[gj * maxV + hl - 8];
        if (hl < 8) reinterpret_cast<float4*>(WA)[hl] = x;
        else reinterpret_cast<float4*>(WB)[hl - 8] = x;
        if (kCull) rkey = abs_coord_key(x);
        ra = P.bodies.dynC[gi]; rb = P.bodies.dynC[gj];
        fa4 = P.S.finvp[gi]; fb4 = P.S.finvp[gj]; ca4 = P.S.cen[gi];
    }
    Fr finvA, finvB;
    finvA.q = conjugate(q4(ra.w, ra.x, ra.y, ra.z)); finvA.p = v3(fa4.x, fa4.y, fa4.z);
    finvB.q = conjugate(q4(rb.w, rb.x, rb.y, rb.z)); finvB.p = v3(fb4.x, fb4.y, fb4.z);
    __syncwarp();
    const float4* planesA = reinterpret_cast<const float4*>(T.words + hA.plane_off);
    const float4* planesB = reinterpret_cast<const float4*>(T.words + hB.plane_off);
    float da, db = 0.f, de = 0.f;
    int fa, fb = -1, ep = -1;
    PairFeature pf; pf.feature = kNone; pf.index = -1;
    face_query_t<6, 8, 16>(planesA, 6, finvA, WB, 8, Lc, hl, da, fa, gshift);
    bool open = valid && !(da >= 0.0f);  // this half's candidate still needs the next query
    if (__any_sync(0xffffffffu, open)) {
        int2 wi = make_int2(0, 0);  // the cull's witnesses: deep vertices of B below A's best face (Lc still holds B in A's space)
        if (kCull) wi = witness_vertices<16>(planesA[fa < 0 ? 0 : fa], Lc, 8, hl, gshift);
        __syncwarp();
        face_query_t<6, 8, 16>(planesB, 6, finvB, WA, 8, Lc, hl, db, fb, gshift);
        open = open && !(db >= 0.0f);
        if (__any_sync(0xffffffffu, open)) {
            const int2* EA = reinterpret_cast<const int2*>(T.words + hA.edge_off);
            const int2* EB = reinterpret_cast<const int2*>(T.words + hB.edge_off);
            if (kCull) {
                const bool hi = gshift != 0;
                const uint32_t r_lo = __reduce_max_sync(0xffffffffu, hi ? 0u : rkey), r_hi = __reduce_max_sync(0xffffffffu, hi ? rkey : 0u);
                const float R = __int_as_float((int)(hi ? r_hi : r_lo));
                edge_query_culled_t<8, 12, 8, 12, 16>(EA, EB, 8, 12, 8, 12, WA, WB, v3(ca4.x, ca4.y, ca4.z), Q, hl,
                                                      make_edge_cull(da, db, R, ld3(WB, wi.x), ld3(WB, wi.y)), de, ep, gshift);
            } else
            edge_query_t<8, 12, 8, 12, 16>(EA, EB, 8, 12, 8, 12, WA, WB, v3(ca4.x, ca4.y, ca4.z), Q, hl, de, ep, gshift);
            if (open && !(de >= 0.0f)) pf = choose_feature(da, fa, db, fb, de, ep);
        }
    }
    if (hl == 0 && valid) { P.S.pairs[d.x].y = pf.feature; P.S.pairs[d.x].z = pf.index; }
}

template <bool kCull>
__global__ void __launch_bounds__(256, XPBD_SAT_MIN_BLOCKS) stage_sat_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const HullTable T{smem};
    const int maxV = P.S.maxV;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float* W = reinterpret_cast<float*>(smem) + P.hull_blob_bytes / 4 + (size_t)warp * sat_words_per_warp(maxV);
    float* Qregion = W + 6 * kVertWords * maxV;
    const unsigned count = *P.S.work_count;
    // Mixed hulls: a candidate costs anything between ~300 instructions (separated by the first face query) and tens of
    // thousands (two 32-vertex hulls in contact), and there are only a few per warp — candidates are handed out one at a
    // time through a cursor (work_count[2], reset with the counter) instead of a fixed stride, so that no warp is left
    // with several heavy ones.  The order in which candidates are taken cannot influence results (own slot each).
    (void)nwarps;
    for (;;) {
        unsigned it = 0;
        if (lane == 0) it = atomicAdd(P.S.work_count + 2, 1u);
        it = __shfl_sync(0xffffffffu, it, 0);
        if (it >= count) break;
        sat_single<kCull>(P, T, P.S.worklist[it], W, W + kVertWords * maxV, W + 2 * kVertWords * maxV, Qregion, lane);
    }
}

// Batches in which every registered hull is cube-shaped (8 vertices, 12 edges, 6 quads): two candidates per warp.
template <bool kCull>
__global__ void __launch_bounds__(256, XPBD_SAT_MIN_BLOCKS) stage_sat_cubes_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const HullTable T{smem};
    const int maxV = P.S.maxV;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    float* W = reinterpret_cast<float*>(smem) + P.hull_blob_bytes / 4 + (size_t)warp * sat_words_per_warp(maxV);
    float* Qregion = W + 6 * kVertWords * maxV;
    const unsigned count = *P.S.work_count;
    for (unsigned it = 2 * (blockIdx.x * nwarps + warp); it < count; it += 2 * gridDim.x * nwarps) {
        const unsigned mine = it + (lane >> 4);
        const bool valid = mine < count;
        int4 d = make_int4(0, 0, 0, 0);
        if (valid) d = P.S.worklist[mine];  // slot, i | j << 16, hull_i | hull_j << 16, world
        sat_dual_cubes<kCull>(P, T, d, valid, W, Qregion, lane);
    }
}

// Manifold of one candidate whose SAT result is in pr (.x = i | j << 16, .y = feature, .z = index): clipping -> <= 4 fresh
// contacts in the pair's slot, count in pairs[slot].w.
template <class Clip>
XD void manifold_candidate(const StagedParams& P, const HullTable& T, uint32_t slot, int4 pr, Clip& clip) {
    const int B = P.B, maxV = P.S.maxV;
    int n_out = 0;
    if (pr.y != kNone) {
        const int64_t world = slot / (uint32_t)P.S.capP;
        const int i = pr.x & 0xffff, j = pr.x >> 16;
        const int64_t gi = world * B + i, gj = world * B + j;
        PairFeature pf; pf.feature = pr.y; pf.index = pr.z;
        const float4 ra = P.bodies.dynC[gi], rb = P.bodies.dynC[gj];
        const float4 pa = P.S.fpos[gi], pb = P.S.fpos[gj];
        Fr FA, FB;
        FA.p = v3(pa.x, pa.y, pa.z); FA.q = q4(ra.w, ra.x, ra.y, ra.z);
        FB.p = v3(pb.x, pb.y, pb.z); FB.q = q4(rb.w, rb.x, rb.y, rb.z);
        const HullDesc& hA = T.desc(P.hull_ids ? P.hull_ids[gi] : 0);
        const HullDesc& hB = T.desc(P.hull_ids ? P.hull_ids[gj] : 0);
        const float* WA = reinterpret_cast<const float*>(P.S.verts + gi * maxV);
        const float* WB = reinterpret_cast<const float*>(P.S.verts + gj * maxV);
        const int incident = feature_incident(pf.feature);
        const int inc = incident == 0 ? i : j, oth = incident == 0 ? j : i;
        Fr finvI, finvO;  // inverse frames of the incident / the other body (their rotations and frame positions are loaded above)
        {
            const float4 fi = P.S.finvp[world * B + inc], fo = P.S.finvp[world * B + oth];
            finvI.q = conjugate(incident == 0 ? FA.q : FB.q); finvI.p = v3(fi.x, fi.y, fi.z);
            finvO.q = conjugate(incident == 0 ? FB.q : FA.q); finvO.p = v3(fo.x, fo.y, fo.z);
        }
        float4* out = P.S.pcons + (size_t)slot * kMaxManifold * 3;
        const float head = __int_as_float(cons_head(kContact, inc, oth));
        // one manifold point -> fresh contact constraint in body-local anchors; dropped when point and target coincide
        auto emit = [&](V3 point, V3 target) {
            const V3 correction = target - point;
            if (magnitude2(correction) == 0.0f) return;  // no direction: dropped
            const V3 a = finvI * point;
            const V3 b = finvO * target;
            const V3 n = normalize(correction);
            out[3 * n_out + 0] = make_float4(head, a.x, a.y, a.z);
            out[3 * n_out + 1] = make_float4(b.x, b.y, b.z, 0.f);
            out[3 * n_out + 2] = make_float4(n.x, n.y, n.z, 0.f);
            ++n_out;
        };
        clip.reset();
        pair_manifold_emit(T, hA, hB, FA, FB, WA, WB, pf, clip, emit);
        if (clip.overflowed()) pair_manifold_local_emit(T, hA, hB, FA, FB, WA, WB, pf, emit);  // never seen; exact by construction
    }
    P.S.pairs[slot].w = n_out;
}

}  // namespace xpbd
#include "xpbd_sat_thread.cuh"  // stage_sat_thread_kernel: thread per candidate (cube-shaped hulls)
namespace xpbd {

// ---- stage 3: manifolds (thread per candidate pair) -> fresh contacts in the pair's slot --------------------
#ifndef XPBD_MANIFOLD_MIN_BLOCKS
#define XPBD_MANIFOLD_MIN_BLOCKS 4
#endif
#ifndef XPBD_MANIFOLD_THREADS
#define XPBD_MANIFOLD_THREADS 128
#endif
constexpr int kManifoldThreads = XPBD_MANIFOLD_THREADS;
template <int CAP>
__global__ void __launch_bounds__(kManifoldThreads, XPBD_MANIFOLD_MIN_BLOCKS) stage_manifold_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint32_t smem[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem, P.hull_blob, P.hull_blob_bytes, &bar, reinterpret_cast<unsigned int*>(P.counters + 2));
    const HullTable T{smem};
    const unsigned count = *P.S.work_count;
    // clip polygons live in shared memory (thread-private arrays would be local memory: ~0.5 KB per candidate of
    // L2/HBM round trips, measured as 3x the stage's algorithmic traffic)
    ClipShared<CAP, kManifoldThreads> clip{reinterpret_cast<float*>(smem) + P.hull_blob_bytes / 4 + threadIdx.x};
    if (P.by_level) {
        // no work list (see stage_prepare_kernel): (world, pair slot) enumerated up to the largest pair count of a world.
        // World-major here — the threads of a warp work on the pairs of the same few worlds and share their vertices and
        // frames through the L1 ((pair slot, world) order, which pays in the SAT kernel, measured 0.251 vs 0.195 ms here).
        const unsigned nw = (unsigned)P.n_worlds, max_pairs = P.S.work_count[1];
        const unsigned long long total = (unsigned long long)max_pairs * nw;
        for (unsigned long long it = blockIdx.x * blockDim.x + threadIdx.x; it < total; it += gridDim.x * blockDim.x) {
            const unsigned w = (unsigned)(it / max_pairs), p = (unsigned)(it - (unsigned long long)w * max_pairs);
            XPBD_CHECK(w < nw && max_pairs <= (unsigned)P.S.capP);
            if (p >= (unsigned)P.S.n_pairs[w]) continue;
            const uint32_t slot = w * (uint32_t)P.S.capP + p;
            manifold_candidate(P, T, slot, P.S.pairs[slot], clip);
        }
        return;
    }
    for (unsigned it = blockIdx.x * blockDim.x + threadIdx.x; it < count; it += gridDim.x * blockDim.x) {
        const uint32_t slot = (uint32_t)P.S.worklist[it].x;
        manifold_candidate(P, T, slot, P.S.pairs[slot], clip);
    }
}

// ---- stage 4: solve (lane pair per world) -----------------------------------------------------------------------
struct GlobalBodies {  // body accessor over the SoA arrays in HBM/L2; g0 = first body of the world
    const BodyArrays& A;
    const float4* gppos;
    const float4* gprot;
    int64_t g0;
    XD V3 pos(int b) const { const float4 a = A.dynA[g0 + b]; return v3(a.x, a.y, a.z); }
    XD Q4 rot(int b) const { const float4 c = A.dynC[g0 + b]; return q4(c.w, c.x, c.y, c.z); }
    XD V3 com(int b) const { const float4 k = __ldg(&A.k0[g0 + b]); return v3(k.y, k.z, k.w); }
    XD float invm(int b) const { return __ldg(&A.k0[g0 + b]).x; }
    XD M3 inertia(int b) const {
        const float4 p = __ldg(&A.k1[g0 + b]), q = __ldg(&A.k2[g0 + b]), r = __ldg(&A.k3[g0 + b]);
        M3 m;
        m.c0 = v3(p.x, p.y, p.z); m.c1 = v3(p.w, q.x, q.y); m.c2 = v3(q.z, q.w, r.x);
        return m;
    }
    XD V3 vel(int b) const { const float4 a = A.dynA[g0 + b], c = A.dynB[g0 + b]; return v3(a.w, c.x, c.y); }
    XD V3 ang(int b) const { const float4 c = A.dynB[g0 + b]; return v3(c.z, c.w, A.dynD[g0 + b]); }
    XD V3 ppos(int b) const { const float4 a = gppos[g0 + b]; return v3(a.x, a.y, a.z); }
    XD Q4 prot(int b) const { const float4 c = gprot[g0 + b]; return q4(c.w, c.x, c.y, c.z); }
    // velocity.x shares dynA with the position; it is stale during the sweep and rewritten by set_vel after it
    XD void set_pose(int b, V3 p, Q4 q) {
        float4* a = &A.dynA[g0 + b];
        *a = make_float4(p.x, p.y, p.z, a->w);
        A.dynC[g0 + b] = make_float4(q.x, q.y, q.z, q.s);
    }
    XD void set_vel(int b, V3 v, V3 w) {
        float4* a = &A.dynA[g0 + b];
        *a = make_float4(a->x, a->y, a->z, v.x);
        A.dynB[g0 + b] = make_float4(v.y, v.z, w.x, w.y);
        A.dynD[g0 + b] = w.z;
    }
};

// One body held in registers by a solver lane.  A manifold applies up to four contacts to the same two bodies in
// a row, and each contact reads the body ~3 times: caching turns ~12 dependent L2 round trips per contact into one
// load when the lane's body changes.  Protocol inside a lane pair, per constraint (`me` = the body this lane
// needs, -1 = none):   release(me);  __syncwarp(pm);  acquire(me);
// A lane writes its dirty body back BEFORE the barrier whenever it moves on to another body, so the other lane
// — which can only need that body after the owner left it (a constraint never has body a == body b) — always
// reads current state.  kVel = false: sweep (pose is written; velocity.x rides along in dynA);
// kVel = true: velocity pass (pose is read-only, velocities are written).
template <bool kVel>
struct CachedBody {
    const BodyArrays& A;
    int64_t g0;
    int cur = -1;
    bool dirty = false;
    float4 a4, b4, c4;  // dynA, dynB, dynC
    float d1;           // dynD
    float invm_; V3 com_; M3 I_;
    XD void release(int me) {
        if (cur >= 0 && cur != me) {
            if (dirty) {
                const int64_t g = g0 + cur;
                A.dynA[g] = a4;
                if (kVel) { A.dynB[g] = b4; A.dynD[g] = d1; } else { A.dynC[g] = c4; }
            }
            cur = -1; dirty = false;
        }
    }
    XD void acquire(int me) {
        if (me >= 0 && cur != me) {
            const int64_t g = g0 + me;
            a4 = A.dynA[g]; c4 = A.dynC[g];
            if (kVel) { b4 = A.dynB[g]; d1 = A.dynD[g]; }
            const float4 k0 = __ldg(&A.k0[g]), p = __ldg(&A.k1[g]), q = __ldg(&A.k2[g]), r = __ldg(&A.k3[g]);
            invm_ = k0.x; com_ = v3(k0.y, k0.z, k0.w);
            I_.c0 = v3(p.x, p.y, p.z); I_.c1 = v3(p.w, q.x, q.y); I_.c2 = v3(q.z, q.w, r.x);
            cur = me; dirty = false;
        }
    }
    XD void flush() { release(-2); }
    // accessor interface of xpbd_solve.cuh (always called with b == cur)
    XD V3 pos(int) const { return v3(a4.x, a4.y, a4.z); }
    XD Q4 rot(int) const { return q4(c4.w, c4.x, c4.y, c4.z); }
    XD V3 com(int) const { return com_; }
    XD float invm(int) const { return invm_; }
    XD M3 inertia(int) const { return I_; }
    XD V3 vel(int) const { return v3(a4.w, b4.x, b4.y); }
    XD V3 ang(int) const { return v3(b4.z, b4.w, d1); }
    XD void set_pose(int, V3 p, Q4 q) { a4 = make_float4(p.x, p.y, p.z, a4.w); c4 = make_float4(q.x, q.y, q.z, q.s); dirty = true; }
    XD void set_vel(int, V3 v, V3 w) { a4.w = v.x; b4 = make_float4(v.y, v.z, w.x, w.y); d1 = w.z; dirty = true; }
};

// (Measured and rejected: prefetch.global.L1 of the next pair's contact records and bodies one pair ahead made this
// stage slower, 0.325 vs 0.300 ms on C3 — it is bound by the L2 working set of the worlds in flight, not by the
// exposed latency of the dependent loads.)
// Which body of a contact this lane holds: 0 = the head's body a, 1 = body b.  When the two indices differ in
// parity the lane keeps the body whose parity equals its side, so that in a chain of contacts (0,1), (1,2), (2,3) ...
// — a stack — every body stays cached in ONE lane across consecutive manifolds instead of being written back by one
// lane and re-read by the other.  The arithmetic per body does not depend on which lane performs it.
XD int contact_role(int side, int head) {
    const int a = head_a(head), b = head_b(head);
    return ((a ^ b) & 1) ? (side ^ (a & 1)) : side;
}

#ifndef XPBD_SOLVE_MIN_BLOCKS
#define XPBD_SOLVE_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(128, XPBD_SOLVE_MIN_BLOCKS) stage_solve_kernel(StagedParams P) {
    const int lane = threadIdx.x & 31;
    const int side = lane & 1;
    const unsigned pm = 3u << (lane & ~1);
    unsigned applied = 0, cands = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // consumed by sat + manifold; the next substep appends again
        *P.S.work_count = 0;
        P.S.work_count[2] = 0;  // stage_sat_kernel's cursor
    }
    const StepConsts sc = P.sc;
    // grid-stride over worlds: the host may launch fewer blocks than worlds / 64 so that the worlds in flight
    // (x ~2.6 KB each) stay inside the L2 instead of cycling through HBM between the sweep, derive and velocity phases
    for (int64_t world = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 1; world < P.n_worlds;
         world += ((int64_t)gridDim.x * blockDim.x) >> 1) {
        const int n_ground = P.S.n_ground[world], n_pairs = P.S.n_pairs[world];
        cands += (unsigned)n_pairs;
        int4* pairs = P.S.pairs + (size_t)world * P.S.capP;
        float4* pc = P.S.pcons + (size_t)world * P.S.capP * kMaxManifold * 3;
        {
            CachedBody<false> C{P.bodies, world * P.B};
            // --- sweep: ground constraints (body-major, hull-vertex order) ...
            const float4* gc = reinterpret_cast<const float4*>(P.S.gcons + (size_t)world * P.S.capG * kGroundWords);
            for (int k = 0; k < n_ground; ++k) {
                const float4 c0 = gc[2 * k], c1 = gc[2 * k + 1];
                const int a = head_a(__float_as_int(c0.x));
                const int owner = a & 1;  // same lane that holds this body in a neighbouring contact
                const int me = side == owner ? a : -1;
                C.release(me); __syncwarp(pm); C.acquire(me);
                if (side == owner) solve_ground(C, me, v3(c0.y, c0.z, c0.w), v3(c1.x, c1.y, c1.z), c1.w, sc);
            }
            applied += n_ground;
            // --- ... then the manifolds of the candidate pairs i<j in lexicographic order ...
            for (int p = 0; p < n_pairs; ++p) {
                const int cnt = pairs[p].w;
                int active_mask = 0;  // contacts of this manifold that were active, for the velocity pass
                for (int k = 0; k < cnt; ++k) {
                    const float4* c = pc + ((size_t)p * kMaxManifold + k) * 3;
                    const float4 c0 = c[0], c1 = c[1], c2 = c[2];
                    const int head = __float_as_int(c0.x);
                    const int role = contact_role(side, head);
                    const int me = role ? head_b(head) : head_a(head);
                    C.release(me); __syncwarp(pm); C.acquire(me);
                    const bool active = solve_contact(C, role, pm, me, role ? v3(c1.x, c1.y, c1.z) : v3(c0.y, c0.z, c0.w),
                                                      v3(c2.x, c2.y, c2.z), sc);
                    active_mask |= (int)active << k;
                }
                // one 4-byte store per pair (stage_manifold rewrites the word every substep) instead of a flag in every
                // contact record: those were four scattered 32-byte sector writes per pair
                if (side == 0 && active_mask) pairs[p].w = cnt | (active_mask << 8);
                applied += cnt;
            }
            // --- ... then the joints in index order
            for (int jn = 0; jn < P.J; ++jn) {
                const DeviceJoint q = P.joints[world * P.J + jn];
                if (q.body_a < 0) continue;
                const int me = side ? q.body_b : q.body_a;
                C.release(me); __syncwarp(pm); C.acquire(me);
                solve_joint(C, side, pm, me, q.body_b >= 0, side ? v3(q.bx, q.by, q.bz) : v3(q.ax, q.ay, q.az), q.distance, sc);
                ++applied;
            }
            C.flush();
        }
        __syncwarp(pm);
        // --- Rigid::derive, bodies split between the two lanes
        {
            GlobalBodies A{P.bodies, P.S.ppos, P.S.prot, world * P.B};
            for (int b = side; b < P.B; b += 2) derive_body(A, b, sc.dt);
        }
        // --- velocity pass over the contacts that were active in the sweep
        {
            CachedBody<true> C{P.bodies, world * P.B};
            for (int p = 0; p < n_pairs; ++p) {
                const int word = pairs[p].w, cnt = word & 0xff, active_mask = word >> 8;
                for (int k = 0; k < cnt; ++k) {
                    if (!((active_mask >> k) & 1)) continue;
                    const float4* c = pc + ((size_t)p * kMaxManifold + k) * 3;
                    const float4 c0 = c[0], c1 = c[1], c2 = c[2];
                    const int head = __float_as_int(c0.x);
                    const int role = contact_role(side, head);
                    const int me = role ? head_b(head) : head_a(head);
                    C.release(me); __syncwarp(pm); C.acquire(me);
                    velocity_contact(C, role, pm, me, role ? v3(c1.x, c1.y, c1.z) : v3(c0.y, c0.z, c0.w), v3(c2.x, c2.y, c2.z),
                                     P.friction);
                }
            }
            C.flush();
        }
        __syncwarp(pm);
    }
    if (side) { applied = 0; cands = 0; }
    // one atomic per warp (lanes leave the loop at different times: reconverge first)
    __syncwarp();
    for (int o = 16; o > 0; o >>= 1) { applied += __shfl_down_sync(0xffffffffu, applied, o); cands += __shfl_down_sync(0xffffffffu, cands, o); }
    if (lane == 0 && applied) atomicAdd(P.counters, (unsigned long long)applied);
    if (lane == 0 && cands && P.cand_stat) atomicAdd(P.cand_stat, (unsigned long long)cands);  // statistics: candidate pairs of this substep
}

// ---- stage 4, wide schedule: LW lanes (LW / 2 lane pairs) per world ---------------------------------------------------
// The lane-pair kernel above walks a world's constraints one after the other; with 64 bodies, 32 joints and a few dozen
// contacts per world (C4) that is one long dependent chain per lane pair and 1 000 warps for the whole device.  But
// Gauss-Seidel only orders constraints that SHARE A BODY: constraint k reads and writes its own one or two bodies, so it
// can run as soon as the last earlier constraint on either body is done, and any schedule that keeps that order per body
// produces bit-identical state.  Per world and substep:
//   ground contacts : every body's contacts are applied, in list order, by the lane that owns the body (body % LW);
//   manifolds       : level[p] = 1 + max(level of the last earlier manifold on body a, on body b), computed serially by
//                     one lane from the pair list (a few instructions per pair); the manifolds of one level touch disjoint
//                     bodies and are dealt to the lane pairs, level after level;
//   joints          : the same with their own levels (C4's 32 joints connect 64 different bodies: one level);
//   derive          : lanes over bodies;
//   velocity pass   : the manifolds again, same levels.
// Between levels every lane writes its cached bodies back and the world's lanes meet (__syncwarp).  Per-contact
// arithmetic is the lane-pair code of xpbd_solve.cuh, unchanged.
constexpr int kSolveWideBlock = 128;
#ifndef XPBD_SOLVE_WIDE_LANES
#define XPBD_SOLVE_WIDE_LANES 16  // C4 with 3 world ranges: 16 lanes 1.31e9, 8 lanes 1.29e9 body-substeps/s
#endif
constexpr int kSolveWideDefaultLanes = XPBD_SOLVE_WIDE_LANES;
#ifndef XPBD_SOLVE_WIDE_MIN_BLOCKS
#define XPBD_SOLVE_WIDE_MIN_BLOCKS 6  // 80 registers: no spills (64 spill a few words)
#endif
__host__ __device__ inline size_t solve_wide_group_words16(int B, int capP, int J) { return (size_t)((2 * B + capP + J + 1) & ~1); }

XD int nth_set_bit(unsigned m, int r) { return (int)__fns(m, 0, r + 1); }  // position of the r-th (0-based) set bit

template <int LW>
__global__ void __launch_bounds__(kSolveWideBlock, XPBD_SOLVE_WIDE_MIN_BLOCKS) stage_solve_wide_kernel(StagedParams P) {
    extern __shared__ __align__(16) uint16_t wide_smem[];
    constexpr int NG = 32 / LW;  // worlds per warp
    constexpr int NP = LW / 2;   // lane pairs per world
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int grp = lane / LW, gl = lane % LW, gshift = grp * LW;
    const unsigned gm = LW == 32 ? 0xffffffffu : (((1u << (LW & 31)) - 1u) << gshift);
    const int side = lane & 1, q = gl >> 1;
    const unsigned pm = 3u << (lane & ~1);
    uint16_t* last_p = wide_smem + (size_t)(warp * NG + grp) * solve_wide_group_words16(P.B, P.S.capP, P.J);
    uint16_t* last_j = last_p + P.B;
    uint16_t* lvp = last_j + P.B;    // level of pair slot p (0: no contacts)
    uint16_t* lvj = lvp + P.S.capP;  // level of joint j (0: inactive)
    unsigned applied = 0, cands = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) {  // consumed by sat + manifold; the next substep appends again
        *P.S.work_count = 0;
        P.S.work_count[2] = 0;
    }
    const StepConsts sc = P.sc;
    for (int64_t w0 = ((int64_t)blockIdx.x * nwarps + warp) * NG; w0 < P.n_worlds; w0 += (int64_t)gridDim.x * nwarps * NG) {
        const int64_t world = w0 + grp;
        const bool valid = world < P.n_worlds;  // the last warp's spare groups run the loops below with zero trip counts
        const int64_t g0 = world * P.B;
        const int nB = valid ? P.B : 0, nJ = valid ? P.J : 0;
        const int n_ground = valid ? P.S.n_ground[world] : 0, n_pairs = valid ? P.S.n_pairs[world] : 0;
        int4* pairs = P.S.pairs + (size_t)world * P.S.capP;
        float4* pc = P.S.pcons + (size_t)world * P.S.capP * kMaxManifold * 3;
        const DeviceJoint* joints = P.joints + world * P.J;
        // --- levels (lane 0: manifolds, lane 1: joints)
        for (int b = gl; b < nB; b += LW) { last_p[b] = 0; last_j[b] = 0; }
        __syncwarp(gm);
        int max_lv = 0;
        if (gl == 0) {
            unsigned n = (unsigned)n_ground;
            cands += (unsigned)n_pairs;
            for (int p = 0; p < n_pairs; ++p) {
                const int4 pr = pairs[p];
                int lv = 0;
                if (pr.w) {
                    const int a = pr.x & 0xffff, b = pr.x >> 16;
                    XPBD_CHECK(a < P.B && b < P.B && a != b && pr.w > 0 && pr.w <= kMaxManifold && p < P.S.capP);
                    lv = 1 + max((int)last_p[a], (int)last_p[b]);
                    last_p[a] = (uint16_t)lv; last_p[b] = (uint16_t)lv;
                    max_lv = max(max_lv, lv);
                    n += (unsigned)pr.w;
                }
                lvp[p] = (uint16_t)lv;
            }
            applied += n;
        } else if (gl == 1) {
            unsigned n = 0;
            for (int j = 0; j < nJ; ++j) {
                const int a = joints[j].body_a, b = joints[j].body_b;
                int lv = 0;
                if (a >= 0) {
                    XPBD_CHECK(a < P.B && b < P.B);
                    lv = 1 + max((int)last_j[a], b >= 0 ? (int)last_j[b] : 0);
                    last_j[a] = (uint16_t)lv;
                    if (b >= 0) last_j[b] = (uint16_t)lv;
                    max_lv = max(max_lv, lv);
                    ++n;
                }
                lvj[j] = (uint16_t)lv;
            }
            applied += n;
        }
        __syncwarp(gm);
        const int max_lp = __shfl_sync(gm, max_lv, gshift), max_lj = __shfl_sync(gm, max_lv, gshift + 1);
        // --- sweep: ground constraints, each body's by its owner lane in list order ...
        {
            CachedBody<false> C{P.bodies, g0};
            const float4* gc = reinterpret_cast<const float4*>(P.S.gcons + (size_t)world * P.S.capG * kGroundWords);
            for (int k = 0; k < n_ground; ++k) {
                const float4 c0 = gc[2 * k];
                const int a = head_a(__float_as_int(c0.x));
                XPBD_CHECK(a >= 0 && a < P.B && k < P.S.capG);
                if (a % LW == gl) {
                    const float4 c1 = gc[2 * k + 1];
                    C.release(a); C.acquire(a);
                    solve_ground(C, a, v3(c0.y, c0.z, c0.w), v3(c1.x, c1.y, c1.z), c1.w, sc);
                }
            }
            C.flush();
        }
        __syncwarp(gm);
        // --- ... then the manifolds, level by level ...
        for (int lv = 1; lv <= max_lp; ++lv) {
            CachedBody<false> C{P.bodies, g0};
            for (int p0 = 0; p0 < n_pairs; p0 += LW) {
                const bool mine = p0 + gl < n_pairs && lvp[p0 + gl] == lv;
                const unsigned m = (__ballot_sync(gm, mine) & gm) >> gshift;
                const int nm = __popc(m);
                for (int r = q; r < nm; r += NP) {
                    const int p = p0 + nth_set_bit(m, r);
                    XPBD_CHECK(p >= p0 && p < n_pairs && lvp[p] == lv);
                    const int cnt = pairs[p].w;
                    int active_mask = 0;
                    for (int k = 0; k < cnt; ++k) {
                        const float4* c = pc + ((size_t)p * kMaxManifold + k) * 3;
                        const float4 c0 = c[0], c1 = c[1], c2 = c[2];
                        const int head = __float_as_int(c0.x);
                        const int role = contact_role(side, head);
                        const int me = role ? head_b(head) : head_a(head);
                        C.release(me); __syncwarp(pm); C.acquire(me);
                        const bool active = solve_contact(C, role, pm, me, role ? v3(c1.x, c1.y, c1.z) : v3(c0.y, c0.z, c0.w),
                                                          v3(c2.x, c2.y, c2.z), sc);
                        active_mask |= (int)active << k;
                    }
                    if (side == 0 && active_mask) pairs[p].w = cnt | (active_mask << 8);
                }
            }
            C.flush();
            __syncwarp(gm);
        }
        // --- ... then the joints, level by level
        for (int lv = 1; lv <= max_lj; ++lv) {
            CachedBody<false> C{P.bodies, g0};
            for (int j0 = 0; j0 < nJ; j0 += LW) {
                const bool mine = j0 + gl < nJ && lvj[j0 + gl] == lv;
                const unsigned m = (__ballot_sync(gm, mine) & gm) >> gshift;
                const int nm = __popc(m);
                for (int r = q; r < nm; r += NP) {
                    XPBD_CHECK(j0 + nth_set_bit(m, r) < nJ && lvj[j0 + nth_set_bit(m, r)] == lv);
                    const DeviceJoint jt = joints[j0 + nth_set_bit(m, r)];
                    const int me = side ? jt.body_b : jt.body_a;
                    C.release(me); __syncwarp(pm); C.acquire(me);
                    solve_joint(C, side, pm, me, jt.body_b >= 0, side ? v3(jt.bx, jt.by, jt.bz) : v3(jt.ax, jt.ay, jt.az), jt.distance, sc);
                }
            }
            C.flush();
            __syncwarp(gm);
        }
        // --- Rigid::derive, lanes over bodies
        {
            GlobalBodies A{P.bodies, P.S.ppos, P.S.prot, g0};
            for (int b = gl; b < nB; b += LW) derive_body(A, b, sc.dt);
        }
        __syncwarp(gm);
        // --- velocity pass over the contacts that were active in the sweep, same levels
        for (int lv = 1; lv <= max_lp; ++lv) {
            CachedBody<true> C{P.bodies, g0};
            for (int p0 = 0; p0 < n_pairs; p0 += LW) {
                const bool mine = p0 + gl < n_pairs && lvp[p0 + gl] == lv;
                const unsigned m = (__ballot_sync(gm, mine) & gm) >> gshift;
                const int nm = __popc(m);
                for (int r = q; r < nm; r += NP) {
                    const int p = p0 + nth_set_bit(m, r);
                    const int word = pairs[p].w, cnt = word & 0xff, active_mask = word >> 8;
                    for (int k = 0; k < cnt; ++k) {
                        if (!((active_mask >> k) & 1)) continue;
                        const float4* c = pc + ((size_t)p * kMaxManifold + k) * 3;
                        const float4 c0 = c[0], c1 = c[1], c2 = c[2];
                        const int head = __float_as_int(c0.x);
                        const int role = contact_role(side, head);
                        const int me = role ? head_b(head) : head_a(head);
                        C.release(me); __syncwarp(pm); C.acquire(me);
                        velocity_contact(C, role, pm, me, role ? v3(c1.x, c1.y, c1.z) : v3(c0.y, c0.z, c0.w), v3(c2.x, c2.y, c2.z),
                                         P.friction);
                    }
                }
            }
            C.flush();
            __syncwarp(gm);
        }
    }
    // one atomic per warp (lanes leave the loop at different times: reconverge first)
    __syncwarp();
    for (int o = 16; o > 0; o >>= 1) { applied += __shfl_down_sync(0xffffffffu, applied, o); cands += __shfl_down_sync(0xffffffffu, cands, o); }
    if (lane == 0 && applied) atomicAdd(P.counters, (unsigned long long)applied);
    if (lane == 0 && cands && P.cand_stat) atomicAdd(P.cand_stat, (unsigned long long)cands);
}

// ---- host side -------------------------------------------------------------------------------------------------
constexpr int kGraphMaxSubsteps = 256;  // longer chains (many frames per call) launch directly
struct RangeGraphKey {    // everything the captured chain of one world range depends on (compared bytewise: zero-filled first)
    StagedParams params;  // of the range, cand_stat = nullptr
    int total;            // substeps in the chain
    int grid[4];
    const void* fn[4];    // prepare, sat, manifold, solve kernels
};
struct RangeGraph {
    cudaGraphExec_t exec = nullptr;
    RangeGraphKey key;
    void release() { if (exec) cudaGraphExecDestroy(exec); exec = nullptr; }
};

struct StagedState {
    StagedBuffers buf;
    RangeGraph graphs[8];
    int64_t n_worlds = 0;
    size_t n_hulls = 0;  // kernel choices depend on the registered hulls
    int B = 0;
    int grid_prepare = 0, grid_sat = 0, grid_manifold = 0;
    size_t smem_prepare = 0, smem_sat = 0, smem_manifold = 0;
    int sat_threads = 256;        // block size of sat_fn
    int sat_cands_per_block = 8;  // candidates one block of sat_fn takes per grid-stride round
    int warps_prepare = 8;
    int clip_cap = kMaxClip;
    void (*manifold_fn)(StagedParams) = nullptr;
    void (*prepare_fn)(StagedParams) = nullptr;
    void (*sat_fn)(StagedParams) = nullptr;
    int prepare_lanes = 32, pair_table_len = 0;
    int n_sm = 148;
    int solve_blocks_per_sm = 0;  // grid-stride over worlds: resident blocks per SM (occupancy), XPBD_SOLVE_BLOCKS_PER_SM overrides
    int solve_lanes_default = 2;  // lanes per world of the solve stage: 2 = stage_solve_kernel, 4 ... 32 = stage_solve_wide_kernel
    size_t smem_solve_wide = 0;   // per block of kSolveWideBlock threads at the default lane count
    int occ_prepare = 1, occ_sat = 1, occ_manifold = 1;
    // Sub-batches: the worlds are split into n_sub contiguous ranges, each stepped by its own chain of launches on its
    // own stream (forked from / joined into the batch stream inside one step call).  Worlds are independent, so the
    // result is the same; the issue-bound SAT kernel of one range then shares the SMs with the latency-bound solve /
    // manifold / prepare kernels of the others (C3: 1.80 -> 1.55 ms per substep with 4 ranges).
    int n_sub = 1;
    std::vector<cudaStream_t> sub_streams;   // n_sub of them when n_sub > 1 (the batch stream only forks and joins)
    std::vector<cudaEvent_t> sub_done;       // n_sub
    cudaEvent_t ev_fork = nullptr;
    unsigned long long* d_range_counters = nullptr;  // kMaxSubBatches x 4: [0] constraints, [1] overflow worlds, [2] tma fallbacks, [3] candidates
    unsigned long long* h_range_counters = nullptr;  // pinned copy, written at the end of every range's chain
    void release_streams() {
        for (cudaStream_t s : sub_streams) cudaStreamDestroy(s);
        for (cudaEvent_t e : sub_done) cudaEventDestroy(e);
        if (ev_fork) cudaEventDestroy(ev_fork);
        sub_streams.clear(); sub_done.clear(); ev_fork = nullptr;
    }
};

constexpr int kWorkCountStride = 32;  // uint32 words between the work counters of two sub-batches (own 128-byte line each)
constexpr int kMaxSubBatches = 8;

static void destroy_staged(StagedState* st) {
    for (RangeGraph& g : st->graphs) g.release();
    st->buf.release();
    st->release_streams();
    if (st->d_range_counters) cudaFree(st->d_range_counters);
    if (st->h_range_counters) cudaFreeHost(st->h_range_counters);
    delete st;
}
void release_world_scratch(xpbd_batch* b) {
    if (!b->world_scratch) return;
    destroy_staged(static_cast<StagedState*>(b->world_scratch));
    b->world_scratch = nullptr;
}

template <class T>
static cudaError_t dev_alloc(T*& p, size_t n, size_t& total) {
    const size_t bytes = n * sizeof(T);
    total += bytes;
    return cudaMalloc(reinterpret_cast<void**>(&p), bytes > 16 ? bytes : 16);
}

// Every registered hull is cube-shaped (8 vertices, 12 edges, 6 quads): the batch can run the cube schedules of the SAT stage.
static bool all_cube_hulls(const xpbd_batch* b) {
    bool all_cubes = !b->hulls.empty();
    for (const HostHull& h : b->hulls) {
        bool quads = h.face_offsets.size() == 7;
        for (size_t f = 0; quads && f + 1 < h.face_offsets.size(); ++f) quads = h.face_offsets[f + 1] - h.face_offsets[f] == 4;
        all_cubes = all_cubes && h.vertices.size() == 8 * 3 && h.edges.size() == 12 * 2 && quads;
    }
    return all_cubes;
}

static int build_staged(xpbd_batch* b, StagedState* st);
// The state is published in b->world_scratch only once it is complete: a failed allocation / capacity check leaves no
// half-built cache behind (the next step call starts over and reports the original error again).
static int ensure_staged(xpbd_batch* b, StagedState*& out) {
    StagedState* st = static_cast<StagedState*>(b->world_scratch);
    if (st && st->buf.maxV == b->dev_hulls.max_verts && st->n_hulls == b->hulls.size()) { out = st; return XPBD_OK; }
    release_world_scratch(b);
    st = new StagedState();
    if (int rc = build_staged(b, st)) { destroy_staged(st); return rc; }
    b->world_scratch = st;
    out = st;
    return XPBD_OK;
}
static int build_staged(xpbd_batch* b, StagedState* st) {
    const int B = b->desc.bodies_per_world;
    const DeviceHulls& H = b->dev_hulls;
    st->n_worlds = b->desc.n_worlds;
    st->B = B;
    st->n_hulls = b->hulls.size();
    StagedBuffers& S = st->buf;
    S.maxV = H.max_verts;
    S.capG = 8 * B < B * S.maxV ? 8 * B : B * S.maxV;
    S.capP = B * (B - 1) / 2 < 4 * B ? B * (B - 1) / 2 + 1 : 4 * B;
    const size_t nb = (size_t)b->n_bodies, nw = (size_t)b->desc.n_worlds;
    if (nw * (size_t)S.capP >= 0x7fffffffull) return fail(XPBD_ERR_CAPACITY, "n_worlds x pair capacity exceeds 2^31");
    XPBD_CUDA_TRY(dev_alloc(S.ppos, nb, S.bytes)); XPBD_CUDA_TRY(dev_alloc(S.prot, nb, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.fpos, nb, S.bytes)); XPBD_CUDA_TRY(dev_alloc(S.finvp, nb, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.cen, nb, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.verts, nb * S.maxV, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.gcons, nw * S.capG * kGroundWords, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.n_ground, nw, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.pairs, nw * S.capP, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.n_pairs, nw, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.pcons, nw * S.capP * kMaxManifold * 3, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.worklist, nw * S.capP, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.work_count, kWorkCountStride * kMaxSubBatches, S.bytes));
    XPBD_CUDA_TRY(dev_alloc(S.overflow, nw, S.bytes));
    XPBD_CUDA_TRY(cudaMemsetAsync(S.work_count, 0, kWorkCountStride * kMaxSubBatches * sizeof(uint32_t), b->stream));
    XPBD_CUDA_TRY(cudaMemsetAsync(S.n_pairs, 0, nw * sizeof(int), b->stream));

    int dev = 0, n_sm = 0, smem_max = 0;
    XPBD_CUDA_TRY(cudaGetDevice(&dev));
    XPBD_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
    XPBD_CUDA_TRY(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    st->n_sm = n_sm;
    {
        int occ_solve = 1;
        XPBD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_solve, stage_solve_kernel, 128, 0));
        st->solve_blocks_per_sm = occ_solve < 6 ? occ_solve : 6;  // measured on C3: 6 resident blocks beat 7, 8 and an unbounded grid (L2 working set)
        if (const char* e = std::getenv("XPBD_SOLVE_BLOCKS_PER_SM")) st->solve_blocks_per_sm = std::atoi(e);
        // Worlds of more than 16 bodies get the wide schedule (several lane pairs per world, constraints on disjoint bodies
        // in parallel); small worlds keep one lane pair each (C3's stack is one dependent chain: nothing to spread).
        st->solve_lanes_default = B > 16 ? kSolveWideDefaultLanes : 2;
        // few large worlds: more lanes per world until the device has about one resident thread per slot (C4: 16 k worlds -> 16)
        while (B > 16 && st->solve_lanes_default < 32 && st->n_worlds * st->solve_lanes_default < (int64_t)n_sm * XPBD_SOLVE_WIDE_MIN_BLOCKS * kSolveWideBlock)
            st->solve_lanes_default *= 2;
    }
    // prepare: warps per block limited by the per-warp body scratch
    st->prepare_lanes = B <= 16 ? 16 : 32;  // lanes per world
    st->prepare_fn = B <= 16 ? stage_prepare_kernel<16> : stage_prepare_kernel<32>;
    const size_t prep_per_warp = (size_t)(32 / st->prepare_lanes) * kPrepWordsPerBody * B * 4;
    const size_t n_pairs_all = (size_t)B * (B - 1) / 2;
    st->pair_table_len = n_pairs_all <= 4096 ? (int)n_pairs_all : 0;  // <= 16 KB of shared memory
    const size_t table_bytes = (size_t)st->pair_table_len * 4;
    int wp = 8;
    while (wp > 1 && H.bytes + wp * prep_per_warp + table_bytes > (size_t)smem_max / 2) wp >>= 1;
    st->warps_prepare = wp;
    st->smem_prepare = H.bytes + wp * prep_per_warp + table_bytes;
    st->smem_sat = H.bytes + 8 * (size_t)sat_words_per_warp(S.maxV) * 4;
    st->clip_cap = 2 * H.max_face_verts <= 8 ? 8 : kMaxClip;  // entries per clip buffer in stage_manifold's scratch
    st->smem_manifold = H.bytes + (size_t)kManifoldThreads * 7 * st->clip_cap * 4;
    if (st->smem_prepare > (size_t)smem_max || st->smem_sat > (size_t)smem_max || st->smem_manifold > (size_t)smem_max)
        return fail(XPBD_ERR_CAPACITY, "hull table / world too large for shared memory (%zu, %zu bytes)", st->smem_prepare, st->smem_sat);
    XPBD_CUDA_TRY(cudaFuncSetAttribute(st->prepare_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)st->smem_prepare));
    // XPBD_SAT_CULL=0: the edge query without the witness cull (every edge pair through the reference's arithmetic) — A/B runs
    bool cull = true;
    if (const char* e = std::getenv("XPBD_SAT_CULL")) cull = e[0] != '0';
    st->sat_fn = cull ? stage_sat_kernel<true> : stage_sat_kernel<false>;
    {
        const bool all_cubes = all_cube_hulls(b);
        // Schedules of the SAT stage for all-cube batches (XPBD_SAT_SCHED, A/B runs): "thread" (default) = thread per
        // candidate, "dual" = two candidates per warp, "warp" = the generic warp-per-candidate kernel.
        const char* e = std::getenv("XPBD_SAT_SCHED");
        const char* dual = std::getenv("XPBD_SAT_DUAL");
        const bool want_warp = (e && !std::strcmp(e, "warp")) || (dual && dual[0] == '0');
        const bool want_dual = e && !std::strcmp(e, "dual");
        if (all_cubes && !want_warp) {
            if (want_dual) {
                st->sat_fn = cull ? stage_sat_cubes_kernel<true> : stage_sat_cubes_kernel<false>;
                st->sat_cands_per_block = 16;
            } else {
                st->sat_fn = cull ? stage_sat_thread_kernel<true> : stage_sat_thread_kernel<false>;
                st->sat_threads = kSatThreadBlock;
                st->sat_cands_per_block = kSatThreadBlock;
                st->smem_sat = sat_thread_smem_bytes(H.bytes);
            }
        }
    }
    XPBD_CUDA_TRY(cudaFuncSetAttribute(st->sat_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)st->smem_sat));
    st->manifold_fn = st->clip_cap == 8 ? stage_manifold_kernel<8> : stage_manifold_kernel<kMaxClip>;
    XPBD_CUDA_TRY(cudaFuncSetAttribute(st->manifold_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)st->smem_manifold));
    int occ = 1;
    XPBD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, st->prepare_fn, 32 * wp, st->smem_prepare));
    st->occ_prepare = occ > 0 ? occ : 1;
    XPBD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, st->sat_fn, st->sat_threads, st->smem_sat));
    st->occ_sat = occ > 0 ? occ : 1;
    XPBD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, st->manifold_fn, kManifoldThreads, st->smem_manifold));
    st->occ_manifold = occ > 0 ? occ : 1;
    return XPBD_OK;
}

// Number of world ranges a step call runs concurrently: the caller's choice (xpbd_batch_set_sub_batches), else
// XPBD_SUB_BATCHES, else one range per 2^18 bodies (rounded) up to 4 for all-cube batches (C3: 4) and up to 3 for
// mixed-hull batches (C4: 3).  The ranges pay where the issue-bound thread-per-candidate SAT kernel of one range shares
// the SMs with the latency-bound stages of the others; the warp-per-candidate kernel of mixed batches is latency-bound
// itself and runs at full occupancy (C4 with the wide solve schedule, measured: 1 range 1.14e9, 2 ranges 1.27e9, 3 ranges
// 1.31e9, 4 ranges 1.28e9 body-substeps/s; the ranges also overlap host copies with kernels in xpbd_batch_step_host_*).
static int pick_sub_batches(const xpbd_batch* b) {
    int k = b->sub_batches;
    if (k <= 0) { if (const char* e = std::getenv("XPBD_SUB_BATCHES")) k = std::atoi(e); }
    if (k <= 0) {
        k = (int)std::min<int64_t>(all_cube_hulls(b) ? 4 : 3, (b->n_bodies + (1 << 17)) >> 18);  // rounded: C3's 10^6 bodies -> 4
        // xpbd_batch_step_host_rigid moves 608 bytes per body over PCIe: the first range's upload and the last range's
        // download are exposed, so twice as many, half as large ranges (C3: e2e_rigid_f64 8.9e8 -> 9.3e8; device time -1 %)
        if (b->io_rigid_in || b->io_rigid_out) k *= 2;
    }
    k = std::max(1, std::min(k, kMaxSubBatches));
    return (int)std::min<int64_t>(k, std::max<int64_t>(1, b->desc.n_worlds));
}

bool staged_takes_host_buffers(const xpbd_batch* b) { return world_path_is_staged(b) && pick_sub_batches(b) > 1; }

static int ensure_sub_streams(xpbd_batch* b, StagedState* st, int n_sub) {
    if (st->n_sub == n_sub && (int)st->sub_streams.size() == (n_sub > 1 ? n_sub : 0)) return XPBD_OK;
    st->release_streams();  // (destroying a stream with pending work is deferred by the runtime until it has drained)
    st->n_sub = n_sub;
    b->fork_needed = true;  // new streams: they must start behind whatever the batch stream already joined
    if (n_sub <= 1) return XPBD_OK;
    XPBD_CUDA_TRY(cudaEventCreateWithFlags(&st->ev_fork, cudaEventDisableTiming));
    for (int k = 0; k < n_sub; ++k) {
        cudaStream_t s = nullptr; cudaEvent_t e = nullptr;
        XPBD_CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        st->sub_streams.push_back(s);
        XPBD_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        st->sub_done.push_back(e);
    }
    if (!st->d_range_counters) {
        XPBD_CUDA_TRY(cudaMalloc(reinterpret_cast<void**>(&st->d_range_counters), kMaxSubBatches * 4 * sizeof(unsigned long long)));
        XPBD_CUDA_TRY(cudaMallocHost(reinterpret_cast<void**>(&st->h_range_counters), kMaxSubBatches * 4 * sizeof(unsigned long long)));
    }
    return XPBD_OK;
}

#ifdef XPBD_CULL_STATS
extern "C" void xpbd_debug_cull_stats(unsigned long long out[43]) {  // debug builds only
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_cull_stats, 43 * sizeof(unsigned long long));
}
#endif

void staged_add_range_counters(const xpbd_batch* b, unsigned long long* acc) {
    const StagedState* st = static_cast<const StagedState*>(b->world_scratch);
    if (!st || !st->h_range_counters) return;
    for (int k = 0; k < st->n_sub; ++k)
        for (int c = 0; c < 4; ++c) acc[c] += st->h_range_counters[4 * k + c];
}

int step_world_staged(xpbd_batch* b, float frame_dt, int substeps, int n_frames) {
    const int B = b->desc.bodies_per_world;
    const DeviceHulls& H = b->dev_hulls;
    if (B > 1024) return fail(XPBD_ERR_CAPACITY, "bodies_per_world > 1024");
    if (H.max_verts > 64 || H.max_face_verts > kMaxClip / 2)
        return fail(XPBD_ERR_CAPACITY, "multi-body path supports hulls with <= 64 vertices and <= %d vertices per face", kMaxClip / 2);
    if (b->desc.body_body_contacts) {
        for (const HostHull& h : b->hulls)
            if (h.edges.empty()) return fail(XPBD_ERR_INVALID, "body-body contacts need hull edges (edge query)");
    }
    StagedState* st = nullptr;
    if (int rc = ensure_staged(b, st)) return rc;
    StagedParams P;
    std::memset(&P, 0, sizeof P);  // (padding included: the ranges' graph keys compare these bytes)
    P.bodies = b->arrays;
    P.n_worlds = b->desc.n_worlds;
    P.B = B;
    P.J = b->desc.joints_per_world;
    P.hull_ids = b->d_hull_ids;
    P.hull_blob = H.blob;
    P.hull_blob_bytes = H.bytes;
    P.joints = b->d_joints;
    P.S = st->buf;
    // same three roundings as make_step_consts on the device (IEEE single division / multiplication)
    P.sc.dt = frame_dt / (float)substeps;
    P.sc.half_dt = P.sc.dt * 0.5f;
    P.sc.compliance = (float)b->desc.compliance / (P.sc.dt * P.sc.dt);
    P.friction = (float)b->desc.friction;
    P.body_body = b->desc.body_body_contacts;
    P.pair_table_len = st->pair_table_len;
    P.compact_verts = 0;
    // (pair slot, world) enumeration instead of a work list: with the thread-per-candidate SAT kernel (XPBD_SAT_ORDER=list|level, A/B runs)
    P.by_level = st->sat_fn == stage_sat_thread_kernel<true> || st->sat_fn == stage_sat_thread_kernel<false>;
    if (const char* e = std::getenv("XPBD_SAT_ORDER")) P.by_level = P.by_level && std::strcmp(e, "list") != 0;
    for (const HostHull& h : b->hulls) if ((int)(h.vertices.size() / 3) != H.max_verts) P.compact_verts = 1;
    P.counters = b->d_counters;
    P.cand_stat = nullptr;
    const int n_sub = pick_sub_batches(b);
    if (int rc = ensure_sub_streams(b, st, n_sub)) return rc;
    // Resident blocks per SM of the persistent SAT kernel: everything the registers allow when it runs alone or on a
    // mixed-hull batch (C4 with 2 ranges: 4 blocks 1.08e9, 3 blocks 1.06e9, 2 blocks 1.01e9 body-substeps/s), 16 warps when
    // several ranges of an all-cube batch share the device (round 1, C3 with 4 ranges of the two-per-warp kernel: 2 blocks
    // of 256 -> 6.47e8, 3 -> 6.42e8, 4 -> 6.33e8; the thread-per-candidate kernel has 128-thread blocks: the same 16 warps
    // are 4 of them).
    int occ_sat = n_sub > 1 && all_cube_hulls(b) ? std::min(st->occ_sat, std::max(2, 512 / st->sat_threads)) : st->occ_sat;
    if (const char* e = std::getenv("XPBD_SAT_BLOCKS_PER_SM")) occ_sat = std::min(st->occ_sat, std::max(1, std::atoi(e)));
    // Schedule of the solve stage: XPBD_SOLVE_LANES = 2 (lane pair per world), 4, 8, 16, 32 (wide) overrides the default.
    int solve_lanes = st->solve_lanes_default;
    if (const char* e = std::getenv("XPBD_SOLVE_LANES")) solve_lanes = std::atoi(e);
    void (*solve_fn)(StagedParams) = stage_solve_kernel;
    switch (solve_lanes) {
        case 4: solve_fn = stage_solve_wide_kernel<4>; break;
        case 8: solve_fn = stage_solve_wide_kernel<8>; break;
        case 16: solve_fn = stage_solve_wide_kernel<16>; break;
        case 32: solve_fn = stage_solve_wide_kernel<32>; break;
        default: solve_lanes = 2; break;
    }
    size_t smem_solve = 0;
    int occ_solve_wide = 1;
    if (solve_fn != stage_solve_kernel) {
        smem_solve = (size_t)(kSolveWideBlock / solve_lanes) * solve_wide_group_words16(B, st->buf.capP, P.J) * sizeof(uint16_t);
        if (smem_solve > 48 * 1024) XPBD_CUDA_TRY(cudaFuncSetAttribute(solve_fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_solve));
        XPBD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_solve_wide, solve_fn, kSolveWideBlock, smem_solve));
        if (occ_solve_wide < 1) return fail(XPBD_ERR_CAPACITY, "worlds too large for the wide solve schedule (%zu bytes of shared memory per block)", smem_solve);
    }
    const int total = substeps * n_frames;
    // The ranges wait for the batch stream only when something was enqueued there since the last step call
    // (b->fork_needed); otherwise each range simply continues behind its own previous frame.
    const bool fork = n_sub > 1 && b->fork_needed;
    if (fork) XPBD_CUDA_TRY(cudaEventRecord(st->ev_fork, b->stream));
    const int64_t worlds_per_block = (int64_t)st->warps_prepare * (32 / st->prepare_lanes);
    // CUDA Graph per world range: with several ranges in flight, the chain of `total` substeps x 4 kernels of a range is
    // captured once per (range, step parameters) and replayed with one cudaGraphLaunch per step call instead of
    // 4 x total launches (C3: 320 launches per frame -> 4 graph launches).  The chain is the same sequence of kernels with
    // the same arguments, so results cannot change.  Single-range calls launch directly: they are the ones whose stages
    // are timed with events between the kernels (xpbd_step_stats.stage_ms).  XPBD_GRAPH=0 switches the graphs off.
    bool use_graph = n_sub > 1 && total <= kGraphMaxSubsteps;
    if (const char* e = std::getenv("XPBD_GRAPH")) use_graph = use_graph && e[0] != '0';
    int ranges_started = 0;  // ranges whose stream has work of this call: they must be joined even when a later range fails
    auto join_started = [&]() {
        for (int k = 0; k < ranges_started && n_sub > 1; ++k) {
            if (cudaEventRecord(st->sub_done[k], st->sub_streams[k]) == cudaSuccess) cudaStreamWaitEvent(b->stream, st->sub_done[k], 0);
        }
    };
    auto run_range = [&](int k) -> int {
        // contiguous world range [w0, w0 + nw) of this sub-batch: every per-world / per-body pointer is advanced to it,
        // so the kernels see an ordinary batch of nw worlds
        const int64_t w0 = P.n_worlds * k / n_sub, nw = P.n_worlds * (k + 1) / n_sub - w0;
        const int64_t g0 = w0 * B;
        StagedParams Q = P;
        Q.n_worlds = nw;
        BodyArrays& A = Q.bodies;
        A.dynA += g0; A.dynB += g0; A.dynC += g0; A.dynD += g0;
        A.k0 += g0; A.k1 += g0; A.k2 += g0; A.k3 += g0; A.k4 += g0; A.k5 += g0; A.k6 += g0;
        if (Q.hull_ids) Q.hull_ids += g0;
        if (Q.joints) Q.joints += w0 * Q.J;
        StagedBuffers& S = Q.S;
        S.ppos += g0; S.prot += g0; S.fpos += g0; S.finvp += g0; S.cen += g0;
        S.verts += g0 * S.maxV;
        S.gcons += w0 * S.capG * kGroundWords;
        S.n_ground += w0; S.n_pairs += w0; S.overflow += w0;
        S.pairs += w0 * S.capP;
        S.pcons += w0 * S.capP * kMaxManifold * 3;
        S.worklist += w0 * S.capP;
        S.work_count += k * kWorkCountStride;
        cudaStream_t stream = b->stream;
        if (n_sub > 1) {  // everything of a range — counters included — lives on the range's stream
            stream = st->sub_streams[k];
            Q.counters = st->d_range_counters + 4 * k;
            if (fork) XPBD_CUDA_TRY(cudaStreamWaitEvent(stream, st->ev_fork, 0));
            ranges_started = k + 1;
            XPBD_CUDA_TRY(cudaMemsetAsync(Q.counters, 0, 4 * sizeof(unsigned long long), stream));
            if (int rc = rigid_io_before(b, g0, nw * B, stream)) return rc;  // xpbd_batch_step_host_rigid, same per-range pipelining
            if (b->io_in) {  // xpbd_batch_step_host_f32: this range's state from the host, under the other ranges' kernels
                XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_dyn_stage + 13 * g0, b->io_in + 13 * g0, (size_t)(nw * B) * 13 * sizeof(float), cudaMemcpyHostToDevice, stream));
                launch_dyn_unpack(b->d_dyn_stage + 13 * g0, b->arrays, g0, nw * B, stream);
            }
        }
        const int grid_prepare = (int)std::min<int64_t>((nw + worlds_per_block - 1) / worlds_per_block, (int64_t)st->n_sm * st->occ_prepare);
        const int grid_sat = (int)std::min<int64_t>((nw * S.capP + st->sat_cands_per_block - 1) / st->sat_cands_per_block, (int64_t)st->n_sm * occ_sat);
        const int grid_manifold = (int)std::min<int64_t>((nw * S.capP + 127) / 128, (int64_t)st->n_sm * st->occ_manifold);
        unsigned grid_solve = (unsigned)((2 * nw + 127) / 128);
        if (st->solve_blocks_per_sm > 0) grid_solve = std::min(grid_solve, (unsigned)(st->solve_blocks_per_sm * st->n_sm));
        if (solve_fn != stage_solve_kernel)
            grid_solve = (unsigned)std::min<int64_t>((nw * solve_lanes + kSolveWideBlock - 1) / kSolveWideBlock, (int64_t)st->n_sm * occ_solve_wide);
        // the chain of this range: overflow flags reset, then `total` substeps of (prepare, sat, manifold, solve)
        auto enqueue_chain = [&](bool with_events) -> int {
            XPBD_CUDA_TRY(cudaMemsetAsync(S.overflow, 0, (size_t)nw * sizeof(int), stream));
            for (int s = 0; s < total; ++s) {
                const bool last = s == total - 1;
                const bool timed = with_events && last && k == 0;  // per-stage device times of the last substep (xpbd_step_stats.stage_ms): range 0
                Q.cand_stat = last ? Q.counters + 3 : nullptr;
                if (timed) cudaEventRecord(b->ev_stage[0], stream);
                st->prepare_fn<<<grid_prepare, 32 * st->warps_prepare, st->smem_prepare, stream>>>(Q);
                if (timed) cudaEventRecord(b->ev_stage[1], stream);
                if (P.body_body) st->sat_fn<<<grid_sat, st->sat_threads, st->smem_sat, stream>>>(Q);
                if (timed) cudaEventRecord(b->ev_stage[2], stream);
                if (P.body_body) st->manifold_fn<<<grid_manifold, kManifoldThreads, st->smem_manifold, stream>>>(Q);
                if (timed) cudaEventRecord(b->ev_stage[3], stream);
                solve_fn<<<grid_solve, 128, smem_solve, stream>>>(Q);
                if (timed) cudaEventRecord(b->ev_stage[4], stream);
            }
            XPBD_CUDA_TRY(cudaGetLastError());
            return XPBD_OK;
        };
        if (use_graph) {
            RangeGraph& G = st->graphs[k];
            RangeGraphKey key;
            std::memset(&key, 0, sizeof key);
            key.params = Q; key.params.cand_stat = nullptr;
            key.total = total; key.grid[0] = grid_prepare; key.grid[1] = grid_sat; key.grid[2] = grid_manifold; key.grid[3] = (int)grid_solve;
            key.fn[0] = (const void*)st->prepare_fn; key.fn[1] = (const void*)st->sat_fn; key.fn[2] = (const void*)st->manifold_fn;
            key.fn[3] = (const void*)solve_fn;
            if (!G.exec || std::memcmp(&G.key, &key, sizeof key) != 0) {
                G.release();
                cudaGraph_t graph = nullptr;
                XPBD_CUDA_TRY(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
                const int rc = enqueue_chain(false);
                const cudaError_t ce = cudaStreamEndCapture(stream, &graph);
                if (rc != XPBD_OK || ce != cudaSuccess) {
                    if (graph) cudaGraphDestroy(graph);
                    return rc != XPBD_OK ? rc : fail(XPBD_ERR_CUDA, "cudaStreamEndCapture failed: %s", cudaGetErrorString(ce));
                }
                const cudaError_t ie = cudaGraphInstantiate(&G.exec, graph, 0);
                cudaGraphDestroy(graph);
                if (ie != cudaSuccess) { G.exec = nullptr; return fail(XPBD_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(ie)); }
                std::memcpy(&G.key, &key, sizeof key);
            }
            XPBD_CUDA_TRY(cudaGraphLaunch(G.exec, stream));
        } else {
            if (int rc = enqueue_chain(true)) return rc;
        }
        if (n_sub > 1 && b->io_out) {  // ... and its result back to the host
            launch_dyn_pack(b->arrays, b->d_dyn_stage + 13 * g0, g0, nw * B, stream);
            XPBD_CUDA_TRY(cudaMemcpyAsync(b->io_out + 13 * g0, b->d_dyn_stage + 13 * g0, (size_t)(nw * B) * 13 * sizeof(float), cudaMemcpyDeviceToHost, stream));
        }
        if (n_sub > 1) { if (int rc = rigid_io_after(b, g0, nw * B, stream)) return rc; }
        if (n_sub > 1)
            XPBD_CUDA_TRY(cudaMemcpyAsync(st->h_range_counters + 4 * k, Q.counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
        return XPBD_OK;
    };
    for (int k = 0; k < n_sub; ++k) {
        if (int rc = run_range(k)) {
            // a failure in the middle: the ranges already forked are joined into the batch stream (the caller's stream sees
            // everything that was enqueued), and the next step call forks again
            join_started();
            b->fork_needed = true;
            b->io_in = nullptr; b->io_out = nullptr; b->io_rigid_in = nullptr; b->io_rigid_out = nullptr;
            return rc;
        }
    }
    if (n_sub > 1) {
        for (int k = 0; k < n_sub; ++k) {  // join: the batch stream sees the step call as one operation
            XPBD_CUDA_TRY(cudaEventRecord(st->sub_done[k], st->sub_streams[k]));
            XPBD_CUDA_TRY(cudaStreamWaitEvent(b->stream, st->sub_done[k], 0));
        }
    }
    b->stage_ms_valid = !use_graph;
    b->fork_needed = false;
    if (n_sub > 1) { b->io_in = nullptr; b->io_out = nullptr; b->io_rigid_in = nullptr; b->io_rigid_out = nullptr; }  // consumed per range
    b->last.sub_batches = n_sub;
    b->stage_timed = true;
    XPBD_CUDA_TRY(cudaGetLastError());
    b->last.kernel_launches = n_sub * total * (P.body_body ? 4 : 2);
    return XPBD_OK;
}

}  // namespace xpbd

// ---- the SAT stage on synthetic candidates (A/B testing of its schedules and of the witness cull) -------------------
namespace xpbd {
// One "world" per candidate: bodies 2i (A) and 2i + 1 (B), pair slot i.  Fills exactly what stage_prepare leaves behind
// for the SAT stage: world-space vertices (frame * vertex, the same expression), inverse frame positions, A's world
// centroid, the rotations, the pair record and the work-list descriptor.
__global__ void sat_fuzz_setup_kernel(const double* fa, const double* fb, int64_t n, const uint32_t* hull_blob, int maxV,
                                      float4* dynC, float4* finvp, float4* cen, float4* verts, int4* pairs, int4* worklist,
                                      int hull_b) {
    const HullTable T{hull_blob};
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < 2 * n; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t >> 1;
        const int side = (int)(t & 1);
        const double* f = (side ? fb : fa) + 7 * i;
        Fr frame;
        frame.p = v3((float)f[0], (float)f[1], (float)f[2]);
        frame.q = q4((float)f[6], (float)f[3], (float)f[4], (float)f[5]);
        const HullDesc& hd = T.desc(side ? hull_b : 0);
        const Fr finv = frame_inverse(frame);
        const V3 c = frame * T.centroid(hd);
        dynC[t] = make_float4(frame.q.x, frame.q.y, frame.q.z, frame.q.s);
        finvp[t] = make_float4(finv.p.x, finv.p.y, finv.p.z, 0.f);
        cen[t] = make_float4(c.x, c.y, c.z, 0.f);
        for (int v = 0; v < hd.n_verts; ++v) {
            const V3 p = frame * T.vertex(hd, v);
            verts[t * maxV + v] = make_float4(p.x, p.y, p.z, 0.f);
        }
        if (!side) {
            pairs[i] = make_int4(0 | (1 << 16), kNone, -1, 0);
            worklist[i] = make_int4((int)i, 0 | (1 << 16), 0 | (hull_b << 16), (int)i);
        }
    }
}
}  // namespace xpbd

extern "C" int xpbd_sat_feature_batch(const xpbd_frame* fa, const xpbd_frame* fb, int64_t n, const xpbd_polytope* ha,
                                      const xpbd_polytope* hb, int32_t schedule, int32_t cull, int32_t* feature, int32_t* index) {
    using namespace xpbd;
    if (!fa || !fb || !ha || !hb || !feature || !index || n < 0) return fail(XPBD_ERR_INVALID, "bad argument");
    if (schedule < 0 || schedule > 2) return fail(XPBD_ERR_INVALID, "schedule must be 0 (warp), 1 (two per warp) or 2 (thread)");
    if (n >= (1ll << 30)) return fail(XPBD_ERR_CAPACITY, "at most 2^30 candidates per call");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(XPBD_ERR_CUDA, "no CUDA device; this library has no CPU fallback");
    std::vector<HostHull> hulls;
    for (const xpbd_polytope* h : {ha, hb}) {
        if (const char* why = xpbd_host::validate_hull(h)) return fail(XPBD_ERR_INVALID, "hull: %s", why);
        if (h->n_vertices > 64 || h->n_edges <= 0) return fail(XPBD_ERR_CAPACITY, "hulls need edges and at most 64 vertices");
        hulls.push_back(copy_hull(h));
        if (schedule != 0 && !(h->n_vertices == 8 && h->n_edges == 12 && h->n_faces == 6))
            return fail(XPBD_ERR_INVALID, "schedules 1 and 2 are for cube-shaped hulls (8 vertices, 12 edges, 6 faces)");
    }
    DeviceHulls dev;
    if (int rc = dev.upload(hulls, nullptr)) return rc;
    struct Bufs {
        DeviceHulls& dev;
        std::vector<void*> ptrs;
        ~Bufs() { for (void* p : ptrs) cudaFree(p); dev.release(); }
        cudaError_t alloc(void** p, size_t bytes) { cudaError_t e = cudaMalloc(p, bytes ? bytes : 16); if (e == cudaSuccess) ptrs.push_back(*p); return e; }
    } bufs{dev, {}};
    if (n == 0) return XPBD_OK;
    const int maxV = dev.max_verts;
    double *d_fa, *d_fb;
    StagedParams P{};
    unsigned long long* d_counters;
    XPBD_CUDA_TRY(bufs.alloc((void**)&d_fa, (size_t)n * 56)); XPBD_CUDA_TRY(bufs.alloc((void**)&d_fb, (size_t)n * 56));
    XPBD_CUDA_TRY(bufs.alloc((void**)&P.bodies.dynC, (size_t)2 * n * 16));
    XPBD_CUDA_TRY(bufs.alloc((void**)&P.S.finvp, (size_t)2 * n * 16)); XPBD_CUDA_TRY(bufs.alloc((void**)&P.S.cen, (size_t)2 * n * 16));
    XPBD_CUDA_TRY(bufs.alloc((void**)&P.S.verts, (size_t)2 * n * maxV * 16));
    XPBD_CUDA_TRY(bufs.alloc((void**)&P.S.pairs, (size_t)n * 16)); XPBD_CUDA_TRY(bufs.alloc((void**)&P.S.worklist, (size_t)n * 16));
    XPBD_CUDA_TRY(bufs.alloc((void**)&P.S.work_count, 64)); XPBD_CUDA_TRY(bufs.alloc((void**)&d_counters, 32));
    XPBD_CUDA_TRY(cudaMemcpy(d_fa, fa, (size_t)n * 56, cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemcpy(d_fb, fb, (size_t)n * 56, cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemset(d_counters, 0, 32));
    const uint32_t wc[4] = {(uint32_t)n, 0, 0, 0};
    XPBD_CUDA_TRY(cudaMemcpy(P.S.work_count, wc, sizeof wc, cudaMemcpyHostToDevice));
    sat_fuzz_setup_kernel<<<1184, 256>>>(d_fa, d_fb, n, dev.blob, maxV, P.bodies.dynC, P.S.finvp, P.S.cen, P.S.verts, P.S.pairs,
                                         P.S.worklist, 1);
    XPBD_CUDA_TRY(cudaGetLastError());
    P.n_worlds = n; P.B = 2; P.J = 0;
    P.hull_blob = dev.blob; P.hull_blob_bytes = dev.bytes;
    P.S.maxV = maxV; P.S.capP = 1;
    P.body_body = 1;
    P.counters = d_counters;
    void (*fn)(StagedParams) = nullptr;
    int threads = 256;
    size_t smem = dev.bytes + 8 * (size_t)sat_words_per_warp(maxV) * 4;
    if (schedule == 0) fn = cull ? stage_sat_kernel<true> : stage_sat_kernel<false>;
    else if (schedule == 1) fn = cull ? stage_sat_cubes_kernel<true> : stage_sat_cubes_kernel<false>;
    else { fn = cull ? stage_sat_thread_kernel<true> : stage_sat_thread_kernel<false>; threads = kSatThreadBlock; smem = sat_thread_smem_bytes(dev.bytes); }
    XPBD_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1, n_sm = 148, devid = 0;
    XPBD_CUDA_TRY(cudaGetDevice(&devid));
    XPBD_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, devid));
    XPBD_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, threads, smem));
    fn<<<n_sm * (occ > 0 ? occ : 1), threads, smem>>>(P);
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaDeviceSynchronize());
    std::vector<int4> host((size_t)n);
    XPBD_CUDA_TRY(cudaMemcpy(host.data(), P.S.pairs, (size_t)n * 16, cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < n; ++i) { feature[i] = host[(size_t)i].y; index[i] = host[(size_t)i].z; }
    return XPBD_OK;
}
