// Shared-memory layout of the multi-body step kernel (xpbd_world.cu).
//
// Two kinds of regions follow the hull table in a block's dynamic shared memory:
//   * one PERSISTENT region per resident world (state of its bodies + its constraint list), alive for the whole
//     step call.  Arrays are indexed [component * B + body] (structure-of-arrays over the bodies of the world);
//     regions are `stride` words apart with `stride` odd so that the solve phase — lanes spread over worlds —
//     does not hit one bank from every lane;
//   * one SCRATCH region per warp, used while that warp generates the contacts of one world (world-space hull
//     vertices, frames, candidate pairs, the edge query's survivor queue).  16-byte aligned: vertices are float4.
#pragma once
#include "xpbd_math.cuh"
#include "xpbd_narrow.cuh"

namespace xpbd {

struct WorldLayout {
    int B;     // bodies per world
    int maxV;  // max hull vertices
    int capC;  // constraint capacity per world
    int capP;  // candidate pair capacity per world
    int J;     // joints per world
    // persistent region: word offsets
    int pos, rot, vel, ang;                          // dynamic: 3B, 4B, 3B, 3B
    int invm, inertia, com, fext, fint, text, tint;  // constants: B, 9B, 3B x 5
    int ppos, prot;                                  // pose at the start of the substep: 3B, 4B
    int hull;                                        // hull id per body
    int cons;                                        // constraint records, kConsWords each
    int counters;                                    // [0] n_constraints of this substep, [1] overflow flag
    int stride;                                      // words per persistent region (odd)
    // scratch region: word offsets (all multiples of 4)
    int s_verts;   // world-space hull vertices: float4 x maxV per body
    int s_local;   // float4 x maxV: the other hull's vertices in the face query's local space
    int s_queue;   // edge query survivor queue
    int s_fpos, s_finvp, s_cen, s_pfpos;  // post-integrate frame position, its inverse's position, centroid, past frame position: 3B each
    int s_pairs;   // candidate pairs: 3 words each (packed ij, feature, index)
    int sstride;   // words per scratch region (multiple of 4)
};

constexpr int kConsWords = 10;

__host__ __device__ inline WorldLayout make_world_layout(int B, int maxV, int capC, int capP, int J) {
    WorldLayout L;
    L.B = B; L.maxV = maxV; L.capC = capC; L.capP = capP; L.J = J;
    int o = 0;
    auto take = [&](int words) { int r = o; o += words; return r; };
    L.pos = take(3 * B); L.rot = take(4 * B); L.vel = take(3 * B); L.ang = take(3 * B);
    L.invm = take(B); L.inertia = take(9 * B); L.com = take(3 * B);
    L.fext = take(3 * B); L.fint = take(3 * B); L.text = take(3 * B); L.tint = take(3 * B);
    L.ppos = take(3 * B); L.prot = take(4 * B);
    L.hull = take(B);
    L.cons = take(kConsWords * capC);
    L.counters = take(2);
    L.stride = o | 1;
    o = 0;
    auto take4 = [&](int words) { int r = o; o += (words + 3) & ~3; return r; };
    L.s_verts = take4(4 * maxV * B);
    L.s_local = take4(4 * maxV);
    L.s_queue = take4(kEdgeQueueWords);
    L.s_fpos = take4(3 * B); L.s_finvp = take4(3 * B); L.s_cen = take4(3 * B); L.s_pfpos = take4(3 * B);
    L.s_pairs = take4(3 * capP);
    L.sstride = o;
    return L;
}

// Constraint record in shared memory (kConsWords words):
//   [0] kind (bits 0-1) | active (bit 2) | rigid_a << 4 (12 bits) | (rigid_b + 1) << 16 (12 bits)
//   kGround : [1..3] contact0 (world)   [4..6] direction (unit)   [7] |contact1 - contact0|
//   kContact: [1..3] a (object space of rigid_a)  [4..6] b (object space of rigid_b)  [7..9] normal (world)
//   kJoint  : [1..3] anchor a  [4..6] anchor b (world space if rigid_b < 0)  [7] rest distance
enum ConstraintKind : int { kGround = 0, kContact = 1, kJoint = 2 };
constexpr int kActiveBit = 4;
XD int cons_head(int kind, int ra, int rb) { return kind | (ra << 4) | ((rb + 1) << 16); }
XD int head_kind(int h) { return h & 3; }
XD int head_a(int h) { return (h >> 4) & 0xfff; }
XD int head_b(int h) { return ((h >> 16) & 0xfff) - 1; }

XD V3 sld3(const float* base, int B, int b) { return v3(base[b], base[B + b], base[2 * B + b]); }
XD void sst3(float* base, int B, int b, V3 v) { base[b] = v.x; base[B + b] = v.y; base[2 * B + b] = v.z; }
XD Q4 sldq(const float* base, int B, int b) { return q4(base[3 * B + b], base[b], base[B + b], base[2 * B + b]); }
XD void sstq(float* base, int B, int b, Q4 q) { base[b] = q.x; base[B + b] = q.y; base[2 * B + b] = q.z; base[3 * B + b] = q.s; }
XD M3 sldm(const float* base, int B, int b) {
    M3 m;
    m.c0 = v3(base[b], base[B + b], base[2 * B + b]);
    m.c1 = v3(base[3 * B + b], base[4 * B + b], base[5 * B + b]);
    m.c2 = v3(base[6 * B + b], base[7 * B + b], base[8 * B + b]);
    return m;
}

}  // namespace xpbd
