// Shared-memory layout of one world inside the multi-body step kernel (xpbd_world.cu).
// Every array is indexed [component * B + body] (structure-of-arrays over the bodies of the world), so the
// per-body lanes of a warp touch consecutive banks; the per-world regions of a block are `stride` words
// apart with `stride` odd, so the solve phase (one lane per world) is conflict-free too.
#pragma once
#include "xpbd_math.cuh"

namespace xpbd {

struct WorldLayout {
    int B;          // bodies per world
    int maxV;       // max hull vertices
    int capC;       // constraint capacity per world
    int capP;       // candidate pair capacity per world
    int J;          // joints per world
    // word offsets inside a world region
    int pos, rot, vel, ang;                      // dynamic: 3B, 4B, 3B, 3B
    int invm, inertia, com, fext, fint, text, tint;  // constants: B, 9B, 3B x 5
    int ppos, prot, pfpos;                       // past position / rotation / frame position: 3B, 4B, 3B
    int fpos, finvp, cen;                        // post-integrate frame position, inverse-frame position, world centroid
    int hull;                                    // int per body
    int verts;                                   // world-space hull vertices: 3 * maxV per body
    int cons;                                    // constraints: 12 words each
    int pairs;                                   // candidate pairs: 3 words each (packed ij, feature, index)
    int scratchL;                                // 3 * maxV: A-local vertices for the face query
    int counters;                                // [0] n_constraints, [1] n_pairs, [2] overflow
    int stride;                                  // words per world region (odd)
};

__host__ __device__ inline WorldLayout make_world_layout(int B, int maxV, int capC, int capP, int J) {
    WorldLayout L;
    L.B = B; L.maxV = maxV; L.capC = capC; L.capP = capP; L.J = J;
    int o = 0;
    auto take = [&](int words) { int r = o; o += words; return r; };
    L.pos = take(3 * B); L.rot = take(4 * B); L.vel = take(3 * B); L.ang = take(3 * B);
    L.invm = take(B); L.inertia = take(9 * B); L.com = take(3 * B);
    L.fext = take(3 * B); L.fint = take(3 * B); L.text = take(3 * B); L.tint = take(3 * B);
    L.ppos = take(3 * B); L.prot = take(4 * B); L.pfpos = take(3 * B);
    L.fpos = take(3 * B); L.finvp = take(3 * B); L.cen = take(3 * B);
    L.hull = take(B);
    L.verts = take(3 * maxV * B);
    L.cons = take(12 * capC);
    L.pairs = take(3 * capP);
    L.scratchL = take(3 * maxV);
    L.counters = take(4);
    L.stride = o | 1;
    return L;
}

// Constraint record in shared memory (12 words):
//   [0] kind | active << 8      [1] rigid_a | (rigid_b + 1) << 16
//   kGround : [2..4] contact0 (world)   [5..7] direction (unit)   [11] |contact1 - contact0|
//   kContact: [2..4] a (object space of rigid_a)  [5..7] b (object space of rigid_b)  [8..10] normal (world)
//   kJoint  : [2..4] anchor a  [5..7] anchor b (world space if rigid_b < 0)  [11] rest distance
enum ConstraintKind : int { kGround = 0, kContact = 1, kJoint = 2 };

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
