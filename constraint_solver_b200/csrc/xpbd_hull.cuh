// Device hull table: every registered Polytope (src/geometry.rs:81-93) packed into one
// 16-byte-aligned blob of 4-byte words so that a block stages it into shared memory with a
// single TMA bulk copy (xpbd_tma.cuh).
//
//   words [0, 16*n_hulls)            HullDesc per hull (64 B each)
//   then, per hull, 16-byte aligned:  vertices float4 (x, y, z, 0)
//                                     face planes float4 (normal, displacement)  [Polytope::plane,
//                                         computed on the device in fp32 by hull_planes_kernel]
//                                     edges int2
//                                     face offsets int (n_faces + 1), face indices int
#pragma once
#include "xpbd_math.cuh"

namespace xpbd {

struct HullDesc {
    int32_t vert_off;   // word offset of the float4 vertex array
    int32_t n_verts;
    int32_t plane_off;  // word offset of the float4 plane array
    int32_t n_faces;
    int32_t edge_off;   // word offset of the int2 edge array
    int32_t n_edges;
    int32_t foff_off;   // word offset of face offsets (n_faces + 1 ints)
    int32_t fidx_off;   // word offset of face vertex indices
    float cx, cy, cz;   // centroid (object space)
    float radius;       // max |vertex - centroid| (bounding sphere for the broadphase)
    int32_t max_face_verts;
    // Box-like hulls (8 vertices, 12 edges whose fp32 object-space edge vectors fall into 3 classes of 4 EXACTLY parallel
    // edges — cubes and axis-scaled boxes): edge_perm = the 12 edge indices sorted by class, 4 bits each (entry k in bits
    // [4k, 4k + 4) of lo | hi << 32); box_classes = 1.  Otherwise box_classes = 0.  Used by the thread-per-candidate SAT
    // kernel to cull four parallel edges of B with one test.
    uint32_t edge_perm_lo, edge_perm_hi;
    int32_t box_classes;
};
static_assert(sizeof(HullDesc) == 64, "HullDesc must be 64 bytes");

struct HullTable {
    const uint32_t* words;  // blob base (global memory or shared memory)
    XD const HullDesc& desc(int h) const { return reinterpret_cast<const HullDesc*>(words)[h]; }
    XD V3 vertex(const HullDesc& d, int i) const {
        const float4 v = reinterpret_cast<const float4*>(words + d.vert_off)[i];
        return v3(v.x, v.y, v.z);
    }
    XD float4 plane(const HullDesc& d, int f) const { return reinterpret_cast<const float4*>(words + d.plane_off)[f]; }
    XD int2 edge(const HullDesc& d, int e) const { return reinterpret_cast<const int2*>(words + d.edge_off)[e]; }
    XD int face_begin(const HullDesc& d, int f) const { return (int)words[d.foff_off + f]; }
    XD int face_vertex(const HullDesc& d, int k) const { return (int)words[d.fidx_off + k]; }
    XD V3 centroid(const HullDesc& d) const { return v3(d.cx, d.cy, d.cz); }
};

}  // namespace xpbd
