// Boundary kernels: Rigid AoS f64 (the Rust struct's layout, include/xpbd_b200.h xpbd_rigid)
// <-> fp32 SoA in HBM (xpbd_state.cuh); packed fp32 dynamic-state snapshots; hull face planes.
#include "xpbd_hull.cuh"
#include "xpbd_state.cuh"

namespace xpbd {

// One thread per body. AoS reads are strided (38 doubles per record) but contiguous per warp
// overall; this runs at setup / snapshot time, not on the step path.
__global__ void aos64_to_soa_kernel(const double* __restrict__ aos, BodyArrays s, int64_t first, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* r = aos + 38 * i;
    int64_t g = first + i;
    auto f = [&](int k) { return __double2float_rn(r[k]); };
    s.k0[g] = make_float4(f(0), f(28), f(29), f(30));
    s.k1[g] = make_float4(f(1), f(2), f(3), f(4));
    s.k2[g] = make_float4(f(5), f(6), f(7), f(8));
    s.k3[g] = make_float4(f(9), f(10), f(11), f(12));
    s.k4[g] = make_float4(f(13), f(14), f(15), f(16));
    s.k5[g] = make_float4(f(17), f(18), f(19), f(20));
    s.k6[g] = f(21);
    s.dynA[g] = make_float4(f(31), f(32), f(33), f(22));
    s.dynB[g] = make_float4(f(23), f(24), f(25), f(26));
    s.dynC[g] = make_float4(f(34), f(35), f(36), f(37));
    s.dynD[g] = f(27);
}

__global__ void soa_to_aos64_kernel(BodyArrays s, double* __restrict__ aos, int64_t first, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double* r = aos + 38 * i;
    int64_t g = first + i;
    float4 k0 = s.k0[g], k1 = s.k1[g], k2 = s.k2[g], k3 = s.k3[g], k4 = s.k4[g], k5 = s.k5[g];
    float4 a = s.dynA[g], b = s.dynB[g], c = s.dynC[g];
    r[0] = k0.x; r[28] = k0.y; r[29] = k0.z; r[30] = k0.w;
    r[1] = k1.x; r[2] = k1.y; r[3] = k1.z; r[4] = k1.w;
    r[5] = k2.x; r[6] = k2.y; r[7] = k2.z; r[8] = k2.w;
    r[9] = k3.x; r[10] = k3.y; r[11] = k3.z; r[12] = k3.w;
    r[13] = k4.x; r[14] = k4.y; r[15] = k4.z; r[16] = k4.w;
    r[17] = k5.x; r[18] = k5.y; r[19] = k5.z; r[20] = k5.w;
    r[21] = s.k6[g];
    r[31] = a.x; r[32] = a.y; r[33] = a.z; r[22] = a.w;
    r[23] = b.x; r[24] = b.y; r[25] = b.z; r[26] = b.w;
    r[34] = c.x; r[35] = c.y; r[36] = c.z; r[37] = c.w;
    r[27] = s.dynD[g];
}

// Only the 13 values a step mutates (velocity, angular_velocity, position, rotation), widened into AoS f64 records whose
// other 25 fields are left as they are: the caller's f64 constants survive a round trip bit for bit.
__global__ void soa_dyn_to_aos64_kernel(BodyArrays s, double* __restrict__ aos, int64_t first, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double* r = aos + 38 * i;
    int64_t g = first + i;
    float4 a = s.dynA[g], b = s.dynB[g], c = s.dynC[g];
    r[31] = a.x; r[32] = a.y; r[33] = a.z; r[22] = a.w;
    r[23] = b.x; r[24] = b.y; r[25] = b.z; r[26] = b.w;
    r[34] = c.x; r[35] = c.y; r[36] = c.z; r[37] = c.w;
    r[27] = s.dynD[g];
}

// Replicate the bodies of one world (already converted at g in [0, B)) into all worlds.
__global__ void broadcast_world_kernel(BodyArrays s, int bodies_per_world, int64_t n_bodies) {
    int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_bodies || g < bodies_per_world) return;
    int64_t src = g % bodies_per_world;
    s.k0[g] = s.k0[src]; s.k1[g] = s.k1[src]; s.k2[g] = s.k2[src]; s.k3[g] = s.k3[src];
    s.k4[g] = s.k4[src]; s.k5[g] = s.k5[src]; s.k6[g] = s.k6[src];
    s.dynA[g] = s.dynA[src]; s.dynB[g] = s.dynB[src]; s.dynC[g] = s.dynC[src]; s.dynD[g] = s.dynD[src];
}

// Packed dynamic snapshot: 13 floats per body: px py pz qx qy qz qs vx vy vz wx wy wz.
__global__ void dyn_pack_kernel(BodyArrays s, float* __restrict__ out, int64_t first, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t g = first + i;
    float4 a = s.dynA[g], b = s.dynB[g], c = s.dynC[g];
    float* o = out + 13 * i;
    o[0] = a.x; o[1] = a.y; o[2] = a.z;
    o[3] = c.x; o[4] = c.y; o[5] = c.z; o[6] = c.w;
    o[7] = a.w; o[8] = b.x; o[9] = b.y;
    o[10] = b.z; o[11] = b.w; o[12] = s.dynD[g];
}
__global__ void dyn_unpack_kernel(const float* __restrict__ in, BodyArrays s, int64_t first, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int64_t g = first + i;
    const float* o = in + 13 * i;
    s.dynA[g] = make_float4(o[0], o[1], o[2], o[7]);
    s.dynB[g] = make_float4(o[8], o[9], o[10], o[11]);
    s.dynC[g] = make_float4(o[3], o[4], o[5], o[6]);
    s.dynD[g] = o[12];
}

// Polytope::plane (src/geometry.rs:262-271) for every face of every hull, in fp32 on the device,
// written into the blob's plane section.  One thread per (hull, face).
__global__ void hull_planes_kernel(uint32_t* blob, int n_hulls) {
    HullTable t{blob};
    int h = blockIdx.x;
    if (h >= n_hulls) return;
    const HullDesc d = t.desc(h);
    for (int f = threadIdx.x; f < d.n_faces; f += blockDim.x) {
        int b = t.face_begin(d, f);
        V3 p0 = t.vertex(d, t.face_vertex(d, b));
        V3 p1 = t.vertex(d, t.face_vertex(d, b + 1));
        V3 p2 = t.vertex(d, t.face_vertex(d, b + 2));
        // Plane::from_points (src/geometry.rs:16-24)
        V3 n = normalize(cross(p1 - p0, p2 - p0));
        float disp = dot(n, p0);
        // Plane::facing(centroid) (src/geometry.rs:66-68): normal . (c - displacement*normal) >= 0
        V3 support = disp * n;
        bool facing = dot(n, t.centroid(d) - support) >= 0.0f;
        if (facing) { n = -n; disp = -disp; }
        reinterpret_cast<float4*>(blob + d.plane_off)[f] = make_float4(n.x, n.y, n.z, disp);
    }
}

// Failure detection (the reference panics; here a diverged world is counted): worlds whose dynamic
// state holds a NaN / Inf after a step call.  Thread per world over its B bodies; one atomic per warp that saw any.
__global__ void count_nan_worlds_kernel(BodyArrays s, int64_t n_worlds, int B, unsigned long long* counter) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = false;
    if (w < n_worlds) {
        for (int b = 0; b < B; ++b) {
            BodyDyn d;
            load_dyn(s, w * B + b, d);
            bad = bad || dyn_nonfinite(d);
        }
    }
    const unsigned n = __popc(__ballot_sync(0xffffffffu, bad));
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(counter, (unsigned long long)n);
}
void launch_count_nan_worlds(const BodyArrays& s, int64_t n_worlds, int B, unsigned long long* counter, cudaStream_t st) {
    if (n_worlds > 0) count_nan_worlds_kernel<<<(unsigned)((n_worlds + 127) / 128), 128, 0, st>>>(s, n_worlds, B, counter);
}

void launch_aos64_to_soa(const double* aos, const BodyArrays& s, int64_t first, int64_t n, cudaStream_t st) {
    if (n > 0) aos64_to_soa_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(aos, s, first, n);
}
void launch_soa_to_aos64(const BodyArrays& s, double* aos, int64_t first, int64_t n, cudaStream_t st) {
    if (n > 0) soa_to_aos64_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s, aos, first, n);
}
void launch_soa_dyn_to_aos64(const BodyArrays& s, double* aos, int64_t first, int64_t n, cudaStream_t st) {
    if (n > 0) soa_dyn_to_aos64_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s, aos, first, n);
}
void launch_broadcast_world(const BodyArrays& s, int bodies_per_world, int64_t n_bodies, cudaStream_t st) {
    if (n_bodies > 0) broadcast_world_kernel<<<(unsigned)((n_bodies + 255) / 256), 256, 0, st>>>(s, bodies_per_world, n_bodies);
}
void launch_dyn_pack(const BodyArrays& s, float* out, int64_t first, int64_t n, cudaStream_t st) {
    if (n > 0) dyn_pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(s, out, first, n);
}
void launch_dyn_unpack(const float* in, const BodyArrays& s, int64_t first, int64_t n, cudaStream_t st) {
    if (n > 0) dyn_unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, s, first, n);
}
void launch_hull_planes(uint32_t* blob, int n_hulls, cudaStream_t st) {
    if (n_hulls > 0) hull_planes_kernel<<<n_hulls, 64, 0, st>>>(blob, n_hulls);
}

}  // namespace xpbd
