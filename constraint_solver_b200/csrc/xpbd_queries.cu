// Batched building blocks behind the per-function entry points of the C ABI: each reference function on
// the path (Rigid::integrate/derive/apply_impulse/frame, collision::ground, solver::solve,
// face/edge_axes_separation, the finished sat) as its own CUDA kernel over n independent items.
// Host records are the reference's AoS f64; the kernels round to fp32 on load and widen on store.
#include <cstring>
#include <mutex>

#include "xpbd_host_math.hpp"
#include "xpbd_internal.hpp"
#include "xpbd_narrow.cuh"
#include "xpbd_rigid.cuh"

namespace xpbd {

XD float f32(double x) { return __double2float_rn(x); }

XD void load_rigid64(const double* r, BodyDyn& d, BodyConst& k) {
    k.inverse_mass = f32(r[0]);
    k.inverse_inertia.c0 = v3(f32(r[1]), f32(r[2]), f32(r[3]));
    k.inverse_inertia.c1 = v3(f32(r[4]), f32(r[5]), f32(r[6]));
    k.inverse_inertia.c2 = v3(f32(r[7]), f32(r[8]), f32(r[9]));
    k.external_force = v3(f32(r[10]), f32(r[11]), f32(r[12]));
    k.internal_force = v3(f32(r[13]), f32(r[14]), f32(r[15]));
    k.external_torque = v3(f32(r[16]), f32(r[17]), f32(r[18]));
    k.internal_torque = v3(f32(r[19]), f32(r[20]), f32(r[21]));
    d.velocity = v3(f32(r[22]), f32(r[23]), f32(r[24]));
    d.angular_velocity = v3(f32(r[25]), f32(r[26]), f32(r[27]));
    k.center_of_mass = v3(f32(r[28]), f32(r[29]), f32(r[30]));
    d.position = v3(f32(r[31]), f32(r[32]), f32(r[33]));
    d.rotation = q4(f32(r[37]), f32(r[34]), f32(r[35]), f32(r[36]));
}
XD void store_dyn64(double* r, const BodyDyn& d) {
    r[22] = d.velocity.x; r[23] = d.velocity.y; r[24] = d.velocity.z;
    r[25] = d.angular_velocity.x; r[26] = d.angular_velocity.y; r[27] = d.angular_velocity.z;
    r[31] = d.position.x; r[32] = d.position.y; r[33] = d.position.z;
    r[34] = d.rotation.x; r[35] = d.rotation.y; r[36] = d.rotation.z; r[37] = d.rotation.s;
}
XD Fr load_frame64(const double* f) {
    Fr o;
    o.p = v3(f32(f[0]), f32(f[1]), f32(f[2]));
    o.q = q4(f32(f[6]), f32(f[3]), f32(f[4]), f32(f[5]));
    return o;
}

__global__ void rigid_integrate_kernel(double* rigids, int64_t n, float dt) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BodyDyn d; BodyConst k;
    load_rigid64(rigids + 38 * i, d, k);
    StepConsts sc; sc.dt = dt; sc.half_dt = fmul(dt, 0.5f); sc.compliance = 0.0f;
    rigid_integrate(d, k, dt * k.inverse_inertia, sc);
    store_dyn64(rigids + 38 * i, d);
}
__global__ void rigid_derive_kernel(double* rigids, const double* past, int64_t n, float dt) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BodyDyn d; BodyConst k;
    load_rigid64(rigids + 38 * i, d, k);
    const double* p = past + 7 * i;
    rigid_derive(d, v3(f32(p[0]), f32(p[1]), f32(p[2])), q4(f32(p[6]), f32(p[3]), f32(p[4]), f32(p[5])), dt);
    store_dyn64(rigids + 38 * i, d);
}
__global__ void rigid_apply_impulse_kernel(double* rigids, const double* ip, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BodyDyn d; BodyConst k;
    load_rigid64(rigids + 38 * i, d, k);
    const double* p = ip + 6 * i;
    rigid_apply_impulse(d.position, d.rotation, k.inverse_mass, k.inverse_inertia, k.center_of_mass,
                        v3(f32(p[0]), f32(p[1]), f32(p[2])), v3(f32(p[3]), f32(p[4]), f32(p[5])));
    store_dyn64(rigids + 38 * i, d);
}
__global__ void rigid_frame_kernel(const double* rigids, int64_t n, double* out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BodyDyn d; BodyConst k;
    load_rigid64(rigids + 38 * i, d, k);
    Fr f = rigid_frame(d.position, d.rotation, k.center_of_mass);
    double* o = out + 7 * i;
    o[0] = f.p.x; o[1] = f.p.y; o[2] = f.p.z; o[3] = f.q.x; o[4] = f.q.y; o[5] = f.q.z; o[6] = f.q.s;
}

// xpbd_constraint device view: 8 doubles {rigid|rigid_b as two int32, contact0[3], contact1[3], distance}
XD void store_constraint64(double* c, int rigid, int rigid_b, V3 c0, V3 c1, float distance) {
    int32_t* ids = reinterpret_cast<int32_t*>(c);
    ids[0] = rigid; ids[1] = rigid_b;
    c[1] = c0.x; c[2] = c0.y; c[3] = c0.z; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = distance;
}

// collision::ground (src/collision.rs:13-35), one thread per body.
__global__ void ground_kernel(const double* rigids, const double* past, int64_t n, const uint32_t* blob, float friction,
                              double* out, int32_t* counts) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    HullTable T{blob};
    const HullDesc hd = T.desc(0);
    BodyDyn d; BodyConst k;
    load_rigid64(rigids + 38 * i, d, k);
    const Fr pastf = load_frame64(past + 7 * i);
    const Fr frame = rigid_frame(d.position, d.rotation, k.center_of_mass);
    const Fr finv = frame_inverse(frame);
    int c = 0;
    for (int v = 0; v < hd.n_verts; ++v) {
        const V3 p = frame * T.vertex(hd, v);
        if (p.z >= 0.0f) continue;
        const V3 target = v3(p.x, p.y, 0.0f);
        const V3 correction = target - p;
        const V3 delta = p - (pastf * (finv * p));
        const V3 delta_t = delta - project_on(delta, correction);
        store_constraint64(out + 8 * (i * hd.n_verts + c), 0, -1, p, target - friction * delta_t, 0.0f);
        ++c;
    }
    counts[i] = c;
}

// solver::solve (src/solver.rs:19-27), one thread per body, constraints in list order.
__global__ void solve_kernel(double* rigids, int64_t n, const double* cons, const int32_t* counts, int stride, float dt,
                             float compliance_raw) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BodyDyn d; BodyConst k;
    load_rigid64(rigids + 38 * i, d, k);
    const float compliance = fdiv(compliance_raw, fmul(dt, dt));
    for (int c = 0; c < counts[i]; ++c) {
        const double* q = cons + 8 * ((int64_t)i * stride + c);
        const V3 c0 = v3(f32(q[1]), f32(q[2]), f32(q[3])), c1 = v3(f32(q[4]), f32(q[5]), f32(q[6]));
        const V3 diff = c1 - c0;
        const float dist = magnitude(diff);
        const V3 dir = diff * frcp(dist);
        const float w = inverse_resistance(d.position, d.rotation, k.inverse_mass, k.inverse_inertia, k.center_of_mass, c0, dir);
        const float lambda = fdiv(fsub(dist, f32(q[7])), fadd(w, compliance));
        rigid_apply_impulse(d.position, d.rotation, k.inverse_mass, k.inverse_inertia, k.center_of_mass, lambda * dir, c0);
    }
    store_dyn64(rigids + 38 * i, d);
}

// ---- pair queries: one warp per frame pair, hull 0 = A, hull 1 = B in the blob ----------------------
constexpr int kQueryMaxVerts = 256;

struct __align__(16) PairScratch {
    float WA[kVertWords * kQueryMaxVerts];
    float WB[kVertWords * kQueryMaxVerts];
    float L[kVertWords * kQueryMaxVerts];
    float Q[kEdgeQueueWords];
};

XD void pair_setup(const HullTable& T, const HullDesc& hA, const HullDesc& hB, const Fr& FA, const Fr& FB, PairScratch& s,
                   int lane) {
    for (int v = lane; v < hA.n_verts; v += 32) st3(s.WA, v, FA * T.vertex(hA, v));
    for (int v = lane; v < hB.n_verts; v += 32) st3(s.WB, v, FB * T.vertex(hB, v));
    __syncwarp();
}

// mode 0: face query (A faces vs B verts); 1: edge query; 2: full sat manifold
__global__ void __launch_bounds__(32) pair_query_kernel(const double* fa, const double* fb, int64_t n, const uint32_t* blob,
                                                        int mode, double* out_d, int32_t* out_i, double* out_cons,
                                                        int32_t* out_counts, int32_t* out_feature) {
    __shared__ PairScratch s;
    const int64_t i = blockIdx.x;
    const int lane = threadIdx.x;
    if (i >= n) return;
    HullTable T{blob};
    const HullDesc hA = T.desc(0), hB = T.desc(1);
    const Fr FA = load_frame64(fa + 7 * i), FB = load_frame64(fb + 7 * i);
    pair_setup(T, hA, hB, FA, FB, s, lane);
    if (mode == 0) {
        float d; int f;
        face_query(T, hA, frame_inverse(FA), s.WB, hB.n_verts, s.L, lane, d, f);
        if (lane == 0) { out_d[i] = d; out_i[i] = f; }
        return;
    }
    const V3 CA = FA * T.centroid(hA);
    if (mode == 1) {
        float d; int e;
        edge_query(T, hA, hB, s.WA, s.WB, CA, s.Q, lane, d, e);
        if (lane == 0) {
            out_d[i] = d;
            out_i[2 * i] = e < 0 ? -1 : e / hB.n_edges;
            out_i[2 * i + 1] = e < 0 ? -1 : e % hB.n_edges;
        }
        return;
    }
    // finished sat
    float da, db, de = 0.0f; int fA, fB, ep = -1;
    int feature = kNone;
    face_query(T, hA, frame_inverse(FA), s.WB, hB.n_verts, s.L, lane, da, fA);
    bool hit = !(da >= 0.0f);
    // the step's edge query: witness cull + the reference arithmetic on the survivors (same outcome as mode 1's full scan)
    int2 wi = make_int2(0, 0);
    if (hit) wi = witness_vertices<32>(reinterpret_cast<const float4*>(T.words + hA.plane_off)[fA < 0 ? 0 : fA], s.L, hB.n_verts, lane, 0);
    __syncwarp();
    if (hit) { face_query(T, hB, frame_inverse(FB), s.WA, hA.n_verts, s.L, lane, db, fB); hit = !(db >= 0.0f); }
    if (hit) {
        uint32_t rkey = 0;
        for (int v = lane; v < hA.n_verts; v += 32) rkey = max(rkey, abs_coord_key(reinterpret_cast<const float4*>(s.WA)[v]));
        for (int v = lane; v < hB.n_verts; v += 32) rkey = max(rkey, abs_coord_key(reinterpret_cast<const float4*>(s.WB)[v]));
        const float R = __int_as_float((int)__reduce_max_sync(0xffffffffu, rkey));
        edge_query_culled_t<0, 0, 0, 0>(reinterpret_cast<const int2*>(T.words + hA.edge_off), reinterpret_cast<const int2*>(T.words + hB.edge_off),
                                        hA.n_verts, hA.n_edges, hB.n_verts, hB.n_edges, s.WA, s.WB, CA, s.Q, lane,
                                        make_edge_cull(da, db, R, ld3(s.WB, wi.x), ld3(s.WB, wi.y)), de, ep);
        hit = !(de >= 0.0f);
    }
    if (lane == 0) {
        Manifold m; m.count = 0;
        int incident = 0;
        if (hit) {
            const PairFeature pf = choose_feature(da, fA, db, fB, de, ep);
            feature = pf.feature;
            pair_manifold(T, hA, hB, FA, FB, s.WA, s.WB, pf, m, incident);
        }
        for (int k = 0; k < m.count; ++k)
            store_constraint64(out_cons + 8 * (i * kMaxManifold + k), incident, 1 - incident, m.point[k], m.target[k], 0.0f);
        out_counts[i] = m.count;
        out_feature[i] = feature;
    }
}

// FP32 issue-rate microbenchmark: independent FADD/FMUL (or FFMA) chains, 8 per thread.
__global__ void fp32_peak_kernel(float* out, int iters, int use_fma) {
    float a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = 1.0f + threadIdx.x * 1e-6f + k;
    const float m = 1.0000001f, c = 1e-7f;
    if (use_fma) {
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = __fmaf_rn(a[k], m, c);
        }
    } else {
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int k = 0; k < 8; ++k) a[k] = __fadd_rn(__fmul_rn(a[k], m), c);
        }
    }
    float s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace xpbd

// ---- host side of the per-function entry points ------------------------------------------------------
using namespace xpbd;

namespace {

struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 16); }
    template <class T> T* as() { return static_cast<T*>(p); }
};

int need_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(XPBD_ERR_CUDA, "no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    return XPBD_OK;
}

#define BLOCKS(n) (unsigned)(((n) + 127) / 128), 128

// upload -> kernel -> download of n rigid records
template <class Launch>
int run_on_rigids(xpbd_rigid* rigids, int64_t n, Launch launch) {
    if (!rigids || n < 0) return fail(XPBD_ERR_INVALID, "bad rigids/n");
    if (int rc = need_device()) return rc;
    if (n == 0) return XPBD_OK;
    DevBuf d;
    XPBD_CUDA_TRY(d.alloc((size_t)n * sizeof(xpbd_rigid)));
    XPBD_CUDA_TRY(cudaMemcpy(d.p, rigids, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyHostToDevice));
    if (int rc = launch(d.as<double>())) return rc;
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaMemcpy(rigids, d.p, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyDeviceToHost));
    return XPBD_OK;
}

int upload_hulls(const xpbd_polytope* a, const xpbd_polytope* b, DeviceHulls& dev, bool need_edges) {
    std::vector<HostHull> hulls;
    for (const xpbd_polytope* h : {a, b}) {
        if (!h) continue;
        if (const char* why = xpbd_host::validate_hull(h)) return fail(XPBD_ERR_INVALID, "hull: %s", why);
        if (h->n_vertices > kQueryMaxVerts) return fail(XPBD_ERR_CAPACITY, "hull has more than %d vertices", kQueryMaxVerts);
        if (need_edges && h->n_edges == 0) return fail(XPBD_ERR_INVALID, "hull has no edges");
        hulls.push_back(copy_hull(h));
    }
    return dev.upload(hulls, nullptr);
}

}  // namespace

extern "C" {

int xpbd_rigid_integrate_batch(xpbd_rigid* rigids, int64_t n, double dt) {
    return run_on_rigids(rigids, n, [&](double* d) { rigid_integrate_kernel<<<BLOCKS(n)>>>(d, n, (float)dt); return 0; });
}

int xpbd_rigid_derive_batch(xpbd_rigid* rigids, const double* past, int64_t n, double dt) {
    if (!past) return fail(XPBD_ERR_INVALID, "past is null");
    return run_on_rigids(rigids, n, [&](double* d) -> int {
        DevBuf p;
        XPBD_CUDA_TRY(p.alloc((size_t)n * 56));
        XPBD_CUDA_TRY(cudaMemcpy(p.p, past, (size_t)n * 56, cudaMemcpyHostToDevice));
        rigid_derive_kernel<<<BLOCKS(n)>>>(d, p.as<double>(), n, (float)dt);
        XPBD_CUDA_TRY(cudaDeviceSynchronize());
        return 0;
    });
}

int xpbd_rigid_apply_impulse_batch(xpbd_rigid* rigids, const double* impulse_point, int64_t n) {
    if (!impulse_point) return fail(XPBD_ERR_INVALID, "impulse_point is null");
    return run_on_rigids(rigids, n, [&](double* d) -> int {
        DevBuf p;
        XPBD_CUDA_TRY(p.alloc((size_t)n * 48));
        XPBD_CUDA_TRY(cudaMemcpy(p.p, impulse_point, (size_t)n * 48, cudaMemcpyHostToDevice));
        rigid_apply_impulse_kernel<<<BLOCKS(n)>>>(d, p.as<double>(), n);
        XPBD_CUDA_TRY(cudaDeviceSynchronize());
        return 0;
    });
}

int xpbd_rigid_frame_batch(const xpbd_rigid* rigids, int64_t n, xpbd_frame* out) {
    if (!rigids || !out || n < 0) return fail(XPBD_ERR_INVALID, "bad argument");
    if (int rc = need_device()) return rc;
    if (n == 0) return XPBD_OK;
    DevBuf d, o;
    XPBD_CUDA_TRY(d.alloc((size_t)n * sizeof(xpbd_rigid)));
    XPBD_CUDA_TRY(o.alloc((size_t)n * sizeof(xpbd_frame)));
    XPBD_CUDA_TRY(cudaMemcpy(d.p, rigids, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyHostToDevice));
    rigid_frame_kernel<<<BLOCKS(n)>>>(d.as<double>(), n, o.as<double>());
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaMemcpy(out, o.p, (size_t)n * sizeof(xpbd_frame), cudaMemcpyDeviceToHost));
    return XPBD_OK;
}

int xpbd_ground_batch(const xpbd_rigid* rigids, const xpbd_frame* past, int64_t n, const xpbd_polytope* hull,
                      xpbd_constraint* out, int32_t* counts) {
    if (!rigids || !past || !out || !counts || n < 0) return fail(XPBD_ERR_INVALID, "bad argument");
    if (int rc = need_device()) return rc;
    DeviceHulls dev;
    if (int rc = upload_hulls(hull, nullptr, dev, false)) return rc;
    struct Release { DeviceHulls& d; ~Release() { d.release(); } } rel{dev};
    if (n == 0) return XPBD_OK;
    const size_t nc = (size_t)n * hull->n_vertices;
    DevBuf d, p, o, c;
    XPBD_CUDA_TRY(d.alloc((size_t)n * sizeof(xpbd_rigid)));
    XPBD_CUDA_TRY(p.alloc((size_t)n * sizeof(xpbd_frame)));
    XPBD_CUDA_TRY(o.alloc(nc * sizeof(xpbd_constraint)));
    XPBD_CUDA_TRY(c.alloc((size_t)n * sizeof(int32_t)));
    XPBD_CUDA_TRY(cudaMemcpy(d.p, rigids, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemcpy(p.p, past, (size_t)n * sizeof(xpbd_frame), cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemset(o.p, 0, nc * sizeof(xpbd_constraint)));
    ground_kernel<<<BLOCKS(n)>>>(d.as<double>(), p.as<double>(), n, dev.blob, 1.0f, o.as<double>(), c.as<int32_t>());
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaMemcpy(out, o.p, nc * sizeof(xpbd_constraint), cudaMemcpyDeviceToHost));
    XPBD_CUDA_TRY(cudaMemcpy(counts, c.p, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost));
    return XPBD_OK;
}

int xpbd_solve_batch(xpbd_rigid* rigids, int64_t n, const xpbd_constraint* constraints, const int32_t* counts,
                     int32_t stride, double dt, double compliance) {
    if (!constraints || !counts || stride <= 0) return fail(XPBD_ERR_INVALID, "bad constraints/counts/stride");
    for (int64_t i = 0; i < n; ++i)
        if (counts[i] < 0 || counts[i] > stride) return fail(XPBD_ERR_INVALID, "counts[%lld] outside [0, stride]", (long long)i);
    return run_on_rigids(rigids, n, [&](double* d) -> int {
        DevBuf cs, ct;
        XPBD_CUDA_TRY(cs.alloc((size_t)n * stride * sizeof(xpbd_constraint)));
        XPBD_CUDA_TRY(ct.alloc((size_t)n * sizeof(int32_t)));
        XPBD_CUDA_TRY(cudaMemcpy(cs.p, constraints, (size_t)n * stride * sizeof(xpbd_constraint), cudaMemcpyHostToDevice));
        XPBD_CUDA_TRY(cudaMemcpy(ct.p, counts, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
        solve_kernel<<<BLOCKS(n)>>>(d, n, cs.as<double>(), ct.as<int32_t>(), stride, (float)dt, (float)compliance);
        XPBD_CUDA_TRY(cudaDeviceSynchronize());
        return 0;
    });
}

static int pair_query(const xpbd_frame* fa, const xpbd_frame* fb, int64_t n, const xpbd_polytope* ha,
                      const xpbd_polytope* hb, int mode, double* distance, int32_t* index, xpbd_constraint* cons,
                      int32_t* counts, int32_t* feature) {
    if (!fa || !fb || !ha || !hb || n < 0) return fail(XPBD_ERR_INVALID, "bad argument");
    if (int rc = need_device()) return rc;
    DeviceHulls dev;
    if (int rc = upload_hulls(ha, hb, dev, mode != 0)) return rc;
    struct Release { DeviceHulls& d; ~Release() { d.release(); } } rel{dev};
    if (n == 0) return XPBD_OK;
    DevBuf a, b, od, oi, oc, on, of;
    const int per = mode == 1 ? 2 : 1;
    XPBD_CUDA_TRY(a.alloc((size_t)n * 56)); XPBD_CUDA_TRY(b.alloc((size_t)n * 56));
    XPBD_CUDA_TRY(od.alloc((size_t)n * 8)); XPBD_CUDA_TRY(oi.alloc((size_t)n * 4 * per));
    XPBD_CUDA_TRY(oc.alloc((size_t)n * kMaxManifold * 64)); XPBD_CUDA_TRY(on.alloc((size_t)n * 4)); XPBD_CUDA_TRY(of.alloc((size_t)n * 4));
    XPBD_CUDA_TRY(cudaMemcpy(a.p, fa, (size_t)n * 56, cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemcpy(b.p, fb, (size_t)n * 56, cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemset(oc.p, 0, (size_t)n * kMaxManifold * 64));
    pair_query_kernel<<<(unsigned)n, 32>>>(a.as<double>(), b.as<double>(), n, dev.blob, mode, od.as<double>(), oi.as<int32_t>(),
                                           oc.as<double>(), on.as<int32_t>(), of.as<int32_t>());
    XPBD_CUDA_TRY(cudaGetLastError());
    if (distance) XPBD_CUDA_TRY(cudaMemcpy(distance, od.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
    if (index) XPBD_CUDA_TRY(cudaMemcpy(index, oi.p, (size_t)n * 4 * per, cudaMemcpyDeviceToHost));
    if (cons) XPBD_CUDA_TRY(cudaMemcpy(cons, oc.p, (size_t)n * kMaxManifold * 64, cudaMemcpyDeviceToHost));
    if (counts) XPBD_CUDA_TRY(cudaMemcpy(counts, on.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    if (feature) XPBD_CUDA_TRY(cudaMemcpy(feature, of.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
    XPBD_CUDA_TRY(cudaDeviceSynchronize());
    return XPBD_OK;
}

int xpbd_face_axes_separation_batch(const xpbd_frame* fa, const xpbd_frame* fb, int64_t n, const xpbd_polytope* ha,
                                    const xpbd_polytope* hb, double* distance, int32_t* face) {
    if (!distance || !face) return fail(XPBD_ERR_INVALID, "null output");
    return pair_query(fa, fb, n, ha, hb, 0, distance, face, nullptr, nullptr, nullptr);
}
int xpbd_edge_axes_separation_batch(const xpbd_frame* fa, const xpbd_frame* fb, int64_t n, const xpbd_polytope* ha,
                                    const xpbd_polytope* hb, double* distance, int32_t* edges) {
    if (!distance || !edges) return fail(XPBD_ERR_INVALID, "null output");
    return pair_query(fa, fb, n, ha, hb, 1, distance, edges, nullptr, nullptr, nullptr);
}
int xpbd_sat_manifold_batch(const xpbd_frame* fa, const xpbd_frame* fb, int64_t n, const xpbd_polytope* ha,
                            const xpbd_polytope* hb, xpbd_constraint* out, int32_t* counts, int32_t* feature) {
    if (!out || !counts || !feature) return fail(XPBD_ERR_INVALID, "null output");
    return pair_query(fa, fb, n, ha, hb, 2, nullptr, nullptr, out, counts, feature);
}

int xpbd_fp32_peak_microbench(int32_t device, int32_t use_fma, double* tflops) {
    if (!tflops) return fail(XPBD_ERR_INVALID, "tflops is null");
    if (int rc = need_device()) return rc;
    XPBD_CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    XPBD_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
    DevBuf o;
    XPBD_CUDA_TRY(o.alloc((size_t)blocks * threads * 4));
    cudaEvent_t e0, e1;
    XPBD_CUDA_TRY(cudaEventCreate(&e0)); XPBD_CUDA_TRY(cudaEventCreate(&e1));
    fp32_peak_kernel<<<blocks, threads>>>(o.as<float>(), iters, use_fma);  // warm-up
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0);
        fp32_peak_kernel<<<blocks, threads>>>(o.as<float>(), iters, use_fma);
        cudaEventRecord(e1);
        XPBD_CUDA_TRY(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    const double flops = (double)blocks * threads * iters * 8 * 2;  // FMA = 2 flops; FMUL + FADD = 2 flops
    *tflops = flops / (best * 1e-3) / 1e12;
    return XPBD_OK;
}

}  // extern "C"

// ---- xpbd_solve_batch without any allocation: working copies in a file-scope device array -----------------------------
// No counterpart in the reference; added for callers that must not call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel (solve_kernel) and same
// arithmetic as xpbd_solve_batch; only where the device-side working copies live differs.
//
// g_static_scratch is part of the library's device image: every GPU that loads the library holds these 3.2 MiB until the
// process exits.  In HOST code the name below is only a host-side stand-in for the symbol, not a device address; the
// address the kernel gets comes from cudaGetSymbolAddress, asked once per device with that device current.
namespace {

constexpr size_t kStaticRigidBytes = (size_t)XPBD_STATIC_MAX_BODIES * sizeof(xpbd_rigid);             // 4096 x 304
constexpr size_t kStaticConsBytes = (size_t)XPBD_STATIC_MAX_CONSTRAINTS * sizeof(xpbd_constraint);    // 32768 x 64
constexpr size_t kStaticCountBytes = (size_t)XPBD_STATIC_MAX_BODIES * sizeof(int32_t);
constexpr size_t kStaticConsOffset = kStaticRigidBytes;
constexpr size_t kStaticCountOffset = kStaticConsOffset + kStaticConsBytes;
constexpr size_t kStaticScratchBytes = kStaticCountOffset + kStaticCountBytes;
static_assert(sizeof(xpbd_rigid) == 304 && sizeof(xpbd_constraint) == 64, "record sizes of include/xpbd_b200.h");
static_assert(kStaticConsOffset % 256 == 0 && kStaticCountOffset % 256 == 0, "sub-buffers stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mutex;                                  // set-up only; the calls themselves are not serialised
unsigned char* g_static_ptr[kStaticMaxDevices] = {};        // device address of g_static_scratch on each device

// Device address of the scratch array on the CURRENT device (looked up once per device).
int static_scratch(unsigned char** out) {
    int dev = 0;
    XPBD_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kStaticMaxDevices) return fail(XPBD_ERR_CAPACITY, "device ordinal %d outside [0, %d)", dev, kStaticMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        XPBD_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail(XPBD_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_static_ptr[dev];
    return XPBD_OK;
}

}  // namespace

extern "C" int xpbd_solve_batch_static(xpbd_rigid* rigids, int64_t n, const xpbd_constraint* constraints, const int32_t* counts,
                                       int32_t stride, double dt, double compliance) {
    if (!rigids || n < 0) return fail(XPBD_ERR_INVALID, "bad rigids/n");
    if (!constraints || !counts || stride <= 0) return fail(XPBD_ERR_INVALID, "bad constraints/counts/stride");
    if (n > XPBD_STATIC_MAX_BODIES)
        return fail(XPBD_ERR_CAPACITY, "xpbd_solve_batch_static: n = %lld bodies, the static scratch holds at most %d "
                    "(use xpbd_solve_batch, which allocates)", (long long)n, XPBD_STATIC_MAX_BODIES);
    if (n * (int64_t)stride > XPBD_STATIC_MAX_CONSTRAINTS)
        return fail(XPBD_ERR_CAPACITY, "xpbd_solve_batch_static: n * stride = %lld constraint slots, the static scratch holds at "
                    "most %d (use xpbd_solve_batch, which allocates)", (long long)(n * (int64_t)stride), XPBD_STATIC_MAX_CONSTRAINTS);
    for (int64_t i = 0; i < n; ++i)
        if (counts[i] < 0 || counts[i] > stride) return fail(XPBD_ERR_INVALID, "counts[%lld] outside [0, stride]", (long long)i);
    if (int rc = need_device()) return rc;
    if (n == 0) return XPBD_OK;
    unsigned char* base = nullptr;
    if (int rc = static_scratch(&base)) return rc;
    double* d_rigids = reinterpret_cast<double*>(base);
    double* d_cons = reinterpret_cast<double*>(base + kStaticConsOffset);
    int32_t* d_counts = reinterpret_cast<int32_t*>(base + kStaticCountOffset);
    // The array keeps the previous call's contents.  Nothing in it has to start at zero: the kernel reads records [0, n),
    // constraint slots [i * stride, i * stride + counts[i]) and counts [0, n) only, and all of those are overwritten by
    // this call's copies before the launch (same default stream, in order).
    XPBD_CUDA_TRY(cudaMemcpy(d_rigids, rigids, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemcpy(d_cons, constraints, (size_t)n * stride * sizeof(xpbd_constraint), cudaMemcpyHostToDevice));
    XPBD_CUDA_TRY(cudaMemcpy(d_counts, counts, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
    solve_kernel<<<BLOCKS(n)>>>(d_rigids, n, d_cons, d_counts, stride, (float)dt, (float)compliance);
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaMemcpy(rigids, d_rigids, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyDeviceToHost));
    return XPBD_OK;
}
