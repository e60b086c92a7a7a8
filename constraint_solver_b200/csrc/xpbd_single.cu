// Fused single-body XPBD step: solver::step (src/solver.rs:3-17) for n independent bodies,
// all substeps of all frames in ONE launch with the body state held in registers.
//
// Mapping: one thread per body (every body is its own world on this path: world.rs:41-42 steps
// a and b independently, each only against the ground plane).  HBM is touched once per step
// call: 38 floats read + 13 written per body, coalesced float4 (xpbd_state.cuh).  The hull
// table is staged into shared memory by one TMA bulk copy per block (xpbd_tma.cuh).
//
// Gauss-Seidel order = hull vertex order (src/collision.rs:16).  The reference materialises a
// Vec<Constraint> before the sweep, with contact points computed from the post-integrate pose
// and NOT refreshed while the sweep moves the body (src/solver.rs:12-13).  Here the
// post-integrate frame is kept in registers and each constraint is generated right before it
// is applied — from that saved frame, so the values are identical to the reference's.
#include <cstdlib>

#include "xpbd_hull.cuh"
#include "xpbd_internal.hpp"
#include "xpbd_rigid.cuh"
#include "xpbd_tma.cuh"

namespace xpbd {

struct SingleStepParams {
    BodyArrays bodies;
    int64_t n_bodies;
    const int32_t* hull_ids;  // per body, or nullptr = hull 0
    const uint32_t* hull_blob;
    uint32_t hull_blob_bytes;  // multiple of 16
    float frame_dt;
    int32_t substeps;
    int32_t n_frames;
    float compliance;
    float friction;
    unsigned long long* constraint_counter;  // += constraints applied
    unsigned int* tma_fallback_flag;
    const int32_t* order;    // thread -> body (nullptr: identity), see regroup_bodies below
    uint8_t* contact_count;  // per body: contacts of the call's last substep (nullptr: not wanted)
};

// One ground constraint generated from the saved post-integrate frame and applied immediately.
// `p` is the (already tested, p.z < 0) world-space vertex position.
XD void ground_contact(BodyDyn& d, const BodyConst& k, const Fr& frame_inv, const Fr& past, V3 p, float compliance,
                       float friction) {
    // collision::ground (src/collision.rs:22-31)
    V3 target = v3(p.x, p.y, 0.0f);
    V3 correction = target - p;
    V3 local = frame_inv * p;  // Frame::delta (src/frame.rs:40-44)
    V3 delta = p - (past * local);
    V3 delta_t = delta - project_on(delta, correction);
    V3 c1 = target - friction * delta_t;
    // solver::solve body (src/solver.rs:22-26), Constraint methods (src/constraint.rs:13-37)
    V3 diff = c1 - p;
    float dist = magnitude(diff);                  // current_distance(); distance == 0
    V3 dir = diff * frcp(dist);              // direction() = normalize
    float w = inverse_resistance(d.position, d.rotation, k.inverse_mass, k.inverse_inertia, k.center_of_mass, p, dir);
    float lambda = fdiv(fsub(dist, 0.0f), fadd(w, compliance));
    rigid_apply_impulse(d.position, d.rotation, k.inverse_mass, k.inverse_inertia, k.center_of_mass, lambda * dir, p);
}

__global__ void __launch_bounds__(128) step_single_kernel(SingleStepParams P) {
    extern __shared__ __align__(16) uint32_t smem_blob[];
    __shared__ __align__(8) uint64_t bar;
    stage_bulk(smem_blob, P.hull_blob, P.hull_blob_bytes, &bar, P.tma_fallback_flag);

    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int applied = 0;
    bool bad = false;  // this body's state is NaN / Inf after the call (xpbd_step_stats.nan_worlds; a world = one body here)
    if (t < P.n_bodies) {
        const int64_t g = P.order ? P.order[t] : t;
        unsigned int last = 0;
        HullTable hulls{smem_blob};
        const HullDesc hd = hulls.desc(P.hull_ids ? P.hull_ids[g] : 0);

        BodyDyn d;
        BodyConst k;
        load_dyn(P.bodies, g, d);
        load_const(P.bodies, g, k);
        const StepConsts sc = make_step_consts(P.frame_dt, P.substeps, P.compliance);
        const M3 dtI = sc.dt * k.inverse_inertia;

        const int total = P.n_frames * P.substeps;
        for (int s = 0; s < total; ++s) {
            const V3 past_position = d.position;
            const Q4 past_rotation = d.rotation;
            const Fr past = rigid_frame(d.position, d.rotation, k.center_of_mass);
            rigid_integrate(d, k, dtI, sc);
            const Fr frame = rigid_frame(d.position, d.rotation, k.center_of_mass);
            const Fr frame_inv = frame_inverse(frame);
            // Pass 1: which vertices are below the ground (src/collision.rs:17-20), as a bit mask per lane.
            // Pass 2: every lane walks ITS OWN set bits in ascending vertex order (= the reference's
            // Gauss-Seidel order), so a warp iterates max-over-lanes(contacts) times instead of once per
            // vertex slot that any lane penetrates: with mixed orientations that is ~2x fewer divergent passes.
            for (int base = 0; base < hd.n_verts; base += 32) {
                const int cnt = min(32, hd.n_verts - base);
                unsigned mask = 0;
                for (int v = 0; v < cnt; ++v) {
                    const V3 p = frame * hulls.vertex(hd, base + v);
                    if (!(p.z >= 0.0f)) mask |= 1u << v;
                }
                applied += __popc(mask);
                if (s == total - 1) last += __popc(mask);
                while (mask) {
                    const int v = __ffs(mask) - 1;
                    mask &= mask - 1;
                    const V3 p = frame * hulls.vertex(hd, base + v);
                    ground_contact(d, k, frame_inv, past, p, sc.compliance, P.friction);
                }
            }
            rigid_derive(d, past_position, past_rotation, sc.dt);
        }
        store_dyn(P.bodies, g, d);
        bad = dyn_nonfinite(d);
        if (P.contact_count) P.contact_count[g] = (uint8_t)(last < 15u ? last : 15u);
    }
    // one atomic per warp
    for (int o = 16; o > 0; o >>= 1) applied += __shfl_down_sync(0xffffffffu, applied, o);
    if ((threadIdx.x & 31) == 0 && applied) atomicAdd(P.constraint_counter, (unsigned long long)applied);
    const unsigned nbad = __popc(__ballot_sync(0xffffffffu, bad));
    if ((threadIdx.x & 31) == 0 && nbad) atomicAdd(P.constraint_counter + 4, (unsigned long long)nbad);
}

// ---- regrouping: bodies with the same number of ground contacts share warps --------------------------------------------
// The step kernel is FP32-issue bound and a warp iterates max-over-lanes(contact count) times per substep: with bodies in
// upload order, a warp mixes airborne, edge-resting and face-resting bodies and 14.6 of 32 lanes are active per
// instruction (ncu).  Bodies are independent, so which thread steps which body is free: after every step call a counting
// sort by the contact count of the call's last substep (16 bins) writes the thread -> body permutation of the next call.
// Loads and stores become 16-byte gathers (the path uses ~1 % of the HBM bandwidth).  Results cannot change.
constexpr int kOrderBins = 16;
__global__ void order_hist_kernel(const uint8_t* __restrict__ cnt, int64_t n, int32_t* bins) {
    __shared__ int h[kOrderBins];
    if (threadIdx.x < kOrderBins) h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n) atomicAdd(&h[cnt[g]], 1);
    __syncthreads();
    if (threadIdx.x < kOrderBins && h[threadIdx.x]) atomicAdd(&bins[threadIdx.x], h[threadIdx.x]);
}
__global__ void order_scan_kernel(int32_t* bins) {  // bins[0..16) counts -> bins[16..32) cursors (exclusive prefix)
    int acc = 0;
    for (int k = 0; k < kOrderBins; ++k) { bins[kOrderBins + k] = acc; acc += bins[k]; }
}
__global__ void order_scatter_kernel(const uint8_t* __restrict__ cnt, int64_t n, int32_t* bins, int32_t* __restrict__ order) {
    __shared__ int h[kOrderBins], base[kOrderBins];
    if (threadIdx.x < kOrderBins) h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int bin = 0, rank = 0;
    if (g < n) { bin = cnt[g]; rank = atomicAdd(&h[bin], 1); }
    __syncthreads();
    if (threadIdx.x < kOrderBins && h[threadIdx.x]) base[threadIdx.x] = atomicAdd(&bins[kOrderBins + threadIdx.x], h[threadIdx.x]);
    __syncthreads();
    if (g < n) order[base[bin] + rank] = (int32_t)g;
}

static int regroup_bodies(xpbd_batch* b) {
    const int64_t n = b->n_bodies;
    const unsigned grid = (unsigned)((n + 255) / 256);
    XPBD_CUDA_TRY(cudaMemsetAsync(b->d_order_bins, 0, 2 * kOrderBins * sizeof(int32_t), b->stream));
    order_hist_kernel<<<grid, 256, 0, b->stream>>>(b->d_contact_count, n, b->d_order_bins);
    order_scan_kernel<<<1, 1, 0, b->stream>>>(b->d_order_bins);
    order_scatter_kernel<<<grid, 256, 0, b->stream>>>(b->d_contact_count, n, b->d_order_bins, b->d_order[b->order_cur ^ 1]);
    XPBD_CUDA_TRY(cudaGetLastError());
    b->order_cur ^= 1;
    b->order_valid = true;
    return XPBD_OK;
}

int step_single(xpbd_batch* b, float frame_dt, int substeps, int n_frames) {
    const uint32_t blob_bytes = b->dev_hulls.bytes;
    if (blob_bytes > 200u * 1024u)
        return fail(XPBD_ERR_CAPACITY, "hull table of %u bytes does not fit in shared memory", blob_bytes);
    if (blob_bytes > 48u * 1024u)  // per-device attribute; cheap, so set on every call that needs it
        XPBD_CUDA_TRY(cudaFuncSetAttribute(step_single_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)blob_bytes));
    SingleStepParams P;
    P.bodies = b->arrays;
    P.n_bodies = b->n_bodies;
    P.hull_ids = b->d_hull_ids;
    P.hull_blob = b->dev_hulls.blob;
    P.hull_blob_bytes = blob_bytes;
    P.frame_dt = frame_dt;
    P.substeps = substeps;
    P.n_frames = n_frames;
    P.compliance = (float)b->desc.compliance;
    P.friction = (float)b->desc.friction;
    P.constraint_counter = b->d_counters;
    P.tma_fallback_flag = reinterpret_cast<unsigned int*>(b->d_counters + 2);
    // regrouping pays once there are enough bodies to fill the device several times over; XPBD_SINGLE_REGROUP=0/1 overrides
    bool regroup = b->n_bodies >= (1 << 16) && b->n_bodies < (int64_t)1 << 31;
    if (const char* e = std::getenv("XPBD_SINGLE_REGROUP")) regroup = e[0] != '0' && b->n_bodies < (int64_t)1 << 31;
    if (regroup && !b->d_contact_count) {
        XPBD_CUDA_TRY(cudaMalloc(&b->d_order[0], (size_t)b->n_bodies * sizeof(int32_t)));
        XPBD_CUDA_TRY(cudaMalloc(&b->d_order[1], (size_t)b->n_bodies * sizeof(int32_t)));
        XPBD_CUDA_TRY(cudaMalloc(&b->d_contact_count, (size_t)b->n_bodies));
        XPBD_CUDA_TRY(cudaMalloc(&b->d_order_bins, 2 * kOrderBins * sizeof(int32_t)));
    }
    P.order = regroup && b->order_valid ? b->d_order[b->order_cur] : nullptr;
    P.contact_count = regroup ? b->d_contact_count : nullptr;
    const int block = 128;
    const int64_t grid = (P.n_bodies + block - 1) / block;
    step_single_kernel<<<(unsigned)grid, block, blob_bytes, b->stream>>>(P);
    XPBD_CUDA_TRY(cudaGetLastError());
    b->last.kernel_launches = 1;
    if (regroup) {
        if (int rc = regroup_bodies(b)) return rc;
        b->last.kernel_launches += 3;
    }
    return XPBD_OK;
}

}  // namespace xpbd
