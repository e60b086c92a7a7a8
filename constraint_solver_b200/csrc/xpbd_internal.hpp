// Internal host-side declarations shared by the translation units of libxpbd_b200.so.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/xpbd_b200.h"
#include "xpbd_hull.cuh"
#include "xpbd_state.cuh"

namespace xpbd {

// ---- error plumbing (thread-local text behind xpbd_last_error) ---------------------------
void set_error(const char* fmt, ...);
int fail(int status, const char* fmt, ...);

#define XPBD_CUDA_TRY(expr)                                                                         \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess)                                                                      \
            return ::xpbd::fail(XPBD_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                                __FILE__, __LINE__);                                                \
    } while (0)

// ---- host copy of a registered hull ---------------------------------------------------------
struct HostHull {
    std::vector<double> vertices;  // n x 3
    std::vector<int32_t> edges;    // m x 2
    std::vector<int32_t> face_offsets, face_indices;
    double centroid[3];
    double radius;  // max |v - centroid| in f64
    xpbd_polytope view() const {
        xpbd_polytope p;
        p.vertices = vertices.data(); p.n_vertices = (int32_t)(vertices.size() / 3);
        p.edges = edges.data(); p.n_edges = (int32_t)(edges.size() / 2);
        p.face_offsets = face_offsets.data(); p.face_indices = face_indices.data();
        p.n_faces = (int32_t)face_offsets.size() - 1;
        p.centroid[0] = centroid[0]; p.centroid[1] = centroid[1]; p.centroid[2] = centroid[2];
        return p;
    }
};
HostHull copy_hull(const xpbd_polytope* h);
// Packs hulls into the device blob format of xpbd_hull.cuh (planes left zero: hull_planes_kernel fills them).
std::vector<uint32_t> pack_hull_blob(const std::vector<HostHull>& hulls);

// Device hull table owned by a batch or a one-shot query.
struct DeviceHulls {
    uint32_t* blob = nullptr;
    uint32_t bytes = 0;
    int n_hulls = 0;
    int max_verts = 0, max_faces = 0, max_edges = 0, max_face_verts = 0;
    int upload(const std::vector<HostHull>& hulls, cudaStream_t st);  // returns xpbd_status
    void release();
};

// ---- kernel launchers (defined in the .cu files) --------------------------------------------
void launch_aos64_to_soa(const double* aos, const BodyArrays& s, int64_t first, int64_t n, cudaStream_t st);
void launch_soa_to_aos64(const BodyArrays& s, double* aos, int64_t first, int64_t n, cudaStream_t st);
void launch_soa_dyn_to_aos64(const BodyArrays& s, double* aos, int64_t first, int64_t n, cudaStream_t st);
void launch_broadcast_world(const BodyArrays& s, int bodies_per_world, int64_t n_bodies, cudaStream_t st);
void launch_dyn_pack(const BodyArrays& s, float* out, int64_t first, int64_t n, cudaStream_t st);
void launch_dyn_unpack(const float* in, const BodyArrays& s, int64_t first, int64_t n, cudaStream_t st);
void launch_hull_planes(uint32_t* blob, int n_hulls, cudaStream_t st);
void launch_count_nan_worlds(const BodyArrays& s, int64_t n_worlds, int B, unsigned long long* counter, cudaStream_t st);

constexpr int kBatchCounters = 8;
struct SingleStepParams;
struct WorldStepParams;

// Device joint record (fp32): anchors in body space (or world space when body_b < 0).
struct DeviceJoint {
    int32_t body_a, body_b;
    float ax, ay, az;
    float bx, by, bz;
    float distance;
    float pad[3];
};
static_assert(sizeof(DeviceJoint) == 48, "DeviceJoint is 48 bytes");

}  // namespace xpbd

// The opaque handle.
struct xpbd_batch {
    xpbd_batch_desc desc{};
    int64_t n_bodies = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    xpbd::BodyArrays arrays{};
    void* state_alloc = nullptr;
    int32_t* d_hull_ids = nullptr;  // nullptr: all hull 0
    std::vector<xpbd::HostHull> hulls;
    xpbd::DeviceHulls dev_hulls;
    bool hulls_dirty = true;
    xpbd::DeviceJoint* d_joints = nullptr;
    // staging for AoS <-> SoA conversion (two slots for copy/convert overlap)
    double* d_stage[2] = {nullptr, nullptr};
    cudaEvent_t stage_free[2] = {nullptr, nullptr};
    int64_t stage_bodies = 0;
    float* d_dyn_stage = nullptr;  // n_bodies x 13 floats, allocated on first use
    // stats
    unsigned long long* d_counters = nullptr;  // kBatchCounters: [0] constraints, [1] overflow worlds, [2] tma fallbacks, [3] candidates, [4] NaN worlds
    unsigned long long* h_counters = nullptr;  // pinned
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    cudaEvent_t ev_stage[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // staged pipeline: around the last substep's kernels
    bool stage_timed = false;
    bool stage_ms_valid = false;   // ev_stage[] were recorded by the last step call (not when the chains ran as CUDA graphs)
    xpbd_step_stats last{};
    bool stepped = false;
    void* world_scratch = nullptr;  // owned by the multi-body path
    bool fused_too_large = false;   // a world of this batch does not fit the fused kernel's shared memory: staged pipeline
    // single-body path: thread -> body permutation that groups bodies by their ground-contact count (xpbd_single.cu)
    int32_t* d_order[2] = {nullptr, nullptr};  // [cur] read by the step kernel, [cur ^ 1] written by the regrouping pass
    int order_cur = 0;
    bool order_valid = false;
    uint8_t* d_contact_count = nullptr;  // per body: ground contacts in the last substep of the last step call
    int32_t* d_order_bins = nullptr;     // 32 ints: histogram, then cursors
    int sub_batches = 0;            // xpbd_batch_set_sub_batches: 0 = automatic
    // Set by every call that enqueues work on `stream` touching the batch's device memory (uploads, downloads, hull /
    // joint tables, a new stream); cleared by a step call.  When the staged pipeline runs several world ranges on its
    // internal streams, those streams wait for `stream` only if this is set — between back-to-back step calls each
    // range just continues behind its own previous frame, so the ranges drift out of phase and kernels of different
    // stages overlap (a barrier per frame would re-align them).
    bool fork_needed = true;
    // xpbd_batch_step_host_f32: host buffers of the step call in flight (packed dynamic state, 13 floats per body)
    const float* io_in = nullptr;
    float* io_out = nullptr;
    // xpbd_batch_step_host_rigid: the same with the reference's own records (xpbd_rigid, 38 f64 per body).  d_rigid_stage
    // mirrors the caller's records on the device (n_bodies x 304 B); rigid_stage_valid = it holds every body's constants
    // as uploaded, so a download only has to refresh the 13 dynamic values.
    const xpbd_rigid* io_rigid_in = nullptr;
    xpbd_rigid* io_rigid_out = nullptr;
    double* d_rigid_stage = nullptr;
    bool rigid_stage_valid = false;
};

namespace xpbd {
int finalize_hulls(xpbd_batch* b);
int step_single(xpbd_batch* b, float frame_dt, int substeps, int n_frames);
int step_world_fused(xpbd_batch* b, float frame_dt, int substeps, int n_frames);   // xpbd_world.cu
int step_world_staged(xpbd_batch* b, float frame_dt, int substeps, int n_frames);  // xpbd_staged.cu
int step_world(xpbd_batch* b, float frame_dt, int substeps, int n_frames);         // picks one of the two
void release_world_scratch(xpbd_batch* b);
// Host-record legs of a step call for bodies [g0, g0 + n) on `st` (xpbd_batch_step_host_rigid): H2D + conversion before the
// kernels, conversion + D2H after them.  Used per world range by the staged pipeline, for the whole batch otherwise.
int rigid_io_before(xpbd_batch* b, int64_t g0, int64_t n, cudaStream_t st);
int rigid_io_after(xpbd_batch* b, int64_t g0, int64_t n, cudaStream_t st);
// true when the next multi-body step call will run the staged pipeline on several world ranges (which then move
// b->io_in / b->io_out per range, on the range streams)
bool staged_takes_host_buffers(const xpbd_batch* b);
bool world_path_is_staged(const xpbd_batch* b);  // XPBD_WORLD_PATH / batch size
// adds the per-range counters of the last staged step call (valid once ev_stop has completed) to acc[0..3]
void staged_add_range_counters(const xpbd_batch* b, unsigned long long* acc);
}  // namespace xpbd
