// C ABI, batch part: lifetime, hulls/joints, state upload/download, the step call.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>

#include "xpbd_host_math.hpp"
#include "xpbd_internal.hpp"

using namespace xpbd;

namespace {

constexpr int64_t kStageBodies = 1 << 16;  // 65 536 bodies x 304 B = 19.9 MB per staging slot

int check_batch(const xpbd_batch* b) {
    if (!b) return fail(XPBD_ERR_INVALID, "batch is null");
    return XPBD_OK;
}
int check_range(const xpbd_batch* b, int64_t first, int64_t n) {
    if (first < 0 || n < 0 || first + n > b->desc.n_worlds)
        return fail(XPBD_ERR_INVALID, "world range [%lld, %lld) outside [0, %lld)", (long long)first,
                    (long long)(first + n), (long long)b->desc.n_worlds);
    return XPBD_OK;
}
int bind_device(const xpbd_batch* b) {
    XPBD_CUDA_TRY(cudaSetDevice(b->desc.device));
    return XPBD_OK;
}
int ensure_stage(xpbd_batch* b) {
    if (b->d_stage[0]) return XPBD_OK;
    b->stage_bodies = std::min<int64_t>(kStageBodies, b->n_bodies);
    for (int s = 0; s < 2; ++s) {
        XPBD_CUDA_TRY(cudaMalloc(&b->d_stage[s], (size_t)b->stage_bodies * sizeof(xpbd_rigid)));
        XPBD_CUDA_TRY(cudaEventCreateWithFlags(&b->stage_free[s], cudaEventDisableTiming));
    }
    return XPBD_OK;
}

}  // namespace

namespace xpbd {
// Multi-body path selection.  XPBD_WORLD_PATH=fused|staged forces one; default: the staged pipeline for batches
// large enough to fill the GPU per stage, the fused single-launch kernel for small ones (latency-bound) — unless a
// world does not fit the fused kernel's shared memory (a few worlds of many bodies), which also go to the staged
// pipeline: xpbd_batch_create's bodies_per_world range is reachable for every n_worlds.
static int forced_world_path() {  // 0 = no override, 1 = fused, 2 = staged
    const char* v = getenv("XPBD_WORLD_PATH");
    if (v && !strcmp(v, "fused")) return 1;
    if (v && !strcmp(v, "staged")) return 2;
    return 0;
}
int rigid_io_before(xpbd_batch* b, int64_t g0, int64_t n, cudaStream_t st) {
    if (!b->io_rigid_in || n <= 0) return XPBD_OK;
    double* dst = b->d_rigid_stage + 38 * g0;
    XPBD_CUDA_TRY(cudaMemcpyAsync(dst, b->io_rigid_in + g0, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyHostToDevice, st));
    launch_aos64_to_soa(dst, b->arrays, g0, n, st);
    return XPBD_OK;
}
int rigid_io_after(xpbd_batch* b, int64_t g0, int64_t n, cudaStream_t st) {
    if (!b->io_rigid_out || n <= 0) return XPBD_OK;
    double* src = b->d_rigid_stage + 38 * g0;
    if (b->rigid_stage_valid) launch_soa_dyn_to_aos64(b->arrays, src, g0, n, st);  // constants stay the caller's f64 bits
    else launch_soa_to_aos64(b->arrays, src, g0, n, st);
    XPBD_CUDA_TRY(cudaMemcpyAsync(b->io_rigid_out + g0, src, (size_t)n * sizeof(xpbd_rigid), cudaMemcpyDeviceToHost, st));
    return XPBD_OK;
}
bool world_path_is_staged(const xpbd_batch* b) {
    const int forced = forced_world_path();
    if (forced) return forced == 2;
    return b->desc.n_worlds >= 4096 || b->fused_too_large;
}
int step_world(xpbd_batch* b, float frame_dt, int substeps, int n_frames) {
    if (world_path_is_staged(b)) return step_world_staged(b, frame_dt, substeps, n_frames);
    const int rc = step_world_fused(b, frame_dt, substeps, n_frames);
    if (rc == XPBD_ERR_CAPACITY && !forced_world_path()) {
        // the fused kernel's capacity checks run before anything is enqueued: nothing to undo
        b->fused_too_large = true;
        return step_world_staged(b, frame_dt, substeps, n_frames);
    }
    return rc;
}
}  // namespace xpbd

extern "C" {

int xpbd_batch_create(const xpbd_batch_desc* desc, xpbd_batch** out) {
    if (!desc || !out) return fail(XPBD_ERR_INVALID, "null argument");
    *out = nullptr;
    if (desc->n_worlds <= 0) return fail(XPBD_ERR_INVALID, "n_worlds must be positive");
    if (desc->bodies_per_world <= 0 || desc->bodies_per_world > 1024)
        return fail(XPBD_ERR_INVALID, "bodies_per_world must be in [1, 1024]");
    if (desc->joints_per_world < 0) return fail(XPBD_ERR_INVALID, "joints_per_world must be >= 0");
    if (!(desc->compliance >= 0.0) || !std::isfinite(desc->friction))
        return fail(XPBD_ERR_INVALID, "compliance must be >= 0 and friction finite");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(XPBD_ERR_CUDA, "no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (desc->device < 0 || desc->device >= ndev) return fail(XPBD_ERR_INVALID, "device %d out of range", desc->device);

    xpbd_batch* b = new (std::nothrow) xpbd_batch();
    if (!b) return fail(XPBD_ERR_INVALID, "out of host memory");
    b->desc = *desc;
    b->n_bodies = desc->n_worlds * desc->bodies_per_world;
    auto bail = [&](int rc) { xpbd_batch_destroy(b); return rc; };
#define TRY_OR_BAIL(expr)                                                                              \
    do {                                                                                               \
        cudaError_t _e = (expr);                                                                       \
        if (_e != cudaSuccess) {                                                                       \
            set_error("%s failed: %s", #expr, cudaGetErrorString(_e));                                 \
            return bail(XPBD_ERR_CUDA);                                                                \
        }                                                                                              \
    } while (0)
    TRY_OR_BAIL(cudaSetDevice(desc->device));
    TRY_OR_BAIL(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
    b->own_stream = true;
    // One allocation, carved into the SoA arrays; every array 256-byte aligned.
    const size_t n = (size_t)b->n_bodies;
    auto pad = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t f4 = pad(n * sizeof(float4)), f1 = pad(n * sizeof(float));
    const size_t total = 9 * f4 + 2 * f1;
    TRY_OR_BAIL(cudaMalloc(&b->state_alloc, total));
    TRY_OR_BAIL(cudaMemsetAsync(b->state_alloc, 0, total, b->stream));
    char* p = static_cast<char*>(b->state_alloc);
    auto take4 = [&]() { float4* r = reinterpret_cast<float4*>(p); p += f4; return r; };
    auto take1 = [&]() { float* r = reinterpret_cast<float*>(p); p += f1; return r; };
    b->arrays.dynA = take4(); b->arrays.dynB = take4(); b->arrays.dynC = take4(); b->arrays.dynD = take1();
    b->arrays.k0 = take4(); b->arrays.k1 = take4(); b->arrays.k2 = take4(); b->arrays.k3 = take4();
    b->arrays.k4 = take4(); b->arrays.k5 = take4(); b->arrays.k6 = take1();
    TRY_OR_BAIL(cudaMalloc(&b->d_counters, kBatchCounters * sizeof(unsigned long long)));
    TRY_OR_BAIL(cudaMemsetAsync(b->d_counters, 0, kBatchCounters * sizeof(unsigned long long), b->stream));
    TRY_OR_BAIL(cudaMallocHost(&b->h_counters, kBatchCounters * sizeof(unsigned long long)));
    std::memset(b->h_counters, 0, kBatchCounters * sizeof(unsigned long long));
    TRY_OR_BAIL(cudaEventCreate(&b->ev_start));
    TRY_OR_BAIL(cudaEventCreate(&b->ev_stop));
    for (int k = 0; k < 5; ++k) TRY_OR_BAIL(cudaEventCreate(&b->ev_stage[k]));
    if (desc->joints_per_world > 0) {
        const size_t jb = (size_t)desc->n_worlds * desc->joints_per_world * sizeof(DeviceJoint);
        TRY_OR_BAIL(cudaMalloc(&b->d_joints, jb));
        TRY_OR_BAIL(cudaMemsetAsync(b->d_joints, 0xff, jb, b->stream));  // body_a = -1: inactive
    }
    TRY_OR_BAIL(cudaStreamSynchronize(b->stream));
#undef TRY_OR_BAIL
    *out = b;
    return XPBD_OK;
}

int xpbd_batch_destroy(xpbd_batch* b) {
    if (!b) return XPBD_OK;
    cudaSetDevice(b->desc.device);
    if (b->stream) cudaStreamSynchronize(b->stream);
    release_world_scratch(b);
    b->dev_hulls.release();
    if (b->state_alloc) cudaFree(b->state_alloc);
    if (b->d_hull_ids) cudaFree(b->d_hull_ids);
    if (b->d_joints) cudaFree(b->d_joints);
    for (int s = 0; s < 2; ++s) {
        if (b->d_stage[s]) cudaFree(b->d_stage[s]);
        if (b->stage_free[s]) cudaEventDestroy(b->stage_free[s]);
    }
    if (b->d_dyn_stage) cudaFree(b->d_dyn_stage);
    if (b->d_rigid_stage) cudaFree(b->d_rigid_stage);
    for (int s = 0; s < 2; ++s)
        if (b->d_order[s]) cudaFree(b->d_order[s]);
    if (b->d_contact_count) cudaFree(b->d_contact_count);
    if (b->d_order_bins) cudaFree(b->d_order_bins);
    if (b->d_counters) cudaFree(b->d_counters);
    if (b->h_counters) cudaFreeHost(b->h_counters);
    if (b->ev_start) cudaEventDestroy(b->ev_start);
    if (b->ev_stop) cudaEventDestroy(b->ev_stop);
    for (int k = 0; k < 5; ++k) if (b->ev_stage[k]) cudaEventDestroy(b->ev_stage[k]);
    if (b->own_stream && b->stream) cudaStreamDestroy(b->stream);
    delete b;
    return XPBD_OK;
}

int xpbd_batch_add_hull(xpbd_batch* b, const xpbd_polytope* hull, int32_t* hull_id) {
    if (int rc = check_batch(b)) return rc;
    if (const char* why = xpbd_host::validate_hull(hull)) return fail(XPBD_ERR_INVALID, "hull: %s", why);
    if (hull->n_vertices > 256 || hull->n_faces > 256 || hull->n_edges > 1024)
        return fail(XPBD_ERR_CAPACITY, "hull too large (max 256 vertices, 256 faces, 1024 edges)");
    b->hulls.push_back(copy_hull(hull));
    b->hulls_dirty = true;
    if (hull_id) *hull_id = (int32_t)b->hulls.size() - 1;
    return XPBD_OK;
}

int xpbd_batch_set_hulls(xpbd_batch* b, int64_t first_world, int64_t n_worlds, const int32_t* ids) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = check_range(b, first_world, n_worlds)) return rc;
    if (int rc = bind_device(b)) return rc;
    const int64_t B = b->desc.bodies_per_world, n = n_worlds * B;
    if (!ids && !b->d_hull_ids) return XPBD_OK;  // already all hull 0
    if (!b->d_hull_ids) {
        XPBD_CUDA_TRY(cudaMalloc(&b->d_hull_ids, (size_t)b->n_bodies * sizeof(int32_t)));
        XPBD_CUDA_TRY(cudaMemsetAsync(b->d_hull_ids, 0, (size_t)b->n_bodies * sizeof(int32_t), b->stream));
    }
    if (!ids) {
        XPBD_CUDA_TRY(cudaMemsetAsync(b->d_hull_ids + first_world * B, 0, (size_t)n * sizeof(int32_t), b->stream));
    } else {
        for (int64_t i = 0; i < n; ++i)
            if (ids[i] < 0 || ids[i] >= (int32_t)b->hulls.size())
                return fail(XPBD_ERR_INVALID, "hull id %d at index %lld is not registered", ids[i], (long long)i);
        XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_hull_ids + first_world * B, ids, (size_t)n * sizeof(int32_t),
                                      cudaMemcpyHostToDevice, b->stream));
    }
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

int xpbd_batch_set_joints(xpbd_batch* b, int64_t first_world, int64_t n_worlds, const xpbd_joint* joints) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = check_range(b, first_world, n_worlds)) return rc;
    if (!joints) return fail(XPBD_ERR_INVALID, "joints is null");
    const int J = b->desc.joints_per_world;
    if (J == 0) return fail(XPBD_ERR_STATE, "batch was created with joints_per_world = 0");
    if (int rc = bind_device(b)) return rc;
    const int64_t n = n_worlds * J;
    std::vector<DeviceJoint> host((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const xpbd_joint& j = joints[i];
        if (j.body_a >= b->desc.bodies_per_world || j.body_b >= b->desc.bodies_per_world)
            return fail(XPBD_ERR_INVALID, "joint %lld: body index out of range", (long long)i);
        if (j.body_a >= 0 && j.body_a == j.body_b)
            return fail(XPBD_ERR_INVALID, "joint %lld: body_a == body_b", (long long)i);
        DeviceJoint& d = host[(size_t)i];
        std::memset(&d, 0, sizeof d);
        d.body_a = j.body_a; d.body_b = j.body_b;   // body_a < 0: inactive slot
        d.ax = (float)j.anchor_a[0]; d.ay = (float)j.anchor_a[1]; d.az = (float)j.anchor_a[2];
        d.bx = (float)j.anchor_b[0]; d.by = (float)j.anchor_b[1]; d.bz = (float)j.anchor_b[2];
        d.distance = (float)j.distance;
    }
    XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_joints + first_world * J, host.data(), (size_t)n * sizeof(DeviceJoint),
                                  cudaMemcpyHostToDevice, b->stream));
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

int xpbd_batch_set_stream(xpbd_batch* b, void* cuda_stream) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = bind_device(b)) return rc;
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    if (b->own_stream) cudaStreamDestroy(b->stream);
    b->stream = static_cast<cudaStream_t>(cuda_stream);
    b->own_stream = false;
    return XPBD_OK;
}

int xpbd_batch_set_sub_batches(xpbd_batch* b, int32_t count) {
    if (int rc = check_batch(b)) return rc;
    if (count < 0 || count > 8) return fail(XPBD_ERR_INVALID, "sub-batch count must be 0 (automatic) .. 8");
    b->sub_batches = count;
    b->fork_needed = true;
    return XPBD_OK;
}

int xpbd_batch_upload(xpbd_batch* b, int64_t first_world, int64_t n_worlds, const xpbd_rigid* rigids) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = check_range(b, first_world, n_worlds)) return rc;
    if (!rigids) return fail(XPBD_ERR_INVALID, "rigids is null");
    if (int rc = bind_device(b)) return rc;
    if (int rc = ensure_stage(b)) return rc;
    const int64_t B = b->desc.bodies_per_world, first = first_world * B, n = n_worlds * B;
    int slot = 0;
    for (int64_t done = 0; done < n; done += b->stage_bodies, slot ^= 1) {
        const int64_t m = std::min(b->stage_bodies, n - done);
        XPBD_CUDA_TRY(cudaEventSynchronize(b->stage_free[slot]));
        XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_stage[slot], rigids + done, (size_t)m * sizeof(xpbd_rigid),
                                      cudaMemcpyHostToDevice, b->stream));
        launch_aos64_to_soa(b->d_stage[slot], b->arrays, first + done, m, b->stream);
        XPBD_CUDA_TRY(cudaGetLastError());
        XPBD_CUDA_TRY(cudaEventRecord(b->stage_free[slot], b->stream));
    }
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

int xpbd_batch_download(xpbd_batch* b, int64_t first_world, int64_t n_worlds, xpbd_rigid* rigids) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = check_range(b, first_world, n_worlds)) return rc;
    if (!rigids) return fail(XPBD_ERR_INVALID, "rigids is null");
    if (int rc = bind_device(b)) return rc;
    if (int rc = ensure_stage(b)) return rc;
    const int64_t B = b->desc.bodies_per_world, first = first_world * B, n = n_worlds * B;
    int slot = 0;
    for (int64_t done = 0; done < n; done += b->stage_bodies, slot ^= 1) {
        const int64_t m = std::min(b->stage_bodies, n - done);
        XPBD_CUDA_TRY(cudaEventSynchronize(b->stage_free[slot]));
        launch_soa_to_aos64(b->arrays, b->d_stage[slot], first + done, m, b->stream);
        XPBD_CUDA_TRY(cudaGetLastError());
        XPBD_CUDA_TRY(cudaMemcpyAsync(rigids + done, b->d_stage[slot], (size_t)m * sizeof(xpbd_rigid),
                                      cudaMemcpyDeviceToHost, b->stream));
        XPBD_CUDA_TRY(cudaEventRecord(b->stage_free[slot], b->stream));
    }
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

int xpbd_batch_upload_broadcast(xpbd_batch* b, const xpbd_rigid* world_bodies) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (!world_bodies) return fail(XPBD_ERR_INVALID, "world_bodies is null");
    if (int rc = bind_device(b)) return rc;
    if (int rc = ensure_stage(b)) return rc;
    const int B = b->desc.bodies_per_world;
    XPBD_CUDA_TRY(cudaEventSynchronize(b->stage_free[0]));
    XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_stage[0], world_bodies, (size_t)B * sizeof(xpbd_rigid), cudaMemcpyHostToDevice,
                                  b->stream));
    launch_aos64_to_soa(b->d_stage[0], b->arrays, 0, B, b->stream);
    launch_broadcast_world(b->arrays, B, b->n_bodies, b->stream);
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaEventRecord(b->stage_free[0], b->stream));
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

static int ensure_dyn_stage(xpbd_batch* b) {
    if (b->d_dyn_stage) return XPBD_OK;
    XPBD_CUDA_TRY(cudaMalloc(&b->d_dyn_stage, (size_t)b->n_bodies * 13 * sizeof(float)));
    return XPBD_OK;
}

// Asynchronous on the batch stream except for the final synchronize (the caller owns `out`).
int xpbd_batch_download_dynamic_f32(xpbd_batch* b, int64_t first_world, int64_t n_worlds, float* out) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = check_range(b, first_world, n_worlds)) return rc;
    if (!out) return fail(XPBD_ERR_INVALID, "out is null");
    if (int rc = bind_device(b)) return rc;
    if (int rc = ensure_dyn_stage(b)) return rc;
    const int64_t B = b->desc.bodies_per_world, first = first_world * B, n = n_worlds * B;
    launch_dyn_pack(b->arrays, b->d_dyn_stage + 13 * first, first, n, b->stream);
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaMemcpyAsync(out, b->d_dyn_stage + 13 * first, (size_t)n * 13 * sizeof(float),
                                  cudaMemcpyDeviceToHost, b->stream));
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

int xpbd_batch_upload_dynamic_f32(xpbd_batch* b, int64_t first_world, int64_t n_worlds, const float* in) {
    if (int rc = check_batch(b)) return rc;
    b->fork_needed = true;
    if (int rc = check_range(b, first_world, n_worlds)) return rc;
    if (!in) return fail(XPBD_ERR_INVALID, "in is null");
    if (int rc = bind_device(b)) return rc;
    if (int rc = ensure_dyn_stage(b)) return rc;
    const int64_t B = b->desc.bodies_per_world, first = first_world * B, n = n_worlds * B;
    XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_dyn_stage + 13 * first, in, (size_t)n * 13 * sizeof(float),
                                  cudaMemcpyHostToDevice, b->stream));
    launch_dyn_unpack(b->d_dyn_stage + 13 * first, b->arrays, first, n, b->stream);
    XPBD_CUDA_TRY(cudaGetLastError());
    return XPBD_OK;  // ordered before any later launch on the batch stream
}

static int step_common(xpbd_batch* b, double dt, int32_t substep_count, int32_t n_frames, const float* in, float* out,
                       const xpbd_rigid* rin = nullptr, xpbd_rigid* rout = nullptr);

int xpbd_batch_step(xpbd_batch* b, double dt, int32_t substep_count, int32_t n_frames) {
    return step_common(b, dt, substep_count, n_frames, nullptr, nullptr);
}

int xpbd_batch_step_host_f32(xpbd_batch* b, double dt, int32_t substep_count, int32_t n_frames, const float* in, float* out) {
    return step_common(b, dt, substep_count, n_frames, in, out);
}

int xpbd_batch_step_host_rigid(xpbd_batch* b, double dt, int32_t substep_count, int32_t n_frames, const xpbd_rigid* in,
                               xpbd_rigid* out) {
    return step_common(b, dt, substep_count, n_frames, nullptr, nullptr, in, out);
}

static int step_common(xpbd_batch* b, double dt, int32_t substep_count, int32_t n_frames, const float* in, float* out,
                       const xpbd_rigid* rin, xpbd_rigid* rout) {
    if (int rc = check_batch(b)) return rc;
    if (substep_count <= 0 || n_frames <= 0) return fail(XPBD_ERR_INVALID, "substep_count and n_frames must be positive");
    if (!(dt > 0.0) || !std::isfinite(dt)) return fail(XPBD_ERR_INVALID, "dt must be positive and finite");
    if (int rc = bind_device(b)) return rc;
    if (int rc = finalize_hulls(b)) return rc;
    if (in || out) { if (int rc = ensure_dyn_stage(b)) return rc; }
    if ((rin || rout) && !b->d_rigid_stage) {
        XPBD_CUDA_TRY(cudaMalloc(&b->d_rigid_stage, (size_t)b->n_bodies * sizeof(xpbd_rigid)));
        b->rigid_stage_valid = false;
    }
    XPBD_CUDA_TRY(cudaEventRecord(b->ev_start, b->stream));
    XPBD_CUDA_TRY(cudaMemsetAsync(b->d_counters, 0, kBatchCounters * sizeof(unsigned long long), b->stream));
    b->stage_timed = false;
    b->stage_ms_valid = false;
    b->last.sub_batches = 1;
    const bool world_path = b->desc.body_body_contacts != 0 || b->desc.joints_per_world > 0;
    // Host buffers: the staged pipeline consumes them per world range (copies on the range's stream) and clears the
    // pointers; every other path copies the whole batch on the batch stream around its kernels.
    b->io_in = in;
    b->io_out = out;
    b->io_rigid_in = rin;
    b->io_rigid_out = rout;
    const bool per_range_io = world_path && staged_takes_host_buffers(b);
    if (rin) b->rigid_stage_valid = true;  // the staging mirror holds every body's record once this call's uploads ran
    if (!per_range_io) { if (int rc = rigid_io_before(b, 0, b->n_bodies, b->stream)) return rc; }
    if (b->io_in && !(world_path && staged_takes_host_buffers(b))) {
        XPBD_CUDA_TRY(cudaMemcpyAsync(b->d_dyn_stage, b->io_in, (size_t)b->n_bodies * 13 * sizeof(float), cudaMemcpyHostToDevice, b->stream));
        launch_dyn_unpack(b->d_dyn_stage, b->arrays, 0, b->n_bodies, b->stream);
        b->io_in = nullptr;
    }
    int rc = world_path ? step_world(b, (float)dt, substep_count, n_frames)
                        : step_single(b, (float)dt, substep_count, n_frames);
    if (rc != XPBD_OK) { b->io_in = nullptr; b->io_out = nullptr; b->io_rigid_in = nullptr; b->io_rigid_out = nullptr; return rc; }
    if (!per_range_io || b->last.sub_batches <= 1) { if (int rc2 = rigid_io_after(b, 0, b->n_bodies, b->stream)) return rc2; }
    b->io_rigid_in = nullptr;
    b->io_rigid_out = nullptr;
    if (b->io_out) {  // not consumed per range
        launch_dyn_pack(b->arrays, b->d_dyn_stage, 0, b->n_bodies, b->stream);
        XPBD_CUDA_TRY(cudaMemcpyAsync(b->io_out, b->d_dyn_stage, (size_t)b->n_bodies * 13 * sizeof(float), cudaMemcpyDeviceToHost, b->stream));
    }
    b->io_in = nullptr;
    b->io_out = nullptr;
    if (b->last.sub_batches <= 1) b->fork_needed = true;  // this step's kernels ran on the batch stream itself
    // multi-body paths: count the worlds that went NaN / Inf (the single-body kernel counts in its epilogue)
    if (world_path) {
        launch_count_nan_worlds(b->arrays, b->desc.n_worlds, b->desc.bodies_per_world, b->d_counters + 4, b->stream);
        b->last.kernel_launches += 1;
    }
    XPBD_CUDA_TRY(cudaMemcpyAsync(b->h_counters, b->d_counters, kBatchCounters * sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                                  b->stream));
    XPBD_CUDA_TRY(cudaEventRecord(b->ev_stop, b->stream));
    b->last.body_substeps = (uint64_t)b->n_bodies * (uint64_t)substep_count * (uint64_t)n_frames;
    b->stepped = true;
    return XPBD_OK;
}

int xpbd_batch_sync(xpbd_batch* b) {
    if (int rc = check_batch(b)) return rc;
    if (int rc = bind_device(b)) return rc;
    XPBD_CUDA_TRY(cudaStreamSynchronize(b->stream));
    return XPBD_OK;
}

int xpbd_batch_last_stats(xpbd_batch* b, xpbd_step_stats* out) {
    if (int rc = check_batch(b)) return rc;
    if (!out) return fail(XPBD_ERR_INVALID, "out is null");
    if (!b->stepped) return fail(XPBD_ERR_STATE, "no step call yet");
    if (int rc = bind_device(b)) return rc;
    XPBD_CUDA_TRY(cudaEventSynchronize(b->ev_stop));
    float ms = 0.f;
    XPBD_CUDA_TRY(cudaEventElapsedTime(&ms, b->ev_start, b->ev_stop));
    b->last.elapsed_ms = ms;
    unsigned long long cnt[4] = {b->h_counters[0], b->h_counters[1], b->h_counters[2], b->h_counters[3]};
    if (b->last.sub_batches > 1) staged_add_range_counters(b, cnt);
    b->last.constraints = cnt[0];
    b->last.overflow_worlds = cnt[1];
    b->last.candidate_pairs = b->stage_timed ? cnt[3] : 0;
    b->last.nan_worlds = (int32_t)std::min<unsigned long long>(b->h_counters[4], 0x7fffffffull);
    for (int k = 0; k < 4; ++k) {
        b->last.stage_ms[k] = 0.f;
        if (b->stage_timed && b->stage_ms_valid) XPBD_CUDA_TRY(cudaEventElapsedTime(&b->last.stage_ms[k], b->ev_stage[k], b->ev_stage[k + 1]));
    }
    if (cnt[2]) set_error("TMA hull staging fell back to plain loads in %llu blocks", cnt[2]);
    *out = b->last;
    return XPBD_OK;
}

}  // extern "C"
