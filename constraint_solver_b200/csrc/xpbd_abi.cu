// C ABI of libxpbd_b200.so (include/xpbd_b200.h): errors, hull table, batch lifetime, state
// upload/download, the step call.  No CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>

#include "xpbd_host_math.hpp"
#include "xpbd_internal.hpp"

namespace xpbd {

static thread_local std::string g_last_error;

void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
}
int fail(int status, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return status;
}

HostHull copy_hull(const xpbd_polytope* h) {
    HostHull o;
    o.vertices.assign(h->vertices, h->vertices + 3 * (size_t)h->n_vertices);
    if (h->n_edges > 0) o.edges.assign(h->edges, h->edges + 2 * (size_t)h->n_edges);
    o.face_offsets.assign(h->face_offsets, h->face_offsets + h->n_faces + 1);
    o.face_indices.assign(h->face_indices, h->face_indices + h->face_offsets[h->n_faces]);
    for (int k = 0; k < 3; ++k) o.centroid[k] = h->centroid[k];
    double r = 0.0;
    for (int i = 0; i < h->n_vertices; ++i) {
        xpbd_host::d3 d = {h->vertices[3 * i] - h->centroid[0], h->vertices[3 * i + 1] - h->centroid[1],
                           h->vertices[3 * i + 2] - h->centroid[2]};
        r = std::max(r, std::sqrt(xpbd_host::dot(d, d)));
    }
    o.radius = r;
    return o;
}

static size_t align4(size_t words) { return (words + 3) & ~(size_t)3; }  // 16-byte sections

std::vector<uint32_t> pack_hull_blob(const std::vector<HostHull>& hulls) {
    const size_t n = hulls.size();
    size_t words = 16 * n;
    std::vector<HullDesc> descs(n);
    for (size_t h = 0; h < n; ++h) {
        const HostHull& H = hulls[h];
        HullDesc& d = descs[h];
        std::memset(&d, 0, sizeof d);
        d.n_verts = (int32_t)(H.vertices.size() / 3);
        d.n_faces = (int32_t)H.face_offsets.size() - 1;
        d.n_edges = (int32_t)(H.edges.size() / 2);
        d.vert_off = (int32_t)words;  words = align4(words + 4 * (size_t)d.n_verts);
        d.plane_off = (int32_t)words; words = align4(words + 4 * (size_t)d.n_faces);
        d.edge_off = (int32_t)words;  words = align4(words + 2 * (size_t)d.n_edges);
        d.foff_off = (int32_t)words;  words = align4(words + (size_t)d.n_faces + 1);
        d.fidx_off = (int32_t)words;  words = align4(words + H.face_indices.size());
        d.cx = (float)H.centroid[0]; d.cy = (float)H.centroid[1]; d.cz = (float)H.centroid[2];
        d.radius = (float)H.radius;
        int mfv = 0;
        for (int f = 0; f < d.n_faces; ++f) mfv = std::max(mfv, H.face_offsets[f + 1] - H.face_offsets[f]);
        d.max_face_verts = mfv;
        // box classes: 3 groups of 4 edges whose fp32 object-space vectors are equal up to sign, bit for bit
        if (d.n_verts == 8 && d.n_edges == 12) {
            double ev[12][3];
            for (int e = 0; e < 12; ++e)
                for (int k = 0; k < 3; ++k)
                    ev[e][k] = (double)(float)H.vertices[3 * H.edges[2 * e + 1] + k] - (double)(float)H.vertices[3 * H.edges[2 * e] + k];
            int cls[12], n_cls = 0, count[12] = {0};
            for (int e = 0; e < 12; ++e) {
                cls[e] = -1;
                for (int f = 0; f < e && cls[e] < 0; ++f) {
                    const bool same = ev[e][0] == ev[f][0] && ev[e][1] == ev[f][1] && ev[e][2] == ev[f][2];
                    const bool opposite = ev[e][0] == -ev[f][0] && ev[e][1] == -ev[f][1] && ev[e][2] == -ev[f][2];
                    if (same || opposite) cls[e] = cls[f];
                }
                if (cls[e] < 0) cls[e] = n_cls++;
                ++count[cls[e]];
            }
            const bool nonzero = [&] { for (int e = 0; e < 12; ++e) if (ev[e][0] == 0 && ev[e][1] == 0 && ev[e][2] == 0) return false; return true; }();
            if (n_cls == 3 && count[0] == 4 && count[1] == 4 && count[2] == 4 && nonzero) {
                unsigned long long perm = 0;
                int k = 0;
                for (int c = 0; c < 3; ++c)
                    for (int e = 0; e < 12; ++e)
                        if (cls[e] == c) perm |= (unsigned long long)e << (4 * k++);
                d.edge_perm_lo = (uint32_t)perm; d.edge_perm_hi = (uint32_t)(perm >> 32);
                d.box_classes = 1;
            }
        }
    }
    std::vector<uint32_t> blob(words, 0u);
    for (size_t h = 0; h < n; ++h) {
        const HostHull& H = hulls[h];
        const HullDesc& d = descs[h];
        std::memcpy(&blob[16 * h], &d, sizeof d);
        float* v = reinterpret_cast<float*>(&blob[d.vert_off]);
        for (int i = 0; i < d.n_verts; ++i) {
            v[4 * i + 0] = (float)H.vertices[3 * i + 0];
            v[4 * i + 1] = (float)H.vertices[3 * i + 1];
            v[4 * i + 2] = (float)H.vertices[3 * i + 2];
            v[4 * i + 3] = 0.0f;
        }
        for (size_t e = 0; e < H.edges.size(); ++e) blob[d.edge_off + e] = (uint32_t)H.edges[e];
        for (size_t f = 0; f < H.face_offsets.size(); ++f) blob[d.foff_off + f] = (uint32_t)H.face_offsets[f];
        for (size_t k = 0; k < H.face_indices.size(); ++k) blob[d.fidx_off + k] = (uint32_t)H.face_indices[k];
    }
    return blob;
}

int DeviceHulls::upload(const std::vector<HostHull>& hulls, cudaStream_t st) {
    release();
    std::vector<uint32_t> host = pack_hull_blob(hulls);
    bytes = (uint32_t)(host.size() * 4);
    n_hulls = (int)hulls.size();
    max_verts = max_faces = max_edges = max_face_verts = 0;
    for (const HostHull& H : hulls) {
        max_verts = std::max(max_verts, (int)(H.vertices.size() / 3));
        max_faces = std::max(max_faces, (int)H.face_offsets.size() - 1);
        max_edges = std::max(max_edges, (int)(H.edges.size() / 2));
        for (size_t f = 0; f + 1 < H.face_offsets.size(); ++f)
            max_face_verts = std::max(max_face_verts, H.face_offsets[f + 1] - H.face_offsets[f]);
    }
    XPBD_CUDA_TRY(cudaMalloc(&blob, bytes));
    XPBD_CUDA_TRY(cudaMemcpyAsync(blob, host.data(), bytes, cudaMemcpyHostToDevice, st));
    launch_hull_planes(blob, n_hulls, st);
    XPBD_CUDA_TRY(cudaGetLastError());
    XPBD_CUDA_TRY(cudaStreamSynchronize(st));  // `host` goes out of scope
    return XPBD_OK;
}
void DeviceHulls::release() {
    if (blob) cudaFree(blob);
    blob = nullptr;
    bytes = 0;
    n_hulls = 0;
}

int finalize_hulls(xpbd_batch* b) {
    if (!b->hulls_dirty) return XPBD_OK;
    if (b->hulls.empty()) return fail(XPBD_ERR_STATE, "no hull registered: call xpbd_batch_add_hull first");
    b->fork_needed = true;
    int rc = b->dev_hulls.upload(b->hulls, b->stream);
    if (rc != XPBD_OK) return rc;
    b->hulls_dirty = false;
    return XPBD_OK;
}

}  // namespace xpbd

using namespace xpbd;

extern "C" {

const char* xpbd_last_error(void) { return g_last_error.c_str(); }
const char* xpbd_version(void) { return "xpbd_b200 0.1 (sm_100a)"; }

int xpbd_device_count(int32_t* count) {
    if (!count) return fail(XPBD_ERR_INVALID, "count is null");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        return fail(XPBD_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    }
    *count = n;
    return XPBD_OK;
}

int xpbd_polytope_rigid_metrics(const xpbd_polytope* hull, double density, xpbd_rigid_metrics* out) {
    if (!out) return fail(XPBD_ERR_INVALID, "out is null");
    if (const char* why = xpbd_host::validate_hull(hull)) return fail(XPBD_ERR_INVALID, "hull: %s", why);
    xpbd_host::rigid_metrics(*hull, density, out);
    return XPBD_OK;
}

int xpbd_rigid_new(const xpbd_rigid_metrics* m, xpbd_rigid* out) {
    if (!m || !out) return fail(XPBD_ERR_INVALID, "null argument");
    std::memset(out, 0, sizeof *out);
    out->inverse_mass = 1.0 / m->mass;
    if (!xpbd_host::invert3(m->inertia_tensor, out->inverse_inertia))
        return fail(XPBD_ERR_SINGULAR, "Inertia tensor is not invertible");
    for (int k = 0; k < 3; ++k) out->center_of_mass[k] = m->center_of_mass[k];
    out->rotation[3] = 1.0;
    return XPBD_OK;
}

int xpbd_polytope_planes(const xpbd_polytope* hull, double* out) {
    if (!out) return fail(XPBD_ERR_INVALID, "out is null");
    if (const char* why = xpbd_host::validate_hull(hull)) return fail(XPBD_ERR_INVALID, "hull: %s", why);
    for (int f = 0; f < hull->n_faces; ++f) {
        xpbd_host::PlaneD p = xpbd_host::face_plane(*hull, f);
        out[4 * f] = p.n[0]; out[4 * f + 1] = p.n[1]; out[4 * f + 2] = p.n[2]; out[4 * f + 3] = p.d;
    }
    return XPBD_OK;
}

}  // extern "C"
