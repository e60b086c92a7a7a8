// QuickHull (Barber, Dobkin, Huhdanpaa 1996; README.md:20 of the reference links the algorithm, the crate never
// implements it): convex hull of a 3-D point cloud -> the arrays of a `Polytope` (src/geometry.rs:81-93):
// vertices, edges, faces as convex polygons (coplanar triangles merged), centroid = mean of the hull vertices
// (what the reference's built-in hulls use: src/geometry.rs:99,116,159).  Setup-time host code, f64, SURVEY.md §8 F3.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <map>
#include <set>
#include <vector>

#include "../../include/xpbd_b200.h"

namespace xpbd {
int fail(int status, const char* fmt, ...);
}

namespace {

struct P3 { double x, y, z; };
P3 operator-(P3 a, P3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
P3 cross(P3 a, P3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
double dot(P3 a, P3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
double norm(P3 a) { return std::sqrt(dot(a, a)); }

struct Tri {
    int v[3];
    P3 n;       // unit outward normal
    double d;   // n . x = d
    bool alive = true;
    std::vector<int> outside;  // points strictly above this face
};

struct Hull {
    const std::vector<P3>& pts;
    std::vector<Tri> tris;
    double eps;
    explicit Hull(const std::vector<P3>& p, double e) : pts(p), eps(e) {}
    double height(const Tri& t, int p) const { return dot(t.n, pts[p]) - t.d; }
    // adds triangle (a, b, c) oriented so that `inside` is below it
    int add(int a, int b, int c, P3 inside) {
        Tri t;
        P3 n = cross(pts[b] - pts[a], pts[c] - pts[a]);
        double len = norm(n);
        n = {n.x / len, n.y / len, n.z / len};
        double d = dot(n, pts[a]);
        if (dot(n, inside) - d > 0) { std::swap(b, c); n = {-n.x, -n.y, -n.z}; d = -d; }
        t.v[0] = a; t.v[1] = b; t.v[2] = c; t.n = n; t.d = d;
        tris.push_back(t);
        return (int)tris.size() - 1;
    }
};

}  // namespace

extern "C" int xpbd_quickhull(const double* points, int32_t n_points, double* vertices, int32_t* n_vertices,
                              int32_t* edges, int32_t* n_edges, int32_t* face_offsets, int32_t* face_indices,
                              int32_t* n_faces, int32_t* n_face_indices, double centroid[3]) {
    using xpbd::fail;
    if (!points || !vertices || !n_vertices || !edges || !n_edges || !face_offsets || !face_indices || !n_faces ||
        !n_face_indices || !centroid)
        return fail(XPBD_ERR_INVALID, "xpbd_quickhull: null argument");
    if (n_points < 4) return fail(XPBD_ERR_INVALID, "xpbd_quickhull: need at least 4 points");
    std::vector<P3> pts((size_t)n_points);
    P3 lo = {1e300, 1e300, 1e300}, hi = {-1e300, -1e300, -1e300};
    for (int i = 0; i < n_points; ++i) {
        pts[i] = {points[3 * i], points[3 * i + 1], points[3 * i + 2]};
        if (!std::isfinite(pts[i].x) || !std::isfinite(pts[i].y) || !std::isfinite(pts[i].z))
            return fail(XPBD_ERR_INVALID, "xpbd_quickhull: point %d is not finite", i);
        lo = {std::min(lo.x, pts[i].x), std::min(lo.y, pts[i].y), std::min(lo.z, pts[i].z)};
        hi = {std::max(hi.x, pts[i].x), std::max(hi.y, pts[i].y), std::max(hi.z, pts[i].z)};
    }
    const double scale = std::max({hi.x - lo.x, hi.y - lo.y, hi.z - lo.z});
    if (!(scale > 0)) return fail(XPBD_ERR_INVALID, "xpbd_quickhull: all points coincide");
    const double eps = 1e-9 * scale;

    // --- initial simplex: extreme pair, farthest from their line, farthest from their plane
    int i0 = 0, i1 = 0;
    for (int i = 0; i < n_points; ++i) {
        if (pts[i].x < pts[i0].x) i0 = i;
        if (pts[i].x > pts[i1].x) i1 = i;
    }
    if (i0 == i1) {  // degenerate in x: take the farthest pair from point 0
        double best = -1;
        for (int i = 0; i < n_points; ++i) { double d = norm(pts[i] - pts[0]); if (d > best) { best = d; i1 = i; } }
        i0 = 0;
    }
    int i2 = -1; double best = eps;
    for (int i = 0; i < n_points; ++i) {
        double d = norm(cross(pts[i] - pts[i0], pts[i1] - pts[i0])) / norm(pts[i1] - pts[i0]);
        if (d > best) { best = d; i2 = i; }
    }
    if (i2 < 0) return fail(XPBD_ERR_INVALID, "xpbd_quickhull: points are collinear");
    P3 n012 = cross(pts[i1] - pts[i0], pts[i2] - pts[i0]);
    const double l012 = norm(n012);
    int i3 = -1; best = eps;
    for (int i = 0; i < n_points; ++i) {
        double d = std::fabs(dot(n012, pts[i] - pts[i0])) / l012;
        if (d > best) { best = d; i3 = i; }
    }
    if (i3 < 0) return fail(XPBD_ERR_INVALID, "xpbd_quickhull: points are coplanar");
    const P3 inside = {(pts[i0].x + pts[i1].x + pts[i2].x + pts[i3].x) / 4, (pts[i0].y + pts[i1].y + pts[i2].y + pts[i3].y) / 4,
                       (pts[i0].z + pts[i1].z + pts[i2].z + pts[i3].z) / 4};
    Hull H(pts, eps);
    H.add(i0, i1, i2, inside); H.add(i0, i1, i3, inside); H.add(i0, i2, i3, inside); H.add(i1, i2, i3, inside);
    for (int p = 0; p < n_points; ++p) {
        if (p == i0 || p == i1 || p == i2 || p == i3) continue;
        int bf = -1; double bh = eps;
        for (int f = 0; f < 4; ++f) { double h = H.height(H.tris[f], p); if (h > bh) { bh = h; bf = f; } }
        if (bf >= 0) H.tris[bf].outside.push_back(p);
    }
    // --- main loop: farthest point of some face's outside set, horizon, cone of new faces
    for (;;) {
        int f = -1;
        for (int k = 0; k < (int)H.tris.size(); ++k)
            if (H.tris[k].alive && !H.tris[k].outside.empty()) { f = k; break; }
        if (f < 0) break;
        int apex = -1; double ah = -1;
        for (int p : H.tris[f].outside) { double h = H.height(H.tris[f], p); if (h > ah) { ah = h; apex = p; } }
        // visible set (all alive faces the apex is above; n <= 256 points, so a linear scan is fine)
        std::vector<int> visible;
        for (int k = 0; k < (int)H.tris.size(); ++k)
            if (H.tris[k].alive && H.height(H.tris[k], apex) > eps) visible.push_back(k);
        // horizon = directed edges of visible faces whose reverse edge is not in a visible face
        std::set<std::pair<int, int>> dir;
        for (int k : visible)
            for (int e = 0; e < 3; ++e) dir.insert({H.tris[k].v[e], H.tris[k].v[(e + 1) % 3]});
        std::vector<std::pair<int, int>> horizon;
        for (auto& e : dir)
            if (!dir.count({e.second, e.first})) horizon.push_back(e);
        std::vector<int> orphans;
        for (int k : visible) {
            H.tris[k].alive = false;
            for (int p : H.tris[k].outside) if (p != apex) orphans.push_back(p);
            H.tris[k].outside.clear();
        }
        std::vector<int> fresh;
        for (auto& e : horizon) fresh.push_back(H.add(e.first, e.second, apex, inside));
        for (int p : orphans) {
            int bf = -1; double bh = eps;
            for (int k : fresh) { double h = H.height(H.tris[k], p); if (h > bh) { bh = h; bf = k; } }
            if (bf >= 0) H.tris[bf].outside.push_back(p);
        }
    }
    // --- merge coplanar, edge-connected triangles into polygons
    std::vector<int> alive;
    for (int k = 0; k < (int)H.tris.size(); ++k) if (H.tris[k].alive) alive.push_back(k);
    const int nt = (int)alive.size();
    std::vector<int> group(nt);
    for (int k = 0; k < nt; ++k) group[k] = k;
    auto find = [&](int x) { while (group[x] != x) x = group[x] = group[group[x]]; return x; };
    std::map<std::pair<int, int>, int> edge_owner;  // directed edge -> triangle (index into alive)
    for (int k = 0; k < nt; ++k)
        for (int e = 0; e < 3; ++e) edge_owner[{H.tris[alive[k]].v[e], H.tris[alive[k]].v[(e + 1) % 3]}] = k;
    const double coplanar = 1e-7;
    for (auto& kv : edge_owner) {
        auto rev = edge_owner.find({kv.first.second, kv.first.first});
        if (rev == edge_owner.end()) continue;
        const Tri& a = H.tris[alive[kv.second]];
        const Tri& b = H.tris[alive[rev->second]];
        if (dot(a.n, b.n) > 1.0 - coplanar) {
            // all three vertices of b on a's plane (relative to the hull size)
            bool on = true;
            for (int e = 0; e < 3; ++e) on = on && std::fabs(H.height(a, b.v[e])) <= 1e-7 * scale;
            if (on) group[find(kv.second)] = find(rev->second);
        }
    }
    std::map<int, std::vector<int>> members;
    for (int k = 0; k < nt; ++k) members[find(k)].push_back(k);
    std::vector<std::vector<int>> polys;
    for (auto& m : members) {
        // boundary = directed edges whose reverse is not inside the same group; walk the loop
        std::map<int, int> next;
        for (int k : m.second)
            for (int e = 0; e < 3; ++e) {
                int a = H.tris[alive[k]].v[e], b = H.tris[alive[k]].v[(e + 1) % 3];
                auto rev = edge_owner.find({b, a});
                if (rev != edge_owner.end() && find(rev->second) == m.first) continue;
                next[a] = b;
            }
        std::vector<int> loop;
        int start = next.begin()->first, cur = start;
        do { loop.push_back(cur); cur = next[cur]; } while (cur != start && (int)loop.size() <= (int)next.size());
        // drop collinear vertices (merged triangles can leave them on polygon edges)
        std::vector<int> clean;
        const int L = (int)loop.size();
        for (int k = 0; k < L; ++k) {
            P3 a = pts[loop[(k + L - 1) % L]], b = pts[loop[k]], c = pts[loop[(k + 1) % L]];
            if (norm(cross(b - a, c - b)) > 1e-9 * scale * scale) clean.push_back(loop[k]);
        }
        if (clean.size() >= 3) polys.push_back(clean);
    }
    // --- vertices in input order, remap, edges
    std::set<int> used;
    for (auto& p : polys) for (int v : p) used.insert(v);
    std::map<int, int> remap;
    int nv = 0;
    for (int v : used) remap[v] = nv++;
    std::set<std::pair<int, int>> uedges;
    int nfi = 0;
    for (auto& p : polys) {
        nfi += (int)p.size();
        for (size_t k = 0; k < p.size(); ++k) {
            int a = remap[p[k]], b = remap[p[(k + 1) % p.size()]];
            uedges.insert({std::min(a, b), std::max(a, b)});
        }
    }
    if (nv > *n_vertices || (int)uedges.size() > *n_edges || (int)polys.size() > *n_faces || nfi > *n_face_indices)
        return fail(XPBD_ERR_CAPACITY, "xpbd_quickhull: output needs %d vertices, %d edges, %d faces, %d face indices", nv,
                    (int)uedges.size(), (int)polys.size(), nfi);
    double cx = 0, cy = 0, cz = 0;
    for (int v : used) {
        const int o = remap[v];
        vertices[3 * o] = pts[v].x; vertices[3 * o + 1] = pts[v].y; vertices[3 * o + 2] = pts[v].z;
        cx += pts[v].x; cy += pts[v].y; cz += pts[v].z;
    }
    centroid[0] = cx / nv; centroid[1] = cy / nv; centroid[2] = cz / nv;
    int k = 0;
    for (auto& e : uedges) { edges[2 * k] = e.first; edges[2 * k + 1] = e.second; ++k; }
    int off = 0, fi = 0;
    // deterministic face order: by smallest remapped vertex sequence
    std::vector<std::vector<int>> out;
    for (auto& p : polys) {
        std::vector<int> q;
        for (int v : p) q.push_back(remap[v]);
        std::rotate(q.begin(), std::min_element(q.begin(), q.end()), q.end());
        out.push_back(q);
    }
    std::sort(out.begin(), out.end());
    for (auto& q : out) {
        face_offsets[fi++] = off;
        for (int v : q) face_indices[off++] = v;
    }
    face_offsets[fi] = off;
    *n_vertices = nv; *n_edges = (int)uedges.size(); *n_faces = (int)out.size(); *n_face_indices = off;
    return XPBD_OK;
}
