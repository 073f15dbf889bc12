// rt_io_data.cpp — particle (Time mode), STL shape and Mesh-mode loaders + the cache format.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <map>
#include <unordered_map>
#include <numeric>
#include <string>
#include <mutex>
#include <thread>
#include <vector>

#include "rt_io_internal.hpp"
#include "rt_io_vtk.hpp"

using rtio::VtkArray;
using rtio::VtkPolyData;

// ================================================================== Time-mode particle file
struct rt_io_particles_s {
    std::vector<float> pos, quat, vel, size;
    std::vector<uint64_t> id, shape;
};

namespace {

const VtkArray &need(const std::map<std::string, VtkArray> &m, const char *name, size_t comps, size_t tuples, const char *what) {
    auto it = m.find(name);
    if (it == m.end()) raise(RT_IO_ERR_FORMAT, std::string("VTK: ") + what + " array '" + name + "' is missing");
    if (it->second.comps < comps || it->second.tuples() < tuples)
        raise(RT_IO_ERR_FORMAT, std::string("VTK: array '") + name + "' has the wrong shape");
    return it->second;
}

// vtk_reader::readVTKTimeFile (src/Util/VTKReaderImpl.cpp:139-251): POINTS, id, quat (4 components in
// file order -> float4{x,y,z,w}), vel, shape_id; doubles narrowed with static_cast<float>.
void load_particles(const std::string &path, rt_io_particles_s &p) {
    VtkPolyData pd;
    rtio::load_polydata(path, pd);
    const size_t n = pd.n_points();
    const VtkArray &id = need(pd.point_data, "id", 1, n, "point data");
    const VtkArray &q = need(pd.point_data, "quat", 4, n, "point data");
    const VtkArray &v = need(pd.point_data, "vel", 3, n, "point data");
    const VtkArray &sh = need(pd.point_data, "shape_id", 1, n, "point data");
    p.pos.resize(3 * n);
    p.quat.resize(4 * n);
    p.vel.resize(3 * n);
    p.id.resize(n);
    p.shape.resize(n);
    for (size_t i = 0; i < n; i++) {
        for (int k = 0; k < 3; k++) p.pos[3 * i + k] = static_cast<float>(pd.points[3 * i + k]);
        p.id[i] = static_cast<uint64_t>(id.v[i * id.comps]);
        for (int k = 0; k < 4; k++) p.quat[4 * i + k] = static_cast<float>(q.v[i * q.comps + k]);
        for (int k = 0; k < 3; k++) p.vel[3 * i + k] = static_cast<float>(v.v[i * v.comps + k]);
        p.shape[i] = static_cast<uint64_t>(sh.v[i * sh.comps]);
    }
    auto sz = pd.point_data.find("size");
    if (sz != pd.point_data.end() && sz->second.tuples() >= n) {
        p.size.resize(n);
        for (size_t i = 0; i < n; i++) p.size[i] = static_cast<float>(sz->second.v[i * sz->second.comps]);
    }
}

}  // namespace

extern "C" {

int rt_io_particles_load(const char *vtk_path, rt_io_particles *out) {
    if (!vtk_path || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    auto *p = new rt_io_particles_s;
    const int rc = guarded([&] { load_particles(vtk_path, *p); });
    if (rc) {
        delete p;
        return rc;
    }
    *out = p;
    return RT_IO_OK;
}
int rt_io_particles_get(rt_io_particles p, rt_io_particles_view *out) {
    if (!p || !out) return fail(RT_IO_ERR_ARG, "null argument");
    out->n = p->id.size();
    out->pos = p->pos.data();
    out->id = p->id.data();
    out->quat = p->quat.data();
    out->vel = p->vel.data();
    out->shape_id = p->shape.data();
    out->size = p->size.empty() ? nullptr : p->size.data();
    return RT_IO_OK;
}
void rt_io_particles_free(rt_io_particles p) { delete p; }

}  // extern "C"

// ================================================================== STL shapes
struct ShapeData {
    std::string path;
    std::vector<float> verts, fnorm, vnorm;
};
struct rt_io_shapes_s {
    std::vector<ShapeData> shapes;
};

namespace {

struct V3d {
    double x, y, z;
};
V3d sub(V3d a, V3d b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
V3d cross(V3d a, V3d b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
double dot(V3d a, V3d b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V3d unit(V3d a) {
    const double l = std::sqrt(dot(a, a));
    return l > 0.0 ? V3d{a.x / l, a.y / l, a.z / l} : V3d{0.0, 0.0, 0.0};
}

struct UnionFind {
    std::vector<size_t> p;
    explicit UnionFind(size_t n) : p(n) { std::iota(p.begin(), p.end(), 0); }
    size_t find(size_t a) {
        while (p[a] != a) a = p[a] = p[p[a]];
        return a;
    }
    void join(size_t a, size_t b) { p[find(a)] = find(b); }
};

// Outward orientation per connected component (stand-in for vtkPolyDataNormals' consistency +
// auto-orient pass): a closed component whose signed volume is negative has its normals flipped;
// open / flat components (|volume| ~ 0, e.g. the 2-triangle wall planes) keep their winding normal.
// tri_pt: 3 point ids per triangle; returns per-triangle sign (+1 / -1).
std::vector<int> orient_components(const std::vector<double> &pts, const std::vector<size_t> &tri_pt) {
    const size_t nt = tri_pt.size() / 3, np = pts.size() / 3;
    UnionFind uf(np);
    for (size_t t = 0; t < nt; t++) {
        uf.join(tri_pt[3 * t], tri_pt[3 * t + 1]);
        uf.join(tri_pt[3 * t], tri_pt[3 * t + 2]);
    }
    std::vector<double> vol(np, 0.0), area(np, 0.0);
    // signed volume about the component's first vertex (translation invariant for closed surfaces)
    std::vector<size_t> anchor(np, SIZE_MAX);
    for (size_t t = 0; t < nt; t++) {
        const size_t c = uf.find(tri_pt[3 * t]);
        if (anchor[c] == SIZE_MAX) anchor[c] = tri_pt[3 * t];
        auto P = [&](size_t i) { return V3d{pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]}; };
        const V3d o = P(anchor[c]);
        const V3d a = sub(P(tri_pt[3 * t]), o), b = sub(P(tri_pt[3 * t + 1]), o), d = sub(P(tri_pt[3 * t + 2]), o);
        vol[c] += dot(a, cross(b, d)) / 6.0;
        const V3d n = cross(sub(b, a), sub(d, a));
        area[c] += 0.5 * std::sqrt(dot(n, n));
    }
    std::vector<int> sign(nt, 1);
    for (size_t t = 0; t < nt; t++) {
        const size_t c = uf.find(tri_pt[3 * t]);
        // |V| must be significant against A^1.5 (a sphere has V = A^1.5 / (6 sqrt(pi)) ~ 0.094 A^1.5)
        if (vol[c] < 0.0 && -vol[c] > 1e-6 * std::pow(area[c], 1.5)) sign[t] = -1;
    }
    return sign;
}

// merges bit-identical vertices (vtkSTLReader merges points by default) -> point ids
std::vector<size_t> weld(const std::vector<float> &verts, std::vector<double> &pts) {
    struct Key {
        uint32_t a, b, c;
        bool operator==(const Key &o) const { return a == o.a && b == o.b && c == o.c; }
    };
    struct Hash {
        size_t operator()(const Key &k) const { return ((size_t)k.a * 0x9E3779B97F4A7C15ull) ^ ((size_t)k.b << 21) ^ ((size_t)k.c * 0xC2B2AE3D27D4EB4Full); }
    };
    std::unordered_map<Key, size_t, Hash> map;
    std::vector<size_t> ids(verts.size() / 3);
    for (size_t i = 0; i < ids.size(); i++) {
        Key k;
        float v[3] = {verts[3 * i] + 0.0f, verts[3 * i + 1] + 0.0f, verts[3 * i + 2] + 0.0f};  // -0 -> +0
        std::memcpy(&k.a, &v[0], 4);
        std::memcpy(&k.b, &v[1], 4);
        std::memcpy(&k.c, &v[2], 4);
        auto it = map.find(k);
        if (it == map.end()) {
            it = map.emplace(k, pts.size() / 3).first;
            pts.insert(pts.end(), {(double)v[0], (double)v[1], (double)v[2]});
        }
        ids[i] = it->second;
    }
    return ids;
}

// vtk_reader::readSTLFile (src/Util/VTKReaderImpl.cpp:254-321): vtkSTLReader ignores the facet normals
// of the file; cell normals come from the triangle winding, oriented outward.
void load_stl(const std::string &path, ShapeData &sd) {
    std::string text;
    if (!read_file(path, text)) raise(RT_IO_ERR_OPEN, "Failed to open file: " + path + "!");
    sd.path = path;
    bool ascii = text.compare(0, 5, "solid") == 0;
    if (ascii && text.size() >= 84) {  // binary files may also start with "solid": check the size equation
        uint32_t n;
        std::memcpy(&n, text.data() + 80, 4);
        if (text.size() == 84 + (size_t)n * 50 && text.find("facet") == std::string::npos) ascii = false;
    }
    if (ascii) {
        size_t p = 0;
        auto next_tok = [&](std::string &t) {
            while (p < text.size() && std::isspace((unsigned char)text[p])) p++;
            const size_t b = p;
            while (p < text.size() && !std::isspace((unsigned char)text[p])) p++;
            t = text.substr(b, p - b);
            return b != p;
        };
        std::string t;
        while (next_tok(t)) {
            if (t != "vertex") continue;
            for (int k = 0; k < 3; k++) {
                if (!next_tok(t)) raise(RT_IO_ERR_PARSE, "STL: truncated vertex in " + path);
                char *e = nullptr;
                const double v = std::strtod(t.c_str(), &e);
                if (e == t.c_str()) raise(RT_IO_ERR_PARSE, "STL: bad number '" + t + "' in " + path);
                sd.verts.push_back(static_cast<float>(v));
            }
        }
    } else {
        if (text.size() < 84) raise(RT_IO_ERR_PARSE, "STL: file too short: " + path);
        uint32_t n;
        std::memcpy(&n, text.data() + 80, 4);
        if (text.size() < 84 + (size_t)n * 50) raise(RT_IO_ERR_PARSE, "STL: truncated binary file: " + path);
        sd.verts.resize((size_t)n * 9);
        for (uint32_t i = 0; i < n; i++) std::memcpy(&sd.verts[(size_t)i * 9], text.data() + 84 + (size_t)i * 50 + 12, 36);
    }
    if (sd.verts.size() % 9) raise(RT_IO_ERR_PARSE, "STL: vertex count is not a multiple of 3 in " + path);
    const size_t nt = sd.verts.size() / 9;
    std::vector<double> pts;
    const std::vector<size_t> ids = weld(sd.verts, pts);
    const std::vector<int> sign = orient_components(pts, ids);
    sd.fnorm.resize(3 * nt);
    sd.vnorm.resize(9 * nt);
    for (size_t t = 0; t < nt; t++) {
        const float *v = &sd.verts[9 * t];
        const V3d a{v[0], v[1], v[2]}, b{v[3], v[4], v[5]}, c{v[6], v[7], v[8]};
        V3d n = unit(cross(sub(b, a), sub(c, a)));
        if (sign[t] < 0) n = {-n.x, -n.y, -n.z};
        const float f[3] = {static_cast<float>(n.x), static_cast<float>(n.y), static_cast<float>(n.z)};
        for (int k = 0; k < 3; k++) {
            sd.fnorm[3 * t + k] = f[k];
            for (int j = 0; j < 3; j++) sd.vnorm[9 * t + 3 * j + k] = f[k];
        }
    }
}

}  // namespace

extern "C" {

int rt_io_shapes_load_file(const char *stl_path, rt_io_shapes *out) {
    if (!stl_path || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    auto *s = new rt_io_shapes_s;
    s->shapes.resize(1);
    const int rc = guarded([&] { load_stl(stl_path, s->shapes[0]); });
    if (rc) {
        delete s;
        return rc;
    }
    *out = s;
    return RT_IO_OK;
}

// VTKTimeReader::readSTLFiles (src/Util/VTKTimeReader.cu:90-124): regular files with extension ".stl",
// sorted by path; the position in that order is the particle's shape_id.
int rt_io_shapes_load_dir(const char *stl_dir, rt_io_shapes *out) {
    if (!stl_dir || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    auto *s = new rt_io_shapes_s;
    const int rc = guarded([&] {
        std::error_code ec;
        std::vector<fs::path> paths;
        fs::directory_iterator it(stl_dir, ec);
        if (ec) raise(RT_IO_ERR_OPEN, std::string("Failed to open directory: ") + stl_dir + "!");
        for (const auto &e : it)
            if (e.is_regular_file() && e.path().extension() == ".stl") paths.push_back(e.path());
        std::sort(paths.begin(), paths.end());
        s->shapes.resize(paths.size());
        for (size_t i = 0; i < paths.size(); i++) load_stl(paths[i].string(), s->shapes[i]);
    });
    if (rc) {
        delete s;
        return rc;
    }
    *out = s;
    return RT_IO_OK;
}
size_t rt_io_shapes_count(rt_io_shapes s) { return s ? s->shapes.size() : 0; }
int rt_io_shapes_get(rt_io_shapes s, size_t i, rt_io_shape_view *out) {
    if (!s || !out || i >= s->shapes.size()) return fail(RT_IO_ERR_ARG, "bad shape index");
    const ShapeData &d = s->shapes[i];
    out->n_tris = d.verts.size() / 9;
    out->vertices = d.verts.data();
    out->face_normals = d.fnorm.data();
    out->vertex_normals = d.vnorm.data();
    out->path = d.path.c_str();
    return RT_IO_OK;
}
void rt_io_shapes_free(rt_io_shapes s) { delete s; }

}  // extern "C"

// ================================================================== Mesh-mode file + cache
struct rt_io_mesh_s {
    std::vector<uint64_t> id, voff;
    std::vector<float> vel, verts, norms;
    size_t n_cells = 0;
};

namespace {

// vtk_reader::readVTKMeshFile (src/Util/VTKReaderImpl.cpp:24-137): every cell must be a triangle strip;
// CELL_DATA id / vel; triangle k of a strip = (p_k, p_k+1, p_k+2) with the last two swapped for odd k;
// point normals = normalised sum of the unit normals of the triangles around a point, oriented
// outward per connected component (stand-in for vtkPolyDataNormals with point normals, no splitting,
// consistency and auto-orient on).
void load_mesh_vtk(const std::string &path, rt_io_mesh_s &m) {
    VtkPolyData pd;
    rtio::load_polydata(path, pd);
    if (pd.n_points() == 0) raise(RT_IO_ERR_FORMAT, "Failed to get poly data pointer or there is no points in file!");
    if (!pd.verts.empty()) raise(RT_IO_ERR_FORMAT, "Found illegal cell type: vtkVertex!");
    if (!pd.lines.empty()) raise(RT_IO_ERR_FORMAT, "Found illegal cell type: vtkLine!");
    if (!pd.polys.empty()) raise(RT_IO_ERR_FORMAT, "Found illegal cell type: vtkPolygon!");
    const size_t nc = pd.strips.size();
    if (pd.cell_data.find("id") == pd.cell_data.end() || pd.cell_data.find("vel") == pd.cell_data.end())
        raise(RT_IO_ERR_FORMAT, "Failed to read cell data!");
    const VtkArray &id = need(pd.cell_data, "id", 1, nc, "cell data");
    const VtkArray &vel = need(pd.cell_data, "vel", 3, nc, "cell data");
    const size_t np = pd.n_points();
    // triangle list over point ids
    std::vector<size_t> tri_pt;
    m.voff.assign(nc + 1, 0);
    for (size_t c = 0; c < nc; c++) {
        const auto &s = pd.strips[c];
        const size_t nt = s.size() >= 3 ? s.size() - 2 : 0;
        for (size_t k = 0; k < nt; k++) {
            size_t a = (size_t)s[k], b = (size_t)s[k + 1], d = (size_t)s[k + 2];
            if (k & 1) std::swap(b, d);
            if (a >= np || b >= np || d >= np) raise(RT_IO_ERR_FORMAT, "VTK: strip references a point out of range");
            tri_pt.insert(tri_pt.end(), {a, b, d});
        }
        m.voff[c + 1] = m.voff[c] + 3 * nt;
    }
    const std::vector<int> sign = orient_components(pd.points, tri_pt);
    std::vector<V3d> acc(np, V3d{0, 0, 0});
    auto P = [&](size_t i) { return V3d{pd.points[3 * i], pd.points[3 * i + 1], pd.points[3 * i + 2]}; };
    for (size_t t = 0; t < tri_pt.size() / 3; t++) {
        const V3d a = P(tri_pt[3 * t]), b = P(tri_pt[3 * t + 1]), d = P(tri_pt[3 * t + 2]);
        V3d n = unit(cross(sub(b, a), sub(d, a)));
        if (sign[t] < 0) n = {-n.x, -n.y, -n.z};
        for (int k = 0; k < 3; k++) {
            V3d &q = acc[tri_pt[3 * t + k]];
            q = {q.x + n.x, q.y + n.y, q.z + n.z};
        }
    }
    for (auto &q : acc) q = unit(q);
    m.n_cells = nc;
    m.id.resize(nc);
    m.vel.resize(3 * nc);
    for (size_t c = 0; c < nc; c++) {
        m.id[c] = static_cast<uint64_t>(id.v[c * id.comps]);
        for (int k = 0; k < 3; k++) m.vel[3 * c + k] = static_cast<float>(vel.v[c * vel.comps + k]);
    }
    m.verts.resize(3 * tri_pt.size());
    m.norms.resize(3 * tri_pt.size());
    for (size_t i = 0; i < tri_pt.size(); i++) {
        const size_t p = tri_pt[i];
        for (int k = 0; k < 3; k++) m.verts[3 * i + k] = static_cast<float>(pd.points[3 * p + k]);
        m.norms[3 * i] = static_cast<float>(acc[p].x);
        m.norms[3 * i + 1] = static_cast<float>(acc[p].y);
        m.norms[3 * i + 2] = static_cast<float>(acc[p].z);
    }
}

// Cache layout (include/Util/VTKMeshReader.cuh:10-24; writer VTKMeshReader.cu:54-72, reader :217-258),
// native little-endian, no header:  u64 particleCount; per particle: u64 id, float3 vel,
// u64 vertexCount (= 3 * triangles), float3 vertices[vertexCount], float3 normals[vertexCount].
void write_cache(const rt_io_mesh_s &m, const std::string &path) {
    std::ofstream out(path, std::ios::out | std::ios::binary);
    if (!out.is_open()) raise(RT_IO_ERR_OPEN, "Failed to open file: " + path + "!");
    const uint64_t n = m.id.size();
    out.write(reinterpret_cast<const char *>(&n), 8);
    for (uint64_t i = 0; i < n; i++) {
        out.write(reinterpret_cast<const char *>(&m.id[i]), 8);
        out.write(reinterpret_cast<const char *>(&m.vel[3 * i]), 12);
        const uint64_t vc = m.voff[i + 1] - m.voff[i];
        out.write(reinterpret_cast<const char *>(&vc), 8);
        out.write(reinterpret_cast<const char *>(&m.verts[3 * m.voff[i]]), (std::streamsize)(vc * 12));
        out.write(reinterpret_cast<const char *>(&m.norms[3 * m.voff[i]]), (std::streamsize)(vc * 12));
    }
    if (!out) raise(RT_IO_ERR_OPEN, "write failed: " + path);
}

void read_cache(const std::string &path, rt_io_mesh_s &m) {
    std::string buf;
    if (!read_file(path, buf)) raise(RT_IO_ERR_OPEN, "Failed to open file: " + path + "!");
    size_t p = 0;
    auto take = [&](void *dst, size_t bytes) {
        if (p + bytes > buf.size()) raise(RT_IO_ERR_PARSE, "cache file truncated: " + path);
        std::memcpy(dst, buf.data() + p, bytes);
        p += bytes;
    };
    uint64_t n = 0;
    take(&n, 8);
    if (n > buf.size()) raise(RT_IO_ERR_PARSE, "cache file corrupt: " + path);
    m.n_cells = n;
    m.id.resize(n);
    m.vel.resize(3 * n);
    m.voff.assign(n + 1, 0);
    for (uint64_t i = 0; i < n; i++) {
        take(&m.id[i], 8);
        take(&m.vel[3 * i], 12);
        uint64_t vc = 0;
        take(&vc, 8);
        if (vc * 24 > buf.size() - p) raise(RT_IO_ERR_PARSE, "cache file truncated: " + path);
        m.voff[i + 1] = m.voff[i] + vc;
        m.verts.resize(3 * m.voff[i + 1]);
        m.norms.resize(3 * m.voff[i + 1]);
        take(&m.verts[3 * m.voff[i]], vc * 12);
        take(&m.norms[3 * m.voff[i]], vc * 12);
    }
}

}  // namespace

extern "C" {

int rt_io_mesh_load_vtk(const char *vtk_path, rt_io_mesh *out) {
    if (!vtk_path || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    auto *m = new rt_io_mesh_s;
    const int rc = guarded([&] { load_mesh_vtk(vtk_path, *m); });
    if (rc) {
        delete m;
        return rc;
    }
    *out = m;
    return RT_IO_OK;
}
int rt_io_mesh_load_cache(const char *cache_path, rt_io_mesh *out) {
    if (!cache_path || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    auto *m = new rt_io_mesh_s;
    const int rc = guarded([&] { read_cache(cache_path, *m); });
    if (rc) {
        delete m;
        return rc;
    }
    *out = m;
    return RT_IO_OK;
}
int rt_io_mesh_write_cache(rt_io_mesh m, const char *cache_path) {
    if (!m || !cache_path) return fail(RT_IO_ERR_ARG, "null argument");
    return guarded([&] { write_cache(*m, cache_path); });
}
int rt_io_mesh_get(rt_io_mesh m, rt_io_mesh_view *out) {
    if (!m || !out) return fail(RT_IO_ERR_ARG, "null argument");
    out->n_particles = m->id.size();
    out->id = m->id.data();
    out->vel = m->vel.data();
    out->vertex_offset = m->voff.data();
    out->vertices = m->verts.data();
    out->normals = m->norms.data();
    return RT_IO_OK;
}
void rt_io_mesh_free(rt_io_mesh m) { delete m; }

// VTKMeshReader::writeVTKDataCache (src/Util/VTKMeshReader.cu:147-215)
int rt_io_cache_generate(const char *series_path, const char *series_name, const char *cache_dir, size_t max_threads,
                         uint64_t *max_cells_out) {
    if (!series_path || !series_name || !cache_dir) return fail(RT_IO_ERR_ARG, "null argument");
    rt_io_series ser = nullptr;
    int rc = rt_io_series_load(series_path, series_name, &ser);
    if (rc) return rc;
    const std::string dir = cache_dir;
    rc = guarded([&] {
        std::error_code ec;
        fs::create_directories(fs::path(dir), ec);
        fs::directory_iterator it(dir, ec);
        if (ec) raise(RT_IO_ERR_OPEN, "Failed to open directory: " + dir + "!");
        std::vector<fs::path> old;
        for (const auto &e : it)
            if (e.is_regular_file() && e.path().extension() == ".cache") old.push_back(e.path());
        for (const auto &p : old) fs::remove(p, ec);
        const size_t n = rt_io_series_count(ser);
        const size_t hw = std::max(1u, std::thread::hardware_concurrency());
        const size_t nthreads = std::max<size_t>(1, std::min<size_t>(hw, max_threads ? max_threads : 1));
        std::atomic<size_t> next(0), max_cells(0);
        std::atomic<int> first_err(0);
        std::string err_msg;
        std::mutex mu;
        auto work = [&] {
            for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= n || first_err.load()) return;
                rt_io_mesh_s m;
                try {
                    load_mesh_vtk(rt_io_series_file(ser, i), m);
                    write_cache(m, dir + "particle" + std::to_string(i) + ".cache");
                    size_t cur = max_cells.load();
                    while (m.n_cells > cur && !max_cells.compare_exchange_weak(cur, m.n_cells)) {}
                } catch (const rtio::IoError &e) {
                    std::lock_guard<std::mutex> g(mu);
                    if (!first_err.exchange(e.code)) err_msg = e.msg;
                }
            }
        };
        std::vector<std::thread> pool;
        for (size_t t = 0; t < nthreads; t++) pool.emplace_back(work);
        for (auto &t : pool) t.join();
        if (first_err.load()) raise(first_err.load(), err_msg);
        std::ofstream meta(dir + "metadata.cache", std::ios::out);
        if (!meta.is_open()) raise(RT_IO_ERR_OPEN, "Failed to open file: " + dir + "metadata.cache!");
        meta << (uint64_t)max_cells.load();
        if (max_cells_out) *max_cells_out = max_cells.load();
    });
    rt_io_series_free(ser);
    return rc;
}

// VTKMeshReader::readMaxCellCountAllVTKFile (src/Util/VTKMeshReader.cu:270-280)
int rt_io_cache_read_max_cells(const char *cache_dir, uint64_t *out) {
    if (!cache_dir || !out) return fail(RT_IO_ERR_ARG, "null argument");
    const std::string path = std::string(cache_dir) + "metadata.cache";
    std::ifstream meta(path);
    if (!meta.is_open()) return fail(RT_IO_ERR_OPEN, "Failed to open file: %s!", path.c_str());
    uint64_t v = 0;
    if (!(meta >> v)) return fail(RT_IO_ERR_PARSE, "metadata.cache does not hold a number: %s", path.c_str());
    *out = v;
    return RT_IO_OK;
}

}  // extern "C"
