// rt_io.cpp — librtb200_io.so: config / series / VTK / STL / cache readers and headless image
// writers (include/rtb200_io.h).  Pure CPU C++17.  The reference does this work with nlohmann/json
// and VTK 9.5 (vtkPolyDataReader, vtkSTLReader, vtkPolyDataNormals); both are replaced by the small
// parsers below, which read exactly the subset of the formats the reference consumes.
#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <filesystem>
#include <fstream>
#include <numeric>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/rtb200.h"
#include "../../include/rtb200_io.h"
#include "rt_json.hpp"

#include "rt_io_internal.hpp"

// ================================================================== config.json
struct rt_io_config_s {
    std::string series_path, series_name, cache_path, stl_path, preset, api, output;
    std::vector<float> roughs, metals, sphere_cr, sphere_tf;
    std::vector<int32_t> sphere_mt;
    std::vector<uint64_t> sphere_mi;
    std::vector<float> tri_v, tri_n;
    std::vector<int32_t> tri_mt;
    std::vector<uint64_t> tri_mi;
    rt_io_config_view v{};
};

namespace {

void vec3(const rtjson::Value &o, const char *key, float out[3]) {
    const auto a = o.at(key).as_floats();
    if (a.size() < 3) raise(RT_IO_ERR_PARSE, std::string("config: '") + key + "' needs 3 numbers");
    out[0] = a[0];
    out[1] = a[1];
    out[2] = a[2];
}

// ProgramArgumentParser::parseProgramArguments (src/Util/ProgramArgumentParser.cu:41-160): every
// key below is required (nlohmann's at()/operator[] on a missing key aborts the reference).
void parse_config(const std::string &text, rt_io_config_s &c) {
    const rtjson::Value d = rtjson::parse(text);
    for (const auto &r : *d.at("roughs").arr) {
        const auto a = r.at("albedo").as_floats();
        if (a.size() < 3) raise(RT_IO_ERR_PARSE, "config: rough albedo needs 3 numbers");
        c.roughs.insert(c.roughs.end(), {a[0], a[1], a[2]});
    }
    for (const auto &m : *d.at("metals").arr) {
        const auto a = m.at("albedo").as_floats();
        if (a.size() < 3) raise(RT_IO_ERR_PARSE, "config: metal albedo needs 3 numbers");
        c.metals.insert(c.metals.end(), {a[0], a[1], a[2], (float)m.at("fuzz").as_number()});
    }
    // parseSphereData (ProgramArgumentParser.cu:4-39)
    for (const auto &s : *d.at("spheres").arr) {
        float ce[3], sh[3], ro[3], sc[3], tf[12];
        vec3(s, "center", ce);
        vec3(s, "shift", sh);
        vec3(s, "rotate", ro);
        vec3(s, "scale", sc);
        c.sphere_cr.insert(c.sphere_cr.end(), {ce[0], ce[1], ce[2], (float)s.at("radius").as_number()});
        c.sphere_mt.push_back(s.at("mat-type").as_string() == "ROUGH" ? RT_MAT_ROUGH : RT_MAT_METAL);
        c.sphere_mi.push_back((uint64_t)s.at("mat-index").as_number());
        rt_construct_transform(sh, ro, sc, tf);
        c.sphere_tf.insert(c.sphere_tf.end(), tf, tf + 12);
    }
    // "triangles" is documented (docs/configuration.md:232-247) but never parsed by the reference, whose upload path
    // for them is dead code with an index bug (RendererMesh.cu:49,52: currentTriangles[i] instead of [j]; SURVEY §C
    // quirk 10).  Implemented behind the documented schema: vertices / normals as 3 x [x,y,z], mat-type, mat-index.
    if (d.contains("triangles")) {
        for (const auto &t : *d.at("triangles").arr) {
            for (const char *key : {"vertices", "normals"}) {
                const auto &rows = *t.at(key).arr;
                if (rows.size() != 3) raise(RT_IO_ERR_PARSE, std::string("config: triangle '") + key + "' needs 3 rows");
                for (const auto &row : rows) {
                    const auto a = row.as_floats();
                    if (a.size() < 3) raise(RT_IO_ERR_PARSE, std::string("config: triangle '") + key + "' rows need 3 numbers");
                    std::vector<float> &dst = key[0] == 'v' ? c.tri_v : c.tri_n;
                    dst.insert(dst.end(), {a[0], a[1], a[2]});
                }
            }
            c.tri_mt.push_back(t.at("mat-type").as_string() == "ROUGH" ? RT_MAT_ROUGH : RT_MAT_METAL);
            c.tri_mi.push_back((uint64_t)t.at("mat-index").as_number());
        }
    }
    const rtjson::Value &l = d.at("loop-data");
    c.api = l.at("api").as_string();
    if (c.api != "OGL" && c.api != "VK" && c.api != "D3D11" && c.api != "D3D12")
        raise(RT_IO_ERR_FORMAT, "config: invalid api type '" + c.api + "', must be \"OGL\", \"VK\", \"D3D11\" or \"D3D12\"");
    rt_io_config_view &v = c.v;
    v.mesh = d.at("mesh").as_bool();
    c.series_path = d.at("series-path").as_string();
    c.series_name = d.at("series-name").as_string();
    c.cache_path = d.at("cache-path").as_string();
    c.stl_path = d.at("stl-path").as_string();
    c.preset = d.at("particle-material-preset").as_string();
    v.window_width = (int)l.at("window-width").as_number();
    v.window_height = (int)l.at("window-height").as_number();
    v.fps = (uint64_t)l.at("fps").as_number();
    vec3(l, "camera-center", v.camera_center);
    vec3(l, "camera-target", v.camera_target);
    vec3(l, "up-direction", v.up_direction);
    v.render_speed_ratio = (uint64_t)l.at("render-speed-ratio").as_number();
    vec3(l, "particle-shift", v.particle_shift);
    vec3(l, "particle-scale", v.particle_scale);
    v.mouse_sensitivity = (float)l.at("mouse-sensitivity").as_number();
    v.camera_pitch_limit_degree = (float)l.at("camera-pitch-limit-degree").as_number();
    v.camera_speed_stride = (float)l.at("camera-speed-stride").as_number();
    v.camera_initial_speed_ratio = (uint64_t)l.at("camera-initial-speed-ratio").as_number();
    v.debug_mode = d.at("debug-mode").as_bool();
    v.cache = d.at("cache").as_bool();
    v.thread_count = (uint64_t)d.at("cache-process-thread-count").as_number();
    // additive keys
    v.spp = d.contains("spp") ? (uint32_t)d.at("spp").as_number() : 1u;
    v.max_depth = d.contains("max-depth") ? (uint32_t)d.at("max-depth").as_number() : 5u;  // Shader.cuh:8
    v.seed = d.contains("seed") ? (uint64_t)d.at("seed").as_number() : 0ull;
    v.frames = d.contains("frames") ? (uint64_t)d.at("frames").as_number() : 0ull;
    if (d.contains("output")) c.output = d.at("output").as_string();
    v.flip_up = c.api != "OGL";  // SDL_GraphicsWindow.cu:7-9
    v.n_rough = (uint32_t)(c.roughs.size() / 3);
    v.n_metal = (uint32_t)(c.metals.size() / 4);
    v.n_sphere = (uint32_t)c.sphere_mt.size();
    v.series_path = c.series_path.c_str();
    v.series_name = c.series_name.c_str();
    v.cache_path = c.cache_path.c_str();
    v.stl_path = c.stl_path.c_str();
    v.preset = c.preset.c_str();
    v.api = c.api.c_str();
    v.output = c.output.c_str();
    v.roughs = c.roughs.data();
    v.metals = c.metals.data();
    v.sphere_center_radius = c.sphere_cr.data();
    v.sphere_mat_type = c.sphere_mt.data();
    v.sphere_mat_index = c.sphere_mi.data();
    v.sphere_transform = c.sphere_tf.data();
    v.n_triangle = (uint32_t)c.tri_mt.size();
    v.triangle_vertices = c.tri_v.data();
    v.triangle_normals = c.tri_n.data();
    v.triangle_mat_type = c.tri_mt.data();
    v.triangle_mat_index = c.tri_mi.data();
}

}  // namespace

extern "C" {

const char *rt_io_last_error(void) { return rtio::tl_err().c_str(); }

int rt_io_config_parse(const char *json_text, rt_io_config *out) {
    if (!json_text || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    auto *c = new rt_io_config_s;
    const int rc = guarded([&] { parse_config(json_text, *c); });
    if (rc) {
        delete c;
        return rc;
    }
    *out = c;
    return RT_IO_OK;
}

int rt_io_config_load(const char *path, rt_io_config *out) {
    if (!path || !out) return fail(RT_IO_ERR_ARG, "null argument");
    std::string text;
    if (!read_file(path, text)) return fail(RT_IO_ERR_OPEN, "Failed to open config: %s!", path);
    return rt_io_config_parse(text.c_str(), out);
}

int rt_io_config_get(rt_io_config cfg, rt_io_config_view *out) {
    if (!cfg || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = cfg->v;
    return RT_IO_OK;
}

void rt_io_config_free(rt_io_config cfg) { delete cfg; }

}  // extern "C"

// ================================================================== .vtk.series
struct rt_io_series_s {
    std::vector<std::string> files;
    std::vector<float> times, durations;
};

extern "C" {

// VTKTimeReader::readSeriesFile (src/Util/VTKTimeReader.cu:31-88)
int rt_io_series_load(const char *series_path, const char *series_name, rt_io_series *out) {
    if (!series_path || !series_name || !out) return fail(RT_IO_ERR_ARG, "null argument");
    *out = nullptr;
    const std::string dir = series_path, path = dir + series_name;
    std::string text;
    if (!read_file(path, text)) return fail(RT_IO_ERR_OPEN, "Failed to open file: %s!", path.c_str());
    auto *s = new rt_io_series_s;
    const int rc = guarded([&] {
        const rtjson::Value d = rtjson::parse(text);
        (void)d.at("file-series-version").as_string();
        if (!(d.contains("files") && d.at("files").is_array())) raise(RT_IO_ERR_PARSE, "Failed to parse files array in series file!");
        for (const auto &it : *d.at("files").arr) {
            s->files.push_back(dir + it.at("name").as_string());
            s->times.push_back((float)it.at("time").as_number());
        }
        const size_t n = s->files.size();
        s->durations.assign(n, 0.0f);
        if (n == 1) {
            s->durations[0] = 1000.0f;
        } else if (n > 1) {
            for (size_t i = 0; i + 1 < n; i++) s->durations[i] = s->times[i + 1] - s->times[i];
            s->durations[n - 1] = s->durations[n - 2];
        }
    });
    if (rc) {
        delete s;
        return rc;
    }
    *out = s;
    return RT_IO_OK;
}
size_t rt_io_series_count(rt_io_series s) { return s ? s->files.size() : 0; }
const char *rt_io_series_file(rt_io_series s, size_t i) { return (s && i < s->files.size()) ? s->files[i].c_str() : ""; }
float rt_io_series_time(rt_io_series s, size_t i) { return (s && i < s->times.size()) ? s->times[i] : 0.0f; }
float rt_io_series_duration(rt_io_series s, size_t i) { return (s && i < s->durations.size()) ? s->durations[i] : 0.0f; }
void rt_io_series_free(rt_io_series s) { delete s; }

// RendererTime.cu:427-428 / RendererMesh.cu:370-371: static_cast<size_t>(duration * float(fps * ratio))
size_t rt_io_frames_for_file(float duration, uint64_t fps, uint64_t render_speed_ratio) {
    const float f = duration * static_cast<float>(fps * render_speed_ratio);
    return f > 0.0f ? static_cast<size_t>(f) : 0;
}

}  // extern "C"
