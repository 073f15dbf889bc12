// rtb200_render.cpp — headless driver: the reference's main() (src/Global/Main.cu:12-46) with
// RendererTime / RendererMesh commitRendererData + startRender on top of the C ABIs (rtb200.h,
// rtb200_io.h).  Reads the reference's config.json unchanged; presentation (SDL window + GL/VK/D3D
// copy) is replaced by image files.  Additive command-line switches exist because the reference has
// no argv parsing at all (its config path is a compile-time constant, ProgramArgumentParser.cuh:41).
//
//   rtb200_render CONFIG.json [--out DIR] [--width W] [--height H] [--spp N] [--depth D] [--seed S]
//                 [--max-files K] [--file-stride K] [--frame-stride S] [--max-frames F] [--extra-series NAME]...
//                 [--spheres] [--drop-shape ID]... [--format pfm|exr|ppm|none] [--device N] [--quiet]
//                 [--ranks N [--rank R]]   frame-parallel over the files: N forked processes, GPU (device + r) % count each
//                 [--aov]     also write the first-hit albedo and normal buffers (the reference's denoiser guides,
//                             shader/Shader.cu:270-272, src/Global/RendererImpl.cu:680-711) as <stem>_albedo / <stem>_normal
//                             images and the primary-hit primitive ids as <stem>_primid.u32 (raw little-endian u32, row-major)
//                 [--compat]  reference shading quirks 1+2 (scene option compat_shading = 3) and the reference's transport
//                             details (RT_FLAG_REF_NO_SELF_SKIP | RT_FLAG_REF_MUL_ON_RETURN): the reference's own pixels
//                 [--fused]   the queue-free persistent path kernel (scene option fused = 1)
#include <cuda_runtime.h>
#include <sys/wait.h>
#include <unistd.h>
#include <cerrno>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rtb200.h"
#include "../../include/rtb200_io.h"

namespace {

struct Options {
    std::string config, out = "out/", format = "pfm";
    int width = 0, height = 0, device = 0;
    int ranks = 1, rank = -1;  // frame-parallel over the files of the series: file k of the rendered set goes to rank k % ranks
    long spp = -1, depth = -1, max_files = -1, file_stride = 1, frame_stride = 1, max_frames = -1;
    long long seed = -1;
    bool spheres = false, quiet = false, aov = false, compat = false, fused = false;
    std::vector<std::string> extra_series;
    std::set<uint64_t> drop_shapes;
};

[[noreturn]] void die(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vfprintf(stderr, fmt, ap);
    va_end(ap);
    fputc('\n', stderr);
    exit(code);
}
#define IO(expr)                                                              \
    do {                                                                      \
        if ((expr) != 0) die(2, "%s: %s", #expr, rt_io_last_error());          \
    } while (0)
#define CU(expr)                                                                             \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) die(3, "%s: %s", #expr, cudaGetErrorString(_e));               \
    } while (0)

rt_context g_ctx = nullptr;
std::chrono::steady_clock::time_point g_t0;  // start of main(): RunStats.load_s = input loading (files, STL shapes, caches)
#define RT(expr)                                                                  \
    do {                                                                          \
        if (int _rc = (expr)) die(4, "%s: error %d: %s", #expr, _rc, rt_last_error(g_ctx)); \
        (void)0;                                                                  \
    } while (0)

Options parse_args(int argc, char **argv) {
    Options o;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto val = [&]() -> const char * {
            if (i + 1 >= argc) die(1, "missing value after %s", a.c_str());
            return argv[++i];
        };
        if (a == "--out") o.out = val();
        else if (a == "--width") o.width = atoi(val());
        else if (a == "--height") o.height = atoi(val());
        else if (a == "--spp") o.spp = atol(val());
        else if (a == "--depth") o.depth = atol(val());
        else if (a == "--seed") o.seed = atoll(val());
        else if (a == "--max-files") o.max_files = atol(val());
        else if (a == "--frame-stride") o.frame_stride = std::max(1L, atol(val()));
        else if (a == "--file-stride") o.file_stride = std::max(1L, atol(val()));  // render files 0, K, 2K, ... only
        else if (a == "--max-frames") o.max_frames = atol(val());
        else if (a == "--extra-series") o.extra_series.push_back(val());
        else if (a == "--drop-shape") o.drop_shapes.insert((uint64_t)atoll(val()));
        else if (a == "--format") o.format = val();
        else if (a == "--device") o.device = atoi(val());
        else if (a == "--ranks") o.ranks = std::max(1, atoi(val()));  // alone: fork one process per rank (GPU (device + r) % count)
        else if (a == "--rank") o.rank = atoi(val());                 // with --ranks: be that rank only (external launcher)
        else if (a == "--spheres") o.spheres = true;
        else if (a == "--quiet") o.quiet = true;
        else if (a == "--aov") o.aov = true;
        else if (a == "--compat") o.compat = true;
        else if (a == "--fused") o.fused = true;
        else if (!a.empty() && a[0] != '-' && o.config.empty()) o.config = a;
        else die(1, "unknown argument %s", a.c_str());
    }
    if (o.config.empty()) die(1, "usage: rtb200_render CONFIG.json [options]  (see the header of rtb200_render.cpp)");
    if (!o.out.empty() && o.out.back() != '/') o.out.push_back('/');
    return o;
}

template <typename T>
T *to_device(const T *h, size_t n) {
    T *d = nullptr;
    CU(cudaMalloc(&d, std::max<size_t>(n, 1) * sizeof(T)));
    if (n) CU(cudaMemcpy(d, h, n * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

// one loaded series (Time mode): per file the particle arrays
struct TimeSeries {
    rt_io_series series = nullptr;
    std::vector<rt_io_particles> files;
    std::vector<rt_io_particles_view> views;
};

// RendererTime::commitRendererData loads every file of the series up front with a pool of
// min(hardware_concurrency, cache-process-thread-count) threads (RendererTime.cu:186-224)
TimeSeries load_time_series(const std::string &dir, const std::string &name, size_t max_files, size_t threads) {
    TimeSeries ts;
    IO(rt_io_series_load(dir.c_str(), name.c_str(), &ts.series));
    size_t n = rt_io_series_count(ts.series);
    if (max_files && n > max_files) n = max_files;
    ts.files.assign(n, nullptr);
    ts.views.resize(n);
    std::atomic<size_t> next(0);
    std::atomic<int> err(0);
    std::string msg;
    std::mutex mu;
    auto work = [&] {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= n || err.load()) return;
            if (int rc = rt_io_particles_load(rt_io_series_file(ts.series, i), &ts.files[i])) {
                std::lock_guard<std::mutex> g(mu);
                if (!err.exchange(rc)) msg = rt_io_last_error();
            }
        }
    };
    std::vector<std::thread> pool;
    const size_t nt = std::max<size_t>(1, std::min<size_t>(threads, std::max(1u, std::thread::hardware_concurrency())));
    for (size_t t = 0; t < nt; t++) pool.emplace_back(work);
    for (auto &t : pool) t.join();
    if (err.load()) die(2, "loading %s%s: %s", dir.c_str(), name.c_str(), msg.c_str());
    for (size_t i = 0; i < n; i++) IO(rt_io_particles_get(ts.files[i], &ts.views[i]));
    return ts;
}

struct Frame {
    float *d_accum = nullptr, *d_color = nullptr;
    float *d_albedo = nullptr, *d_normal = nullptr;  // --aov
    uint32_t *d_primid = nullptr;
    uint8_t *d_rgba8 = nullptr;
    std::vector<float> h_color;
    std::vector<uint8_t> h_rgba8;
    uint32_t w = 0, h = 0;
};

// first-hit albedo / normal (float4 per pixel; the normal is written as n * 0.5 + 0.5 into PPM, raw into PFM / EXR)
// and primary-hit primitive ids
void write_aovs(const Options &o, Frame &fr, int flip_y, const std::string &stem, cudaStream_t st) {
    const size_t n = (size_t)fr.w * fr.h;
    for (int which = 0; which < 2; which++) {
        const std::string name = o.out + stem + (which ? "_normal" : "_albedo");
        CU(cudaMemcpyAsync(fr.h_color.data(), which ? fr.d_normal : fr.d_albedo, 16 * n, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (o.format == "ppm") {
            for (size_t i = 0; i < n; i++) {
                for (int c = 0; c < 3; c++) {
                    float v = fr.h_color[4 * i + c];
                    if (which) v = v * 0.5f + 0.5f;
                    fr.h_rgba8[4 * i + c] = (uint8_t)std::min(255.0f, std::max(0.0f, v * 256.0f));
                }
                fr.h_rgba8[4 * i + 3] = 255;
            }
            IO(rt_io_write_ppm((name + ".ppm").c_str(), fr.w, fr.h, fr.h_rgba8.data(), flip_y));
        } else if (o.format == "exr") {
            IO(rt_io_write_exr((name + ".exr").c_str(), fr.w, fr.h, fr.h_color.data(), flip_y));
        } else {
            IO(rt_io_write_pfm((name + ".pfm").c_str(), fr.w, fr.h, fr.h_color.data(), flip_y));
        }
    }
    std::vector<uint32_t> ids(n);
    CU(cudaMemcpyAsync(ids.data(), fr.d_primid, 4 * n, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const std::string name = o.out + stem + "_primid.u32";
    FILE *f = fopen(name.c_str(), "wb");
    if (!f || fwrite(ids.data(), 4, n, f) != n) die(2, "cannot write %s", name.c_str());
    fclose(f);
}

void write_frame(const Options &o, Frame &fr, rt_context ctx, int flip_y, const std::string &stem, cudaStream_t st) {
    if (o.format == "none") return;
    const size_t n = (size_t)fr.w * fr.h;
    if (o.aov) write_aovs(o, fr, flip_y, stem, st);
    if (o.format == "ppm") {
        RT(rt_resolve(ctx, fr.d_accum, n, nullptr, fr.d_rgba8, st));
        CU(cudaMemcpyAsync(fr.h_rgba8.data(), fr.d_rgba8, 4 * n, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        IO(rt_io_write_ppm((o.out + stem + ".ppm").c_str(), fr.w, fr.h, fr.h_rgba8.data(), flip_y));
        return;
    }
    // PFM / EXR hold LINEAR radiance: mean of the samples (accum.rgb / accum.w), no sRGB encoding
    CU(cudaMemcpyAsync(fr.h_color.data(), fr.d_accum, 16 * n, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    for (size_t i = 0; i < n; i++) {
        float *p = &fr.h_color[4 * i];
        const float inv = p[3] > 0.0f ? 1.0f / p[3] : 0.0f;
        p[0] *= inv;
        p[1] *= inv;
        p[2] *= inv;
        p[3] = 1.0f;
    }
    if (o.format == "exr") IO(rt_io_write_exr((o.out + stem + ".exr").c_str(), fr.w, fr.h, fr.h_color.data(), flip_y));
    else IO(rt_io_write_pfm((o.out + stem + ".pfm").c_str(), fr.w, fr.h, fr.h_color.data(), flip_y));
}

}  // namespace

namespace {

struct RunStats {
    size_t frames = 0;
    double build_ms = 0, render_ms = 0, load_s = 0;
    unsigned long long segments = 0, prims = 0;
};

// materials: config roughs ++ bakeColorRamp(preset, max particle count) (RendererTime.cu:235-243,
// RendererMesh.cu:230-233); config spheres come first in instance order (SBT order)
void set_materials(rt_scene scene, const rt_io_config_view &cfg, size_t max_particles) {
    std::vector<float> roughs(cfg.roughs, cfg.roughs + 3 * (size_t)cfg.n_rough);
    std::vector<float> ramp(3 * std::max<size_t>(max_particles, 1));
    rt_bake_color_ramp(cfg.preset, max_particles, ramp.data());
    roughs.insert(roughs.end(), ramp.begin(), ramp.begin() + 3 * max_particles);
    RT(rt_scene_set_materials(scene, roughs.data(), roughs.size() / 3, cfg.metals, cfg.n_metal));
}

struct ConfigSpheres {
    std::vector<float *> d_center, d_radius;
    float *d_tri_v = nullptr, *d_tri_n = nullptr;  // config "triangles"
};
// GeometryData.spheres: one single-sphere group + one instance each (RendererTime.cu:74-85,116-119).
// Every sphere gets its own JSON transform (the reference's user hook only refreshes instance 0,
// Main.cu:5-9, so there spheres 1.. stay at identity).
void add_config_spheres(rt_scene scene, const rt_io_config_view &cfg, ConfigSpheres &cs) {
    if (cs.d_center.empty())
        for (uint32_t i = 0; i < cfg.n_sphere; i++) {
            cs.d_center.push_back(to_device(cfg.sphere_center_radius + 4 * i, 3));
            cs.d_radius.push_back(to_device(cfg.sphere_center_radius + 4 * i + 3, 1));
        }
    for (uint32_t i = 0; i < cfg.n_sphere; i++) {
        uint32_t g = 0;
        RT(rt_scene_add_spheres(scene, cs.d_center[i], cs.d_radius[i], 1, &g));
        RT(rt_scene_add_instance(scene, g, cfg.sphere_transform + 12 * i, cfg.sphere_mat_type[i], (uint32_t)cfg.sphere_mat_index[i], 0, nullptr));
    }
    // GeometryData.triangles follow the spheres in SBT order (RendererMesh.cu:110-118): one world-space triangle
    // group + identity instance each (the reference's upload of them is dead code)
    if (cfg.n_triangle && !cs.d_tri_v) {
        cs.d_tri_v = to_device(cfg.triangle_vertices, 9 * (size_t)cfg.n_triangle);
        cs.d_tri_n = to_device(cfg.triangle_normals, 9 * (size_t)cfg.n_triangle);
    }
    for (uint32_t i = 0; i < cfg.n_triangle; i++) {
        uint32_t g = 0;
        RT(rt_scene_add_triangles(scene, cs.d_tri_v + 9 * (size_t)i, cs.d_tri_n + 9 * (size_t)i, 1, &g));
        RT(rt_scene_add_instance(scene, g, nullptr, cfg.triangle_mat_type[i], (uint32_t)cfg.triangle_mat_index[i], 0, nullptr));
    }
}

// Frame-parallel partition (north_star: "frame-parallel over the VTK time series"; the reference's loop over the files,
// RendererTime.cu:420-515 / RendererMesh.cu:363-444, is serial on one GPU, docs/technical-details.md:327).  The k-th file of
// the rendered set belongs to rank k % ranks.  A frame is a pure function of (scene of its file, sub-frame, seed, sample
// offset = global frame index * spp), and the global index of a file's first frame follows from the durations of the
// files before it, so every rank writes exactly the bytes the serial run writes for its files: no exchange of any kind.
bool file_is_mine(const Options &o, size_t fi) {
    return o.ranks <= 1 || (int)((fi / (size_t)o.file_stride) % (size_t)o.ranks) == o.rank;
}

void render_and_write(const Options &o, const rt_io_config_view &cfg, rt_scene scene, Frame &fr, const rt_camera &cam,
                      cudaStream_t st, uint32_t spp, uint32_t depth, uint64_t seed, size_t frame_total, const std::string &stem,
                      RunStats &rs) {
    RT(rt_accel_build(scene, st));
    rt_frame_params fp{};
    fp.width = fr.w;
    fp.height = fr.h;
    fp.spp = spp;
    fp.max_depth = depth;
    fp.seed = seed;
    fp.sample_offset = (uint32_t)(frame_total * spp);  // streams continue from frame to frame
    fp.background[0] = 0.7f;                            // RendererTime.cu:245-247 / RendererMesh.cu:260-262
    fp.background[1] = 0.8f;
    fp.background[2] = 0.9f;
    if (o.compat) fp.flags |= RT_FLAG_REF_NO_SELF_SKIP | (depth <= 5 ? RT_FLAG_REF_MUL_ON_RETURN : 0u);
    rt_outputs out{fr.d_accum, o.aov ? fr.d_albedo : nullptr, o.aov ? fr.d_normal : nullptr, o.aov ? fr.d_primid : nullptr};
    RT(rt_render(scene, &cam, &fp, &out, st));
    write_frame(o, fr, g_ctx, cfg.flip_up ? 0 : 1, stem, st);
    rt_stats stt;
    RT(rt_get_stats(scene, &stt));
    rs.frames++;
    rs.build_ms += stt.build_ms;
    rs.render_ms += stt.render_ms;
    rs.segments += stt.segments;
    rs.prims += stt.n_prims;
    if (!o.quiet)
        fprintf(stderr, "%s: %llu prims, build %.3f ms, render %.3f ms, %.1f Mrays/s\n", stem.c_str(), (unsigned long long)stt.n_prims,
                stt.build_ms, stt.render_ms, stt.render_ms > 0 ? stt.segments / stt.render_ms * 1e-3 : 0.0);
}

// ---------------------------------------------------------------------------------------------
// Time mode: RendererTime::commitRendererData + startRender (src/Global/RendererTime.cu:153-535)
int run_time_mode(const Options &o, const rt_io_config_view &cfg, rt_scene scene, Frame &fr, const rt_camera &cam, cudaStream_t st,
                  RunStats &rs) {
    const size_t threads = std::max<size_t>(1, cfg.thread_count);
    const size_t max_files = o.max_files > 0 ? (size_t)o.max_files : 0;
    std::vector<TimeSeries> all;
    all.push_back(load_time_series(cfg.series_path, cfg.series_name, max_files, threads));
    for (const auto &e : o.extra_series) all.push_back(load_time_series(cfg.series_path, e, max_files, threads));
    const TimeSeries &main_ts = all[0];
    const size_t n_files = main_ts.files.size();
    if (!n_files) die(2, "series %s%s has no files", cfg.series_path, cfg.series_name);

    // STL shapes -> one triangle group per shape (was: one GAS per shape, RendererTime.cu:180-182)
    rt_io_shapes shapes = nullptr;
    std::vector<float *> d_sv, d_sn;
    std::vector<size_t> s_nt;
    if (!o.spheres) {
        IO(rt_io_shapes_load_dir(cfg.stl_path, &shapes));
        for (size_t i = 0; i < rt_io_shapes_count(shapes); i++) {
            rt_io_shape_view sv;
            IO(rt_io_shapes_get(shapes, i, &sv));
            d_sv.push_back(to_device(sv.vertices, 9 * sv.n_tris));
            d_sn.push_back(to_device(sv.vertex_normals, 9 * sv.n_tris));
            s_nt.push_back(sv.n_tris);
        }
    }
    // material table sized by the largest particle id (+1) over every loaded file, >= max particle count
    size_t max_particles = 0;
    for (const auto &ts : all)
        for (const auto &v : ts.views) {
            max_particles = std::max(max_particles, v.n);
            for (size_t k = 0; k < v.n; k++) max_particles = std::max<size_t>(max_particles, v.id[k] + 1);
        }
    set_materials(scene, cfg, max_particles);
    rs.load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count();

    const uint32_t spp = o.spp > 0 ? (uint32_t)o.spp : cfg.spp, depth = o.depth > 0 ? (uint32_t)o.depth : cfg.max_depth;
    const uint64_t seed = o.seed >= 0 ? (uint64_t)o.seed : cfg.seed;
    ConfigSpheres cs;
    std::vector<float> tf;
    std::vector<float> sph_c, sph_r;
    float *d_pc = nullptr, *d_pr = nullptr;
    size_t d_cap = 0;
    size_t frame_total = 0;
    for (size_t fi = 0; fi < n_files; fi += (size_t)o.file_stride) {
        const float dur = rt_io_series_duration(main_ts.series, fi);
        const size_t frames = rt_io_frames_for_file(dur, cfg.fps, cfg.render_speed_ratio);
        if (!file_is_mine(o, fi)) {  // another rank's file: only the global frame counter (sample offsets) moves on
            frame_total += (frames + (size_t)o.frame_stride - 1) / (size_t)o.frame_stride;
            if (o.max_frames > 0 && (long)frame_total >= o.max_frames) return 0;
            continue;
        }
        // ---- scene of this file (the reference switches to this file's IAS + SBT, RendererTime.cu:420-424)
        RT(rt_scene_clear_geometry(scene));
        add_config_spheres(scene, cfg, cs);
        std::vector<uint32_t> shape_group(s_nt.size());
        for (size_t s = 0; s < s_nt.size(); s++) RT(rt_scene_add_triangles(scene, d_sv[s], d_sn[s], s_nt[s], &shape_group[s]));
        struct Ref {
            const rt_io_particles_view *cur, *next;
            size_t k;
        };
        std::vector<Ref> parts;
        for (const auto &ts : all) {
            if (fi >= ts.views.size()) continue;
            const rt_io_particles_view *cur = &ts.views[fi];
            const rt_io_particles_view *next = (fi + 1 < ts.views.size()) ? &ts.views[fi + 1] : cur;  // RendererTime.cu:442-447
            for (size_t k = 0; k < cur->n; k++)
                if (!o.drop_shapes.count(cur->shape_id[k])) parts.push_back({cur, next, k});
        }
        uint32_t first_particle = cfg.n_sphere + cfg.n_triangle, sphere_group = 0;
        if (o.spheres) {
            // BASELINE config 1 mapping: particle -> analytic sphere, centre = POINTS, r = size / 2 (0.05 when absent)
            sph_c.assign(3 * parts.size(), 0.0f);
            sph_r.resize(parts.size());
            for (size_t i = 0; i < parts.size(); i++) sph_r[i] = parts[i].cur->size ? parts[i].cur->size[parts[i].k] * 0.5f : 0.05f;
            if (parts.size() > d_cap) {
                if (d_pc) CU(cudaFree(d_pc));
                if (d_pr) CU(cudaFree(d_pr));
                d_cap = parts.size() + parts.size() / 4 + 16;
                CU(cudaMalloc(&d_pc, 3 * d_cap * sizeof(float)));
                CU(cudaMalloc(&d_pr, d_cap * sizeof(float)));
            }
            CU(cudaMemcpy(d_pc, sph_c.data(), 3 * parts.size() * sizeof(float), cudaMemcpyHostToDevice));
            CU(cudaMemcpy(d_pr, sph_r.data(), parts.size() * sizeof(float), cudaMemcpyHostToDevice));
        }
        for (size_t i = 0; i < parts.size(); i++) {
            const Ref &p = parts[i];
            const uint32_t mat = cfg.n_rough + (uint32_t)p.cur->id[p.k];  // particle.id + materialOffset (RendererTime.cu:261)
            if (o.spheres) {
                // one single-sphere group per particle keeps the instance <-> transform mapping of Time mode
                RT(rt_scene_add_spheres(scene, d_pc + 3 * i, d_pr + i, 1, &sphere_group));
                RT(rt_scene_add_instance(scene, sphere_group, nullptr, RT_MAT_ROUGH, mat, 0, nullptr));
            } else {
                const uint64_t sid = p.cur->shape_id[p.k];
                if (sid >= shape_group.size()) die(2, "particle shape_id %llu has no STL file", (unsigned long long)sid);
                RT(rt_scene_add_instance(scene, shape_group[sid], nullptr, RT_MAT_ROUGH, mat, 0, nullptr));
            }
        }
        // ---- frames of this file
        tf.resize(12 * parts.size());
        for (size_t f = 0; f < frames; f += (size_t)o.frame_stride) {
            for (size_t i = 0; i < parts.size(); i++) {
                const Ref &p = parts[i];
                // same local index in the next file (RendererTime.cu:447), guarded against a shorter next file
                const float *qn = (p.k < p.next->n) ? p.next->quat + 4 * p.k : p.cur->quat + 4 * p.k;
                rt_transform_time(p.cur->pos + 3 * p.k, p.cur->vel + 3 * p.k, p.cur->quat + 4 * p.k, qn, dur, frames, f,
                                  cfg.particle_shift, cfg.particle_scale, &tf[12 * i]);
                if (o.spheres) {  // spheres ignore rotation and scale: keep the translation only
                    float *m = &tf[12 * i];
                    const float t[3] = {m[3], m[7], m[11]};
                    const float id[12] = {1, 0, 0, t[0], 0, 1, 0, t[1], 0, 0, 1, t[2]};
                    memcpy(m, id, sizeof id);
                }
            }
            if (!parts.empty()) RT(rt_scene_set_transforms(scene, tf.data(), first_particle, parts.size(), st));
            char stem[64];
            snprintf(stem, sizeof stem, "frame_%05zu_%03zu", fi, f);
            render_and_write(o, cfg, scene, fr, cam, st, spp, depth, seed, frame_total, stem, rs);
            frame_total++;
            if (o.max_frames > 0 && (long)frame_total >= o.max_frames) return 0;
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Mesh mode: RendererMesh::commitRendererData + startRender (src/Global/RendererMesh.cu:150-454); particles come
// from the cache files when they exist, else straight from the VTK files.
int run_mesh_mode(const Options &o, const rt_io_config_view &cfg, rt_scene scene, Frame &fr, const rt_camera &cam, cudaStream_t st,
                  RunStats &rs) {
    rt_io_series series = nullptr;
    IO(rt_io_series_load(cfg.series_path, cfg.series_name, &series));
    size_t n_files = rt_io_series_count(series);
    if (o.max_files > 0) n_files = std::min<size_t>(n_files, (size_t)o.max_files);
    std::vector<rt_io_mesh> files(n_files, nullptr);
    std::vector<rt_io_mesh_view> views(n_files);
    size_t max_particles = 0;
    for (size_t i = 0; i < n_files; i++) {
        const std::string cache = std::string(cfg.cache_path) + "particle" + std::to_string(i) + ".cache";
        if (std::filesystem::exists(cache)) IO(rt_io_mesh_load_cache(cache.c_str(), &files[i]));
        else IO(rt_io_mesh_load_vtk(rt_io_series_file(series, i), &files[i]));
        IO(rt_io_mesh_get(files[i], &views[i]));
        max_particles = std::max(max_particles, views[i].n_particles);
        for (size_t k = 0; k < views[i].n_particles; k++) max_particles = std::max<size_t>(max_particles, views[i].id[k] + 1);
    }
    set_materials(scene, cfg, max_particles);
    rs.load_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count();
    const uint32_t spp = o.spp > 0 ? (uint32_t)o.spp : cfg.spp, depth = o.depth > 0 ? (uint32_t)o.depth : cfg.max_depth;
    const uint64_t seed = o.seed >= 0 ? (uint64_t)o.seed : cfg.seed;
    ConfigSpheres cs;
    std::vector<float> tf;
    float *d_v = nullptr, *d_n = nullptr;
    size_t cap = 0, frame_total = 0;
    for (size_t fi = 0; fi < n_files; fi += (size_t)o.file_stride) {
        const rt_io_mesh_view &mv = views[fi];
        const float dur = rt_io_series_duration(series, fi);
        const size_t frames = rt_io_frames_for_file(dur, cfg.fps, cfg.render_speed_ratio);
        if (!file_is_mine(o, fi)) {
            frame_total += (frames + (size_t)o.frame_stride - 1) / (size_t)o.frame_stride;
            if (o.max_frames > 0 && (long)frame_total >= o.max_frames) return 0;
            continue;
        }
        const size_t nv = mv.n_particles ? mv.vertex_offset[mv.n_particles] : 0;
        if (nv > cap) {
            if (d_v) CU(cudaFree(d_v));
            if (d_n) CU(cudaFree(d_n));
            cap = nv + nv / 4 + 48;
            CU(cudaMalloc(&d_v, 3 * cap * sizeof(float)));
            CU(cudaMalloc(&d_n, 3 * cap * sizeof(float)));
        }
        if (nv) {
            CU(cudaMemcpyAsync(d_v, mv.vertices, 3 * nv * sizeof(float), cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(d_n, mv.normals, 3 * nv * sizeof(float), cudaMemcpyHostToDevice, st));
        }
        RT(rt_scene_clear_geometry(scene));
        add_config_spheres(scene, cfg, cs);
        // one triangle group + one instance per particle (was: one GAS per particle, RendererMesh.cu:113-117)
        for (size_t k = 0; k < mv.n_particles; k++) {
            const size_t v0 = mv.vertex_offset[k], nt = (mv.vertex_offset[k + 1] - v0) / 3;
            uint32_t g = 0;
            RT(rt_scene_add_triangles(scene, d_v + 3 * v0, d_n + 3 * v0, nt, &g));
            RT(rt_scene_add_instance(scene, g, nullptr, RT_MAT_ROUGH, cfg.n_rough + (uint32_t)mv.id[k], 0, nullptr));  // RendererMesh.cu:276
        }
        tf.resize(12 * mv.n_particles);
        for (size_t f = 0; f < frames; f += (size_t)o.frame_stride) {
            for (size_t k = 0; k < mv.n_particles; k++)
                rt_transform_mesh(mv.vel + 3 * k, dur, frames, f, cfg.particle_shift, cfg.particle_scale, &tf[12 * k]);
            if (mv.n_particles) RT(rt_scene_set_transforms(scene, tf.data(), cfg.n_sphere + cfg.n_triangle, mv.n_particles, st));
            char stem[64];
            snprintf(stem, sizeof stem, "frame_%05zu_%03zu", fi, f);
            render_and_write(o, cfg, scene, fr, cam, st, spp, depth, seed, frame_total, stem, rs);
            frame_total++;
            if (o.max_frames > 0 && (long)frame_total >= o.max_frames) return 0;
        }
    }
    return 0;
}

}  // namespace

int main(int argc, char **argv) {
    Options o = parse_args(argc, argv);
    if (o.ranks > 1 && o.rank >= o.ranks) die(1, "--rank %d outside --ranks %d", o.rank, o.ranks);
    if (o.ranks > 1 && o.rank < 0) {
        // one child process per rank, forked before this process has touched CUDA; the parent only waits
        std::vector<pid_t> kids;
        for (int r = 0; r < o.ranks; r++) {
            fflush(stdout);
            fflush(stderr);
            const pid_t pid = fork();
            if (pid < 0) die(5, "fork: %s", strerror(errno));
            if (pid == 0) {
                o.rank = r;
                kids.clear();
                break;
            }
            kids.push_back(pid);
        }
        if (o.rank < 0) {
            int failed = 0;
            for (pid_t pid : kids) {
                int status = 0;
                if (waitpid(pid, &status, 0) < 0 || !WIFEXITED(status) || WEXITSTATUS(status) != 0) failed++;
            }
            printf("{\"ranks\": %d, \"failed\": %d, \"partition\": \"frame-parallel: file k of the rendered set -> rank k %% ranks\"}\n", o.ranks, failed);
            return failed ? 6 : 0;
        }
        int n_dev = 0;
        CU(cudaGetDeviceCount(&n_dev));
        if (n_dev < 1) die(3, "no CUDA device");
        o.device = (o.device + o.rank) % n_dev;  // fewer GPUs than ranks: ranks share devices
    }
    if (o.rank < 0) o.rank = 0;
    rt_io_config cfgh = nullptr;
    if (int rc = rt_io_config_load(o.config.c_str(), &cfgh)) die(-2 & 0xFF, "config: %s (code %d)", rt_io_last_error(), rc);
    rt_io_config_view cfg;
    IO(rt_io_config_get(cfgh, &cfg));
    // "cache": true with mesh input = write the cache files and exit (Main.cu:23-25, RendererMesh.cu:502-510)
    if (cfg.mesh && cfg.cache) {
        uint64_t mx = 0;
        IO(rt_io_cache_generate(cfg.series_path, cfg.series_name, cfg.cache_path, std::max<size_t>(1, cfg.thread_count), &mx));
        fprintf(stderr, "VTK data cache files are written to cache directory: %s (max cell count %llu)\n", cfg.cache_path, (unsigned long long)mx);
        return 0;
    }
    if (o.format != "none") std::filesystem::create_directories(o.out);
    CU(cudaSetDevice(o.device));
    if (int rc = rt_context_create(o.device, &g_ctx)) die(4, "rt_context_create: error %d: %s", rc, rt_last_error(nullptr));
    rt_scene scene = nullptr;
    RT(rt_scene_create(g_ctx, &scene));
    cudaStream_t st;
    CU(cudaStreamCreate(&st));
    Frame fr;
    fr.w = (uint32_t)(o.width > 0 ? o.width : cfg.window_width);
    fr.h = (uint32_t)(o.height > 0 ? o.height : cfg.window_height);
    const size_t n = (size_t)fr.w * fr.h;
    CU(cudaMalloc(&fr.d_accum, 16 * n));
    CU(cudaMalloc(&fr.d_rgba8, 4 * n));
    if (o.aov) {
        CU(cudaMalloc(&fr.d_albedo, 16 * n));
        CU(cudaMalloc(&fr.d_normal, 16 * n));
        CU(cudaMalloc(&fr.d_primid, 4 * n));
    }
    if (o.compat) RT(rt_scene_set_option(scene, "compat_shading", 3));
    if (o.fused) RT(rt_scene_set_option(scene, "fused", 1));
    fr.h_color.resize(4 * n);
    fr.h_rgba8.resize(4 * n);
    rt_camera cam;
    rt_camera_configure(cfg.camera_center, cfg.camera_target, cfg.up_direction, cfg.flip_up, &cam);  // SDL_GraphicsWindow.cu:4-14
    RunStats rs;
    const auto t0 = g_t0 = std::chrono::steady_clock::now();
    const int rc = cfg.mesh ? run_mesh_mode(o, cfg, scene, fr, cam, st, rs) : run_time_mode(o, cfg, scene, fr, cam, st, rs);
    const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    // frames_per_s / mrays_per_s count the frame loop (transform update + upload + rebuild + render + resolve/readback/file
    // output as selected), not the one-off input loading
    const double loop_s = std::max(wall - rs.load_s, 1e-9);
    printf("{\"frames\": %zu, \"width\": %u, \"height\": %u, \"segments\": %llu, \"prims\": %llu, \"build_ms\": %.3f, \"render_ms\": %.3f, "
           "\"wall_s\": %.3f, \"load_s\": %.3f, \"loop_s\": %.3f, \"frames_per_s\": %.3f, \"mrays_per_s\": %.2f, "
           "\"mrays_per_s_render\": %.2f, \"build_mprims_per_s\": %.2f, \"mode\": \"%s\", \"rank\": %d, \"ranks\": %d, \"device\": %d}\n",
           rs.frames, fr.w, fr.h, rs.segments, rs.prims, rs.build_ms, rs.render_ms, wall, rs.load_s, loop_s, rs.frames / loop_s,
           rs.segments / loop_s * 1e-6, rs.render_ms > 0 ? rs.segments / rs.render_ms * 1e-3 : 0.0,
           rs.build_ms > 0 ? rs.prims / rs.build_ms * 1e-3 : 0.0, cfg.mesh ? "mesh" : "time", o.rank, o.ranks, o.device);
    RT(rt_scene_destroy(scene));
    rt_context_destroy(g_ctx);
    return rc;
}
