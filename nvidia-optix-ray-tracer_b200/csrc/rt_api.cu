// rt_api.cu — C ABI entry points (include/rtb200.h): context, scene assembly, stats.
#include <algorithm>
#include <cstring>
#include <new>

#include "rt_internal.h"

namespace rtb {

static thread_local std::string tl_error;

int set_error(Context *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    tl_error = buf;
    if (ctx) {
        std::lock_guard<std::mutex> g(ctx->mu);
        ctx->last_error = buf;
    }
    return code;
}

static bool linear_is_identity(const float *m) {
    return m[0] == 1.0f && m[1] == 0.0f && m[2] == 0.0f && m[4] == 0.0f && m[5] == 1.0f && m[6] == 0.0f &&
           m[8] == 0.0f && m[9] == 0.0f && m[10] == 1.0f;
}

struct DeviceGuard {  // bind the calling thread to the context's device for the duration of a call
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace rtb

using namespace rtb;

struct rt_context_s {
    Context c;
};
struct rt_scene_s {
    Scene s;
};

extern "C" {

int rt_abi_version(void) { return RTB200_ABI_VERSION; }

int rt_context_create(int device, rt_context *out) {
    if (!out) return RT_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return set_error(nullptr, RT_ERR_CUDA, "no CUDA device: %s (this library has no CPU fallback)",
                         e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= n) return set_error(nullptr, RT_ERR_INVALID, "device %d out of range [0,%d)", device, n);
    rt_context_s *c = new (std::nothrow) rt_context_s;
    if (!c) return RT_ERR_OOM;
    c->c.device = device;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) {
        delete c;
        return set_error(nullptr, RT_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    }
    if (prop.major < 10) {
        delete c;
        return set_error(nullptr, RT_ERR_UNSUPPORTED, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                         prop.major, prop.minor);
    }
    c->c.sm_count = prop.multiProcessorCount;
    e = cudaSetDevice(device);
    if (e != cudaSuccess) {
        delete c;
        return set_error(nullptr, RT_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    }
    *out = c;
    return RT_OK;
}

int rt_context_destroy(rt_context ctx) {
    if (!ctx) return RT_ERR_INVALID;
    delete ctx;
    return RT_OK;
}

const char *rt_last_error(rt_context ctx) {
    if (!tl_error.empty() || !ctx) return tl_error.c_str();
    std::lock_guard<std::mutex> g(ctx->c.mu);
    tl_error = ctx->c.last_error;
    return tl_error.c_str();
}

int rt_scene_create(rt_context ctx, rt_scene *out) {
    if (!ctx || !out) return RT_ERR_INVALID;
    rt_scene_s *s = new (std::nothrow) rt_scene_s;
    if (!s) return RT_ERR_OOM;
    s->s.ctx = &ctx->c;
    if (const char *e = getenv("RTB200_LEAF_MAX")) s->s.leaf_max = (uint32_t)std::min(8, std::max(1, atoi(e)));
    if (const char *e = getenv("RTB200_BATCH_SPP")) s->s.batch_spp = (uint32_t)std::min(64, std::max(1, atoi(e)));
    *out = s;
    return RT_OK;
}

int rt_scene_set_option(rt_scene scene, const char *name, int64_t value) {
    if (!scene || !name) return RT_ERR_INVALID;
    Scene &s = scene->s;
    const std::string k = name;
    if (k == "profile") s.profile = value != 0;
    else if (k == "leaf_max") { s.leaf_max = (uint32_t)std::min<int64_t>(8, std::max<int64_t>(0, value)); s.built = false; }
    else if (k == "batch_spp") s.batch_spp = (uint32_t)std::min<int64_t>(64, std::max<int64_t>(0, value));
    else if (k == "wide") { s.wide = value != 0; s.built = false; }
    else if (k == "fused") s.fused = value != 0;
    else if (k == "compact_queues") s.compact_queues = value != 0;
    else if (k == "chunk_spp") s.chunk_spp = (uint32_t)std::min<int64_t>(1024, std::max<int64_t>(1, value));
    else if (k == "compat_shading") { s.compat = (uint32_t)(value & 3); s.built = false; }
    else if (k == "t1") s.t1 = (uint32_t)std::min<int64_t>(33, std::max<int64_t>(0, value));
    else if (k == "t2") s.t2 = (uint32_t)std::min<int64_t>(33, std::max<int64_t>(0, value));
    else if (k == "t3") s.t3 = (uint32_t)std::min<int64_t>(33, std::max<int64_t>(0, value));
    else if (k == "fetch_min") s.fetch_min = (uint32_t)std::min<int64_t>(32, std::max<int64_t>(1, value));
    else if (k == "ray_bins") s.ray_bins = value >= 24 ? 24u : 8u;
    else if (k == "shade_threads") s.shade_threads = value >= 1024 ? 1024u : (value <= 256 ? 256u : 512u);
    else if (k == "collapse_mode") { s.collapse_mode = (uint32_t)std::min<int64_t>(1, std::max<int64_t>(0, value)); s.built = false; }
    else if (k == "refit_mode") { s.refit_mode = (uint32_t)std::min<int64_t>(2, std::max<int64_t>(0, value)); s.built = false; }
    else if (k == "extend_blocks_per_sm") s.extend_blocks_per_sm = (uint32_t)std::min<int64_t>(32, std::max<int64_t>(1, value));
    else if (k == "rng_max_lag") s.rng_max_lag = (uint32_t)std::min<int64_t>(2048, std::max<int64_t>(0, value));
    else return set_error(s.ctx, RT_ERR_INVALID, "unknown option '%s'", name);
    return RT_OK;
}

int rt_scene_destroy(rt_scene scene) {
    if (!scene) return RT_ERR_INVALID;
    Scene &s = scene->s;
    DeviceGuard g(s.ctx->device);
    cudaDeviceSynchronize();
    s.d_groups.release(); s.d_insts.release(); s.d_inst_base.release(); s.d_roughs.release(); s.d_metals.release();
    s.w_geom.release(); s.w_mat.release(); s.w_shade.release(); s.tri_v12.release(); s.tri_nrm.release();
    for (int k = 0; k < 2; k++) {
        s.keys[k].release(); s.vals[k].release();
        s.q_o[k].release(); s.q_d[k].release(); s.q_t[k].release(); s.q_r[k].release(); s.q_r4[k].release(); s.q_m[k].release();
    }
    s.sort_hist.release(); s.bucket_tab.release(); s.nodes.release(); s.parent_int.release(); s.parent_leaf.release();
    s.visit.release(); s.slot_geom.release(); s.slot_prim.release(); s.slot_mat.release(); s.info.release();
    s.wnodes.release(); s.wslot_geom.release(); s.wslot_prim.release(); s.wslot_mat.release();
    s.phit.release(); s.hit.release(); s.u_skip.release(); s.prim_fetch.release(); s.rng_cur.release(); s.result.release(); s.hit_pix.release(); s.rng_start.release(); s.rng_pix_off.release(); s.counters.release(); s.seg_total.release();
    for (auto &e : s.ev)
        if (e) cudaEventDestroy(e);
    for (auto &e : s.prof_ev) cudaEventDestroy(e);
    if (s.ev_hit) cudaEventDestroy(s.ev_hit);
    if (s.h_hit) cudaFreeHost(s.h_hit);
    if (s.ev_fork) cudaEventDestroy(s.ev_fork);
    if (s.ev_join) cudaEventDestroy(s.ev_join);
    if (s.aux_stream) cudaStreamDestroy(s.aux_stream);
    delete scene;
    return RT_OK;
}

int rt_scene_set_materials(rt_scene scene, const float *roughs, size_t n_rough, const float *metals, size_t n_metal) {
    if (!scene || (n_rough && !roughs) || (n_metal && !metals)) return RT_ERR_INVALID;
    Scene &s = scene->s;
    Context *ctx = s.ctx;
    DeviceGuard g(ctx->device);
    // expand float3 -> float4 rows so that a material is one 128-bit load
    RT_CUDA(ctx, s.d_roughs.reserve(n_rough ? n_rough : 1, &s.bytes));
    RT_CUDA(ctx, s.d_metals.reserve(n_metal ? n_metal : 1, &s.bytes));
    if (n_rough) {
        RT_CUDA(ctx, cudaMemset(s.d_roughs.p, 0, n_rough * sizeof(float4)));
        RT_CUDA(ctx, cudaMemcpy2D(s.d_roughs.p, sizeof(float4), roughs, 3 * sizeof(float), 3 * sizeof(float), n_rough, cudaMemcpyDefault));
    }
    if (n_metal) RT_CUDA(ctx, cudaMemcpy(s.d_metals.p, metals, n_metal * sizeof(float4), cudaMemcpyDefault));
    // rt_accel_build checked every instance's material range against the tables as they stood then; a table that
    // shrinks could leave indices pointing at stale rows of the grow-only buffers, so the check has to run again
    if (n_rough < s.n_rough || n_metal < s.n_metal) s.built = false;
    s.n_rough = n_rough;
    s.n_metal = n_metal;
    return RT_OK;
}

static int add_group(rt_scene scene, const float *a, const float *b, size_t count, uint32_t kind, uint32_t *out) {
    if (!scene || (count && (!a || !b))) return RT_ERR_INVALID;
    Scene &s = scene->s;
    if (count >= (1ull << 31)) return set_error(s.ctx, RT_ERR_INVALID, "group too large");
    GroupDev g{a, b, (uint32_t)count, kind};
    s.groups.push_back(g);
    s.built = false;
    if (out) *out = (uint32_t)(s.groups.size() - 1);
    return RT_OK;
}
int rt_scene_add_spheres(rt_scene scene, const float *dev_centers, const float *dev_radii, size_t count, uint32_t *group_out) {
    return add_group(scene, dev_centers, dev_radii, count, 0, group_out);
}
int rt_scene_add_triangles(rt_scene scene, const float *dev_vertices, const float *dev_normals, size_t count, uint32_t *group_out) {
    return add_group(scene, dev_vertices, dev_normals, count, 1, group_out);
}
int rt_scene_clear_geometry(rt_scene scene) {
    if (!scene) return RT_ERR_INVALID;
    Scene &s = scene->s;
    s.groups.clear();
    s.insts.clear();
    s.dev_tf.clear();
    s.built = false;
    return RT_OK;
}
int rt_scene_update_group(rt_scene scene, uint32_t group, const float *a, const float *b, size_t count) {
    if (!scene) return RT_ERR_INVALID;
    Scene &s = scene->s;
    if (group >= s.groups.size() || (count && (!a || !b)) || count >= (1ull << 31))
        return set_error(s.ctx, RT_ERR_INVALID, "bad group %u", group);
    s.groups[group].a = a;
    s.groups[group].b = b;
    s.groups[group].count = (uint32_t)count;
    s.built = false;
    return RT_OK;
}

int rt_scene_add_instance(rt_scene scene, uint32_t group, const float *transform_h, int mat_type, uint32_t mat_index,
                          uint32_t mat_stride, uint32_t *instance_out) {
    if (!scene) return RT_ERR_INVALID;
    Scene &s = scene->s;
    if (group >= s.groups.size()) return set_error(s.ctx, RT_ERR_INVALID, "bad group %u", group);
    if (mat_type != RT_MAT_ROUGH && mat_type != RT_MAT_METAL) return set_error(s.ctx, RT_ERR_INVALID, "bad material type %d", mat_type);
    if (mat_index >= (1u << 31)) return set_error(s.ctx, RT_ERR_INVALID, "material index too large");
    InstDev in;
    memset(&in, 0, sizeof in);
    static const float ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    memcpy(in.m, transform_h ? transform_h : ident, sizeof in.m);
    in.group = group;
    in.mat = ((uint32_t)mat_type << 31) | mat_index;
    in.mat_stride = mat_stride;
    in.identity = linear_is_identity(in.m) ? 1u : 0u;
    s.insts.push_back(in);
    s.built = false;
    if (instance_out) *instance_out = (uint32_t)(s.insts.size() - 1);
    return RT_OK;
}

int rt_scene_set_transforms(rt_scene scene, const float *transforms, size_t first, size_t n, void *stream) {
    if (!scene || (n && !transforms)) return RT_ERR_INVALID;
    Scene &s = scene->s;
    Context *ctx = s.ctx;
    if (first + n > s.insts.size()) return set_error(ctx, RT_ERR_INVALID, "instances [%zu,%zu) out of range", first, first + n);
    DeviceGuard g(ctx->device);
    // The instance table is small and host-resident (it is re-uploaded by rt_accel_build together
    // with the prefix table); device-resident transform arrays are fetched here.
    cudaPointerAttributes attr;
    std::vector<float> tmp;
    const float *src = transforms;
    if (cudaPointerGetAttributes(&attr, transforms) == cudaSuccess && attr.type == cudaMemoryTypeDevice) {
        tmp.resize(12 * n);
        RT_CUDA(ctx, cudaMemcpyAsync(tmp.data(), transforms, 12 * n * sizeof(float), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
        RT_CUDA(ctx, cudaStreamSynchronize((cudaStream_t)stream));
        src = tmp.data();
    } else {
        cudaGetLastError();
    }
    for (size_t i = 0; i < n; i++) {
        InstDev &in = s.insts[first + i];
        memcpy(in.m, src + 12 * i, sizeof in.m);
        in.identity = linear_is_identity(in.m) ? 1u : 0u;
    }
    // values set here replace any device-resident range that covers the same instances
    for (size_t k = s.dev_tf.size(); k-- > 0;)
        if (s.dev_tf[k].first < first + n && first < s.dev_tf[k].first + s.dev_tf[k].n) s.dev_tf.erase(s.dev_tf.begin() + k);
    s.built = false;
    return RT_OK;
}

int rt_scene_set_transforms_device(rt_scene scene, const float *dev_transforms, size_t first, size_t n) {
    if (!scene || (n && !dev_transforms)) return RT_ERR_INVALID;
    Scene &s = scene->s;
    if (first + n > s.insts.size()) return set_error(s.ctx, RT_ERR_INVALID, "instances [%zu,%zu) out of range", first, first + n);
    for (size_t k = s.dev_tf.size(); k-- > 0;)
        if (s.dev_tf[k].first < first + n && first < s.dev_tf[k].first + s.dev_tf[k].n) s.dev_tf.erase(s.dev_tf.begin() + k);
    if (n) s.dev_tf.push_back(rtb::DevTransforms{dev_transforms, first, n});
    s.built = false;
    return RT_OK;
}

int rt_accel_build(rt_scene scene, void *stream) {
    if (!scene) return RT_ERR_INVALID;
    DeviceGuard g(scene->s.ctx->device);
    return build_accel(&scene->s, (cudaStream_t)stream);
}

int rt_render(rt_scene scene, const rt_camera *camera, const rt_frame_params *params, const rt_outputs *out, void *stream) {
    if (!scene || !camera || !params || !out) return RT_ERR_INVALID;
    DeviceGuard g(scene->s.ctx->device);
    return render_frame(&scene->s, camera, params, out, (cudaStream_t)stream);
}

int rt_render_primary(rt_scene scene, const rt_camera *camera, uint32_t width, uint32_t height, size_t first_pixel,
                      size_t n_pixels, float *dev_hits, void *stream) {
    if (!scene || !camera || (n_pixels && !dev_hits)) return RT_ERR_INVALID;
    DeviceGuard g(scene->s.ctx->device);
    return render_primary(&scene->s, camera, width, height, first_pixel, n_pixels, dev_hits, (cudaStream_t)stream);
}

int rt_scene_set_primary_hits(rt_scene scene, const float *dev_hits) {
    if (!scene) return RT_ERR_INVALID;
    scene->s.ext_phit = reinterpret_cast<const float4 *>(dev_hits);
    return RT_OK;
}

int rt_resolve(rt_context ctx, const float *dev_accum, size_t n_pixels, float *dev_color_srgb, uint8_t *dev_rgba8, void *stream) {
    if (!ctx || !dev_accum) return RT_ERR_INVALID;
    DeviceGuard g(ctx->c.device);
    return resolve(&ctx->c, dev_accum, n_pixels, dev_color_srgb, dev_rgba8, (cudaStream_t)stream);
}

int rt_spheres_advect(rt_context ctx, const float *dev_base_centers, const float *dev_base_radii, const float *dev_velocities,
                      size_t n, float duration, size_t frames_this_file, size_t frame, const float particle_shift[3],
                      const float particle_scale[3], float *dev_centers_out, float *dev_radii_out, void *stream) {
    if (!ctx || !particle_shift || !particle_scale || !frames_this_file) return RT_ERR_INVALID;
    if (n && (!dev_base_centers || !dev_base_radii || !dev_velocities || !dev_centers_out || !dev_radii_out)) return RT_ERR_INVALID;
    if (particle_scale[0] != particle_scale[1] || particle_scale[0] != particle_scale[2])
        return rtb::set_error(&ctx->c, RT_ERR_UNSUPPORTED, "rt_spheres_advect: non-uniform particle scale on spheres");
    DeviceGuard g(ctx->c.device);
    return rtb::spheres_advect(&ctx->c, dev_base_centers, dev_base_radii, dev_velocities, n, duration, frames_this_file, frame,
                               particle_shift, particle_scale, dev_centers_out, dev_radii_out, (cudaStream_t)stream);
}

int rt_transforms_time(rt_context ctx, const float *dev_positions, const float *dev_velocities, const float *dev_quat_cur,
                       const float *dev_quat_next, size_t n, float duration, size_t frames_this_file, size_t frame,
                       const float particle_shift[3], const float particle_scale[3], float *dev_transforms_out, void *stream) {
    if (!ctx || !particle_shift || !particle_scale || !frames_this_file) return RT_ERR_INVALID;
    if (n && (!dev_positions || !dev_velocities || !dev_quat_cur || !dev_quat_next || !dev_transforms_out)) return RT_ERR_INVALID;
    DeviceGuard g(ctx->c.device);
    return rtb::transforms_time(&ctx->c, dev_positions, dev_velocities, dev_quat_cur, dev_quat_next, n, duration, frames_this_file,
                                frame, particle_shift, particle_scale, dev_transforms_out, (cudaStream_t)stream);
}

int rt_peer_alloc(rt_context ctx, size_t bytes, void **dev_ptr_out, unsigned char handle_out[RT_PEER_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_PEER_HANDLE_BYTES, "IPC handle size");
    if (!ctx || !dev_ptr_out || !handle_out || !bytes) return RT_ERR_INVALID;
    DeviceGuard g(ctx->c.device);
    void *p = nullptr;
    RT_CUDA(&ctx->c, cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return rtb::set_error(&ctx->c, RT_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    }
    memcpy(handle_out, &h, sizeof(h));
    *dev_ptr_out = p;
    return RT_OK;
}

int rt_peer_free(rt_context ctx, void *dev_ptr) {
    if (!ctx) return RT_ERR_INVALID;
    if (!dev_ptr) return RT_OK;
    DeviceGuard g(ctx->c.device);
    RT_CUDA(&ctx->c, cudaFree(dev_ptr));
    return RT_OK;
}

int rt_peer_open(rt_context ctx, const unsigned char handle[RT_PEER_HANDLE_BYTES], void **dev_ptr_out) {
    if (!ctx || !handle || !dev_ptr_out) return RT_ERR_INVALID;
    DeviceGuard g(ctx->c.device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *p = nullptr;
    RT_CUDA(&ctx->c, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *dev_ptr_out = p;
    return RT_OK;
}

int rt_peer_close(rt_context ctx, void *dev_ptr) {
    if (!ctx) return RT_ERR_INVALID;
    if (!dev_ptr) return RT_OK;
    DeviceGuard g(ctx->c.device);
    RT_CUDA(&ctx->c, cudaIpcCloseMemHandle(dev_ptr));
    return RT_OK;
}

int rt_resolve_parts(rt_context ctx, const float *const *dev_parts, int n_parts, size_t first_pixel, size_t n_pixels,
                     float *dev_sum_out, float *dev_color_srgb, uint8_t *dev_rgba8, void *stream) {
    if (!ctx || !dev_parts || n_parts < 1 || n_parts > RT_MAX_PARTS) return RT_ERR_INVALID;
    for (int j = 0; j < n_parts; j++)
        if (!dev_parts[j]) return RT_ERR_INVALID;
    DeviceGuard g(ctx->c.device);
    return resolve_parts(&ctx->c, dev_parts, n_parts, first_pixel, n_pixels, dev_sum_out, dev_color_srgb, dev_rgba8,
                         (cudaStream_t)stream);
}

int rt_trace_rays(rt_scene scene, const float *dev_rays, const uint32_t *dev_skip, size_t n, float *dev_t, uint32_t *dev_id, void *stream) {
    if (!scene || (n && (!dev_rays || !dev_t || !dev_id))) return RT_ERR_INVALID;
    DeviceGuard g(scene->s.ctx->device);
    return trace_rays(&scene->s, dev_rays, dev_skip, n, dev_t, dev_id, (cudaStream_t)stream);
}

int rt_trace_rays_static(rt_scene scene, const float *dev_rays, const uint32_t *dev_skip, size_t n, float *dev_t, uint32_t *dev_id, void *stream) {
    if (!scene || (n && (!dev_rays || !dev_t || !dev_id))) return RT_ERR_INVALID;
    DeviceGuard g(scene->s.ctx->device);
    return trace_rays_static(&scene->s, dev_rays, dev_skip, n, dev_t, dev_id, (cudaStream_t)stream);
}

int rt_get_stats(rt_scene scene, rt_stats *out) {
    if (!scene || !out) return RT_ERR_INVALID;
    Scene &s = scene->s;
    Context *ctx = s.ctx;
    DeviceGuard g(ctx->device);
    memset(out, 0, sizeof *out);
    RT_CUDA(ctx, cudaStreamSynchronize(s.last_stream));
    out->n_prims = s.n_prims;
    out->device_bytes = s.bytes;
    out->kernel_launches = s.launches;
    if (s.built) {
        BuildInfo bi;
        RT_CUDA(ctx, cudaMemcpy(&bi, s.info.p, sizeof bi, cudaMemcpyDeviceToHost));
        out->n_bvh_prims = bi.n_bvh;
        out->n_huge = bi.n_huge;
        out->n_nodes = s.wide ? bi.n_wnodes : (bi.n_bvh >= 2 ? bi.n_bvh - 1 : 1);
        for (int a = 0; a < 3; a++) {
            out->scene_lo[a] = bi.scene_lo[a];
            out->scene_hi[a] = bi.scene_hi[a];
        }
    }
    if (s.have_build_ev) RT_CUDA(ctx, cudaEventElapsedTime(&out->build_ms, s.ev[0], s.ev[1]));
    if (s.have_render_ev) {
        RT_CUDA(ctx, cudaEventElapsedTime(&out->render_ms, s.ev[2], s.ev[3]));
        std::vector<uint32_t> cnt(s.last_n_cnt);
        if (!cnt.empty()) RT_CUDA(ctx, cudaMemcpy(cnt.data(), s.counters.p, cnt.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost));
        uint64_t seg = s.last_primary;
        for (uint32_t c : cnt) seg += c;
        out->segments = seg;
        unsigned long long tot = 0;
        if (s.seg_total.p) RT_CUDA(ctx, cudaMemcpy(&tot, s.seg_total.p, sizeof tot, cudaMemcpyDeviceToHost));
        out->segments_total = tot;
        for (size_t i = 0; i + 1 < s.prof_used; i += 2) {
            float ms = 0.0f;
            RT_CUDA(ctx, cudaEventElapsedTime(&ms, s.prof_ev[i], s.prof_ev[i + 1]));
            out->trace_ms += ms;
            out->trace_launches += 1;
        }
        s.prof_used = 0;
    }
    return RT_OK;
}

int rt_accel_download(rt_scene scene, float *nodes_h, float *slot_geom_h, uint32_t *slot_prim_h, uint32_t *sorted_keys_h) {
    if (!scene) return RT_ERR_INVALID;
    Scene &s = scene->s;
    Context *ctx = s.ctx;
    if (!s.built) return set_error(ctx, RT_ERR_STATE, "rt_accel_download before rt_accel_build");
    DeviceGuard g(ctx->device);
    RT_CUDA(ctx, cudaStreamSynchronize(s.last_stream));
    BuildInfo bi;
    RT_CUDA(ctx, cudaMemcpy(&bi, s.info.p, sizeof bi, cudaMemcpyDeviceToHost));
    const size_t nn = bi.n_bvh >= 2 ? bi.n_bvh - 1 : 1;
    if (nodes_h) RT_CUDA(ctx, cudaMemcpy(nodes_h, s.nodes.p, nn * sizeof(Node), cudaMemcpyDeviceToHost));
    if (s.n_prims) {
        if (slot_geom_h) RT_CUDA(ctx, cudaMemcpy(slot_geom_h, s.slot_geom.p, s.n_prims * sizeof(float4), cudaMemcpyDeviceToHost));
        if (slot_prim_h) RT_CUDA(ctx, cudaMemcpy(slot_prim_h, s.slot_prim.p, s.n_prims * sizeof(uint32_t), cudaMemcpyDeviceToHost));
        if (sorted_keys_h) RT_CUDA(ctx, cudaMemcpy(sorted_keys_h, s.keys[0].p, s.n_prims * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    }
    return RT_OK;
}

int rt_accel_download_wide(rt_scene scene, float *wnodes_h, float *slot_geom_h, uint32_t *slot_prim_h, uint64_t *n_wnodes_out) {
    if (!scene) return RT_ERR_INVALID;
    Scene &s = scene->s;
    Context *ctx = s.ctx;
    if (!s.built || !s.wide) return set_error(ctx, RT_ERR_STATE, "rt_accel_download_wide needs a built scene with option wide=1");
    DeviceGuard g(ctx->device);
    RT_CUDA(ctx, cudaStreamSynchronize(s.last_stream));
    BuildInfo bi;
    RT_CUDA(ctx, cudaMemcpy(&bi, s.info.p, sizeof bi, cudaMemcpyDeviceToHost));
    if (n_wnodes_out) *n_wnodes_out = bi.n_wnodes;
    if (wnodes_h && bi.n_wnodes) RT_CUDA(ctx, cudaMemcpy(wnodes_h, s.wnodes.p, (size_t)bi.n_wnodes * sizeof(WNode), cudaMemcpyDeviceToHost));
    if (s.n_prims) {
        if (slot_geom_h) RT_CUDA(ctx, cudaMemcpy(slot_geom_h, s.wslot_geom.p, s.n_prims * sizeof(float4), cudaMemcpyDeviceToHost));
        if (slot_prim_h) RT_CUDA(ctx, cudaMemcpy(slot_prim_h, s.wslot_prim.p, s.n_prims * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    }
    return RT_OK;
}

}  // extern "C"
