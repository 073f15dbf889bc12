/*
 * ref_shader.cpp — TEST INFRASTRUCTURE ONLY (oracle/_ref/librt_ref_shader.so).
 *
 * Compiles the UNMODIFIED reference device programs, where they lie under /root/reference, as HOST
 * code with g++ and runs them:
 *   shader/Shader.cu                         (whole file: setPayload/getPayload/rayTrace :8-92,
 *                                             closesthitImpl :94-242, __raygen__raygenProgram :246-273,
 *                                             __miss__missProgram :276-287, the four __closesthit__* PGs)
 *   include/Global/Shader.cuh                (rayTraceDepth, GlobalParams, RayGenParams, MissParams,
 *                                             HitGroupParams)
 *   include/Global/DeviceFunctions.cuh       (vector maths, normalize, colorToFloat4, randomSpaceVector)
 *   include/Util/ColorRamp.cuh               (colorStopsForPreset, bakeColorRamp, resolveColorRampPreset)
 * No reference source is stored in this repository: the files are #included from the reference
 * checkout through -I paths (oracle/Makefile).  OptiX itself is absent and closed source; its device
 * API is emulated by oracle/ref_stub/optix_device.h (thread-local invocation frames) and
 * Global/HostFunctions.cuh (which drags in SDL2/Vulkan) by the stub of the same name.
 *
 * What is the reference's and what is ours in a run of ref_render():
 *   reference: ray generation, payload passing, the recursion, depth cut-off, hit point, sphere and
 *              triangle shading normals (object-space data: quirks 1, 2), rough / metal scatter,
 *              direction guards, albedo products, AOV handling, miss colour, sRGB encoding, XORWOW
 *              draws (cuRAND's own header), colour ramps.
 *   ours:      the closest-hit QUERY behind optixTrace (OptiX's traversal + built-in intersectors are
 *              not source): the callback below calls the oracle's orc_closest_hit() with tmin/tmax as
 *              passed by the reference and NO self-exclusion; the SBT record lookup (instance ->
 *              HitGroupParams, src/Global/RendererImpl.cu:472-575); the RNG seeding (the reference
 *              seeds with tid ^ clock64(), src/Global/HostFunctions.cu:126: quirk 7) = the oracle's
 *              stream(p, s) = curand_init(seed, p, (sample_offset + s) * 32).
 */
#define QUALIFIERS static inline
#include <cuda_runtime.h>
#include <curand_kernel.h>

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include <optix_device.h>  // oracle/ref_stub

namespace ref_optix {
int g_rsqrt_mode = 0;
thread_local Frame *t_cur = nullptr;
}  // namespace ref_optix

// ---- the reference, unmodified
#include <Shader.cu>            // -I $(REF)/shader
#include <Util/ColorRamp.cuh>   // -I $(REF)/include (its <Global/HostFunctions.cuh> resolves to ref_stub/)

extern "C" {
project::GlobalParams params;  // the launch-parameter block Shader.cu declares (OPTIX_GLOBAL_PARAM)
}

#include "rt_oracle.h"

namespace {

typedef int (*closest_hit_fn)(const orc_scene *, const void *, const float *, const float *, float, float, uint32_t,
                              float *, uint32_t *, float *, float *);

struct RefScene {
    const orc_scene *sc;
    const void *accel;
    closest_hit_fn hit;
    // per-GAS arrays as the reference uploads them (RendererImpl.cu:124-135, 149-159): float3 centres +
    // float radii of all spheres, float3 vertex normals (3 per triangle) of all triangles.  With
    // orc_scene.sphere_shade these are the OBJECT-space values the shader really reads.
    std::vector<float3> centers;
    std::vector<float> radii;
    std::vector<float3> normals;
    project::MissParams miss;
};
thread_local const RefScene *t_scene = nullptr;
thread_local float t_top_result[4];  // payload p0..p3 returned by the depth-0 optixTrace (linear radiance)
thread_local uint64_t t_segments = 0;

}  // namespace

// optixTrace: closest hit, then the hit group's closest-hit program or the miss program runs on the
// caller's payload registers (OptiX programming model; SBT record = instance, RendererImpl.cu:472-575).
void ref_optix::trace(OptixTraversableHandle, float3 o, float3 d, float tmin, float tmax, unsigned int (&p)[12]) {
    const RefScene *rs = t_scene;
    Frame *caller = t_cur;
    Frame f;
    std::memcpy(f.p, p, sizeof(f.p));
    f.launch_index = caller->launch_index;
    f.launch_dims = caller->launch_dims;
    f.ray_o = o;
    f.ray_d = d;
    f.depth = caller->depth + 1;
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    float t = 0.0f, u = 0.0f, v = 0.0f;
    uint32_t id = ORC_MISS;
    const int got = rs->hit(rs->sc, rs->accel, oo, dd, tmin, tmax, ORC_MISS, &t, &id, &u, &v);
    t_segments += 1;
    t_cur = &f;
    if (!got) {
        f.sbt_data = &rs->miss;
        f.tmax = tmax;
        __miss__missProgram();
    } else {
        const orc_scene *sc = rs->sc;
        const uint32_t ki = sc->kind_index[id], idx = ki & 0x7FFFFFFFu;
        const uint32_t m = sc->mat[id], mi = m & 0x7FFFFFFFu;
        const bool sphere = (ki >> 31) == ORC_KIND_SPHERE, rough = (m >> 31) == ORC_MAT_ROUGH;
        project::HitGroupParams hg;
        std::memset(&hg, 0, sizeof(hg));
        if (sphere) {
            hg.sphere.centers = const_cast<float3 *>(rs->centers.data());
            hg.sphere.radii = const_cast<float *>(rs->radii.data());
        } else {
            hg.triangles.vertexNormals = const_cast<float3 *>(rs->normals.data());
        }
        if (rough) {
            hg.rough.albedo = make_float3(sc->roughs[3 * (size_t)mi], sc->roughs[3 * (size_t)mi + 1], sc->roughs[3 * (size_t)mi + 2]);
        } else {
            const float *mt = sc->metals + 4 * (size_t)mi;
            hg.metal.albedo = make_float3(mt[0], mt[1], mt[2]);
            hg.metal.fuzz = mt[3];
        }
        f.sbt_data = &hg;
        f.tmax = t;
        f.prim_index = idx;
        f.bary = make_float2(u, v);
        if (sphere) {
            if (rough) __closesthit__closesthitProgramSphereRough(); else __closesthit__closesthitProgramSphereMetal();
        } else {
            if (rough) __closesthit__closesthitProgramTriangleRough(); else __closesthit__closesthitProgramTriangleMetal();
        }
    }
    std::memcpy(p, f.p, sizeof(f.p));
    t_cur = caller;
    if (caller->depth == 0) {
        for (int i = 0; i < 4; i++) t_top_result[i] = __uint_as_float(p[i]);
    }
}

extern "C" {

int ref_shader_trace_depth(void) { return (int)project::rayTraceDepth; }
void ref_shader_set_rsqrt_mode(int mode) { ref_optix::g_rsqrt_mode = mode; }

/* One launch of the reference's pipeline per sample s in [0, spp): optixLaunch(w, h, 1) =
 * __raygen__raygenProgram for every pixel.  Outputs (any may be NULL), all float4 per pixel:
 *   linear_sum  sum over the samples of the linear radiance the top-level optixTrace returned (w = spp)
 *   color_last  colorBuffer of the last launch (sRGB-encoded by raygen, Shader.cu:270)
 *   albedo_last / normal_last  the AOV buffers of the last launch (Shader.cu:271-272)
 * hit_fn = address of orc_closest_hit, accel = orc_accel_create() handle or NULL. */
int ref_shader_render(const orc_scene *sc, const void *accel, void *hit_fn, const orc_camera *cam, const orc_params *p,
                      float *linear_sum, float *color_last, float *albedo_last, float *normal_last, uint64_t *segments) {
    if (!sc || !cam || !p || !hit_fn || !p->width || !p->height) return 1;
    if (p->max_depth != project::rayTraceDepth) return 2;  // the reference's depth is a compile-time constant
    RefScene rs;
    rs.sc = sc;
    rs.accel = accel;
    rs.hit = reinterpret_cast<closest_hit_fn>(hit_fn);
    uint32_t n_sph = 0, n_tri = 0;
    for (uint32_t g = 0; g < sc->n_prims; g++) {
        const uint32_t ki = sc->kind_index[g], idx = ki & 0x7FFFFFFFu;
        if ((ki >> 31) == ORC_KIND_SPHERE) n_sph = std::max(n_sph, idx + 1); else n_tri = std::max(n_tri, idx + 1);
    }
    const float *sph = sc->sphere_shade ? sc->sphere_shade : sc->spheres;
    rs.centers.resize(n_sph);
    rs.radii.resize(n_sph);
    for (uint32_t i = 0; i < n_sph; i++) {
        rs.centers[i] = make_float3(sph[4 * (size_t)i], sph[4 * (size_t)i + 1], sph[4 * (size_t)i + 2]);
        rs.radii[i] = sph[4 * (size_t)i + 3];
    }
    rs.normals.resize(3 * (size_t)n_tri);
    for (size_t i = 0; i < 3 * (size_t)n_tri; i++)
        rs.normals[i] = make_float3(sc->tri_normals[3 * i], sc->tri_normals[3 * i + 1], sc->tri_normals[3 * i + 2]);
    rs.miss.backgroundColor = make_float3(p->background[0], p->background[1], p->background[2]);

    const size_t npix = (size_t)p->width * p->height;
    std::vector<curandState> states(npix);
    std::vector<float4> color(npix), albedo(npix), normal(npix);
    params.handle = 1;
    params.stateArray = states.data();
    project::RayGenParams rg;
    rg.width = p->width;
    rg.height = p->height;
    rg.colorBuffer = color.data();
    rg.albedoBuffer = albedo.data();
    rg.normalBuffer = normal.data();
    rg.cameraCenter = make_float3(cam->center[0], cam->center[1], cam->center[2]);
    rg.cameraU = make_float3(cam->U[0], cam->U[1], cam->U[2]);
    rg.cameraV = make_float3(cam->V[0], cam->V[1], cam->V[2]);
    rg.cameraW = make_float3(cam->W[0], cam->W[1], cam->W[2]);
    if (linear_sum) std::memset(linear_sum, 0, npix * 16);
    uint64_t total_segments = 0;
    const long H = (long)p->height;
    for (uint32_t s = 0; s < p->spp; s++) {
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : total_segments)
        for (long y = 0; y < H; y++) {
            t_scene = &rs;
            for (uint32_t x = 0; x < p->width; x++) {
                const size_t pix = (size_t)y * p->width + x;
                curand_init(p->seed, pix, (unsigned long long)(p->sample_offset + s) * ORC_RNG_SAMPLE_STRIDE, &states[pix]);
                ref_optix::Frame f;
                std::memset(&f, 0, sizeof(f));
                f.launch_index = make_uint3(x, (unsigned)y, 0);
                f.launch_dims = make_uint3(p->width, p->height, 1);
                f.sbt_data = &rg;
                f.depth = 0;
                ref_optix::t_cur = &f;
                t_segments = 0;
                __raygen__raygenProgram();
                ref_optix::t_cur = nullptr;
                total_segments += t_segments;
                if (linear_sum) {
                    linear_sum[4 * pix + 0] += t_top_result[0];
                    linear_sum[4 * pix + 1] += t_top_result[1];
                    linear_sum[4 * pix + 2] += t_top_result[2];
                    linear_sum[4 * pix + 3] += 1.0f;
                }
            }
        }
    }
    if (color_last) std::memcpy(color_last, color.data(), npix * 16);
    if (albedo_last) std::memcpy(albedo_last, albedo.data(), npix * 16);
    if (normal_last) std::memcpy(normal_last, normal.data(), npix * 16);
    if (segments) *segments = total_segments;
    return 0;
}

/* include/Util/ColorRamp.cuh: bakeColorRamp(colorStopsForPreset(resolveColorRampPreset(name)), count) */
int ref_bake_color_ramp(const char *preset, size_t count, float *out_rgb) {
    const auto colors = project::bakeColorRamp(project::colorStopsForPreset(project::resolveColorRampPreset(std::string(preset))), count);
    for (size_t i = 0; i < colors.size(); i++) {
        out_rgb[3 * i] = colors[i].x;
        out_rgb[3 * i + 1] = colors[i].y;
        out_rgb[3 * i + 2] = colors[i].z;
    }
    return (int)colors.size() == (int)count ? 0 : 1;
}
const char *ref_preset_name(const char *preset) { return project::presetName(project::resolveColorRampPreset(std::string(preset))); }

}  // extern "C"
