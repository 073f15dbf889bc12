/*
 * ref_shim.cu — TEST INFRASTRUCTURE ONLY (oracle/_ref).
 *
 * Compiles pieces of the UNMODIFIED reference, where they lie under /root/reference, as HOST code
 * and exports them with a C ABI so tests can pin the oracle's restatement against the real thing:
 *   include/Global/DeviceFunctions.cuh             (whole header: maths, colour, transforms, RNG use)
 *   src/GraphicsAPI/SDL_GraphicsWindow.cu:4-14     SDL_GraphicsWindowConfigureCamera
 *   include/GraphicsAPI/SDL_GraphicsWindow.cuh:15-29 (camera struct + API enum)
 *   src/Global/RendererTime.cu:294-370             slerp, quatToEuler
 * The line ranges are extracted at build time by oracle/Makefile into oracle/_ref/gen/ (git-ignored);
 * no reference source is stored in this repository.  The reference's __device__-only helpers
 * (randomDouble, randomSpaceVector) are made host-callable by re-defining __device__ for this TU.
 */
#include <cuda_runtime.h>
#include <cmath>
#include <cstring>
#include <cstdint>
#include <algorithm>

// make the reference's __device__ functions (and cuRAND's) callable on the host in this TU
#undef __device__
#define __device__ __location__(host) __location__(device)
#define QUALIFIERS static __forceinline__ __host__ __location__(device)
#include <curand_kernel.h>
#include <Global/DeviceFunctions.cuh>

namespace project {
#include "_ref/gen/camera_types.inc"
#include "_ref/gen/camera_fn.inc"
#include "_ref/gen/slerp_euler.inc"
}  // namespace project

using namespace project;

extern "C" {
void ref_construct_transform(const float s[3], const float r[3], const float sc[3], float out[12]) {
    MathHelper::constructTransformMatrix(float3{s[0], s[1], s[2]}, float3{r[0], r[1], r[2]},
                                         float3{sc[0], sc[1], sc[2]}, out);
}
void ref_color_to_float4(const float c[3], float out[4]) {
    const float4 r = colorToFloat4(float3{c[0], c[1], c[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void ref_color_to_uchar4(const float c[3], uint8_t out[4]) {
    const uchar4 r = colorToUchar4(float3{c[0], c[1], c[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void ref_normalize(const float v[3], float out[3]) {
    const float3 r = normalize(float3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
void ref_camera_configure(const float c[3], const float t[3], const float up[3], int is_opengl, float out[12]) {
    const auto cam = SDL_GraphicsWindowConfigureCamera(
        float3{c[0], c[1], c[2]}, float3{t[0], t[1], t[2]}, float3{up[0], up[1], up[2]},
        is_opengl ? SDL_GraphicsWindowAPIType::OPENGL : SDL_GraphicsWindowAPIType::VULKAN);
    out[0] = cam.cameraCenter.x; out[1] = cam.cameraCenter.y; out[2] = cam.cameraCenter.z;
    out[3] = cam.cameraU.x; out[4] = cam.cameraU.y; out[5] = cam.cameraU.z;
    out[6] = cam.cameraV.x; out[7] = cam.cameraV.y; out[8] = cam.cameraV.z;
    out[9] = cam.cameraW.x; out[10] = cam.cameraW.y; out[11] = cam.cameraW.z;
}
void ref_slerp(const float q1[4], const float q2[4], float t, float out[4]) {
    const float4 r = slerp(float4{q1[0], q1[1], q1[2], q1[3]}, float4{q2[0], q2[1], q2[2], q2[3]}, t);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void ref_quat_to_euler(const float q[4], float out[3]) {
    const float3 r = quatToEuler(float4{q[0], q[1], q[2], q[3]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// randomSpaceVector driven by a host-side XORWOW state initialised like HostFunctions.cu:126
void ref_random_space_vectors(uint64_t seed, uint64_t subsequence, uint64_t offset, size_t n, float *out) {
    curandState s;
    curand_init(seed, subsequence, offset, &s);
    for (size_t i = 0; i < n; i++) {
        const float3 v = randomSpaceVector(&s, 1.0f);
        out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
    }
}
void ref_constants(float out[3]) {
    out[0] = FLOAT_ZERO_VALUE; out[1] = FLOAT_INFINITY_VALUE; out[2] = PI;
}
}
