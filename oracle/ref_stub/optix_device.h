/*
 * optix_device.h — HOST stand-in for the OptiX 9 device API.  TEST INFRASTRUCTURE ONLY (oracle/_ref).
 *
 * The OptiX SDK is absent from this image and closed source.  The reference's device programs
 * (shader/Shader.cu, include/Global/Shader.cuh, include/Global/DeviceFunctions.cuh) touch OptiX only
 * through the intrinsics below (Shader.cu:8-92, 95-118, 148, 246-287), all of which are plain
 * accessors of "the current program invocation".  This header gives them a thread-local invocation
 * frame so that the UNMODIFIED Shader.cu compiles with g++ as host code (oracle/ref_shader.cpp,
 * recipe in oracle/Makefile) and its closest-hit / miss / raygen programs can be executed and
 * compared with the oracle's restatement.  The one thing OptiX does that is NOT reference source —
 * BVH traversal + the built-in sphere/triangle intersectors behind optixTrace — is delegated to a
 * callback (ref_optix::g_trace) which ref_shader.cpp points at the oracle's closest_hit().
 *
 * Under nvcc (oracle/ref_shim.cu compiles DeviceFunctions.cuh that way) this header stays empty: that
 * translation unit uses no OptiX call.
 */
#pragma once
#ifndef __CUDACC__

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>

// ---- CUDA device built-ins the reference's headers use, as host functions
using std::abs;
using std::isfinite;
using std::max;
using std::min;
static inline unsigned int __float_as_uint(float f) { unsigned int u; std::memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(unsigned int u) { float f; std::memcpy(&f, &u, 4); return f; }
namespace ref_optix {
// 0: IEEE single 1/sqrt(x) (what the oracle and the CUDA kernels use: DESIGN.md §2 item 6)
// 1: (float)(1.0 / sqrt((double)x)), CUDA's host rsqrtf (crt/math_functions.hpp) and within the
//    <= 2 ulp of the device's MUFU.RSQ-based rsqrtf — used to measure the size of that deviation
extern int g_rsqrt_mode;
}  // namespace ref_optix
static inline float rsqrtf(float x) {
    if (ref_optix::g_rsqrt_mode == 1) return static_cast<float>(1.0 / std::sqrt(static_cast<double>(x)));
    return 1.0f / std::sqrt(x);
}

// ---- OptiX types used by Shader.cu / Shader.cuh
typedef unsigned long long OptixTraversableHandle;
typedef unsigned int OptixVisibilityMask;
typedef unsigned long long CUdeviceptr_ref;
enum OptixRayFlags { OPTIX_RAY_FLAG_NONE = 0u };

namespace ref_optix {

// one program invocation (what the optixGet*/optixSet* intrinsics of that program see)
struct Frame {
    unsigned int p[12];         // payload registers p0..p11 (shared with the optixTrace caller)
    uint3 launch_index, launch_dims;
    const void *sbt_data;       // the SBT record's data block of the running program
    float3 ray_o, ray_d;        // world-space ray of the optixTrace call
    float tmax;                 // closest-hit: the hit distance (optixGetRayTmax)
    unsigned int prim_index;    // index inside the hit GAS
    float2 bary;                // triangle barycentrics (u, v) of vertices 1 and 2
    int depth;                  // 0 = raygen
};
extern thread_local Frame *t_cur;

// traversal + program dispatch, implemented by oracle/ref_shader.cpp
void trace(OptixTraversableHandle handle, float3 o, float3 d, float tmin, float tmax, unsigned int (&p)[12]);

}  // namespace ref_optix

#define REF_OPTIX_PAYLOAD(N)                                                                  \
    static inline unsigned int optixGetPayload_##N() { return ref_optix::t_cur->p[N]; }      \
    static inline void optixSetPayload_##N(unsigned int v) { ref_optix::t_cur->p[N] = v; }
REF_OPTIX_PAYLOAD(0) REF_OPTIX_PAYLOAD(1) REF_OPTIX_PAYLOAD(2) REF_OPTIX_PAYLOAD(3)
REF_OPTIX_PAYLOAD(4) REF_OPTIX_PAYLOAD(5) REF_OPTIX_PAYLOAD(6) REF_OPTIX_PAYLOAD(7)
REF_OPTIX_PAYLOAD(8) REF_OPTIX_PAYLOAD(9) REF_OPTIX_PAYLOAD(10) REF_OPTIX_PAYLOAD(11)
#undef REF_OPTIX_PAYLOAD

static inline uint3 optixGetLaunchIndex() { return ref_optix::t_cur->launch_index; }
static inline uint3 optixGetLaunchDimensions() { return ref_optix::t_cur->launch_dims; }
static inline CUdeviceptr_ref optixGetSbtDataPointer() { return reinterpret_cast<CUdeviceptr_ref>(ref_optix::t_cur->sbt_data); }
static inline float optixGetRayTmax() { return ref_optix::t_cur->tmax; }
static inline float3 optixGetWorldRayOrigin() { return ref_optix::t_cur->ray_o; }
static inline float3 optixGetWorldRayDirection() { return ref_optix::t_cur->ray_d; }
static inline unsigned int optixGetPrimitiveIndex() { return ref_optix::t_cur->prim_index; }
static inline float2 optixGetTriangleBarycentrics() { return ref_optix::t_cur->bary; }

// optixTrace(handle, origin, direction, tmin, tmax, rayTime, visibilityMask, rayFlags, SBToffset, SBTstride,
//            missSBTIndex, payload...) — Shader.cu:70-76 passes 12 payload registers, mask 1, no flags, 0/0/0.
template <typename... P>
static inline void optixTrace(OptixTraversableHandle handle, float3 o, float3 d, float tmin, float tmax, float /*rayTime*/,
                              OptixVisibilityMask /*mask*/, unsigned int /*flags*/, unsigned int /*sbtOffset*/,
                              unsigned int /*sbtStride*/, unsigned int /*missIndex*/, P &...payload) {
    static_assert(sizeof...(P) == 12, "the reference traces with 12 payload registers");
    unsigned int *ptr[12] = {&payload...};
    unsigned int regs[12];
    for (int i = 0; i < 12; i++) regs[i] = *ptr[i];
    ref_optix::trace(handle, o, d, tmin, tmax, regs);
    for (int i = 0; i < 12; i++) *ptr[i] = regs[i];
}

#endif  // !__CUDACC__
