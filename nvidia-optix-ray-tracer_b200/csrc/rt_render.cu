// rt_render.cu — wavefront path tracer kernels for sm_100a.
//
// Replaces the OptiX programs of the reference:
//   __raygen__raygenProgram            shader/Shader.cu:246-273   -> k_raygen / k_prepare / k_shade_first
//   closesthitImpl (4 closest-hit PGs) shader/Shader.cu:94-242    -> scatter()
//   __miss__missProgram                shader/Shader.cu:276-287   -> miss branches
//   __initializeDeviceRandomGenerators src/Global/HostFunctions.cu:122-136 -> k_rng_init
//   colorToFloat4 / convertFloat4ToUchar4Kernel  DeviceFunctions.cuh:153-212,
//                                      src/Global/RendererImpl.cu:672-678 -> k_resolve
// The reference's recursion (optixTrace inside closest-hit, depth <= rayTraceDepth) becomes an
// iterative wavefront: one kernel per bounce over a compacted SoA ray queue.
#include <curand_kernel.h>

#include <algorithm>

#include "rt_device.cuh"

namespace rtb {

struct CamDev {
    F3 c, U, V, W;
};

struct Queue {
    float4 *o;    // origin.xyz, bits(path id)
    float4 *d;    // dir.xyz, bits(skip slot)
    float4 *t;    // throughput.rgb, bits(rng.d)
    uint4 *r;     // rng v0..v3
    uint32_t *r4; // rng v4
    uint4 *m;     // RT_FLAG_REF_MUL_ON_RETURN only (else NULL): material words of the path's bounces 1..4
};

// ------------------------------------------------------------------ pixel <-> thread mapping (8x4 tiles)
__device__ __forceinline__ bool tile_pixel(uint32_t tid, uint32_t W, uint32_t H, uint32_t &x, uint32_t &y) {
    const uint32_t tiles_x = (W + 7u) >> 3;
    const uint32_t tile = tid >> 5, lane = tid & 31u;
    x = (tile % tiles_x) * 8u + (lane & 7u);
    y = (tile / tiles_x) * 4u + (lane >> 3);
    return x < W && y < H;
}

// shader/Shader.cu:249-261
__device__ __forceinline__ Ray primary_ray(const CamDev &cam, uint32_t x, uint32_t y, uint32_t W, uint32_t H) {
    const float nx = (((float)x + 0.5f) / (float)W) * 2.0f - 1.0f;
    const float ny = (((float)y + 0.5f) / (float)H) * 2.0f - 1.0f;
    const float aspect = (float)W / (float)H;
    Ray r;
    r.o = cam.c;
    r.d = normalize3(nx * aspect * cam.U + ny * cam.V + cam.W);
    r.tmin = RT_ZERO;
    r.tmax = RT_INF;
    return r;
}

// ------------------------------------------------------------------ RNG state init (HostFunctions.cu:122-127)
// seed fixed by the caller (the reference xors clock64() in: quirk 7), subsequence = pixel,
// offset = sample_offset * RT_RNG_SAMPLE_STRIDE; bounds-checked and unpadded (quirk 8).
// cuRAND's sequence skip costs ~20 GF(2) matrix-vector products (160x160 bits each) per state.  The
// states of consecutive pixels differ by ONE application of the sequence-skip matrix
// S = precalc_xorwow_matrix[0] (state(p+1) = S * state(p): all skips are powers of the same one-step
// matrix and commute), so each thread pays curand_init once, for the first pixel of a run of
// kRngRun, and then walks the run with one matrix-vector product per pixel.  GF(2) arithmetic is
// exact: the states are bit-identical to curand_init(seed, p, offset) (checked against the oracle,
// which calls the real thing per pixel).  d (the Weyl counter) depends on seed and offset only.
constexpr uint32_t kRngRun = 64;
constexpr uint32_t kRngMaxLag = 2048;  // samples; beyond it curand_init (sequence skip by matrices) is cheaper than stepping
__global__ void __launch_bounds__(128)
k_rng_init(uint32_t *__restrict__ rng, uint32_t *__restrict__ rng_off, uint32_t off_samples, uint32_t npix, unsigned long long seed,
           unsigned long long offset) {
    __shared__ uint4 s_row[160][2];  // row i of S padded to 8 words: one 128-bit + one 32-bit shared load
    for (uint32_t i = threadIdx.x; i < 160; i += blockDim.x) {
        const unsigned int *m = &precalc_xorwow_matrix[0][5 * i];
        s_row[i][0] = make_uint4(m[0], m[1], m[2], m[3]);
        s_row[i][1] = make_uint4(m[4], 0u, 0u, 0u);
    }
    __syncthreads();
    const uint32_t p0 = (blockIdx.x * blockDim.x + threadIdx.x) * kRngRun;
    if (p0 >= npix) return;
    curandStateXORWOW st;
    curand_init(seed, (unsigned long long)p0, offset, &st);
    uint32_t v[5] = {st.v[0], st.v[1], st.v[2], st.v[3], st.v[4]};
    const uint32_t end = min(p0 + kRngRun, npix);
    for (uint32_t p = p0; p < end; p++) {
        rng[0 * (size_t)npix + p] = st.d;
        rng[1 * (size_t)npix + p] = v[0];
        rng[2 * (size_t)npix + p] = v[1];
        rng[3 * (size_t)npix + p] = v[2];
        rng[4 * (size_t)npix + p] = v[3];
        rng[5 * (size_t)npix + p] = v[4];
        rng_off[p] = off_samples;  // the sample this pixel's state stands at (states advance lazily, per pixel)
        if (p + 1 == end) break;
        uint32_t y0 = 0, y1 = 0, y2 = 0, y3 = 0, y4 = 0;
#pragma unroll
        for (int w = 0; w < 5; w++) {
            const uint32_t x = v[w];
#pragma unroll 8
            for (int j = 0; j < 32; j++) {
                const uint32_t m = 0u - ((x >> j) & 1u);
                const uint4 r0 = s_row[32 * w + j][0];
                const uint32_t r1 = s_row[32 * w + j][1].x;
                y0 ^= r0.x & m; y1 ^= r0.y & m; y2 ^= r0.z & m; y3 ^= r0.w & m; y4 ^= r1 & m;
            }
        }
        v[0] = y0; v[1] = y1; v[2] = y2; v[3] = y3; v[4] = y4;
    }
}

// ------------------------------------------------------------------ closest-hit shading (Shader.cu:94-242)
struct Scatter {
    F3 P, dir, alb, N;
    uint32_t mat;
};
__device__ __forceinline__ Scatter scatter(const SceneDev &sc, const Ray &ray, const Hit &hit, Rng &rng) {
    Scatter s;
    s.P = ray.o + hit.t * ray.d;  // :114
    const float4 g = ldg4(sc.slot_geom + hit.slot);
    const uint32_t wb = __float_as_uint(g.w);
    F3 N;
    if (!(wb & 0x80000000u)) {  // :122-136, world-space centre (compat: the object-space centre/radius of :124-125)
        float4 c = g;
        if (sc.shade_geom) c = ldg4(sc.shade_geom + __ldg(sc.slot_prim + hit.slot));
        const F3 on = (s.P - f3(c.x, c.y, c.z)) / c.w;
        N = (dot(ray.d, on) < 0.0f) ? on : -on;
    } else {  // :137-155
        const uint32_t ti = wb & 0x7FFFFFFFu;
        const float4 n1 = ldg4(sc.tri_nrm + 3 * (size_t)ti);
        const float4 n2 = ldg4(sc.tri_nrm + 3 * (size_t)ti + 1);
        const float4 n3 = ldg4(sc.tri_nrm + 3 * (size_t)ti + 2);
        const float u = hit.u, v = hit.v;
        const float w = 1.0f - u - v;
        const F3 nn = w * f3(n1.x, n1.y, n1.z) + u * f3(n2.x, n2.y, n2.z) + v * f3(n3.x, n3.y, n3.z);
        N = (dot(ray.d, nn) < 0.0f) ? nn : -nn;
    }
    const uint32_t m = __ldg(sc.slot_mat + hit.slot);
    const uint32_t mi = m & 0x7FFFFFFFu;
    F3 dir;
    if (!(m >> 31)) {  // ROUGH :169-179
        const float4 a = ldg4(sc.roughs + mi);
        dir = N + random_space_vector(rng);
        if (fabsf(dot(dir, dir) - RT_ZERO * RT_ZERO) < RT_ZERO) dir = N;
        s.alb = f3(a.x, a.y, a.z);
    } else {  // METAL :180-192
        const float4 a = ldg4(sc.metals + mi);
        const F3 vv = ray.d;
        dir = normalize3(vv - 2 * dot(vv, N) * N);
        if (a.w > 0.0f) dir = dir + a.w * random_space_vector(rng);
        s.alb = f3(a.x, a.y, a.z);
    }
    // :202-213
    if (!finite3(dir) || dot(dir, dir) <= RT_ZERO * RT_ZERO) {
        dir = N;
        if (dot(dir, dir) <= RT_ZERO * RT_ZERO || !finite3(dir)) dir = f3(0.0f, 0.0f, 1.0f);
    }
    s.dir = dir;
    s.N = N;
    s.mat = m;
    return s;
}

// RT_FLAG_REF_MUL_ON_RETURN: radiance of a finished path in the reference's rounding order — the recursion
// multiplies the albedos innermost first while it unwinds (Shader.cu:236-238): ((bg * a_n) * a_{n-1}) ... * a_1
__device__ __forceinline__ F3 albedo_of(const SceneDev &sc, uint32_t m) {
    const float4 a = (m >> 31) ? ldg4(sc.metals + (m & 0x7FFFFFFFu)) : ldg4(sc.roughs + (m & 0x7FFFFFFFu));
    return f3(a.x, a.y, a.z);
}
__device__ __forceinline__ F3 unwind_product(const SceneDev &sc, F3 bg, uint4 mats, uint32_t n) {
    F3 r = bg;
    const uint32_t mm[4] = {mats.x, mats.y, mats.z, mats.w};
#pragma unroll
    for (int k = 3; k >= 0; k--)
        if ((uint32_t)k < n) {
            const F3 a = albedo_of(sc, mm[k]);
            r = f3(r.x * a.x, r.y * a.y, r.z * a.z);
        }
    return r;
}

// warp-aggregated queue append; returns the slot for lanes with emit == true
__device__ __forceinline__ uint32_t queue_push(uint32_t *counter, bool emit) {
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, emit);
    if (!mask) return 0;
    const uint32_t lane = threadIdx.x & 31u;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, __popc(mask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    return base + __popc(mask & ((1u << lane) - 1u));
}

__device__ __forceinline__ void queue_write(const Queue &q, uint32_t i, F3 o, F3 d, F3 thr, uint32_t pid,
                                            uint32_t skip, const Rng &r) {
    q.o[i] = make_float4(o.x, o.y, o.z, __uint_as_float(pid));
    q.d[i] = make_float4(d.x, d.y, d.z, __uint_as_float(skip));
    q.t[i] = make_float4(thr.x, thr.y, thr.z, __uint_as_float(r.d));
    q.r[i] = make_uint4(r.v0, r.v1, r.v2, r.v3);
    q.r4[i] = r.v4;
}

// Block-wide binned append: the rays a block emits in one step are written to the queue grouped by the
// octant of their direction, so that the 32 lanes of a traversal warp mostly share one child-visiting
// order (extend_warp_wide walks children in slot ^ octant order) while staying close in origin (one
// block = a few neighbouring pixel tiles).  Results do not depend on queue order (closest hit is
// order-free, radiance is written by path id), so this is a pure scheduling choice.
// Must be called by every thread of the block; NW = warps per block.
constexpr uint32_t kBins = 24;  // capacity; nb = 8 (octant) or 24 (octant x dominant axis) at run time
__device__ __forceinline__ uint32_t dir_octant(F3 d) {
    return (d.x >= 0.0f ? 1u : 0u) | (d.y >= 0.0f ? 2u : 0u) | (d.z >= 0.0f ? 4u : 0u);
}
__device__ __forceinline__ uint32_t dir_bin(F3 d, uint32_t nb) {
    const uint32_t o = dir_octant(d);
    if (nb <= 8u) return o;
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    const uint32_t m = (ax >= ay && ax >= az) ? 0u : (ay >= az ? 1u : 2u);
    return o * 3u + m;
}
template <int NW>
__device__ __forceinline__ uint32_t binned_push(uint32_t *counter, bool emit, uint32_t bin, uint32_t nb, uint32_t (*s_cnt)[kBins + 1],
                                                uint32_t *s_base) {
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t key = emit ? bin : kBins;
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
    if (lane <= kBins) s_cnt[w][lane] = 0;
    __syncwarp();
    if ((int)lane == __ffs(peers) - 1) s_cnt[w][key] = __popc(peers);
    __syncthreads();
    if (w == 0) {
        uint32_t tot = 0;
        if (lane < nb) {
#pragma unroll
            for (int ww = 0; ww < NW; ww++) {
                const uint32_t c = s_cnt[ww][lane];
                s_cnt[ww][lane] = tot;
                tot += c;
            }
        }
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        uint32_t base = 0;
        if (lane == 31u && inc) base = atomicAdd(counter, inc);
        base = __shfl_sync(0xFFFFFFFFu, base, 31);
        if (lane < nb) s_base[lane] = base + inc - tot;
    }
    __syncthreads();
    const uint32_t pos = emit ? s_base[key] + s_cnt[w][key] + __popc(peers & ((1u << lane) - 1u)) : 0u;
    __syncthreads();  // s_cnt / s_base are reused by the next step
    return pos;
}

// ------------------------------------------------------------------ closest hit for a ray buffer (persistent warps)
__global__ void __launch_bounds__(128, 8)
k_extend(SceneDev sc, ExtendIn in, const uint32_t *__restrict__ n_ptr, uint32_t n_fixed, uint32_t *fetch_ctr,
         float4 *__restrict__ hits, int t1, int t2, int fetch_min) {
    const uint32_t n = n_ptr ? *n_ptr : n_fixed;
    extend_warp(sc, in, n, fetch_ctr, hits, t1, t2, fetch_min);
}

template <int kMinBlocks>
__global__ void __launch_bounds__(RT_EXTEND_THREADS, kMinBlocks)
k_extend_wide(SceneDev sc, ExtendIn in, const uint32_t *__restrict__ n_ptr, uint32_t n_fixed, uint32_t *fetch_ctr,
              float4 *__restrict__ hits, int t1, int t2, int t3, int fetch_min, uint32_t magic) {
    __shared__ uint8_t s_perm[8][256];
    __shared__ uint2 s_stack[RT_SH_STACK][RT_EXTEND_THREADS];
    fill_perm_table(s_perm);
    const uint32_t n = n_ptr ? *n_ptr : n_fixed;
    extend_warp_wide(sc, in, n, fetch_ctr, hits, t1, t2, t3, fetch_min, magic, s_perm, s_stack);
}

// ------------------------------------------------------------------ primary rays (raygen, Shader.cu:246-267)
// One ray per pixel, written in pixel order into the path-queue layout (path id = pixel, no skip).
__global__ void __launch_bounds__(128)
k_raygen(CamDev cam, uint32_t W, uint32_t H, float4 *__restrict__ qo, float4 *__restrict__ qd) {
    uint32_t x, y;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, W, H, x, y)) return;
    const uint32_t p = y * W + x;  // Shader.cu:270
    const Ray r = primary_ray(cam, x, y, W, H);
    qo[p] = make_float4(r.o.x, r.o.y, r.o.z, __uint_as_float(p));
    qd[p] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(kNoSlot));
}

// primary-hit ids and first-hit AOVs (Shader.cu:216-227, 284-286)
__global__ void __launch_bounds__(128)
k_primary_aov(SceneDev sc, CamDev cam, uint32_t W, uint32_t H, const float4 *__restrict__ phit,
              uint32_t *__restrict__ prim_id, float4 *__restrict__ albedo, float4 *__restrict__ normal) {
    uint32_t x, y;
    if (!tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, W, H, x, y)) return;
    const size_t p = (size_t)y * W + x;
    const float4 hh = phit[p];
    Hit h;
    h.t = hh.x;
    h.slot = __float_as_uint(hh.y);
    h.u = hh.z;
    h.v = hh.w;
    if (prim_id) prim_id[p] = (h.slot == kNoSlot) ? RT_MISS_ID : __ldg(sc.slot_prim + h.slot);
    if (albedo || normal) {
        float4 a = make_float4(0, 0, 0, 0), n = make_float4(0, 0, 0, 0);  // Shader.cu:284-286
        if (h.slot != kNoSlot) {
            // first-hit AOVs never consume RNG (fuzz/rough draws excluded)
            Rng dummy = {0, 1, 2, 3, 4, 5};
            const Ray r = primary_ray(cam, x, y, W, H);
            const Scatter s = scatter(sc, r, h, dummy);
            const F3 nn = normalize3(s.N);
            a = make_float4(s.alb.x, s.alb.y, s.alb.z, 1.0f);
            n = make_float4(nn.x, nn.y, nn.z, 0.0f);
        }
        if (albedo) albedo[p] = a;
        if (normal) normal[p] = n;
    }
}

// ------------------------------------------------------------------ first bounce: one thread per (hit pixel, sample)
// The reference has no sub-pixel jitter (Shader.cu:249-252), so all samples of a pixel share the primary hit; each
// sample continues with its own XORWOW stream(p, s).  k_prepare has compacted the pixels that continue a path
// (hit_pix / n_hit), settled the others in the accumulation buffer and stored the stream state at the start of every
// sample (rng_start), so the samples are independent here: every lane has work (the per-pixel loop this replaces kept
// 76 % of C4's lanes idle through 32 block-wide queue appends and wrote 3.2 GB of background colour per batch).
// Path id = i * S + s: result[] holds hit pixels only and is summed in sample order by k_accumulate_hits.
#ifndef RT_FIRST_LOG
#define RT_FIRST_LOG 8
#endif
constexpr int kFirstThreads = 1 << RT_FIRST_LOG;
__global__ void __launch_bounds__(kFirstThreads, 1024 / kFirstThreads)
k_shade_first(SceneDev sc, CamDev cam, uint32_t W, uint32_t H, uint32_t S, const float4 *__restrict__ phit,
              const uint32_t *__restrict__ hit_pix, const uint32_t *__restrict__ n_hit, const uint32_t *__restrict__ rng_start,
              uint32_t cap, Queue out, uint32_t *out_count, uint32_t nb, uint32_t flags) {
    __shared__ uint32_t s_cnt[kFirstThreads / 32][kBins + 1];
    __shared__ uint32_t s_base[kBins];
    const uint32_t n = *n_hit;
    // a block pass = 8 consecutive hit pixels x 32 samples (256 threads, S >= 32): 8 lanes read one 32-byte sector of
    // every rng_start word, and all samples of a pixel enter the queue together, next to their neighbours' (rays that
    // start from the same hit point walk the same nodes: queueing sample-major over the whole frame cost 2.5 % of
    // the traversal time)
    // (fewer samples per launch: SP = the power of two >= min(S, 32) sample slots, 512 / SP pixels per pass)
    uint32_t sp_log = 0;
    while ((1u << sp_log) < S && sp_log < 5u) sp_log++;
    const uint32_t pi_log = (uint32_t)RT_FIRST_LOG - sp_log;
    const uint32_t il = threadIdx.x & ((1u << pi_log) - 1u), sl = threadIdx.x >> pi_log;
    const uint32_t n_groups = (n + (1u << pi_log) - 1u) >> pi_log, s_passes = (S + (1u << sp_log) - 1u) >> sp_log;
    for (uint32_t w = blockIdx.x; w < n_groups * s_passes; w += gridDim.x) {  // block-uniform trip count
        const uint32_t g = w / s_passes, smp = ((w - g * s_passes) << sp_log) + sl, i = (g << pi_log) + il;
        const bool valid = i < n && smp < S;
        Scatter sc1;
        Rng r = {0, 0, 0, 0, 0, 0};
        uint32_t pid = 0, slot = kNoSlot;
        if (valid) {
            const uint32_t p = hit_pix[i];
            const Ray ray = primary_ray(cam, p % W, p / W, W, H);
            const float4 h = phit[p];
            Hit hit;
            hit.t = h.x;
            hit.slot = __float_as_uint(h.y);
            hit.u = h.z;
            hit.v = h.w;
            const uint32_t *rs = rng_start + (size_t)smp * 6u * cap + i;
            r.d = rs[0];
            r.v0 = rs[1 * (size_t)cap]; r.v1 = rs[2 * (size_t)cap]; r.v2 = rs[3 * (size_t)cap];
            r.v3 = rs[4 * (size_t)cap]; r.v4 = rs[5 * (size_t)cap];
            sc1 = scatter(sc, ray, hit, r);
            pid = i * S + smp;
            // Shader.cu:230-233: the reference relies on tmin alone; by default the ray skips the primitive it leaves
            slot = (flags & RT_FLAG_REF_NO_SELF_SKIP) ? kNoSlot : hit.slot;
        }
        const uint32_t qi = binned_push<kFirstThreads / 32>(out_count, valid, valid ? dir_bin(sc1.dir, nb) : 0u, nb, s_cnt, s_base);
        if (valid) {
            queue_write(out, qi, sc1.P, sc1.dir, sc1.alb, pid, slot, r);
            if (out.m) out.m[qi] = make_uint4(sc1.mat, 0u, 0u, 0u);
        }
    }
}

// ------------------------------------------------------------------ closest-hit / miss programs over the ray queue
// One thread per queue entry: reads the ray, its hit record (written by k_extend) and the path's RNG
// state, applies Shader.cu:94-242 / :276-287 and appends the scattered ray to the next queue.
template <int kShadeThreads, bool kRefMul>
__global__ void __launch_bounds__(kShadeThreads, kShadeThreads == 512 ? 3 : (kShadeThreads == 256 ? 6 : 1))
k_shade(SceneDev sc, Queue in, const uint32_t *__restrict__ in_count, const float4 *__restrict__ hits, Queue out,
        uint32_t *out_count, float4 *__restrict__ result, uint32_t depth, uint32_t max_depth, F3 bg, uint32_t nb, uint32_t flags) {
    __shared__ uint32_t s_cnt[kShadeThreads / 32][kBins + 1];
    __shared__ uint32_t s_base[kBins];
    const uint32_t n = *in_count;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t b0 = blockIdx.x * blockDim.x; b0 < n; b0 += stride) {  // block-uniform trip count
        const uint32_t i = b0 + threadIdx.x;
        const bool valid = i < n;
        bool emit = false;
        Scatter s;
        Rng r;
        F3 thr;
        uint32_t pid = 0, slot = kNoSlot;
        uint4 mats = make_uint4(0u, 0u, 0u, 0u);
        if (valid) {
            const float4 o = in.o[i], d = in.d[i], t = in.t[i], hh = hits[i];
            if (kRefMul) mats = in.m[i];
            Ray ray;
            ray.o = f3(o.x, o.y, o.z);
            ray.d = f3(d.x, d.y, d.z);
            ray.tmin = RT_ZERO;  // Shader.cu:231-233
            ray.tmax = RT_INF;
            pid = __float_as_uint(o.w);
            thr = f3(t.x, t.y, t.z);
            Hit h;
            h.t = hh.x;
            h.slot = __float_as_uint(hh.y);
            h.u = hh.z;
            h.v = hh.w;
            if (h.slot == kNoSlot) {
                if (kRefMul) {
                    const F3 rr = unwind_product(sc, bg, mats, depth - 1u);
                    result[pid] = make_float4(rr.x, rr.y, rr.z, 0.0f);
                } else {
                    result[pid] = make_float4(thr.x * bg.x, thr.y * bg.y, thr.z * bg.z, 0.0f);
                }
            } else if (depth >= max_depth) {
                result[pid] = make_float4(0, 0, 0, 0);
            } else {
                const uint4 rv = in.r[i];
                r.d = __float_as_uint(t.w);
                r.v0 = rv.x; r.v1 = rv.y; r.v2 = rv.z; r.v3 = rv.w;
                r.v4 = in.r4[i];
                s = scatter(sc, ray, h, r);
                thr = f3(thr.x * s.alb.x, thr.y * s.alb.y, thr.z * s.alb.z);  // Shader.cu:236-238
                slot = (flags & RT_FLAG_REF_NO_SELF_SKIP) ? kNoSlot : h.slot;
                if (kRefMul) {
                    if (depth == 2u) mats.y = s.mat; else if (depth == 3u) mats.z = s.mat; else mats.w = s.mat;
                }
                emit = true;
            }
        }
        const uint32_t qi = binned_push<kShadeThreads / 32>(out_count, emit, emit ? dir_bin(s.dir, nb) : 0u, nb, s_cnt, s_base);
        if (emit) {
            queue_write(out, qi, s.P, s.dir, thr, pid, slot, r);
            if (kRefMul) out.m[qi] = mats;
        }
    }
}

// ==================================================================== fused path kernel (option "fused" = 1; the default is the wavefront form)
// The reference's recursion (optixTrace inside closest-hit, Shader.cu:229-241) as ONE persistent kernel: a lane
// owns a path from its first scatter to its end — closest hit over the 8-wide tree, then the closest-hit / miss
// program, then the next segment — so no ray queues, no hit records and no per-bounce launches (and drains of
// the persistent grid) exist: the wavefront form above moves ~200 bytes per segment through HBM and spends 17 %
// of a C4 frame outside the traversal kernel.  What stays of the wavefront design is the warp-phased traversal
// (extend_warp_wide): lanes of a warp run wide-node tests, primitive tests and the shading step as three
// warp-uniform phases, and a lane that has finished a path takes the next work unit from a global counter.
//
// Work unit = (hit pixel i, chunk c of kChunk consecutive samples): the unit's samples are traced one after the
// other by the same lane (the primary hit is shared, Shader.cu:249-261 has no jitter), each with its own XORWOW
// stream(p, s); k_prepare stores the stream state at the start of every chunk.  Radiance goes to
// result[i * S + s] and is summed per pixel in sample order by k_accumulate_hits, so the frame is bit-identical
// to the wavefront form and to the oracle whatever the scheduling.
// Per-lane path state that is not needed during traversal lives in shared memory ([word][thread]).
enum PathWord { PW_OX, PW_OY, PW_OZ, PW_DX, PW_DY, PW_DZ, PW_TR, PW_TG, PW_TB, PW_RD, PW_R0, PW_R1, PW_R2, PW_R3, PW_R4,
                PW_PX, PW_PY, PW_PZ, PW_PIX, PW_IDX, PW_S, PW_SEND, PW_DEPTH, PW_M0, PW_M1, PW_M2, PW_M3, PW_COUNT };
constexpr int kPathWords = PW_M0;          // words per lane without the reference-order albedo list
constexpr int kPathWordsRef = PW_COUNT;    // with it (RT_FLAG_REF_MUL_ON_RETURN)

struct PathArgs {
    CamDev cam;
    uint32_t W, H, S, chunk, n_chunks, max_depth, flags;
    F3 bg;
    const float4 *phit;           // primary hit per pixel
    const uint32_t *hit_pix;      // compacted list of pixels whose primary ray hit something
    const uint32_t *n_hit;        // its length (device)
    const uint32_t *rng_start;    // [S][6][cap]: XORWOW state of stream(p, s) for hit pixel i (written by k_prepare)
    uint32_t cap;                 // stride of rng_start (= W*H)
    float4 *result;               // [i * S + s]
    uint32_t *unit_ctr;           // work-unit counter
    uint32_t *seg_ctr;            // segments traced by this launch
};

template <bool kRefMul>
__device__ __forceinline__ void path_warp(const SceneDev &sc, const PathArgs &a, int t1, int t2, int t3, int fetch_min, uint32_t magic,
                                          const uint8_t (*s_perm)[256], uint2 (*s_stack)[RT_EXTEND_THREADS],
                                          uint32_t (*s_w)[RT_EXTEND_THREADS]) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u, tx = threadIdx.x;
    const uint32_t lt = (1u << lane) - 1u;
    const float4 *wn = reinterpret_cast<const float4 *>(sc.wnodes);
    const BuildInfo *info = sc.info;
    const uint32_t n_huge = info->n_huge;
    const bool have_tree = info->n_bvh != 0;
    const uint32_t n_hit = *a.n_hit;
    const uint32_t n_units = n_hit * a.n_chunks;  // host guarantees npix * n_chunks < 2^32

    WideStack stack;
    stack.sh = s_stack;
    WideTrav t;
    trav_clear(t);
    bool alive = false;      // the lane holds a ray whose closest hit is being searched
    bool have_unit = false;  // the lane owns a work unit
    bool exhausted = false;  // the unit counter has run out (warp-uniform)
    uint32_t nseg = 0;

#define PWF(k) __uint_as_float(s_w[k][tx])
    for (;;) {
        // ---- phase 0: lanes whose ray is finished (or that own nothing) run the closest-hit / miss program,
        //      start the next sample or fetch the next unit.  Entered when at least fetch_min lanes wait, or
        //      when nobody has traversal work left.
        const bool fin = alive && trav_finished(t);
        const bool idle = !alive && !have_unit && !exhausted;
        const uint32_t m0 = __ballot_sync(FULL, fin || idle);
        if (m0 && (__popc(m0) >= fetch_min || !__any_sync(FULL, alive && !fin))) {
            bool scatter_now = false, start_sample = false;
            Ray ray;
            Hit hit;
            F3 thr = f3(1.0f, 1.0f, 1.0f);
            Rng r;
            uint32_t depth = 0;
            if (fin) {
                alive = false;
                depth = s_w[PW_DEPTH][tx];
                bool path_done = false;
                F3 rgb = f3(0.0f, 0.0f, 0.0f);
                if (t.best.slot == kNoSlot) {  // __miss__missProgram, Shader.cu:276-287
                    if (kRefMul) {
                        rgb = unwind_product(sc, a.bg, make_uint4(s_w[PW_M0][tx], s_w[PW_M1][tx], s_w[PW_M2][tx], s_w[PW_M3][tx]), depth - 1u);
                    } else {
                        rgb = f3(PWF(PW_TR) * a.bg.x, PWF(PW_TG) * a.bg.y, PWF(PW_TB) * a.bg.z);
                    }
                    path_done = true;
                } else if (depth >= a.max_depth) {  // Shader.cu:102-107
                    path_done = true;
                } else {
                    ray.o = f3(PWF(PW_OX), PWF(PW_OY), PWF(PW_OZ));
                    ray.d = f3(PWF(PW_DX), PWF(PW_DY), PWF(PW_DZ));
                    ray.tmin = RT_ZERO;
                    ray.tmax = RT_INF;
                    hit = t.best;
                    thr = f3(PWF(PW_TR), PWF(PW_TG), PWF(PW_TB));
                    scatter_now = true;
                }
                if (path_done) {
                    const uint32_t sidx = s_w[PW_S][tx];
                    a.result[(size_t)s_w[PW_IDX][tx] * a.S + sidx] = make_float4(rgb.x, rgb.y, rgb.z, 0.0f);
                    s_w[PW_S][tx] = sidx + 1u;
                    if (sidx + 1u < s_w[PW_SEND][tx]) start_sample = true; else have_unit = false;
                }
            }
            // ---- unit fetch (warp-aggregated)
            {
                const uint32_t want = __ballot_sync(FULL, !alive && !have_unit && !scatter_now && !start_sample);
                if (want && !exhausted) {
                    uint32_t base = 0;
                    if (lane == 0) base = atomicAdd(a.unit_ctr, (uint32_t)__popc(want));
                    base = __shfl_sync(FULL, base, 0);
                    if (base + __popc(want) >= n_units) exhausted = true;
                    const uint32_t u = base + __popc(want & lt);
                    if (((want >> lane) & 1u) && u < n_units) {
                        const uint32_t c = u / n_hit, i = u - c * n_hit;
                        const uint32_t p = a.hit_pix[i];
                        const Ray pr = primary_ray(a.cam, p % a.W, p / a.W, a.W, a.H);  // shared by the unit's samples
                        s_w[PW_PX][tx] = __float_as_uint(pr.d.x); s_w[PW_PY][tx] = __float_as_uint(pr.d.y); s_w[PW_PZ][tx] = __float_as_uint(pr.d.z);
                        s_w[PW_PIX][tx] = p;
                        s_w[PW_IDX][tx] = i;
                        s_w[PW_S][tx] = c * a.chunk;
                        s_w[PW_SEND][tx] = min(a.S, (c + 1u) * a.chunk);
                        have_unit = true;
                        start_sample = true;
                    }
                }
            }
            if (start_sample) {
                // stream(p, s): state written by k_prepare (one 32-draw stride after stream(p, s-1))
                const uint32_t *rs = a.rng_start + (size_t)s_w[PW_S][tx] * 6u * a.cap + s_w[PW_IDX][tx];
                r.d = rs[0];
                r.v0 = rs[1 * (size_t)a.cap]; r.v1 = rs[2 * (size_t)a.cap]; r.v2 = rs[3 * (size_t)a.cap];
                r.v3 = rs[4 * (size_t)a.cap]; r.v4 = rs[5 * (size_t)a.cap];
                ray.o = a.cam.c;
                ray.d = f3(PWF(PW_PX), PWF(PW_PY), PWF(PW_PZ));
                ray.tmin = RT_ZERO;
                ray.tmax = RT_INF;
                const float4 h = a.phit[s_w[PW_PIX][tx]];
                hit.t = h.x;
                hit.slot = __float_as_uint(h.y);
                hit.u = h.z;
                hit.v = h.w;
                depth = 1u;
                thr = f3(1.0f, 1.0f, 1.0f);
                scatter_now = true;
            } else if (scatter_now) {
                r.d = s_w[PW_RD][tx];
                r.v0 = s_w[PW_R0][tx]; r.v1 = s_w[PW_R1][tx]; r.v2 = s_w[PW_R2][tx]; r.v3 = s_w[PW_R3][tx]; r.v4 = s_w[PW_R4][tx];
            }
            if (scatter_now) {  // closesthitImpl, Shader.cu:94-242
                const Scatter s1 = scatter(sc, ray, hit, r);
                thr = f3(thr.x * s1.alb.x, thr.y * s1.alb.y, thr.z * s1.alb.z);  // :236-238 (on the way down)
                s_w[PW_OX][tx] = __float_as_uint(s1.P.x); s_w[PW_OY][tx] = __float_as_uint(s1.P.y); s_w[PW_OZ][tx] = __float_as_uint(s1.P.z);
                s_w[PW_DX][tx] = __float_as_uint(s1.dir.x); s_w[PW_DY][tx] = __float_as_uint(s1.dir.y); s_w[PW_DZ][tx] = __float_as_uint(s1.dir.z);
                s_w[PW_TR][tx] = __float_as_uint(thr.x); s_w[PW_TG][tx] = __float_as_uint(thr.y); s_w[PW_TB][tx] = __float_as_uint(thr.z);
                s_w[PW_RD][tx] = r.d;
                s_w[PW_R0][tx] = r.v0; s_w[PW_R1][tx] = r.v1; s_w[PW_R2][tx] = r.v2; s_w[PW_R3][tx] = r.v3; s_w[PW_R4][tx] = r.v4;
                if (kRefMul) s_w[PW_M0 + (depth - 1u)][tx] = s1.mat;  // depth <= 4 here (max_depth <= 5)
                s_w[PW_DEPTH][tx] = depth + 1u;
                // Shader.cu:230-233: tmin 1e-6, tmax 1e16; by default the ray skips the primitive it leaves
                trav_start(t, s1.P, s1.dir, RT_ZERO, RT_INF, (a.flags & RT_FLAG_REF_NO_SELF_SKIP) ? kNoSlot : hit.slot, n_huge, have_tree);
                alive = true;
                nseg += 1u;
            }
        }
        if (!__any_sync(FULL, alive)) {
            if (exhausted) break;
            continue;
        }

        // ---- phase 1: wide nodes
        bool first = true;
        for (;;) {
            const bool want = alive && trav_wants_node(t);
            const uint32_t mw = __ballot_sync(FULL, want);
            if (!mw) break;
            if (!first && __popc(mw) < t1 && __any_sync(FULL, alive && trav_wants_prim(t))) break;
            first = false;
            if (want) trav_node(t, stack, wn, magic, s_perm);
        }
        // ---- phase 2: postponed primitives, one per iteration
        first = true;
        for (;;) {
            const bool hl = alive && trav_wants_prim(t);
            const uint32_t ml = __ballot_sync(FULL, hl);
            if (!ml) break;
            if (!first && __popc(ml) < t2 && __any_sync(FULL, alive && trav_wants_node(t))) break;
            first = false;
            if (hl) trav_prim(t, sc, f3(PWF(PW_OX), PWF(PW_OY), PWF(PW_OZ)), f3(PWF(PW_DX), PWF(PW_DY), PWF(PW_DZ)));
        }
        // ---- outlier primitives
        trav_outliers(t, alive, info, n_huge, t3, f3(PWF(PW_OX), PWF(PW_OY), PWF(PW_OZ)), f3(PWF(PW_DX), PWF(PW_DY), PWF(PW_DZ)));
    }
#undef PWF
    // segments traced by this warp
    for (int o = 16; o > 0; o >>= 1) nseg += __shfl_xor_sync(FULL, nseg, o);
    if (lane == 0 && nseg) atomicAdd(a.seg_ctr, nseg);
}

template <int kMinBlocks, bool kRefMul>
__global__ void __launch_bounds__(RT_EXTEND_THREADS, kMinBlocks)
k_path(SceneDev sc, PathArgs a, int t1, int t2, int t3, int fetch_min, uint32_t magic) {
    __shared__ uint8_t s_perm[8][256];
    __shared__ uint2 s_stack[RT_SH_STACK][RT_EXTEND_THREADS];
    __shared__ uint32_t s_w[kRefMul ? kPathWordsRef : kPathWords][RT_EXTEND_THREADS];
    fill_perm_table(s_perm);
    path_warp<kRefMul>(sc, a, t1, t2, t3, fetch_min, magic, s_perm, s_stack, s_w);
}

// Per pixel, before k_path.  Pixels whose primary ray hit something (and may scatter: max_depth > 1) are appended to
// hit_pix and get the XORWOW state of each of their S streams, stream(p, s) = curand_init(seed, p, (offset + s) * 32),
// by stepping the pixel's state (full warps: the hit pixels of an 8x4 tile mostly come together).  All other
// pixels receive their S samples here (miss: 1 * background, Shader.cu:276-287; depth cut-off: 0, :102-107), added one by
// one in sample order, and do NOT step their state: rng_off[p] records the sample offset a pixel's state stands
// at, and a pixel that starts to hit after a rebuild catches up first (or is re-initialised when it lags far).
// The list order is block-ordered (tile order inside a block): neighbouring lanes of k_path work on neighbouring pixels.
constexpr int kPrepThreads = 256;
__global__ void __launch_bounds__(kPrepThreads)
k_prepare(uint32_t W, uint32_t H, uint32_t S, uint32_t max_depth, F3 bg, const float4 *__restrict__ phit, uint32_t *__restrict__ rng,
          uint32_t *__restrict__ rng_off, unsigned long long seed, uint32_t sample_offset, uint32_t *__restrict__ hit_pix,
          uint32_t *__restrict__ n_hit, uint32_t *__restrict__ rng_start, uint32_t cap, float4 *__restrict__ accum, int fresh) {
    __shared__ uint32_t s_warp[kPrepThreads / 32];
    __shared__ uint32_t s_base;
    uint32_t x, y;
    const bool valid = tile_pixel(blockIdx.x * blockDim.x + threadIdx.x, W, H, x, y);
    const uint32_t npix = W * H;
    const uint32_t p = valid ? y * W + x : 0u;
    bool is_hit = false;
    if (valid) is_hit = __float_as_uint(phit[p].y) != kNoSlot;
    const bool go = is_hit && max_depth > 1u;
    // block-ordered append
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t bal = __ballot_sync(0xFFFFFFFFu, go);
    if (lane == 0) s_warp[w] = __popc(bal);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int k = 0; k < kPrepThreads / 32; k++) {
            const uint32_t c = s_warp[k];
            s_warp[k] = tot;
            tot += c;
        }
        s_base = tot ? atomicAdd(n_hit, tot) : 0u;
    }
    __syncthreads();
    const uint32_t i = s_base + s_warp[w] + __popc(bal & ((1u << lane) - 1u));
    if (!valid) return;
    if (!go) {
        if (accum) {
            float4 acc = fresh ? make_float4(0, 0, 0, 0) : accum[p];
            const float4 add = is_hit ? make_float4(0, 0, 0, 0) : make_float4(1.0f * bg.x, 1.0f * bg.y, 1.0f * bg.z, 0.0f);
            for (uint32_t s = 0; s < S; s++) {
                acc.x += add.x; acc.y += add.y; acc.z += add.z;
            }
            acc.w += (float)S;
            accum[p] = acc;
        }
        return;
    }
    hit_pix[i] = p;
    Rng st;
    const uint32_t have = rng_off[p];
    if (have > sample_offset || sample_offset - have > kRngMaxLag) {
        curandStateXORWOW cs;
        curand_init(seed, (unsigned long long)p, (unsigned long long)sample_offset * RT_RNG_SAMPLE_STRIDE, &cs);
        st.d = cs.d; st.v0 = cs.v[0]; st.v1 = cs.v[1]; st.v2 = cs.v[2]; st.v3 = cs.v[3]; st.v4 = cs.v[4];
    } else {
        st.d = rng[0 * (size_t)npix + p];
        st.v0 = rng[1 * (size_t)npix + p];
        st.v1 = rng[2 * (size_t)npix + p];
        st.v2 = rng[3 * (size_t)npix + p];
        st.v3 = rng[4 * (size_t)npix + p];
        st.v4 = rng[5 * (size_t)npix + p];
        for (uint32_t k = (sample_offset - have) * RT_RNG_SAMPLE_STRIDE; k > 0; k--) (void)rng_next(st);
    }
    for (uint32_t s = 0; s < S; s++) {
        uint32_t *rs = rng_start + (size_t)s * 6u * cap + i;  // [sample][word][hit pixel], stride cap >= number of hit pixels
        rs[0] = st.d;
        rs[1 * (size_t)cap] = st.v0;
        rs[2 * (size_t)cap] = st.v1;
        rs[3 * (size_t)cap] = st.v2;
        rs[4 * (size_t)cap] = st.v3;
        rs[5 * (size_t)cap] = st.v4;
#pragma unroll 8
        for (uint32_t k = 0; k < RT_RNG_SAMPLE_STRIDE; k++) (void)rng_next(st);
    }
    rng[0 * (size_t)npix + p] = st.d;
    rng[1 * (size_t)npix + p] = st.v0;
    rng[2 * (size_t)npix + p] = st.v1;
    rng[3 * (size_t)npix + p] = st.v2;
    rng[4 * (size_t)npix + p] = st.v3;
    rng[5 * (size_t)npix + p] = st.v4;
    rng_off[p] = sample_offset + S;
}

// ordered sum of the S path results of every hit pixel (after k_path)
__global__ void __launch_bounds__(256)
k_accumulate_hits(const float4 *__restrict__ result, const uint32_t *__restrict__ hit_pix, const uint32_t *__restrict__ n_hit, uint32_t S,
                  float4 *__restrict__ accum, int fresh) {
    const uint32_t n = *n_hit;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t p = hit_pix[i];
        float4 a = fresh ? make_float4(0, 0, 0, 0) : accum[p];
        for (uint32_t s = 0; s < S; s++) {
            const float4 r = result[(size_t)i * S + s];
            a.x += r.x;
            a.y += r.y;
            a.z += r.z;
        }
        a.w += (float)S;
        accum[p] = a;
    }
}

// ------------------------------------------------------------------ resolve: mean, clamp, sRGB, quantise
// DeviceFunctions.cuh:153-212
__device__ __forceinline__ float srgb(float c) {
    const float cx = fmaxf(0.0f, fminf(c, 1.0f));
    const float px = powf(cx, 1.0f / 2.4f);
    const float sx = cx < 0.0031308f ? 12.92f * cx : 1.055f * px - 0.055f;
    return fmaxf(0.0f, fminf(sx, 1.0f));
}
__global__ void __launch_bounds__(256)
k_resolve(const float4 *__restrict__ accum, size_t n, float4 *__restrict__ color, uchar4 *__restrict__ rgba8) {
    const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const float4 a = accum[p];
    const float inv = a.w > 0.0f ? 1.0f / a.w : 0.0f;
    const float sx = srgb(a.x * inv), sy = srgb(a.y * inv), sz = srgb(a.z * inv);
    if (color) color[p] = make_float4(sx, sy, sz, 1.0f);
    if (rgba8)
        rgba8[p] = make_uchar4((unsigned char)min((unsigned int)(sx * 256.0f), 255u),
                               (unsigned char)min((unsigned int)(sy * 256.0f), 255u),
                               (unsigned char)min((unsigned int)(sz * 256.0f), 255u), 255);
}

// ------------------------------------------------------------------ validation: caller-supplied rays
__global__ void __launch_bounds__(128)
k_pack_rays(const float *__restrict__ rays, const uint32_t *__restrict__ skip_gid, const uint32_t *__restrict__ gid_to_slot,
            uint32_t n, float4 *__restrict__ qo, float4 *__restrict__ qd, uint32_t *__restrict__ qskip) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *r = rays + 8 * (size_t)i;
    qo[i] = make_float4(r[0], r[1], r[2], r[3]);
    qd[i] = make_float4(r[4], r[5], r[6], r[7]);
    uint32_t skip = kNoSlot;
    if (skip_gid && skip_gid[i] != RT_MISS_ID) skip = gid_to_slot[skip_gid[i]];
    qskip[i] = skip;
}
__global__ void __launch_bounds__(128)
k_unpack_hits(SceneDev sc, const float4 *__restrict__ hits, uint32_t n, float *__restrict__ out_t, uint32_t *__restrict__ out_id) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 h = hits[i];
    const uint32_t slot = __float_as_uint(h.y);
    out_t[i] = (slot == kNoSlot) ? __int_as_float(0x7F800000) : h.x;
    out_id[i] = (slot == kNoSlot) ? RT_MISS_ID : sc.slot_prim[slot];
}
__global__ void k_invert_slots(const uint32_t *__restrict__ slot_prim, uint32_t n, uint32_t *__restrict__ gid_to_slot) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) gid_to_slot[slot_prim[i]] = i;
}

// primary hit records in their rank-independent form: slot (position in THIS build's re-packed primitive
// arrays; the collapse allocates with atomics, so it differs from build to build) <-> global primitive id
__global__ void k_export_hits(const uint32_t *__restrict__ slot_prim, float4 *__restrict__ hits, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 h = hits[i];
    const uint32_t slot = __float_as_uint(h.y);
    h.y = __uint_as_float(slot == kNoSlot ? RT_MISS_ID : __ldg(slot_prim + slot));
    hits[i] = h;
}
__global__ void k_import_hits(const float4 *__restrict__ ext, const uint32_t *__restrict__ gid_to_slot, uint32_t n,
                              float4 *__restrict__ phit) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 h = ext[i];
    const uint32_t gid = __float_as_uint(h.y);
    h.y = __uint_as_float(gid == RT_MISS_ID ? kNoSlot : __ldg(gid_to_slot + gid));
    phit[i] = h;
}

// pixels whose primary ray hit something: what the wavefront queues of this frame have to hold per sample in flight
__global__ void __launch_bounds__(1024)
k_count_hits(const float4 *__restrict__ phit, uint32_t npix, uint32_t *__restrict__ out) {
    __shared__ uint32_t s_w[32];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool hit = i < npix && __float_as_uint(phit[i].y) != kNoSlot;
    const uint32_t c = __popc(__ballot_sync(0xFFFFFFFFu, hit));
    if ((threadIdx.x & 31u) == 0) s_w[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t v = s_w[threadIdx.x];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
        if (threadIdx.x == 0 && v) atomicAdd(out, v);
    }
}

// cumulative segment counter (device side, so that no host read is needed between frames)
__global__ void k_tally(const uint32_t *__restrict__ counters, uint32_t n_cnt, uint32_t primary, unsigned long long *total) {
    if (threadIdx.x || blockIdx.x) return;
    unsigned long long s = primary;
    for (uint32_t i = 0; i < n_cnt; i++) s += counters[i];
    *total += s;
}

// ------------------------------------------------------------------ host drivers
static inline uint32_t div_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }

// optional per-launch timing of the traversal kernels (rt_scene_set_option "profile")
static cudaError_t prof_mark(Scene *s, cudaStream_t st) {
    if (!s->profile) return cudaSuccess;
    if (s->prof_used == s->prof_ev.size()) {
        cudaEvent_t e;
        cudaError_t err = cudaEventCreate(&e);
        if (err != cudaSuccess) return err;
        s->prof_ev.push_back(e);
    }
    return cudaEventRecord(s->prof_ev[s->prof_used++], st);
}

static Queue make_queue(Scene *s, int k, bool ref_mul = false) {
    return Queue{s->q_o[k].p, s->q_d[k].p, s->q_t[k].p, s->q_r[k].p, s->q_r4[k].p, ref_mul ? s->q_m[k].p : nullptr};
}

// one closest-hit pass over a ray buffer; timed when option "profile" is on
static int launch_extend(Scene *s, const SceneDev &sd, const ExtendIn &in, const uint32_t *n_ptr, uint32_t n_fixed,
                         uint32_t *fetch_ctr, float4 *hits, cudaStream_t st) {
    Context *ctx = s->ctx;
    uint32_t grid = (uint32_t)ctx->sm_count * s->extend_blocks_per_sm;
    if (!n_ptr) grid = std::min(grid, std::max(1u, div_up(n_fixed, 128u)));
    RT_CUDA(ctx, prof_mark(s, st));
    if (s->wide) {
        if (s->extend_blocks_per_sm >= 8)
            k_extend_wide<8><<<grid, 128, 0, st>>>(sd, in, n_ptr, n_fixed, fetch_ctr, hits, (int)s->t1, (int)s->t2, (int)s->t3, (int)s->fetch_min, RT_DEQ_MAGIC);
        else if (s->extend_blocks_per_sm == 7)
            k_extend_wide<7><<<grid, 128, 0, st>>>(sd, in, n_ptr, n_fixed, fetch_ctr, hits, (int)s->t1, (int)s->t2, (int)s->t3, (int)s->fetch_min, RT_DEQ_MAGIC);
        else if (s->extend_blocks_per_sm >= 6)
            k_extend_wide<6><<<grid, 128, 0, st>>>(sd, in, n_ptr, n_fixed, fetch_ctr, hits, (int)s->t1, (int)s->t2, (int)s->t3, (int)s->fetch_min, RT_DEQ_MAGIC);
        else
            k_extend_wide<5><<<grid, 128, 0, st>>>(sd, in, n_ptr, n_fixed, fetch_ctr, hits, (int)s->t1, (int)s->t2, (int)s->t3, (int)s->fetch_min, RT_DEQ_MAGIC);
    }
    else
        k_extend<<<grid, 128, 0, st>>>(sd, in, n_ptr, n_fixed, fetch_ctr, hits, (int)s->t1, (int)s->t2, (int)s->fetch_min);
    RT_CUDA(ctx, prof_mark(s, st));
    s->launches += 1;
    return RT_OK;
}

int render_frame(Scene *s, const rt_camera *cam, const rt_frame_params *fp, const rt_outputs *out, cudaStream_t st) {
    Context *ctx = s->ctx;
    if (!s->built) return set_error(ctx, RT_ERR_STATE, "rt_render before rt_accel_build");
    const uint32_t W = fp->width, H = fp->height;
    if (!W || !H || (uint64_t)W * H >= (1ull << 31)) return set_error(ctx, RT_ERR_INVALID, "bad image size %ux%u", W, H);
    if (fp->max_depth > 255) return set_error(ctx, RT_ERR_INVALID, "max_depth %u too large", fp->max_depth);
    if (fp->max_depth > 1 && 3ull * (fp->max_depth - 1) > RT_RNG_SAMPLE_STRIDE)
        return set_error(ctx, RT_ERR_INVALID, "max_depth %u exceeds the RNG stride budget", fp->max_depth);
    const bool ref_mul = (fp->flags & RT_FLAG_REF_MUL_ON_RETURN) != 0;
    if (ref_mul && fp->max_depth > 5)
        return set_error(ctx, RT_ERR_INVALID, "RT_FLAG_REF_MUL_ON_RETURN keeps 4 albedos per path: max_depth <= 5 (the reference's rayTraceDepth)");
    const uint32_t npix = W * H;
    const uint32_t spp = out->accum_linear ? fp->spp : 0u;
    // samples per pixel in flight per wavefront batch: larger batches amortise the drain of the persistent
    // traversal grid (C4: 4 -> 32 spp per batch = +5 %); 168 bytes of queue per (pixel, sample).  Auto: up to
    // 32, within a third of the free device memory (and 48 GB).  Results do not depend on it.
    // Fused form (option "fused" = 1, not the default; needs the 8-wide tree): no queues.  Per batch of Sb samples: 16 bytes of
    // result and 24 bytes of stream state per (pixel, sample), sized for the worst case (every pixel hits).
    const bool fused = s->fused && s->wide;
    const uint32_t chunk = std::max(1u, s->chunk_spp);
    uint32_t Sb = s->batch_spp;
    if (fused && !Sb) {
        if (!s->auto_batch || s->auto_batch_npix != npix) {
            size_t free_b = 0, total_b = 0;
            RT_CUDA(ctx, cudaMemGetInfo(&free_b, &total_b));
            const size_t budget = std::min<size_t>((size_t)12 << 30, (free_b + s->bytes) / 4);
            uint32_t b = 64;
            while (b > 1 && ((size_t)npix * b * 40 > budget || (uint64_t)npix * b >= (1ull << 32))) b >>= 1;
            s->auto_batch = b;
            s->auto_batch_npix = npix;
        }
        Sb = s->auto_batch;
    }
    if (!Sb) {
        if (!s->auto_batch || s->auto_batch_npix != npix) {
            size_t free_b = 0, total_b = 0;
            RT_CUDA(ctx, cudaMemGetInfo(&free_b, &total_b));
            const size_t budget = std::min<size_t>((size_t)48 << 30, (free_b + s->bytes) / 3);
            uint32_t b = 32;
            while (b > 1 && (size_t)npix * b * 168 > budget) b >>= 1;
            s->auto_batch = b;
            s->auto_batch_npix = npix;
        }
        Sb = s->auto_batch;
    }
    if ((uint64_t)npix * Sb >= (1ull << 32)) return set_error(ctx, RT_ERR_INVALID, "batch too large");
    const uint32_t n_batches = spp ? div_up(spp, Sb) : 0u;
    const uint32_t D = fp->max_depth;

    RT_CUDA(ctx, s->phit.reserve(npix, &s->bytes));
    // Wavefront queues hold one entry per (continuing pixel, sample in flight).  With option "compact_queues" (default) they
    // are sized from THIS frame's primary-hit count, read back after the primary-visibility pass (one 4-byte copy and a
    // stream synchronisation per rt_render: the host waits for the build + primary pass it has just enqueued, ~4 ms on C4,
    // and then enqueues the bounces; nothing is idle but the few microseconds of launch latency).  Without it they are
    // sized for every pixel (C4: 44.6 GB instead of 9.7 GB) and rt_render never blocks.
    const bool compact = spp && !fused && s->compact_queues;
    const size_t qn_full = std::max((size_t)npix, spp ? (size_t)npix * std::min(Sb, spp) : (size_t)0);
    const size_t qn = (fused || compact) ? (size_t)npix : qn_full;
    const size_t rn = fused ? (size_t)npix : qn_full;  // result: one entry per (pixel, sample in flight)
    for (int k = 0; k < 2; k++) {
        const size_t qk = (k == 0 || (spp && !fused && !compact)) ? qn : 1;  // primary rays use queue 0's origin/direction arrays
        RT_CUDA(ctx, s->q_o[k].reserve(qk, &s->bytes));
        RT_CUDA(ctx, s->q_d[k].reserve(qk, &s->bytes));
    }
    if (spp && fused) {
        RT_CUDA(ctx, s->rng_cur.reserve(6 * (size_t)npix, &s->bytes));
        RT_CUDA(ctx, s->hit_pix.reserve(npix, &s->bytes));
        RT_CUDA(ctx, s->rng_pix_off.reserve(npix, &s->bytes));
        RT_CUDA(ctx, s->rng_start.reserve((size_t)std::min(Sb, spp) * 6 * npix, &s->bytes));
        RT_CUDA(ctx, s->result.reserve((size_t)npix * std::min(Sb, spp), &s->bytes));
    } else if (spp) {
        RT_CUDA(ctx, s->rng_cur.reserve(6 * (size_t)npix, &s->bytes));
        RT_CUDA(ctx, s->rng_pix_off.reserve(npix, &s->bytes));
        RT_CUDA(ctx, s->hit_pix.reserve(npix, &s->bytes));
        if (!compact) {
            for (int k = 0; k < 2; k++) {
                RT_CUDA(ctx, s->q_t[k].reserve(qn, &s->bytes));
                RT_CUDA(ctx, s->q_r[k].reserve(qn, &s->bytes));
                RT_CUDA(ctx, s->q_r4[k].reserve(qn, &s->bytes));
                if (ref_mul) RT_CUDA(ctx, s->q_m[k].reserve(qn, &s->bytes));
            }
            RT_CUDA(ctx, s->hit.reserve(qn, &s->bytes));
            RT_CUDA(ctx, s->result.reserve(rn, &s->bytes));
            RT_CUDA(ctx, s->rng_start.reserve((size_t)std::min(Sb, spp) * 6 * npix, &s->bytes));
        } else {
            RT_CUDA(ctx, s->prim_fetch.reserve(1, &s->bytes));
        }
    }
    // counters: [n_batches][D+3] queue sizes, then one fetch counter per k_extend launch
    // fused: [n_batches] segment counters (summed by k_tally / rt_get_stats like the queue sizes), then [n_batches]
    // hit-pixel counts, [n_batches] unit counters and the primary pass's fetch counter
    const size_t n_cnt = fused ? (size_t)(n_batches ? n_batches : 1) : (size_t)(n_batches ? n_batches : 1) * (D + 3);
    // (wavefront: then one hit-pixel count per batch)
    const size_t n_fetch = fused ? 1 + 2 * (size_t)n_batches : 1 + (size_t)n_batches * (D + 1) + n_batches;
    RT_CUDA(ctx, s->counters.reserve(n_cnt + n_fetch, &s->bytes));
    if (!s->seg_total.p) {
        RT_CUDA(ctx, s->seg_total.reserve(1, &s->bytes));
        RT_CUDA(ctx, cudaMemsetAsync(s->seg_total.p, 0, sizeof(unsigned long long), st));
    }
    if (!s->ev[0])
        for (int k = 0; k < 4; k++) RT_CUDA(ctx, cudaEventCreate(&s->ev[k]));

    RT_CUDA(ctx, cudaEventRecord(s->ev[2], st));
    RT_CUDA(ctx, cudaMemsetAsync(s->counters.p, 0, (n_cnt + n_fetch) * sizeof(uint32_t), st));
    uint32_t *fetch = s->counters.p + n_cnt + (fused ? 2 * (size_t)n_batches : 0);

    const SceneDev sd = scene_dev(s);
    CamDev cd;
    cd.c = F3{cam->center[0], cam->center[1], cam->center[2]};
    cd.U = F3{cam->U[0], cam->U[1], cam->U[2]};
    cd.V = F3{cam->V[0], cam->V[1], cam->V[2]};
    cd.W = F3{cam->W[0], cam->W[1], cam->W[2]};
    const F3 bg = F3{fp->background[0], fp->background[1], fp->background[2]};

    const uint32_t tiles = div_up(W, 8) * div_up(H, 4);
    const uint32_t pix_blocks = div_up(tiles * 32u, 128u);
    // RNG: stream(p, s) = curand_init(seed, p, (sample_offset + s) * 32).  When the per-pixel states have to be
    // (re)positioned (first frame, or a rank of the spp split whose sample blocks are not contiguous from
    // frame to frame) that runs on a side stream next to the primary-visibility pass, which does not need them.
    bool rng_forked = false;
    if (spp) {
        // Both forms keep one state per pixel (rng_cur) plus the sample it stands at (rng_pix_off) and advance it lazily.
        // A new seed / size, a request behind the states or far ahead of them re-creates all of them at this launch's
        // first sample (k_rng_init, ~1 ms at 4K) rather than leaving it to the per-pixel curand_init of the kernels.
        const bool same = s->rng_valid && s->rng_seed == fp->seed && s->rng_w == W && s->rng_h == H;
        const bool reinit = !same || (uint64_t)fp->sample_offset < s->rng_cur_off || (uint64_t)fp->sample_offset - s->rng_cur_off > std::min(s->rng_max_lag, kRngMaxLag);
        if (reinit) {
            if (!s->aux_stream) {
                RT_CUDA(ctx, cudaStreamCreateWithFlags(&s->aux_stream, cudaStreamNonBlocking));
                RT_CUDA(ctx, cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
                RT_CUDA(ctx, cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
            }
            RT_CUDA(ctx, cudaEventRecord(s->ev_fork, st));
            RT_CUDA(ctx, cudaStreamWaitEvent(s->aux_stream, s->ev_fork, 0));
            k_rng_init<<<div_up(div_up(npix, kRngRun), 128), 128, 0, s->aux_stream>>>(s->rng_cur.p, s->rng_pix_off.p, fp->sample_offset, npix, fp->seed,
                                                                                    (unsigned long long)fp->sample_offset * RT_RNG_SAMPLE_STRIDE);
            s->launches += 1;
            s->rng_seed = fp->seed; s->rng_w = W; s->rng_h = H;
            s->rng_valid = true;
            RT_CUDA(ctx, cudaEventRecord(s->ev_join, s->aux_stream));
            rng_forked = true;
        }
    }
    // primary visibility: raygen -> closest hit -> ids / AOVs; skipped when the caller supplied the hit
    // records (rt_scene_set_primary_hits: the multi-GPU split traces each pixel's primary ray on ONE rank)
    const float4 *phit = s->phit.p;
    if (s->ext_phit) {
        // vals[1] is free after the build: global id -> slot of this build
        if (s->n_prims) k_invert_slots<<<div_up(s->n_prims, 256), 256, 0, st>>>(sd.slot_prim, s->n_prims, s->vals[1].p);
        k_import_hits<<<div_up(npix, 256), 256, 0, st>>>(s->ext_phit, s->vals[1].p, npix, s->phit.p);
        s->launches += 2;
    } else {
        k_raygen<<<pix_blocks, 128, 0, st>>>(cd, W, H, s->q_o[0].p, s->q_d[0].p);
        s->launches += 1;
        const ExtendIn in{s->q_o[0].p, s->q_d[0].p, nullptr};
        const int rc = launch_extend(s, sd, in, nullptr, npix, fetch++, s->phit.p, st);
        if (rc) return rc;
    }
    if (out->prim_id || out->albedo || out->normal) {
        k_primary_aov<<<pix_blocks, 128, 0, st>>>(sd, cd, W, H, phit, out->prim_id, reinterpret_cast<float4 *>(out->albedo),
                                                  reinterpret_cast<float4 *>(out->normal));
        s->launches += 1;
    }

    uint32_t hit_cap = npix;  // stride of rng_start = capacity in hit pixels
    if (compact) {
        uint32_t n_hit = 0;
        if (D > 1) {
            RT_CUDA(ctx, cudaMemsetAsync(s->prim_fetch.p, 0, sizeof(uint32_t), st));
            k_count_hits<<<div_up(npix, 1024), 1024, 0, st>>>(phit, npix, s->prim_fetch.p);
            s->launches += 1;
            // read back through pinned memory and wait by polling an event: cudaStreamSynchronize may put the thread to
            // sleep, and with one process per GPU on a shared host its wake-up cost 8.6 ms per frame on C5 at N = 8 (r7h)
            if (!s->h_hit) {
                RT_CUDA(ctx, cudaHostAlloc(reinterpret_cast<void **>(&s->h_hit), sizeof(uint32_t), cudaHostAllocDefault));
                RT_CUDA(ctx, cudaEventCreateWithFlags(&s->ev_hit, cudaEventDisableTiming));
            }
            RT_CUDA(ctx, cudaMemcpyAsync(s->h_hit, s->prim_fetch.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
            RT_CUDA(ctx, cudaEventRecord(s->ev_hit, st));
            for (;;) {
                const cudaError_t q = cudaEventQuery(s->ev_hit);
                if (q == cudaSuccess) break;
                if (q != cudaErrorNotReady) RT_CUDA(ctx, q);
            }
            n_hit = *s->h_hit;
        }
        const size_t need = (size_t)std::max(n_hit, 1u) * std::min(Sb, spp);
        const size_t qc = need + need / 32;  // a little head room; a scene whose hit count drifts further regrows by half (DevBuf)
        for (int k = 0; k < 2; k++) {
            RT_CUDA(ctx, s->q_o[k].reserve(k == 0 ? std::max(qc, (size_t)npix) : qc, &s->bytes));
            RT_CUDA(ctx, s->q_d[k].reserve(k == 0 ? std::max(qc, (size_t)npix) : qc, &s->bytes));
            RT_CUDA(ctx, s->q_t[k].reserve(qc, &s->bytes));
            RT_CUDA(ctx, s->q_r[k].reserve(qc, &s->bytes));
            RT_CUDA(ctx, s->q_r4[k].reserve(qc, &s->bytes));
            if (ref_mul) RT_CUDA(ctx, s->q_m[k].reserve(qc, &s->bytes));
        }
        RT_CUDA(ctx, s->hit.reserve(qc, &s->bytes));
        RT_CUDA(ctx, s->result.reserve(qc, &s->bytes));
        RT_CUDA(ctx, s->rng_start.reserve(6 * qc, &s->bytes));
        hit_cap = (uint32_t)std::min<size_t>(qc / std::min(Sb, spp), npix);  // hit pixels the buffers can hold
    }
    if (spp && fused) {
        if (rng_forked) RT_CUDA(ctx, cudaStreamWaitEvent(st, s->ev_join, 0));
        const uint32_t grid = (uint32_t)ctx->sm_count * s->extend_blocks_per_sm;
        uint32_t done = 0;
        for (uint32_t b = 0; b < n_batches; b++) {
            const uint32_t S = std::min(Sb, spp - done);
            const int fresh = (b == 0 && !(fp->flags & RT_FLAG_ACCUMULATE)) ? 1 : 0;
            uint32_t *seg_ctr = s->counters.p + b, *n_hit = s->counters.p + n_cnt + b, *unit_ctr = s->counters.p + n_cnt + n_batches + b;
            k_prepare<<<div_up(tiles * 32u, kPrepThreads), kPrepThreads, 0, st>>>(W, H, S, D, bg, phit, s->rng_cur.p, s->rng_pix_off.p, fp->seed, fp->sample_offset + done,
                                                                                  s->hit_pix.p, n_hit, s->rng_start.p, npix, reinterpret_cast<float4 *>(out->accum_linear), fresh);
            s->launches += 1;
            if (D > 1) {
                PathArgs pa;
                pa.cam = cd; pa.W = W; pa.H = H; pa.S = S; pa.chunk = chunk; pa.n_chunks = div_up(S, chunk); pa.max_depth = D; pa.flags = fp->flags;
                pa.bg = bg; pa.phit = phit; pa.hit_pix = s->hit_pix.p; pa.n_hit = n_hit; pa.rng_start = s->rng_start.p; pa.cap = npix;
                pa.result = s->result.p; pa.unit_ctr = unit_ctr; pa.seg_ctr = seg_ctr;
                RT_CUDA(ctx, prof_mark(s, st));
                const int t1 = (int)s->t1, t2 = (int)s->t2, t3 = (int)s->t3, fm = (int)s->fetch_min;
                if (ref_mul)
                    k_path<6, true><<<grid, RT_EXTEND_THREADS, 0, st>>>(sd, pa, t1, t2, t3, fm, RT_DEQ_MAGIC);
                else if (s->extend_blocks_per_sm >= 8)
                    k_path<8, false><<<grid, RT_EXTEND_THREADS, 0, st>>>(sd, pa, t1, t2, t3, fm, RT_DEQ_MAGIC);
                else if (s->extend_blocks_per_sm == 7)
                    k_path<7, false><<<grid, RT_EXTEND_THREADS, 0, st>>>(sd, pa, t1, t2, t3, fm, RT_DEQ_MAGIC);
                else
                    k_path<6, false><<<grid, RT_EXTEND_THREADS, 0, st>>>(sd, pa, t1, t2, t3, fm, RT_DEQ_MAGIC);
                RT_CUDA(ctx, prof_mark(s, st));
                k_accumulate_hits<<<(uint32_t)ctx->sm_count * 8u, 256, 0, st>>>(s->result.p, s->hit_pix.p, n_hit, S, reinterpret_cast<float4 *>(out->accum_linear), fresh);
                s->launches += 2;
            }
            done += S;
        }
        s->rng_cur_off = (uint64_t)fp->sample_offset + spp;
    } else if (spp) {
        if (rng_forked) RT_CUDA(ctx, cudaStreamWaitEvent(st, s->ev_join, 0));
        const uint32_t shade_grid = (uint32_t)ctx->sm_count * 3u;  // 3 resident blocks of 512 threads (40 registers)
        uint32_t done = 0;
        for (uint32_t b = 0; b < n_batches; b++) {
            const uint32_t S = std::min(Sb, spp - done);
            uint32_t *cnt = s->counters.p + (size_t)b * (D + 3);
            const int fresh = (b == 0 && !(fp->flags & RT_FLAG_ACCUMULATE)) ? 1 : 0;
            uint32_t *n_hit = s->counters.p + n_cnt + 1 + (size_t)n_batches * (D + 1) + b;  // behind the fetch counters: not a segment count
            k_prepare<<<div_up(tiles * 32u, kPrepThreads), kPrepThreads, 0, st>>>(W, H, S, D, bg, phit, s->rng_cur.p, s->rng_pix_off.p, fp->seed, fp->sample_offset + done,
                                                                                  s->hit_pix.p, n_hit, s->rng_start.p, hit_cap, reinterpret_cast<float4 *>(out->accum_linear), fresh);
            if (D > 1)
                k_shade_first<<<(uint32_t)ctx->sm_count * (1024u / kFirstThreads), kFirstThreads, 0, st>>>(sd, cd, W, H, S, phit, s->hit_pix.p, n_hit, s->rng_start.p, hit_cap,
                                                                                      make_queue(s, 0, ref_mul), cnt + 2, s->ray_bins, fp->flags);
            s->launches += 2;
            for (uint32_t d = 2; d <= D; d++) {
                const Queue qin = make_queue(s, d & 1, ref_mul), qout = make_queue(s, (d + 1) & 1, ref_mul);
                const ExtendIn in{qin.o, qin.d, nullptr};
                const int rc = launch_extend(s, sd, in, cnt + d, 0, fetch++, s->hit.p, st);
                if (rc) return rc;
                if (ref_mul)
                    k_shade<512, true><<<shade_grid, 512, 0, st>>>(sd, qin, cnt + d, s->hit.p, qout, cnt + d + 1, s->result.p, d, D, bg, s->ray_bins, fp->flags);
                else if (s->shade_threads <= 256)  // small blocks (40 registers): one fits next to six resident traversal blocks
                    k_shade<256, false><<<(uint32_t)ctx->sm_count * 6u, 256, 0, st>>>(sd, qin, cnt + d, s->hit.p, qout, cnt + d + 1, s->result.p, d, D, bg, s->ray_bins, fp->flags);
                else if (s->shade_threads >= 1024)
                    k_shade<1024, false><<<(uint32_t)ctx->sm_count * 2u, 1024, 0, st>>>(sd, qin, cnt + d, s->hit.p, qout, cnt + d + 1, s->result.p, d, D, bg, s->ray_bins, fp->flags);
                else
                    k_shade<512, false><<<shade_grid, 512, 0, st>>>(sd, qin, cnt + d, s->hit.p, qout, cnt + d + 1, s->result.p, d, D, bg, s->ray_bins, fp->flags);
                s->launches += 1;
            }
            // hit pixels: ordered sum of their S path results on top of what k_prepare left (fresh / previous sum)
            if (D > 1) k_accumulate_hits<<<(uint32_t)ctx->sm_count * 8u, 256, 0, st>>>(s->result.p, s->hit_pix.p, n_hit, S, reinterpret_cast<float4 *>(out->accum_linear), fresh);
            s->launches += 1;
            done += S;
        }
        s->rng_cur_off = (uint64_t)fp->sample_offset + spp;
    }
    const uint32_t n_primary = s->ext_phit ? 0u : npix;  // supplied primaries were counted by rt_render_primary
    k_tally<<<1, 32, 0, st>>>(s->counters.p, (uint32_t)n_cnt, n_primary, s->seg_total.p);
    s->launches += 1;
    RT_CUDA(ctx, cudaEventRecord(s->ev[3], st));
    RT_CUDA(ctx, cudaGetLastError());
    s->have_render_ev = true;
    s->last_stream = st;
    s->last_primary = n_primary;
    s->last_n_cnt = n_cnt;
#ifdef RT_TRAV_STATS
    {
        unsigned long long h[8] = {0};
        cudaDeviceSynchronize();
        cudaMemcpyFromSymbol(h, g_trav_stats, sizeof h);
        fprintf(stderr, "trav stats: rays %llu visits %llu (%.2f/ray) none-hit %llu (%.3f) box-miss loose %llu (%.3f) tight %llu (%.3f) prim tests %llu (%.2f/ray)\n",
                h[5], h[0], (double)h[0] / (double)(h[5] ? h[5] : 1), h[1], (double)h[1] / (double)(h[0] ? h[0] : 1), h[2],
                (double)h[2] / (double)(h[0] ? h[0] : 1), h[3], (double)h[3] / (double)(h[0] ? h[0] : 1), h[4], (double)h[4] / (double)(h[5] ? h[5] : 1));
        memset(h, 0, sizeof h);
        cudaMemcpyToSymbol(g_trav_stats, h, sizeof h);
    }
#endif
    return RT_OK;
}

// Primary visibility of the pixel range [first, first + count) only: hit records (t, slot, u, v) into
// hits_out[first ...], indexed by absolute pixel.  Used by the multi-GPU spp split, where every rank
// would otherwise repeat the same primary pass.
int render_primary(Scene *s, const rt_camera *cam, uint32_t W, uint32_t H, size_t first, size_t count, float *hits_out, cudaStream_t st) {
    Context *ctx = s->ctx;
    if (!s->built) return set_error(ctx, RT_ERR_STATE, "rt_render_primary before rt_accel_build");
    if (!W || !H || (uint64_t)W * H >= (1ull << 31)) return set_error(ctx, RT_ERR_INVALID, "bad image size %ux%u", W, H);
    const uint32_t npix = W * H;
    if (first > npix || count > npix - first) return set_error(ctx, RT_ERR_INVALID, "pixel range outside the frame");
    if (!count) return RT_OK;
    for (int k = 0; k < 2; k++) {
        const size_t qk = k == 0 ? (size_t)npix : (s->q_o[1].cap ? s->q_o[1].cap : 1);
        RT_CUDA(ctx, s->q_o[k].reserve(qk, &s->bytes));
        RT_CUDA(ctx, s->q_d[k].reserve(qk, &s->bytes));
    }
    RT_CUDA(ctx, s->prim_fetch.reserve(1, &s->bytes));
    if (!s->seg_total.p) {
        RT_CUDA(ctx, s->seg_total.reserve(1, &s->bytes));
        RT_CUDA(ctx, cudaMemsetAsync(s->seg_total.p, 0, sizeof(unsigned long long), st));
    }
    RT_CUDA(ctx, cudaMemsetAsync(s->prim_fetch.p, 0, sizeof(uint32_t), st));
    const SceneDev sd = scene_dev(s);
    CamDev cd;
    cd.c = F3{cam->center[0], cam->center[1], cam->center[2]};
    cd.U = F3{cam->U[0], cam->U[1], cam->U[2]};
    cd.V = F3{cam->V[0], cam->V[1], cam->V[2]};
    cd.W = F3{cam->W[0], cam->W[1], cam->W[2]};
    const uint32_t tiles = div_up(W, 8) * div_up(H, 4);
    k_raygen<<<div_up(tiles * 32u, 128u), 128, 0, st>>>(cd, W, H, s->q_o[0].p, s->q_d[0].p);
    s->launches += 1;
    const ExtendIn in{s->q_o[0].p + first, s->q_d[0].p + first, nullptr};
    const int rc = launch_extend(s, sd, in, nullptr, (uint32_t)count, s->prim_fetch.p, reinterpret_cast<float4 *>(hits_out) + first, st);
    if (rc) return rc;
    k_export_hits<<<div_up((uint32_t)count, 256), 256, 0, st>>>(sd.slot_prim, reinterpret_cast<float4 *>(hits_out) + first, (uint32_t)count);
    k_tally<<<1, 32, 0, st>>>(s->prim_fetch.p, 0u, (uint32_t)count, s->seg_total.p);
    s->launches += 2;
    RT_CUDA(ctx, cudaGetLastError());
    s->last_stream = st;
    return RT_OK;
}

int resolve(Context *ctx, const float *accum, size_t n, float *color, uint8_t *rgba8, cudaStream_t st) {
    if (!n) return RT_OK;
    k_resolve<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(reinterpret_cast<const float4 *>(accum), n,
                                                         reinterpret_cast<float4 *>(color), reinterpret_cast<uchar4 *>(rgba8));
    RT_CUDA(ctx, cudaGetLastError());
    return RT_OK;
}

// ------------------------------------------------------------------ multi-GPU: sum of peer partial sums + resolve
// One pass over [first, first + n): 128-bit loads from up to RT_MAX_PARTS accumulation buffers (local
// HBM or peer memory over NVLink), fixed-order sum, display transform, stores to (possibly peer) outputs.
struct PartPtrs {
    const float4 *p[RT_MAX_PARTS];
};
__global__ void __launch_bounds__(256)
k_resolve_parts(PartPtrs parts, int n_parts, size_t first, size_t n, float4 *__restrict__ sum_out,
                float4 *__restrict__ color, uchar4 *__restrict__ rgba8) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const size_t p = first + k;
        float4 v[RT_MAX_PARTS];
#pragma unroll
        for (int j = 0; j < RT_MAX_PARTS; j++)
            if (j < n_parts) v[j] = __ldcg(parts.p[j] + p);  // all loads in flight before the first add
        float4 a = v[0];
#pragma unroll
        for (int j = 1; j < RT_MAX_PARTS; j++)
            if (j < n_parts) {
                a.x += v[j].x; a.y += v[j].y; a.z += v[j].z; a.w += v[j].w;
            }
        if (sum_out) sum_out[p] = a;
        if (color || rgba8) {
            const float inv = a.w > 0.0f ? 1.0f / a.w : 0.0f;
            const float sx = srgb(a.x * inv), sy = srgb(a.y * inv), sz = srgb(a.z * inv);
            if (color) color[p] = make_float4(sx, sy, sz, 1.0f);
            if (rgba8)
                rgba8[p] = make_uchar4((unsigned char)min((unsigned int)(sx * 256.0f), 255u),
                                       (unsigned char)min((unsigned int)(sy * 256.0f), 255u),
                                       (unsigned char)min((unsigned int)(sz * 256.0f), 255u), 255);
        }
    }
}

int resolve_parts(Context *ctx, const float *const *parts, int n_parts, size_t first, size_t n, float *sum_out, float *color,
                  uint8_t *rgba8, cudaStream_t st) {
    if (!n) return RT_OK;
    PartPtrs pp;
    for (int j = 0; j < RT_MAX_PARTS; j++) pp.p[j] = reinterpret_cast<const float4 *>(parts[j < n_parts ? j : 0]);
    const unsigned grid = (unsigned)std::min<size_t>((n + 255) / 256, (size_t)ctx->sm_count * 16);
    k_resolve_parts<<<grid, 256, 0, st>>>(pp, n_parts, first, n, reinterpret_cast<float4 *>(sum_out),
                                          reinterpret_cast<float4 *>(color), reinterpret_cast<uchar4 *>(rgba8));
    RT_CUDA(ctx, cudaGetLastError());
    return RT_OK;
}

int trace_rays(Scene *s, const float *rays, const uint32_t *skip, size_t n, float *t, uint32_t *id, cudaStream_t st) {
    Context *ctx = s->ctx;
    if (!s->built) return set_error(ctx, RT_ERR_STATE, "rt_trace_rays before rt_accel_build");
    if (!n) return RT_OK;
    if (n >= (1ull << 31)) return set_error(ctx, RT_ERR_INVALID, "too many rays");
    RT_CUDA(ctx, s->q_o[0].reserve(n, &s->bytes));
    RT_CUDA(ctx, s->q_d[0].reserve(n, &s->bytes));
    RT_CUDA(ctx, s->hit.reserve(n, &s->bytes));
    RT_CUDA(ctx, s->u_skip.reserve(n + 1, &s->bytes));
    // vals[1] is free after the build: reuse it as the gid -> slot map
    if (skip && s->n_prims) {
        k_invert_slots<<<div_up(s->n_prims, 256), 256, 0, st>>>(scene_dev(s).slot_prim, s->n_prims, s->vals[1].p);
        s->launches += 1;
    }
    k_pack_rays<<<div_up((uint32_t)n, 128), 128, 0, st>>>(rays, skip, s->vals[1].p, (uint32_t)n, s->q_o[0].p, s->q_d[0].p, s->u_skip.p);
    uint32_t *fetch = s->u_skip.p + n;
    RT_CUDA(ctx, cudaMemsetAsync(fetch, 0, sizeof(uint32_t), st));
    const ExtendIn in{s->q_o[0].p, s->q_d[0].p, s->u_skip.p};
    const int rc = launch_extend(s, scene_dev(s), in, nullptr, (uint32_t)n, fetch, s->hit.p, st);
    if (rc) return rc;
    k_unpack_hits<<<div_up((uint32_t)n, 128), 128, 0, st>>>(scene_dev(s), s->hit.p, (uint32_t)n, t, id);
    s->launches += 2;
    RT_CUDA(ctx, cudaGetLastError());
    s->last_stream = st;
    return RT_OK;
}

// ------------------------------------------------------------------ closest hits without any allocation
// rt_trace_rays_static (rtb200.h): the same launches as trace_rays() above, but the packed ray queue, the skip
// slots, the hit records and the fetch counter of the persistent traversal kernel live in ONE fixed array that is
// part of the library image instead of in DevBuf storage grown with cudaMalloc.  Meant for hosts that are not
// allowed to allocate at the moment they need a visibility query: a renderer thread inside a frame whose memory
// budget was fixed at start-up, a plugin whose host application owns the allocator, a process sharing the GPU
// with a job that has already taken the rest of the memory.
//
// g_trace_static is reserved on every device that loads librtb200.so, from load until the process exits.  All calls on a
// device use the same bytes: two calls whose GPU work overlaps (different streams, different host threads,
// different scenes) overwrite each other's rays and hits.  Contents persist between calls, so the one word that has
// to start at zero — the fetch counter — is cleared on the stream before each use; every other byte a call
// reads it has written itself first.
constexpr size_t kStaticRays = RT_TRACE_STATIC_MAX_RAYS;
constexpr size_t kStaticOffD = kStaticRays * sizeof(float4);                // packed directions + tmax
constexpr size_t kStaticOffHit = 2 * kStaticRays * sizeof(float4);          // hit records
constexpr size_t kStaticOffSkip = 3 * kStaticRays * sizeof(float4);         // skip slot per ray
constexpr size_t kStaticOffFetch = kStaticOffSkip + kStaticRays * sizeof(uint32_t);  // fetch counter (own 256-byte line)
constexpr size_t kStaticBytes = kStaticOffFetch + 256;
static __device__ __align__(256) unsigned char g_trace_static[kStaticBytes];

// In host code `g_trace_static` is only the handle the runtime registered for the symbol; the address the kernels need
// is looked up once per device (with that device current) and kept here.
constexpr int kStaticMaxDevices = 64;
static unsigned char *g_trace_static_addr[kStaticMaxDevices] = {};
static std::mutex g_trace_static_mu;

static int trace_static_base(Context *ctx, unsigned char **base) {
    int dev = -1;
    RT_CUDA(ctx, cudaGetDevice(&dev));
    if (dev != ctx->device) return set_error(ctx, RT_ERR_STATE, "rt_trace_rays_static: device %d is current, the context is bound to %d", dev, ctx->device);
    if (dev < 0 || dev >= kStaticMaxDevices) return set_error(ctx, RT_ERR_UNSUPPORTED, "rt_trace_rays_static: device ordinal %d beyond %d", dev, kStaticMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_trace_static_mu);
    if (!g_trace_static_addr[dev]) {
        void *p = nullptr;
        RT_CUDA(ctx, cudaGetSymbolAddress(&p, g_trace_static));
        if (!p || (reinterpret_cast<uintptr_t>(p) & 255u)) return set_error(ctx, RT_ERR_CUDA, "rt_trace_rays_static: bad address for the static scratch array");
        g_trace_static_addr[dev] = static_cast<unsigned char *>(p);
    }
    *base = g_trace_static_addr[dev];
    return RT_OK;
}

int trace_rays_static(Scene *s, const float *rays, const uint32_t *skip, size_t n, float *t, uint32_t *id, cudaStream_t st) {
    Context *ctx = s->ctx;
    if (!s->built) return set_error(ctx, RT_ERR_STATE, "rt_trace_rays_static before rt_accel_build");
    if (n > kStaticRays)
        return set_error(ctx, RT_ERR_INVALID, "rt_trace_rays_static: %zu rays exceed the static scratch array (RT_TRACE_STATIC_MAX_RAYS = %zu); split the batch or use rt_trace_rays",
                         n, kStaticRays);
    // first call on this device: resolves the array's device address (do it before any stream capture)
    unsigned char *base = nullptr;
    const int rc0 = trace_static_base(ctx, &base);
    if (rc0) return rc0;
    if (!n) return RT_OK;
    float4 *qo = reinterpret_cast<float4 *>(base);
    float4 *qd = reinterpret_cast<float4 *>(base + kStaticOffD);
    float4 *hit = reinterpret_cast<float4 *>(base + kStaticOffHit);
    uint32_t *qskip = reinterpret_cast<uint32_t *>(base + kStaticOffSkip);
    uint32_t *fetch = reinterpret_cast<uint32_t *>(base + kStaticOffFetch);
    const SceneDev sd = scene_dev(s);
    // gid -> slot map in vals[1], a buffer the build already owns (as in trace_rays)
    if (skip && s->n_prims) {
        k_invert_slots<<<div_up(s->n_prims, 256), 256, 0, st>>>(sd.slot_prim, s->n_prims, s->vals[1].p);
        s->launches += 1;
    }
    k_pack_rays<<<div_up((uint32_t)n, 128), 128, 0, st>>>(rays, skip, s->vals[1].p, (uint32_t)n, qo, qd, qskip);
    RT_CUDA(ctx, cudaMemsetAsync(fetch, 0, sizeof(uint32_t), st));
    const ExtendIn in{qo, qd, qskip};
    // the dispatch of launch_extend without its optional event timing (cudaEventCreate allocates)
    const uint32_t grid = std::min((uint32_t)ctx->sm_count * s->extend_blocks_per_sm, std::max(1u, div_up((uint32_t)n, 128u)));
    const int t1 = (int)s->t1, t2 = (int)s->t2, t3 = (int)s->t3, fm = (int)s->fetch_min;
    if (!s->wide)
        k_extend<<<grid, 128, 0, st>>>(sd, in, nullptr, (uint32_t)n, fetch, hit, t1, t2, fm);
    else if (s->extend_blocks_per_sm >= 8)
        k_extend_wide<8><<<grid, 128, 0, st>>>(sd, in, nullptr, (uint32_t)n, fetch, hit, t1, t2, t3, fm, RT_DEQ_MAGIC);
    else if (s->extend_blocks_per_sm == 7)
        k_extend_wide<7><<<grid, 128, 0, st>>>(sd, in, nullptr, (uint32_t)n, fetch, hit, t1, t2, t3, fm, RT_DEQ_MAGIC);
    else if (s->extend_blocks_per_sm >= 6)
        k_extend_wide<6><<<grid, 128, 0, st>>>(sd, in, nullptr, (uint32_t)n, fetch, hit, t1, t2, t3, fm, RT_DEQ_MAGIC);
    else
        k_extend_wide<5><<<grid, 128, 0, st>>>(sd, in, nullptr, (uint32_t)n, fetch, hit, t1, t2, t3, fm, RT_DEQ_MAGIC);
    k_unpack_hits<<<div_up((uint32_t)n, 128), 128, 0, st>>>(sd, hit, (uint32_t)n, t, id);
    s->launches += 3;
    RT_CUDA(ctx, cudaGetLastError());
    s->last_stream = st;
    return RT_OK;
}

}  // namespace rtb
