// rt_build.cu — per-frame LBVH build for sm_100a.
//
// Replaces the reference's closed-source acceleration-structure path: buildASImpl /
// buildGASForSpheres / buildGASForTriangles / buildGASForParticle / buildIAS / updateIAS
// (src/Global/RendererImpl.cu:30-242, all thin wrappers over optixAccelBuild).  Instead of a
// two-level GAS/IAS that is refitted per frame, instances are flattened to world space and one
// LBVH is rebuilt from scratch every frame:
//   flatten+bounds -> decide (huge prims, Morton frame) -> 30-bit Morton keys -> LSD radix sort
//   (8-bit digits, 4 passes) -> gather to Morton order -> Karras hierarchy -> bottom-up refit.
// All kernels are HBM-bound streaming passes; algorithmic bytes per primitive are listed in
// DESIGN.md §5.
#include <algorithm>
#include <cfloat>

#include <cooperative_groups.h>

#include "rt_device.cuh"

namespace cg = cooperative_groups;

// occupancy targets of the two register-heavy build kernels (min blocks per SM in __launch_bounds__);
// overridable at build time for experiments: make EXTRA=-DRT_COLLAPSE_MIN_BLOCKS=12
#ifndef RT_SORT_MIN_BLOCKS
#define RT_SORT_MIN_BLOCKS 4
#endif
#ifndef RT_COLLAPSE_MIN_BLOCKS
#define RT_COLLAPSE_MIN_BLOCKS 6
#endif

namespace rtb {

// ------------------------------------------------------------------ small device helpers
__device__ __forceinline__ uint32_t f2ord(float f) {  // order-preserving float -> uint
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k);
}

struct Box3 {
    F3 lo, hi;
};

// Unpadded world-space AABB of the primitive whose packed geometry word is g.
__device__ __forceinline__ Box3 prim_box(float4 g, const float4 *tri_v12) {
    Box3 b;
    const uint32_t wb = __float_as_uint(g.w);
    if (!(wb & 0x80000000u)) {
        b.lo = f3(g.x - g.w, g.y - g.w, g.z - g.w);
        b.hi = f3(g.x + g.w, g.y + g.w, g.z + g.w);
    } else {
        const uint32_t ti = wb & 0x7FFFFFFFu;
        const float4 v1 = tri_v12[2 * (size_t)ti];
        const float4 v2 = tri_v12[2 * (size_t)ti + 1];
        b.lo = f3(fminf(g.x, fminf(v1.x, v2.x)), fminf(g.y, fminf(v1.y, v2.y)), fminf(g.z, fminf(v1.z, v2.z)));
        b.hi = f3(fmaxf(g.x, fmaxf(v1.x, v2.x)), fmaxf(g.y, fmaxf(v1.y, v2.y)), fmaxf(g.z, fmaxf(v1.z, v2.z)));
    }
    return b;
}
// Conservative padding (2^-15 relative per coordinate): together with the 2e-6 widening in the
// slab test it dominates every rounding error of the box test and the intersectors.
__device__ __forceinline__ Box3 pad_box(Box3 b) {
    const float k = 3.0517578125e-5f;
    const float px = fmaxf(fabsf(b.lo.x), fabsf(b.hi.x)) * k + 1e-30f;
    const float py = fmaxf(fabsf(b.lo.y), fabsf(b.hi.y)) * k + 1e-30f;
    const float pz = fmaxf(fabsf(b.lo.z), fabsf(b.hi.z)) * k + 1e-30f;
    b.lo = f3(b.lo.x - px, b.lo.y - py, b.lo.z - pz);
    b.hi = f3(b.hi.x + px, b.hi.y + py, b.hi.z + pz);
    return b;
}
__device__ __forceinline__ int extent_bucket(const Box3 &b) {
    const float e = fmaxf(b.hi.x - b.lo.x, fmaxf(b.hi.y - b.lo.y, b.hi.z - b.lo.z));
    if (!(e > 0.0f)) return 0;
    int ex = (int)((__float_as_uint(e) >> 23) & 0xFF) - 127 + 64;
    return min(max(ex, 0), (int)kExtentBuckets - 1);
}

// ------------------------------------------------------------------ K1: flatten instances to world space
// One thread per world primitive.  Spec A.1: p' = ((m0*x + m1*y) + m2*z) + m3.
__device__ __forceinline__ F3 xform_point(const float *m, F3 p) {
    return f3(((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3], ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7],
              ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11]);
}
__device__ __forceinline__ F3 xform_normal(const float *m, F3 n) {
    F3 v = f3((m[0] * n.x + m[1] * n.y) + m[2] * n.z, (m[4] * n.x + m[5] * n.y) + m[6] * n.z,
              (m[8] * n.x + m[9] * n.y) + m[10] * n.z);
    return normalize3(v);
}

constexpr uint32_t kFlatItems = 4;  // primitives per thread of k_flatten
__global__ void __launch_bounds__(256)
k_flatten(const InstDev *__restrict__ insts, const uint32_t *__restrict__ inst_base, uint32_t n_inst,
          const GroupDev *__restrict__ groups, uint32_t n_prims, float4 *__restrict__ w_geom,
          uint32_t *__restrict__ w_mat, float4 *__restrict__ tri_v12, float4 *__restrict__ tri_nrm,
          uint32_t *__restrict__ bucket_tab, float4 *__restrict__ w_shade, uint32_t compat) {
    __shared__ uint32_t s_tab[kExtentBuckets * 8];
    __shared__ InstDev s_inst;
    __shared__ GroupDev s_group;
    __shared__ uint32_t s_uniform;
    for (uint32_t i = threadIdx.x; i < kExtentBuckets * 8; i += blockDim.x) {
        const uint32_t f = i & 7u;
        s_tab[i] = (f >= 1 && f <= 3) ? 0xFFFFFFFFu : 0u;
    }
    // The block's kFlatItems * 256 primitives almost always belong to ONE instance: look it up once and keep
    // its records in shared memory, so that a thread's first data load does not sit behind a chain of
    // four dependent global loads (search, instance, group, geometry).
    const uint32_t blk0 = blockIdx.x * (kFlatItems * 256u);
    if (threadIdx.x == 0) {
        uint32_t lo = 0, hi = n_inst;
        while (hi - lo > 1) {
            const uint32_t mid = (lo + hi) >> 1;
            if (__ldg(inst_base + mid) <= blk0) lo = mid; else hi = mid;
        }
        const uint32_t last = min(blk0 + kFlatItems * 256u, n_prims) - 1u;
        s_uniform = (__ldg(inst_base + lo + 1) > last) ? 1u : 0u;  // inst_base[n_inst] = n_prims
        s_inst = insts[lo];
        s_group = groups[insts[lo].group];
    }
    __syncthreads();
    const bool uniform = s_uniform != 0u;

#pragma unroll 1
  for (uint32_t item = 0; item < kFlatItems; item++) {
    const uint32_t gid = blk0 + item * 256u + threadIdx.x;
    const bool valid = gid < n_prims;
    int bucket = -1;
    F3 c = f3(0, 0, 0);
    if (valid) {
        const InstDev *in = &s_inst;
        if (!uniform) {
            // last instance whose prim_base <= gid
            uint32_t lo = 0, hi = n_inst;
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (__ldg(inst_base + mid) <= gid) lo = mid; else hi = mid;
            }
            in = insts + lo;
        }
        const uint32_t k = gid - in->prim_base;
        const GroupDev g = uniform ? s_group : groups[in->group];
        const uint32_t mat = in->mat + k * in->mat_stride;
        float4 wg;
        Box3 b;
        if (g.kind == 0) {
            const F3 oc = f3(g.a[3 * (size_t)k], g.a[3 * (size_t)k + 1], g.a[3 * (size_t)k + 2]);
            const float rr = g.b[k];
            if (w_shade) w_shade[gid] = make_float4(oc.x, oc.y, oc.z, rr);  // what Shader.cu:124-125 reads (quirk 1)
            if (in->identity) {
                // identical to xform_point for an identity linear part: ((1*x + 0*y) + 0*z) + t
                wg = make_float4(oc.x + in->m[3], oc.y + in->m[7], oc.z + in->m[11], rr);
            } else {
                const F3 wc = xform_point(in->m, oc);
                const float s = sqrtf((in->m[0] * in->m[0] + in->m[4] * in->m[4]) + in->m[8] * in->m[8]);
                wg = make_float4(wc.x, wc.y, wc.z, rr * s);
            }
            b = prim_box(wg, nullptr);
        } else {
            const uint32_t ti = in->tri_base + k;
            const float *v = g.a + 9 * (size_t)k;
            const F3 p0 = xform_point(in->m, f3(v[0], v[1], v[2]));
            const F3 p1 = xform_point(in->m, f3(v[3], v[4], v[5]));
            const F3 p2 = xform_point(in->m, f3(v[6], v[7], v[8]));
            wg = make_float4(p0.x, p0.y, p0.z, __uint_as_float(0x80000000u | ti));
            tri_v12[2 * (size_t)ti] = make_float4(p1.x, p1.y, p1.z, 0.0f);
            tri_v12[2 * (size_t)ti + 1] = make_float4(p2.x, p2.y, p2.z, 0.0f);
            const float *nn = g.b + 9 * (size_t)k;
            for (int j = 0; j < 3; j++) {
                F3 n = f3(nn[3 * j], nn[3 * j + 1], nn[3 * j + 2]);
                if (!in->identity && !(compat & 2u)) n = xform_normal(in->m, n);  // compat: Shader.cu:139-142 reads them unrotated (quirk 2)
                tri_nrm[3 * (size_t)ti + j] = make_float4(n.x, n.y, n.z, 0.0f);
            }
            b.lo = f3(fminf(p0.x, fminf(p1.x, p2.x)), fminf(p0.y, fminf(p1.y, p2.y)), fminf(p0.z, fminf(p1.z, p2.z)));
            b.hi = f3(fmaxf(p0.x, fmaxf(p1.x, p2.x)), fmaxf(p0.y, fmaxf(p1.y, p2.y)), fmaxf(p0.z, fmaxf(p1.z, p2.z)));
        }
        w_geom[gid] = wg;
        w_mat[gid] = mat;
        bucket = extent_bucket(b);
        c = f3(0.5f * (b.lo.x + b.hi.x), 0.5f * (b.lo.y + b.hi.y), 0.5f * (b.lo.z + b.hi.z));
    }

    // per-extent-bucket centroid bounds: warp-aggregate when the whole warp shares a bucket
    const uint32_t act = __ballot_sync(0xFFFFFFFFu, valid);
    if (valid) {
        const uint32_t peers = __match_any_sync(act, bucket);
        uint32_t kx0 = f2ord(c.x), ky0 = f2ord(c.y), kz0 = f2ord(c.z);
        uint32_t kx1 = kx0, ky1 = ky0, kz1 = kz0;
        uint32_t cnt = 1;
        if (peers == act && act == 0xFFFFFFFFu) {
            for (int o = 16; o > 0; o >>= 1) {
                kx0 = min(kx0, __shfl_xor_sync(0xFFFFFFFFu, kx0, o));
                ky0 = min(ky0, __shfl_xor_sync(0xFFFFFFFFu, ky0, o));
                kz0 = min(kz0, __shfl_xor_sync(0xFFFFFFFFu, kz0, o));
                kx1 = max(kx1, __shfl_xor_sync(0xFFFFFFFFu, kx1, o));
                ky1 = max(ky1, __shfl_xor_sync(0xFFFFFFFFu, ky1, o));
                kz1 = max(kz1, __shfl_xor_sync(0xFFFFFFFFu, kz1, o));
            }
            cnt = 32;
            if ((threadIdx.x & 31) != 0) cnt = 0;
        }
        if (cnt) {
            uint32_t *t = s_tab + bucket * 8;
            atomicAdd(t + 0, cnt);
            atomicMin(t + 1, kx0); atomicMin(t + 2, ky0); atomicMin(t + 3, kz0);
            atomicMax(t + 4, kx1); atomicMax(t + 5, ky1); atomicMax(t + 6, kz1);
        }
    }
  }
    __syncthreads();
    for (uint32_t b = threadIdx.x; b < kExtentBuckets; b += blockDim.x) {
        const uint32_t *t = s_tab + b * 8;
        if (t[0]) {
            uint32_t *g = bucket_tab + b * 8;
            atomicAdd(g + 0, t[0]);
            atomicMin(g + 1, t[1]); atomicMin(g + 2, t[2]); atomicMin(g + 3, t[3]);
            atomicMax(g + 4, t[4]); atomicMax(g + 5, t[5]); atomicMax(g + 6, t[6]);
        }
    }
}

__global__ void k_init_buckets(uint32_t *bucket_tab) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < kExtentBuckets * 8) {
        const uint32_t f = i & 7u;
        bucket_tab[i] = (f >= 1 && f <= 3) ? 0xFFFFFFFFu : 0u;
    }
}

// ------------------------------------------------------------------ K2: decide huge prims + Morton frame
// Prims whose extent bucket exceeds the median bucket by more than 6 (>= 64x the median extent,
// e.g. the r=1000 ground sphere of files/config.json:20-30 among r=0.05 particles) are kept out of
// the tree (at most kMaxHuge of them) and tested linearly by every ray.
__global__ void k_decide(const uint32_t *__restrict__ bucket_tab, uint32_t n_total, BuildInfo *info) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint32_t run = 0;
    int median = (int)kExtentBuckets - 1;
    for (int b = 0; b < (int)kExtentBuckets; b++) {
        run += bucket_tab[b * 8];
        if (2 * (uint64_t)run >= (uint64_t)n_total) { median = b; break; }
    }
    int huge_bucket = median + 6;
    uint32_t n_huge = 0;
    for (int b = huge_bucket + 1; b < (int)kExtentBuckets; b++) n_huge += bucket_tab[b * 8];
    if (n_huge > kMaxHuge || n_total < 2) {
        huge_bucket = (int)kExtentBuckets;
        n_huge = 0;
    }
    uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0, 0, 0};
    for (int b = 0; b < (int)kExtentBuckets && b <= huge_bucket; b++) {
        if (!bucket_tab[b * 8]) continue;
        for (int a = 0; a < 3; a++) {
            lo[a] = min(lo[a], bucket_tab[b * 8 + 1 + a]);
            hi[a] = max(hi[a], bucket_tab[b * 8 + 4 + a]);
        }
    }
    info->n_total = n_total;
    info->n_huge = n_huge;
    info->n_bvh = n_total - n_huge;
    info->huge_bucket = huge_bucket;
    for (int a = 0; a < 3; a++) {
        const float l = (n_total - n_huge) ? ord2f(lo[a]) : 0.0f;
        const float h = (n_total - n_huge) ? ord2f(hi[a]) : 0.0f;
        info->lo[a] = l;
        info->scale[a] = (h > l) ? 1024.0f / (h - l) : 0.0f;
        info->scene_lo[a] = 0.0f;
        info->scene_hi[a] = 0.0f;
    }
    for (uint32_t i = 0; i < kMaxHuge; i++) info->huge_slot[i] = (n_total - n_huge) + i;
    info->n_wnodes = 1;  // the root's slot; k_collapse allocates the rest
    info->n_wprims = 0;
    info->level_end = 0;
    info->pad_ = 0;
    info->pad2_[0] = 0;  // k_collapse's grid barrier: arrivals, generation
    info->pad2_[1] = 0;
}

// ------------------------------------------------------------------ K3: 30-bit Morton keys
__device__ __forceinline__ uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__global__ void __launch_bounds__(256)
k_morton(const float4 *__restrict__ w_geom, const float4 *__restrict__ tri_v12, uint32_t n,
         const BuildInfo *__restrict__ info, uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= n) return;
    const Box3 b = prim_box(w_geom[gid], tri_v12);
    uint32_t key;
    if (extent_bucket(b) > info->huge_bucket) {
        key = 0xFFFFFFFFu;
    } else {
        const float cx = 0.5f * (b.lo.x + b.hi.x), cy = 0.5f * (b.lo.y + b.hi.y), cz = 0.5f * (b.lo.z + b.hi.z);
        const float qx = fminf(fmaxf((cx - info->lo[0]) * info->scale[0], 0.0f), 1023.0f);
        const float qy = fminf(fmaxf((cy - info->lo[1]) * info->scale[1], 0.0f), 1023.0f);
        const float qz = fminf(fmaxf((cz - info->lo[2]) * info->scale[2], 0.0f), 1023.0f);
        key = (expand10((uint32_t)qx) << 2) | (expand10((uint32_t)qy) << 1) | expand10((uint32_t)qz);
    }
    keys[gid] = key;
    vals[gid] = gid;
}

// ------------------------------------------------------------------ K4: LSD radix sort, 8-bit digits
// One-sweep organisation (Adinets & Merrill 2022): ONE histogram kernel reads the keys once and
// counts all four digits; each of the four passes is then a single kernel in which a tile of 4096
// keys ranks its keys, obtains the number of equal-digit keys in the tiles before it by a decoupled
// look-back over per-tile status words (flag + count in one 32-bit word), and scatters.  Tiles take
// their index from an atomic ticket so that a tile only ever waits for tiles that already run.
constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 keys per block
constexpr uint32_t kFlagAggregate = 1u << 30, kFlagPrefix = 2u << 30, kFlagMask = 3u << 30;
// layout of the sort scratch buffer (u32 words): [4][256] digit histograms, [4] tickets, [4][ntiles][256] status
__host__ __device__ inline size_t sort_scratch_words(uint32_t ntiles) { return 4 * 256 + 4 + 4 * (size_t)ntiles * 256; }

__global__ void __launch_bounds__(kSortThreads)
k_sort_hist(const uint32_t *__restrict__ keys, uint32_t n, uint32_t *__restrict__ ghist) {
    __shared__ uint32_t h[4][256];
    for (int i = threadIdx.x; i < 4 * 256; i += kSortThreads) (&h[0][0])[i] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * kSortTile;
#pragma unroll 4
    for (int j = 0; j < kSortItems; j++) {
        const uint32_t i = base + j * kSortThreads + threadIdx.x;
        if (i < n) {
            const uint32_t k = keys[i];
            atomicAdd(&h[0][k & 255u], 1u);
            atomicAdd(&h[1][(k >> 8) & 255u], 1u);
            atomicAdd(&h[2][(k >> 16) & 255u], 1u);
            atomicAdd(&h[3][k >> 24], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 4 * 256; i += kSortThreads) {
        const uint32_t v = (&h[0][0])[i];
        if (v) atomicAdd(ghist + i, v);
    }
}

__global__ void __launch_bounds__(kSortThreads, RT_SORT_MIN_BLOCKS)
k_sort_pass(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
            uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, uint32_t n, int pass,
            uint32_t *__restrict__ scratch, uint32_t ntiles) {
    __shared__ uint32_t cnt[8][256];
    __shared__ uint32_t s_digit_base[256];  // global position of the tile's first key of each digit, minus its tile-local start
    __shared__ uint32_t s_local[256];       // tile-local start of each digit
    __shared__ uint32_t s_wsum[8];
    __shared__ uint32_t s_tile;
    __shared__ uint32_t s_key[kSortTile];   // the tile sorted by digit: staged so that the global writes are runs
    __shared__ uint32_t s_val[kSortTile];
    const uint32_t *ghist = scratch + 256 * pass;
    uint32_t *ticket = scratch + 4 * 256 + pass;
    volatile uint32_t *status = scratch + 4 * 256 + 4 + (size_t)pass * ntiles * 256;
    const int shift = 8 * pass;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = threadIdx.x; i < 8 * 256; i += kSortThreads) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t base = tile * kSortTile + w * (kSortTile / 8);
    const uint32_t tile_n = min((uint32_t)kSortTile, n - tile * kSortTile);
    uint32_t key[kSortItems], val[kSortItems];
    uint16_t rank[kSortItems];
#pragma unroll
    for (int c = 0; c < kSortItems; c++) {
        const uint32_t i = base + c * 32 + lane;
        const bool valid = i < n;
        key[c] = valid ? keys_in[i] : 0xFFFFFFFFu;
        val[c] = valid ? vals_in[i] : 0u;
    }
#pragma unroll
    for (int c = 0; c < kSortItems; c++) {
        const bool valid = base + c * 32 + lane < n;
        const uint32_t dg = valid ? ((key[c] >> shift) & 255u) : 256u;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, dg);
        const int leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (valid && (int)lane == leader) {
            pre = cnt[w][dg];
            cnt[w][dg] = pre + __popc(peers);
        }
        pre = __shfl_sync(0xFFFFFFFFu, pre, leader);
        rank[c] = (uint16_t)(pre + __popc(peers & lt));
        __syncwarp();
    }
    __syncthreads();
    {
        // thread dg: per-warp exclusive offsets inside the tile, tile total, look-back, global base
        const uint32_t dg = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int ww = 0; ww < 8; ww++) {
            const uint32_t t = cnt[ww][dg];
            cnt[ww][dg] = run;
            run += t;
        }
        const uint32_t total = run;
        if (tile + 1 < ntiles) status[(size_t)tile * 256 + dg] = (tile == 0 ? kFlagPrefix : kFlagAggregate) | total;
        // tile-local exclusive prefix over digits (where each digit's run starts in the staged tile)
        uint32_t linc = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, linc, o);
            if (lane >= (uint32_t)o) linc += t;
        }
        if (lane == 31) s_wsum[w] = linc;
        __syncthreads();
        uint32_t lbase = 0;
        for (uint32_t ww = 0; ww < w; ww++) lbase += s_wsum[ww];
        const uint32_t lstart = lbase + linc - total;
        s_local[dg] = lstart;
        __syncthreads();
        // stage the tile in digit order while the look-back below waits on earlier tiles
#pragma unroll
        for (int c = 0; c < kSortItems; c++) {
            if (base + c * 32 + lane < n) {
                const uint32_t d2 = (key[c] >> shift) & 255u;
                const uint32_t lp = s_local[d2] + cnt[w][d2] + rank[c];
                s_key[lp] = key[c];
                s_val[lp] = val[c];
            }
        }
        uint32_t excl = 0;
        for (int t = (int)tile - 1; t >= 0;) {
            const uint32_t v = status[(size_t)t * 256 + dg];
            if (!(v & kFlagMask)) continue;  // not published yet: tile t runs (it holds an earlier ticket)
            excl += v & ~kFlagMask;
            if (v & kFlagPrefix) break;
            t--;
        }
        if (tile > 0 && tile + 1 < ntiles) status[(size_t)tile * 256 + dg] = kFlagPrefix | (excl + total);
        // exclusive prefix of the global digit histogram (256 bins: one warp-scan per 32, then 8 partial sums)
        uint32_t g = ghist[dg], inc = g;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, inc, o);
            if (lane >= (uint32_t)o) inc += t;
        }
        __syncthreads();
        if (lane == 31) s_wsum[w] = inc;
        __syncthreads();
        uint32_t wbase = 0;
        for (uint32_t ww = 0; ww < w; ww++) wbase += s_wsum[ww];
        s_digit_base[dg] = wbase + inc - g + excl - lstart;
    }
    __syncthreads();
    // consecutive threads write consecutive staged keys: each digit's run lands contiguously in global memory
    for (uint32_t j = threadIdx.x; j < tile_n; j += kSortThreads) {
        const uint32_t k = s_key[j];
        const uint32_t pos = s_digit_base[(k >> shift) & 255u] + j;
        keys_out[pos] = k;
        vals_out[pos] = s_val[j];
    }
}

// ------------------------------------------------------------------ K5: gather primitives into Morton order
__global__ void __launch_bounds__(256)
k_gather(const uint32_t *__restrict__ vals, const float4 *__restrict__ w_geom, const uint32_t *__restrict__ w_mat,
         uint32_t n, float4 *__restrict__ slot_geom, uint32_t *__restrict__ slot_prim, uint32_t *__restrict__ slot_mat) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t gid = vals[i];
    slot_geom[i] = w_geom[gid];
    slot_prim[i] = gid;
    slot_mat[i] = w_mat[gid];
}

// ------------------------------------------------------------------ K6: Karras (2012) hierarchy
// delta(i,j) = length of the common prefix of keys i and j; equal keys fall back to the index so
// the tree is well defined for duplicate Morton codes.
__device__ __forceinline__ int delta(const uint32_t *__restrict__ keys, int n, int i, uint32_t key_i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint32_t b = __ldg(keys + j);
    if (key_i == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clz(key_i ^ b);
}
__device__ __forceinline__ int encode_child(int first, int last, int internal_idx, int leaf_max) {
    const int cnt = last - first + 1;
    if (cnt <= leaf_max) return (int)(kLeafFlag | ((uint32_t)(cnt - 1) << kLeafCountShift) | (uint32_t)first);
    return internal_idx;
}
constexpr uint32_t kRefitBlock = 512;  // leaves per block of k_refit_local
constexpr uint32_t kLinkParentCross = 0x80000000u, kLinkSelfCross = 0x40000000u, kLinkIndexMask = 0x0FFFFFFFu;
__global__ void __launch_bounds__(256)
k_hierarchy(const uint32_t *__restrict__ keys, const BuildInfo *__restrict__ info, Node *__restrict__ nodes,
            uint32_t *__restrict__ parent_int, uint32_t *__restrict__ parent_leaf, int leaf_max) {
    const int n = (int)info->n_bvh;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < 2 || i >= n - 1) return;
    const uint32_t ki = __ldg(keys + i);
    const int d = (delta(keys, n, i, ki, i + 1) - delta(keys, n, i, ki, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, ki, i - d);
    int lmax = 2;
    while (delta(keys, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, ki, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, ki, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (delta(keys, n, i, ki, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int split = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    // topology links (always the full Karras tree: the refit climbs through collapsed nodes too)
    // link = parent index << 1 | which child, plus two flags for the two-kernel refit: bit 31 = the parent's
    // leaf range crosses a boundary of kRefitBlock consecutive leaves, bit 30 = the child's own range does
    const uint32_t pc = ((uint32_t)first / kRefitBlock != (uint32_t)last / kRefitBlock) ? kLinkParentCross : 0u;
    const uint32_t c0c = ((uint32_t)first / kRefitBlock != (uint32_t)split / kRefitBlock) ? kLinkSelfCross : 0u;
    const uint32_t c1c = ((uint32_t)(split + 1) / kRefitBlock != (uint32_t)last / kRefitBlock) ? kLinkSelfCross : 0u;
    if (split == first) parent_leaf[split] = ((uint32_t)i << 1) | 0u | pc; else parent_int[split] = ((uint32_t)i << 1) | 0u | pc | c0c;
    if (split + 1 == last) parent_leaf[split + 1] = ((uint32_t)i << 1) | 1u | pc; else parent_int[split + 1] = ((uint32_t)i << 1) | 1u | pc | c1c;
    if (i == 0) parent_int[0] = 0xFFFFFFFFu;
    // child pointers seen by traversal: ranges of <= leaf_max prims become one leaf
    const int c0 = encode_child(first, split, split, leaf_max);
    const int c1 = encode_child(split + 1, last, split + 1, leaf_max);
    // .z/.w: the node's leaf range, used by the refit to tell warp-local subtrees from shared ones
    nodes[i].n3 = make_float4(__int_as_float(c0), __int_as_float(c1), __int_as_float(first), __int_as_float(last));
}

// ------------------------------------------------------------------ K7: bottom-up refit
__device__ __forceinline__ void store_child_box(Node *nd, uint32_t which, const Box3 &b) {
    float4 *p = reinterpret_cast<float4 *>(nd);
    if (which == 0) {
        __stcg(p + 0, make_float4(b.lo.x, b.hi.x, b.lo.y, b.hi.y));
        __stcg(reinterpret_cast<float2 *>(p + 2), make_float2(b.lo.z, b.hi.z));
    } else {
        __stcg(p + 1, make_float4(b.lo.x, b.hi.x, b.lo.y, b.hi.y));
        __stcg(reinterpret_cast<float2 *>(p + 2) + 1, make_float2(b.lo.z, b.hi.z));
    }
}
// Bottom-up with "second arriver continues" counters (Karras 2012).  Two variants that hand the boxes of
// warp-local / block-local subtrees over through shuffles or shared memory were measured slower on
// B200 (1.07 ms and 6.1 ms against 0.75 ms at 10 M primitives, profiles/r1l, r1m): the climb is a chain
// of dependent DRAM reads either way and the variants keep mostly-empty warps resident for longer.
__global__ void __launch_bounds__(256)
k_refit(const float4 *__restrict__ slot_geom, const float4 *__restrict__ tri_v12, BuildInfo *info,
        Node *nodes, const uint32_t *__restrict__ parent_int, const uint32_t *__restrict__ parent_leaf,
        uint32_t *visit) {
    const uint32_t n = info->n_bvh;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < 2 || i >= n) return;
    Box3 b = pad_box(prim_box(slot_geom[i], tri_v12));
    uint32_t link = parent_leaf[i];
    while (true) {
        const uint32_t p = (link >> 1) & kLinkIndexMask, which = link & 1u;
        store_child_box(nodes + p, which, b);
        __threadfence();
        if (atomicAdd(visit + p, 1u) == 0u) return;  // sibling not done yet
        // second arriver: read the sibling's box (written before its fence) from L2
        const float4 *q = reinterpret_cast<const float4 *>(nodes + p);
        const float4 sxy = __ldcg(q + (which ^ 1u));
        const float2 sz = __ldcg(reinterpret_cast<const float2 *>(q + 2) + (which ^ 1u));
        b.lo = f3(fminf(b.lo.x, sxy.x), fminf(b.lo.y, sxy.z), fminf(b.lo.z, sz.x));
        b.hi = f3(fmaxf(b.hi.x, sxy.y), fmaxf(b.hi.y, sxy.w), fmaxf(b.hi.z, sz.y));
        link = parent_int[p];
        if (link == 0xFFFFFFFFu) {
            info->scene_lo[0] = b.lo.x; info->scene_lo[1] = b.lo.y; info->scene_lo[2] = b.lo.z;
            info->scene_hi[0] = b.hi.x; info->scene_hi[1] = b.hi.y; info->scene_hi[2] = b.hi.z;
            return;
        }
    }
}

// Two-kernel variant (option "refit_mode" = 1).  k_refit_local: one block per kRefitBlock consecutive
// leaves climbs every subtree that lies inside the block, handing boxes over through shared memory
// (shared counter + block-level fence, no global atomics, no device-wide fences) and stops at the first
// parent whose range crosses a block boundary.  k_refit_top: the nodes that stopped there (a leaf or an
// in-block subtree root whose parent crosses) carry on with the global counter scheme; that is ~n/50
// starters instead of n.
__global__ void __launch_bounds__(kRefitBlock)
k_refit_local(const float4 *__restrict__ slot_geom, const float4 *__restrict__ tri_v12, BuildInfo *info, Node *nodes,
              const uint32_t *__restrict__ parent_int, const uint32_t *__restrict__ parent_leaf) {
    __shared__ float s_box[2][6][kRefitBlock];  // [which child][lo.xyz hi.xyz][node slot]
    __shared__ uint32_t s_visit[kRefitBlock];
    __shared__ uint32_t s_link[kRefitBlock];  // parent links of the block's internal nodes: one coalesced load
    const uint32_t n = info->n_bvh;           // instead of a dependent global load per level of the climb
    const uint32_t b0 = blockIdx.x * kRefitBlock;
    const uint32_t i = b0 + threadIdx.x;
    s_visit[threadIdx.x] = 0;
    s_link[threadIdx.x] = (i + 1u < n) ? parent_int[i] : 0xFFFFFFFFu;  // internal nodes are 0 .. n-2
    __syncthreads();
    if (n < 2 || i >= n) return;
    Box3 b = pad_box(prim_box(slot_geom[i], tri_v12));
    uint32_t link = parent_leaf[i];
    for (;;) {
        if (link == 0xFFFFFFFFu) {  // the whole tree sits in this block
            info->scene_lo[0] = b.lo.x; info->scene_lo[1] = b.lo.y; info->scene_lo[2] = b.lo.z;
            info->scene_hi[0] = b.hi.x; info->scene_hi[1] = b.hi.y; info->scene_hi[2] = b.hi.z;
            return;
        }
        const uint32_t p = (link >> 1) & kLinkIndexMask, which = link & 1u;
        if (link & kLinkParentCross) {  // k_refit_top continues from here
            store_child_box(nodes + p, which, b);
            return;
        }
        const uint32_t l = p - b0;  // a Karras node's index is one end of its range: it lies in this block
        const uint32_t next = s_link[l];
        float(*mine)[kRefitBlock] = s_box[which];
        mine[0][l] = b.lo.x; mine[1][l] = b.lo.y; mine[2][l] = b.lo.z;
        mine[3][l] = b.hi.x; mine[4][l] = b.hi.y; mine[5][l] = b.hi.z;
        __threadfence_block();
        if (atomicAdd(&s_visit[l], 1u) == 0u) return;  // the sibling will pick this box up
        __threadfence_block();
        volatile float(*sib)[kRefitBlock] = s_box[which ^ 1u];
        Box3 o;
        o.lo = f3(sib[0][l], sib[1][l], sib[2][l]);
        o.hi = f3(sib[3][l], sib[4][l], sib[5][l]);
        {   // the second arriver holds both child boxes: three full 128-bit stores instead of four partial ones
            const Box3 &c0 = which ? o : b, &c1 = which ? b : o;
            float4 *q = reinterpret_cast<float4 *>(nodes + p);
            __stcg(q + 0, make_float4(c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y));
            __stcg(q + 1, make_float4(c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y));
            __stcg(q + 2, make_float4(c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z));
        }
        b.lo = f3(fminf(b.lo.x, o.lo.x), fminf(b.lo.y, o.lo.y), fminf(b.lo.z, o.lo.z));
        b.hi = f3(fmaxf(b.hi.x, o.hi.x), fmaxf(b.hi.y, o.hi.y), fmaxf(b.hi.z, o.hi.z));
        link = next;
    }
}

// Range-query variant of the block-local part (option "refit_mode" = 2).  A Karras node's box is the union of
// the leaf boxes of its range [first, last] (k_hierarchy keeps the range in n3.zw), and the 511 internal nodes whose
// index falls into a block of kRefitBlock consecutive leaves either have their whole range inside that block or cross
// it.  So the block builds a sparse table of windowed box unions over its 512 leaf boxes in shared memory (window
// 2^k at level k: 4 107 entries x 24 B = 96 KB) and every non-crossing node reads the boxes of its two children as
// two range queries of two table look-ups each: no climb, no counters, no fences, no idle lanes (k_refit_local keeps
// 9.5 of 32 lanes busy, ncu r4f), and the node's three 128-bit words are written once.  What k_refit_top expects
// stays the same: a leaf or non-crossing node whose PARENT crosses also stores its own box into the parent's slot.
// Measured (10 M spheres, r4i): 0.43 ms, the same as k_refit_local — with 96 KB of table per block only two blocks fit
// an SM and each runs load -> 9 synchronised table levels -> store with nothing to overlap the DRAM latency of the
// next tile, so both variants sit at ~2.7 TB/s.  Deterministic and atomic-free, kept as an option.
constexpr uint32_t kRmqLevels = 10;  // windows 1 .. 512
__host__ __device__ constexpr uint32_t rmq_offset(uint32_t k) {  // first entry of level k: sum_{j<k} (513 - 2^j)
    return k * (kRefitBlock + 1u) - ((1u << k) - 1u);
}
constexpr uint32_t kRmqEntries = rmq_offset(kRmqLevels);  // 4 107
__device__ __forceinline__ Box3 rmq_query(const float *__restrict__ t, uint32_t l, uint32_t r) {
    const uint32_t k = 31u - (uint32_t)__clz(r - l + 1u);
    const uint32_t a = rmq_offset(k) + l, b = rmq_offset(k) + r + 1u - (1u << k);
    Box3 o;
    o.lo = f3(fminf(t[0 * kRmqEntries + a], t[0 * kRmqEntries + b]), fminf(t[1 * kRmqEntries + a], t[1 * kRmqEntries + b]),
              fminf(t[2 * kRmqEntries + a], t[2 * kRmqEntries + b]));
    o.hi = f3(fmaxf(t[3 * kRmqEntries + a], t[3 * kRmqEntries + b]), fmaxf(t[4 * kRmqEntries + a], t[4 * kRmqEntries + b]),
              fmaxf(t[5 * kRmqEntries + a], t[5 * kRmqEntries + b]));
    return o;
}
__global__ void __launch_bounds__(kRefitBlock, 2)
k_refit_rmq(const float4 *__restrict__ slot_geom, const float4 *__restrict__ tri_v12, BuildInfo *info, Node *nodes,
            const uint32_t *__restrict__ parent_int, const uint32_t *__restrict__ parent_leaf) {
    extern __shared__ float s_t[];  // [6][kRmqEntries]: lo.xyz hi.xyz
    const uint32_t n = info->n_bvh;
    const uint32_t b0 = blockIdx.x * kRefitBlock, tx = threadIdx.x;
    const uint32_t i = b0 + tx;
    if (n < 2 || b0 >= n) return;  // block-uniform
    const float inf = __int_as_float(0x7F800000);
    Box3 b;
    b.lo = f3(inf, inf, inf);
    b.hi = f3(-inf, -inf, -inf);
    if (i < n) b = pad_box(prim_box(slot_geom[i], tri_v12));
    s_t[0 * kRmqEntries + tx] = b.lo.x; s_t[1 * kRmqEntries + tx] = b.lo.y; s_t[2 * kRmqEntries + tx] = b.lo.z;
    s_t[3 * kRmqEntries + tx] = b.hi.x; s_t[4 * kRmqEntries + tx] = b.hi.y; s_t[5 * kRmqEntries + tx] = b.hi.z;
    __syncthreads();
#pragma unroll 1
    for (uint32_t k = 1; k < kRmqLevels; k++) {
        const uint32_t h = 1u << (k - 1u), src = rmq_offset(k - 1u), dst = rmq_offset(k);
        if (tx + (h << 1) <= kRefitBlock) {
#pragma unroll
            for (uint32_t c = 0; c < 3; c++) {
                s_t[c * kRmqEntries + dst + tx] = fminf(s_t[c * kRmqEntries + src + tx], s_t[c * kRmqEntries + src + tx + h]);
                s_t[(c + 3u) * kRmqEntries + dst + tx] = fmaxf(s_t[(c + 3u) * kRmqEntries + src + tx], s_t[(c + 3u) * kRmqEntries + src + tx + h]);
            }
        }
        __syncthreads();
    }
    // leaves whose parent crosses the block: their box goes into the parent's slot
    if (i < n) {
        const uint32_t link = parent_leaf[i];
        if (link != 0xFFFFFFFFu && (link & kLinkParentCross)) store_child_box(nodes + ((link >> 1) & kLinkIndexMask), link & 1u, b);
    }
    // internal node i (0 .. n-2): one end of its range is leaf i
    if (i + 1u < n) {
        const float4 n3 = __ldg(reinterpret_cast<const float4 *>(nodes + i) + 3);
        const uint32_t first = __float_as_uint(n3.z), last = __float_as_uint(n3.w);
        if (first >= b0 && last < b0 + kRefitBlock) {
            const int c1 = __float_as_int(n3.y);
            const uint32_t split1 = c1 >= 0 ? (uint32_t)c1 : ((uint32_t)c1 & kLeafFirstMask);  // first leaf of child 1
            const Box3 x0 = rmq_query(s_t, first - b0, split1 - 1u - b0), x1 = rmq_query(s_t, split1 - b0, last - b0);
            float4 *q = reinterpret_cast<float4 *>(nodes + i);
            __stcg(q + 0, make_float4(x0.lo.x, x0.hi.x, x0.lo.y, x0.hi.y));
            __stcg(q + 1, make_float4(x1.lo.x, x1.hi.x, x1.lo.y, x1.hi.y));
            __stcg(q + 2, make_float4(x0.lo.z, x0.hi.z, x1.lo.z, x1.hi.z));
            const uint32_t link = parent_int[i];
            if (link == 0xFFFFFFFFu || (link & kLinkParentCross)) {
                Box3 u;
                u.lo = f3(fminf(x0.lo.x, x1.lo.x), fminf(x0.lo.y, x1.lo.y), fminf(x0.lo.z, x1.lo.z));
                u.hi = f3(fmaxf(x0.hi.x, x1.hi.x), fmaxf(x0.hi.y, x1.hi.y), fmaxf(x0.hi.z, x1.hi.z));
                if (link == 0xFFFFFFFFu) {  // the whole tree sits in this block
                    info->scene_lo[0] = u.lo.x; info->scene_lo[1] = u.lo.y; info->scene_lo[2] = u.lo.z;
                    info->scene_hi[0] = u.hi.x; info->scene_hi[1] = u.hi.y; info->scene_hi[2] = u.hi.z;
                } else {
                    store_child_box(nodes + ((link >> 1) & kLinkIndexMask), link & 1u, u);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(256)
k_refit_top(const float4 *__restrict__ slot_geom, const float4 *__restrict__ tri_v12, BuildInfo *info, Node *nodes,
            const uint32_t *__restrict__ parent_int, const uint32_t *__restrict__ parent_leaf, uint32_t *visit) {
    const uint32_t n = info->n_bvh;
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < 2 || t >= 2 * n - 1) return;
    uint32_t link;
    Box3 b;
    if (t < n) {
        link = parent_leaf[t];
        if (!(link & kLinkParentCross)) return;
        b = pad_box(prim_box(slot_geom[t], tri_v12));
    } else {
        const uint32_t q = t - n;
        link = parent_int[q];
        if (link == 0xFFFFFFFFu || !(link & kLinkParentCross) || (link & kLinkSelfCross)) return;
        const float4 *c = reinterpret_cast<const float4 *>(nodes + q);  // both child boxes, written by k_refit_local
        const float4 c0 = c[0], c1 = c[1], cz = c[2];
        b.lo = f3(fminf(c0.x, c1.x), fminf(c0.z, c1.z), fminf(cz.x, cz.z));
        b.hi = f3(fmaxf(c0.y, c1.y), fmaxf(c0.w, c1.w), fmaxf(cz.y, cz.w));
    }
    bool stored = true;  // k_refit_local already wrote this box into its parent's slot
    while (true) {
        const uint32_t p = (link >> 1) & kLinkIndexMask, which = link & 1u;
        if (!stored) store_child_box(nodes + p, which, b);
        stored = false;
        __threadfence();
        if (atomicAdd(visit + p, 1u) == 0u) return;  // sibling not done yet
        const float4 *q = reinterpret_cast<const float4 *>(nodes + p);
        const float4 sxy = __ldcg(q + (which ^ 1u));
        const float2 sz = __ldcg(reinterpret_cast<const float2 *>(q + 2) + (which ^ 1u));
        b.lo = f3(fminf(b.lo.x, sxy.x), fminf(b.lo.y, sxy.z), fminf(b.lo.z, sz.x));
        b.hi = f3(fmaxf(b.hi.x, sxy.y), fmaxf(b.hi.y, sxy.w), fmaxf(b.hi.z, sz.y));
        link = parent_int[p];
        if (link == 0xFFFFFFFFu) {
            info->scene_lo[0] = b.lo.x; info->scene_lo[1] = b.lo.y; info->scene_lo[2] = b.lo.z;
            info->scene_hi[0] = b.hi.x; info->scene_hi[1] = b.hi.y; info->scene_hi[2] = b.hi.z;
            return;
        }
    }
}

// padded boxes of the outlier primitives (tested after the tree walk, culled by these boxes) and the
// "free box": the tree's bounds clipped, one outlier at a time, along the axis that keeps most of it,
// so that no outlier box reaches inside.  Segments inside the free box skip the outlier list.
__global__ void k_huge_boxes(const float4 *__restrict__ slot_geom, const float4 *__restrict__ tri_v12, BuildInfo *info) {
    const uint32_t k = threadIdx.x;
    const uint32_t nh = info->n_huge;
    if (k < nh) {
        const Box3 b = pad_box(prim_box(slot_geom[info->n_bvh + k], tri_v12));
        info->huge_lo[k][0] = b.lo.x; info->huge_lo[k][1] = b.lo.y; info->huge_lo[k][2] = b.lo.z;
        info->huge_hi[k][0] = b.hi.x; info->huge_hi[k][1] = b.hi.y; info->huge_hi[k][2] = b.hi.z;
    }
    __syncthreads();
    if (k != 0) return;
    float lo[3], hi[3];
    bool empty = info->n_bvh == 0;
    for (int a = 0; a < 3; a++) {
        lo[a] = info->scene_lo[a];
        hi[a] = info->scene_hi[a];
    }
    for (uint32_t h = 0; h < nh && !empty; h++) {
        const float *bl = info->huge_lo[h], *bh = info->huge_hi[h];
        bool overlap = true;
        for (int a = 0; a < 3; a++) overlap = overlap && bl[a] < hi[a] && bh[a] > lo[a];
        if (!overlap) continue;
        int best_axis = -1, best_side = 0;
        float best_keep = 0.0f;
        for (int a = 0; a < 3; a++) {
            const float ext = hi[a] - lo[a];
            if (!(ext > 0.0f)) continue;
            if (bl[a] > lo[a]) {  // keep [lo, box.lo]
                const float keep = (bl[a] - lo[a]) / ext;
                if (keep > best_keep) { best_keep = keep; best_axis = a; best_side = 0; }
            }
            if (bh[a] < hi[a]) {  // keep [box.hi, hi]
                const float keep = (hi[a] - bh[a]) / ext;
                if (keep > best_keep) { best_keep = keep; best_axis = a; best_side = 1; }
            }
        }
        if (best_axis < 0) empty = true;
        else if (best_side == 0) hi[best_axis] = bl[best_axis];
        else lo[best_axis] = bh[best_axis];
    }
    for (int a = 0; a < 3; a++) {
        // margin: dominates the fp32 rounding of the segment end point o + t*d
        const float m = 6.1035156e-5f * fmaxf(fmaxf(fabsf(lo[a]), fabsf(hi[a])), hi[a] - lo[a]) + 1e-30f;
        info->free_lo[a] = empty ? 1.0f : lo[a] + m;
        info->free_hi[a] = empty ? -1.0f : hi[a] - m;
    }
}

// n_bvh <= 1: the root is the only node
__global__ void k_small_tree(const float4 *__restrict__ slot_geom, const float4 *__restrict__ tri_v12, BuildInfo *info, Node *nodes) {
    if (threadIdx.x || blockIdx.x) return;
    const uint32_t n = info->n_bvh;
    if (n >= 2) return;
    const float inf = __int_as_float(0x7F800000);
    Box3 e;
    e.lo = f3(inf, inf, inf);
    e.hi = f3(-inf, -inf, -inf);
    Box3 b = e;
    const int invalid = (int)(kLeafFlag | 0u);
    int c0 = invalid;
    if (n == 1) {
        b = pad_box(prim_box(slot_geom[0], tri_v12));
        c0 = (int)(kLeafFlag | 0u);
        for (int a = 0; a < 3; a++) {
            info->scene_lo[a] = a == 0 ? b.lo.x : (a == 1 ? b.lo.y : b.lo.z);
            info->scene_hi[a] = a == 0 ? b.hi.x : (a == 1 ? b.hi.y : b.hi.z);
        }
    }
    store_child_box(nodes, 0, b);
    store_child_box(nodes, 1, e);
    nodes[0].n3 = make_float4(__int_as_float(c0), __int_as_float(invalid), 0.0f, 0.0f);
}

// ------------------------------------------------------------------ K8: collapse to the 8-wide compressed BVH
// Top-down, one thread per wide node, level-synchronous inside ONE cooperative kernel (grid.sync
// between levels).  The task queue is the wide-node array itself: wtask[w] = binary node that wide
// node w collapses.  A wide node starts from the two children of its binary node and keeps opening
// the internal child with the largest surface area until it has 8 children or only leaves.  Leaf
// primitives are re-packed so that every wide node's primitives are contiguous (prim_base).
struct CChild {
    int enc;  // binary child encoding: >= 0 internal node, < 0 leaf (first slot, count)
    float lo[3], hi[3];
};
__device__ __forceinline__ void read_children(const Node *__restrict__ nodes, int b, CChild &c0, CChild &c1) {
    const float4 *q = reinterpret_cast<const float4 *>(nodes + b);
    const float4 n0 = q[0], n1 = q[1], n2 = q[2], n3 = q[3];
    c0.enc = __float_as_int(n3.x);
    c0.lo[0] = n0.x; c0.hi[0] = n0.y; c0.lo[1] = n0.z; c0.hi[1] = n0.w; c0.lo[2] = n2.x; c0.hi[2] = n2.y;
    c1.enc = __float_as_int(n3.y);
    c1.lo[0] = n1.x; c1.hi[0] = n1.y; c1.lo[1] = n1.z; c1.hi[1] = n1.w; c1.lo[2] = n2.z; c1.hi[2] = n2.w;
}
__device__ __forceinline__ float child_area(const CChild &c) {
    if (c.enc < 0) return -1.0f;
    const float dx = c.hi[0] - c.lo[0], dy = c.hi[1] - c.lo[1], dz = c.hi[2] - c.lo[2];
    return dx * dy + dy * dz + dz * dx;
}
// smallest biased exponent e >= 1 with 255 * 2^(e-127) >= ext
__device__ __forceinline__ uint32_t quant_exp(float ext) {
    const uint32_t b = __float_as_uint(ext * (1.0f / 255.0f));
    uint32_t e = (b >> 23) & 0xFFu;
    if (b & 0x7FFFFFu) e += 1;
    return min(max(e, 1u), 254u);
}

// Wide-node and primitive slots of a whole block's wide nodes with ONE pair of atomics.  Same-address atomics are
// serialised at L2 at roughly 4 ns per request (ncu r4f: the allocating thread's wait for its atomicAdd result was
// 29 % of k_collapse8's stall samples, and 1.43 M nodes x 2 counters is where the one-thread-per-node kernel's
// 0.88 ms went as well), so the requests are summed over the block first.  BLOCK-COLLECTIVE: every thread calls it
// (n_int = n_pr = 0 for threads without a node); sums per block are < 2^16.
__device__ __forceinline__ void block_alloc(BuildInfo *info, uint32_t n_int, uint32_t n_pr, uint32_t &child_base, uint32_t &prim_base) {
    __shared__ uint32_t s_warp[4], s_base[2];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t v = n_int | (n_pr << 16);
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xFFFFFFFFu, inc, o);
        if ((int)lane >= o) inc += u;
    }
    if (lane == 31u) s_warp[wid] = inc;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t tot = 0;
        for (int k = 0; k < 4; k++) {
            const uint32_t x = s_warp[k];
            s_warp[k] = tot;
            tot += x;
        }
        s_base[0] = (tot & 0xFFFFu) ? atomicAdd(&info->n_wnodes, tot & 0xFFFFu) : 0u;
        s_base[1] = (tot >> 16) ? atomicAdd(&info->n_wprims, tot >> 16) : 0u;
    }
    __syncthreads();
    const uint32_t excl = s_warp[wid] + inc - v;
    child_base = s_base[0] + (excl & 0xFFFFu);
    prim_base = s_base[1] + (excl >> 16);
    __syncthreads();  // the scratch is reused by the next call
}

// BLOCK-COLLECTIVE (block_alloc): every thread of the block calls it, `active` = this thread has a wide node
__device__ void collapse_one(bool active, uint32_t w, int b, const Node *__restrict__ nodes, BuildInfo *info, uint32_t *__restrict__ wtask,
                             WNode *__restrict__ wnodes, const float4 *__restrict__ slot_geom,
                             const uint32_t *__restrict__ slot_prim, const uint32_t *__restrict__ slot_mat,
                             float4 *__restrict__ wslot_geom, uint32_t *__restrict__ wslot_prim, uint32_t *__restrict__ wslot_mat) {
    CChild ch[8];
    int cnt = 0;
    if (active) {
        CChild c0, c1;
        read_children(nodes, b, c0, c1);
        if (c0.lo[0] <= c0.hi[0]) ch[cnt++] = c0;  // the 1-primitive tree has an empty second child
        if (c1.lo[0] <= c1.hi[0]) ch[cnt++] = c1;
    }
    float area[8];  // half surface area of the internal children, -1 for leaves (never opened)
    for (int i = 0; i < cnt; i++) area[i] = child_area(ch[i]);
    while (cnt < 8) {
        int bi = -1;
        float ba = -1.0f;
        for (int i = 0; i < cnt; i++) {
            if (area[i] > ba) {
                ba = area[i];
                bi = i;
            }
        }
        if (bi < 0) break;
        CChild c0, c1;
        read_children(nodes, ch[bi].enc, c0, c1);
        ch[bi] = c0;
        area[bi] = child_area(c0);
        area[cnt] = child_area(c1);
        ch[cnt++] = c1;
    }
    // node box and per-axis quantisation grid
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) {
        lo[a] = ch[0].lo[a];
        hi[a] = ch[0].hi[a];
        for (int i = 1; i < cnt; i++) {
            lo[a] = fminf(lo[a], ch[i].lo[a]);
            hi[a] = fmaxf(hi[a], ch[i].hi[a]);
        }
    }
    // slots: octant of the child's centre relative to the node centre (bit0 = +x, bit1 = +y, bit2 = +z);
    // a taken slot falls back to the free slot with the fewest differing bits (lowest index on ties)
    uint32_t slot_of = 0;  // 4 bits per child
    uint32_t used = 0;
    for (int i = 0; i < cnt; i++) {
        uint32_t want = 0;
        for (int a = 0; a < 3; a++)
            if (ch[i].lo[a] + ch[i].hi[a] > lo[a] + hi[a]) want |= 1u << a;
        uint32_t sl = want;
        if ((used >> sl) & 1u) {
            // free slot at Hamming distance 1, else 2, else 3 from `want`, lowest index first
            const uint32_t fr = ~used & 0xFFu;
            const uint32_t m1 = (1u << (want ^ 1u)) | (1u << (want ^ 2u)) | (1u << (want ^ 4u));
            const uint32_t m2 = (1u << (want ^ 3u)) | (1u << (want ^ 5u)) | (1u << (want ^ 6u));
            const uint32_t c = (fr & m1) ? (fr & m1) : ((fr & m2) ? (fr & m2) : fr);
            sl = (uint32_t)__ffs(c) - 1u;
        }
        used |= 1u << sl;
        slot_of |= sl << (4 * i);
    }
    // every leaf of the binary tree holds exactly one primitive under the wide tree (build_accel)
    uint32_t imask = 0, leafmask = 0;
    for (int i = 0; i < cnt; i++) {
        const uint32_t k = (slot_of >> (4 * i)) & 7u;
        if (ch[i].enc >= 0) imask |= 1u << k; else leafmask |= 1u << k;
    }
    // The grid origin sits one step below the node box and every child interval is widened by one
    // step per side: the traversal's folded FMA may move a plane by half a step (rt_device.cuh qfloat).
    // Quantised planes go straight into 64-bit words (byte k = slot k); empty slots keep an inverted
    // box (lo = 255, hi = 0) and are never hit.
    unsigned long long empty_lo = 0;
    for (uint32_t k = 0; k < 8; k++)
        if (!((used >> k) & 1u)) empty_lo |= 255ull << (8 * k);
    uint32_t eb[3];
    unsigned long long qlo64[3], qhi64[3];
    float org[3];
    for (int a = 0; a < 3; a++) {
        eb[a] = quant_exp((hi[a] - lo[a]) * (255.0f / 252.0f));
        for (;;) {
            const float sc = __uint_as_float(eb[a] << 23);
            // 1 / sc is a power of two as well (2^-127 for eb = 254 is a subnormal, still exact), so the
            // products below are the correctly rounded quotients: identical to dividing, without the
            // ~15-instruction IEEE division sequence 48 times per node
            const float isc = __uint_as_float(eb[a] >= 254u ? 0x00400000u : ((254u - eb[a]) << 23));
            org[a] = lo[a] - sc;
            bool ok = true;
            unsigned long long wl = empty_lo, wh = 0;
            for (int i = 0; i < cnt; i++) {
                // conservative: decoded lo (rounded up) <= child lo, decoded hi (rounded down) >= child hi
                float ql = fmaxf(floorf((ch[i].lo[a] - org[a]) * isc), 0.0f);
                if (ql > 0.0f && __fmaf_ru(ql, sc, org[a]) > ch[i].lo[a]) ql -= 1.0f;
                float qh = fmaxf(ceilf((ch[i].hi[a] - org[a]) * isc), 0.0f);
                if (__fmaf_rd(qh, sc, org[a]) < ch[i].hi[a]) qh += 1.0f;
                ql = fmaxf(ql - 1.0f, 0.0f);
                qh += 1.0f;
                if (qh > 255.0f || ql > 255.0f) {
                    if (eb[a] < 254u) {
                        ok = false;
                        break;
                    }
                    ql = 0.0f;  // largest grid and still out of range (non-finite extent): the whole axis, conservative
                    qh = 255.0f;
                }
                const uint32_t sh = 8u * ((slot_of >> (4 * i)) & 7u);
                wl |= (unsigned long long)(uint32_t)ql << sh;
                wh |= (unsigned long long)(uint32_t)qh << sh;
            }
            qlo64[a] = wl;
            qhi64[a] = wh;
            if (ok || eb[a] >= 254u) break;
            eb[a] += 1;
        }
    }
    const uint32_t n_int = __popc(imask), n_pr = __popc(leafmask);
    uint32_t child_base, prim_base;
    block_alloc(info, n_int, n_pr, child_base, prim_base);
    if (!active) return;
    // internal children and leaf primitives are stored in slot order
    for (int i = 0; i < cnt; i++) {
        const uint32_t k = (slot_of >> (4 * i)) & 7u;
        const uint32_t below = (1u << k) - 1u;
        if (ch[i].enc >= 0) {
            wtask[child_base + __popc(imask & below)] = (uint32_t)ch[i].enc;
        } else {
            const uint32_t f = (uint32_t)ch[i].enc & kLeafFirstMask;
            const uint32_t d = prim_base + __popc(leafmask & below);
            wslot_geom[d] = slot_geom[f];
            wslot_prim[d] = slot_prim[f];
            wslot_mat[d] = slot_mat[f];
        }
    }
    const uint32_t px[2] = {(uint32_t)qlo64[0], (uint32_t)(qlo64[0] >> 32)}, py[2] = {(uint32_t)qlo64[1], (uint32_t)(qlo64[1] >> 32)};
    const uint32_t pz[2] = {(uint32_t)qlo64[2], (uint32_t)(qlo64[2] >> 32)}, hx[2] = {(uint32_t)qhi64[0], (uint32_t)(qhi64[0] >> 32)};
    const uint32_t hy[2] = {(uint32_t)qhi64[1], (uint32_t)(qhi64[1] >> 32)}, hz[2] = {(uint32_t)qhi64[2], (uint32_t)(qhi64[2] >> 32)};
    float4 *o = reinterpret_cast<float4 *>(wnodes + w);
    o[0] = make_float4(org[0], org[1], org[2], __uint_as_float(eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24)));
    o[1] = make_float4(__uint_as_float(child_base), __uint_as_float(prim_base), __uint_as_float(leafmask), 0.0f);
    o[2] = make_float4(__uint_as_float(px[0]), __uint_as_float(px[1]), __uint_as_float(py[0]), __uint_as_float(py[1]));
    o[3] = make_float4(__uint_as_float(pz[0]), __uint_as_float(pz[1]), __uint_as_float(hx[0]), __uint_as_float(hx[1]));
    o[4] = make_float4(__uint_as_float(hy[0]), __uint_as_float(hy[1]), __uint_as_float(hz[0]), __uint_as_float(hz[1]));
}

// Grid-wide barrier between the levels of k_collapse (the kernel is launched cooperatively, so all blocks
// are co-resident).  One barrier per level: the LAST block to arrive publishes the end of the next level
// (BuildInfo.level_end = n_wnodes) before it releases the others, so nobody can observe allocations made
// by blocks that already run ahead into the next level.  bar[0] counts arrivals (monotonic), bar[1] is
// the generation; both are zeroed by k_decide at the start of every build.
__device__ __forceinline__ uint32_t collapse_barrier(BuildInfo *info, uint32_t gen) {
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile uint32_t *vgen = &info->pad2_[1];
        __threadfence();
        const uint32_t arrived = atomicAdd(&info->pad2_[0], 1u) + 1u;
        if (arrived == gridDim.x * (gen + 1u)) {
            info->level_end = *reinterpret_cast<volatile uint32_t *>(&info->n_wnodes);
            __threadfence();
            *vgen = gen + 1u;
        } else {
            while (*vgen <= gen) __nanosleep(64);
        }
        __threadfence();
    }
    __syncthreads();
    return *reinterpret_cast<volatile uint32_t *>(&info->level_end);
}

__global__ void __launch_bounds__(128, RT_COLLAPSE_MIN_BLOCKS)
k_collapse(const Node *__restrict__ nodes, BuildInfo *info, uint32_t *__restrict__ wtask, WNode *__restrict__ wnodes,
           const float4 *__restrict__ slot_geom, const uint32_t *__restrict__ slot_prim, const uint32_t *__restrict__ slot_mat,
           float4 *__restrict__ wslot_geom, uint32_t *__restrict__ wslot_prim, uint32_t *__restrict__ wslot_mat) {
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    const uint32_t n = info->n_bvh;
    // outlier primitives keep their positions behind the tree's primitives
    for (uint32_t i = gtid; i < info->n_huge; i += gsize) {
        wslot_geom[n + i] = slot_geom[n + i];
        wslot_prim[n + i] = slot_prim[n + i];
        wslot_mat[n + i] = slot_mat[n + i];
    }
    if (n == 0) {  // uniform over the grid
        if (gtid == 0) info->n_wnodes = 0;
        return;
    }
    // k_decide left n_wnodes = 1 (the root's slot: wide node 0 collapses binary node 0).  Task entries are
    // read with __ldcg: a 128-byte line of wtask can sit in L1 from an earlier level, when the entries of
    // this level were not written yet.
    uint32_t begin = 0, end = 1, gen = 0;
    while (begin < end) {
        for (uint32_t w0 = begin + blockIdx.x * blockDim.x; w0 < end; w0 += gsize) {  // block-uniform trip count
            const uint32_t w = w0 + threadIdx.x;
            const bool active = w < end;
            collapse_one(active, w, (active && w != 0u) ? (int)__ldcg(wtask + w) : 0, nodes, info, wtask, wnodes, slot_geom, slot_prim, slot_mat,
                         wslot_geom, wslot_prim, wslot_mat);
        }
        begin = end;
        end = collapse_barrier(info, gen++);
    }
}

// ------------------------------------------------------------------ K8': group-cooperative collapse (option "collapse_mode" = 1)
// The same job as k_collapse with EIGHT LANES PER WIDE NODE (four wide nodes per warp): lane k of a group holds one
// child (binary encoding + box) in registers.  Per round every internal child is opened at once — its lane reads the
// binary node, keeps child 0 and hands child 1 to the next free lane — so a balanced subtree reaches 8 children in
// 3 dependent reads instead of 6; when fewer slots than internal children are left, the largest (half surface area,
// lane index on ties) open first.  Node box, slot assignment, quantisation and the plane words are group reductions
// over shuffles; the up-to-8 primitive copies of a bottom-level node are one parallel step.  No per-thread child
// array (the one-thread-per-node version keeps 8 x 28 bytes in local memory and walks it with dynamic indices).
// All four groups of a warp run in lockstep (shuffles are warp-wide), so every loop below is warp-uniform.
// Measured (10 M spheres, r4e-r4g): bit-identical images and the same traversal time as k_collapse's trees, but
// 1.2 ms against 0.88 ms: 3.5 x the warp instructions (shuffle reductions, unrolled 8-step loops) at the same ~20 %
// issue utilisation; before block_alloc it was 3.3 ms, bound by the serialised same-address atomics.  Kept as an
// option; the one-thread-per-node kernel stays the default.
__device__ __forceinline__ float gsh(float v, int src) { return __shfl_sync(0xFFFFFFFFu, v, src, 8); }
__device__ __forceinline__ int gshi(int v, int src) { return __shfl_sync(0xFFFFFFFFu, v, src, 8); }
__device__ __forceinline__ uint32_t gshu(uint32_t v, int src) { return __shfl_sync(0xFFFFFFFFu, v, src, 8); }
__device__ __forceinline__ float gmin8(float v) {
    v = fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, 1));
    v = fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, 2));
    return fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, 4));
}
__device__ __forceinline__ float gmax8(float v) {
    v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, 1));
    v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, 2));
    return fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, 4));
}
__device__ __forceinline__ uint32_t gor8(uint32_t v) {
    v |= __shfl_xor_sync(0xFFFFFFFFu, v, 1);
    v |= __shfl_xor_sync(0xFFFFFFFFu, v, 2);
    return v | __shfl_xor_sync(0xFFFFFFFFu, v, 4);
}

__device__ __forceinline__ void collapse_group(bool active, uint32_t w, int b, const Node *__restrict__ nodes, BuildInfo *info,
                                               uint32_t *__restrict__ wtask, WNode *__restrict__ wnodes,
                                               const float4 *__restrict__ slot_geom, const uint32_t *__restrict__ slot_prim,
                                               const uint32_t *__restrict__ slot_mat, float4 *__restrict__ wslot_geom,
                                               uint32_t *__restrict__ wslot_prim, uint32_t *__restrict__ wslot_mat) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u, gl = lane & 7u, gbit = lane & 24u;
    int enc = 0;
    float lo0 = 0, lo1 = 0, lo2 = 0, hi0 = 0, hi1 = 0, hi2 = 0;
    bool valid = false;
    if (active && gl < 2u) {
        const float4 *q = reinterpret_cast<const float4 *>(nodes + b);
        const float4 n0 = q[0], n1 = q[1], n2 = q[2], n3 = q[3];
        if (gl == 0u) {
            enc = __float_as_int(n3.x);
            lo0 = n0.x; hi0 = n0.y; lo1 = n0.z; hi1 = n0.w; lo2 = n2.x; hi2 = n2.y;
        } else {
            enc = __float_as_int(n3.y);
            lo0 = n1.x; hi0 = n1.y; lo1 = n1.z; hi1 = n1.w; lo2 = n2.z; hi2 = n2.w;
        }
        valid = lo0 <= hi0;  // the 1-primitive tree has an empty second child
    }
    int cnt = __popc((__ballot_sync(FULL, valid) >> gbit) & 0xFFu);
    // ---- open internal children, all at once per round
    for (;;) {
        const bool internal = valid && enc >= 0;
        const uint32_t im = (__ballot_sync(FULL, internal) >> gbit) & 0xFFu;
        const int room = 8 - cnt;
        const bool more = im != 0u && room > 0;
        if (!__any_sync(FULL, more)) break;
        const float dx = hi0 - lo0, dy = hi1 - lo1, dz = hi2 - lo2;
        const float area = internal ? dx * dy + dy * dz + dz * dx : -1.0f;
        int rank = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const float aj = gsh(area, j);
            rank += (aj > area || (aj == area && j < (int)gl)) ? 1 : 0;
        }
        const bool open = more && internal && rank < room;
        const uint32_t om = (__ballot_sync(FULL, open) >> gbit) & 0xFFu;
        int e1 = 0;
        float a0 = 0, a1 = 0, a2 = 0, b0 = 0, b1 = 0, b2 = 0;  // child 1 of the node this lane opens
        if (open) {
            const float4 *q = reinterpret_cast<const float4 *>(nodes + enc);
            const float4 n0 = q[0], n1 = q[1], n2 = q[2], n3 = q[3];
            enc = __float_as_int(n3.x);
            lo0 = n0.x; hi0 = n0.y; lo1 = n0.z; hi1 = n0.w; lo2 = n2.x; hi2 = n2.y;
            e1 = __float_as_int(n3.y);
            a0 = n1.x; b0 = n1.y; a1 = n1.z; b1 = n1.w; a2 = n2.z; b2 = n2.w;
        }
        // lane cnt + k takes child 1 of the k-th opening lane
        const int k = (int)gl - cnt, n_open = __popc(om);
        int src = 0, c = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if ((om >> j) & 1u) {
                if (c == k) src = j;
                c++;
            }
        }
        const int r_e = gshi(e1, src);
        const float r_a0 = gsh(a0, src), r_a1 = gsh(a1, src), r_a2 = gsh(a2, src);
        const float r_b0 = gsh(b0, src), r_b1 = gsh(b1, src), r_b2 = gsh(b2, src);
        if (k >= 0 && k < n_open) {
            enc = r_e;
            lo0 = r_a0; lo1 = r_a1; lo2 = r_a2; hi0 = r_b0; hi1 = r_b1; hi2 = r_b2;
            valid = true;
        }
        cnt += n_open;
    }
    // ---- node box
    const float inf = __int_as_float(0x7F800000);
    const float nlo0 = gmin8(valid ? lo0 : inf), nlo1 = gmin8(valid ? lo1 : inf), nlo2 = gmin8(valid ? lo2 : inf);
    const float nhi0 = gmax8(valid ? hi0 : -inf), nhi1 = gmax8(valid ? hi1 : -inf), nhi2 = gmax8(valid ? hi2 : -inf);
    // ---- slots: octant of the child's centre relative to the node centre; a taken slot falls back to the free slot
    // with the fewest differing bits (lowest index on ties); children pick in lane order
    uint32_t want = 0;
    if (lo0 + hi0 > nlo0 + nhi0) want |= 1u;
    if (lo1 + hi1 > nlo1 + nhi1) want |= 2u;
    if (lo2 + hi2 > nlo2 + nhi2) want |= 4u;
    uint32_t used = 0, my_slot = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t wi = gshu(want, i);
        if (i < cnt) {
            uint32_t sl = wi;
            if ((used >> sl) & 1u) {
                const uint32_t fr = ~used & 0xFFu;
                const uint32_t m1 = (1u << (wi ^ 1u)) | (1u << (wi ^ 2u)) | (1u << (wi ^ 4u));
                const uint32_t m2 = (1u << (wi ^ 3u)) | (1u << (wi ^ 5u)) | (1u << (wi ^ 6u));
                const uint32_t cc = (fr & m1) ? (fr & m1) : ((fr & m2) ? (fr & m2) : fr);
                sl = (uint32_t)__ffs(cc) - 1u;
            }
            used |= 1u << sl;
            if ((int)gl == i) my_slot = sl;
        }
    }
    const uint32_t imask = gor8((valid && enc >= 0) ? (1u << my_slot) : 0u);
    const uint32_t leafmask = gor8((valid && enc < 0) ? (1u << my_slot) : 0u);
    // ---- quantisation grid per axis (same rules as collapse_one) and the plane bytes of this lane's child
    uint32_t empty_lo_a = 0, empty_lo_b = 0;  // slots 0..3 / 4..7
#pragma unroll
    for (uint32_t kk = 0; kk < 4; kk++) {
        if (!((used >> kk) & 1u)) empty_lo_a |= 255u << (8 * kk);
        if (!((used >> (kk + 4)) & 1u)) empty_lo_b |= 255u << (8 * kk);
    }
    const float nlo[3] = {nlo0, nlo1, nlo2}, nhi[3] = {nhi0, nhi1, nhi2};
    const float clo[3] = {lo0, lo1, lo2}, chi[3] = {hi0, hi1, hi2};
    uint32_t eb[3], wla[3], wlb[3], wha[3], whb[3];
    float org[3];
    const uint32_t sh = 8u * (my_slot & 3u);
    const bool upper = my_slot >= 4u;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        eb[a] = quant_exp((nhi[a] - nlo[a]) * (255.0f / 252.0f));
        float ql = 0.0f, qh = 0.0f;
        for (;;) {
            const float sc = __uint_as_float(eb[a] << 23);
            const float isc = __uint_as_float(eb[a] >= 254u ? 0x00400000u : ((254u - eb[a]) << 23));
            org[a] = nlo[a] - sc;
            bool ok = true;
            if (valid) {
                ql = fmaxf(floorf((clo[a] - org[a]) * isc), 0.0f);
                if (ql > 0.0f && __fmaf_ru(ql, sc, org[a]) > clo[a]) ql -= 1.0f;
                qh = fmaxf(ceilf((chi[a] - org[a]) * isc), 0.0f);
                if (__fmaf_rd(qh, sc, org[a]) < chi[a]) qh += 1.0f;
                ql = fmaxf(ql - 1.0f, 0.0f);
                qh += 1.0f;
                if (qh > 255.0f || ql > 255.0f) {
                    if (eb[a] < 254u) ok = false;
                    else { ql = 0.0f; qh = 255.0f; }  // largest grid and still out of range (non-finite extent)
                }
            }
            // the grid is the node's: every lane of the group takes the same decision
            const uint32_t bad = (__ballot_sync(FULL, !ok) >> gbit) & 0xFFu;
            if (bad) eb[a] += 1;
            if (!__any_sync(FULL, bad != 0u)) break;
            // groups that were fine recompute the same values next round (eb unchanged)
        }
        const uint32_t bl = valid ? ((uint32_t)ql << sh) : 0u, bh = valid ? ((uint32_t)qh << sh) : 0u;
        wla[a] = gor8(upper ? 0u : bl) | empty_lo_a;
        wlb[a] = gor8(upper ? bl : 0u) | empty_lo_b;
        wha[a] = gor8(upper ? 0u : bh);
        whb[a] = gor8(upper ? bh : 0u);
    }
    // ---- allocation (one lane per group), children tasks and primitive re-pack (one lane per child)
    const uint32_t n_int = __popc(imask), n_pr = __popc(leafmask);
    uint32_t child_base = 0, prim_base = 0;
    {
        const bool mine = active && gl == 0u;
        block_alloc(info, mine ? n_int : 0u, mine ? n_pr : 0u, child_base, prim_base);
    }
    child_base = gshu(child_base, 0);
    prim_base = gshu(prim_base, 0);
    if (active && valid) {
        const uint32_t below = (1u << my_slot) - 1u;
        if (enc >= 0) {
            wtask[child_base + __popc(imask & below)] = (uint32_t)enc;
        } else {
            const uint32_t f = (uint32_t)enc & kLeafFirstMask;
            const uint32_t d = prim_base + __popc(leafmask & below);
            wslot_geom[d] = slot_geom[f];
            wslot_prim[d] = slot_prim[f];
            wslot_mat[d] = slot_mat[f];
        }
    }
    if (active && gl < 5u) {
        float4 v;
        if (gl == 0u) v = make_float4(org[0], org[1], org[2], __uint_as_float(eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24)));
        else if (gl == 1u) v = make_float4(__uint_as_float(child_base), __uint_as_float(prim_base), __uint_as_float(leafmask), 0.0f);
        else if (gl == 2u) v = make_float4(__uint_as_float(wla[0]), __uint_as_float(wlb[0]), __uint_as_float(wla[1]), __uint_as_float(wlb[1]));
        else if (gl == 3u) v = make_float4(__uint_as_float(wla[2]), __uint_as_float(wlb[2]), __uint_as_float(wha[0]), __uint_as_float(whb[0]));
        else v = make_float4(__uint_as_float(wha[1]), __uint_as_float(whb[1]), __uint_as_float(wha[2]), __uint_as_float(whb[2]));
        reinterpret_cast<float4 *>(wnodes + w)[gl] = v;
    }
}

#ifndef RT_COLLAPSE8_MIN_BLOCKS
#define RT_COLLAPSE8_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(128, RT_COLLAPSE8_MIN_BLOCKS)
k_collapse8(const Node *__restrict__ nodes, BuildInfo *info, uint32_t *__restrict__ wtask, WNode *__restrict__ wnodes,
            const float4 *__restrict__ slot_geom, const uint32_t *__restrict__ slot_prim, const uint32_t *__restrict__ slot_mat,
            float4 *__restrict__ wslot_geom, uint32_t *__restrict__ wslot_prim, uint32_t *__restrict__ wslot_mat) {
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    const uint32_t n = info->n_bvh;
    for (uint32_t i = gtid; i < info->n_huge; i += gsize) {  // outlier primitives keep their positions behind the tree's
        wslot_geom[n + i] = slot_geom[n + i];
        wslot_prim[n + i] = slot_prim[n + i];
        wslot_mat[n + i] = slot_mat[n + i];
    }
    if (n == 0) {  // uniform over the grid
        if (gtid == 0) info->n_wnodes = 0;
        return;
    }
    const uint32_t group = gtid >> 3, n_groups = gsize >> 3;
    uint32_t begin = 0, end = 1, gen = 0;
    while (begin < end) {
        // block-uniform trip count (block_alloc): the 16 groups of a block hold consecutive tasks
        for (uint32_t w0 = begin + (group & ~15u); w0 < end; w0 += n_groups) {
            const uint32_t w = w0 + (group & 15u);
            const bool active = w < end;
            const int b = (active && w != 0u) ? (int)__ldcg(wtask + w) : 0;
            collapse_group(active, w, b, nodes, info, wtask, wnodes, slot_geom, slot_prim, slot_mat, wslot_geom, wslot_prim, wslot_mat);
        }
        begin = end;
        end = collapse_barrier(info, gen++);
    }
}

// ------------------------------------------------------------------ per-frame particle update (Mesh-mode translation)
// src/Global/RendererMesh.cu:378-391 + DeviceFunctions.cuh:133-148 with zero rotation, one thread per
// coordinate so that the float3 arrays are read and written fully coalesced (12 B per particle each way).
__device__ __forceinline__ float advect1(float c, float v, float duration, float frames, float frame, float sh, float sc) {
    const float total = v * duration;            // RendererMesh.cu:383
    const float per_frame = total / frames;      // :386
    const float shift = sh + per_frame * frame;  // :388
    return sc * c + shift;                       // row of the 3x4 matrix applied to the centre
}
// 4 particles (12 coordinates) per thread: three 128-bit loads per input array, three 128-bit stores
__global__ void __launch_bounds__(256)
k_spheres_advect(const float *__restrict__ base_c, const float *__restrict__ base_r, const float *__restrict__ vel, size_t n,
                 float duration, float frames, float frame, F3 pshift, F3 pscale, float *__restrict__ out_c,
                 float *__restrict__ out_r) {
    const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // quad of particles
    const size_t p0 = 4 * q;
    if (p0 >= n) return;
    const float rs = sqrtf((pscale.x * pscale.x + 0.0f) + 0.0f);  // k_flatten's radius scale
    const float sh[3] = {pshift.x, pshift.y, pshift.z}, sc[3] = {pscale.x, pscale.y, pscale.z};
    if (p0 + 4 <= n) {
        const float4 *c4 = reinterpret_cast<const float4 *>(base_c) + 3 * q;
        const float4 *v4 = reinterpret_cast<const float4 *>(vel) + 3 * q;
        float4 *o4 = reinterpret_cast<float4 *>(out_c) + 3 * q;
        float c[12], v[12];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const float4 a = c4[j], b = v4[j];
            c[4 * j] = a.x; c[4 * j + 1] = a.y; c[4 * j + 2] = a.z; c[4 * j + 3] = a.w;
            v[4 * j] = b.x; v[4 * j + 1] = b.y; v[4 * j + 2] = b.z; v[4 * j + 3] = b.w;
        }
#pragma unroll
        for (int j = 0; j < 12; j++) c[j] = advect1(c[j], v[j], duration, frames, frame, sh[j % 3], sc[j % 3]);
#pragma unroll
        for (int j = 0; j < 3; j++) o4[j] = make_float4(c[4 * j], c[4 * j + 1], c[4 * j + 2], c[4 * j + 3]);
        const float4 r = reinterpret_cast<const float4 *>(base_r)[q];
        reinterpret_cast<float4 *>(out_r)[q] = make_float4(r.x * rs, r.y * rs, r.z * rs, r.w * rs);
    } else {
        for (size_t p = p0; p < n; p++) {
            for (int a = 0; a < 3; a++)
                out_c[3 * p + a] = advect1(base_c[3 * p + a], vel[3 * p + a], duration, frames, frame, sh[a], sc[a]);
            out_r[p] = base_r[p] * rs;
        }
    }
}

int spheres_advect(Context *ctx, const float *base_c, const float *base_r, const float *vel, size_t n, float duration,
                   size_t frames, size_t frame, const float *pshift, const float *pscale, float *out_c, float *out_r,
                   cudaStream_t st) {
    if (!n) return RT_OK;
    const size_t blocks = ((n + 3) / 4 + 255) / 256;
    if (blocks > 0x7FFFFFFFull) return set_error(ctx, RT_ERR_INVALID, "too many particles");
    if ((((uintptr_t)base_c | (uintptr_t)base_r | (uintptr_t)vel | (uintptr_t)out_c | (uintptr_t)out_r) & 15u) != 0)
        return set_error(ctx, RT_ERR_INVALID, "rt_spheres_advect: arrays must be 16-byte aligned");
    k_spheres_advect<<<(unsigned)blocks, 256, 0, st>>>(base_c, base_r, vel, n, duration, (float)frames, (float)frame,
                                                      F3{pshift[0], pshift[1], pshift[2]}, F3{pscale[0], pscale[1], pscale[2]},
                                                      out_c, out_r);
    RT_CUDA(ctx, cudaGetLastError());
    return RT_OK;
}

// ------------------------------------------------------------------ per-frame particle transforms (Time mode)
// src/Global/RendererTime.cu:283-293 (position + velocity * time, slerp of the two key-frame quaternions, :297-340,
// quatToEuler :343-370) followed by constructTransformMatrix (DeviceFunctions.cuh:133-148: T * (Rx*Ry*Rz) * S with the
// 4x4 products accumulated from 0 in ascending n).  Same operation order as the host chain in rt_host.cpp; the
// translation column is bit-identical to it, the rotation entries differ by the CUDA libm's sinf / cosf / acosf /
// asinf / atan2f against glibc's (<= 2 ulp each: <= 4e-6 absolute on a matrix entry, tests/test_gpu_build.py).
struct M4 {
    float v[4][4];
};
__device__ __forceinline__ M4 m4_identity() {
    M4 m;
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) m.v[i][j] = (i == j) ? 1.0f : 0.0f;
    return m;
}
__device__ __forceinline__ M4 m4_mul(const M4 &a, const M4 &b) {
    M4 r;
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float acc = 0.0f;
#pragma unroll
            for (int n = 0; n < 4; n++) acc += a.v[i][n] * b.v[n][j];
            r.v[i][j] = acc;
        }
    return r;
}
template <int kAxis>
__device__ __forceinline__ M4 m4_rotation(float degree) {
    constexpr float kPiRef = 3.1415926;  // DeviceFunctions.cuh:20
    const float theta = degree * kPiRef / 180.0f;
    const float c = cosf(theta), s = sinf(theta);
    M4 m = m4_identity();
    constexpr int a = (kAxis + 1) % 3, b = (kAxis + 2) % 3;
    m.v[a][a] = c;
    m.v[a][b] = -s;
    m.v[b][a] = s;
    m.v[b][b] = c;
    return m;
}
__global__ void __launch_bounds__(128)
k_transforms_time(const float *__restrict__ pos, const float *__restrict__ vel, const float4 *__restrict__ quat_cur,
                  const float4 *__restrict__ quat_next, size_t n, float duration, float frames, float frame, float factor,
                  F3 pshift, F3 pscale, float4 *__restrict__ out) {
    constexpr float kPiRef = 3.1415926;
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float sh[3] = {pshift.x, pshift.y, pshift.z};
    float shift[3];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        const float total = vel[3 * i + a] * duration;
        const float per_frame = total / frames;
        const float here = pos[3 * i + a] + per_frame * frame;
        shift[a] = sh[a] + here;
    }
    // slerp with the reference's positional return: q = {x, y, z, w} in, {w-, x-, y-, z-expression} out
    const float4 qa = quat_cur[i];
    float4 qb = quat_next[i];
    float dp = qa.w * qb.w + qa.x * qb.x + qa.y * qb.y + qa.z * qb.z;
    if (dp < 0.0f) {
        qb = make_float4(-qb.x, -qb.y, -qb.z, -qb.w);
        dp = -dp;
    }
    float q0, q1, q2, q3;
    if (dp > 0.9995f) {
        float rx = qa.w + factor * (qb.w - qa.w);
        float ry = qa.x + factor * (qb.x - qa.x);
        float rz = qa.y + factor * (qb.y - qa.y);
        float rw = qa.z + factor * (qb.z - qa.z);
        const float mag = sqrtf(rw * rw + rx * rx + ry * ry + rz * rz);
        if (mag > 0.0f) {
            rw /= mag; rx /= mag; ry /= mag; rz /= mag;
        }
        q0 = rx; q1 = ry; q2 = rz; q3 = rw;
    } else {
        const float theta0 = acosf(dp);
        const float theta = theta0 * factor;
        const float st = sinf(theta), st0 = sinf(theta0);
        const float k0 = cosf(theta) - dp * st / st0;
        const float k1 = st / st0;
        q0 = (k0 * qa.w) + (k1 * qb.w);
        q1 = (k0 * qa.x) + (k1 * qb.x);
        q2 = (k0 * qa.y) + (k1 * qb.y);
        q3 = (k0 * qa.z) + (k1 * qb.z);
    }
    // quatToEuler on {x, y, z, w} = {q0, q1, q2, q3}, degrees
    const float roll = atan2f(2 * (q3 * q0 + q1 * q2), 1 - 2 * (q0 * q0 + q1 * q1));
    const float sp = 2 * (q3 * q1 - q2 * q0);
    const float pitch = (fabsf(sp) >= 1) ? copysignf(kPiRef / 2, sp) : asinf(sp);
    const float yaw = atan2f(2 * (q3 * q2 + q0 * q1), 1 - 2 * (q1 * q1 + q2 * q2));
    const float ex = roll * 180.0f / kPiRef, ey = pitch * 180.0f / kPiRef, ez = yaw * 180.0f / kPiRef;

    M4 T = m4_identity();
    T.v[0][3] = shift[0]; T.v[1][3] = shift[1]; T.v[2][3] = shift[2];
    const M4 R = m4_mul(m4_mul(m4_rotation<0>(ex), m4_rotation<1>(ey)), m4_rotation<2>(ez));
    M4 S = m4_identity();
    S.v[0][0] = pscale.x; S.v[1][1] = pscale.y; S.v[2][2] = pscale.z;
    const M4 M = m4_mul(m4_mul(T, R), S);
#pragma unroll
    for (int r = 0; r < 3; r++) out[3 * i + r] = make_float4(M.v[r][0], M.v[r][1], M.v[r][2], M.v[r][3]);
}

int transforms_time(Context *ctx, const float *pos, const float *vel, const float *quat_cur, const float *quat_next, size_t n,
                    float duration, size_t frames, size_t frame, const float *pshift, const float *pscale, float *out12,
                    cudaStream_t st) {
    if (!n) return RT_OK;
    const size_t blocks = (n + 127) / 128;
    if (blocks > 0x7FFFFFFFull) return set_error(ctx, RT_ERR_INVALID, "too many particles");
    if ((((uintptr_t)quat_cur | (uintptr_t)quat_next | (uintptr_t)out12) & 15u) != 0)
        return set_error(ctx, RT_ERR_INVALID, "rt_transforms_time: quaternion and output arrays must be 16-byte aligned");
    float factor = 1.0f;  // RendererTime.cu:283-286
    if (frames > 1) factor = (float)frame / (float)(frames - 1);
    k_transforms_time<<<(unsigned)blocks, 128, 0, st>>>(pos, vel, reinterpret_cast<const float4 *>(quat_cur),
                                                       reinterpret_cast<const float4 *>(quat_next), n, duration, (float)frames,
                                                       (float)frame, factor, F3{pshift[0], pshift[1], pshift[2]},
                                                       F3{pscale[0], pscale[1], pscale[2]}, reinterpret_cast<float4 *>(out12));
    RT_CUDA(ctx, cudaGetLastError());
    return RT_OK;
}

// Device-resident instance transforms (rt_scene_set_transforms_device): written over the uploaded instance table at
// the start of every build, on the build's stream — no host round trip.
__global__ void __launch_bounds__(256)
k_apply_transforms(InstDev *__restrict__ insts, const float *__restrict__ tf, uint32_t first, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float m[12];
#pragma unroll
    for (int k = 0; k < 12; k++) m[k] = tf[12 * (size_t)i + k];
    InstDev *in = insts + first + i;
#pragma unroll
    for (int k = 0; k < 12; k++) in->m[k] = m[k];
    const bool ident = m[0] == 1.0f && m[1] == 0.0f && m[2] == 0.0f && m[4] == 0.0f && m[5] == 1.0f && m[6] == 0.0f &&
                       m[8] == 0.0f && m[9] == 0.0f && m[10] == 1.0f;
    in->identity = ident ? 1u : 0u;
}

// ------------------------------------------------------------------ host driver
static inline uint32_t div_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }

int build_accel(Scene *s, cudaStream_t st) {
    Context *ctx = s->ctx;
    // ---- tables
    uint64_t np = 0, nt = 0;
    std::vector<uint32_t> base(s->insts.size() + 1);
    for (size_t i = 0; i < s->insts.size(); i++) {
        InstDev &in = s->insts[i];
        const GroupDev &g = s->groups[in.group];
        in.prim_base = (uint32_t)np;
        in.tri_base = (uint32_t)nt;
        base[i] = (uint32_t)np;
        np += g.count;
        if (g.kind == 1) nt += g.count;
        if (g.kind == 0 && !in.identity) {
            // spheres only support similarity transforms (uniform scale)
            const float *m = in.m;
            const float c0 = m[0] * m[0] + m[4] * m[4] + m[8] * m[8];
            const float c1 = m[1] * m[1] + m[5] * m[5] + m[9] * m[9];
            const float c2 = m[2] * m[2] + m[6] * m[6] + m[10] * m[10];
            const float mx = fmaxf(c0, fmaxf(c1, c2)), mn = fminf(c0, fminf(c1, c2));
            if (!(mx > 0.0f) || (mx - mn) > 1e-4f * mx)
                return set_error(ctx, RT_ERR_UNSUPPORTED, "instance %zu: non-uniform scale on a sphere group", i);
        }
    }
    base[s->insts.size()] = (uint32_t)np;
    for (size_t i = 0; i < s->insts.size(); i++) {
        const InstDev &in = s->insts[i];
        const GroupDev &g = s->groups[in.group];
        if (g.count) {
            // material of primitive k = (type, index + k * stride): every index must lie inside the table of its
            // type (scatter() reads roughs[] / metals[] unchecked) and must not carry into the type bit
            const uint64_t mi = in.mat & 0x7FFFFFFFu;
            const uint64_t last = mi + (uint64_t)(g.count - 1) * in.mat_stride;
            const uint64_t table = (in.mat >> 31) ? s->n_metal : s->n_rough;
            if (last >= table || last >= (1ull << 31))
                return set_error(ctx, RT_ERR_INVALID, "instance %zu: material index %llu (first %llu, stride %u, %u primitives) outside the %s table of %llu entries",
                                 i, (unsigned long long)last, (unsigned long long)mi, in.mat_stride, g.count, (in.mat >> 31) ? "metal" : "rough",
                                 (unsigned long long)table);
        }
    }
    if (np >= (1ull << 27))
        return set_error(ctx, RT_ERR_UNSUPPORTED, "more than 2^27 primitives (%llu)", (unsigned long long)np);
    if (np >= 0x7FFFFFFFull) return set_error(ctx, RT_ERR_UNSUPPORTED, "too many primitives");
    const uint32_t n = (uint32_t)np;
    s->n_prims = n;
    s->n_tris = (uint32_t)nt;

    const size_t ng = s->groups.size() ? s->groups.size() : 1, ni = s->insts.size() ? s->insts.size() : 1;
    RT_CUDA(ctx, s->d_groups.reserve(ng, &s->bytes));
    RT_CUDA(ctx, s->d_insts.reserve(ni, &s->bytes));
    RT_CUDA(ctx, s->d_inst_base.reserve(ni + 1, &s->bytes));
    if (!s->groups.empty())
        RT_CUDA(ctx, cudaMemcpyAsync(s->d_groups.p, s->groups.data(), s->groups.size() * sizeof(GroupDev), cudaMemcpyHostToDevice, st));
    if (!s->insts.empty())
        RT_CUDA(ctx, cudaMemcpyAsync(s->d_insts.p, s->insts.data(), s->insts.size() * sizeof(InstDev), cudaMemcpyHostToDevice, st));
    RT_CUDA(ctx, cudaMemcpyAsync(s->d_inst_base.p, base.data(), base.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    for (const DevTransforms &d : s->dev_tf)
        if (d.n) k_apply_transforms<<<div_up((uint32_t)d.n, 256u), 256, 0, st>>>(s->d_insts.p, d.ptr, (uint32_t)d.first, (uint32_t)d.n);
    // pageable sources: the copies above are staged before returning, so `base` may die here.

    const size_t n1 = n ? n : 1;
    const uint32_t sort_blocks = div_up(n ? n : 1, kSortTile);
    RT_CUDA(ctx, s->w_geom.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->w_mat.reserve(n1, &s->bytes));
    if (s->compat & 1u) RT_CUDA(ctx, s->w_shade.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->tri_v12.reserve(2 * (nt ? nt : 1), &s->bytes));
    RT_CUDA(ctx, s->tri_nrm.reserve(3 * (nt ? nt : 1), &s->bytes));
    for (int k = 0; k < 2; k++) {
        RT_CUDA(ctx, s->keys[k].reserve(n1, &s->bytes));
        RT_CUDA(ctx, s->vals[k].reserve(n1, &s->bytes));
    }
    RT_CUDA(ctx, s->sort_hist.reserve(sort_scratch_words(sort_blocks), &s->bytes));
    RT_CUDA(ctx, s->bucket_tab.reserve(kExtentBuckets * 8, &s->bytes));
    RT_CUDA(ctx, s->nodes.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->parent_int.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->parent_leaf.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->visit.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->slot_geom.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->slot_prim.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->slot_mat.reserve(n1, &s->bytes));
    RT_CUDA(ctx, s->info.reserve(1, &s->bytes));
    if (s->wide) {
        RT_CUDA(ctx, s->wnodes.reserve(n1, &s->bytes));
        RT_CUDA(ctx, s->wslot_geom.reserve(n1, &s->bytes));
        RT_CUDA(ctx, s->wslot_prim.reserve(n1, &s->bytes));
        RT_CUDA(ctx, s->wslot_mat.reserve(n1, &s->bytes));
    }

    if (!s->ev[0]) {
        for (int k = 0; k < 4; k++) RT_CUDA(ctx, cudaEventCreate(&s->ev[k]));
    }
    RT_CUDA(ctx, cudaEventRecord(s->ev[0], st));

    const uint32_t gb = div_up(n ? n : 1, 256);
    k_init_buckets<<<div_up(kExtentBuckets * 8, 256), 256, 0, st>>>(s->bucket_tab.p);
    RT_CUDA(ctx, cudaMemsetAsync(s->visit.p, 0, n1 * sizeof(uint32_t), st));
    s->launches += 1;
    if (n) {
        k_flatten<<<div_up(n, kFlatItems * 256u), 256, 0, st>>>(s->d_insts.p, s->d_inst_base.p, (uint32_t)s->insts.size(), s->d_groups.p, n,
                                      s->w_geom.p, s->w_mat.p, s->tri_v12.p, s->tri_nrm.p, s->bucket_tab.p,
                                      (s->compat & 1u) ? s->w_shade.p : nullptr, s->compat);
        s->launches += 1;
    }
    k_decide<<<1, 32, 0, st>>>(s->bucket_tab.p, n, s->info.p);
    s->launches += 1;
    if (n) {
        k_morton<<<gb, 256, 0, st>>>(s->w_geom.p, s->tri_v12.p, n, s->info.p, s->keys[0].p, s->vals[0].p);
        s->launches += 1;
        RT_CUDA(ctx, cudaMemsetAsync(s->sort_hist.p, 0, sort_scratch_words(sort_blocks) * sizeof(uint32_t), st));
        k_sort_hist<<<sort_blocks, kSortThreads, 0, st>>>(s->keys[0].p, n, s->sort_hist.p);
        s->launches += 1;
        int src = 0;
        for (int pass = 0; pass < 4; pass++) {
            k_sort_pass<<<sort_blocks, kSortThreads, 0, st>>>(s->keys[src].p, s->vals[src].p, s->keys[src ^ 1].p,
                                                              s->vals[src ^ 1].p, n, pass, s->sort_hist.p, sort_blocks);
            src ^= 1;
            s->launches += 1;
        }
        // 4 passes: result back in buffer 0
        k_gather<<<gb, 256, 0, st>>>(s->vals[0].p, s->w_geom.p, s->w_mat.p, n, s->slot_geom.p, s->slot_prim.p, s->slot_mat.p);
        k_hierarchy<<<gb, 256, 0, st>>>(s->keys[0].p, s->info.p, s->nodes.p, s->parent_int.p, s->parent_leaf.p, (int)(s->wide ? 1u : (s->leaf_max ? s->leaf_max : 4u)));
        if (s->refit_mode >= 1) {
            if (s->refit_mode == 2) {
                const size_t smem = (size_t)6 * kRmqEntries * sizeof(float);
                if (!ctx->rmq_attr_set) {
                    RT_CUDA(ctx, cudaFuncSetAttribute(k_refit_rmq, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                    ctx->rmq_attr_set = true;
                }
                k_refit_rmq<<<div_up(n, kRefitBlock), kRefitBlock, smem, st>>>(s->slot_geom.p, s->tri_v12.p, s->info.p, s->nodes.p, s->parent_int.p, s->parent_leaf.p);
            } else
                k_refit_local<<<div_up(n, kRefitBlock), kRefitBlock, 0, st>>>(s->slot_geom.p, s->tri_v12.p, s->info.p, s->nodes.p, s->parent_int.p, s->parent_leaf.p);
            k_refit_top<<<div_up(2 * n, 256), 256, 0, st>>>(s->slot_geom.p, s->tri_v12.p, s->info.p, s->nodes.p, s->parent_int.p, s->parent_leaf.p, s->visit.p);
            s->launches += 1;
        } else {
            k_refit<<<gb, 256, 0, st>>>(s->slot_geom.p, s->tri_v12.p, s->info.p, s->nodes.p, s->parent_int.p, s->parent_leaf.p, s->visit.p);
        }
        s->launches += 3;
    }
    k_small_tree<<<1, 32, 0, st>>>(s->slot_geom.p, s->tri_v12.p, s->info.p, s->nodes.p);
    s->launches += 1;
    if (n) {
        k_huge_boxes<<<1, 32, 0, st>>>(s->slot_geom.p, s->tri_v12.p, s->info.p);
        s->launches += 1;
    }
    if (s->wide) {
        const void *collapse_fn = s->collapse_mode == 0 ? (const void *)k_collapse : (const void *)k_collapse8;
        int &blocks = s->collapse_mode == 0 ? ctx->collapse_blocks : ctx->collapse8_blocks;
        if (!blocks) {
            int per_sm = 0;
            if (s->collapse_mode == 0) RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_collapse, 128, 0));
            else RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_collapse8, 128, 0));
            if (per_sm < 1) return set_error(ctx, RT_ERR_CUDA, "k_collapse cannot be resident");
            blocks = per_sm * ctx->sm_count;
        }
        const Node *a_nodes = s->nodes.p;
        BuildInfo *a_info = s->info.p;
        uint32_t *a_task = s->parent_leaf.p;  // free after the refit
        WNode *a_wn = s->wnodes.p;
        const float4 *a_sg = s->slot_geom.p;
        const uint32_t *a_sp = s->slot_prim.p, *a_sm = s->slot_mat.p;
        float4 *a_wg = s->wslot_geom.p;
        uint32_t *a_wp = s->wslot_prim.p, *a_wm = s->wslot_mat.p;
        void *args[] = {&a_nodes, &a_info, &a_task, &a_wn, &a_sg, &a_sp, &a_sm, &a_wg, &a_wp, &a_wm};
        RT_CUDA(ctx, cudaLaunchCooperativeKernel(collapse_fn, dim3((unsigned)blocks), dim3(128), args, 0, st));
        s->launches += 1;
    }
    RT_CUDA(ctx, cudaEventRecord(s->ev[1], st));
    RT_CUDA(ctx, cudaGetLastError());
    s->have_build_ev = true;
    s->built = true;
    s->last_stream = st;
    return RT_OK;
}

SceneDev scene_dev(const Scene *s) {
    SceneDev d;
    d.nodes = s->nodes.p;
    d.wnodes = s->wnodes.p;
    d.slot_geom = s->wide ? s->wslot_geom.p : s->slot_geom.p;
    d.slot_prim = s->wide ? s->wslot_prim.p : s->slot_prim.p;
    d.slot_mat = s->wide ? s->wslot_mat.p : s->slot_mat.p;
    d.tri_v12 = s->tri_v12.p;
    d.tri_nrm = s->tri_nrm.p;
    d.roughs = s->d_roughs.p;
    d.metals = s->d_metals.p;
    d.info = s->info.p;
    d.shade_geom = (s->compat & 1u) ? s->w_shade.p : nullptr;
    return d;
}

}  // namespace rtb
