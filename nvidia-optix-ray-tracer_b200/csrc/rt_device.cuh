// rt_device.cuh — device maths, intersectors and BVH traversal.
//
// Floating-point contract: every translation unit that includes this header is
// compiled with --fmad=false, so `a*b+c` is two IEEE roundings exactly as in the CPU oracle; IEEE
// division and square root (nvcc defaults -prec-div=true -prec-sqrt=true).  Where a fused
// multiply-add is wanted for speed and is parity-neutral (conservative box tests) it is requested
// explicitly with __fmaf_rn.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_internal.h"

namespace rtb {

// include/Global/DeviceFunctions.cuh:18-19
#define RT_ZERO 1e-6f
#define RT_INF 1e16f

struct F3 {
    float x, y, z;
};
__device__ __forceinline__ F3 f3(float x, float y, float z) { return F3{x, y, z}; }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ F3 operator-(F3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ F3 operator*(F3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ F3 operator*(float s, F3 a) { return {a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ F3 operator/(F3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// DeviceFunctions.cuh:407-409
__device__ __forceinline__ float dot(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// DeviceFunctions.cuh:410-416
__device__ __forceinline__ F3 cross(F3 a, F3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// DeviceFunctions.cuh:397-404 with IEEE 1/sqrt instead of the approximate rsqrtf
__device__ __forceinline__ F3 normalize3(F3 a) {
    const float l2 = dot(a, a);
    if (l2 <= RT_ZERO * RT_ZERO) return f3(0.0f, 0.0f, 1.0f);
    const float inv = 1.0f / sqrtf(l2);
    return a * inv;
}
__device__ __forceinline__ bool finite3(F3 a) { return isfinite(a.x) && isfinite(a.y) && isfinite(a.z); }

// ------------------------------------------------------------------ XORWOW in registers
// Bit-identical to curand() / curand_uniform() of curand_kernel.h on a curandStateXORWOW whose
// (d, v[0..4]) are these six words.
struct Rng {
    uint32_t d, v0, v1, v2, v3, v4;
};
__device__ __forceinline__ uint32_t rng_next(Rng &s) {
    uint32_t t = s.v0 ^ (s.v0 >> 2);
    s.v0 = s.v1;
    s.v1 = s.v2;
    s.v2 = s.v3;
    s.v3 = s.v4;
    s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
    s.d += 362437u;
    return s.v4 + s.d;
}
// curand_uniform: x * 2^-32 + 2^-33 (two roundings; curand_uniform.h _curand_uniform)
__device__ __forceinline__ float rng_uniform(Rng &s) {
    return (float)rng_next(s) * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}
// DeviceFunctions.cuh:220-222 + 570-582
__device__ __forceinline__ F3 random_space_vector(Rng &s) {
    F3 r;
    float l2;
    do {
        r.x = -1.0f + (1.0f - (-1.0f)) * rng_uniform(s);
        r.y = -1.0f + (1.0f - (-1.0f)) * rng_uniform(s);
        r.z = -1.0f + (1.0f - (-1.0f)) * rng_uniform(s);
        l2 = dot(r, r);
    } while (l2 < RT_ZERO * RT_ZERO);
    r = normalize3(r);
    return r * 1.0f;
}

// ------------------------------------------------------------------ rays and hits
struct Ray {
    F3 o, d;
    float tmin, tmax;
};
struct Hit {
    float t;
    uint32_t slot;  // kNoSlot = miss
    float u, v;
};

// Robust ray/sphere (Haines et al., RT Gems ch.7); same operation order as oracle hit_sphere().
__device__ __forceinline__ float hit_sphere(const Ray &r, float a, float inv_a, float4 s, bool self) {
    const F3 f = r.o - f3(s.x, s.y, s.z);
    const float bp = -dot(f, r.d);
    const float k = bp * inv_a;
    const F3 l = f + r.d * k;
    const float disc = s.w * s.w - dot(l, l);
    if (!(disc >= 0.0f)) return -1.0f;
    const float cc = dot(f, f) - s.w * s.w;
    const float sq = sqrtf(a * disc);
    const float q = (bp >= 0.0f) ? (bp + sq) : (bp - sq);
    const float t_cq = cc / q;
    const float t_qa = q * inv_a;
    const float t_near = (bp >= 0.0f) ? t_cq : t_qa;
    const float t_far = (bp >= 0.0f) ? t_qa : t_cq;
    if (self) {
        const float t = (fabsf(t_near) <= fabsf(t_far)) ? t_far : t_near;
        return (t > r.tmin && t < r.tmax) ? t : -1.0f;
    }
    if (t_near > r.tmin && t_near < r.tmax) return t_near;
    if (t_far > r.tmin && t_far < r.tmax) return t_far;
    return -1.0f;
}

// Moeller-Trumbore, two-sided; same operation order as oracle hit_triangle().
__device__ __forceinline__ float hit_triangle(const Ray &r, F3 v0, F3 v1, F3 v2, float &u_out, float &v_out) {
    const F3 e1 = v1 - v0;
    const F3 e2 = v2 - v0;
    const F3 p = cross(r.d, e2);
    const float det = dot(e1, p);
    if (!(det != 0.0f)) return -1.0f;
    const float inv = 1.0f / det;
    const F3 tv = r.o - v0;
    const float u = dot(tv, p) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return -1.0f;
    const F3 q = cross(tv, e1);
    const float v = dot(r.d, q) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return -1.0f;
    const float t = dot(e2, q) * inv;
    if (!(t > r.tmin && t < r.tmax)) return -1.0f;
    u_out = u;
    v_out = v;
    return t;
}

__device__ __forceinline__ float4 ldg4(const float4 *p) { return __ldg(p); }

// Test the primitive stored in `slot` and keep the closest hit; ties -> smaller global prim id.
__device__ __forceinline__ void test_slot(const SceneDev &sc, const Ray &r, float a, float inv_a,
                                          uint32_t slot, uint32_t skip_slot, Hit &best) {
    const float4 g = ldg4(sc.slot_geom + slot);
    float t, u = 0.0f, v = 0.0f;
    const uint32_t wb = __float_as_uint(g.w);
    if (!(wb & 0x80000000u)) {
        t = hit_sphere(r, a, inv_a, g, slot == skip_slot);
    } else {
        if (slot == skip_slot) return;
        const uint32_t ti = wb & 0x7FFFFFFFu;
        const float4 a1 = ldg4(sc.tri_v12 + 2 * (size_t)ti);
        const float4 a2 = ldg4(sc.tri_v12 + 2 * (size_t)ti + 1);
        t = hit_triangle(r, f3(g.x, g.y, g.z), f3(a1.x, a1.y, a1.z), f3(a2.x, a2.y, a2.z), u, v);
    }
    if (t < 0.0f) return;
    bool take = t < best.t;
    if (t == best.t && best.slot != kNoSlot) take = __ldg(sc.slot_prim + slot) < __ldg(sc.slot_prim + best.slot);
    if (take) {
        best.t = t;
        best.slot = slot;
        best.u = u;
        best.v = v;
    }
}

#define RT_STACK 64

// ------------------------------------------------------------------ warp-cooperative closest hit
// One lane = one ray, but the 32 lanes of a warp move through the tree in two warp-uniform phases
// (Aila & Laine 2009 "while-while" with postponed leaves):
//   phase 1: lanes that hold an internal node visit it (one 64-byte node = both child boxes); the
//            first leaf a lane meets is postponed and the lane keeps descending speculatively;
//   phase 2: lanes that hold a postponed leaf test its primitives, one primitive per iteration.
// A phase is left when fewer than t1 / t2 lanes still want it and other lanes have work of the
// other kind.  Finished lanes are refilled from a global ray counter (dynamic fetch) once at least
// `fetch_min` lanes are idle, so a long ray does not hold 31 idle lanes.
// The result is independent of all three thresholds: boxes are conservative and ties are resolved
// by (t, global primitive id), i.e. it equals brute force over all primitives.
constexpr int kSent = 0x7FFFFFFF;  // "no node": stack empty
__device__ __forceinline__ bool is_internal(int n) { return (uint32_t)n < 0x7FFFFFFFu; }

struct ExtendIn {
    const float4 *o;       // path queue: origin.xyz, bits(path id)   | user rays: origin.xyz, tmin
    const float4 *d;       // path queue: dir.xyz, bits(skip slot)    | user rays: dir.xyz, tmax
    const uint32_t *skip;  // NULL = path queue layout, else user-ray layout with skip slot per ray
};

__device__ __forceinline__ void extend_warp(const SceneDev &sc, const ExtendIn &in, uint32_t n, uint32_t *fetch_ctr,
                                            float4 *__restrict__ hits, int t1, int t2, int fetch_min) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt = (1u << lane) - 1u;
    const float4 *nodes = reinterpret_cast<const float4 *>(sc.nodes);
    const BuildInfo *info = sc.info;
    const uint32_t n_huge = info->n_huge;
    const bool have_tree = info->n_bvh != 0;

    int stack[RT_STACK];
    int sp = 0, node = kSent, leaf = 0;
    uint32_t li = 0, idx = 0, skip_slot = kNoSlot;
    Ray r;
    r.o = r.d = f3(0, 0, 0);
    r.tmin = RT_ZERO;
    r.tmax = RT_INF;
    Hit best;
    best.t = 0.0f;
    best.slot = kNoSlot;
    best.u = best.v = 0.0f;
    float a = 1.0f, inv_a = 1.0f, ix = 0, iy = 0, iz = 0, ox = 0, oy = 0, oz = 0;
    bool alive = false, exhausted = false;

    for (;;) {
        // ---- dynamic fetch
        const uint32_t dead = __ballot_sync(FULL, !alive);
        if (dead && !exhausted && (dead == FULL || __popc(dead) >= fetch_min)) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(fetch_ctr, (uint32_t)__popc(dead));
            base = __shfl_sync(FULL, base, 0);
            if (base + __popc(dead) >= n) exhausted = true;
            const uint32_t i = base + __popc(dead & lt);
            if (!alive && i < n) {
                const float4 o4 = in.o[i], d4 = in.d[i];
                idx = i;
                r.o = f3(o4.x, o4.y, o4.z);
                r.d = f3(d4.x, d4.y, d4.z);
                if (in.skip) {
                    r.tmin = o4.w;
                    r.tmax = d4.w;
                    skip_slot = in.skip[i];
                } else {
                    r.tmin = RT_ZERO;  // Shader.cu:231-233, 264-267
                    r.tmax = RT_INF;
                    skip_slot = __float_as_uint(d4.w);
                }
                best.t = r.tmax;
                best.slot = kNoSlot;
                best.u = best.v = 0.0f;
                a = dot(r.d, r.d);
                inv_a = 1.0f / a;
                for (uint32_t k = 0; k < n_huge; k++) test_slot(sc, r, a, inv_a, info->huge_slot[k], skip_slot, best);
                // reciprocal direction, clamped so that 0 * inf never appears
                const float ex = fabsf(r.d.x) < 1e-30f ? copysignf(1e-30f, r.d.x) : r.d.x;
                const float ey = fabsf(r.d.y) < 1e-30f ? copysignf(1e-30f, r.d.y) : r.d.y;
                const float ez = fabsf(r.d.z) < 1e-30f ? copysignf(1e-30f, r.d.z) : r.d.z;
                ix = 1.0f / ex; iy = 1.0f / ey; iz = 1.0f / ez;
                ox = -r.o.x * ix; oy = -r.o.y * iy; oz = -r.o.z * iz;
                sp = 0;
                leaf = 0;
                li = 0;
                node = have_tree ? 0 : kSent;
                alive = true;
            }
        }
        if (!__any_sync(FULL, alive)) break;

        // ---- phase 1: internal nodes
        bool first = true;
        for (;;) {
            const bool want = alive && is_internal(node);
            const uint32_t mw = __ballot_sync(FULL, want);
            if (!mw) break;
            if (!first && __popc(mw) < t1 && __any_sync(FULL, alive && leaf != 0)) break;
            first = false;
            if (want) {
                const float4 n0 = ldg4(nodes + 4 * (size_t)node);
                const float4 n1 = ldg4(nodes + 4 * (size_t)node + 1);
                const float4 n2 = ldg4(nodes + 4 * (size_t)node + 2);
                const float2 n3 = __ldg(reinterpret_cast<const float2 *>(nodes + 4 * (size_t)node + 3));
                const float c0x0 = __fmaf_rn(n0.x, ix, ox), c0x1 = __fmaf_rn(n0.y, ix, ox);
                const float c0y0 = __fmaf_rn(n0.z, iy, oy), c0y1 = __fmaf_rn(n0.w, iy, oy);
                const float c0z0 = __fmaf_rn(n2.x, iz, oz), c0z1 = __fmaf_rn(n2.y, iz, oz);
                const float c1x0 = __fmaf_rn(n1.x, ix, ox), c1x1 = __fmaf_rn(n1.y, ix, ox);
                const float c1y0 = __fmaf_rn(n1.z, iy, oy), c1y1 = __fmaf_rn(n1.w, iy, oy);
                const float c1z0 = __fmaf_rn(n2.z, iz, oz), c1z1 = __fmaf_rn(n2.w, iz, oz);
                const float lo0 = fmaxf(fmaxf(fminf(c0x0, c0x1), fminf(c0y0, c0y1)), fmaxf(fminf(c0z0, c0z1), r.tmin));
                const float hi0 = fminf(fminf(fmaxf(c0x0, c0x1), fmaxf(c0y0, c0y1)), fminf(fmaxf(c0z0, c0z1), best.t));
                const float lo1 = fmaxf(fmaxf(fminf(c1x0, c1x1), fminf(c1y0, c1y1)), fmaxf(fminf(c1z0, c1z1), r.tmin));
                const float hi1 = fminf(fminf(fmaxf(c1x0, c1x1), fmaxf(c1y0, c1y1)), fminf(fmaxf(c1z0, c1z1), best.t));
                // an empty child (k_small_tree) keeps an inverted +-inf box, which the min/max slab test would pass
                const bool h0 = lo0 <= hi0 * 1.000002f && n0.x <= n0.y;
                const bool h1 = lo1 <= hi1 * 1.000002f && n1.x <= n1.y;
                const int ch0 = __float_as_int(n3.x), ch1 = __float_as_int(n3.y);
                if (h0 && h1) {
                    const bool swap = lo1 < lo0;
                    node = swap ? ch1 : ch0;
                    stack[sp++] = swap ? ch0 : ch1;  // depth <= 62 < RT_STACK (30 key bits + 32 index bits)
                } else if (h0) {
                    node = ch0;
                } else if (h1) {
                    node = ch1;
                } else {
                    node = sp ? stack[--sp] : kSent;
                }
                if (node < 0 && leaf == 0) {  // postpone the first leaf, keep descending
                    leaf = node;
                    node = sp ? stack[--sp] : kSent;
                }
            }
        }

        // ---- phase 2: postponed leaves, one primitive per iteration
        first = true;
        for (;;) {
            const bool hl = alive && leaf != 0;
            const uint32_t ml = __ballot_sync(FULL, hl);
            if (!ml) break;
            if (!first && __popc(ml) < t2 && __any_sync(FULL, alive && is_internal(node))) break;
            first = false;
            if (hl) {
                const uint32_t enc = (uint32_t)leaf;
                const uint32_t cnt = ((enc >> kLeafCountShift) & 7u) + 1u;
                test_slot(sc, r, a, inv_a, (enc & kLeafFirstMask) + li, skip_slot, best);
                if (++li == cnt) {
                    li = 0;
                    leaf = 0;
                    if (node < 0) {
                        leaf = node;
                        node = sp ? stack[--sp] : kSent;
                    }
                }
            }
        }

        // ---- retire finished rays
        if (alive && node == kSent && leaf == 0) {
            hits[idx] = make_float4(best.slot == kNoSlot ? RT_INF : best.t, __uint_as_float(best.slot), best.u, best.v);
            alive = false;
        }
    }
}

// ------------------------------------------------------------------ 8-wide compressed BVH traversal
// Same two-phase warp scheme as extend_warp, over WNode (rt_internal.h).  A lane keeps
//   (gbase, gmask): current node group = child_base, [7:0] imask of the node, [15:8] pending internal-child
//                   hits at their traversal priority (slot ^ octant, nearer octant = higher bit);
//   (pbase, pmask), (pbase2, pmask2): up to two postponed primitive groups, [15:0] = which slots hold a
//                   primitive (primitive of slot b = pbase + popc(present below b)), [31:16] pending hits.
// The stack holds node groups only (8 bytes each).  Children are visited in octant order, which is
// a heuristic only: every hit child is eventually visited, so the closest hit equals brute force.
#define RT_WSTACK 64
// Traversal stack = shared-memory short stack + local-memory overflow.  The first RT_SH_STACK entries of a lane's
// stack live in shared memory, laid out [entry][thread] so that the 32 lanes of a warp touch 32 consecutive
// 8-byte words (no bank conflicts, no tag lookups, no L1 write-back traffic); deeper entries — the 8-wide tree
// over 10^7 primitives is ~8 levels deep and a lane pushes at most one group per level, so they are rare — spill
// to a per-thread local array.  (Round 1 kept all 64 entries in local memory: 23 % of the kernel's L1 bytes.)
#ifndef RT_SH_STACK
#define RT_SH_STACK 12
#endif
#define RT_EXTEND_THREADS 128
struct WideStack {
    uint2 (*sh)[RT_EXTEND_THREADS];  // [RT_SH_STACK][threads of the block]
    uint2 ovf[RT_WSTACK - RT_SH_STACK];
    __device__ __forceinline__ void push(int sp, uint2 e) {
        if (sp < RT_SH_STACK) sh[sp][threadIdx.x] = e; else ovf[sp - RT_SH_STACK] = e;
    }
    __device__ __forceinline__ uint2 pop(int sp) const {
        return sp < RT_SH_STACK ? sh[sp][threadIdx.x] : ovf[sp - RT_SH_STACK];
    }
};

// byte j of w as the float 2^23 + q (exact): the 2^23 is folded into the FMA addend of the slab test.
// That addend is rounded once, which can move a child plane by up to half a quantisation step; the
// build widens every quantised box by one step per side to stay conservative (collapse_one).
template <int J>
__device__ __forceinline__ float qfloat(uint32_t w, uint32_t magic) {
    // prmt with the selector as the immediate and 0x4B000000 in a register (the compiler otherwise
    // inlines the constant and spends a register move per selector)
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(magic), "n"(0x7650 + J));
    return __uint_as_float(r);
}

// Two bytes per PRMT (default; -DRT_BYTE_DEQ restores one byte per PRMT: 123.5 instead of 121.4 ms of traversal per C4
// frame).  ncu (profiles/r5b_k_extend_wide_ncu.txt) shows the traversal kernel bound by the ALU pipe (70 % busy: PRMT,
// FMNMX, LOP3, SEL, FSETP all live there) with the FMA pipe at 26 %.  Bytes j and j+1 of w, each with 0x64 above it, are
// the two halves 1024 + q of one register (fp16 0x6400 | q = 1024 + q exactly); the two half -> float conversions
// (HADD2.F32) run on the FMA pipe: 24 PRMT + 48 HADD2.F32 per node instead of 48 PRMT.  1024 is folded into the FMA
// addend like 2^23 is for qfloat (and rounds that addend 2^13 times finer).  Moving more ALU work over was tried and is
// slower: the hit mask from the sign bit of fma(thi, 1.000002, -tlo) via mul.hi / mad.lo (124.3 ms) and the near / far
// word selects as integer multiply-adds (127.5 ms); both together 129.0 ms (r5d).
template <int J>
__device__ __forceinline__ void qfloat2(uint32_t w, uint32_t magic, float &f0, float &f1) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(magic), "n"(0x5140 + J * 0x101));  // result bytes: w[J], magic[0], w[J+1], magic[1]
    asm("{ .reg .b16 lo, hi; mov.b32 {lo, hi}, %2; cvt.f32.f16 %0, lo; cvt.f32.f16 %1, hi; }" : "=f"(f0), "=f"(f1) : "r"(r));
}
#ifndef RT_BYTE_DEQ
#define RT_HALF_DEQ 1
#endif
#ifdef RT_HALF_DEQ
#define RT_DEQ_MAGIC 0x64646464u
#define RT_DEQ_BIAS 1024.0f
#else
#define RT_DEQ_MAGIC 0x4B000000u
#define RT_DEQ_BIAS 8388608.0f
#endif

// Blackwell packed fp32: two fused multiply-adds per issue slot (FFMA2).  (d0, d1) = (a0, a1) * s + b
// with s and b given as register pairs holding the same value twice.
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void ffma2(float &d0, float &d1, float a0, float a1, unsigned long long s2, unsigned long long b2) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(pack2(a0, a1)), "l"(s2), "l"(b2));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

// s_perm[o][m]: the 8-bit child mask m (slot order) permuted to traversal priority for ray octant o,
// bit (j ^ o) of the result = bit j of m.  2 KB of shared memory, filled by every block at start.
__device__ __forceinline__ void fill_perm_table(uint8_t (*s_perm)[256]) {
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x) {
        const uint32_t o = i >> 8, m = i & 255u;
        uint32_t r = 0;
#pragma unroll
        for (uint32_t j = 0; j < 8; j++) r |= ((m >> j) & 1u) << (j ^ o);
        s_perm[o][m] = (uint8_t)r;
    }
    __syncthreads();
}

#ifdef RT_TRAV_STATS
// Experiment build only (make EXTRA=-DRT_TRAV_STATS OUT=...): per-visit counters, printed by rt_scene_destroy.
//   0 node visits, 1 visits where no valid child was hit, 2 visits whose node box (origin .. origin + 255 steps) misses
//   [tmin, best.t], 3 the same with the tight box (min/max of the child planes), 4 primitive tests, 5 rays
static __device__ unsigned long long g_trav_stats[8];
__device__ __forceinline__ uint32_t bmax4(uint32_t a, uint32_t b) { return __vmaxu4(a, b); }
__device__ __forceinline__ uint32_t bmin4(uint32_t a, uint32_t b) { return __vminu4(a, b); }
__device__ __forceinline__ uint32_t hmax8(uint32_t a, uint32_t b) { uint32_t m = bmax4(a, b); m = bmax4(m, m >> 16); m = bmax4(m, m >> 8); return m & 255u; }
__device__ __forceinline__ uint32_t hmin8(uint32_t a, uint32_t b) { uint32_t m = bmin4(a, b); m = bmin4(m, m >> 16); m = bmin4(m, m >> 8); return m & 255u; }
#endif

// ---- per-lane state of one closest-hit query over the 8-wide tree (kept in registers)
struct WideTrav {
    uint32_t gbase, gmask, pbase, pmask, pbase2, pmask2;
    int sp;
    bool huge_done, huge_wait;
    float tmin, tmax, a, inv_a, ix, iy, iz, ox, oy, oz;
    uint32_t oct, skip_slot;
    Hit best;
};
__device__ __forceinline__ void trav_clear(WideTrav &t) {
    t.gbase = t.gmask = t.pbase = t.pmask = t.pbase2 = t.pmask2 = 0u;
    t.sp = 0;
    t.huge_done = true;
    t.huge_wait = false;
    t.tmin = RT_ZERO;
    t.tmax = RT_INF;
    t.a = t.inv_a = 1.0f;
    t.ix = t.iy = t.iz = t.ox = t.oy = t.oz = 0.0f;
    t.oct = 0u;
    t.skip_slot = kNoSlot;
    t.best.t = 0.0f;
    t.best.slot = kNoSlot;
    t.best.u = t.best.v = 0.0f;
}
// origin and direction are NOT kept in registers: the node test only needs 1/d and -o/d, the primitive
// tests re-read the ray (queue entry: an L1/L2 hit; fused path kernel: shared memory)
__device__ __forceinline__ void trav_start(WideTrav &t, F3 ro, F3 rd, float tmin, float tmax, uint32_t skip_slot,
                                           uint32_t n_huge, bool have_tree) {
    t.tmin = tmin;
    t.tmax = tmax;
    t.skip_slot = skip_slot;
    t.best.t = tmax;
    t.best.slot = kNoSlot;
    t.best.u = t.best.v = 0.0f;
    t.a = dot(rd, rd);
    t.inv_a = 1.0f / t.a;
    // reciprocal direction, clamped so that 0 * inf never appears
    const float ex = fabsf(rd.x) < 1e-30f ? copysignf(1e-30f, rd.x) : rd.x;
    const float ey = fabsf(rd.y) < 1e-30f ? copysignf(1e-30f, rd.y) : rd.y;
    const float ez = fabsf(rd.z) < 1e-30f ? copysignf(1e-30f, rd.z) : rd.z;
    t.ix = 1.0f / ex; t.iy = 1.0f / ey; t.iz = 1.0f / ez;
    t.ox = -ro.x * t.ix; t.oy = -ro.y * t.iy; t.oz = -ro.z * t.iz;
    t.oct = (t.ix >= 0.0f ? 1u : 0u) | (t.iy >= 0.0f ? 2u : 0u) | (t.iz >= 0.0f ? 4u : 0u);
    t.huge_done = n_huge == 0;
    t.huge_wait = false;
    t.sp = 0;
    t.pmask = t.pmask2 = 0;
#ifdef RT_TRAV_STATS
    atomicAdd(&g_trav_stats[5], 1ull);
#endif
    t.gbase = 0;
    t.gmask = have_tree ? 0x8000u : 0u;  // root: slot 7^oct of an imask-0 group -> index 0
}
__device__ __forceinline__ bool trav_has_nodes(const WideTrav &t) { return (t.gmask >> 8) != 0 || t.sp > 0; }
__device__ __forceinline__ bool trav_wants_node(const WideTrav &t) { return (t.pmask2 >> 16) == 0 && trav_has_nodes(t); }
__device__ __forceinline__ bool trav_wants_prim(const WideTrav &t) { return (t.pmask >> 16) != 0; }
__device__ __forceinline__ bool trav_tree_done(const WideTrav &t) { return !trav_has_nodes(t) && (t.pmask >> 16) == 0; }


// One wide-node visit of a lane: pop the next node group entry, fetch the node (five 128-bit loads), slab-test
// its 8 quantised child boxes, queue the hit internal children (octant order) and the hit leaf primitives.
// `magic` must be RT_DEQ_MAGIC; it arrives as a kernel argument so that it lives in a register (qfloat / qfloat2).
__device__ __forceinline__ void trav_node(WideTrav &t, WideStack &stack, const float4 *__restrict__ wn, uint32_t magic,
                                          const uint8_t (*s_perm)[256]) {
    if ((t.gmask >> 8) == 0) {
        const uint2 e = stack.pop(--t.sp);
        t.gbase = e.x;
        t.gmask = e.y;
    }
    const uint32_t bit = 31u - (uint32_t)__clz(t.gmask);
    t.gmask &= ~(1u << bit);
    const uint32_t slot = (bit - 8u) ^ t.oct;
    const uint32_t child = t.gbase + __popc(t.gmask & ((1u << slot) - 1u));
    if (t.gmask >> 8) stack.push(t.sp++, make_uint2(t.gbase, t.gmask));

    const float4 *q = wn + 5 * (size_t)child;
    const float4 w0 = ldg4(q), w1 = ldg4(q + 1), w2 = ldg4(q + 2), w3 = ldg4(q + 3), w4 = ldg4(q + 4);
    const uint32_t eim = __float_as_uint(w0.w);
    const float sx = __uint_as_float((eim & 0xFFu) << 23) * t.ix;
    const float sy = __uint_as_float(((eim >> 8) & 0xFFu) << 23) * t.iy;
    const float sz = __uint_as_float(((eim >> 16) & 0xFFu) << 23) * t.iz;
    // t = (2^23 + q) * s + (b - 2^23 * s), b = (origin - o) / d
    const float bx = __fmaf_rn(-RT_DEQ_BIAS, sx, __fmaf_rn(w0.x, t.ix, t.ox));
    const float by = __fmaf_rn(-RT_DEQ_BIAS, sy, __fmaf_rn(w0.y, t.iy, t.oy));
    const float bz = __fmaf_rn(-RT_DEQ_BIAS, sz, __fmaf_rn(w0.z, t.iz, t.oz));
    const unsigned long long sx2 = pack2(sx, sx), sy2 = pack2(sy, sy), sz2 = pack2(sz, sz);
    const unsigned long long bx2 = pack2(bx, bx), by2 = pack2(by, by), bz2 = pack2(bz, bz);
    const bool px = t.ix >= 0.0f, py = t.iy >= 0.0f, pz = t.iz >= 0.0f;
    uint32_t hm = 0;  // bit j = child slot j hit
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t lox = __float_as_uint(h ? w2.y : w2.x), loy = __float_as_uint(h ? w2.w : w2.z);
        const uint32_t loz = __float_as_uint(h ? w3.y : w3.x), hix = __float_as_uint(h ? w3.w : w3.z);
        const uint32_t hiy = __float_as_uint(h ? w4.y : w4.x), hiz = __float_as_uint(h ? w4.w : w4.z);
        const uint32_t nx = px ? lox : hix, fx = px ? hix : lox;
        const uint32_t ny = py ? loy : hiy, fy = py ? hiy : loy;
        const uint32_t nz = pz ? loz : hiz, fz = pz ? hiz : loz;
#pragma unroll
        for (int jp = 0; jp < 2; jp++) {
            float tnx[2], tfx[2], tny[2], tfy[2], tnz[2], tfz[2];
#ifdef RT_HALF_DEQ
            {
                float a0, a1;
                if (jp == 0) {
                    qfloat2<0>(nx, magic, a0, a1); ffma2(tnx[0], tnx[1], a0, a1, sx2, bx2);
                    qfloat2<0>(fx, magic, a0, a1); ffma2(tfx[0], tfx[1], a0, a1, sx2, bx2);
                    qfloat2<0>(ny, magic, a0, a1); ffma2(tny[0], tny[1], a0, a1, sy2, by2);
                    qfloat2<0>(fy, magic, a0, a1); ffma2(tfy[0], tfy[1], a0, a1, sy2, by2);
                    qfloat2<0>(nz, magic, a0, a1); ffma2(tnz[0], tnz[1], a0, a1, sz2, bz2);
                    qfloat2<0>(fz, magic, a0, a1); ffma2(tfz[0], tfz[1], a0, a1, sz2, bz2);
                } else {
                    qfloat2<2>(nx, magic, a0, a1); ffma2(tnx[0], tnx[1], a0, a1, sx2, bx2);
                    qfloat2<2>(fx, magic, a0, a1); ffma2(tfx[0], tfx[1], a0, a1, sx2, bx2);
                    qfloat2<2>(ny, magic, a0, a1); ffma2(tny[0], tny[1], a0, a1, sy2, by2);
                    qfloat2<2>(fy, magic, a0, a1); ffma2(tfy[0], tfy[1], a0, a1, sy2, by2);
                    qfloat2<2>(nz, magic, a0, a1); ffma2(tnz[0], tnz[1], a0, a1, sz2, bz2);
                    qfloat2<2>(fz, magic, a0, a1); ffma2(tfz[0], tfz[1], a0, a1, sz2, bz2);
                }
            }
#else
            if (jp == 0) {
                ffma2(tnx[0], tnx[1], qfloat<0>(nx, magic), qfloat<1>(nx, magic), sx2, bx2);
                ffma2(tfx[0], tfx[1], qfloat<0>(fx, magic), qfloat<1>(fx, magic), sx2, bx2);
                ffma2(tny[0], tny[1], qfloat<0>(ny, magic), qfloat<1>(ny, magic), sy2, by2);
                ffma2(tfy[0], tfy[1], qfloat<0>(fy, magic), qfloat<1>(fy, magic), sy2, by2);
                ffma2(tnz[0], tnz[1], qfloat<0>(nz, magic), qfloat<1>(nz, magic), sz2, bz2);
                ffma2(tfz[0], tfz[1], qfloat<0>(fz, magic), qfloat<1>(fz, magic), sz2, bz2);
            } else {
                ffma2(tnx[0], tnx[1], qfloat<2>(nx, magic), qfloat<3>(nx, magic), sx2, bx2);
                ffma2(tfx[0], tfx[1], qfloat<2>(fx, magic), qfloat<3>(fx, magic), sx2, bx2);
                ffma2(tny[0], tny[1], qfloat<2>(ny, magic), qfloat<3>(ny, magic), sy2, by2);
                ffma2(tfy[0], tfy[1], qfloat<2>(fy, magic), qfloat<3>(fy, magic), sy2, by2);
                ffma2(tnz[0], tnz[1], qfloat<2>(nz, magic), qfloat<3>(nz, magic), sz2, bz2);
                ffma2(tfz[0], tfz[1], qfloat<2>(fz, magic), qfloat<3>(fz, magic), sz2, bz2);
            }
#endif
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const float tlo = fmaxf(fmaxf(tnx[e], tny[e]), fmaxf(tnz[e], t.tmin));
                const float thi = fminf(fminf(tfx[e], tfy[e]), fminf(tfz[e], t.best.t));
                if (tlo <= thi * 1.000002f) hm |= 1u << (4 * h + 2 * jp + e);  // empty slots hold an inverted box
            }
        }
    }
    const uint32_t imask = eim >> 24, leafmask = __float_as_uint(w1.z) & 0xFFu;
#ifdef RT_TRAV_STATS
    {
        atomicAdd(&g_trav_stats[0], 1ull);
        if ((hm & (imask | leafmask)) == 0) atomicAdd(&g_trav_stats[1], 1ull);
        const float b0x = __fmaf_rn(w0.x, t.ix, t.ox), b0y = __fmaf_rn(w0.y, t.iy, t.oy), b0z = __fmaf_rn(w0.z, t.iz, t.oz);
        for (int tight = 0; tight < 2; tight++) {
            float qlx = 0, qly = 0, qlz = 0, qhx = 255, qhy = 255, qhz = 255;
            if (tight) {
                qlx = (float)hmin8(__float_as_uint(w2.x), __float_as_uint(w2.y)); qly = (float)hmin8(__float_as_uint(w2.z), __float_as_uint(w2.w));
                qlz = (float)hmin8(__float_as_uint(w3.x), __float_as_uint(w3.y)); qhx = (float)hmax8(__float_as_uint(w3.z), __float_as_uint(w3.w));
                qhy = (float)hmax8(__float_as_uint(w4.x), __float_as_uint(w4.y)); qhz = (float)hmax8(__float_as_uint(w4.z), __float_as_uint(w4.w));
            }
            const float x0 = __fmaf_rn(qlx, sx, b0x), x1 = __fmaf_rn(qhx, sx, b0x);
            const float y0 = __fmaf_rn(qly, sy, b0y), y1 = __fmaf_rn(qhy, sy, b0y);
            const float z0 = __fmaf_rn(qlz, sz, b0z), z1 = __fmaf_rn(qhz, sz, b0z);
            const float tlo = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), t.tmin));
            const float thi = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), t.best.t));
            if (!(tlo <= thi * 1.000002f)) atomicAdd(&g_trav_stats[2 + tight], 1ull);
        }
    }
#endif
    t.gbase = __float_as_uint(w1.x);
    t.gmask = imask | ((uint32_t)s_perm[t.oct][hm & imask] << 8);
    const uint32_t pm = hm & leafmask;
    if (pm) {
        if ((t.pmask >> 16) == 0) {
            t.pbase = __float_as_uint(w1.y);
            t.pmask = leafmask | (pm << 16);
        } else {
            t.pbase2 = __float_as_uint(w1.y);
            t.pmask2 = leafmask | (pm << 16);
        }
    }
}

// One postponed primitive test of a lane (ray origin / direction supplied by the caller).
__device__ __forceinline__ void trav_prim(WideTrav &t, const SceneDev &sc, F3 ro, F3 rd) {
    const uint32_t b = (uint32_t)__ffs(t.pmask >> 16) - 1u;
    t.pmask &= ~(0x10000u << b);
#ifdef RT_TRAV_STATS
    atomicAdd(&g_trav_stats[4], 1ull);
#endif
    Ray r;
    r.o = ro;
    r.d = rd;
    r.tmin = t.tmin;
    r.tmax = t.tmax;
    test_slot(sc, r, t.a, t.inv_a, t.pbase + __popc(t.pmask & ((1u << b) - 1u)), t.skip_slot, t.best);
    if ((t.pmask >> 16) == 0 && (t.pmask2 >> 16) != 0) {
        t.pbase = t.pbase2;
        t.pmask = t.pmask2;
        t.pmask2 = 0;
    }
}

// Rays whose tree walk is over: outlier primitives (BuildInfo.huge_*).  A segment with both ends inside the
// outlier-free box is done at once; the others wait (huge_wait) until at least t3 lanes of the warp are in the
// same state or nothing else is left to do, then cull the outliers by their boxes against the closest hit so
// far and queue the survivors for the primitive phase.  WARP-COLLECTIVE: every lane must call it; `alive` =
// the lane holds a ray, (ro, rd) its origin and direction (read only when the tree walk has just ended).
__device__ __forceinline__ void trav_outliers(WideTrav &t, bool alive, const BuildInfo *__restrict__ info, uint32_t n_huge,
                                              int t3, F3 ro, F3 rd) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    if (alive && !t.huge_done && !t.huge_wait && trav_tree_done(t)) {
        bool need = t.best.slot == kNoSlot;
        if (!need) {
            const float ex = ro.x + t.best.t * rd.x, ey = ro.y + t.best.t * rd.y, ez = ro.z + t.best.t * rd.z;
            const float l0 = info->free_lo[0], l1 = info->free_lo[1], l2 = info->free_lo[2];
            const float h0 = info->free_hi[0], h1 = info->free_hi[1], h2 = info->free_hi[2];
            const bool in_o = ro.x > l0 && ro.x < h0 && ro.y > l1 && ro.y < h1 && ro.z > l2 && ro.z < h2;
            const bool in_e = ex > l0 && ex < h0 && ey > l1 && ey < h1 && ez > l2 && ez < h2;
            need = !(in_o && in_e);
        }
        t.huge_wait = need;
        t.huge_done = !need;
    }
    const uint32_t m3 = __ballot_sync(FULL, alive && t.huge_wait);
    if (m3 && (__popc(m3) >= t3 || !__any_sync(FULL, alive && !t.huge_wait && !trav_tree_done(t)))) {
        // warp-cooperative cull: the waiting rays take turns, lane k tests outlier box k
        uint32_t mine = 0;
        float hl0 = 0, hl1 = 0, hl2 = 0, hh0 = 0, hh1 = 0, hh2 = 0;
        if (lane < n_huge) {
            hl0 = info->huge_lo[lane][0]; hl1 = info->huge_lo[lane][1]; hl2 = info->huge_lo[lane][2];
            hh0 = info->huge_hi[lane][0]; hh1 = info->huge_hi[lane][1]; hh2 = info->huge_hi[lane][2];
        }
        for (uint32_t rem = m3; rem; rem &= rem - 1u) {
            const int src = __ffs(rem) - 1;
            const float rix = __shfl_sync(FULL, t.ix, src), riy = __shfl_sync(FULL, t.iy, src), riz = __shfl_sync(FULL, t.iz, src);
            const float rox = __shfl_sync(FULL, t.ox, src), roy = __shfl_sync(FULL, t.oy, src), roz = __shfl_sync(FULL, t.oz, src);
            const float rt0 = __shfl_sync(FULL, t.tmin, src), rt1 = __shfl_sync(FULL, t.best.t, src);
            const float x0 = __fmaf_rn(hl0, rix, rox), x1 = __fmaf_rn(hh0, rix, rox);
            const float y0 = __fmaf_rn(hl1, riy, roy), y1 = __fmaf_rn(hh1, riy, roy);
            const float z0 = __fmaf_rn(hl2, riz, roz), z1 = __fmaf_rn(hh2, riz, roz);
            const float tlo = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), rt0));
            const float thi = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), rt1));
            const uint32_t mm = __ballot_sync(FULL, lane < n_huge && tlo <= thi * 1.000002f);
            if ((int)lane == src) mine = mm;
        }
        if (alive && t.huge_wait) {
            t.pbase = info->huge_slot[0];
            t.pmask = 0xFFFFu | (mine << 16);  // all kMaxHuge <= 16 slots present: primitive k = pbase + k
            t.huge_wait = false;
            t.huge_done = true;
        }
    }
}
__device__ __forceinline__ bool trav_finished(const WideTrav &t) { return t.huge_done && trav_tree_done(t); }

// Closest hit for a buffer of rays (wavefront form): persistent warps, dynamic ray fetch.
__device__ __forceinline__ void extend_warp_wide(const SceneDev &sc, const ExtendIn &in, uint32_t n, uint32_t *fetch_ctr,
                                                 float4 *__restrict__ hits, int t1, int t2, int t3, int fetch_min, uint32_t magic,
                                                 const uint8_t (*s_perm)[256], uint2 (*s_stack)[RT_EXTEND_THREADS]) {
    const uint32_t FULL = 0xFFFFFFFFu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt = (1u << lane) - 1u;
    const float4 *wn = reinterpret_cast<const float4 *>(sc.wnodes);
    const BuildInfo *info = sc.info;
    const uint32_t n_huge = info->n_huge;
    const bool have_tree = info->n_bvh != 0;

    WideStack stack;
    stack.sh = s_stack;
    WideTrav t;
    trav_clear(t);
    uint32_t idx = 0;
    bool alive = false, exhausted = false;

    for (;;) {
        // ---- dynamic fetch
        const uint32_t dead = __ballot_sync(FULL, !alive);
        if (dead && !exhausted && (dead == FULL || __popc(dead) >= fetch_min)) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(fetch_ctr, (uint32_t)__popc(dead));
            base = __shfl_sync(FULL, base, 0);
            if (base + __popc(dead) >= n) exhausted = true;
            const uint32_t i = base + __popc(dead & lt);
            if (!alive && i < n) {
                const float4 o4 = in.o[i], d4 = in.d[i];
                idx = i;
                // path queue: Shader.cu:231-233, 264-267 (tmin 1e-6, tmax 1e16); user rays carry their own range
                trav_start(t, f3(o4.x, o4.y, o4.z), f3(d4.x, d4.y, d4.z), in.skip ? o4.w : RT_ZERO, in.skip ? d4.w : RT_INF,
                           in.skip ? in.skip[i] : __float_as_uint(d4.w), n_huge, have_tree);
                alive = true;
            }
        }
        if (!__any_sync(FULL, alive)) break;

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
            if (hl) {
                const float4 o4 = in.o[idx], d4 = in.d[idx];
                trav_prim(t, sc, f3(o4.x, o4.y, o4.z), f3(d4.x, d4.y, d4.z));
            }
        }

        // ---- outlier primitives
        {
            F3 ro = f3(0, 0, 0), rd = f3(0, 0, 0);
            if (alive && !t.huge_done && !t.huge_wait && trav_tree_done(t) && t.best.slot != kNoSlot) {
                const float4 o4 = in.o[idx], d4 = in.d[idx];
                ro = f3(o4.x, o4.y, o4.z);
                rd = f3(d4.x, d4.y, d4.z);
            }
            trav_outliers(t, alive, info, n_huge, t3, ro, rd);
        }
        // ---- retire finished rays
        if (alive && trav_finished(t)) {
            hits[idx] = make_float4(t.best.slot == kNoSlot ? RT_INF : t.best.t, __uint_as_float(t.best.slot), t.best.u, t.best.v);
            alive = false;
        }
    }
}

}  // namespace rtb
