/*
 * rt_oracle.cpp — CPU ORACLE for the B200 path tracer.  TEST INFRASTRUCTURE ONLY.
 *
 * A scalar C++17 restatement (OpenMP over pixel rows) of the reference's hot path:
 *   raygen / closest-hit / miss programs      shader/Shader.cu:94-287
 *   trace depth constant                      include/Global/Shader.cuh:8
 *   vector maths, normalize, colour, RNG use  include/Global/DeviceFunctions.cuh:18-20,133-148,
 *                                             153-212,216-222,397-416,570-582
 *   RNG state init                            src/Global/HostFunctions.cu:122-127
 *   camera basis                              src/GraphicsAPI/SDL_GraphicsWindow.cu:4-14
 *   per-frame transforms                      src/Global/RendererTime.cu:297-370,436-472,
 *                                             src/Global/RendererMesh.cu:378-391
 *
 * PARITY STATUS
 *   - Host maths (camera, transforms, slerp, quat->Euler, sRGB) and XORWOW: PINNED.
 *     Checked in tests/ against the reference's own header compiled from
 *     /root/reference (oracle/_ref, see oracle/Makefile) and against cuRAND's header.
 *   - Intersection + traversal semantics: PARITY UNPINNED.  In the reference these live in
 *     closed-source NVIDIA OptiX 9.0.0 (optixAccelBuild / optixTrace / built-in sphere and
 *     triangle intersectors; call sites src/Global/RendererImpl.cu:56,202,232,295-314 and
 *     shader/Shader.cu:70).  The reference has no tests or golden vectors.  The intersectors
 *     below are OUR definition of "closest hit in (tmin,tmax), two-sided, (t, prim id)
 *     lexicographic tie rule"; the CUDA path must match them bit-exactly.
 *   - Shading (closesthitImpl) follows Shader.cu line by line but cannot be executed against the
 *     reference (OptiX device API) -> restated, unpinned.
 *
 * Floating point contract (shared with the CUDA path): IEEE fp32, round-to-nearest, NO fused
 * multiply-add contraction (build with -ffp-contract=off), IEEE sqrt and division, dot(a,b) =
 * (ax*bx + ay*by) + az*bz evaluated left to right.  normalize() uses 1/sqrt(len2) in IEEE
 * arithmetic instead of the reference's approximate device rsqrtf (documented <=2 ulp deviation).
 */
#define QUALIFIERS static inline
#include <cuda_runtime.h>
#include <curand_kernel.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "rt_oracle.h"

namespace {

// ---------------------------------------------------------------- constants
// include/Global/DeviceFunctions.cuh:18-20
constexpr float kZero = 1e-6f;
constexpr float kInf = 1e16f;
constexpr float kPi = 3.1415926;

struct V3 {
    float x, y, z;
};
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return a * s; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// DeviceFunctions.cuh:407-409 — (x*x + y*y) + z*z
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float len2(V3 a) { return dot(a, a); }
// DeviceFunctions.cuh:410-416
inline V3 cross(V3 a, V3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// DeviceFunctions.cuh:397-404 (device path; rsqrtf replaced by IEEE 1/sqrt, see header)
inline V3 normalize_dev(V3 a) {
    const float l2 = len2(a);
    if (l2 <= kZero * kZero) return v3(0.0f, 0.0f, 1.0f);
    const float inv = 1.0f / std::sqrt(l2);
    return a * inv;
}
// Same function executed on the HOST by the reference (camera code): CUDA's host rsqrtf is
// (float)(1.0 / sqrt((double)a))  (crt/math_functions.hpp:2953-2956,3333-3336 of CUDA 12.9).
inline V3 normalize_host(V3 a) {
    const float l2 = len2(a);
    if (l2 <= kZero * kZero) return v3(0.0f, 0.0f, 1.0f);
    const float inv = static_cast<float>(1.0 / std::sqrt(static_cast<double>(l2)));
    return a * inv;
}
inline bool finite3(V3 a) { return std::isfinite(a.x) && std::isfinite(a.y) && std::isfinite(a.z); }
// DeviceFunctions.cuh:38-40
inline bool float_equals(float a, float b) { return std::fabs(a - b) < kZero; }

// ---------------------------------------------------------------- RNG (cuRAND XORWOW, host-compiled header)
// DeviceFunctions.cuh:216-222 randomDouble(state,min,max) = min + (max-min)*curand_uniform
inline float rnd(curandState *s, float lo, float hi) { return lo + (hi - lo) * curand_uniform(s); }
// DeviceFunctions.cuh:570-582 randomSpaceVector (cube sample, normalised — kept on purpose)
inline V3 random_space_vector(curandState *s) {
    V3 r;
    float l2;
    do {
        r.x = rnd(s, -1.0f, 1.0f);
        r.y = rnd(s, -1.0f, 1.0f);
        r.z = rnd(s, -1.0f, 1.0f);
        l2 = len2(r);
    } while (l2 < kZero * kZero);
    r = normalize_dev(r);
    return r * 1.0f;
}

// ---------------------------------------------------------------- intersection (our definition, see header)
struct Ray {
    V3 o, d;
    float tmin, tmax;
    float inv_a;  // 1 / dot(d,d)
};
inline Ray make_ray(V3 o, V3 d, float tmin, float tmax) {
    Ray r{o, d, tmin, tmax, 0.0f};
    r.inv_a = 1.0f / dot(d, d);
    return r;
}

// Robust ray/sphere (Haines et al., Ray Tracing Gems ch.7).  Returns t or -1.
// self: the ray starts on this sphere -> drop the root nearer to zero.
inline float hit_sphere(const Ray &r, const float *s, bool self) {
    const V3 c = v3(s[0], s[1], s[2]);
    const float rad = s[3];
    const V3 f = r.o - c;
    const float bp = -dot(f, r.d);
    const float k = bp * r.inv_a;
    const V3 l = f + r.d * k;
    const float disc = rad * rad - dot(l, l);
    if (!(disc >= 0.0f)) return -1.0f;
    const float cc = dot(f, f) - rad * rad;
    const float a = dot(r.d, r.d);
    const float sq = std::sqrt(a * disc);
    const float q = (bp >= 0.0f) ? (bp + sq) : (bp - sq);
    const float t_cq = cc / q;
    const float t_qa = q * r.inv_a;
    const float t_near = (bp >= 0.0f) ? t_cq : t_qa;
    const float t_far = (bp >= 0.0f) ? t_qa : t_cq;
    if (self) {
        const float t = (std::fabs(t_near) <= std::fabs(t_far)) ? t_far : t_near;
        return (t > r.tmin && t < r.tmax) ? t : -1.0f;
    }
    if (t_near > r.tmin && t_near < r.tmax) return t_near;
    if (t_far > r.tmin && t_far < r.tmax) return t_far;
    return -1.0f;
}

// Moeller-Trumbore, two-sided, no culling.  Returns t or -1, barycentrics (u,v) for v1,v2.
inline float hit_triangle(const Ray &r, const float *tv, float *u_out, float *v_out) {
    const V3 v0 = v3(tv[0], tv[1], tv[2]);
    const V3 e1 = v3(tv[3], tv[4], tv[5]) - v0;
    const V3 e2 = v3(tv[6], tv[7], tv[8]) - v0;
    const V3 p = cross(r.d, e2);
    const float det = dot(e1, p);
    if (!(det != 0.0f)) return -1.0f;
    const float inv = 1.0f / det;
    const V3 tvec = r.o - v0;
    const float u = dot(tvec, p) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return -1.0f;
    const V3 q = cross(tvec, e1);
    const float v = dot(r.d, q) * inv;
    if (!(v >= 0.0f && u + v <= 1.0f)) return -1.0f;
    const float t = dot(e2, q) * inv;
    if (!(t > r.tmin && t < r.tmax)) return -1.0f;
    *u_out = u;
    *v_out = v;
    return t;
}

struct Hit {
    float t;
    uint32_t id;
    float u, v;
};

inline void test_prim(const orc_scene *sc, const Ray &r, uint32_t gid, uint32_t skip, Hit &best) {
    const uint32_t ki = sc->kind_index[gid];
    const uint32_t idx = ki & 0x7FFFFFFFu;
    float t, u = 0.0f, v = 0.0f;
    if ((ki >> 31) == ORC_KIND_SPHERE) {
        t = hit_sphere(r, sc->spheres + 4 * (size_t)idx, gid == skip);
    } else {
        if (gid == skip) return;
        t = hit_triangle(r, sc->tri_verts + 9 * (size_t)idx, &u, &v);
    }
    if (t < 0.0f) return;
    // closest hit, ties broken by smaller global prim id
    if (t < best.t || (t == best.t && gid < best.id)) {
        best.t = t;
        best.id = gid;
        best.u = u;
        best.v = v;
    }
}

// ---------------------------------------------------------------- oracle BVH (median split; conservative boxes)
struct Box {
    float lo[3], hi[3];
};
struct BNode {
    Box box;
    uint32_t left, right;  // internal: children; leaf: first, count | 0x80000000
};
struct Bvh {
    std::vector<BNode> nodes;
    std::vector<uint32_t> order;
};

inline Box prim_box(const orc_scene *sc, uint32_t gid) {
    const uint32_t ki = sc->kind_index[gid];
    const uint32_t idx = ki & 0x7FFFFFFFu;
    Box b;
    if ((ki >> 31) == ORC_KIND_SPHERE) {
        const float *s = sc->spheres + 4 * (size_t)idx;
        for (int a = 0; a < 3; a++) {
            b.lo[a] = s[a] - s[3];
            b.hi[a] = s[a] + s[3];
        }
    } else {
        const float *t = sc->tri_verts + 9 * (size_t)idx;
        for (int a = 0; a < 3; a++) {
            b.lo[a] = std::min(t[a], std::min(t[3 + a], t[6 + a]));
            b.hi[a] = std::max(t[a], std::max(t[3 + a], t[6 + a]));
        }
    }
    // pad so that fp32 rounding in the slab test can never cull a valid hit
    for (int a = 0; a < 3; a++) {
        const float m = std::max(std::fabs(b.lo[a]), std::fabs(b.hi[a]));
        const float pad = m * 3.0517578125e-5f + 1e-30f;  // 2^-15 relative
        b.lo[a] -= pad;
        b.hi[a] += pad;
    }
    return b;
}

void build_rec(const orc_scene *sc, Bvh &bvh, std::vector<Box> &boxes, std::vector<V3> &cent,
               uint32_t node, uint32_t first, uint32_t count) {
    Box b;
    for (int a = 0; a < 3; a++) {
        b.lo[a] = INFINITY;
        b.hi[a] = -INFINITY;
    }
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = first; i < first + count; i++) {
        const uint32_t g = bvh.order[i];
        const float c[3] = {cent[g].x, cent[g].y, cent[g].z};
        for (int a = 0; a < 3; a++) {
            b.lo[a] = std::min(b.lo[a], boxes[g].lo[a]);
            b.hi[a] = std::max(b.hi[a], boxes[g].hi[a]);
            clo[a] = std::min(clo[a], c[a]);
            chi[a] = std::max(chi[a], c[a]);
        }
    }
    bvh.nodes[node].box = b;
    if (count <= 4) {
        bvh.nodes[node].left = first;
        bvh.nodes[node].right = count | 0x80000000u;
        return;
    }
    int axis = 0;
    if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
    if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
    const uint32_t mid = count / 2;
    auto key = [&](uint32_t g) { return axis == 0 ? cent[g].x : (axis == 1 ? cent[g].y : cent[g].z); };
    std::nth_element(bvh.order.begin() + first, bvh.order.begin() + first + mid,
                     bvh.order.begin() + first + count,
                     [&](uint32_t a, uint32_t c) { return key(a) < key(c); });
    const uint32_t l = (uint32_t)bvh.nodes.size();
    bvh.nodes.push_back(BNode{});
    bvh.nodes.push_back(BNode{});
    bvh.nodes[node].left = l;
    bvh.nodes[node].right = l + 1;
    build_rec(sc, bvh, boxes, cent, l, first, mid);
    build_rec(sc, bvh, boxes, cent, l + 1, first + mid, count - mid);
}

// Prims whose box is enormous relative to the rest (the r=1000 ground sphere of files/config.json)
// are tested linearly so they do not wreck the tree.
struct Accel {
    Bvh bvh;
    std::vector<uint32_t> linear;
};

void build_accel(const orc_scene *sc, Accel &ac) {
    const uint32_t n = sc->n_prims;
    std::vector<Box> boxes(n);
    std::vector<V3> cent(n);
    std::vector<float> ext(n);
    for (uint32_t g = 0; g < n; g++) {
        boxes[g] = prim_box(sc, g);
        cent[g] = v3(0.5f * (boxes[g].lo[0] + boxes[g].hi[0]), 0.5f * (boxes[g].lo[1] + boxes[g].hi[1]),
                     0.5f * (boxes[g].lo[2] + boxes[g].hi[2]));
        ext[g] = std::max(boxes[g].hi[0] - boxes[g].lo[0],
                          std::max(boxes[g].hi[1] - boxes[g].lo[1], boxes[g].hi[2] - boxes[g].lo[2]));
    }
    std::vector<float> sorted(ext);
    float med = 0.0f;
    if (n) {
        std::nth_element(sorted.begin(), sorted.begin() + n / 2, sorted.end());
        med = sorted[n / 2];
    }
    ac.bvh.order.clear();
    for (uint32_t g = 0; g < n; g++) {
        if (n > 8 && ext[g] > 64.0f * med && ac.linear.size() < 64)
            ac.linear.push_back(g);
        else
            ac.bvh.order.push_back(g);
    }
    ac.bvh.nodes.clear();
    if (!ac.bvh.order.empty()) {
        ac.bvh.nodes.reserve(ac.bvh.order.size());
        ac.bvh.nodes.push_back(BNode{});
        build_rec(sc, ac.bvh, boxes, cent, 0, 0, (uint32_t)ac.bvh.order.size());
    }
}

// Slab test returning the entry distance (or -1 when the box is missed).  Boxes are padded by 2^-15
// relative per coordinate (prim_box) and the far bound is widened by 2e-6 relative, which together
// dominate every fp32 rounding error of this test and of the primitive intersectors: the BVH can
// never cull a hit that brute force would find.  Axes with d == 0 are handled exactly (no 0*inf).
inline float slab(const Box &b, const float o[3], const float d[3], const float inv[3], float tmin, float tfar) {
    float t0 = tmin, t1 = tfar;
    for (int a = 0; a < 3; a++) {
        if (d[a] == 0.0f) {
            if (o[a] < b.lo[a] || o[a] > b.hi[a]) return -1.0f;
            continue;
        }
        float ta = (b.lo[a] - o[a]) * inv[a];
        float tb = (b.hi[a] - o[a]) * inv[a];
        if (ta > tb) { const float t = ta; ta = tb; tb = t; }
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
    }
    return (t0 <= t1 * 1.000002f) ? t0 : -1.0f;
}

inline Hit closest_hit(const orc_scene *sc, const Accel *ac, const Ray &r, uint32_t skip) {
    Hit best{INFINITY, ORC_MISS, 0.0f, 0.0f};
    if (!ac) {
        for (uint32_t g = 0; g < sc->n_prims; g++) test_prim(sc, r, g, skip, best);
        return best;
    }
    for (uint32_t g : ac->linear) test_prim(sc, r, g, skip, best);
    if (ac->bvh.nodes.empty()) return best;
    const float o[3] = {r.o.x, r.o.y, r.o.z};
    const float d[3] = {r.d.x, r.d.y, r.d.z};
    const float inv[3] = {1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z};
    const BNode *nodes = ac->bvh.nodes.data();
    struct Entry {
        uint32_t node;
        float tnear;
    };
    Entry stack[128];
    int sp = 0;
    if (slab(nodes[0].box, o, d, inv, r.tmin, r.tmax) < 0.0f) return best;
    stack[sp++] = Entry{0, 0.0f};
    while (sp) {
        const Entry e = stack[--sp];
        if (e.tnear > best.t * 1.000002f) continue;  // culled by a hit found since it was pushed
        const BNode &n = nodes[e.node];
        if (n.right & 0x80000000u) {
            const uint32_t cnt = n.right & 0x7FFFFFFFu;
            for (uint32_t i = 0; i < cnt; i++) test_prim(sc, r, ac->bvh.order[n.left + i], skip, best);
            continue;
        }
        const float tfar = std::min(r.tmax, best.t);
        const float tl = slab(nodes[n.left].box, o, d, inv, r.tmin, tfar);
        const float tr = slab(nodes[n.right].box, o, d, inv, r.tmin, tfar);
        if (tl >= 0.0f && tr >= 0.0f) {  // near child on top of the stack
            if (tl <= tr) {
                stack[sp++] = Entry{n.right, tr};
                stack[sp++] = Entry{n.left, tl};
            } else {
                stack[sp++] = Entry{n.left, tl};
                stack[sp++] = Entry{n.right, tr};
            }
        } else if (tl >= 0.0f) {
            stack[sp++] = Entry{n.left, tl};
        } else if (tr >= 0.0f) {
            stack[sp++] = Entry{n.right, tr};
        }
    }
    return best;
}

// ---------------------------------------------------------------- one path (Shader.cu:94-287 made iterative)
struct Aov {
    float albedo[4], normal[4];
};

// Shader.cu:122-155: shading normal at the hit, flipped to face the incoming ray
inline V3 hit_normal(const orc_scene *sc, const Ray &ray, const Hit &hit, V3 P) {
    const uint32_t ki = sc->kind_index[hit.id];
    const uint32_t idx = ki & 0x7FFFFFFFu;
    if ((ki >> 31) == ORC_KIND_SPHERE) {  // Shader.cu:122-136 (world-space centre: quirk 1 fixed,
        // unless the scene carries the object-space values the reference's shader reads)
        const float *s = (sc->sphere_shade ? sc->sphere_shade : sc->spheres) + 4 * (size_t)idx;
        const V3 on = (P - v3(s[0], s[1], s[2])) / s[3];
        return (dot(ray.d, on) < 0.0f) ? on : -on;
    }
    // Shader.cu:137-155
    const float *n = sc->tri_normals + 9 * (size_t)idx;
    const float u = hit.u, v = hit.v;
    const float w = 1.0f - u - v;
    const V3 nn = w * v3(n[0], n[1], n[2]) + u * v3(n[3], n[4], n[5]) + v * v3(n[6], n[7], n[8]);
    return (dot(ray.d, nn) < 0.0f) ? nn : -nn;
}

// state after the primary hit is shared by all samples of a pixel (the reference has no jitter:
// Shader.cu:249-261 uses the pixel centre), so the primary ray is traced once per pixel.
inline void shade_path(const orc_scene *sc, const Accel *ac, const orc_params *p, const Ray &primary,
                       const Hit &first, curandState *rng, float out_rgb[3],
                       uint64_t *segments) {
    V3 thr = v3(1.0f, 1.0f, 1.0f);
    Ray ray = primary;
    Hit hit = first;
    uint32_t depth = 1;  // Shader.cu:264 payload.w = 1
    const bool self_skip = !(p->flags & ORC_FLAG_NO_SELF_SKIP);
    const bool mul_on_return = (p->flags & ORC_FLAG_MUL_ON_RETURN) != 0;
    V3 albs[256];  // max_depth <= 255 (checked by the caller)
    uint32_t n_alb = 0;
    for (;;) {
        if (hit.id == ORC_MISS) {
            // __miss__missProgram, Shader.cu:276-287; albedo products applied on the way back :236-238
            if (mul_on_return) {  // the reference's order: ((bg * a_n) * a_{n-1}) ... * a_1
                V3 r = v3(p->background[0], p->background[1], p->background[2]);
                for (uint32_t k = n_alb; k-- > 0;) r = v3(r.x * albs[k].x, r.y * albs[k].y, r.z * albs[k].z);
                out_rgb[0] = r.x; out_rgb[1] = r.y; out_rgb[2] = r.z;
                return;
            }
            out_rgb[0] = thr.x * p->background[0];
            out_rgb[1] = thr.y * p->background[1];
            out_rgb[2] = thr.z * p->background[2];
            return;
        }
        if (depth >= p->max_depth) {  // Shader.cu:102-107
            out_rgb[0] = out_rgb[1] = out_rgb[2] = 0.0f;
            return;
        }
        const V3 P = ray.o + hit.t * ray.d;  // Shader.cu:114
        const V3 N = hit_normal(sc, ray, hit, P);
        const uint32_t m = sc->mat[hit.id];
        const uint32_t mi = m & 0x7FFFFFFFu;
        V3 dir, alb;
        if ((m >> 31) == ORC_MAT_ROUGH) {  // Shader.cu:169-179
            dir = N + random_space_vector(rng);
            if (float_equals(len2(dir), kZero * kZero)) dir = N;
            alb = v3(sc->roughs[3 * (size_t)mi], sc->roughs[3 * (size_t)mi + 1], sc->roughs[3 * (size_t)mi + 2]);
        } else {  // Shader.cu:180-192
            const float *mt = sc->metals + 4 * (size_t)mi;
            const V3 vv = ray.d;
            dir = normalize_dev(vv - 2 * dot(vv, N) * N);
            if (mt[3] > 0.0f) dir = dir + mt[3] * random_space_vector(rng);
            alb = v3(mt[0], mt[1], mt[2]);
        }
        // Shader.cu:202-213
        if (!finite3(dir) || len2(dir) <= kZero * kZero) {
            dir = N;
            if (len2(dir) <= kZero * kZero || !finite3(dir)) dir = v3(0.0f, 0.0f, 1.0f);
        }
        // Shader.cu:236-238 — the reference multiplies on the way back up the recursion; we
        // multiply on the way down (same product, rounding order differs by <= depth ulp).
        thr = v3(thr.x * alb.x, thr.y * alb.y, thr.z * alb.z);
        albs[n_alb++] = alb;
        const uint32_t skip = self_skip ? hit.id : ORC_MISS;
        ray = make_ray(P, dir, kZero, kInf);  // Shader.cu:231-233
        depth += 1;
        hit = closest_hit(sc, ac, ray, skip);
        if (segments) *segments += 1;
    }
}

inline Ray primary_ray(const orc_camera *cam, const orc_params *p, uint32_t x, uint32_t y) {
    // Shader.cu:249-261
    const float nx = ((static_cast<float>(x) + 0.5f) / static_cast<float>(p->width)) * 2.0f - 1.0f;
    const float ny = ((static_cast<float>(y) + 0.5f) / static_cast<float>(p->height)) * 2.0f - 1.0f;
    const float aspect = static_cast<float>(p->width) / static_cast<float>(p->height);
    const V3 U = v3(cam->U[0], cam->U[1], cam->U[2]);
    const V3 V = v3(cam->V[0], cam->V[1], cam->V[2]);
    const V3 W = v3(cam->W[0], cam->W[1], cam->W[2]);
    const V3 d = normalize_dev(nx * aspect * U + ny * V + W);
    return make_ray(v3(cam->center[0], cam->center[1], cam->center[2]), d, kZero, kInf);
}

// ---------------------------------------------------------------- 5x5 one-indexed matrix of the reference
// DeviceFunctions.cuh:43-148
struct M5 {
    float d[5][5];
    size_t row, col;
};
inline M5 mul(const M5 &a, const M5 &b) {
    M5 r{};
    r.row = a.row;
    r.col = b.col;
    for (size_t i = 1; i <= r.row; i++)
        for (size_t j = 1; j <= r.col; j++) {
            float s = 0.0f;
            for (size_t n = 1; n <= a.col; n++) s += a.d[i][n] * b.d[n][j];
            r.d[i][j] = s;
        }
    return r;
}
inline M5 ident() {
    M5 m{};
    m.row = m.col = 4;
    for (int i = 1; i <= 4; i++) m.d[i][i] = 1.0f;
    return m;
}
inline float deg2rad(float d) { return d * kPi / 180.0f; }
inline float rad2deg(float r) { return r * 180.0f / kPi; }
inline M5 rot_axis(float degree, int axis) {
    const float th = deg2rad(degree);
    const float c = std::cos(th), s = std::sin(th);
    M5 m = ident();
    switch (axis) {
        case 0: m.d[2][2] = c; m.d[2][3] = -s; m.d[3][2] = s; m.d[3][3] = c; break;
        case 1: m.d[1][1] = c; m.d[1][3] = s; m.d[3][1] = -s; m.d[3][3] = c; break;
        default: m.d[1][1] = c; m.d[1][2] = -s; m.d[2][1] = s; m.d[2][2] = c; break;
    }
    return m;
}

}  // namespace

// =================================================================== C interface
extern "C" {

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// bench.py's CPU arms: use every host core whatever OMP_NUM_THREADS says (torchrun exports 1)
int orc_hw_threads(void) {
#ifdef _OPENMP
    return omp_get_num_procs();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

void *orc_accel_create(const orc_scene *sc) {
    Accel *a = new Accel;
    build_accel(sc, *a);
    return a;
}
void orc_accel_destroy(void *h) { delete static_cast<Accel *>(h); }

static int render_impl(const orc_scene *sc, const Accel *ac, const orc_camera *cam, const orc_params *p, float *accum,
                       float *albedo, float *normal, uint32_t *prim_id, uint64_t *segments);

int orc_render_accel(const orc_scene *sc, const void *accel_handle, const orc_camera *cam, const orc_params *p,
                     float *accum, float *albedo, float *normal, uint32_t *prim_id, uint64_t *segments) {
    if (!sc || !cam || !p || p->width == 0 || p->height == 0 || p->max_depth > 255) return 1;
    return render_impl(sc, static_cast<const Accel *>(accel_handle), cam, p, accum, albedo, normal, prim_id, segments);
}

int orc_render(const orc_scene *sc, const orc_camera *cam, const orc_params *p, int accel, float *accum,
               float *albedo, float *normal, uint32_t *prim_id, uint64_t *segments) {
    if (!sc || !cam || !p || p->width == 0 || p->height == 0 || p->max_depth > 255) return 1;
    Accel acs;
    const Accel *ac = nullptr;
    if (accel) {
        build_accel(sc, acs);
        ac = &acs;
    }
    return render_impl(sc, ac, cam, p, accum, albedo, normal, prim_id, segments);
}

static int render_impl(const orc_scene *sc, const Accel *ac, const orc_camera *cam, const orc_params *p, float *accum,
                       float *albedo, float *normal, uint32_t *prim_id, uint64_t *segments) {
    uint64_t total_segments = 0;
    const long H = (long)p->height;
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : total_segments)
    for (long y = 0; y < H; y++) {
        for (uint32_t x = 0; x < p->width; x++) {
            const size_t pix = (size_t)y * p->width + x;  // Shader.cu:270
            const Ray pr = primary_ray(cam, p, x, (uint32_t)y);
            const Hit first = closest_hit(sc, ac, pr, ORC_MISS);
            total_segments += 1;
            if (prim_id) prim_id[pix] = first.id;
            Aov aov{};
            if (first.id != ORC_MISS) {  // Shader.cu:216-227 (real first-hit AOVs: quirk 9 fixed)
                const V3 P = pr.o + first.t * pr.d;
                const V3 nn = normalize_dev(hit_normal(sc, pr, first, P));
                const uint32_t m = sc->mat[first.id];
                const uint32_t mi = m & 0x7FFFFFFFu;
                const float *al = (m >> 31) == ORC_MAT_ROUGH ? sc->roughs + 3 * (size_t)mi : sc->metals + 4 * (size_t)mi;
                aov.albedo[0] = al[0]; aov.albedo[1] = al[1]; aov.albedo[2] = al[2]; aov.albedo[3] = 1.0f;
                aov.normal[0] = nn.x; aov.normal[1] = nn.y; aov.normal[2] = nn.z; aov.normal[3] = 0.0f;
            }
            // stream(pixel, sample) = curand_init(seed, pixel, (sample_offset + s) * 32):
            // generator/subsequence per pixel as HostFunctions.cu:122-127, seed fixed (quirk 7).
            curandState base;
            curand_init(p->seed, pix, (unsigned long long)p->sample_offset * ORC_RNG_SAMPLE_STRIDE, &base);
            float acc[3] = {0.0f, 0.0f, 0.0f};
            for (uint32_t s = 0; s < p->spp; s++) {
                curandState rng = base;
                float rgb[3];
                uint64_t seg = 0;
                shade_path(sc, ac, p, pr, first, &rng, rgb, &seg);
                total_segments += seg;
                acc[0] += rgb[0];
                acc[1] += rgb[1];
                acc[2] += rgb[2];
                for (uint32_t k = 0; k < ORC_RNG_SAMPLE_STRIDE; k++) (void)curand(&base);
            }
            if (accum) {
                accum[4 * pix + 0] = acc[0];
                accum[4 * pix + 1] = acc[1];
                accum[4 * pix + 2] = acc[2];
                accum[4 * pix + 3] = (float)p->spp;
            }
            if (albedo) std::memcpy(albedo + 4 * pix, aov.albedo, 16);
            if (normal) std::memcpy(normal + 4 * pix, aov.normal, 16);
        }
    }
    if (segments) *segments = total_segments;
    return 0;
}

int orc_trace(const orc_scene *sc, int accel, size_t n_rays, const float *rays, const uint32_t *skip,
              float *out_t, uint32_t *out_id) {
    Accel acs;
    const Accel *ac = nullptr;
    if (accel) {
        build_accel(sc, acs);
        ac = &acs;
    }
#pragma omp parallel for schedule(dynamic, 256)
    for (long i = 0; i < (long)n_rays; i++) {
        const float *r = rays + 8 * i;
        const Ray ray = make_ray(v3(r[0], r[1], r[2]), v3(r[4], r[5], r[6]), r[3], r[7]);
        const Hit h = closest_hit(sc, ac, ray, skip ? skip[i] : ORC_MISS);
        out_t[i] = h.t;
        out_id[i] = h.id;
    }
    return 0;
}

int orc_closest_hit(const orc_scene *sc, const void *accel, const float o[3], const float d[3], float tmin, float tmax,
                    uint32_t skip, float *t, uint32_t *id, float *u, float *v) {
    const Ray ray = make_ray(v3(o[0], o[1], o[2]), v3(d[0], d[1], d[2]), tmin, tmax);
    const Hit h = closest_hit(sc, static_cast<const Accel *>(accel), ray, skip);
    *t = h.t; *id = h.id; *u = h.u; *v = h.v;
    return h.id != ORC_MISS;
}

// src/GraphicsAPI/SDL_GraphicsWindow.cu:4-14 (host code: host rsqrtf semantics)
void orc_camera_configure(const float center[3], const float target[3], const float up[3], int flip_up,
                          orc_camera *out) {
    V3 upn = normalize_host(v3(up[0], up[1], up[2]));
    if (flip_up) upn = -upn;
    const V3 c = v3(center[0], center[1], center[2]);
    const V3 W = v3(target[0], target[1], target[2]) - c;
    const V3 U = normalize_host(cross(W, upn));
    const V3 V = normalize_host(cross(U, W));
    out->center[0] = c.x; out->center[1] = c.y; out->center[2] = c.z;
    out->U[0] = U.x; out->U[1] = U.y; out->U[2] = U.z;
    out->V[0] = V.x; out->V[1] = V.y; out->V[2] = V.z;
    out->W[0] = W.x; out->W[1] = W.y; out->W[2] = W.z;
}

// DeviceFunctions.cuh:133-148: transform = Shift * (Rx*Ry*Rz) * Scale, first three rows
void orc_construct_transform(const float shift[3], const float rot[3], const float scale[3], float out[12]) {
    M5 s = ident();
    s.d[1][4] = shift[0]; s.d[2][4] = shift[1]; s.d[3][4] = shift[2];
    const M5 r = mul(mul(rot_axis(rot[0], 0), rot_axis(rot[1], 1)), rot_axis(rot[2], 2));
    M5 sc = ident();
    sc.d[1][1] = scale[0]; sc.d[2][2] = scale[1]; sc.d[3][3] = scale[2];
    const M5 t = mul(mul(s, r), sc);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 4; j++) out[i * 4 + j] = t.d[i + 1][j + 1];
}

// src/Global/RendererTime.cu:297-340.  q = {x,y,z,w} as float4; the reference returns its result
// positionally as {w-expr, x-expr, y-expr, z-expr} into a float4 {x,y,z,w} (quirk 4) — kept.
void orc_slerp(const float q1[4], const float q2in[4], float t, float out[4]) {
    float q2[4] = {q2in[0], q2in[1], q2in[2], q2in[3]};
    float d = q1[3] * q2[3] + q1[0] * q2[0] + q1[1] * q2[1] + q1[2] * q2[2];
    if (d < 0.0f) {
        for (float &c : q2) c = -c;
        d = -d;
    }
    if (d > 0.9995f) {
        // result.{x,y,z,w} <- {w,x,y,z} expressions
        float r[4] = {q1[3] + t * (q2[3] - q1[3]), q1[0] + t * (q2[0] - q1[0]),
                      q1[1] + t * (q2[1] - q1[1]), q1[2] + t * (q2[2] - q1[2])};
        const float mag = std::sqrt(r[3] * r[3] + r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
        if (mag > 0.0) {
            r[3] /= mag; r[0] /= mag; r[1] /= mag; r[2] /= mag;
        }
        std::memcpy(out, r, 16);
        return;
    }
    const float th0 = std::acos(d);
    const float th = th0 * t;
    const float sth = std::sin(th);
    const float sth0 = std::sin(th0);
    const float s0 = std::cos(th) - d * sth / sth0;
    const float s1 = sth / sth0;
    out[0] = (s0 * q1[3]) + (s1 * q2[3]);
    out[1] = (s0 * q1[0]) + (s1 * q2[0]);
    out[2] = (s0 * q1[1]) + (s1 * q2[1]);
    out[3] = (s0 * q1[2]) + (s1 * q2[2]);
}

// src/Global/RendererTime.cu:343-370
void orc_quat_to_euler(const float q[4], float out[3]) {
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float sinr_cosp = 2 * (w * x + y * z);
    const float cosr_cosp = 1 - 2 * (x * x + y * y);
    float ax = std::atan2(sinr_cosp, cosr_cosp);
    const float sinp = 2 * (w * y - z * x);
    float ay;
    if (std::abs(sinp) >= 1)
        ay = std::copysign(kPi / 2, sinp);
    else
        ay = std::asin(sinp);
    const float siny_cosp = 2 * (w * z + x * y);
    const float cosy_cosp = 1 - 2 * (y * y + z * z);
    float az = std::atan2(siny_cosp, cosy_cosp);
    out[0] = rad2deg(ax);
    out[1] = rad2deg(ay);
    out[2] = rad2deg(az);
}

// src/Global/RendererTime.cu:436-472
void orc_transform_time(const float pos[3], const float vel[3], const float qc[4], const float qn[4],
                        float duration, size_t frames, size_t frame, const float pshift[3],
                        const float pscale[3], float out[12]) {
    float factor = 1.0f;
    if (frames > 1) factor = static_cast<float>(frame) / static_cast<float>(frames - 1);
    const V3 total = v3(vel[0], vel[1], vel[2]) * duration;
    const V3 per = total / static_cast<float>(frames);
    const V3 sh = v3(pos[0], pos[1], pos[2]) + per * static_cast<float>(frame);
    float q[4], e[3];
    orc_slerp(qc, qn, factor, q);
    orc_quat_to_euler(q, e);
    const V3 full = v3(pshift[0], pshift[1], pshift[2]) + sh;
    const float s[3] = {full.x, full.y, full.z};
    orc_construct_transform(s, e, pscale, out);
}

// src/Global/RendererMesh.cu:378-391
void orc_transform_mesh(const float vel[3], float duration, size_t frames, size_t frame,
                        const float pshift[3], const float pscale[3], float out[12]) {
    const V3 total = v3(vel[0], vel[1], vel[2]) * duration;
    const V3 per = total / static_cast<float>(frames);
    const V3 full = v3(pshift[0], pshift[1], pshift[2]) + per * static_cast<float>(frame);
    const float s[3] = {full.x, full.y, full.z};
    const float r[3] = {0.0f, 0.0f, 0.0f};
    orc_construct_transform(s, r, pscale, out);
}

// DeviceFunctions.cuh:188-212
void orc_color_to_float4(const float c[3], float out[4]) {
    for (int i = 0; i < 3; i++) {
        const float cx = std::max(0.0f, std::min(c[i], 1.0f));
        constexpr float invGamma = 1.0f / 2.4f;
        const float px = powf(cx, invGamma);
        const float sx = cx < 0.0031308f ? 12.92f * cx : 1.055f * px - 0.055f;
        out[i] = std::max(0.0f, std::min(sx, 1.0f));
    }
    out[3] = 1.0f;
}
// DeviceFunctions.cuh:153-186
void orc_color_to_uchar4(const float c[3], uint8_t out[4]) {
    float f[4];
    orc_color_to_float4(c, f);
    for (int i = 0; i < 3; i++)
        out[i] = (uint8_t)std::min((unsigned int)(f[i] * 256.0f), 255u);
    out[3] = 255;
}

void orc_xorwow_uniforms(uint64_t seed, uint64_t subsequence, uint64_t offset, size_t n, float *out) {
    curandState s;
    curand_init(seed, subsequence, offset, &s);
    for (size_t i = 0; i < n; i++) out[i] = curand_uniform(&s);
}
void orc_xorwow_state(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t out6[6]) {
    curandState s;
    curand_init(seed, subsequence, offset, &s);
    out6[0] = s.d;
    for (int i = 0; i < 5; i++) out6[1 + i] = s.v[i];
}

// Instance flattening (spec A.1): p' = ((m0*x + m1*y) + m2*z) + m3, row-major 3x4.
void orc_xform_point(const float m[12], const float p[3], float out[3]) {
    for (int r = 0; r < 3; r++)
        out[r] = ((m[4 * r] * p[0] + m[4 * r + 1] * p[1]) + m[4 * r + 2] * p[2]) + m[4 * r + 3];
}
static bool linear_is_identity(const float m[12]) {
    return m[0] == 1.0f && m[1] == 0.0f && m[2] == 0.0f && m[4] == 0.0f && m[5] == 1.0f && m[6] == 0.0f &&
           m[8] == 0.0f && m[9] == 0.0f && m[10] == 1.0f;
}
void orc_xform_sphere(const float m[12], const float c[3], float r, float out4[4]) {
    orc_xform_point(m, c, out4);
    if (linear_is_identity(m)) {
        out4[3] = r;
    } else {
        const float s = std::sqrt((m[0] * m[0] + m[4] * m[4]) + m[8] * m[8]);
        out4[3] = r * s;
    }
}
void orc_xform_normal(const float m[12], const float n[3], float out[3]) {
    if (linear_is_identity(m)) {
        out[0] = n[0]; out[1] = n[1]; out[2] = n[2];
        return;
    }
    V3 v;
    v.x = (m[0] * n[0] + m[1] * n[1]) + m[2] * n[2];
    v.y = (m[4] * n[0] + m[5] * n[1]) + m[6] * n[2];
    v.z = (m[8] * n[0] + m[9] * n[1]) + m[10] * n[2];
    v = normalize_dev(v);
    out[0] = v.x; out[1] = v.y; out[2] = v.z;
}

static inline uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
uint32_t orc_morton30(float x, float y, float z) {
    const float fx = std::min(std::max(x * 1024.0f, 0.0f), 1023.0f);
    const float fy = std::min(std::max(y * 1024.0f, 0.0f), 1023.0f);
    const float fz = std::min(std::max(z * 1024.0f, 0.0f), 1023.0f);
    return (expand10((uint32_t)fx) << 2) | (expand10((uint32_t)fy) << 1) | expand10((uint32_t)fz);
}

}  // extern "C"
