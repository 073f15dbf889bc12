// rt_internal.h — host-side structures shared by the C-ABI implementation files.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rtb200.h"

namespace rtb {

// ------------------------------------------------------------------ device-visible records
struct GroupDev {
    const float *a;  // spheres: float3 centers[n]    triangles: float3 vertices[3n]
    const float *b;  // spheres: float radii[n]       triangles: float3 normals[3n]
    uint32_t count;
    uint32_t kind;  // 0 sphere, 1 triangle
};

struct InstDev {
    float m[12];         // row-major 3x4 (OptixInstance.transform)
    uint32_t group;
    uint32_t mat;        // (mat_type << 31) | mat_index
    uint32_t mat_stride;
    uint32_t prim_base;  // global id of this instance's first primitive
    uint32_t tri_base;   // index of its first triangle in the world triangle arrays
    uint32_t identity;   // linear part is exactly I (spheres keep r, normals untouched)
    uint32_t pad0, pad1;
};
static_assert(sizeof(InstDev) == 80, "InstDev layout");

// 64-byte BVH2 node: both children's boxes live in the parent (one 4x128-bit fetch per visit).
//   n0 = c0.lo.x c0.hi.x c0.lo.y c0.hi.y
//   n1 = c1.lo.x c1.hi.x c1.lo.y c1.hi.y
//   n2 = c0.lo.z c0.hi.z c1.lo.z c1.hi.z
//   n3 = child0 child1 (as int bits) ; unused ; unused
// child >= 0: internal node index.  child < 0: leaf, bits 0..26 first slot, bits 27..29 count-1.
struct alignas(16) Node {
    float4 n0, n1, n2, n3;
};

// 80-byte 8-wide node with 8-bit quantised child boxes (layout after Ylitie, Karras, Laine 2017,
// "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs"), collapsed from the
// binary LBVH every frame.  Five 128-bit words:
//   w0 = origin.x origin.y origin.z  bits(ex | ey<<8 | ez<<16 | imask<<24)   (scale_a = 2^(e_a-127))
//   w1 = bits(child_base) bits(prim_base) bits(leafmask) 0
//   w2 = qlo_x[0..3] qlo_x[4..7] qlo_y[0..3] qlo_y[4..7]
//   w3 = qlo_z[0..3] qlo_z[4..7] qhi_x[0..3] qhi_x[4..7]
//   w4 = qhi_y[0..3] qhi_y[4..7] qhi_z[0..3] qhi_z[4..7]
// child box a-axis = origin_a + q * scale_a.  Slot k is an internal child (bit k of imask), a leaf of
// exactly ONE primitive (bit k of leafmask) or empty (neither; its quantised box is inverted).
// Internal children are stored contiguously from child_base in slot order; the leaf primitives of a
// node are contiguous from prim_base in slot order (primitives are re-packed per wide node), so the
// child / primitive of slot k sits at base + popc(mask & ((1 << k) - 1)).
struct alignas(16) WNode {
    float4 w0, w1, w2, w3, w4;
};

constexpr uint32_t kLeafFlag = 0x80000000u;
constexpr uint32_t kLeafFirstMask = 0x07FFFFFFu;
constexpr int kLeafCountShift = 27;
constexpr uint32_t kMaxHuge = 16;
constexpr uint32_t kExtentBuckets = 96;  // floor(log2(extent)) + 64 clamped to [0,95]
constexpr uint32_t kNoSlot = 0xFFFFFFFFu;

// Written by the build kernels, read by later build kernels and by traversal (device memory).
struct BuildInfo {
    float lo[3];          // centroid bounds of the non-huge prims (Morton frame)
    float scale[3];       // 1024 / extent per axis
    uint32_t n_total;
    uint32_t n_bvh;       // prims inside the BVH ( = n_total - n_huge )
    uint32_t n_huge;
    int32_t huge_bucket;  // prims whose extent bucket is > this are "huge"
    float scene_lo[3], scene_hi[3];  // root bounds after refit
    uint32_t huge_slot[kMaxHuge];    // slots (sorted order positions) of huge prims
    float huge_lo[kMaxHuge][3], huge_hi[kMaxHuge][3];  // padded boxes of the huge prims
    uint32_t n_wnodes;               // wide nodes allocated by the collapse (also its task-queue tail)
    uint32_t n_wprims;               // primitives re-packed by the collapse
    uint32_t level_end;              // collapse: end of the current level in the task queue
    uint32_t pad_;
    // Box that no outlier primitive's padded box reaches into (shrunk by a safety margin): a ray
    // segment whose two end points lie inside it cannot hit an outlier, so their tests are skipped.
    float free_lo[3], free_hi[3];
    uint32_t pad2_[2];
};

struct SceneDev {  // everything a traversal needs, passed by value to kernels
    const Node *nodes;          // binary LBVH (traversed when option "wide" is 0)
    const WNode *wnodes;        // 8-wide compressed BVH
    const float4 *slot_geom;    // sphere: c.xyz r   | triangle: v0.xyz, bits(0x80000000|tri index)
    const uint32_t *slot_prim;  // global primitive id per slot
    const uint32_t *slot_mat;   // (mat_type<<31)|index per slot
    const float4 *tri_v12;      // 2 per triangle: v1.xyz _, v2.xyz _
    const float4 *tri_nrm;      // 3 per triangle: world-space vertex normals
    const float4 *roughs;       // rgb _
    const float4 *metals;       // rgb fuzz
    const BuildInfo *info;
    // compat (option "compat_shading" bit 0): per GLOBAL primitive id, the OBJECT-space
    // sphere centre + radius the reference's closest-hit program reads (shader/Shader.cu:124-125); NULL = off
    const float4 *shade_geom;
};

// ------------------------------------------------------------------ host-side helpers
template <typename T>
struct DevBuf {  // grow-only device buffer
    T *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n, size_t *total_bytes) {
        if (n <= cap) return cudaSuccess;
        const size_t had = cap;
        if (p) {
            cudaError_t e = cudaFree(p);
            if (e != cudaSuccess) return e;
            *total_bytes -= cap * sizeof(T);
            p = nullptr;
            cap = 0;
        }
        // first allocation: the request plus a little; a buffer that has to GROW again (per-frame sizes that drift, e.g. the
        // hit-count-sized ray queues of a moving scene) grows by at least half, so that cudaFree + cudaMalloc — a device-wide
        // synchronisation — happens a couple of times, not every frame
        size_t want = n + n / 16 + 64;
        if (had) want = std::max(want, had + had / 2);
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e != cudaSuccess) return e;
        cap = want;
        *total_bytes += cap * sizeof(T);
        return cudaSuccess;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

struct Context {
    int device = 0;
    int sm_count = 148;
    int collapse_blocks = 0;  // co-resident grid of the cooperative collapse kernel (one thread per wide node)
    bool rmq_attr_set = false;  // k_refit_rmq's dynamic shared memory limit has been raised
    int collapse8_blocks = 0; // the same for k_collapse8 (eight lanes per wide node)
    std::mutex mu;
    std::string last_error;
};

struct GroupHost {
    GroupDev d;
};
struct InstHost {
    InstDev d;
};

struct DevTransforms {  // rt_scene_set_transforms_device: float[12*n] on the device, applied at every build
    const float *ptr;
    size_t first, n;
};
struct Scene {
    Context *ctx = nullptr;
    std::vector<GroupDev> groups;
    std::vector<InstDev> insts;
    std::vector<DevTransforms> dev_tf;
    bool topo_dirty = true;   // groups/instances changed -> re-upload tables
    bool built = false;
    uint32_t n_prims = 0, n_tris = 0;
    uint32_t leaf_max = 0;        // 0 = auto: 1 primitive per leaf under the 8-wide tree, 4 under the binary LBVH
    uint32_t batch_spp = 0;       // 0 = auto (render_frame)
    uint32_t auto_batch = 0, auto_batch_npix = 0;
    size_t bytes = 0;
    uint64_t launches = 0;

    // tables
    DevBuf<GroupDev> d_groups;
    DevBuf<InstDev> d_insts;
    DevBuf<uint32_t> d_inst_base;
    DevBuf<float4> d_roughs, d_metals;
    size_t n_rough = 0, n_metal = 0;

    // build
    DevBuf<float4> w_geom;     // per global prim
    DevBuf<uint32_t> w_mat;    // per global prim
    DevBuf<float4> w_shade;    // per global prim, compat_shading bit 0 only: object-space sphere centre + radius
    uint32_t compat = 0;       // option "compat_shading": bit 0 object-space sphere centre (quirk 1), bit 1 unrotated normals (quirk 2)
    DevBuf<float4> tri_v12;    // 2 per tri
    DevBuf<float4> tri_nrm;    // 3 per tri
    DevBuf<uint32_t> keys[2], vals[2];
    DevBuf<uint32_t> sort_hist;  // 256 * blocks
    DevBuf<uint32_t> bucket_tab; // kExtentBuckets * 8
    DevBuf<Node> nodes;
    DevBuf<uint32_t> parent_int, parent_leaf, visit;
    DevBuf<float4> slot_geom;
    DevBuf<uint32_t> slot_prim, slot_mat;
    DevBuf<WNode> wnodes;         // wide BVH + primitives re-packed per wide node
    DevBuf<float4> wslot_geom;
    DevBuf<uint32_t> wslot_prim, wslot_mat;
    bool wide = true;             // traverse the 8-wide BVH (default) or the binary LBVH
    DevBuf<BuildInfo> info;

    // render
    DevBuf<float4> phit;          // per pixel: t, bits(slot), u, v
    DevBuf<uint32_t> rng_cur;     // XORWOW state per pixel, 6 words SoA: [6][npix]; pixel p stands at sample rng_pix_off[p]
    uint64_t rng_seed = 0;
    uint32_t rng_w = 0, rng_h = 0;               // (seed, size) the states belong to
    uint64_t rng_cur_off = ~0ull;                // sample offset after the last launch (what an up-to-date pixel stands at)
    bool rng_valid = false;
    DevBuf<float4> q_o[2], q_d[2], q_t[2];
    DevBuf<uint4> q_r[2];
    DevBuf<uint32_t> q_r4[2];
    DevBuf<uint4> q_m[2];         // RT_FLAG_REF_MUL_ON_RETURN only: material words of the path's bounces 1..4
    DevBuf<float4> result;        // per path
    DevBuf<uint32_t> hit_pix;     // fused form: pixels whose primary ray hit something
    DevBuf<uint32_t> rng_start;   // fused form: [sample][6][npix] XORWOW state of stream(p, s) per hit pixel
    DevBuf<uint32_t> rng_pix_off; // sample offset the state of pixel p in rng_cur stands at (lazy stepping: only pixels that hit draw)
    uint32_t *h_hit = nullptr;    // pinned word the primary-hit count is read back into (compact_queues)
    cudaEvent_t ev_hit = nullptr;
    bool compact_queues = true;   // option "compact_queues": size the wavefront queues from the frame's primary-hit count (one stream sync per rt_render)
    bool fused = false;           // option "fused": one persistent path kernel instead of per-bounce wavefront launches
    uint32_t chunk_spp = 16;      // option "chunk_spp": samples per work unit of the fused kernel
    DevBuf<float4> hit;           // per queue entry: t, bits(slot), u, v (written by k_extend)
    const float4 *ext_phit = nullptr;  // caller-supplied primary hit records (rt_scene_set_primary_hits)
    DevBuf<uint32_t> prim_fetch;        // fetch counter of rt_render_primary
    DevBuf<uint32_t> u_skip;      // rt_trace_rays: skip slot per ray (+1 fetch counter)
    uint32_t t1 = 16, t2 = 4, t3 = 0, fetch_min = 4;  // traversal thresholds (see extend_warp / extend_warp_wide)
    uint32_t extend_blocks_per_sm = 7;
    uint32_t ray_bins = 8, shade_threads = 256;  // queue grouping of scattered rays (rt_render.cu binned_push)
    uint32_t collapse_mode = 0;   // 0: k_collapse (one thread per wide node); 1: k_collapse8 (eight lanes per wide node)
    // samples: a frame further ahead of the per-pixel stream states than this re-creates them (k_rng_init, +1.4 ms at 4K)
    // instead of stepping every hit pixel's state (0.0155 ms per sample of gap at 4K): break-even 92, profiles/r9c_rng_lag.txt
    uint32_t rng_max_lag = 96;
    uint32_t refit_mode = 1;      // 0: one kernel, global counters; 1: k_refit_local (shared-memory climb) + k_refit_top; 2: k_refit_rmq (range queries) + k_refit_top
    DevBuf<uint32_t> counters;    // queue counters
    DevBuf<unsigned long long> seg_total;

    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    bool have_build_ev = false, have_render_ev = false;
    cudaStream_t last_stream = nullptr;
    cudaStream_t aux_stream = nullptr;      // side stream for the RNG (re)positioning (render_frame)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    uint64_t last_primary = 0;
    size_t last_n_cnt = 0;
    bool profile = false;
    std::vector<cudaEvent_t> prof_ev;  // pairs around traversal launches
    size_t prof_used = 0;
};

int set_error(Context *ctx, int code, const char *fmt, ...);

#define RT_CUDA(ctx, expr)                                                                   \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess)                                                               \
            return rtb::set_error((ctx), _e == cudaErrorMemoryAllocation ? RT_ERR_OOM : RT_ERR_CUDA, \
                                  "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

// implemented in rt_build.cu / rt_render.cu
int build_accel(Scene *s, cudaStream_t stream);
int render_frame(Scene *s, const rt_camera *cam, const rt_frame_params *p, const rt_outputs *out,
                 cudaStream_t stream);
int render_primary(Scene *s, const rt_camera *cam, uint32_t W, uint32_t H, size_t first, size_t count, float *hits_out,
                   cudaStream_t stream);
int resolve(Context *ctx, const float *accum, size_t n, float *color, uint8_t *rgba8, cudaStream_t stream);
int resolve_parts(Context *ctx, const float *const *parts, int n_parts, size_t first, size_t n, float *sum_out, float *color,
                  uint8_t *rgba8, cudaStream_t stream);
int trace_rays(Scene *s, const float *rays, const uint32_t *skip, size_t n, float *t, uint32_t *id,
               cudaStream_t stream);
int trace_rays_static(Scene *s, const float *rays, const uint32_t *skip, size_t n, float *t, uint32_t *id,
                      cudaStream_t stream);
SceneDev scene_dev(const Scene *s);
int spheres_advect(Context *ctx, const float *base_c, const float *base_r, const float *vel, size_t n, float duration,
                   size_t frames, size_t frame, const float *pshift, const float *pscale, float *out_c, float *out_r,
                   cudaStream_t stream);

int transforms_time(Context *ctx, const float *pos, const float *vel, const float *quat_cur, const float *quat_next, size_t n,
                    float duration, size_t frames, size_t frame, const float *pshift, const float *pscale, float *out12,
                    cudaStream_t stream);

}  // namespace rtb
