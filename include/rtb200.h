/*
 * rtb200.h — C ABI of the B200-native path tracer that replaces the OptiX seam of
 * 3169651074/nvidia-optix-ray-tracer ("RendererOptiX").
 *
 * The reference has no plugin/FFI surface; the seam we cut is its OptiX wrapper layer
 * (include/Global/RendererImpl.cuh:153-253) plus the optixLaunch call sites
 * (src/Global/RendererTime.cu:484-500, src/Global/RendererMesh.cu:403-419).  Every entry point
 * below names the reference interface it replaces.  Citations are relative to the reference tree.
 *
 * Conventions
 *   - All functions return an int status (RT_OK = 0).  Nothing ever calls exit(); the reference
 *     logs and exit(-200/-300) on any failure (src/Global/HostFunctions.cu:30-50).
 *   - Pointers named dev_* are CUDA device pointers owned by the CALLER and must stay valid until
 *     the scene is destroyed or the group is replaced (same ownership as the reference's
 *     dev_centers/dev_radii/dev_normals, src/Global/RendererImpl.cu:508-509,533).  Pointers named
 *     *_h are host pointers and are copied before the call returns (or stream-ordered if pinned).
 *   - `stream` is a cudaStream_t passed as void*; NULL = the legacy default stream.  All device
 *     work is stream-ordered; no entry point performs a device-wide synchronisation except
 *     rt_get_stats / rt_context_destroy / rt_scene_destroy.
 *   - A context is bound to one CUDA device.  Calls on different scenes may come from different
 *     host threads (the reference builds from worker threads on private streams,
 *     src/Global/RendererMesh.cu:93-100,205-219); calls on the same scene must be serialised.
 *   - There is NO CPU fallback: with no usable CUDA device rt_context_create fails.
 */
#ifndef RTB200_H
#define RTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB200_ABI_VERSION 1

/* status codes */
#define RT_OK 0
#define RT_ERR_INVALID 1     /* bad argument */
#define RT_ERR_CUDA 2        /* CUDA runtime error (see rt_last_error) */
#define RT_ERR_OOM 3         /* device allocation failed */
#define RT_ERR_UNSUPPORTED 4 /* e.g. non-uniform scale on a sphere instance */
#define RT_ERR_STATE 5       /* call order violated (render before build, ...) */

/* material types — include/Global/Shader.cuh:16-18 (MaterialType::ROUGH, METAL) */
#define RT_MAT_ROUGH 0
#define RT_MAT_METAL 1

/* primitive id written for a primary miss */
#define RT_MISS_ID 0xFFFFFFFFu

/* rt_frame_params.flags */
#define RT_FLAG_ACCUMULATE 1u /* add this call's samples to accum_linear instead of overwriting */
/* The two transport details in which the default path deviates from shader/Shader.cu on purpose,
 * switched back to the reference's behaviour.  With both set (and scene option "compat_shading" = 3 for
 * translated / rotated instances) rt_render reproduces the reference's own closest-hit / miss programs
 * bit for bit, given the same closest hits and an IEEE 1/sqrt (tests/test_gpu_refparity.py). */
#define RT_FLAG_REF_NO_SELF_SKIP 2u  /* scattered rays rely on tmin = 1e-6 alone (Shader.cu:230-233); default: a ray
                                        never re-hits the primitive it leaves (triangles) / takes the far root (spheres) */
#define RT_FLAG_REF_MUL_ON_RETURN 4u /* albedo products applied innermost first as the recursion unwinds
                                        (Shader.cu:236-238); default: throughput multiplied on the way down.  max_depth <= 5 */

/* XORWOW draws reserved per (pixel, sample): stream(p, s) = curand_init(seed, p, (offset+s)*32).
 * A path draws at most 3 per bounce, so max_depth <= 11 (the reference's rayTraceDepth is 5,
 * include/Global/Shader.cuh:8); rt_render rejects deeper paths.  Advancing a pixel's state from one
 * sample to the next costs one XORWOW step per reserved draw, hence no larger than needed. */
#define RT_RNG_SAMPLE_STRIDE 32u

typedef struct rt_context_s *rt_context;
typedef struct rt_scene_s *rt_scene;

/* Camera basis — RayGenParams.cameraCenter/U/V/W, include/Global/Shader.cuh:27-35.
 * W = target - center is NOT normalised; U, V are unit (SDL_GraphicsWindow.cu:10-12). */
typedef struct {
    float center[3], U[3], V[3], W[3];
} rt_camera;

/* Per-launch parameters — replaces RayGenParams.width/height, rayTraceDepth
 * (include/Global/Shader.cuh:8) and MissParams.backgroundColor (Shader.cuh:38-40).
 * spp / seed / sample_offset are additive: the reference renders 1 spp per frame and seeds
 * its RNG from clock64() (src/Global/HostFunctions.cu:126). */
typedef struct {
    uint32_t width, height;
    uint32_t spp;           /* samples per pixel rendered by this call */
    uint32_t max_depth;     /* reference constant: 5 */
    uint64_t seed;          /* XORWOW seed (subsequence = pixel index y*width+x) */
    uint32_t sample_offset; /* first sample index of this call (spp split across GPUs / frames) */
    float background[3];    /* reference: 0.7 0.8 0.9 (src/Global/RendererMesh.cu:260-262) */
    uint32_t flags;
} rt_frame_params;

/* Output buffers (device pointers, any may be NULL) — replaces RayGenParams.colorBuffer /
 * albedoBuffer / normalBuffer (float4[w*h], index y*w+x, shader/Shader.cu:270-272).
 * accum_linear holds the SUM over this call's samples of linear radiance, w = sample count
 * (the reference stores one sRGB-encoded sample; encode with rt_resolve). */
typedef struct {
    float *accum_linear; /* float4[w*h] */
    float *albedo;       /* float4[w*h], first-hit albedo, w = 1 on hit */
    float *normal;       /* float4[w*h], first-hit unit normal, w = 0 */
    uint32_t *prim_id;   /* u32[w*h], global primitive id of the primary hit or RT_MISS_ID */
} rt_outputs;

typedef struct {
    uint64_t n_prims;        /* world-space primitives after instance flattening */
    uint64_t n_bvh_prims;    /* primitives inside the LBVH */
    uint64_t n_huge;         /* outlier primitives kept in the linear list */
    uint64_t n_nodes;        /* internal BVH nodes */
    uint64_t device_bytes;   /* bytes owned by the scene on the device */
    float build_ms;          /* device time of the last rt_accel_build (CUDA events) */
    float render_ms;         /* device time of the last rt_render */
    uint64_t segments;       /* path segments traversed by the last rt_render */
    uint64_t kernel_launches;/* kernels launched by this scene since creation */
    float scene_lo[3], scene_hi[3];
    float trace_ms;          /* sum of traversal-kernel durations since the previous rt_get_stats (option "profile") */
    uint32_t trace_launches; /* number of traversal-kernel launches behind trace_ms */
    uint64_t segments_total; /* path segments traversed by this scene since creation */
} rt_stats;

/* ---- context: replaces createContext (RendererImpl.cu:6-28) -------------------------------- */
int rt_context_create(int device, rt_context *out);
int rt_context_destroy(rt_context ctx);
/* last error text of the calling thread's most recent failing call on this context */
const char *rt_last_error(rt_context ctx);
int rt_abi_version(void);

/* ---- scene assembly: replaces RendererSphere/RendererTriangle/RendererParticle geometry records
 *      (RendererImpl.cuh:60-90), buildGASForSpheres/Triangles/Particle (RendererImpl.cu:113-172)
 *      and the OptixInstance fill (RendererTime.cu:116-130, RendererMesh.cu:130-148) ---------- */
int rt_scene_create(rt_context ctx, rt_scene *out);
int rt_scene_destroy(rt_scene scene);
/* Tuning knobs; none of them changes a result except "compat_shading":
 *   "profile"              0/1: time every traversal launch with CUDA events (rt_stats.trace_ms)
 *   "wide"                 1 (default) = traverse the 8-wide compressed BVH, 0 = the binary LBVH
 *   "leaf_max"             primitives per leaf of the binary tree: 1..8, 0 = 4 (leaves of the 8-wide tree hold one primitive)
 *   "batch_spp"            samples per pixel in flight per wavefront batch; 0 = auto: up to 32 within a third of free memory
 *   "compact_queues"       1 (default) = wavefront queues sized from the frame's primary-hit count (rt_render waits for its
 *                          primary-visibility pass, see rt_render); 0 = sized for every pixel, rt_render never blocks
 *   "fused"                0 (default) = wavefront form: one closest-hit launch + one shading launch per bounce over ray
 *                          queues (16.3 GB on the 4K / 10 M-sphere workload); 1 = ONE persistent kernel in which a lane
 *                          follows a path from its first scatter to its end, no queues (14 GB, ~0.9 of the throughput);
 *                          bit-identical images
 *   "chunk_spp"            fused form: samples per work unit
 *   "compat_shading"       bit 0: sphere normals from the OBJECT-space centre and radius as shader/Shader.cu:124-132 does
 *                          although the hit point is in world space, bit 1: triangle vertex normals not rotated with their
 *                          instance as Shader.cu:139-142 (SURVEY C quirks 1 and 2); default 0 = correct world-space shading
 *   "t1" "t2" "t3" "fetch_min"   warp-phase thresholds of the traversal kernel
 *   "extend_blocks_per_sm" persistent grid size of the traversal kernel (5..8, default 7)
 *   "ray_bins"             8 or 24: key of the per-block grouping of scattered rays (octant, or octant x dominant axis)
 *   "shade_threads"        256 (default), 512 or 1024: block size of the per-bounce shading kernel
 *   "rng_max_lag"          0..2048 samples, default 96: a frame whose sample_offset lies further ahead of the scene's per-pixel
 *                          stream states than this re-creates all states in one launch instead of stepping each hit pixel's
 *                          state forward (break-even measured at 4K: 92 samples, profiles/r9c_rng_lag.txt); same samples either way
 *   "refit_mode"           1 (default) = shared-memory local climb + global top kernel, 0 = one kernel with global
 *                          counters, 2 = range queries over a per-block sparse table + the same top kernel
 *   "collapse_mode"        0 (default) = one thread per wide node, 1 = eight lanes per wide node */
int rt_scene_set_option(rt_scene scene, const char *name, int64_t value);

/* Material tables: roughs = float[n][3] albedo; metals = float[m][4] albedo + fuzz
 * (HitGroupParams.rough/metal, Shader.cuh:57-69).  Host or device pointers.
 * Call it BEFORE rt_accel_build: the build checks every instance's material range (mat_index + k*mat_stride) against
 * the tables as they stand at that moment and fails with RT_ERR_INVALID for a scene whose tables are still empty.
 * Replacing the tables afterwards by ones with at least as many entries needs no rebuild (the values are read at
 * render time); a table that gets SMALLER invalidates the build: rt_render returns RT_ERR_STATE until
 * rt_accel_build has run (and re-checked the ranges) again. */
int rt_scene_set_materials(rt_scene scene, const float *roughs, size_t n_rough,
                           const float *metals, size_t n_metal);

/* Geometry groups (object space; the GAS equivalent).  Sphere SoA: float3 centers[n] + float
 * radii[n] (RendererImpl.cu:124-135).  Triangles: non-indexed float3 vertices[3n] and per-vertex
 * float3 normals[3n] (RendererImpl.cu:149-159; shader/Shader.cu:140-142). */
int rt_scene_add_spheres(rt_scene scene, const float *dev_centers, const float *dev_radii,
                         size_t count, uint32_t *group_out);
int rt_scene_add_triangles(rt_scene scene, const float *dev_vertices, const float *dev_normals,
                           size_t count, uint32_t *group_out);
/* Drop every geometry group and instance (materials and options stay): the scene of the next VTK
 * file of a series is assembled from scratch while the device buffers of the accel are reused.  The
 * reference keeps one IAS + SBT per file instead (RendererTime.cu:186-224, RendererMesh.cu:186-222). */
int rt_scene_clear_geometry(rt_scene scene);
/* Re-point an existing group at new buffers / a new count (per-frame particle data). */
int rt_scene_update_group(rt_scene scene, uint32_t group, const float *dev_a, const float *dev_b,
                          size_t count);

/* Instances, in SBT order (config spheres -> config triangles -> particles).  transform_h is the
 * row-major 3x4 of OptixInstance.transform; NULL = identity.  Material of primitive k of the
 * instance = (mat_type, mat_index + k*mat_stride); mat_stride 0 is the reference's one hit record
 * per instance (RendererImpl.cu:472-575), 1 gives one ramp colour per particle of a big group. */
int rt_scene_add_instance(rt_scene scene, uint32_t group, const float *transform_h, int mat_type,
                          uint32_t mat_index, uint32_t mat_stride, uint32_t *instance_out);
/* Per-frame transform update: replaces memcpy into pin_instances + cudaMemcpy H2D
 * (RendererTime.cu:436-480).  transforms = float[12*n], host (pinned or pageable) or device. */
int rt_scene_set_transforms(rt_scene scene, const float *transforms, size_t first_instance,
                            size_t n, void *stream);
/* The same without a host round trip, for particle counts where RendererTime.cu:436-480's per-frame host loop
 * + H2D copy is the bottleneck: dev_transforms = float[12*n] in device memory (e.g. written by
 * rt_transforms_time).  The array is READ BY EVERY FOLLOWING rt_accel_build, on that build's stream, over the
 * uploaded instance table — it must stay allocated, and hold this frame's values in stream order, until a later
 * call for the same instances (either form) or rt_scene_clear_geometry replaces it.  Sphere groups need
 * similarity transforms (uniform scale); that is checked for host-side transforms only. */
int rt_scene_set_transforms_device(rt_scene scene, const float *dev_transforms, size_t first_instance, size_t n);

/* ---- accel: replaces buildASImpl + buildIAS + updateIAS (RendererImpl.cu:30-88,174-242) -----
 * Flattens instances to world space and rebuilds the LBVH (30-bit Morton, radix sort, Karras
 * hierarchy, bottom-up refit).  Called once per frame. */
int rt_accel_build(rt_scene scene, void *stream);

/* ---- per-frame particle update on the device: the translation-only transform loop of Mesh mode
 *      (src/Global/RendererMesh.cu:378-391: shift = particleOffset + (velocity * duration / frames) * frame,
 *      constructTransformMatrix(shift, 0, particleScale)) applied to sphere particles in place of the
 *      reference's O(n) host loop + H2D copy of OptixInstance[] (RendererMesh.cu:393-399).  For
 *      particle i:  center_out = particle_scale * base_center + shift_i,  radius_out = particle_scale[0]
 *      * base_radius (uniform scale required for spheres).  Same operation order as rt_transform_mesh
 *      followed by the instance flattening of rt_accel_build, so the results are bit-identical to the
 *      host path.  All pointers are device pointers (float3 / float per particle), 16-byte aligned
 *      (any cudaMalloc / torch allocation is); outputs may alias nothing else.  Feed the outputs to rt_scene_add_spheres / rt_scene_update_group. */
int rt_spheres_advect(rt_context ctx, const float *dev_base_centers, const float *dev_base_radii,
                      const float *dev_velocities, size_t n, float duration, size_t frames_this_file,
                      size_t frame, const float particle_shift[3], const float particle_scale[3],
                      float *dev_centers_out, float *dev_radii_out, void *stream);

/* Time-mode per-frame instance transforms on the device: replaces the host loop of RendererTime.cu:283-293
 *      (position + velocity * t, slerp :297-340 of the file's two key-frame quaternions, quatToEuler :343-370,
 *      constructTransformMatrix DeviceFunctions.cuh:133-148) for n particles at once.  Inputs are device arrays:
 *      positions / velocities float[3n], quaternions float[4n] {x,y,z,w} (16-byte aligned); output float[12n]
 *      row-major 3x4 (16-byte aligned), ready for rt_scene_set_transforms_device.  Same operation order as
 *      rt_transform_time: the translation column is bit-identical to it; the rotation entries go through the CUDA
 *      libm's sinf/cosf/acosf/asinf/atan2f instead of glibc's and differ by <= 4e-6 absolute (tests/test_gpu_advect.py).
 *      Parity configs (C1-C3) therefore keep rt_transform_time on the host; this is the at-scale path. */
int rt_transforms_time(rt_context ctx, const float *dev_positions, const float *dev_velocities,
                       const float *dev_quat_cur, const float *dev_quat_next, size_t n, float duration,
                       size_t frames_this_file, size_t frame, const float particle_shift[3],
                       const float particle_scale[3], float *dev_transforms_out, void *stream);

/* ---- launch: replaces optixLaunch(pipeline, stream, params, sbt, w, h, 1) and the raygen /
 *      closest-hit / miss programs of shader/Shader.cu.  Stream-ordered.  One host-side wait by default: the ray
 *      queues of the wavefront form are sized from the frame's own primary-hit count, which rt_render reads back (4 bytes)
 *      after the primary-visibility pass it has enqueued, i.e. the call returns once `stream` has reached that point
 *      (build + primary pass; the reference synchronises the whole device after every launch, RendererTime.cu:500).
 *      rt_scene_set_option(scene, "compact_queues", 0) sizes the queues for every pixel instead (3x the device memory
 *      on BASELINE config 4) and never blocks; option "fused" uses no queues at all. ----------------------------- */
int rt_render(rt_scene scene, const rt_camera *camera, const rt_frame_params *params,
              const rt_outputs *out, void *stream);

/* ---- primary visibility for a pixel range (multi-GPU spp split; no counterpart in the reference):
 *      rt_render_primary traces the camera rays (Shader.cu:246-267) of pixels [first_pixel, first_pixel +
 *      n_pixels) and stores one hit record per pixel (float4: t, bits(global primitive id or RT_MISS_ID),
 *      u, v: independent of the rank and of the build) into dev_hits[first_pixel ...] of a
 *      float4[width*height] buffer.
 *      rt_scene_set_primary_hits makes the following rt_render calls take their primary hits from such a
 *      buffer (complete for the whole frame, same camera and size) instead of tracing them; NULL restores
 *      the default.  The image is bit-identical either way. */
int rt_render_primary(rt_scene scene, const rt_camera *camera, uint32_t width, uint32_t height,
                      size_t first_pixel, size_t n_pixels, float *dev_hits, void *stream);
int rt_scene_set_primary_hits(rt_scene scene, const float *dev_hits);

/* ---- display conversion: replaces colorToFloat4 in raygen (Shader.cu:270) and
 *      convertFloat4ToUchar4Kernel (RendererImpl.cu:672-678).  color = accum.rgb / accum.w,
 *      clamp, sRGB OETF.  color_srgb float4[n] and/or rgba8 uchar4[n]; either may be NULL. */
int rt_resolve(rt_context ctx, const float *dev_accum, size_t n_pixels, float *dev_color_srgb,
               uint8_t *dev_rgba8, void *stream);

/* ---- multi-GPU exchange step of the samples-per-pixel split (no counterpart in the reference, which
 *      is single-GPU: docs/technical-details.md:327; north_star names the split and the sum).
 *      One process per GPU.  Each rank renders its block of samples into an accumulation buffer
 *      obtained from rt_peer_alloc, publishes the 64-byte handle (any host transport; the tests and
 *      bench.py use torch.distributed.all_gather) and maps the other ranks' buffers with rt_peer_open
 *      (CUDA IPC: NVLink / NVSwitch peer memory on a multi-GPU box, same-device mapping when two
 *      processes share a GPU).  rt_resolve_parts is then ONE kernel that, for the pixel range
 *      [first_pixel, first_pixel + n_pixels), loads the n_parts partial sums straight from peer
 *      memory, adds them in index order (fixed order => the frame is bit-reproducible and equal to
 *      the same parts summed on one GPU), divides by the total sample count, applies the display
 *      transform of rt_resolve and stores into dev_sum_out / dev_color_srgb / dev_rgba8 (each
 *      float4[/uchar4] per pixel of the WHOLE frame, indexed by absolute pixel, any may be NULL and
 *      any may itself be peer memory, e.g. rank 0's frame buffer).  With every rank resolving its own
 *      slice this is reduce-scatter + resolve + gather in one pass over NVLink.  The caller orders
 *      the ranks (all parts complete before the call; parts not overwritten until every rank's call
 *      has finished), e.g. with a stream-ordered NCCL barrier ENQUEUED ON THE SAME STREAM as this call and as
 *      the rt_render that filled the part (multigpu.PeerExchange refuses any other stream).  n_parts <= RT_MAX_PARTS. */
#define RT_MAX_PARTS 16
#define RT_PEER_HANDLE_BYTES 64
int rt_peer_alloc(rt_context ctx, size_t bytes, void **dev_ptr_out, unsigned char handle_out[RT_PEER_HANDLE_BYTES]);
int rt_peer_free(rt_context ctx, void *dev_ptr);
int rt_peer_open(rt_context ctx, const unsigned char handle[RT_PEER_HANDLE_BYTES], void **dev_ptr_out);
int rt_peer_close(rt_context ctx, void *dev_ptr);
int rt_resolve_parts(rt_context ctx, const float *const *dev_parts, int n_parts, size_t first_pixel,
                     size_t n_pixels, float *dev_sum_out, float *dev_color_srgb, uint8_t *dev_rgba8,
                     void *stream);

int rt_get_stats(rt_scene scene, rt_stats *out); /* synchronises the scene's last stream */

/* Debug/validation export of the built BVH (device -> host copies; synchronises).
 * nodes: 16 floats per node; slot_geom: 4 floats per slot; slot_prim: u32 per slot. Any NULL. */
int rt_accel_download(rt_scene scene, float *nodes_h, float *slot_geom_h, uint32_t *slot_prim_h,
                      uint32_t *sorted_keys_h);
/* Same for the 8-wide compressed BVH that traversal actually walks (option "wide" = 1, the default):
 * wnodes_h: 20 floats (80 bytes) per wide node, layout in DESIGN.md; slot arrays in the re-packed
 * per-wide-node primitive order.  *n_wnodes_out receives the node count.  Any pointer may be NULL;
 * capacity needed: n_prims entries each (query with all-NULL first). */
int rt_accel_download_wide(rt_scene scene, float *wnodes_h, float *slot_geom_h, uint32_t *slot_prim_h,
                           uint64_t *n_wnodes_out);
/* Closest hit for caller-supplied rays through the same traversal kernel (validation):
 * dev_rays float[8*n] = o.xyz tmin d.xyz tmax; dev_skip u32[n] global prim id or RT_MISS_ID
 * (may be NULL); outputs dev_t float[n], dev_id u32[n]. */
int rt_trace_rays(rt_scene scene, const float *dev_rays, const uint32_t *dev_skip, size_t n,
                  float *dev_t, uint32_t *dev_id, void *stream);

/* rt_trace_rays for callers that must not allocate.  No counterpart in the reference: optixTrace is only reachable
 * from inside its pipeline, and the reference's host side allocates whenever it needs to (e.g. RendererImpl.cu:39-41).
 *
 * Who needs it: code that wants closest hits from an already built scene at a point where device allocation is off
 * the table — a frame loop whose memory plan was frozen at start-up, a plugin inside a host application that owns
 * the device allocator, a process sharing the GPU with something that has claimed the remaining memory (cudaMalloc
 * would fail, or would stall every stream on the device while it runs).  rt_trace_rays grows scene-owned buffers
 * with cudaMalloc when a batch is larger than any before it; this call never does.  It performs no cudaMalloc,
 * cudaFree or other allocating call, whatever the batch size.
 *
 * Same arguments, same launches and bit-identical dev_t / dev_id as rt_trace_rays.  What differs is where the
 * intermediate data lives (packed rays, skip slots, hit records, the traversal kernel's fetch counter): in a fixed
 * array of 52 bytes per ray x RT_TRACE_STATIC_MAX_RAYS + 256 = 3.25 MiB that is part of the library image.  Every
 * GPU that loads librtb200.so holds that array until the process exits, used or not.
 *
 * Limits and rules
 *   - n <= RT_TRACE_STATIC_MAX_RAYS, otherwise RT_ERR_INVALID with a message naming the limit (split the batch).
 *   - The scene must be built (rt_accel_build allocates; do that while allocation is still allowed).  With dev_skip
 *     the id -> slot table is written into a buffer the build already owns, as rt_trace_rays does.
 *   - One array per device, shared by all scenes, contexts and threads of the process.  It keeps whatever the previous
 *     call left; the call clears the one word that must start at zero itself.  Calls whose GPU work can overlap
 *     (different streams without an ordering between them) corrupt each other: keep them on one stream or order
 *     them with events.
 *   - The first call on each device looks up the array's device address (once, under a lock, with the context's
 *     device current).  Make that first call — n = 0 is enough — before starting any stream capture.
 *   - Option "profile" is ignored here (its event pool grows on demand). */
#define RT_TRACE_STATIC_MAX_RAYS 65536
int rt_trace_rays_static(rt_scene scene, const float *dev_rays, const uint32_t *dev_skip, size_t n,
                         float *dev_t, uint32_t *dev_id, void *stream);

/* ---- host-side maths of the path (pure CPU, no CUDA) ----------------------------------------
 * rt_camera_configure: SDL_GraphicsWindowConfigureCamera, src/GraphicsAPI/SDL_GraphicsWindow.cu:4-14
 *   (flip_up != 0 for every API except OpenGL).
 * rt_construct_transform: MathHelper::constructTransformMatrix, DeviceFunctions.cuh:133-148.
 * rt_transform_time: per-particle transform of Time mode, src/Global/RendererTime.cu:436-472
 *   (slerp :297-340, quatToEuler :343-370; reference component order kept).
 * rt_transform_mesh: Mesh mode, src/Global/RendererMesh.cu:378-391. */
void rt_camera_configure(const float center[3], const float target[3], const float up[3],
                         int flip_up, rt_camera *out);
void rt_construct_transform(const float shift[3], const float rotate_deg[3], const float scale[3],
                            float out12[12]);
void rt_transform_time(const float pos[3], const float vel[3], const float quat_cur[4],
                       const float quat_next[4], float duration, size_t frames_this_file,
                       size_t frame, const float particle_shift[3], const float particle_scale[3],
                       float out12[12]);
void rt_transform_mesh(const float vel[3], float duration, size_t frames_this_file, size_t frame,
                       const float particle_shift[3], const float particle_scale[3],
                       float out12[12]);

/* Particle colour ramps -> rough albedo table: resolveColorRampPreset / colorStopsForPreset /
 * bakeColorRamp, include/Util/ColorRamp.cuh:31-152.  preset is case-insensitive (viridis, plasma,
 * spectral, terrain, heatmap, grayscale; anything else = viridis).  out_rgb = float[3*count]. */
int rt_bake_color_ramp(const char *preset, size_t count, float *out_rgb);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_H */
