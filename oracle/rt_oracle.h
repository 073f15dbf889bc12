/*
 * rt_oracle.h — C interface of the CPU ORACLE.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker / the reported CPU baseline.
 *
 * The oracle is a scalar (OpenMP over pixel rows) restatement of the reference's hot
 * path; every function cites the reference file:line it follows (paths relative to the
 * reference checkout).  Parity status per piece is recorded in rt_oracle.cpp's header.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MISS 0xFFFFFFFFu
#define ORC_KIND_SPHERE 0u
#define ORC_KIND_TRIANGLE 1u
#define ORC_MAT_ROUGH 0u
#define ORC_MAT_METAL 1u
#define ORC_RNG_SAMPLE_STRIDE 32u
/* orc_params.flags: switch the two transport deviations of DESIGN.md §2 back to the reference's behaviour
 * (used to pin the oracle against the reference's own shader/Shader.cu, oracle/ref_shader.cpp) */
#define ORC_FLAG_NO_SELF_SKIP 1u  /* scattered rays rely on tmin = 1e-6 alone (Shader.cu:230-233) */
#define ORC_FLAG_MUL_ON_RETURN 2u /* albedo products applied innermost first (Shader.cu:236-238)   */

/* World-space (flattened) scene.  Global primitive id = index into kind_index[]. */
typedef struct {
    uint32_t n_prims;
    const uint32_t *kind_index; /* (kind << 31) | index into spheres[] / tris[]            */
    const uint32_t *mat;        /* (mat_type << 31) | index into roughs[] / metals[]       */
    const float *spheres;       /* 4 floats each: cx cy cz r                                */
    const float *tri_verts;     /* 9 floats each: v0 v1 v2                                  */
    const float *tri_normals;   /* 9 floats each: n0 n1 n2 (world space)                    */
    const float *roughs;        /* 3 floats each                                            */
    const float *metals;        /* 4 floats each: r g b fuzz                                */
    uint32_t n_rough, n_metal;
    /* Optional (NULL = spheres): 4 floats per sphere, the centre and radius the SHADER sees.  The
     * reference reads the object-space centre/radius of the hit GAS (shader/Shader.cu:124-125) although
     * the hit point is in world space (quirk 1); a `compat` scene passes the object-space values here. */
    const float *sphere_shade;
} orc_scene;

typedef struct {
    float center[3], U[3], V[3], W[3];
} orc_camera;

typedef struct {
    uint32_t width, height, spp, max_depth;
    uint64_t seed;
    uint32_t sample_offset;
    float background[3];
    uint32_t flags;
} orc_params;

/* accel: 0 = brute force over all prims, 1 = oracle's own median-split BVH.
 * Outputs (any may be NULL): accum = sum over samples of linear radiance (float4, w = spp),
 * albedo / normal = first-hit AOVs (float4), prim_id = primary-hit global prim id or ORC_MISS,
 * segments = number of path segments traced (primary rays counted once per pixel). */
int orc_render(const orc_scene *scene, const orc_camera *cam, const orc_params *p, int accel,
               float *accum, float *albedo, float *normal, uint32_t *prim_id,
               uint64_t *segments);

/* Same with a pre-built oracle BVH (so the CPU baseline can time build and render separately). */
void *orc_accel_create(const orc_scene *scene);
void orc_accel_destroy(void *accel);
int orc_render_accel(const orc_scene *scene, const void *accel, const orc_camera *cam,
                     const orc_params *p, float *accum, float *albedo, float *normal,
                     uint32_t *prim_id, uint64_t *segments);

/* Closest hit of arbitrary rays (for traversal-vs-brute-force property tests).
 * rays: 8 floats each (o.xyz, tmin, d.xyz, tmax); skip[i] = prim id to self-exclude or ORC_MISS.
 * out_t / out_id per ray. */
int orc_trace(const orc_scene *scene, int accel, size_t n_rays, const float *rays,
              const uint32_t *skip, float *out_t, uint32_t *out_id);

/* One closest-hit query (the oracle's stand-in for optixTrace's traversal, shader/Shader.cu:70): used
 * by oracle/ref_shader.cpp to drive the reference's own closest-hit / miss programs.  accel may be
 * NULL (brute force).  Returns 1 on a hit. */
int orc_closest_hit(const orc_scene *scene, const void *accel, const float o[3], const float d[3],
                    float tmin, float tmax, uint32_t skip, float *t, uint32_t *id, float *u, float *v);

/* ---- host-side maths of the path (each mirrors reference host code) ---- */
void orc_camera_configure(const float center[3], const float target[3], const float up[3],
                          int flip_up, orc_camera *out);
void orc_construct_transform(const float shift[3], const float rotate_deg[3],
                             const float scale[3], float out12[12]);
void orc_slerp(const float q1[4], const float q2[4], float t, float out[4]);
void orc_quat_to_euler(const float q[4], float out_deg[3]);
/* Time-mode particle transform (compat chain of the reference) */
void orc_transform_time(const float pos[3], const float vel[3], const float quat_cur[4],
                        const float quat_next[4], float duration, size_t frames_this_file,
                        size_t frame, const float particle_shift[3],
                        const float particle_scale[3], float out12[12]);
void orc_transform_mesh(const float vel[3], float duration, size_t frames_this_file,
                        size_t frame, const float particle_shift[3],
                        const float particle_scale[3], float out12[12]);
void orc_color_to_float4(const float rgb[3], float out[4]);
void orc_color_to_uchar4(const float rgb[3], uint8_t out[4]);

/* XORWOW: curand_init(seed, subsequence, offset) then n uniforms */
void orc_xorwow_uniforms(uint64_t seed, uint64_t subsequence, uint64_t offset, size_t n,
                         float *out);
/* raw state words after curand_init: d, v[0..4] */
void orc_xorwow_state(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t out6[6]);

/* instance flattening: object-space prim -> world space (spec A.1) */
void orc_xform_point(const float m12[12], const float p[3], float out[3]);
void orc_xform_sphere(const float m12[12], const float c[3], float r, float out4[4]);
void orc_xform_normal(const float m12[12], const float n[3], float out[3]);

uint32_t orc_morton30(float x, float y, float z); /* x,y,z in [0,1] */

int orc_num_threads(void);
int orc_hw_threads(void);           /* cores visible to this process */
void orc_set_num_threads(int n);    /* overrides OMP_NUM_THREADS for the following renders */

#ifdef __cplusplus
}
#endif
#endif
