/*
 * rtb200_io.h — C ABI of the host-side data ingest / output library (librtb200_io.so): the callers
 * and data formats on either side of the hot path.  Pure CPU code, no CUDA,
 * no VTK / nlohmann dependency.  Citations are relative to the reference tree.
 *
 *   rt_io_config_*      config.json            src/Util/ProgramArgumentParser.cu:41-160 (schema docs/configuration.md)
 *   rt_io_series_*      .vtk.series            src/Util/VTKTimeReader.cu:31-88 (= VTKMeshReader.cu:86-145)
 *   rt_io_particles_*   Time-mode VTK          src/Util/VTKReaderImpl.cpp:139-251 (vtkPolyDataReader, point data)
 *   rt_io_shape_*       ASCII / binary STL     src/Util/VTKReaderImpl.cpp:254-321, VTKTimeReader.cu:90-124
 *   rt_io_mesh_*        Mesh-mode VTK + cache  src/Util/VTKReaderImpl.cpp:24-137, VTKMeshReader.cu:38-82,147-283
 *                                              (format include/Util/VTKMeshReader.cuh:10-24)
 *   rt_io_write_*       headless outputs       replace the src/GraphicsAPI presentation layer (PFM, PPM, EXR)
 *
 * All functions return 0 on success or an RT_IO_ERR_* code; rt_io_last_error() gives the message of
 * the calling thread's last failure.  The reference logs and exit(-1) (VTK_READER_ERROR_EXIT_CODE) /
 * exit(-2) (COMMAND_PARSER_ERROR_EXIT_CODE) instead; nothing here exits.  Views returned by *_get
 * point into the handle and stay valid until the matching *_free.
 */
#ifndef RTB200_IO_H
#define RTB200_IO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_IO_OK 0
#define RT_IO_ERR_OPEN 1    /* file / directory cannot be opened */
#define RT_IO_ERR_PARSE 2   /* malformed JSON / VTK / STL text, missing key */
#define RT_IO_ERR_FORMAT 3  /* well-formed but unsupported or inconsistent content */
#define RT_IO_ERR_ARG 4

const char *rt_io_last_error(void);

/* ---------------------------------------------------------------- config.json */
typedef struct rt_io_config_s *rt_io_config;
typedef struct {
    /* top level (ProgramArgumentParser.cu:108-131) */
    int mesh, cache, debug_mode;
    uint64_t thread_count;                 /* cache-process-thread-count */
    const char *series_path, *series_name, *cache_path, *stl_path, *preset;
    /* materials and extra geometry (ProgramArgumentParser.cu:4-39,63-80) */
    uint32_t n_rough, n_metal, n_sphere;
    const float *roughs;                   /* [n_rough][3] */
    const float *metals;                   /* [n_metal][4] albedo + fuzz */
    const float *sphere_center_radius;     /* [n_sphere][4] */
    const int32_t *sphere_mat_type;        /* 0 ROUGH ("ROUGH"), 1 METAL (anything else) */
    const uint64_t *sphere_mat_index;
    const float *sphere_transform;         /* [n_sphere][12] constructTransformMatrix(shift, rotate, scale) */
    /* loop-data (ProgramArgumentParser.cu:85-128) */
    const char *api;                       /* "OGL" | "VK" | "D3D11" | "D3D12" */
    int flip_up;                           /* 1 unless api == "OGL" (SDL_GraphicsWindow.cu:7-9) */
    int window_width, window_height;
    uint64_t fps, render_speed_ratio, camera_initial_speed_ratio;
    float camera_center[3], camera_target[3], up_direction[3];
    float particle_shift[3], particle_scale[3];
    float mouse_sensitivity, camera_pitch_limit_degree, camera_speed_stride;
    /* additive keys (optional): defaults 1 / 5 / 0 / "" / 0 */
    uint32_t spp, max_depth;
    uint64_t seed, frames;
    const char *output;
    /* "triangles" (docs/configuration.md:232-247; documented, never parsed by the reference: SURVEY §C quirk 10):
     * world-space triangles that follow the config spheres in instance (SBT) order */
    uint32_t n_triangle;
    const float *triangle_vertices;        /* [n_triangle][9] */
    const float *triangle_normals;         /* [n_triangle][9] per-vertex normals */
    const int32_t *triangle_mat_type;
    const uint64_t *triangle_mat_index;
} rt_io_config_view;
int rt_io_config_load(const char *path, rt_io_config *out);
int rt_io_config_parse(const char *json_text, rt_io_config *out);
int rt_io_config_get(rt_io_config cfg, rt_io_config_view *out);
void rt_io_config_free(rt_io_config cfg);

/* ---------------------------------------------------------------- .vtk.series */
typedef struct rt_io_series_s *rt_io_series;
int rt_io_series_load(const char *series_path, const char *series_name, rt_io_series *out);
size_t rt_io_series_count(rt_io_series s);
const char *rt_io_series_file(rt_io_series s, size_t i);   /* series_path + name */
float rt_io_series_time(rt_io_series s, size_t i);
float rt_io_series_duration(rt_io_series s, size_t i);     /* t[i+1]-t[i]; last = previous; single file = 1000 */
void rt_io_series_free(rt_io_series s);
/* frames rendered for file i: (size_t)(duration * fps * render_speed_ratio)  (RendererTime.cu:427-428) */
size_t rt_io_frames_for_file(float duration, uint64_t fps, uint64_t render_speed_ratio);

/* ---------------------------------------------------------------- Time-mode particle files */
typedef struct rt_io_particles_s *rt_io_particles;
typedef struct {
    size_t n;
    const float *pos;          /* [n][3]  POINTS */
    const uint64_t *id;        /* [n]     SCALARS id */
    const float *quat;         /* [n][4]  SCALARS quat, file component order */
    const float *vel;          /* [n][3]  VECTORS vel */
    const uint64_t *shape_id;  /* [n]     SCALARS shape_id */
    const float *size;         /* [n] SCALARS size, or NULL when the file has none (walls) — unused by the
                                  reference; the spheres-only mapping of BASELINE config 1 takes r = size/2 */
} rt_io_particles_view;
int rt_io_particles_load(const char *vtk_path, rt_io_particles *out);
int rt_io_particles_get(rt_io_particles p, rt_io_particles_view *out);
void rt_io_particles_free(rt_io_particles p);

/* ---------------------------------------------------------------- STL shapes */
typedef struct rt_io_shapes_s *rt_io_shapes;
typedef struct {
    size_t n_tris;
    const float *vertices;        /* [3 n_tris][3] */
    const float *face_normals;    /* [n_tris][3]   what the reference uploads (one per triangle) */
    const float *vertex_normals;  /* [3 n_tris][3] face normal repeated per vertex: what the shader indexes */
    const char *path;
} rt_io_shape_view;
/* every *.stl of a directory, sorted by file name; index = shape_id (VTKTimeReader.cu:94-102) */
int rt_io_shapes_load_dir(const char *stl_dir, rt_io_shapes *out);
int rt_io_shapes_load_file(const char *stl_path, rt_io_shapes *out);
size_t rt_io_shapes_count(rt_io_shapes s);
int rt_io_shapes_get(rt_io_shapes s, size_t i, rt_io_shape_view *out);
void rt_io_shapes_free(rt_io_shapes s);

/* ---------------------------------------------------------------- Mesh-mode files and cache */
typedef struct rt_io_mesh_s *rt_io_mesh;
typedef struct {
    size_t n_particles;            /* cells (triangle strips) of the VTK file */
    const uint64_t *id;            /* [n] CELL_DATA id */
    const float *vel;              /* [n][3] CELL_DATA vel */
    const uint64_t *vertex_offset; /* [n+1] prefix of vertex counts (3 per triangle) */
    const float *vertices;         /* [total][3] strip triangles, odd ones with the last two swapped */
    const float *normals;          /* [total][3] smooth point normals */
} rt_io_mesh_view;
int rt_io_mesh_load_vtk(const char *vtk_path, rt_io_mesh *out);
int rt_io_mesh_load_cache(const char *cache_path, rt_io_mesh *out);
int rt_io_mesh_write_cache(rt_io_mesh m, const char *cache_path);
int rt_io_mesh_get(rt_io_mesh m, rt_io_mesh_view *out);
void rt_io_mesh_free(rt_io_mesh m);
/* writeVTKDataCache: cache_dir/particle<i>.cache for every series entry + cache_dir/metadata.cache
 * (ASCII max cell count); existing *.cache files are removed first. */
int rt_io_cache_generate(const char *series_path, const char *series_name, const char *cache_dir,
                         size_t max_threads, uint64_t *max_cells_out);
int rt_io_cache_read_max_cells(const char *cache_dir, uint64_t *out);

/* ---------------------------------------------------------------- headless outputs
 * rgba = float[4*w*h], row 0 first.  flip_y != 0 writes rows bottom-up -> top-down (the reference's
 * OGL convention has row 0 at the bottom). */
int rt_io_write_pfm(const char *path, uint32_t w, uint32_t h, const float *rgba, int flip_y);
int rt_io_write_ppm(const char *path, uint32_t w, uint32_t h, const uint8_t *rgba8, int flip_y);
int rt_io_write_exr(const char *path, uint32_t w, uint32_t h, const float *rgba, int flip_y);
int rt_io_read_pfm(const char *path, uint32_t *w, uint32_t *h, float *rgb_out, size_t capacity_floats);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_IO_H */
