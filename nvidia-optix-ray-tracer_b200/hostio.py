"""ctypes binding of librtb200_io.so (include/rtb200_io.h): config.json, .vtk.series, legacy-VTK,
STL and cache readers + PFM / PPM / EXR writers — the callers and data formats either side of the
hot path (reference: src/Util/ProgramArgumentParser.cu, VTKTimeReader.cu, VTKMeshReader.cu,
VTKReaderImpl.cpp).  All parsing is done by the C++ library; this file only marshals."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librtb200_io.so")

IO_SYMBOLS = [
    "rt_io_last_error", "rt_io_config_load", "rt_io_config_parse", "rt_io_config_get", "rt_io_config_free",
    "rt_io_series_load", "rt_io_series_count", "rt_io_series_file", "rt_io_series_time", "rt_io_series_duration",
    "rt_io_series_free", "rt_io_frames_for_file", "rt_io_particles_load", "rt_io_particles_get",
    "rt_io_particles_free", "rt_io_shapes_load_dir", "rt_io_shapes_load_file", "rt_io_shapes_count",
    "rt_io_shapes_get", "rt_io_shapes_free", "rt_io_mesh_load_vtk", "rt_io_mesh_load_cache",
    "rt_io_mesh_write_cache", "rt_io_mesh_get", "rt_io_mesh_free", "rt_io_cache_generate",
    "rt_io_cache_read_max_cells", "rt_io_write_pfm", "rt_io_write_ppm", "rt_io_write_exr", "rt_io_read_pfm",
]

_f3 = C.c_float * 3
_fp = C.POINTER(C.c_float)
_u64p = C.POINTER(C.c_uint64)


class ConfigView(C.Structure):
    _fields_ = [("mesh", C.c_int), ("cache", C.c_int), ("debug_mode", C.c_int), ("thread_count", C.c_uint64),
                ("series_path", C.c_char_p), ("series_name", C.c_char_p), ("cache_path", C.c_char_p),
                ("stl_path", C.c_char_p), ("preset", C.c_char_p),
                ("n_rough", C.c_uint32), ("n_metal", C.c_uint32), ("n_sphere", C.c_uint32),
                ("roughs", _fp), ("metals", _fp), ("sphere_center_radius", _fp),
                ("sphere_mat_type", C.POINTER(C.c_int32)), ("sphere_mat_index", _u64p), ("sphere_transform", _fp),
                ("api", C.c_char_p), ("flip_up", C.c_int), ("window_width", C.c_int), ("window_height", C.c_int),
                ("fps", C.c_uint64), ("render_speed_ratio", C.c_uint64), ("camera_initial_speed_ratio", C.c_uint64),
                ("camera_center", _f3), ("camera_target", _f3), ("up_direction", _f3),
                ("particle_shift", _f3), ("particle_scale", _f3),
                ("mouse_sensitivity", C.c_float), ("camera_pitch_limit_degree", C.c_float), ("camera_speed_stride", C.c_float),
                ("spp", C.c_uint32), ("max_depth", C.c_uint32), ("seed", C.c_uint64), ("frames", C.c_uint64),
                ("output", C.c_char_p),
                ("n_triangle", C.c_uint32), ("triangle_vertices", _fp), ("triangle_normals", _fp),
                ("triangle_mat_type", C.POINTER(C.c_int32)), ("triangle_mat_index", _u64p)]


class ParticlesView(C.Structure):
    _fields_ = [("n", C.c_size_t), ("pos", _fp), ("id", _u64p), ("quat", _fp), ("vel", _fp), ("shape_id", _u64p),
                ("size", _fp)]


class ShapeView(C.Structure):
    _fields_ = [("n_tris", C.c_size_t), ("vertices", _fp), ("face_normals", _fp), ("vertex_normals", _fp),
                ("path", C.c_char_p)]


class MeshView(C.Structure):
    _fields_ = [("n_particles", C.c_size_t), ("id", _u64p), ("vel", _fp), ("vertex_offset", _u64p), ("vertices", _fp),
                ("normals", _fp)]


class IoError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rtb200_io error {code}: {msg}")
        self.code = code


_lib = None


def build(verbose: bool = False) -> str:
    r = subprocess.run(["make", "-C", os.path.join(HERE, "host"), os.path.join(HERE, "librtb200_io.so")],
                       capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode:
        raise RuntimeError("building librtb200_io.so failed")
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run __graft_entry__.build() / make -C host")
        L = C.CDLL(LIB_PATH)
        vp, sz, cs = C.c_void_p, C.c_size_t, C.c_char_p
        L.rt_io_last_error.restype = cs
        L.rt_io_config_load.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_config_parse.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_config_get.argtypes = [vp, C.POINTER(ConfigView)]
        L.rt_io_config_free.argtypes = [vp]
        L.rt_io_config_free.restype = None
        L.rt_io_series_load.argtypes = [cs, cs, C.POINTER(vp)]
        L.rt_io_series_count.argtypes = [vp]
        L.rt_io_series_count.restype = sz
        L.rt_io_series_file.argtypes = [vp, sz]
        L.rt_io_series_file.restype = cs
        L.rt_io_series_time.argtypes = [vp, sz]
        L.rt_io_series_time.restype = C.c_float
        L.rt_io_series_duration.argtypes = [vp, sz]
        L.rt_io_series_duration.restype = C.c_float
        L.rt_io_series_free.argtypes = [vp]
        L.rt_io_series_free.restype = None
        L.rt_io_frames_for_file.argtypes = [C.c_float, C.c_uint64, C.c_uint64]
        L.rt_io_frames_for_file.restype = sz
        L.rt_io_particles_load.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_particles_get.argtypes = [vp, C.POINTER(ParticlesView)]
        L.rt_io_particles_free.argtypes = [vp]
        L.rt_io_particles_free.restype = None
        L.rt_io_shapes_load_dir.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_shapes_load_file.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_shapes_count.argtypes = [vp]
        L.rt_io_shapes_count.restype = sz
        L.rt_io_shapes_get.argtypes = [vp, sz, C.POINTER(ShapeView)]
        L.rt_io_shapes_free.argtypes = [vp]
        L.rt_io_shapes_free.restype = None
        L.rt_io_mesh_load_vtk.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_mesh_load_cache.argtypes = [cs, C.POINTER(vp)]
        L.rt_io_mesh_write_cache.argtypes = [vp, cs]
        L.rt_io_mesh_get.argtypes = [vp, C.POINTER(MeshView)]
        L.rt_io_mesh_free.argtypes = [vp]
        L.rt_io_mesh_free.restype = None
        L.rt_io_cache_generate.argtypes = [cs, cs, cs, sz, _u64p]
        L.rt_io_cache_read_max_cells.argtypes = [cs, _u64p]
        L.rt_io_write_pfm.argtypes = [cs, C.c_uint32, C.c_uint32, vp, C.c_int]
        L.rt_io_write_ppm.argtypes = [cs, C.c_uint32, C.c_uint32, vp, C.c_int]
        L.rt_io_write_exr.argtypes = [cs, C.c_uint32, C.c_uint32, vp, C.c_int]
        L.rt_io_read_pfm.argtypes = [cs, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), vp, sz]
        _lib = L
    return _lib


def _check(rc):
    if rc:
        raise IoError(rc, lib().rt_io_last_error().decode(errors="replace"))


def _arr(ptr, shape, dtype):
    n = int(np.prod(shape))
    if n == 0 or not ptr:
        return np.zeros(shape, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype).reshape(shape).copy()


def _config_dict(h):
    v = ConfigView()
    _check(lib().rt_io_config_get(h, C.byref(v)))
    d = {k: getattr(v, k) for k in ("mesh", "cache", "debug_mode", "thread_count", "flip_up", "window_width", "window_height",
                                    "fps", "render_speed_ratio", "camera_initial_speed_ratio", "mouse_sensitivity",
                                    "camera_pitch_limit_degree", "camera_speed_stride", "spp", "max_depth", "seed", "frames")}
    for k in ("series_path", "series_name", "cache_path", "stl_path", "preset", "api", "output"):
        d[k] = getattr(v, k).decode()
    for k in ("camera_center", "camera_target", "up_direction", "particle_shift", "particle_scale"):
        d[k] = np.array(list(getattr(v, k)), np.float32)
    d["roughs"] = _arr(v.roughs, (v.n_rough, 3), np.float32)
    d["metals"] = _arr(v.metals, (v.n_metal, 4), np.float32)
    d["sphere_center_radius"] = _arr(v.sphere_center_radius, (v.n_sphere, 4), np.float32)
    d["sphere_mat_type"] = _arr(v.sphere_mat_type, (v.n_sphere,), np.int32)
    d["sphere_mat_index"] = _arr(v.sphere_mat_index, (v.n_sphere,), np.uint64)
    d["sphere_transform"] = _arr(v.sphere_transform, (v.n_sphere, 12), np.float32)
    d["triangle_vertices"] = _arr(v.triangle_vertices, (v.n_triangle, 3, 3), np.float32)
    d["triangle_normals"] = _arr(v.triangle_normals, (v.n_triangle, 3, 3), np.float32)
    d["triangle_mat_type"] = _arr(v.triangle_mat_type, (v.n_triangle,), np.int32)
    d["triangle_mat_index"] = _arr(v.triangle_mat_index, (v.n_triangle,), np.uint64)
    return d


def load_config(path: str) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_config_load(path.encode(), C.byref(h)))
    try:
        return _config_dict(h)
    finally:
        lib().rt_io_config_free(h)


def parse_config(text: str) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_config_parse(text.encode(), C.byref(h)))
    try:
        return _config_dict(h)
    finally:
        lib().rt_io_config_free(h)


def load_series(series_path: str, series_name: str) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_series_load(series_path.encode(), series_name.encode(), C.byref(h)))
    try:
        n = lib().rt_io_series_count(h)
        return dict(files=[lib().rt_io_series_file(h, i).decode() for i in range(n)],
                    times=np.array([lib().rt_io_series_time(h, i) for i in range(n)], np.float32),
                    durations=np.array([lib().rt_io_series_duration(h, i) for i in range(n)], np.float32))
    finally:
        lib().rt_io_series_free(h)


def frames_for_file(duration, fps, ratio) -> int:
    return int(lib().rt_io_frames_for_file(float(duration), int(fps), int(ratio)))


def load_particles(path: str) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_particles_load(path.encode(), C.byref(h)))
    try:
        v = ParticlesView()
        _check(lib().rt_io_particles_get(h, C.byref(v)))
        n = v.n
        return dict(pos=_arr(v.pos, (n, 3), np.float32), id=_arr(v.id, (n,), np.uint64), quat=_arr(v.quat, (n, 4), np.float32),
                    vel=_arr(v.vel, (n, 3), np.float32), shape_id=_arr(v.shape_id, (n,), np.uint64),
                    size=_arr(v.size, (n,), np.float32) if v.size else None)
    finally:
        lib().rt_io_particles_free(h)


def _shapes(h):
    out = []
    for i in range(lib().rt_io_shapes_count(h)):
        v = ShapeView()
        _check(lib().rt_io_shapes_get(h, i, C.byref(v)))
        n = v.n_tris
        out.append(dict(path=v.path.decode(), vertices=_arr(v.vertices, (3 * n, 3), np.float32),
                        face_normals=_arr(v.face_normals, (n, 3), np.float32),
                        vertex_normals=_arr(v.vertex_normals, (3 * n, 3), np.float32)))
    return out


def load_shapes_dir(stl_dir: str) -> list:
    h = C.c_void_p()
    _check(lib().rt_io_shapes_load_dir(stl_dir.encode(), C.byref(h)))
    try:
        return _shapes(h)
    finally:
        lib().rt_io_shapes_free(h)


def load_stl(path: str) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_shapes_load_file(path.encode(), C.byref(h)))
    try:
        return _shapes(h)[0]
    finally:
        lib().rt_io_shapes_free(h)


def _mesh(h):
    v = MeshView()
    _check(lib().rt_io_mesh_get(h, C.byref(v)))
    n = v.n_particles
    off = _arr(v.vertex_offset, (n + 1,), np.uint64)
    tot = int(off[-1]) if n else 0
    return dict(id=_arr(v.id, (n,), np.uint64), vel=_arr(v.vel, (n, 3), np.float32), vertex_offset=off,
                vertices=_arr(v.vertices, (tot, 3), np.float32), normals=_arr(v.normals, (tot, 3), np.float32))


def load_mesh_vtk(path: str, write_cache_to: str | None = None) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_mesh_load_vtk(path.encode(), C.byref(h)))
    try:
        if write_cache_to:
            _check(lib().rt_io_mesh_write_cache(h, write_cache_to.encode()))
        return _mesh(h)
    finally:
        lib().rt_io_mesh_free(h)


def load_mesh_cache(path: str) -> dict:
    h = C.c_void_p()
    _check(lib().rt_io_mesh_load_cache(path.encode(), C.byref(h)))
    try:
        return _mesh(h)
    finally:
        lib().rt_io_mesh_free(h)


def cache_generate(series_path: str, series_name: str, cache_dir: str, max_threads: int = 8) -> int:
    m = C.c_uint64()
    _check(lib().rt_io_cache_generate(series_path.encode(), series_name.encode(), cache_dir.encode(), max_threads, C.byref(m)))
    return m.value


def cache_read_max_cells(cache_dir: str) -> int:
    m = C.c_uint64()
    _check(lib().rt_io_cache_read_max_cells(cache_dir.encode(), C.byref(m)))
    return m.value


def write_pfm(path, rgba: np.ndarray, flip_y=False):
    h, w = rgba.shape[:2]
    a = np.ascontiguousarray(rgba, np.float32)
    _check(lib().rt_io_write_pfm(path.encode(), w, h, a.ctypes.data, int(flip_y)))


def write_exr(path, rgba: np.ndarray, flip_y=False):
    h, w = rgba.shape[:2]
    a = np.ascontiguousarray(rgba, np.float32)
    _check(lib().rt_io_write_exr(path.encode(), w, h, a.ctypes.data, int(flip_y)))


def write_ppm(path, rgba8: np.ndarray, flip_y=False):
    h, w = rgba8.shape[:2]
    a = np.ascontiguousarray(rgba8, np.uint8)
    _check(lib().rt_io_write_ppm(path.encode(), w, h, a.ctypes.data, int(flip_y)))


def read_pfm(path) -> np.ndarray:
    w, h = C.c_uint32(), C.c_uint32()
    _check(lib().rt_io_read_pfm(path.encode(), C.byref(w), C.byref(h), None, 0))
    out = np.zeros((h.value, w.value, 3), np.float32)
    _check(lib().rt_io_read_pfm(path.encode(), C.byref(w), C.byref(h), out.ctypes.data, out.size))
    return out
