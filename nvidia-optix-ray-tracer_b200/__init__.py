"""ctypes binding of librtb200.so (include/rtb200.h) — the B200-native replacement of the
reference's OptiX seam (include/Global/RendererImpl.cuh:153-253 + optixLaunch call sites).

Python is test/bench orchestration only: every function here forwards to the C ABI.  There is no
fallback path: if the CUDA library is missing or fails to load, importing the GPU entry points
raises.  The directory name carries a hyphen, so import it with
``importlib.import_module("nvidia-optix-ray-tracer_b200")`` (see tests/conftest.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
# RTB200_LIB: load another build of the same library (A/B experiments: make -C csrc OUT=... EXTRA=-D...)
LIB_PATH = os.environ.get("RTB200_LIB") or os.path.join(HERE, "librtb200.so")
HOST_LIB_PATH = os.path.join(HERE, "librtb200_io.so")

RT_OK, RT_ERR_INVALID, RT_ERR_CUDA, RT_ERR_OOM, RT_ERR_UNSUPPORTED, RT_ERR_STATE = range(6)
RT_MAT_ROUGH, RT_MAT_METAL = 0, 1
RT_MISS_ID = 0xFFFFFFFF
RT_FLAG_ACCUMULATE = 1
RT_FLAG_REF_NO_SELF_SKIP = 2   # rtb200.h: the reference's transport details (Shader.cu:230-233, 236-238)
RT_FLAG_REF_MUL_ON_RETURN = 4
RT_RNG_SAMPLE_STRIDE = 32

# every symbol include/rtb200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "rt_abi_version", "rt_context_create", "rt_context_destroy", "rt_last_error",
    "rt_scene_create", "rt_scene_destroy", "rt_scene_set_option", "rt_scene_set_materials", "rt_scene_add_spheres",
    "rt_scene_add_triangles", "rt_scene_clear_geometry", "rt_scene_update_group", "rt_scene_add_instance",
    "rt_scene_set_transforms", "rt_accel_build", "rt_render", "rt_resolve", "rt_get_stats",
    "rt_accel_download", "rt_accel_download_wide", "rt_trace_rays", "rt_camera_configure", "rt_construct_transform",
    "rt_transform_time", "rt_transform_mesh", "rt_bake_color_ramp",
    "rt_spheres_advect", "rt_transforms_time", "rt_scene_set_transforms_device", "rt_render_primary", "rt_scene_set_primary_hits", "rt_peer_alloc", "rt_peer_free", "rt_peer_open", "rt_peer_close", "rt_resolve_parts",
    "rt_trace_rays_static",
]
RT_TRACE_STATIC_MAX_RAYS = 65536
RT_MAX_PARTS = 16
RT_PEER_HANDLE_BYTES = 64


class Camera(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("U", C.c_float * 3), ("V", C.c_float * 3), ("W", C.c_float * 3)]


class FrameParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32), ("max_depth", C.c_uint32),
                ("seed", C.c_uint64), ("sample_offset", C.c_uint32), ("background", C.c_float * 3),
                ("flags", C.c_uint32)]


class Outputs(C.Structure):
    _fields_ = [("accum_linear", C.c_void_p), ("albedo", C.c_void_p), ("normal", C.c_void_p),
                ("prim_id", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [("n_prims", C.c_uint64), ("n_bvh_prims", C.c_uint64), ("n_huge", C.c_uint64),
                ("n_nodes", C.c_uint64), ("device_bytes", C.c_uint64), ("build_ms", C.c_float),
                ("render_ms", C.c_float), ("segments", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("scene_lo", C.c_float * 3), ("scene_hi", C.c_float * 3), ("trace_ms", C.c_float),
                ("trace_launches", C.c_uint32), ("segments_total", C.c_uint64)]


class RtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rtb200 error {code}: {msg}")
        self.code = code


_lib = None


def build(verbose: bool = False) -> str:
    """Compile librtb200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", os.path.join(HERE, "csrc")], capture_output=True, text=True)
    if verbose or r.returncode:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode:
        raise RuntimeError("building librtb200.so failed")
    return LIB_PATH


def lib() -> C.CDLL:
    """Load the C-ABI library.  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run __graft_entry__.build() / make -C csrc (no CPU fallback exists)")
        L = C.CDLL(LIB_PATH)
        vp, u32, u64, sz, f_p = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t, C.POINTER(C.c_float)
        L.rt_abi_version.restype = C.c_int
        L.rt_context_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.rt_context_destroy.argtypes = [vp]
        L.rt_last_error.argtypes = [vp]
        L.rt_last_error.restype = C.c_char_p
        L.rt_scene_create.argtypes = [vp, C.POINTER(vp)]
        L.rt_scene_destroy.argtypes = [vp]
        L.rt_scene_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
        L.rt_scene_set_materials.argtypes = [vp, vp, sz, vp, sz]
        L.rt_scene_add_spheres.argtypes = [vp, vp, vp, sz, C.POINTER(u32)]
        L.rt_scene_add_triangles.argtypes = [vp, vp, vp, sz, C.POINTER(u32)]
        L.rt_scene_clear_geometry.argtypes = [vp]
        L.rt_scene_update_group.argtypes = [vp, u32, vp, vp, sz]
        L.rt_scene_add_instance.argtypes = [vp, u32, vp, C.c_int, u32, u32, C.POINTER(u32)]
        L.rt_scene_set_transforms.argtypes = [vp, vp, sz, sz, vp]
        L.rt_accel_build.argtypes = [vp, vp]
        L.rt_render.argtypes = [vp, C.POINTER(Camera), C.POINTER(FrameParams), C.POINTER(Outputs), vp]
        L.rt_resolve.argtypes = [vp, vp, sz, vp, vp, vp]
        L.rt_spheres_advect.argtypes = [vp, vp, vp, vp, sz, C.c_float, sz, sz, f_p, f_p, vp, vp, vp]
        L.rt_transforms_time.argtypes = [vp, vp, vp, vp, vp, sz, C.c_float, sz, sz, f_p, f_p, vp, vp]
        L.rt_scene_set_transforms_device.argtypes = [vp, vp, sz, sz]
        L.rt_render_primary.argtypes = [vp, C.POINTER(Camera), u32, u32, sz, sz, vp, vp]
        L.rt_scene_set_primary_hits.argtypes = [vp, vp]
        L.rt_peer_alloc.argtypes = [vp, sz, C.POINTER(vp), C.c_char_p]
        L.rt_peer_free.argtypes = [vp, vp]
        L.rt_peer_open.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
        L.rt_peer_close.argtypes = [vp, vp]
        L.rt_resolve_parts.argtypes = [vp, C.POINTER(vp), C.c_int, sz, sz, vp, vp, vp, vp]
        L.rt_get_stats.argtypes = [vp, C.POINTER(Stats)]
        L.rt_accel_download.argtypes = [vp, vp, vp, vp, vp]
        L.rt_accel_download_wide.argtypes = [vp, vp, vp, vp, C.POINTER(u64)]
        L.rt_trace_rays.argtypes = [vp, vp, vp, sz, vp, vp, vp]
        L.rt_trace_rays_static.argtypes = [vp, vp, vp, sz, vp, vp, vp]
        L.rt_camera_configure.argtypes = [f_p, f_p, f_p, C.c_int, C.POINTER(Camera)]
        L.rt_camera_configure.restype = None
        L.rt_construct_transform.argtypes = [f_p, f_p, f_p, f_p]
        L.rt_construct_transform.restype = None
        L.rt_transform_time.argtypes = [f_p, f_p, f_p, f_p, C.c_float, sz, sz, f_p, f_p, f_p]
        L.rt_transform_time.restype = None
        L.rt_transform_mesh.argtypes = [f_p, C.c_float, sz, sz, f_p, f_p, f_p]
        L.rt_transform_mesh.restype = None
        L.rt_bake_color_ramp.argtypes = [C.c_char_p, sz, vp]
        L.rt_bake_color_ramp.restype = C.c_int
        _lib = L
    return _lib


def _fa(values):
    return (C.c_float * len(values))(*[float(v) for v in values])


# ------------------------------------------------------------------ host maths (no GPU needed)
def camera_configure(center, target, up, flip_up: bool) -> Camera:
    cam = Camera()
    lib().rt_camera_configure(_fa(center), _fa(target), _fa(up), int(bool(flip_up)), C.byref(cam))
    return cam


def construct_transform(shift, rotate_deg, scale):
    out = (C.c_float * 12)()
    lib().rt_construct_transform(_fa(shift), _fa(rotate_deg), _fa(scale), out)
    return list(out)


def transform_time(pos, vel, quat_cur, quat_next, duration, frames, frame, pshift, pscale):
    out = (C.c_float * 12)()
    lib().rt_transform_time(_fa(pos), _fa(vel), _fa(quat_cur), _fa(quat_next), float(duration), frames, frame,
                            _fa(pshift), _fa(pscale), out)
    return list(out)


def transform_mesh(vel, duration, frames, frame, pshift, pscale):
    out = (C.c_float * 12)()
    lib().rt_transform_mesh(_fa(vel), float(duration), frames, frame, _fa(pshift), _fa(pscale), out)
    return list(out)


def bake_color_ramp(preset: str, count: int):
    import numpy as np
    out = np.zeros((count, 3), dtype=np.float32)
    lib().rt_bake_color_ramp(preset.encode(), count, out.ctypes.data)
    return out


# ------------------------------------------------------------------ GPU objects
class Context:
    def __init__(self, device: int = 0):
        self._h = C.c_void_p()
        rc = lib().rt_context_create(device, C.byref(self._h))
        if rc:
            raise RtError(rc, lib().rt_last_error(None).decode())
        self.device = device

    def check(self, rc):
        if rc:
            raise RtError(rc, lib().rt_last_error(self._h).decode())

    def resolve(self, accum_ptr, n_pixels, color_ptr=None, rgba8_ptr=None, stream=None):
        self.check(lib().rt_resolve(self._h, accum_ptr, n_pixels, color_ptr, rgba8_ptr, stream))

    def spheres_advect(self, base_c_ptr, base_r_ptr, vel_ptr, n, duration, frames, frame, pshift, pscale, out_c_ptr, out_r_ptr,
                       stream=None):
        self.check(lib().rt_spheres_advect(self._h, base_c_ptr, base_r_ptr, vel_ptr, n, duration, frames, frame, _fa(pshift),
                                           _fa(pscale), out_c_ptr, out_r_ptr, stream))

    # ---- multi-GPU exchange step (include/rtb200.h "rt_peer_*", "rt_resolve_parts")
    def transforms_time(self, pos_ptr, vel_ptr, quat_cur_ptr, quat_next_ptr, n, duration, frames, frame, pshift, pscale, out_ptr,
                        stream=None):
        """Device form of transform_time for n particles (include/rtb200.h rt_transforms_time)."""
        self.check(lib().rt_transforms_time(self._h, pos_ptr, vel_ptr, quat_cur_ptr, quat_next_ptr, n, duration, frames, frame,
                                            _fa(pshift), _fa(pscale), out_ptr, stream))

    def peer_alloc(self, nbytes: int):
        """(device pointer, 64-byte handle) of a buffer other processes can map with peer_open."""
        ptr = C.c_void_p()
        handle = C.create_string_buffer(RT_PEER_HANDLE_BYTES)
        self.check(lib().rt_peer_alloc(self._h, nbytes, C.byref(ptr), handle))
        return ptr.value, handle.raw

    def peer_free(self, ptr):
        self.check(lib().rt_peer_free(self._h, ptr))

    def peer_open(self, handle: bytes) -> int:
        ptr = C.c_void_p()
        self.check(lib().rt_peer_open(self._h, C.create_string_buffer(bytes(handle), RT_PEER_HANDLE_BYTES), C.byref(ptr)))
        return ptr.value

    def peer_close(self, ptr):
        self.check(lib().rt_peer_close(self._h, ptr))

    def resolve_parts(self, part_ptrs, first_pixel, n_pixels, sum_ptr=None, color_ptr=None, rgba8_ptr=None, stream=None):
        arr = (C.c_void_p * len(part_ptrs))(*part_ptrs)
        self.check(lib().rt_resolve_parts(self._h, arr, len(part_ptrs), first_pixel, n_pixels, sum_ptr, color_ptr,
                                          rgba8_ptr, stream))

    def close(self):
        if self._h:
            lib().rt_context_destroy(self._h)
            self._h = C.c_void_p()


class Scene:
    """Mirror of the reference's scene assembly + frame calls (RendererTime/RendererMesh
    commitRendererData / startRender) on top of the C ABI.  Pointers are raw device addresses
    (e.g. torch.Tensor.data_ptr()); the caller keeps the tensors alive."""

    def __init__(self, ctx: Context):
        self.ctx = ctx
        self._h = C.c_void_p()
        ctx.check(lib().rt_scene_create(ctx._h, C.byref(self._h)))

    def set_option(self, name: str, value: int):
        self.ctx.check(lib().rt_scene_set_option(self._h, name.encode(), int(value)))

    def set_materials(self, roughs_ptr, n_rough, metals_ptr, n_metal):
        self.ctx.check(lib().rt_scene_set_materials(self._h, roughs_ptr, n_rough, metals_ptr, n_metal))

    def add_spheres(self, centers_ptr, radii_ptr, count) -> int:
        g = C.c_uint32()
        self.ctx.check(lib().rt_scene_add_spheres(self._h, centers_ptr, radii_ptr, count, C.byref(g)))
        return g.value

    def add_triangles(self, verts_ptr, normals_ptr, count) -> int:
        g = C.c_uint32()
        self.ctx.check(lib().rt_scene_add_triangles(self._h, verts_ptr, normals_ptr, count, C.byref(g)))
        return g.value

    def clear_geometry(self):
        self.ctx.check(lib().rt_scene_clear_geometry(self._h))

    def update_group(self, group, a_ptr, b_ptr, count):
        self.ctx.check(lib().rt_scene_update_group(self._h, group, a_ptr, b_ptr, count))

    def add_instance(self, group, transform=None, mat_type=RT_MAT_ROUGH, mat_index=0, mat_stride=0) -> int:
        i = C.c_uint32()
        t = _fa(transform) if transform is not None else None
        self.ctx.check(lib().rt_scene_add_instance(self._h, group, t, mat_type, mat_index, mat_stride, C.byref(i)))
        return i.value

    def set_transforms(self, transforms_ptr, first, n, stream=None):
        self.ctx.check(lib().rt_scene_set_transforms(self._h, transforms_ptr, first, n, stream))

    def set_transforms_device(self, transforms_ptr, first, n):
        self.ctx.check(lib().rt_scene_set_transforms_device(self._h, transforms_ptr, first, n))

    def build(self, stream=None):
        self.ctx.check(lib().rt_accel_build(self._h, stream))

    def render(self, cam: Camera, width, height, spp, max_depth, seed, background=(0.7, 0.8, 0.9), sample_offset=0,
               flags=0, accum=None, albedo=None, normal=None, prim_id=None, stream=None):
        fp = FrameParams(width, height, spp, max_depth, seed, sample_offset, (C.c_float * 3)(*background), flags)
        out = Outputs(accum, albedo, normal, prim_id)
        self.ctx.check(lib().rt_render(self._h, C.byref(cam), C.byref(fp), C.byref(out), stream))

    def render_primary(self, cam: Camera, width, height, first_pixel, n_pixels, hits_ptr, stream=None):
        self.ctx.check(lib().rt_render_primary(self._h, C.byref(cam), width, height, first_pixel, n_pixels, hits_ptr, stream))

    def set_primary_hits(self, hits_ptr):
        self.ctx.check(lib().rt_scene_set_primary_hits(self._h, hits_ptr))

    def trace_rays(self, rays_ptr, skip_ptr, n, t_ptr, id_ptr, stream=None):
        self.ctx.check(lib().rt_trace_rays(self._h, rays_ptr, skip_ptr, n, t_ptr, id_ptr, stream))

    def trace_rays_static(self, rays_ptr, skip_ptr, n, t_ptr, id_ptr, stream=None):
        """rt_trace_rays through the library's fixed scratch array: allocates nothing, n <= RT_TRACE_STATIC_MAX_RAYS."""
        self.ctx.check(lib().rt_trace_rays_static(self._h, rays_ptr, skip_ptr, n, t_ptr, id_ptr, stream))

    def stats(self) -> Stats:
        s = Stats()
        self.ctx.check(lib().rt_get_stats(self._h, C.byref(s)))
        return s

    def download(self, nodes_ptr=None, slot_geom_ptr=None, slot_prim_ptr=None, keys_ptr=None):
        self.ctx.check(lib().rt_accel_download(self._h, nodes_ptr, slot_geom_ptr, slot_prim_ptr, keys_ptr))

    def download_wide(self, wnodes_ptr=None, slot_geom_ptr=None, slot_prim_ptr=None) -> int:
        n = C.c_uint64()
        self.ctx.check(lib().rt_accel_download_wide(self._h, wnodes_ptr, slot_geom_ptr, slot_prim_ptr, C.byref(n)))
        return n.value

    def close(self):
        if self._h:
            lib().rt_scene_destroy(self._h)
            self._h = C.c_void_p()
