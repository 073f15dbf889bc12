"""ctypes wrapper of the CPU oracle (oracle/rt_oracle.h).  TEST INFRASTRUCTURE: only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB = os.path.join(ORACLE_DIR, "_build", "librt_oracle.so")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "librt_ref.so")
REF_SHADER_LIB = os.path.join(ORACLE_DIR, "_ref", "librt_ref_shader.so")
FLAG_NO_SELF_SKIP = 1   # rt_oracle.h ORC_FLAG_*: the reference's behaviour for the two transport deviations
FLAG_MUL_ON_RETURN = 2

MISS = 0xFFFFFFFF


class OrcScene(C.Structure):
    _fields_ = [("n_prims", C.c_uint32), ("kind_index", C.c_void_p), ("mat", C.c_void_p), ("spheres", C.c_void_p),
                ("tri_verts", C.c_void_p), ("tri_normals", C.c_void_p), ("roughs", C.c_void_p),
                ("metals", C.c_void_p), ("n_rough", C.c_uint32), ("n_metal", C.c_uint32), ("sphere_shade", C.c_void_p)]


class OrcCamera(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("U", C.c_float * 3), ("V", C.c_float * 3), ("W", C.c_float * 3)]


class OrcParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("spp", C.c_uint32), ("max_depth", C.c_uint32),
                ("seed", C.c_uint64), ("sample_offset", C.c_uint32), ("background", C.c_float * 3),
                ("flags", C.c_uint32)]


def build():
    r = subprocess.run(["make", "-C", ORACLE_DIR, "all"], capture_output=True, text=True)
    if r.returncode:
        raise RuntimeError("oracle build failed:\n" + r.stdout[-3000:] + r.stderr[-3000:])


def _fa(v):
    return (C.c_float * len(v))(*[float(x) for x in v])


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        f_p, sz, u64 = C.POINTER(C.c_float), C.c_size_t, C.c_uint64
        lib.orc_render.argtypes = [C.POINTER(OrcScene), C.POINTER(OrcCamera), C.POINTER(OrcParams), C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(u64)]
        lib.orc_trace.argtypes = [C.POINTER(OrcScene), C.c_int, sz, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.orc_camera_configure.argtypes = [f_p, f_p, f_p, C.c_int, C.POINTER(OrcCamera)]
        lib.orc_construct_transform.argtypes = [f_p, f_p, f_p, f_p]
        lib.orc_slerp.argtypes = [f_p, f_p, C.c_float, f_p]
        lib.orc_quat_to_euler.argtypes = [f_p, f_p]
        lib.orc_transform_time.argtypes = [f_p, f_p, f_p, f_p, C.c_float, sz, sz, f_p, f_p, f_p]
        lib.orc_transform_mesh.argtypes = [f_p, C.c_float, sz, sz, f_p, f_p, f_p]
        lib.orc_color_to_float4.argtypes = [f_p, f_p]
        lib.orc_color_to_uchar4.argtypes = [f_p, C.c_void_p]
        lib.orc_xorwow_uniforms.argtypes = [u64, u64, u64, sz, C.c_void_p]
        lib.orc_xorwow_state.argtypes = [u64, u64, u64, C.c_void_p]
        lib.orc_xform_point.argtypes = [f_p, f_p, f_p]
        lib.orc_xform_sphere.argtypes = [f_p, f_p, C.c_float, f_p]
        lib.orc_xform_normal.argtypes = [f_p, f_p, f_p]
        lib.orc_morton30.argtypes = [C.c_float, C.c_float, C.c_float]
        lib.orc_morton30.restype = C.c_uint32
        lib.orc_num_threads.restype = C.c_int

    # ---- scene packing -------------------------------------------------------------------
    @staticmethod
    def pack_scene(world):
        """world: dict from scenes.flatten_world() -> (OrcScene, keepalive list)."""
        keep = [np.ascontiguousarray(world[k]) for k in
                ("kind_index", "mat", "spheres", "tri_verts", "tri_normals", "roughs", "metals")]
        shade = world.get("sphere_shade")
        if shade is not None:
            shade = np.ascontiguousarray(shade, np.float32)
            assert shade.shape == keep[2].shape
            keep.append(shade)
        sc = OrcScene(len(keep[0]), *[a.ctypes.data for a in keep[:7]], len(world["roughs"]), len(world["metals"]),
                      shade.ctypes.data if shade is not None else None)
        return sc, keep

    def render(self, world, cam, width, height, spp, max_depth, seed, background=(0.7, 0.8, 0.9), sample_offset=0,
               accel=0, want=("accum", "albedo", "normal", "prim_id"), flags=0):
        sc, keep = self.pack_scene(world)
        oc = OrcCamera(tuple(cam.center), tuple(cam.U), tuple(cam.V), tuple(cam.W))
        p = OrcParams(width, height, spp, max_depth, seed, sample_offset, (C.c_float * 3)(*background), flags)
        n = width * height
        out = {
            "accum": np.zeros((n, 4), np.float32) if "accum" in want else None,
            "albedo": np.zeros((n, 4), np.float32) if "albedo" in want else None,
            "normal": np.zeros((n, 4), np.float32) if "normal" in want else None,
            "prim_id": np.zeros(n, np.uint32) if "prim_id" in want else None,
        }
        seg = C.c_uint64(0)
        rc = self.lib.orc_render(C.byref(sc), C.byref(oc), C.byref(p), accel,
                                 *[(out[k].ctypes.data if out[k] is not None else None)
                                   for k in ("accum", "albedo", "normal", "prim_id")], C.byref(seg))
        assert rc == 0
        out["segments"] = seg.value
        return out

    # ---- pre-built BVH (CPU baseline timing: build and render separately) -----------------
    def accel_create(self, world):
        sc, keep = self.pack_scene(world)
        self.lib.orc_accel_create.restype = C.c_void_p
        self.lib.orc_accel_create.argtypes = [C.POINTER(OrcScene)]
        h = self.lib.orc_accel_create(C.byref(sc))
        return dict(handle=C.c_void_p(h), scene=sc, keep=keep)

    def accel_destroy(self, acc):
        self.lib.orc_accel_destroy.argtypes = [C.c_void_p]
        self.lib.orc_accel_destroy(acc["handle"])

    def render_accel(self, acc, cam, width, height, spp, max_depth, seed, background=(0.7, 0.8, 0.9), sample_offset=0,
                     want=("accum",)):
        self.lib.orc_render_accel.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.POINTER(OrcCamera), C.POINTER(OrcParams),
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        oc = OrcCamera(tuple(cam.center), tuple(cam.U), tuple(cam.V), tuple(cam.W))
        p = OrcParams(width, height, spp, max_depth, seed, sample_offset, (C.c_float * 3)(*background), 0)
        n = width * height
        out = {"accum": np.zeros((n, 4), np.float32) if "accum" in want else None,
               "prim_id": np.zeros(n, np.uint32) if "prim_id" in want else None}
        seg = C.c_uint64(0)
        rc = self.lib.orc_render_accel(C.byref(acc["scene"]), acc["handle"], C.byref(oc), C.byref(p),
                                       out["accum"].ctypes.data if out["accum"] is not None else None, None, None,
                                       out["prim_id"].ctypes.data if out["prim_id"] is not None else None, C.byref(seg))
        assert rc == 0
        out["segments"] = seg.value
        return out

    def trace(self, world, rays, skip=None, accel=0):
        sc, keep = self.pack_scene(world)
        rays = np.ascontiguousarray(rays, np.float32)
        n = len(rays)
        t = np.zeros(n, np.float32)
        ids = np.zeros(n, np.uint32)
        sk = np.ascontiguousarray(skip, np.uint32) if skip is not None else None
        self.lib.orc_trace(C.byref(sc), accel, n, rays.ctypes.data, sk.ctypes.data if sk is not None else None,
                           t.ctypes.data, ids.ctypes.data)
        return t, ids

    # ---- host maths ----------------------------------------------------------------------
    def camera_configure(self, center, target, up, flip_up):
        cam = OrcCamera()
        self.lib.orc_camera_configure(_fa(center), _fa(target), _fa(up), int(bool(flip_up)), C.byref(cam))
        return cam

    def construct_transform(self, shift, rot, scale):
        out = (C.c_float * 12)()
        self.lib.orc_construct_transform(_fa(shift), _fa(rot), _fa(scale), out)
        return list(out)

    def slerp(self, q1, q2, t):
        out = (C.c_float * 4)()
        self.lib.orc_slerp(_fa(q1), _fa(q2), float(t), out)
        return list(out)

    def quat_to_euler(self, q):
        out = (C.c_float * 3)()
        self.lib.orc_quat_to_euler(_fa(q), out)
        return list(out)

    def transform_time(self, pos, vel, qc, qn, duration, frames, frame, pshift, pscale):
        out = (C.c_float * 12)()
        self.lib.orc_transform_time(_fa(pos), _fa(vel), _fa(qc), _fa(qn), float(duration), frames, frame,
                                    _fa(pshift), _fa(pscale), out)
        return list(out)

    def transform_mesh(self, vel, duration, frames, frame, pshift, pscale):
        out = (C.c_float * 12)()
        self.lib.orc_transform_mesh(_fa(vel), float(duration), frames, frame, _fa(pshift), _fa(pscale), out)
        return list(out)

    def color_to_float4(self, rgb):
        out = (C.c_float * 4)()
        self.lib.orc_color_to_float4(_fa(rgb), out)
        return list(out)

    def color_to_uchar4(self, rgb):
        out = (C.c_uint8 * 4)()
        self.lib.orc_color_to_uchar4(_fa(rgb), out)
        return list(out)

    def xorwow_uniforms(self, seed, subsequence, offset, n):
        out = np.zeros(n, np.float32)
        self.lib.orc_xorwow_uniforms(seed, subsequence, offset, n, out.ctypes.data)
        return out

    def xorwow_state(self, seed, subsequence, offset):
        out = np.zeros(6, np.uint32)
        self.lib.orc_xorwow_state(seed, subsequence, offset, out.ctypes.data)
        return out

    def xform_point(self, m, p):
        out = (C.c_float * 3)()
        self.lib.orc_xform_point(_fa(m), _fa(p), out)
        return list(out)

    def xform_sphere(self, m, c, r):
        out = (C.c_float * 4)()
        self.lib.orc_xform_sphere(_fa(m), _fa(c), float(r), out)
        return list(out)

    def xform_normal(self, m, n):
        out = (C.c_float * 3)()
        self.lib.orc_xform_normal(_fa(m), _fa(n), out)
        return list(out)

    def morton30(self, x, y, z):
        return self.lib.orc_morton30(float(x), float(y), float(z))

    def num_threads(self):
        return self.lib.orc_num_threads()

    def hw_threads(self):
        self.lib.orc_hw_threads.restype = C.c_int
        return self.lib.orc_hw_threads()

    def set_num_threads(self, n):
        self.lib.orc_set_num_threads.argtypes = [C.c_int]
        self.lib.orc_set_num_threads(int(n))

    def closest_hit_address(self):
        """Address of orc_closest_hit: the traversal callback of oracle/ref_shader.cpp."""
        return C.cast(self.lib.orc_closest_hit, C.c_void_p)


class Ref:
    """oracle/_ref/librt_ref.so — the reference's own header + host functions (see oracle/ref_shim.cu)."""

    def __init__(self, lib):
        self.lib = lib
        f_p = C.POINTER(C.c_float)
        lib.ref_construct_transform.argtypes = [f_p, f_p, f_p, f_p]
        lib.ref_color_to_float4.argtypes = [f_p, f_p]
        lib.ref_color_to_uchar4.argtypes = [f_p, C.c_void_p]
        lib.ref_normalize.argtypes = [f_p, f_p]
        lib.ref_camera_configure.argtypes = [f_p, f_p, f_p, C.c_int, f_p]
        lib.ref_slerp.argtypes = [f_p, f_p, C.c_float, f_p]
        lib.ref_quat_to_euler.argtypes = [f_p, f_p]
        lib.ref_random_space_vectors.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_size_t, C.c_void_p]
        lib.ref_constants.argtypes = [f_p]

    def construct_transform(self, shift, rot, scale):
        out = (C.c_float * 12)()
        self.lib.ref_construct_transform(_fa(shift), _fa(rot), _fa(scale), out)
        return list(out)

    def color_to_float4(self, rgb):
        out = (C.c_float * 4)()
        self.lib.ref_color_to_float4(_fa(rgb), out)
        return list(out)

    def color_to_uchar4(self, rgb):
        out = (C.c_uint8 * 4)()
        self.lib.ref_color_to_uchar4(_fa(rgb), out)
        return list(out)

    def camera_configure(self, center, target, up, is_opengl):
        out = (C.c_float * 12)()
        self.lib.ref_camera_configure(_fa(center), _fa(target), _fa(up), int(bool(is_opengl)), out)
        return list(out)

    def slerp(self, q1, q2, t):
        out = (C.c_float * 4)()
        self.lib.ref_slerp(_fa(q1), _fa(q2), float(t), out)
        return list(out)

    def quat_to_euler(self, q):
        out = (C.c_float * 3)()
        self.lib.ref_quat_to_euler(_fa(q), out)
        return list(out)

    def random_space_vectors(self, seed, subsequence, offset, n):
        out = np.zeros((n, 3), np.float32)
        self.lib.ref_random_space_vectors(seed, subsequence, offset, n, out.ctypes.data)
        return out

    def constants(self):
        out = (C.c_float * 3)()
        self.lib.ref_constants(out)
        return list(out)


class RefShader:
    """oracle/_ref/librt_ref_shader.so — the reference's UNMODIFIED shader/Shader.cu (raygen, closest-hit, miss
    programs) and include/Util/ColorRamp.cuh compiled as host code (oracle/ref_shader.cpp); its optixTrace
    delegates the closest-hit query to the oracle."""

    def __init__(self, lib, oracle: Oracle):
        self.lib = lib
        self.orc = oracle
        lib.ref_shader_render.argtypes = [C.POINTER(OrcScene), C.c_void_p, C.c_void_p, C.POINTER(OrcCamera), C.POINTER(OrcParams),
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        lib.ref_bake_color_ramp.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p]
        lib.ref_preset_name.argtypes = [C.c_char_p]
        lib.ref_preset_name.restype = C.c_char_p
        lib.ref_shader_set_rsqrt_mode.argtypes = [C.c_int]

    def trace_depth(self):
        return self.lib.ref_shader_trace_depth()

    def render(self, world, cam, width, height, spp, seed, background=(0.7, 0.8, 0.9), sample_offset=0, accel=0, rsqrt_mode=0):
        """spp launches of the reference pipeline; returns linear radiance sum, last launch's colour / AOV buffers."""
        sc, keep = Oracle.pack_scene(world)
        acc = None
        if accel:
            self.orc.lib.orc_accel_create.restype = C.c_void_p
            self.orc.lib.orc_accel_create.argtypes = [C.POINTER(OrcScene)]
            acc = C.c_void_p(self.orc.lib.orc_accel_create(C.byref(sc)))
        oc = OrcCamera(tuple(cam.center), tuple(cam.U), tuple(cam.V), tuple(cam.W))
        p = OrcParams(width, height, spp, self.trace_depth(), seed, sample_offset, (C.c_float * 3)(*background), 0)
        n = width * height
        out = {k: np.zeros((n, 4), np.float32) for k in ("accum", "color", "albedo", "normal")}
        seg = C.c_uint64(0)
        self.lib.ref_shader_set_rsqrt_mode(int(rsqrt_mode))
        rc = self.lib.ref_shader_render(C.byref(sc), acc, self.orc.closest_hit_address(), C.byref(oc), C.byref(p),
                                        out["accum"].ctypes.data, out["color"].ctypes.data, out["albedo"].ctypes.data,
                                        out["normal"].ctypes.data, C.byref(seg))
        self.lib.ref_shader_set_rsqrt_mode(0)
        if acc:
            self.orc.lib.orc_accel_destroy.argtypes = [C.c_void_p]
            self.orc.lib.orc_accel_destroy(acc)
        assert rc == 0, rc
        out["segments"] = seg.value
        return out

    def bake_color_ramp(self, preset, count):
        out = np.zeros((count, 3), np.float32)
        rc = self.lib.ref_bake_color_ramp(preset.encode(), count, out.ctypes.data)
        assert rc == 0
        return out

    def preset_name(self, preset):
        return self.lib.ref_preset_name(preset.encode()).decode()


_oracle = None


def load() -> Oracle:
    global _oracle
    if _oracle is None:
        if not os.path.exists(LIB):
            build()
        _oracle = Oracle(C.CDLL(LIB))
    return _oracle


def load_ref():
    if not os.path.exists(REF_LIB):
        if os.path.isdir("/root/reference/include/Global"):
            build()
    if not os.path.exists(REF_LIB):
        return None
    return Ref(C.CDLL(REF_LIB))


def load_ref_shader():
    if not os.path.exists(REF_SHADER_LIB):
        if os.path.isdir("/root/reference/shader"):
            build()
    if not os.path.exists(REF_SHADER_LIB):
        return None
    return RefShader(C.CDLL(REF_SHADER_LIB), load())
