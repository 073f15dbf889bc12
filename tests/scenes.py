"""Scene descriptions shared by the parity tests: the same description is (a) flattened to world
space with numpy fp32 following spec A.1 and handed to the CPU oracle, and (b) uploaded through the
C ABI to the CUDA path."""
from __future__ import annotations

import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
IDENT = np.array([1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0], np.float32)
f32 = np.float32


class SceneDesc:
    def __init__(self, roughs, metals):
        self.groups = []
        self.instances = []
        self.roughs = np.ascontiguousarray(roughs, np.float32).reshape(-1, 3)
        self.metals = np.ascontiguousarray(metals, np.float32).reshape(-1, 4)

    def add_spheres(self, centers, radii):
        self.groups.append(dict(kind=0, a=np.ascontiguousarray(centers, np.float32).reshape(-1, 3),
                                b=np.ascontiguousarray(radii, np.float32).reshape(-1)))
        return len(self.groups) - 1

    def add_triangles(self, verts, normals):
        self.groups.append(dict(kind=1, a=np.ascontiguousarray(verts, np.float32).reshape(-1, 3),
                                b=np.ascontiguousarray(normals, np.float32).reshape(-1, 3)))
        return len(self.groups) - 1

    def add_instance(self, group, transform=None, mat_type=0, mat_index=0, mat_stride=0):
        t = IDENT.copy() if transform is None else np.asarray(transform, np.float32).reshape(12).copy()
        self.instances.append(dict(group=group, m=t, mat_type=mat_type, mat_index=mat_index, mat_stride=mat_stride))
        return len(self.instances) - 1

    def prim_count(self, inst):
        g = self.groups[inst["group"]]
        return len(g["b"]) if g["kind"] == 0 else len(g["a"]) // 3


def _is_identity(m):
    return bool(m[0] == 1 and m[1] == 0 and m[2] == 0 and m[4] == 0 and m[5] == 1 and m[6] == 0 and m[8] == 0 and
                m[9] == 0 and m[10] == 1)


def _xform_points(m, p):
    # ((m0*x + m1*y) + m2*z) + m3 in fp32, two roundings per multiply-add (spec A.1)
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    out = np.empty_like(p)
    for r in range(3):
        out[:, r] = ((m[4 * r] * x + m[4 * r + 1] * y) + m[4 * r + 2] * z) + m[4 * r + 3]
    return out


def _xform_normals(m, n):
    x, y, z = n[:, 0], n[:, 1], n[:, 2]
    v = np.empty_like(n)
    for r in range(3):
        v[:, r] = (m[4 * r] * x + m[4 * r + 1] * y) + m[4 * r + 2] * z
    l2 = (v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]) + v[:, 2] * v[:, 2]
    inv = f32(1.0) / np.sqrt(l2)
    out = v * inv[:, None]
    tiny = l2 <= f32(1e-6) * f32(1e-6)
    out[tiny] = np.array([0, 0, 1], np.float32)
    return out


def flatten_world(desc: SceneDesc, compat=False):
    """World-space scene for the oracle; global prim id = running index in instance order.
    compat=True reproduces what the reference's shader reads: sphere shading uses
    the OBJECT-space centre/radius (`sphere_shade`), triangle vertex normals are not rotated."""
    kind_index, mat, spheres, tv, tn, shade = [], [], [], [], [], []
    ns = nt = 0
    for inst in desc.instances:
        g = desc.groups[inst["group"]]
        m = inst["m"]
        ident = _is_identity(m)
        if g["kind"] == 0:
            n = len(g["b"])
            if ident:
                c = g["a"] + np.array([m[3], m[7], m[11]], np.float32)[None, :]
                r = g["b"].copy()
            else:
                c = _xform_points(m, g["a"])
                s = np.sqrt((m[0] * m[0] + m[4] * m[4]) + m[8] * m[8])
                r = g["b"] * f32(s)
            spheres.append(np.concatenate([c, r[:, None]], 1).astype(np.float32))
            shade.append(np.concatenate([g["a"], g["b"][:, None]], 1).astype(np.float32))
            kind_index.append(np.arange(ns, ns + n, dtype=np.uint32))
            ns += n
        else:
            n = len(g["a"]) // 3
            v = _xform_points(m, g["a"])
            nn = g["b"] if (ident or compat) else _xform_normals(m, g["b"])
            tv.append(v.reshape(n, 9))
            tn.append(nn.reshape(n, 9).astype(np.float32))
            kind_index.append(np.arange(nt, nt + n, dtype=np.uint32) | np.uint32(0x80000000))
            nt += n
        k = np.arange(n, dtype=np.uint64)
        mi = (np.uint64(inst["mat_index"]) + k * np.uint64(inst["mat_stride"])).astype(np.uint32)
        mat.append(mi | np.uint32(inst["mat_type"] << 31))
    cat = lambda xs, shape, dt: (np.concatenate(xs).astype(dt) if xs else np.zeros(shape, dt))
    return dict(kind_index=cat(kind_index, (0,), np.uint32), mat=cat(mat, (0,), np.uint32),
                sphere_shade=cat(shade, (0, 4), np.float32) if compat else None,
                spheres=cat(spheres, (0, 4), np.float32), tri_verts=cat(tv, (0, 9), np.float32),
                tri_normals=cat(tn, (0, 9), np.float32),
                roughs=desc.roughs if len(desc.roughs) else np.zeros((1, 3), np.float32),
                metals=desc.metals if len(desc.metals) else np.zeros((1, 4), np.float32))


class GpuScene:
    """Uploads a SceneDesc through the C ABI (device memory held by torch tensors)."""

    def __init__(self, pkg, desc: SceneDesc, ctx=None, device="cuda:0", options=None):
        import torch
        self.torch = torch
        self.pkg = pkg
        self.dev = torch.device(device)
        self.ctx = ctx or pkg.Context(self.dev.index or 0)
        self.scene = pkg.Scene(self.ctx)
        self.keep = []
        r = torch.from_numpy(desc.roughs).to(self.dev)
        m = torch.from_numpy(desc.metals).to(self.dev)
        self.keep += [r, m]
        self.scene.set_materials(r.data_ptr() if len(desc.roughs) else None, len(desc.roughs),
                                 m.data_ptr() if len(desc.metals) else None, len(desc.metals))
        for g in desc.groups:
            a = torch.from_numpy(g["a"]).to(self.dev).contiguous()
            b = torch.from_numpy(g["b"]).to(self.dev).contiguous()
            self.keep += [a, b]
            n = len(g["b"]) if g["kind"] == 0 else len(g["a"]) // 3
            if g["kind"] == 0:
                self.scene.add_spheres(a.data_ptr() if n else None, b.data_ptr() if n else None, n)
            else:
                self.scene.add_triangles(a.data_ptr() if n else None, b.data_ptr() if n else None, n)
        for inst in desc.instances:
            self.scene.add_instance(inst["group"], list(inst["m"]), inst["mat_type"], inst["mat_index"], inst["mat_stride"])
        for k, v in (options or {}).items():
            self.scene.set_option(k, v)
        self.scene.build()

    def render(self, cam, width, height, spp, max_depth, seed, background=(0.7, 0.8, 0.9), sample_offset=0, flags=0,
               want=("accum", "albedo", "normal", "prim_id"), accum_init=None):
        torch = self.torch
        n = width * height
        bufs = {
            "accum": torch.zeros((n, 4), dtype=torch.float32, device=self.dev) if "accum" in want else None,
            "albedo": torch.zeros((n, 4), dtype=torch.float32, device=self.dev) if "albedo" in want else None,
            "normal": torch.zeros((n, 4), dtype=torch.float32, device=self.dev) if "normal" in want else None,
            "prim_id": torch.zeros(n, dtype=torch.int32, device=self.dev) if "prim_id" in want else None,
        }
        if accum_init is not None:
            bufs["accum"].copy_(torch.from_numpy(accum_init).to(self.dev))
        ptr = lambda k: bufs[k].data_ptr() if bufs[k] is not None else None
        self.scene.render(cam, width, height, spp, max_depth, seed, background, sample_offset, flags,
                          ptr("accum"), ptr("albedo"), ptr("normal"), ptr("prim_id"),
                          stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        out = {k: (v.cpu().numpy() if v is not None else None) for k, v in bufs.items()}
        if out["prim_id"] is not None:
            out["prim_id"] = out["prim_id"].view(np.uint32)
        st = self.scene.stats()
        out["segments"] = st.segments
        out["stats"] = st
        return out

    def trace(self, rays, skip=None):
        torch = self.torch
        r = torch.from_numpy(np.ascontiguousarray(rays, np.float32)).to(self.dev)
        n = len(rays)
        t = torch.zeros(n, dtype=torch.float32, device=self.dev)
        ids = torch.zeros(n, dtype=torch.int32, device=self.dev)
        sk = torch.from_numpy(np.ascontiguousarray(skip, np.uint32).view(np.int32)).to(self.dev) if skip is not None else None
        self.scene.trace_rays(r.data_ptr(), sk.data_ptr() if sk is not None else None, n, t.data_ptr(), ids.data_ptr(),
                              stream=torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        return t.cpu().numpy(), ids.cpu().numpy().view(np.uint32)

    def close(self):
        self.scene.close()


# ------------------------------------------------------------------ canned scenes
CONFIG_ROUGHS = [[0.65, 0.05, 0.05], [0.73, 0.73, 0.73], [0.12, 0.45, 0.15], [0.70, 0.60, 0.50]]
CONFIG_METALS = [[0.8, 0.85, 0.88, 0.0]]


def load_particle_frame0():
    with open(os.path.join(GOLDEN, "particle_frame0.json")) as f:
        return json.load(f)


def config1_scene(pkg, with_ground=True):
    """BASELINE config 1: files/particle-short.vtk.series frame 0 as spheres r=0.05 (+ the r=1000
    ground sphere of files/config.json:20-30, shift z = -1000.5), terrain ramp materials."""
    fr = load_particle_frame0()
    pts = np.asarray(fr["points"], np.float32)
    ids = np.asarray(fr["id"], np.int64)
    n = len(pts)
    ramp = pkg.bake_color_ramp("terrain", int(fr["max_particle_count"]))
    roughs = np.concatenate([np.asarray(CONFIG_ROUGHS, np.float32), ramp])
    d = SceneDesc(roughs, CONFIG_METALS)
    if with_ground:
        g = d.add_spheres([[0, 0, 0]], [1000.0])
        d.add_instance(g, pkg.construct_transform([0, 0, -1000.5], [0, 0, 0], [1, 1, 1]), 0, 3, 0)
    for i in range(n):  # one instance per particle (RendererTime.cu:116-130), material id + #config roughs
        g = d.add_spheres([[0, 0, 0]], [0.05])
        t = pkg.construct_transform(list(pts[i]), [0, 0, 0], [1, 1, 1])
        d.add_instance(g, t, 0, int(ids[i]) + len(CONFIG_ROUGHS), 0)
    return d


def config_camera(pkg, api="VK"):
    """files/config.json loop-data camera: centre (5,0,0) -> target 0, up (0,0,1); api VK flips up."""
    return pkg.camera_configure([5.0, 0.0, 0.0], [0.0, 0.0, 0.0], [0.0, 0.0, 1.0], api != "OGL")


def random_scene(pkg, synth, n_spheres, n_tris=0, seed=1, radius=None, metal_every=0, with_walls=True,
                 with_ground=False, fuzz=0.0):
    """Spheres uniform in the unit cube (+ optional random triangles, cube walls, huge ground)."""
    r = radius if radius is not None else 0.35 / max(1.0, n_spheres ** (1.0 / 3.0))
    centers, radii = synth.uniform_spheres(n_spheres, r, seed)
    ramp = pkg.bake_color_ramp("terrain", max(n_spheres, 1))
    roughs = np.concatenate([np.asarray(CONFIG_ROUGHS, np.float32), ramp])
    d = SceneDesc(roughs, [[0.8, 0.85, 0.88, fuzz], [0.9, 0.6, 0.2, 0.3]])
    if with_ground:
        g = d.add_spheres([[0, 0, 0]], [1000.0])
        d.add_instance(g, pkg.construct_transform([0, 0, -1000.5], [0, 0, 0], [1, 1, 1]), 0, 3, 0)
    if n_spheres:
        if metal_every:
            k = np.arange(n_spheres)
            gm = d.add_spheres(centers[k % metal_every == 0], radii[k % metal_every == 0])
            d.add_instance(gm, None, 1, 0, 0)
            gr = d.add_spheres(centers[k % metal_every != 0], radii[k % metal_every != 0])
            d.add_instance(gr, None, 0, 4, 1)
        else:
            g = d.add_spheres(centers, radii)
            d.add_instance(g, None, 0, 4, 1)
    if with_walls:
        v, nn = synth.box_walls(True)
        g = d.add_triangles(v, nn)
        d.add_instance(g, None, 0, 1, 0)
    if n_tris:
        idx = np.arange(n_tris * 9, dtype=np.uint64)
        u = synth.uniform01(seed + 77, idx).reshape(n_tris, 3, 3)
        base = (u[:, 0, :] - f32(0.5)) * f32(0.9)
        e = (u[:, 1:, :] - f32(0.5)) * f32(4.0 * r)
        verts = np.stack([base, base + e[:, 0], base + e[:, 1]], 1).reshape(-1, 3).astype(np.float32)
        fn = np.cross(e[:, 0], e[:, 1])
        fn = fn / np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-20)
        norms = np.repeat(fn.astype(np.float32), 3, axis=0)
        g = d.add_triangles(verts, norms)
        # rotated + translated instance of the soup exercises the non-identity path
        d.add_instance(g, pkg.construct_transform([0.01, -0.02, 0.03], [10, 20, 30], [1, 1, 1]), 1, 1, 0)
    return d


# ------------------------------------------------------------------ scenes of the reference pin (tests/golden/ref_shader.npz)
REF_PIN_SIZE = (64, 36, 4, 20261104)  # width, height, spp, seed; depth = the reference's rayTraceDepth (5)


def instanced_scene(pkg, synth):
    """Translated / rotated / uniformly scaled sphere instances and a rotated triangle instance: the scene on
    which the reference's object-space shading differs from world-space shading."""
    ramp = pkg.bake_color_ramp("plasma", 64)
    d = SceneDesc(np.concatenate([np.asarray(CONFIG_ROUGHS, np.float32), ramp]), [[0.8, 0.85, 0.88, 0.0], [0.9, 0.6, 0.2, 0.25]])
    c, r = synth.uniform_spheres(40, 0.06, 11)
    g = d.add_spheres(c * f32(0.8), r)
    d.add_instance(g, pkg.construct_transform([0.05, -0.1, 0.02], [0, 0, 0], [1, 1, 1]), 0, 4, 1)
    d.add_instance(g, pkg.construct_transform([-0.2, 0.15, 0.3], [25, -40, 10], [0.5, 0.5, 0.5]), 1, 1, 0)
    g1 = d.add_spheres([[0, 0, 0]], [0.2])
    d.add_instance(g1, pkg.construct_transform([0.1, 0.3, -0.25], [0, 0, 0], [1.5, 1.5, 1.5]), 1, 0, 0)
    v, nn = synth.box_walls(True)
    gw = d.add_triangles(v, nn)
    d.add_instance(gw, pkg.construct_transform([0, 0, 0], [0, 0, 15], [1.2, 1.2, 1.2]), 0, 1, 0)
    return d


def ref_pin_cases(pkg, synth):
    """(name, SceneDesc, camera, compat): compat=True feeds the reference's shader the object-space data it
    really reads; False feeds it the world-space data of our default shading."""
    cam1 = config_camera(pkg, "VK")
    cam2 = pkg.camera_configure([2.2, 0.3, 0.4], [0, 0, 0], [0, 0, 1], False)
    c1 = config1_scene(pkg)
    rnd = random_scene(pkg, synth, 300, n_tris=40, seed=3, metal_every=5, fuzz=0.2, with_ground=False)
    inst = instanced_scene(pkg, synth)
    return [("c1_compat", c1, cam1, True), ("c1_world", c1, cam1, False), ("rnd_world", rnd, cam2, False),
            ("rnd_compat", rnd, cam2, True), ("inst_compat", inst, cam2, True), ("inst_world", inst, cam2, False)]


# ------------------------------------------------------------------ loader-independent readers (tests only)
def parse_vtk_particles_py(path):
    """A second, pure-Python reader of NetDEM's legacy-ASCII particle / wall files (POINTS + POINT_DATA arrays), written
    from the file format alone and sharing no code with librtb200_io.so: the pin for rt_io_particles_load and the source of
    the frames tests/test_gpu_configs.py assembles without the product loaders.  Values go text -> double -> float32, as
    the reference's vtkDataArray::GetTuple + static_cast<float> does (src/Util/VTKReaderImpl.cpp:176-214)."""
    tok = open(path).read().split("\n")
    i, n, out = 0, 0, {}

    def block(first, count, width):
        vals, j = [], first
        while len(vals) < count * width:
            vals += tok[j].split()
            j += 1
        return np.array(vals[:count * width], dtype=np.float64).reshape(count, width), j

    while i < len(tok):
        w = tok[i].split()
        if len(w) >= 3 and w[0] == "POINTS":
            n = int(w[1])
            out["pos"], i = block(i + 1, n, 3)
        elif len(w) >= 3 and w[0] == "SCALARS":
            width = int(w[3]) if len(w) > 3 else 1
            out[w[1]], i = block(i + 2, n, width)      # skips the LOOKUP_TABLE line
        elif len(w) >= 3 and w[0] == "VECTORS":
            out[w[1]], i = block(i + 1, n, 3)
        else:
            i += 1
    res = dict(pos=out["pos"].astype(np.float32), vel=out["vel"].astype(np.float32), quat=out["quat"].astype(np.float32),
               id=out["id"][:, 0].astype(np.uint64), shape_id=out["shape_id"][:, 0].astype(np.uint64),
               size=out["size"][:, 0].astype(np.float32) if "size" in out else None)
    return res


def parse_series_py(directory, name):
    """.vtk.series (JSON) without the product's JSON parser: file paths and the per-file durations the frame count is
    computed from (time[i+1] - time[i] in float32; the last file repeats the previous duration, RendererTime.cu:60-75)."""
    d = json.load(open(os.path.join(directory, name)))
    files = [os.path.join(directory, f["name"]) for f in d["files"]]
    times = np.float32([f["time"] for f in d["files"]])
    return files, times
