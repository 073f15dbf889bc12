"""Host-side ingest / output library: config.json, .vtk.series, legacy VTK,
STL, cache format, image writers.  Inputs are written by the tests themselves in the reference's
formats; when the reference checkout is present (build container only) the shipped files are parsed
too and compared with the committed golden values."""
import importlib
import json
import os
import struct

import numpy as np
import pytest

REF = "/root/reference/files"
HERE = os.path.dirname(os.path.abspath(__file__))

CONFIG = {
    "mesh": False, "series-path": "../files/", "series-name": "particle.vtk.series", "cache-path": "../cache/",
    "stl-path": "../files/shape/separated/", "cache": False, "debug-mode": False, "cache-process-thread-count": 8,
    "particle-material-preset": "terrain",
    "roughs": [{"albedo": [0.65, 0.05, 0.05]}, {"albedo": [0.73, 0.73, 0.73]}],
    "metals": [{"albedo": [0.8, 0.85, 0.88], "fuzz": 0.25}],
    "spheres": [{"center": [0.0, 0.0, 0.0], "radius": 1000.0, "mat-type": "ROUGH", "mat-index": 1,
                 "shift": [0.0, 0.0, -1000.5], "rotate": [0.0, 0.0, 0.0], "scale": [1.0, 1.0, 1.0]},
                {"center": [1.0, 2.0, 3.0], "radius": 0.5, "mat-type": "METAL", "mat-index": 0,
                 "shift": [0.1, 0.2, 0.3], "rotate": [10.0, 20.0, 30.0], "scale": [2.0, 2.0, 2.0]}],
    "triangles": [{"vertices": [[-0.6, -0.6, -0.45], [0.6, -0.6, -0.45], [0.6, 0.6, -0.45]],
                   "normals": [[0, 0, 1], [0, 0, 1], [0, 0, 1]], "mat-type": "METAL", "mat-index": 0},
                  {"vertices": [[-0.6, -0.6, -0.45], [0.6, 0.6, -0.45], [-0.6, 0.6, -0.45]],
                   "normals": [[0, 0, 1], [0.1, 0, 0.9], [0, 0.1, 0.9]], "mat-type": "ROUGH", "mat-index": 1}],
    "loop-data": {"api": "VK", "window-width": 1200, "window-height": 800, "fps": 240, "camera-center": [5.0, 0.0, 0.0],
                  "camera-target": [0.0, 0.0, 0.0], "up-direction": [0.0, 0.0, 1.0], "camera-pitch-limit-degree": 85.0,
                  "camera-speed-stride": 0.002, "camera-initial-speed-ratio": 10, "mouse-sensitivity": 0.002,
                  "render-speed-ratio": 4, "particle-shift": [0.0, 0.1, 0.0], "particle-scale": [1.0, 1.0, 1.0]},
}


@pytest.fixture(scope="module")
def hio(pkg):
    m = importlib.import_module("nvidia-optix-ray-tracer_b200.hostio")
    if not os.path.exists(m.LIB_PATH):
        m.build()
    return m


def test_io_library_exports_every_declared_symbol(hio):
    import re
    hdr = open(os.path.join(os.path.dirname(HERE), "include", "rtb200_io.h")).read()
    declared = set(re.findall(r"\b(rt_io_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(hio.IO_SYMBOLS)
    for s in declared:
        assert hasattr(hio.lib(), s), s


def test_config_schema(hio, pkg, built_lib):
    c = hio.parse_config(json.dumps(CONFIG))
    assert (c["mesh"], c["cache"], c["debug_mode"], c["thread_count"]) == (0, 0, 0, 8)
    assert c["series_name"] == "particle.vtk.series" and c["stl_path"] == "../files/shape/separated/" and c["preset"] == "terrain"
    assert c["api"] == "VK" and c["flip_up"] == 1 and (c["window_width"], c["window_height"], c["fps"]) == (1200, 800, 240)
    assert c["render_speed_ratio"] == 4 and c["camera_initial_speed_ratio"] == 10
    assert np.allclose(c["roughs"], [[0.65, 0.05, 0.05], [0.73, 0.73, 0.73]]) and np.allclose(c["metals"], [[0.8, 0.85, 0.88, 0.25]])
    assert np.array_equal(c["sphere_mat_type"], [pkg.RT_MAT_ROUGH, pkg.RT_MAT_METAL]) and np.array_equal(c["sphere_mat_index"], [1, 0])
    assert np.allclose(c["sphere_center_radius"], [[0, 0, 0, 1000], [1, 2, 3, 0.5]])
    # the sphere transform is constructTransformMatrix(shift, rotate, scale), bit for bit
    for i, s in enumerate(CONFIG["spheres"]):
        want = np.float32(pkg.construct_transform(s["shift"], s["rotate"], s["scale"]))
        assert np.array_equal(c["sphere_transform"][i].view(np.uint32), want.view(np.uint32))
    # "triangles": documented (docs/configuration.md:232-247), dead in the reference, parsed here
    assert c["triangle_vertices"].shape == (2, 3, 3) and np.array_equal(c["triangle_mat_type"], [pkg.RT_MAT_METAL, pkg.RT_MAT_ROUGH])
    assert np.array_equal(c["triangle_mat_index"], [0, 1])
    assert np.allclose(c["triangle_vertices"][1], CONFIG["triangles"][1]["vertices"]) and np.allclose(c["triangle_normals"][1], CONFIG["triangles"][1]["normals"])
    no_tri = {k: v for k, v in CONFIG.items() if k != "triangles"}
    assert len(hio.parse_config(json.dumps(no_tri))["triangle_mat_type"]) == 0 and len(hio.parse_config(json.dumps(dict(CONFIG, triangles=[])))["triangle_mat_type"]) == 0
    for bad_tri in ({"vertices": [[0, 0, 0]], "normals": [[0, 0, 1]] * 3, "mat-type": "ROUGH", "mat-index": 0},
                    {"vertices": [[0, 0, 0]] * 3, "normals": [[0, 0]] * 3, "mat-type": "ROUGH", "mat-index": 0},
                    {"vertices": [[0, 0, 0]] * 3, "normals": [[0, 0, 1]] * 3, "mat-index": 0}):
        with pytest.raises(hio.IoError):
            hio.parse_config(json.dumps(dict(CONFIG, triangles=[bad_tri])))
    assert np.allclose(c["particle_shift"], [0, 0.1, 0]) and np.allclose(c["camera_center"], [5, 0, 0])
    assert (c["spp"], c["max_depth"], c["seed"], c["frames"], c["output"]) == (1, 5, 0, 0, "")   # additive keys default
    c2 = hio.parse_config(json.dumps(dict(CONFIG, spp=16, seed=7, output="out/", frames=3, **{"max-depth": 8})))
    assert (c2["spp"], c2["max_depth"], c2["seed"], c2["frames"], c2["output"]) == (16, 8, 7, 3, "out/")
    ogl = json.loads(json.dumps(CONFIG))
    ogl["loop-data"]["api"] = "OGL"
    assert hio.parse_config(json.dumps(ogl))["flip_up"] == 0


def test_config_errors(hio, tmp_path):
    bad = dict(CONFIG)
    del bad["series-name"]                       # every key is required (missing key aborts the reference)
    with pytest.raises(hio.IoError) as e:
        hio.parse_config(json.dumps(bad))
    assert e.value.code == 2 and "series-name" in str(e.value)
    with pytest.raises(hio.IoError) as e:
        hio.parse_config('{"mesh": tru')
    assert e.value.code == 2
    inv = json.loads(json.dumps(CONFIG))
    inv["loop-data"]["api"] = "METAL"
    with pytest.raises(hio.IoError) as e:
        hio.parse_config(json.dumps(inv))
    assert e.value.code == 3
    with pytest.raises(hio.IoError) as e:
        hio.load_config(str(tmp_path / "nope.json"))
    assert e.value.code == 1


def test_series_durations_and_frame_counts(hio, tmp_path):
    d = str(tmp_path) + "/"
    times = [0.0, 0.01, 0.03, 0.035]
    json.dump({"file-series-version": "1.0", "files": [{"name": f"p/p_{i}.vtk", "time": t} for i, t in enumerate(times)]},
              open(d + "s.vtk.series", "w"))
    s = hio.load_series(d, "s.vtk.series")
    assert s["files"] == [d + f"p/p_{i}.vtk" for i in range(4)]
    t32 = np.float32(times)
    assert np.array_equal(s["durations"], np.float32([t32[1] - t32[0], t32[2] - t32[1], t32[3] - t32[2], t32[3] - t32[2]]))
    json.dump({"file-series-version": "1.0", "files": [{"name": "only.vtk", "time": 3.0}]}, open(d + "one.series", "w"))
    assert hio.load_series(d, "one.series")["durations"][0] == 1000.0
    # (size_t)(0.01f * float(240 * 4)) = 9 (RendererTime.cu:427-428)
    assert hio.frames_for_file(np.float32(0.01), 240, 4) == 9
    assert hio.frames_for_file(0.0, 240, 4) == 0
    with pytest.raises(hio.IoError):
        hio.load_series(d, "missing.series")


def write_time_vtk(path, pts, ids, quat, vel, shape, size=None, extra_first=True):
    n = len(pts)
    with open(path, "w") as f:
        f.write("# vtk DataFile Version 2.0\nGenerated by test\nASCII \nDATASET POLYDATA\n")
        f.write(f"POINTS {n} double\n" + "".join("%r %r %r \n" % tuple(map(float, p)) for p in pts))
        f.write(f"VERTICES {n} {2 * n}\n" + "".join(f"1 {i}\n" for i in range(n)))
        f.write(f"POINT_DATA {n}\n")
        f.write("SCALARS id int 1\nLOOKUP_TABLE default\n" + "".join(f"{i}\n" for i in ids))
        if extra_first:
            f.write("SCALARS damp_numerical double 1\nLOOKUP_TABLE default\n" + "0.5\n" * n)
        f.write("SCALARS quat double 4\nLOOKUP_TABLE default\n" + "".join("%r %r %r %r\n" % tuple(map(float, q)) for q in quat))
        f.write("VECTORS vel double\n" + "".join("%r %r %r\n" % tuple(map(float, v)) for v in vel))
        f.write("VECTORS spin double\n" + "0 0 1e-3\n" * n)
        f.write("SCALARS shape_id int 1\nLOOKUP_TABLE default\n" + "".join(f"{s}\n" for s in shape))
        if size is not None:
            f.write("SCALARS size double 1\nLOOKUP_TABLE default\n" + "".join(f"{s}\n" for s in size))
        f.write("SCALARS mask int 1\nLOOKUP_TABLE default\n" + "0\n" * n)


def test_time_mode_particle_file(hio, tmp_path):
    rng = np.random.default_rng(3)
    n = 7
    pts, quat, vel = rng.uniform(-1, 1, (n, 3)), rng.normal(size=(n, 4)), rng.normal(size=(n, 3))
    ids, shape = [5, 0, 3, 1, 6, 2, 4], [0, 1, 0, 1, 1, 7, 2]
    p = str(tmp_path / "t.vtk")
    write_time_vtk(p, pts, ids, quat, vel, shape, size=[0.1] * n)
    d = hio.load_particles(p)
    assert np.array_equal(d["pos"], pts.astype(np.float32)) and np.array_equal(d["quat"], quat.astype(np.float32))
    assert np.array_equal(d["vel"], vel.astype(np.float32))
    assert list(d["id"]) == ids and list(d["shape_id"]) == shape and np.allclose(d["size"], 0.1)
    # a wall-style file has no 'size' array; a file without 'quat' is an error
    write_time_vtk(p, pts, ids, quat, vel, shape, size=None)
    assert hio.load_particles(p)["size"] is None
    txt = open(p).read().replace("SCALARS quat double 4", "SCALARS quat_renamed double 4")
    open(p, "w").write(txt)
    with pytest.raises(hio.IoError) as e:
        hio.load_particles(p)
    assert e.value.code == 3 and "quat" in str(e.value)


def test_binary_vtk_equals_ascii(hio, tmp_path):
    rng = np.random.default_rng(5)
    n = 4
    pts, quat, vel = rng.uniform(-1, 1, (n, 3)), rng.normal(size=(n, 4)), rng.normal(size=(n, 3))
    ids, shape = [3, 1, 0, 2], [1, 0, 1, 0]
    a = str(tmp_path / "a.vtk")
    write_time_vtk(a, pts, ids, quat, vel, shape, extra_first=False)
    b = str(tmp_path / "b.vtk")
    with open(b, "wb") as f:
        f.write(b"# vtk DataFile Version 3.0\nbinary twin\nBINARY\nDATASET POLYDATA\n")
        f.write(b"POINTS %d double\n" % n + pts.astype(">f8").tobytes() + b"\n")
        cells = np.array([[1, i] for i in range(n)], ">i4")
        f.write(b"VERTICES %d %d\n" % (n, 2 * n) + cells.tobytes() + b"\n")
        f.write(b"POINT_DATA %d\n" % n)
        f.write(b"SCALARS id int 1\nLOOKUP_TABLE default\n" + np.array(ids, ">i4").tobytes() + b"\n")
        f.write(b"SCALARS quat double 4\nLOOKUP_TABLE default\n" + quat.astype(">f8").tobytes() + b"\n")
        f.write(b"VECTORS vel float\n" + vel.astype(">f4").tobytes() + b"\n")
        f.write(b"SCALARS shape_id int 1\nLOOKUP_TABLE default\n" + np.array(shape, ">i4").tobytes() + b"\n")
    da, db = hio.load_particles(a), hio.load_particles(b)
    for k in ("pos", "quat", "id", "shape_id"):
        assert np.array_equal(da[k], db[k]), k
    assert np.array_equal(db["vel"], vel.astype(np.float32))


OCTA = np.float64([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]) * 0.05
OCTA_FACES = [(0, 2, 4), (2, 1, 4), (1, 3, 4), (3, 0, 4), (2, 0, 5), (1, 2, 5), (3, 1, 5), (0, 3, 5)]  # outward winding


def write_stl(path, tris, name="netdem"):
    with open(path, "w") as f:
        f.write(f"solid {name}\n")
        for t in tris:
            f.write("facet normal 0 0 0\nouter loop\n" + "".join("vertex %r %r %r\n" % tuple(map(float, v)) for v in t) + "endloop\nendfacet\n")
        f.write(f"endsolid {name}\n")


def test_stl_shapes_sorted_and_outward_normals(hio, tmp_path):
    d = tmp_path / "shapes"
    d.mkdir()
    out_tris = [OCTA[list(f)] + [0.3, 0.2, 0.1] for f in OCTA_FACES]
    in_tris = [t[[0, 2, 1]] for t in out_tris]            # inward winding: normals must still point outward
    plane = [np.float64([[0, -1, -1], [0, 1, -1], [0, 1, 1]]), np.float64([[0, -1, -1], [0, 1, 1], [0, -1, 1]])]
    write_stl(str(d / "shape_0000000001.stl"), in_tris)
    write_stl(str(d / "shape_0000000000.stl"), out_tris)
    write_stl(str(d / "shape_0000000002.stl"), plane)
    (d / "notes.txt").write_text("not an stl")
    s = hio.load_shapes_dir(str(d) + "/")
    assert [os.path.basename(x["path"]) for x in s] == [f"shape_000000000{i}.stl" for i in range(3)]   # index = shape_id
    assert [len(x["face_normals"]) for x in s] == [8, 8, 2]
    c = np.float32([0.3, 0.2, 0.1])
    for x, tris in ((s[0], out_tris), (s[1], in_tris)):
        assert np.array_equal(x["vertices"], np.concatenate(tris).astype(np.float32))
        cen = x["vertices"].reshape(-1, 3, 3).mean(axis=1) - c
        assert np.all((x["face_normals"] * cen).sum(axis=1) > 0)                    # outward
        assert np.allclose(np.linalg.norm(x["face_normals"], axis=1), 1, atol=1e-6)
        assert np.array_equal(x["vertex_normals"].reshape(-1, 3, 3), np.repeat(x["face_normals"][:, None], 3, 1))
    assert np.allclose(np.abs(s[2]["face_normals"]), [[1, 0, 0], [1, 0, 0]])        # open surface keeps its winding normal
    # binary twin of shape 0
    tri = np.concatenate(out_tris).astype("<f4").reshape(-1, 9)
    with open(d / "bin.stl", "wb") as f:
        f.write(b"binary stl".ljust(80, b" ") + struct.pack("<I", len(tri)))
        for t in tri:
            f.write(np.zeros(3, "<f4").tobytes() + t.tobytes() + b"\0\0")
    b = hio.load_stl(str(d / "bin.stl"))
    assert np.array_equal(b["vertices"], s[0]["vertices"]) and np.array_equal(b["face_normals"], s[0]["face_normals"])


def write_mesh_vtk(path, particles, ids, vels):
    """Mesh mode: each particle is ONE triangle-strip cell (docs/usage.md:15-20)."""
    pts, strips = [], []
    for verts, strip in particles:
        base = len(pts)
        pts += list(verts)
        strips.append([base + i for i in strip])
    with open(path, "w") as f:
        f.write("# vtk DataFile Version 3.0\nmesh\nASCII\nDATASET POLYDATA\n")
        f.write(f"POINTS {len(pts)} float\n" + "".join("%r %r %r\n" % tuple(float(x) for x in p) for p in pts))
        f.write(f"TRIANGLE_STRIPS {len(strips)} {sum(len(s) + 1 for s in strips)}\n")
        f.write("".join(" ".join(map(str, [len(s)] + s)) + "\n" for s in strips))
        f.write(f"CELL_DATA {len(strips)}\nSCALARS id int 1\nLOOKUP_TABLE default\n" + "".join(f"{i}\n" for i in ids))
        f.write("VECTORS vel float\n" + "".join("%r %r %r\n" % tuple(float(x) for x in v) for v in vels))


# an octahedron as one strip: 4 5 ... walks all 8 faces (degenerate-free strip over vertices 0..5)
OCTA_STRIP = [0, 2, 4, 1, 3, 5, 0, 2]  # triangles (0,2,4) (2,1,4)* (4,1,3) (1,5,3)* (3,5,0) (5,2,0)* -> 6 of 8 faces


def test_mesh_mode_strips_normals_and_cache_roundtrip(hio, tmp_path):
    p = str(tmp_path / "m0.vtk")
    parts = [(OCTA + [0.5, 0, 0], OCTA_STRIP), (OCTA * 2 + [-0.5, 0.1, 0], OCTA_STRIP[::-1])]
    write_mesh_vtk(p, parts, ids=[4, 9], vels=[[0, 0, -2], [0.5, 0, 0]])
    m = hio.load_mesh_vtk(p, write_cache_to=str(tmp_path / "particle0.cache"))
    assert list(m["id"]) == [4, 9] and np.array_equal(m["vel"], np.float32([[0, 0, -2], [0.5, 0, 0]]))
    assert list(m["vertex_offset"]) == [0, 18, 36]                                  # 6 triangles per 8-vertex strip
    # strip -> triangles with the odd ones flipped (VTKReaderImpl.cpp:103-110)
    v0 = (OCTA + [0.5, 0, 0]).astype(np.float32)
    want = []
    for k in range(6):
        a, b, c = OCTA_STRIP[k], OCTA_STRIP[k + 1], OCTA_STRIP[k + 2]
        want += [v0[a], v0[c], v0[b]] if k & 1 else [v0[a], v0[b], v0[c]]
    assert np.array_equal(m["vertices"][:18], np.float32(want))
    # smooth point normals: unit length, identical for every occurrence of a point, pointing away from the centre
    nrm = m["normals"]
    assert np.allclose(np.linalg.norm(nrm, axis=1), 1, atol=1e-5)
    for k in range(2):
        sl = slice(18 * k, 18 * k + 18)
        cen = np.float32([0.5, 0, 0] if k == 0 else [-0.5, 0.1, 0])
        assert np.all(((m["vertices"][sl] - cen) * nrm[sl]).sum(axis=1) > 0)
        for v in np.unique(m["vertices"][sl], axis=0):
            same = nrm[sl][(m["vertices"][sl] == v).all(axis=1)]
            assert np.all(same == same[0])
    # cache bytes: u64 n | per particle u64 id, float3 vel, u64 vertexCount, float3 v[], float3 n[]
    raw = open(tmp_path / "particle0.cache", "rb").read()
    assert struct.unpack_from("<Q", raw, 0)[0] == 2
    assert struct.unpack_from("<Q3fQ", raw, 8) == (4, 0.0, 0.0, -2.0, 18)
    assert np.array_equal(np.frombuffer(raw, "<f4", 54, 36).reshape(18, 3), m["vertices"][:18])
    assert len(raw) == 8 + 2 * (8 + 12 + 8 + 18 * 12 * 2)
    c = hio.load_mesh_cache(str(tmp_path / "particle0.cache"))
    for k in m:
        assert np.array_equal(m[k], c[k]), k
    open(tmp_path / "cut.cache", "wb").write(raw[:100])
    with pytest.raises(hio.IoError):
        hio.load_mesh_cache(str(tmp_path / "cut.cache"))


def test_mesh_mode_rejects_non_strip_cells(hio, tmp_path):
    p = str(tmp_path / "poly.vtk")
    open(p, "w").write("# vtk DataFile Version 3.0\nx\nASCII\nDATASET POLYDATA\nPOINTS 3 float\n0 0 0\n1 0 0\n0 1 0\n"
                       "POLYGONS 1 4\n3 0 1 2\nCELL_DATA 1\nSCALARS id int 1\nLOOKUP_TABLE default\n0\nVECTORS vel float\n0 0 0\n")
    with pytest.raises(hio.IoError) as e:
        hio.load_mesh_vtk(p)
    assert "illegal cell type" in str(e.value)


def test_cache_generate_whole_series(hio, tmp_path):
    d = str(tmp_path) + "/"
    os.makedirs(d + "mesh")
    files = []
    for i in range(3):
        parts = [(OCTA + [0.2 * k, 0, 0.01 * i], OCTA_STRIP) for k in range(2 + i)]
        write_mesh_vtk(d + f"mesh/m_{i}.vtk", parts, ids=list(range(2 + i)), vels=[[0, 0, -1]] * (2 + i))
        files.append({"name": f"mesh/m_{i}.vtk", "time": 0.01 * i})
    json.dump({"file-series-version": "1.0", "files": files}, open(d + "mesh.vtk.series", "w"))
    os.makedirs(d + "cache")
    open(d + "cache/stale.cache", "w").write("old")
    assert hio.cache_generate(d, "mesh.vtk.series", d + "cache/", 2) == 4
    assert sorted(os.listdir(d + "cache")) == ["metadata.cache", "particle0.cache", "particle1.cache", "particle2.cache"]
    assert open(d + "cache/metadata.cache").read() == "4" and hio.cache_read_max_cells(d + "cache/") == 4
    for i in range(3):
        a, b = hio.load_mesh_cache(d + f"cache/particle{i}.cache"), hio.load_mesh_vtk(d + f"mesh/m_{i}.vtk")
        assert all(np.array_equal(a[k], b[k]) for k in a)


def test_image_writers(hio, tmp_path):
    rng = np.random.default_rng(0)
    img = rng.uniform(0, 2, (5, 7, 4)).astype(np.float32)
    hio.write_pfm(str(tmp_path / "a.pfm"), img)
    raw = open(tmp_path / "a.pfm", "rb").read()
    assert raw.startswith(b"PF\n7 5\n-1.0\n")
    body = np.frombuffer(raw[len(b"PF\n7 5\n-1.0\n"):], "<f4").reshape(5, 7, 3)
    assert np.array_equal(body[::-1], img[:, :, :3])            # PFM stores the bottom row first
    assert np.array_equal(hio.read_pfm(str(tmp_path / "a.pfm")), img[:, :, :3])
    hio.write_pfm(str(tmp_path / "b.pfm"), img, flip_y=True)     # buffer row 0 = bottom (OGL convention)
    assert np.array_equal(hio.read_pfm(str(tmp_path / "b.pfm")), img[::-1, :, :3])
    u8 = rng.integers(0, 256, (5, 7, 4), dtype=np.uint8)
    hio.write_ppm(str(tmp_path / "a.ppm"), u8)
    raw = open(tmp_path / "a.ppm", "rb").read()
    assert raw.startswith(b"P6\n7 5\n255\n") and np.array_equal(np.frombuffer(raw[11:], np.uint8).reshape(5, 7, 3), u8[:, :, :3])
    hio.write_exr(str(tmp_path / "a.exr"), img)
    raw = open(tmp_path / "a.exr", "rb").read()
    assert struct.unpack_from("<II", raw, 0) == (20000630, 2)
    assert b"channels\0chlist\0" in raw and b"compression\0compression\0\x01\0\0\0\0" in raw and b"dataWindow\0box2i\0" in raw
    hdr_end = raw.index(b"screenWindowWidth\0float\0") + len(b"screenWindowWidth\0float\0") + 4 + 4 + 1
    offs = struct.unpack_from("<5Q", raw, hdr_end)
    assert offs[0] == hdr_end + 40 and all(offs[i + 1] - offs[i] == 8 + 16 * 7 for i in range(4))
    y, nbytes = struct.unpack_from("<iI", raw, offs[2])
    assert (y, nbytes) == (2, 16 * 7)
    line = np.frombuffer(raw, "<f4", 28, offs[2] + 8).reshape(4, 7)      # A, B, G, R planes
    assert np.array_equal(line[3], img[2, :, 0]) and np.array_equal(line[0], img[2, :, 3])
    assert len(raw) == offs[4] + 8 + 16 * 7


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout absent (GPU box)")
def test_shipped_reference_files(hio):
    c = hio.load_config(REF + "/config.json")
    assert c["mesh"] == 0 and c["preset"] == "terrain" and c["api"] == "VK" and len(c["roughs"]) == 4 and len(c["metals"]) == 1
    assert np.allclose(c["sphere_center_radius"], [[0, 0, 0, 1000]]) and np.allclose(c["sphere_transform"][0][[3, 7, 11]], [0, 0, -1000.5])
    s = hio.load_series(REF + "/", "particle-short.vtk.series")
    g = json.load(open(os.path.join(HERE, "golden", "particle_frame0.json")))
    assert len(s["files"]) == 71 and np.allclose(s["times"], np.float32(g["series_times"]))
    assert all(hio.frames_for_file(d, c["fps"], c["render_speed_ratio"]) in (9, 10) for d in s["durations"])
    p = hio.load_particles(s["files"][0])
    assert np.array_equal(p["pos"], np.float32(g["points"])) and np.array_equal(p["quat"], np.float32(g["quat"]))
    assert np.array_equal(p["vel"], np.float32(g["vel"])) and list(p["id"]) == g["id"] and list(p["shape_id"]) == g["shape_id"]
    assert np.allclose(p["size"], 0.1)
    w = hio.load_particles(REF + "/wall/wall_000000000000000.vtk")
    assert list(w["shape_id"]) == [2, 3, 4, 5, 6, 7] and w["size"] is None
    shapes = hio.load_shapes_dir(REF + "/shape/separated/")
    assert [len(x["face_normals"]) for x in shapes] == [252, 396, 2, 2, 2, 2, 2, 2]
    sph = shapes[1]
    cen = sph["vertices"].reshape(-1, 3, 3).mean(axis=1)
    assert np.all((sph["face_normals"] * cen).sum(axis=1) > 0)          # outward on the tessellated sphere
    assert abs(np.linalg.norm(sph["vertices"], axis=1).max() - 0.05) < 1e-3
    assert max(len(hio.load_particles(f)["id"]) for f in s["files"][::10]) <= g["max_particle_count"]


DATA = os.path.join(os.path.dirname(HERE), "data", "ref_files")


@pytest.mark.skipif(not os.path.isfile(os.path.join(DATA, "config_c2.json")), reason="data/ref_files absent: run scripts/pack_ref_data.py")
def test_particle_and_wall_loaders_equal_an_independent_parser(hio):
    """tests/test_gpu_configs.py assembles its oracle scenes with the product's own loaders, which pins
    driver <-> render but not the loaders.  Here every array rt_io_particles_load returns for the files that test renders
    (particle and wall files 0 / 1, 200 / 201, 400 of BASELINE config 2, plus every 40th file) is compared bit for bit with a
    pure-Python parser of the same text (tests/scenes.py), and the series' file list / times with json.load."""
    import scenes
    for name, sub in (("particle.vtk.series", "particle"), ("wall.vtk.series", "wall")):
        s = hio.load_series(DATA + "/", name)
        files, times = scenes.parse_series_py(DATA, name)
        assert [os.path.normpath(f) for f in s["files"]] == [os.path.normpath(f) for f in files]
        assert np.array_equal(np.float32(s["times"]), times)
        for fi in sorted(set([0, 1, 200, 201, 400] + list(range(0, 401, 40)))):
            a, b = hio.load_particles(s["files"][fi]), scenes.parse_vtk_particles_py(files[fi])
            for k in ("pos", "vel", "quat", "id", "shape_id"):
                assert a[k].dtype == b[k].dtype and np.array_equal(a[k], b[k]), (sub, fi, k)
            assert (a["size"] is None) == (b["size"] is None)
            if b["size"] is not None:
                assert np.array_equal(a["size"], b["size"])
