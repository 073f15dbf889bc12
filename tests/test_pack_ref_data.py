"""scripts/pack_ref_data.py: the one-strip encoding used to synthesise BASELINE config 3's Mesh-mode series (every closed
particle surface as ONE TRIANGLE_STRIPS cell, joined by index-degenerate bridge triangles) keeps every triangle with its
winding, and the binary VTK it writes is read back by the product's Mesh-mode loader (src/Util/VTKReaderImpl.cpp:24-137)."""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import pack_ref_data as prd  # noqa: E402

OCTA = np.float32([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]) * 0.05
FACES = [(0, 2, 4), (2, 1, 4), (1, 3, 4), (3, 0, 4), (2, 0, 5), (1, 2, 5), (3, 1, 5), (0, 3, 5)]  # outward winding


def strip_triangles(strip):
    out = []
    for k in range(len(strip) - 2):
        a, b, c = strip[k], strip[k + 1], strip[k + 2]
        t = (a, c, b) if k & 1 else (a, b, c)
        if len(set(t)) == 3:
            out.append(prd._canon(t))
    return sorted(out)


def test_stripify_keeps_every_triangle_and_its_winding():
    s = prd.stripify(FACES)
    assert strip_triangles(s) == sorted(prd._canon(f) for f in FACES)
    # two disconnected triangles and a duplicate: bridges only, nothing lost
    tris = [(0, 1, 2), (3, 4, 5), (0, 1, 2)]
    assert strip_triangles(prd.stripify(tris)) == sorted(prd._canon(t) for t in tris)
    rng = np.random.default_rng(3)
    soup = [tuple(int(x) for x in rng.choice(12, 3, replace=False)) for _ in range(40)]
    assert strip_triangles(prd.stripify(soup)) == sorted(prd._canon(t) for t in soup)


def test_weld_merges_bit_identical_vertices_only():
    v = np.float32([[0, 0, 0], [1, 0, 0], [0, 0, 0], [-0.0, 0, 0], [1, 0, np.float32(1e-30)]])
    pts, idx = prd.weld(v)
    assert len(pts) == 3 and idx[0] == idx[2] == idx[3] and idx[1] != idx[4]


def test_written_mesh_file_is_read_back_by_the_loader(tmp_path, pkg):
    hio = importlib.import_module("nvidia-optix-ray-tracer_b200.hostio")
    strip = np.asarray(prd.stripify(FACES))
    pts = [OCTA + np.float32([0.2 * k, 0, 0]) for k in range(3)]
    strips = [strip + 6 * k for k in range(3)]
    path = str(tmp_path / "m.vtk")
    prd.write_mesh_vtk_binary(path, np.concatenate(pts), strips, [4, 0, 2], np.float32([[0, 0, -1], [0, 0, -2], [0, 0, -3]]))
    m = hio.load_mesh_vtk(path)
    assert list(m["id"]) == [4, 0, 2] and np.array_equal(m["vel"][:, 2], np.float32([-1, -2, -3]))
    off = m["vertex_offset"]
    for k in range(3):
        v = m["vertices"][int(off[k]):int(off[k + 1])].reshape(-1, 3, 3)
        area = np.linalg.norm(np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0]), axis=1)
        real = v[area > 0]
        assert len(real) == 8                                   # the 8 faces; the rest are zero-area bridges the reader keeps
        c = real.mean(axis=(0, 1))
        n = m["normals"][int(off[k]):int(off[k + 1])].reshape(-1, 3, 3)[area > 0]
        assert np.all(((real - c) * n).sum(axis=2) > 0)        # smooth point normals point outward
