"""End-to-end through the headless driver (rtb200_render: config.json -> loaders -> C ABI -> PFM): the
frames it writes equal the CPU oracle's render of the same scene assembled independently in Python
(reference flow: Main.cu + RendererTime / RendererMesh commitRendererData / startRender)."""
import importlib
import json
import os
import subprocess

import numpy as np
import pytest

import scenes
import test_io as tio

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
DRIVER = os.path.join(ROOT, "nvidia-optix-ray-tracer_b200", "rtb200_render")
W, H, SPP, DEPTH, SEED = 200, 120, 2, 5, 9


@pytest.fixture(scope="module")
def hio(pkg):
    m = importlib.import_module("nvidia-optix-ray-tracer_b200.hostio")
    if not os.path.exists(m.LIB_PATH):
        m.build()
    if not os.path.exists(DRIVER):
        subprocess.run(["make", "-C", os.path.join(ROOT, "nvidia-optix-ray-tracer_b200", "host")], check=True)
    return m


def config_for(d, mesh, series):
    c = json.loads(json.dumps(tio.CONFIG))
    c.update({"mesh": mesh, "series-path": d, "series-name": series, "cache-path": d + "cache/", "stl-path": d + "shapes/"})
    c["spheres"] = c["spheres"][:1]
    c["spheres"][0]["mat-type"] = "METAL"
    c["spheres"][0]["mat-index"] = 0
    c["metals"][0]["fuzz"] = 0.0
    c["loop-data"].update({"camera-center": [2.0, 0.0, 0.3], "particle-shift": [0.0, 0.05, 0.0]})
    return c


def add_config_triangles(sd, c):
    """config "triangles": one world-space group + identity instance each, after the config spheres (SBT order)."""
    for k in range(len(c["triangle_mat_type"])):
        g = sd.add_triangles(c["triangle_vertices"][k], c["triangle_normals"][k])
        sd.add_instance(g, None, int(c["triangle_mat_type"][k]), int(c["triangle_mat_index"][k]), 0)


def run_driver(cfg_path, out, *extra):
    cmd = [DRIVER, cfg_path, "--out", out, "--width", str(W), "--height", str(H), "--spp", str(SPP), "--depth", str(DEPTH),
           "--seed", str(SEED), "--format", "pfm", "--quiet"] + list(extra)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def make_time_inputs(d):
    rng = np.random.default_rng(11)
    os.makedirs(d + "p")
    os.makedirs(d + "shapes")
    tio.write_stl(d + "shapes/shape_0000000000.stl", [tio.OCTA[list(f)] * 2.0 for f in tio.OCTA_FACES])
    tio.write_stl(d + "shapes/shape_0000000001.stl", [tio.OCTA[list(f)][[0, 2, 1]] * 3.0 for f in tio.OCTA_FACES])
    files, parts = [], []
    for i, n in enumerate((5, 6)):
        pts = rng.uniform(-0.45, 0.45, (n, 3))
        q = rng.normal(size=(n, 4))
        q /= np.linalg.norm(q, axis=1, keepdims=True)
        vel = rng.normal(size=(n, 3)) * 2
        ids, shape = list(rng.permutation(n + 2)[:n]), list(rng.integers(0, 2, n))
        tio.write_time_vtk(d + f"p/p_{i}.vtk", pts, ids, q, vel, shape, size=[0.2] * n)
        files.append({"name": f"p/p_{i}.vtk", "time": 0.01 * i})
        parts.append(dict(pos=pts.astype(np.float32), quat=q.astype(np.float32), vel=vel.astype(np.float32), id=ids, shape=shape))
    json.dump({"file-series-version": "1.0", "files": files}, open(d + "t.vtk.series", "w"))
    return parts


def test_time_mode_driver_matches_oracle(tmp_path, pkg, hio, oracle):
    d = str(tmp_path) + "/"
    parts = make_time_inputs(d)
    cfg = config_for(d, False, "t.vtk.series")
    json.dump(cfg, open(d + "config.json", "w"))
    summary = run_driver(d + "config.json", d + "out", "--frame-stride", "4")
    assert summary["mode"] == "time" and summary["frames"] == 6          # 2 files x frames {0, 4, 8} of 9
    c = hio.load_config(d + "config.json")
    shapes = hio.load_shapes_dir(d + "shapes/")
    max_id = max(max(p["id"]) for p in parts) + 1
    roughs = np.concatenate([c["roughs"], pkg.bake_color_ramp("terrain", max_id)])
    cam = pkg.camera_configure(c["camera_center"], c["camera_target"], c["up_direction"], True)
    dur = np.float32(0.01)
    frames = hio.frames_for_file(dur, 240, 4)
    ft = 0
    for fi, p in enumerate(parts):
        nxt = parts[min(fi + 1, len(parts) - 1)]
        for f in (0, 4, 8):
            sd = scenes.SceneDesc(roughs, c["metals"])
            g0 = sd.add_spheres([[0, 0, 0]], [1000.0])
            sd.add_instance(g0, c["sphere_transform"][0], 1, 0, 0)
            add_config_triangles(sd, c)
            gs = [sd.add_triangles(s["vertices"], s["vertex_normals"]) for s in shapes]
            for k in range(len(p["id"])):
                qn = nxt["quat"][k] if k < len(nxt["id"]) else p["quat"][k]
                t = pkg.transform_time(p["pos"][k], p["vel"][k], p["quat"][k], qn, dur, frames, f, c["particle_shift"], c["particle_scale"])
                sd.add_instance(gs[p["shape"][k]], t, 0, len(c["roughs"]) + int(p["id"][k]), 0)
            o = oracle.render(scenes.flatten_world(sd), cam, W, H, SPP, DEPTH, SEED, sample_offset=ft * SPP)
            got = hio.read_pfm(d + f"out/frame_{fi:05d}_{f:03d}.pfm")
            want = (o["accum"][:, :3] * np.float32(1.0 / SPP)).reshape(H, W, 3)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (fi, f)
            assert (o["prim_id"] != 0xFFFFFFFF).mean() > 0.2
            ft += 1


def test_spheres_mapping_driver_matches_oracle(tmp_path, pkg, hio, oracle):
    """BASELINE config 1 mapping: particles as analytic spheres (centre = POINTS + motion, r = size / 2)."""
    d = str(tmp_path) + "/"
    parts = make_time_inputs(d)
    cfg = config_for(d, False, "t.vtk.series")
    json.dump(cfg, open(d + "config.json", "w"))
    run_driver(d + "config.json", d + "out", "--spheres", "--max-frames", "1", "--aov")
    c = hio.load_config(d + "config.json")
    p = parts[0]
    roughs = np.concatenate([c["roughs"], pkg.bake_color_ramp("terrain", max(max(q["id"]) for q in parts) + 1)])
    sd = scenes.SceneDesc(roughs, c["metals"])
    g0 = sd.add_spheres([[0, 0, 0]], [1000.0])
    sd.add_instance(g0, c["sphere_transform"][0], 1, 0, 0)
    add_config_triangles(sd, c)
    for k in range(len(p["id"])):
        t = pkg.transform_time(p["pos"][k], p["vel"][k], p["quat"][k], p["quat"][k], np.float32(0.01), 9, 0, c["particle_shift"], c["particle_scale"])
        g = sd.add_spheres([[0, 0, 0]], [0.1])
        sd.add_instance(g, [1, 0, 0, t[3], 0, 1, 0, t[7], 0, 0, 1, t[11]], 0, len(c["roughs"]) + int(p["id"][k]), 0)
    cam = pkg.camera_configure(c["camera_center"], c["camera_target"], c["up_direction"], True)
    o = oracle.render(scenes.flatten_world(sd), cam, W, H, SPP, DEPTH, SEED)
    got = hio.read_pfm(d + "out/frame_00000_000.pfm")
    want = (o["accum"][:, :3] * np.float32(1.0 / SPP)).reshape(H, W, 3)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    # --aov: the denoiser guide buffers of the reference (Shader.cu:270-272) and the primary-hit ids, as files
    alb, nrm = hio.read_pfm(d + "out/frame_00000_000_albedo.pfm"), hio.read_pfm(d + "out/frame_00000_000_normal.pfm")
    assert np.array_equal(alb.view(np.uint32), o["albedo"][:, :3].reshape(H, W, 3).view(np.uint32))
    assert np.array_equal(nrm.view(np.uint32), o["normal"][:, :3].reshape(H, W, 3).view(np.uint32))
    ids = np.fromfile(d + "out/frame_00000_000_primid.u32", dtype="<u4")
    assert np.array_equal(ids, o["prim_id"]) and len(set(ids.tolist())) > 4
    assert (ids == 1).any() or (ids == 2).any()   # the config triangles (global ids 1 and 2) are visible


def test_mesh_mode_driver_cache_then_render(tmp_path, pkg, hio, oracle):
    d = str(tmp_path) + "/"
    os.makedirs(d + "mesh")
    files, metas = [], []
    for i in range(2):
        parts = [(tio.OCTA * 3 + [0.3 * k - 0.3, 0.1 * i, 0.0], tio.OCTA_STRIP) for k in range(3)]
        vels = [[0, 0, -1.0 - k] for k in range(3)]
        tio.write_mesh_vtk(d + f"mesh/m_{i}.vtk", parts, ids=[2, 0, 4], vels=vels)
        files.append({"name": f"mesh/m_{i}.vtk", "time": 0.01 * i})
    json.dump({"file-series-version": "1.0", "files": files}, open(d + "m.vtk.series", "w"))
    cfg = config_for(d, True, "m.vtk.series")
    cfg["cache"] = True                                   # first run: write the cache files and exit (Main.cu:23-25)
    json.dump(cfg, open(d + "config.json", "w"))
    r = subprocess.run([DRIVER, d + "config.json"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and sorted(os.listdir(d + "cache")) == ["metadata.cache", "particle0.cache", "particle1.cache"]
    cfg["cache"] = False
    json.dump(cfg, open(d + "config.json", "w"))
    summary = run_driver(d + "config.json", d + "out", "--frame-stride", "8")
    assert summary["mode"] == "mesh" and summary["frames"] == 4
    c = hio.load_config(d + "config.json")
    roughs = np.concatenate([c["roughs"], pkg.bake_color_ramp("terrain", 5)])
    cam = pkg.camera_configure(c["camera_center"], c["camera_target"], c["up_direction"], True)
    ft = 0
    for fi in range(2):
        m = hio.load_mesh_cache(d + f"cache/particle{fi}.cache")
        for f in (0, 8):
            sd = scenes.SceneDesc(roughs, c["metals"])
            g0 = sd.add_spheres([[0, 0, 0]], [1000.0])
            sd.add_instance(g0, c["sphere_transform"][0], 1, 0, 0)
            add_config_triangles(sd, c)
            for k in range(3):
                a, b = int(m["vertex_offset"][k]), int(m["vertex_offset"][k + 1])
                g = sd.add_triangles(m["vertices"][a:b], m["normals"][a:b])
                t = pkg.transform_mesh(m["vel"][k], np.float32(0.01), 9, f, c["particle_shift"], c["particle_scale"])
                sd.add_instance(g, t, 0, len(c["roughs"]) + int(m["id"][k]), 0)
            o = oracle.render(scenes.flatten_world(sd), cam, W, H, SPP, DEPTH, SEED, sample_offset=ft * SPP)
            got = hio.read_pfm(d + f"out/frame_{fi:05d}_{f:03d}.pfm")
            want = (o["accum"][:, :3] * np.float32(1.0 / SPP)).reshape(H, W, 3)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (fi, f)
            ft += 1


def _dir_bytes(path):
    return {f: open(os.path.join(path, f), "rb").read() for f in sorted(os.listdir(path))}


def test_frame_parallel_ranks_write_the_serial_runs_bytes(tmp_path, pkg, hio):
    """rtb200_render --ranks N: the files of the series are dealt to N forked processes (file k -> rank k % N, one GPU each, or
    a shared GPU when the box has fewer).  Every output file must be byte-identical to the serial run's (frame-
    parallel output files must be byte-identical to the serial run) — sample offsets follow the global frame index, which a
    rank derives from the durations of the files it skips.  Also through --rank R with an external launcher."""
    d = str(tmp_path) + "/"
    make_time_inputs(d)
    json.dump(config_for(d, False, "t.vtk.series"), open(d + "config.json", "w"))
    serial = run_driver(d + "config.json", d + "out1", "--frame-stride", "4")
    assert serial["frames"] == 6 and serial["ranks"] == 1
    want = _dir_bytes(d + "out1")
    assert len(want) == 6
    # forked by the driver itself: one summary line per rank, then the parent's
    cmd = [DRIVER, d + "config.json", "--out", d + "out2", "--width", str(W), "--height", str(H), "--spp", str(SPP), "--depth", str(DEPTH),
           "--seed", str(SEED), "--format", "pfm", "--quiet", "--frame-stride", "4", "--ranks", "2"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [json.loads(x) for x in r.stdout.strip().splitlines()]
    assert lines[-1]["ranks"] == 2 and lines[-1]["failed"] == 0
    per_rank = sorted((x["rank"], x["frames"]) for x in lines[:-1])
    assert per_rank == [(0, 3), (1, 3)]
    assert _dir_bytes(d + "out2") == want
    # three ranks for two files (one rank idles), launched one by one as an external launcher would
    for rk in range(3):
        s = run_driver(d + "config.json", d + "out3", "--frame-stride", "4", "--ranks", "3", "--rank", str(rk))
        assert s["rank"] == rk and s["frames"] == (3 if rk < 2 else 0)
    assert _dir_bytes(d + "out3") == want
