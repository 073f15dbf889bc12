"""BASELINE configs 2 and 3 on the reference's SHIPPED data (packed by scripts/pack_ref_data.py into the git-ignored
data/ref_files/, which travels to the GPU box): three frames of each series rendered by the headless driver
(config.json -> loaders -> C ABI) equal the CPU oracle's render of the same scene assembled independently here —
linear radiance bit for bit, at reduced resolution so that the oracle finishes in seconds.  The full-size runs
(1080p / 16 spp / depth 8 over all 401 files; 4K / 64 spp over the 71 Mesh-mode files) are `scripts/gpu.sh configs`,
recorded under profiles/.  Reference flow: src/Global/RendererTime.cu:420-525, src/Global/RendererMesh.cu:363-444."""
import importlib
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
DATA = os.path.join(ROOT, "data", "ref_files")
DRIVER = os.path.join(ROOT, "nvidia-optix-ray-tracer_b200", "rtb200_render")
W, H = 320, 180

needs_data = pytest.mark.skipif(not os.path.isfile(os.path.join(DATA, "config_c2.json")),
                                reason="data/ref_files absent: run scripts/pack_ref_data.py in the build container")


@pytest.fixture(scope="module")
def hio(pkg):
    m = importlib.import_module("nvidia-optix-ray-tracer_b200.hostio")
    if not os.path.exists(m.LIB_PATH):
        m.build()
    if not os.path.exists(DRIVER):
        subprocess.run(["make", "-C", os.path.join(ROOT, "nvidia-optix-ray-tracer_b200", "host")], check=True)
    return m


def run_driver(cfg, out, *extra):
    # the packed configs hold paths relative to the repository root, like the reference's "../files/"
    r = subprocess.run([DRIVER, cfg, "--out", out, "--width", str(W), "--height", str(H), "--format", "pfm", "--quiet"] + list(extra),
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def base_scene(pkg, c, roughs):
    sd = scenes.SceneDesc(roughs, c["metals"])
    for k in range(len(c["sphere_center_radius"])):
        g = sd.add_spheres([c["sphere_center_radius"][k][:3]], [c["sphere_center_radius"][k][3]])
        sd.add_instance(g, c["sphere_transform"][k], int(c["sphere_mat_type"][k]), int(c["sphere_mat_index"][k]), 0)
    return sd


@needs_data
def test_c2_time_series_with_walls_matches_oracle(tmp_path, pkg, hio, oracle):
    """files/particle.vtk.series + wall.vtk.series, rough ramp + metal ground, 16 spp, depth 8, BVH rebuilt per frame:
    files 0, 200, 400.  The camera-facing wall is dropped so that the particles are visible.  That wall is
    shape 2, not shape 3: the wall files carry NetDEM's identity quaternion as (w, x, y, z) = (1, 0, 0, 0); the reference
    reads the components as (x, y, z, w) (VTKReaderImpl.cpp:194-199) and its slerp returns them shifted by one place
    (RendererTime.cu:297-340, aggregate order w, x, y, z into a float4), so every wall is turned by 180 degrees about y
    and the x = -0.5 plane (shape 2) ends up at x = +0.5, facing the camera at (5, 0, 0).  Reproduced bit-exactly."""
    SPP, DEPTH, SEED, STRIDE = 16, 8, 7, 200
    out = str(tmp_path) + "/out"
    s = run_driver("data/ref_files/config_c2.json", out, "--spp", str(SPP), "--depth", str(DEPTH), "--seed", str(SEED),
                   "--extra-series", "wall.vtk.series", "--drop-shape", "2", "--file-stride", str(STRIDE), "--frame-stride", "9")
    assert s["mode"] == "time" and s["frames"] == 3
    c = hio.load_config(os.path.join(DATA, "config_c2.json"))
    assert int(c["sphere_mat_type"][0]) == 1                      # the ground sphere is the metal
    shapes = hio.load_shapes_dir(DATA + "/shape/separated/")
    series = [hio.load_series(DATA + "/", n) for n in ("particle.vtk.series", "wall.vtk.series")]
    assert len(series[0]["files"]) == 401 and len(series[1]["files"]) == 401
    # the driver sizes the ramp by the largest id (+1) over EVERY file of every series
    max_id = 0
    for sr in series:
        for f in sr["files"]:
            p = hio.load_particles(f)
            max_id = max(max_id, len(p["id"]), int(p["id"].max()) + 1 if len(p["id"]) else 0)
    roughs = np.concatenate([c["roughs"], pkg.bake_color_ramp(c["preset"], max_id)])
    cam = pkg.camera_configure(c["camera_center"], c["camera_target"], c["up_direction"], c["api"] != "OGL")
    n_tri_total = 0
    for ft, fi in enumerate((0, 200, 400)):
        dur = series[0]["durations"][fi]
        frames = hio.frames_for_file(dur, c["fps"], c["render_speed_ratio"])
        sd = base_scene(pkg, c, roughs)
        gs = [sd.add_triangles(sh["vertices"], sh["vertex_normals"]) for sh in shapes]
        n_inst = 0
        for sr in series:
            # file 0 of both series comes from the pure-Python parser of tests/scenes.py, not from the product loaders
            # (tests/test_io.py pins the loaders against the same parser on every file used here)
            rd = scenes.parse_vtk_particles_py if fi == 0 else hio.load_particles
            p = rd(sr["files"][fi])
            nxt = rd(sr["files"][min(fi + 1, 400)])
            for k in range(len(p["id"])):
                if int(p["shape_id"][k]) == 2:
                    continue
                qn = nxt["quat"][k] if k < len(nxt["id"]) else p["quat"][k]
                t = pkg.transform_time(p["pos"][k], p["vel"][k], p["quat"][k], qn, dur, frames, 0, c["particle_shift"], c["particle_scale"])
                sd.add_instance(gs[int(p["shape_id"][k])], t, 0, len(c["roughs"]) + int(p["id"][k]), 0)
                n_inst += 1
        world = scenes.flatten_world(sd)
        n_tri_total += n_inst
        o = oracle.render(world, cam, W, H, SPP, DEPTH, SEED, sample_offset=ft * SPP, accel=1)
        got = hio.read_pfm(out + f"/frame_{fi:05d}_000.pfm")
        want = (o["accum"][:, :3] * np.float32(1.0 / SPP)).reshape(H, W, 3)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), fi
        assert (o["prim_id"] != 0xFFFFFFFF).mean() > 0.3
    assert n_tri_total > 25 + 200 + 400                           # particle counts grow 25 -> 500 over the series


@needs_data
def test_c3_mesh_series_cache_then_render_matches_oracle(tmp_path, pkg, hio, oracle):
    """particle_mesh-short.vtk.series (synthesised, see scripts/pack_ref_data.py): "cache": true writes the cache files,
    the second run renders from them (Main.cu:23-25); 64 spp, depth 5: files 0, 35, 70."""
    SPP, DEPTH, SEED = 64, 5, 3
    d = str(tmp_path) + "/"
    cfg = json.load(open(os.path.join(DATA, "config_c3.json")))
    cfg["cache-path"] = d + "cache/"
    cfg["cache"] = True
    json.dump(cfg, open(d + "c3.json", "w"))
    r = subprocess.run([DRIVER, d + "c3.json"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    assert len(os.listdir(d + "cache")) == 72                     # 71 particle<i>.cache + metadata.cache
    cfg["cache"] = False
    json.dump(cfg, open(d + "c3.json", "w"))
    s = run_driver(d + "c3.json", d + "out", "--spp", str(SPP), "--depth", str(DEPTH), "--seed", str(SEED), "--file-stride", "35",
                   "--frame-stride", "9")
    assert s["mode"] == "mesh" and s["frames"] == 3
    c = hio.load_config(d + "c3.json")
    metas = [hio.load_mesh_cache(d + f"cache/particle{i}.cache") for i in range(71)]
    max_id = max(max(len(m["id"]), int(m["id"].max()) + 1) for m in metas)
    roughs = np.concatenate([c["roughs"], pkg.bake_color_ramp(c["preset"], max_id)])
    cam = pkg.camera_configure(c["camera_center"], c["camera_target"], c["up_direction"], c["api"] != "OGL")
    series = hio.load_series(DATA + "/", "particle_mesh-short.vtk.series")
    for ft, fi in enumerate((0, 35, 70)):
        m = metas[fi]
        dur = series["durations"][fi]
        frames = hio.frames_for_file(dur, c["fps"], c["render_speed_ratio"])
        sd = base_scene(pkg, c, roughs)
        for k in range(len(m["id"])):
            a, b = int(m["vertex_offset"][k]), int(m["vertex_offset"][k + 1])
            g = sd.add_triangles(m["vertices"][a:b], m["normals"][a:b])
            t = pkg.transform_mesh(m["vel"][k], dur, frames, 0, c["particle_shift"], c["particle_scale"])
            sd.add_instance(g, t, 0, len(c["roughs"]) + int(m["id"][k]), 0)
        o = oracle.render(scenes.flatten_world(sd), cam, W, H, SPP, DEPTH, SEED, sample_offset=ft * SPP, accel=1)
        got = hio.read_pfm(d + f"out/frame_{fi:05d}_000.pfm")
        want = (o["accum"][:, :3] * np.float32(1.0 / SPP)).reshape(H, W, 3)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), fi
        assert (o["prim_id"] != 0xFFFFFFFF).mean() > 0.3
    shutil.rmtree(d + "cache")
