#!/usr/bin/env python
"""Generates the committed fixtures under tests/golden/ from the reference checkout.

Run HERE (the build container, /root/reference present):  python tests/golden/make_golden.py
The GPU box has no /root/reference; tests only read the JSON/NPZ files this script wrote.

  particle_frame0.json   — files/particle/particle_000000000000000.vtk (BASELINE config 1 input):
                           POINTS, id, quat, vel, shape_id parsed with a throw-away parser, plus the
                           maximum particle count over files/particle-short.vtk.series.
  ref_host_math.json     — outputs of the REFERENCE's own code (oracle/_ref, compiled from
                           /root/reference by oracle/Makefile) on fixed inputs: constructTransformMatrix,
                           colorToFloat4/Uchar4, SDL_GraphicsWindowConfigureCamera, slerp, quatToEuler,
                           randomSpaceVector over host-compiled cuRAND XORWOW, constants.
  ref_shader.npz         — outputs of the REFERENCE's own device programs (shader/Shader.cu, unmodified, compiled
                           as host code by oracle/Makefile -> oracle/_ref/librt_ref_shader.so; its optixTrace asks
                           the oracle for the closest hit): per scene of scenes.ref_pin_cases() the sum over 4
                           launches of the linear radiance returned to raygen, plus the sRGB colour buffer and the
                           (all-zero: SURVEY §C quirk 9) albedo / normal buffers of the last launch for one scene.
  ref_color_ramp.npz     — include/Util/ColorRamp.cuh bakeColorRamp for every preset (same library).
  oracle_images.npz      — oracle renders of small canned scenes (prim ids + linear accum) so that
                           a change in the oracle itself is caught by the CPU test suite.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def parse_vtk_points(path):
    toks = open(path).read().split()
    i = toks.index("POINTS")
    n = int(toks[i + 1])
    pts = np.array(toks[i + 3:i + 3 + 3 * n], dtype=np.float64).reshape(n, 3)

    def field(name, comps, kind):
        j = [k for k in range(len(toks) - 1) if toks[k] in ("SCALARS", "VECTORS") and toks[k + 1] == name][0]
        if toks[j] == "SCALARS":
            start = j + 4 + 2  # SCALARS name type ncomp LOOKUP_TABLE default
        else:
            start = j + 3      # VECTORS name type
        return np.array(toks[start:start + comps * n], dtype=kind).reshape(n, comps)

    return dict(points=pts, id=field("id", 1, np.int64)[:, 0], quat=field("quat", 4, np.float64),
                vel=field("vel", 3, np.float64), shape_id=field("shape_id", 1, np.int64)[:, 0])


def main():
    import oracle_lib
    oracle_lib.build()
    ref = oracle_lib.load_ref()
    assert ref is not None, "oracle/_ref missing (need /root/reference)"

    # ---- particle frame 0
    series = json.load(open(os.path.join(REF, "files/particle-short.vtk.series")))
    first = os.path.join(REF, "files", series["files"][0]["name"])
    fr = parse_vtk_points(first)
    max_count = 0
    for e in series["files"]:
        toks = open(os.path.join(REF, "files", e["name"])).read(400).split()
        max_count = max(max_count, int(toks[toks.index("POINTS") + 1]))
    json.dump(dict(source="files/" + series["files"][0]["name"], points=fr["points"].tolist(), id=fr["id"].tolist(),
                   quat=fr["quat"].tolist(), vel=fr["vel"].tolist(), shape_id=fr["shape_id"].tolist(),
                   max_particle_count=max_count, series_times=[e["time"] for e in series["files"]]),
              open(os.path.join(HERE, "particle_frame0.json"), "w"))

    # ---- reference host maths
    rng = np.random.default_rng(20261104)
    cases = dict(transform=[], color=[], camera=[], slerp=[], euler=[], rsv=[])
    tf_inputs = [([0, 0, -1000.5], [0, 0, 0], [1, 1, 1]), ([0.1, -0.2, 0.3], [10, 20, 30], [1, 1, 1]),
                 ([1, 2, 3], [-45, 90, 180], [2, 2, 2]), ([0, 0, 0], [33.3, -12.5, 271.0], [0.5, 1.5, 2.5])]
    for _ in range(12):
        tf_inputs.append((rng.uniform(-2, 2, 3).round(4).tolist(), rng.uniform(-180, 180, 3).round(3).tolist(),
                          rng.uniform(0.2, 3, 3).round(3).tolist()))
    for s, r, sc in tf_inputs:
        cases["transform"].append(dict(shift=s, rotate=r, scale=sc, out=[float(np.float32(x)) for x in ref.construct_transform(s, r, sc)]))
    for c in [[0.7, 0.8, 0.9], [0, 0, 0], [1, 1, 1], [0.002, 0.0031308, 0.004], [-0.5, 1.5, 0.5], [0.18, 0.5, 0.999]] + \
            rng.uniform(0, 1, (10, 3)).round(5).tolist():
        cases["color"].append(dict(rgb=c, f4=[float(x) for x in ref.color_to_float4(c)], u4=ref.color_to_uchar4(c)))
    cams = [([5, 0, 0], [0, 0, 0], [0, 0, 1]), ([2.2, 0, 0], [0, 0, 0], [0, 0, 1]), ([1, 2, 3], [0.1, -0.2, 0.3], [0, 1, 0]),
            ([-3, 0.5, 2], [0, 0, 0.4], [0.1, 0.2, 0.97])]
    for c, t, u in cams:
        for ogl in (0, 1):
            cases["camera"].append(dict(center=c, target=t, up=u, opengl=ogl, out=[float(x) for x in ref.camera_configure(c, t, u, ogl)]))
    quats = np.asarray(fr["quat"], np.float32)
    pairs = [(quats[i].tolist(), quats[(i * 7 + 3) % len(quats)].tolist(), float(t)) for i, t in
             zip(range(10), [0, 0.1, 0.25, 0.5, 0.75, 1.0, 0.33, 0.9, 0.01, 0.66])]
    pairs += [([0, 0, 0, 1], [0, 0, 0, 1], 0.5), ([0, 0, 0, 1], [0, 0, 0, -1], 0.5), ([1, 0, 0, 0], [0, 1, 0, 0], 0.3),
              ([0.5, 0.5, 0.5, 0.5], [0.5, -0.5, 0.5, -0.5], 0.8)]
    for q1, q2, t in pairs:
        cases["slerp"].append(dict(q1=q1, q2=q2, t=t, out=[float(x) for x in ref.slerp(q1, q2, t)]))
    for q in quats[:10].tolist() + [[0, 0, 0, 1], [0, 0.7071068, 0, 0.7071068], [0.5, 0.5, 0.5, 0.5], [0, 1, 0, 0]]:
        cases["euler"].append(dict(q=q, out=[float(x) for x in ref.quat_to_euler(q)]))
    for seed, sub, off in [(1234, 0, 0), (1234, 12345, 64), (1, 2073599, 0), (0xDEADBEEF, 8294399, 640), (7, 77, 777)]:
        cases["rsv"].append(dict(seed=seed, subsequence=sub, offset=off, out=ref.random_space_vectors(seed, sub, off, 4).astype(float).tolist()))
    cases["constants"] = [float(x) for x in ref.constants()]
    json.dump(cases, open(os.path.join(HERE, "ref_host_math.json"), "w"), indent=0)

    import importlib
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    synth = importlib.import_module("nvidia-optix-ray-tracer_b200.synth")
    import scenes
    orc = oracle_lib.load()

    # ---- the reference's shader programs and colour ramps, executed
    rs = oracle_lib.load_ref_shader()
    assert rs is not None, "oracle/_ref/librt_ref_shader.so missing (need /root/reference)"
    W, H, spp, seed = scenes.REF_PIN_SIZE
    out = {"depth": np.array([rs.trace_depth()], np.uint32)}
    for name, desc, cam, compat in scenes.ref_pin_cases(pkg, synth):
        r = rs.render(scenes.flatten_world(desc, compat=compat), cam, W, H, spp, seed)
        out[name + "_accum"] = r["accum"][:, :3].copy()
        out[name + "_segments"] = np.array([r["segments"]], np.uint64)
        if name == "rnd_world":
            out[name + "_color"] = r["color"]
        assert not r["albedo"].any() and not r["normal"].any()  # quirk 9, by execution
    np.savez_compressed(os.path.join(HERE, "ref_shader.npz"), **out)
    ramps = {}
    for preset in ("viridis", "plasma", "spectral", "terrain", "heatmap", "grayscale", "TeRRain", "nonsense"):
        for n in (1, 2, 7, 500):
            ramps[f"{preset}_{n}"] = rs.bake_color_ramp(preset, n)
    np.savez_compressed(os.path.join(HERE, "ref_color_ramp.npz"), **ramps)

    # ---- oracle renders of canned scenes
    out = {}
    d = scenes.config1_scene(pkg)
    cam = scenes.config_camera(pkg, "VK")
    r = orc.render(scenes.flatten_world(d), cam, 240, 135, 2, 5, 1234)
    out["c1_prim_id"] = r["prim_id"]
    out["c1_accum"] = r["accum"]
    d = scenes.random_scene(pkg, synth, 300, n_tris=40, seed=3, metal_every=5, fuzz=0.2, with_ground=True)
    cam2 = pkg.camera_configure([2.2, 0.3, 0.4], [0, 0, 0], [0, 0, 1], False)
    r = orc.render(scenes.flatten_world(d), cam2, 96, 64, 3, 6, 99)
    out["rnd_prim_id"] = r["prim_id"]
    out["rnd_accum"] = r["accum"]
    out["rnd_segments"] = np.array([r["segments"]], np.uint64)
    np.savez_compressed(os.path.join(HERE, "oracle_images.npz"), **out)
    print("golden fixtures written")


if __name__ == "__main__":
    main()
