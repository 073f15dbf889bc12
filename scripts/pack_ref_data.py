#!/usr/bin/env python
"""Pack the reference's shipped DATA files for BASELINE configs 2 and 3 into data/ref_files/.

data/ is git-ignored (data files of the reference are not product source) but NOT gpurun-ignored, so the
directory travels to the GPU box where /root/reference does not exist.  Run in the build container:

    python scripts/pack_ref_data.py [--ref /root/reference/files] [--out data/ref_files]

What it writes:
  - copies of files/config.json, particle.vtk.series, particle-short.vtk.series, wall.vtk.series, particle/*.vtk,
    wall/*.vtk, shape/separated/*.stl (C2: the full 401-file Time-mode sequence);
  - particle_mesh/particle_*.vtk + particle_mesh-short.vtk.series (C3): the reference repository git-ignores its
    files/particle_mesh/ directory, so the Mesh-mode series is SYNTHESISED as SURVEY 8(d) prescribes: every
    Time-mode particle of particle-short.vtk.series (71 files) is baked (STL shape x the file's sub-frame-0
    transform, rt_transform_time) into one world-space TRIANGLE_STRIPS cell with CELL_DATA id / vel — the Mesh-mode
    input format of docs/usage.md:15-20 and src/Util/VTKReaderImpl.cpp:24-137.  A particle's closed surface is one
    strip: greedy strips joined by index-degenerate bridge triangles (kept by the reference reader as zero-area
    triangles, never hit);
  - config_c2.json / config_c3.json: files/config.json with the paths re-pointed; the ground sphere is the metal
    (metals[0], fuzz 0) — SURVEY 8(d) C2 asks to declare where the metal goes.
Nothing here is read by the product at run time except through the public config / loader API.
"""
import argparse
import importlib
import json
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def weld(verts):
    """bit-identical vertices -> point ids (the STL readers merge points)"""
    v = np.ascontiguousarray(verts, np.float32) + np.float32(0.0)
    uniq, inv = np.unique(v.view(np.uint32).reshape(-1, 3), axis=0, return_inverse=True)
    return uniq.view(np.float32).reshape(-1, 3), inv.reshape(-1)


def _emitted(strip, k):
    a, b, c = strip[k], strip[k + 1], strip[k + 2]
    return (a, c, b) if k & 1 else (a, b, c)


def _canon(t):
    i = min(range(3), key=lambda j: t[j])
    return (t[i], t[(i + 1) % 3], t[(i + 2) % 3])


def stripify(tris):
    """One triangle strip over every triangle of `tris` (point-id triples, consistent winding).  Triangle k of a
    strip is (s[k], s[k+1], s[k+2]), last two swapped for odd k (VTKReaderImpl.cpp:97-110); winding is preserved.
    Greedy strips are joined with repeated indices (index-degenerate triangles)."""
    tris = [tuple(int(x) for x in t) for t in tris]
    want = {}
    for i, t in enumerate(tris):
        want.setdefault(_canon(t), []).append(i)
    edge = {}
    for i, t in enumerate(tris):
        for j in range(3):
            edge.setdefault(frozenset((t[j], t[(j + 1) % 3])), []).append(i)
    used = [False] * len(tris)

    def grow(start, rot, mark):
        t = tris[start]
        s = [t[rot], t[(rot + 1) % 3], t[(rot + 2) % 3]]
        taken = [start]
        seen = {start}
        while True:
            nxt = None
            for cand in edge.get(frozenset((s[-2], s[-1])), ()):
                if used[cand] or cand in seen:
                    continue
                third = [p for p in tris[cand] if p != s[-2] and p != s[-1]]
                if len(third) != 1:
                    continue
                if _canon(_emitted(s + third, len(s) - 2)) == _canon(tris[cand]):
                    nxt = (cand, third[0])
                    break
            if nxt is None:
                break
            s.append(nxt[1])
            taken.append(nxt[0])
            seen.add(nxt[0])
        if mark:
            for i in taken:
                used[i] = True
        return s, taken

    out = []
    for start in range(len(tris)):
        if used[start]:
            continue
        best = max(range(3), key=lambda r: len(grow(start, r, False)[1]))
        s, _ = grow(start, best, True)
        if out:
            out += [out[-1], s[0]]
            if len(out) & 1:
                out.append(s[0])
        out += s
    # verify: the non-degenerate triangles of the strip are exactly the input triangles, winding included
    got = {}
    for k in range(len(out) - 2):
        t = _emitted(out, k)
        if len(set(t)) == 3:
            got[_canon(t)] = got.get(_canon(t), 0) + 1
    assert got == {k: len(v) for k, v in want.items()}, "stripify lost or flipped a triangle"
    return out


def write_mesh_vtk_binary(path, points, strips, ids, vels):
    with open(path, "wb") as f:
        f.write(b"# vtk DataFile Version 3.0\nsynthesised Mesh-mode frame (scripts/pack_ref_data.py)\nBINARY\nDATASET POLYDATA\n")
        f.write(b"POINTS %d float\n" % len(points) + np.asarray(points, ">f4").tobytes() + b"\n")
        flat = np.concatenate([np.concatenate([[len(s)], s]) for s in strips]).astype(">i4")
        f.write(b"TRIANGLE_STRIPS %d %d\n" % (len(strips), len(flat)) + flat.tobytes() + b"\n")
        f.write(b"CELL_DATA %d\n" % len(strips))
        f.write(b"SCALARS id int 1\nLOOKUP_TABLE default\n" + np.asarray(ids, ">i4").tobytes() + b"\n")
        f.write(b"VECTORS vel float\n" + np.asarray(vels, ">f4").tobytes() + b"\n")


def synth_mesh_series(out, pkg, hio, cfg):
    series = hio.load_series(out + "/", "particle-short.vtk.series")
    shapes = hio.load_shapes_dir(out + "/shape/separated/")
    welded = []
    for s in shapes:
        pts, idx = weld(s["vertices"])
        welded.append((pts, stripify(idx.reshape(-1, 3))))
    os.makedirs(out + "/particle_mesh", exist_ok=True)
    files = []
    lens = []
    for i, path in enumerate(series["files"]):
        p = hio.load_particles(path)
        nxt = hio.load_particles(series["files"][min(i + 1, len(series["files"]) - 1)])
        dur = series["durations"][i]
        frames = hio.frames_for_file(dur, cfg["loop-data"]["fps"], cfg["loop-data"]["render-speed-ratio"])
        points, strips = [], []
        base = 0
        for k in range(len(p["id"])):
            qn = nxt["quat"][k] if k < len(nxt["id"]) else p["quat"][k]
            m = np.float32(pkg.transform_time(p["pos"][k], p["vel"][k], p["quat"][k], qn, dur, frames, 0, [0, 0, 0], [1, 1, 1])).reshape(3, 4)
            pts, strip = welded[int(p["shape_id"][k])]
            points.append(pts @ m[:, :3].T + m[:, 3])
            strips.append(np.asarray(strip) + base)
            base += len(pts)
            lens.append(len(strip) - 2)
        name = "particle_mesh/" + os.path.basename(path)
        write_mesh_vtk_binary(out + "/" + name, np.concatenate(points), strips, p["id"].astype(np.int64), p["vel"])
        files.append({"name": name, "time": float(series["times"][i])})
    json.dump({"file-series-version": "1.0", "files": files}, open(out + "/particle_mesh-short.vtk.series", "w"), indent=1)
    return len(files), int(np.mean(lens))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference/files")
    ap.add_argument("--out", default=os.path.join(ROOT, "data", "ref_files"))
    a = ap.parse_args()
    if not os.path.isdir(a.ref):
        sys.exit(f"{a.ref} not found: run this in the build container")
    out = os.path.abspath(a.out)
    os.makedirs(out, exist_ok=True)
    for name in ("config.json", "particle.vtk.series", "particle-short.vtk.series", "wall.vtk.series"):
        shutil.copy(os.path.join(a.ref, name), out)
    for d in ("particle", "wall", "shape/separated"):
        shutil.copytree(os.path.join(a.ref, d), os.path.join(out, d), dirs_exist_ok=True)
    cfg = json.load(open(os.path.join(out, "config.json")))
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    hio = importlib.import_module("nvidia-optix-ray-tracer_b200.hostio")
    n, avg = synth_mesh_series(out, pkg, hio, cfg)
    for tag, mesh, series in (("c2", False, "particle.vtk.series"), ("c3", True, "particle_mesh-short.vtk.series")):
        c = json.loads(json.dumps(cfg))
        # relative to the working directory, like the reference's "../files/": run from the repository root
        c.update({"mesh": mesh, "series-path": "data/ref_files/", "series-name": series, "cache-path": "data/ref_files/cache/",
                  "stl-path": "data/ref_files/shape/separated/", "cache": False})
        c["spheres"][0].update({"mat-type": "METAL", "mat-index": 0})
        json.dump(c, open(os.path.join(out, f"config_{tag}.json"), "w"), indent=1)
    size = sum(os.path.getsize(os.path.join(r, f)) for r, _, fs in os.walk(out) for f in fs)
    print(f"packed {out}: {size / 1e6:.1f} MB; synthesised {n} Mesh-mode files, ~{avg} triangles (incl. bridges) per particle")


if __name__ == "__main__":
    main()
