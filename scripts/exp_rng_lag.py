#!/usr/bin/env python
"""Break-even of the host-side stream re-positioning threshold (scene option "rng_max_lag").

A frame whose first sample lies G samples ahead of where the per-pixel XORWOW states stand either steps every hit
pixel's state G*32 draws inside k_prepare (G <= rng_max_lag) or re-creates all states with k_rng_init on the side stream
(G > rng_max_lag).  This renders a 4K frame sequence with a fixed gap between frames under both policies and prints the
device time per frame (CUDA events around rt_render, median of the timed frames).

    python scripts/exp_rng_lag.py [--n 1000000] [--spp 16] > gpurun_out/<tag>_rng_lag.txt
"""
import argparse
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--gaps", default="0,2,4,8,16,32,56,112,256,1024")
    a = ap.parse_args()
    import torch
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    synth = importlib.import_module("nvidia-optix-ray-tracer_b200.synth")
    dev = torch.device("cuda", 0)
    c, r = synth.uniform_spheres(a.n, radius=1.3e-3 * (10_000_000 / a.n) ** (1 / 3))
    wv, wn = synth.box_walls()
    d_c, d_r = torch.from_numpy(c).to(dev), torch.from_numpy(r).to(dev)
    d_wv, d_wn = torch.from_numpy(wv).to(dev), torch.from_numpy(wn).to(dev)
    roughs = np.full((a.n + 1, 3), 0.6, np.float32)
    ctx = pkg.Context(0)
    cam = pkg.camera_configure(**synth.c4_camera())
    npix = a.width * a.height
    acc = torch.zeros((npix, 4), device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    print(f"# {a.n} spheres, {a.width}x{a.height}, {a.spp} spp per frame, depth 5; ms per frame (rt_render only), median of {a.frames}")
    print("# gap = samples skipped between consecutive frames' sample ranges")
    print("gap  step_ms  reinit_ms")
    for gap in [int(g) for g in a.gaps.split(",")]:
        row = []
        for lag in (2048, 0):
            sc = pkg.Scene(ctx)
            sc.set_option("rng_max_lag", lag)
            sc.set_materials(roughs.ctypes.data, len(roughs), None, 0)
            g_s = sc.add_spheres(d_c.data_ptr(), d_r.data_ptr(), a.n)
            g_w = sc.add_triangles(d_wv.data_ptr(), d_wn.data_ptr(), len(wv) // 3)
            sc.add_instance(g_s, None, pkg.RT_MAT_ROUGH, 1, 1)
            sc.add_instance(g_w, None, pkg.RT_MAT_ROUGH, 0, 0)
            sc.build(stream)
            ms = []
            for f in range(a.frames + 2):
                off = f * (a.spp + gap)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                sc.render(cam, a.width, a.height, a.spp, 5, 99, (0.7, 0.8, 0.9), off, 0, acc.data_ptr(), None, None, None, stream)
                e1.record()
                torch.cuda.synchronize()
                if f >= 2:
                    ms.append(e0.elapsed_time(e1))
            row.append(float(np.median(ms)))
            sc.close()
        print(f"{gap:4d} {row[0]:8.3f} {row[1]:9.3f}", flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
