#!/usr/bin/env python
"""Experiment (GPU): how much does the ORDER of the bounce rays in the queue matter to k_extend_wide?
Traces the same set of scattered rays of the C4 scene through rt_trace_rays in several orders and prints
Mrays/s for each.  Not part of the product or the tests; results are summarised in profiles/README.md."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def morton3(q, bits):
    import torch
    out = torch.zeros_like(q[:, 0])
    for b in range(bits):
        for a in range(3):
            out |= ((q[:, a] >> b) & 1) << (3 * b + a)
    return out


def main():
    import torch
    pkg = importlib.import_module("nvidia-optix-ray-tracer_b200")
    synth = importlib.import_module("nvidia-optix-ray-tracer_b200.synth")
    dev = torch.device("cuda", 0)
    n = int(os.environ.get("N_SPHERES", 10_000_000))
    W, H, S = 3840, 2160, int(os.environ.get("SPP", 8))
    centers, radii = synth.uniform_spheres(n, 0.0013, 0x5EED0001)
    wv, wn = synth.box_walls(True)
    d_c = torch.from_numpy(centers).to(dev)
    d_r = torch.from_numpy(radii).to(dev)
    d_wv = torch.from_numpy(wv).to(dev)
    d_wn = torch.from_numpy(wn).to(dev)
    roughs = np.float32([[0.73, 0.73, 0.73]])
    metals = np.float32([[0.8, 0.85, 0.88, 0.0]])
    ctx = pkg.Context(0)
    sc = pkg.Scene(ctx)
    sc.set_materials(roughs.ctypes.data, 1, metals.ctypes.data, 1)
    g_s = sc.add_spheres(d_c.data_ptr(), d_r.data_ptr(), n)
    g_w = sc.add_triangles(d_wv.data_ptr(), d_wn.data_ptr(), len(wv) // 3)
    sc.add_instance(g_s, None, pkg.RT_MAT_ROUGH, 0, 0)
    sc.add_instance(g_w, None, pkg.RT_MAT_ROUGH, 0, 0)
    stream = torch.cuda.current_stream().cuda_stream
    sc.build(stream)
    cam = pkg.camera_configure(**synth.c4_camera())
    npix = W * H

    # camera rays (Shader.cu:249-261, float32 in torch: close enough for an ordering experiment)
    ys, xs = torch.meshgrid(torch.arange(H, device=dev), torch.arange(W, device=dev), indexing="ij")
    nx = ((xs.float() + 0.5) / W) * 2 - 1
    ny = ((ys.float() + 0.5) / H) * 2 - 1
    U = torch.tensor(list(cam.U), device=dev)
    V = torch.tensor(list(cam.V), device=dev)
    Wv = torch.tensor(list(cam.W), device=dev)
    C0 = torch.tensor(list(cam.center), device=dev)
    d = (nx * (W / H))[..., None] * U + ny[..., None] * V + Wv
    d = d / d.norm(dim=-1, keepdim=True)
    d = d.reshape(-1, 3)

    def trace(o, dd, reps=3):
        m = o.shape[0]
        rays = torch.empty((m, 8), dtype=torch.float32, device=dev)
        rays[:, 0:3] = o
        rays[:, 3] = 1e-6
        rays[:, 4:7] = dd
        rays[:, 7] = 1e16
        t = torch.empty(m, dtype=torch.float32, device=dev)
        pid = torch.empty(m, dtype=torch.int32, device=dev)
        best = 1e9
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            sc.trace_rays(rays.data_ptr(), None, m, t.data_ptr(), pid.data_ptr(), stream)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return t, pid, best

    o0 = C0.expand(npix, 3).contiguous()
    t0, id0, ms0 = trace(o0, d)
    print(json.dumps({"set": "primary", "rays": npix, "ms": ms0, "mrays_s": npix / ms0 / 1e3}), flush=True)

    gen = torch.Generator(device=dev)
    gen.manual_seed(1234)

    def scatter(o, dd, t, pid, copies):
        hit = (pid >= 0) & (pid < n) & (t < 1e15)
        idx = hit.nonzero()[:, 0]
        P = o[idx] + t[idx, None] * dd[idx]
        N = (P - d_c[pid[idx].long()]) / d_r[pid[idx].long(), None]
        P = P.repeat_interleave(copies, 0)
        N = N.repeat_interleave(copies, 0)
        r = torch.randn((P.shape[0], 3), device=dev, generator=gen)
        r = r / r.norm(dim=-1, keepdim=True)
        nd = N + r
        nd = nd / nd.norm(dim=-1, keepdim=True).clamp_min(1e-6)
        return P + N * 1e-5, nd, idx.repeat_interleave(copies)

    def orders(o, dd, pix, label):
        m = o.shape[0]
        lo, hi = o.min(0).values, o.max(0).values
        octant = (dd[:, 0] >= 0).long() | ((dd[:, 1] >= 0).long() << 1) | ((dd[:, 2] >= 0).long() << 2)
        res = {}

        def run(name, perm):
            oo, d2 = (o, dd) if perm is None else (o[perm].contiguous(), dd[perm].contiguous())
            _, _, ms = trace(oo, d2)
            res[name] = m / ms / 1e3
            print(json.dumps({"set": label, "order": name, "rays": m, "ms": round(ms, 3), "mrays_s": round(m / ms / 1e3, 1)}), flush=True)

        run("as generated (pixel-major, copies adjacent)", None)
        run("random permutation", torch.randperm(m, device=dev, generator=gen))
        # the renderer's queue order: sample-major inside 512-pixel strips, grouped by octant
        tile = pix // 512 if pix is not None else torch.arange(m, device=dev) // 4096
        run("octant within blocks of 512 sources (renderer-like)", torch.sort(tile * 8 + octant, stable=True).indices)
        for bits in (5, 6, 7, 8):
            q = ((o - lo) / (hi - lo).clamp_min(1e-9) * ((1 << bits) - 1)).long()
            mc = morton3(q, bits)
            run(f"octant | morton{bits}", torch.sort(octant * (1 << (3 * bits)) + mc, stable=True).indices)
            run(f"morton{bits} | octant", torch.sort(mc * 8 + octant, stable=True).indices)
        q = ((o - lo) / (hi - lo).clamp_min(1e-9) * 63).long()
        mc = morton3(q, 6)
        # 6x6x... direction cells on the octahedral map
        ad = dd.abs().sum(-1, keepdim=True)
        pq = dd[:, :2] / ad
        neg = dd[:, 2] < 0
        px = torch.where(neg, (1 - pq[:, 1].abs()) * torch.sign(pq[:, 0]), pq[:, 0])
        py = torch.where(neg, (1 - pq[:, 0].abs()) * torch.sign(pq[:, 1]), pq[:, 1])
        cell = ((px * 0.5 + 0.5) * 7.999).long() * 8 + ((py * 0.5 + 0.5) * 7.999).long()
        run("dircell64 | morton6", torch.sort(cell * (1 << 18) + mc, stable=True).indices)
        run("morton6 | dircell64", torch.sort(mc * 64 + cell, stable=True).indices)
        return res

    o1, d1, pix1 = scatter(o0, d, t0, id0, S)
    r1 = orders(o1, d1, pix1, "bounce1")
    t1, id1, _ = trace(o1, d1, reps=1)
    o2, d2, _ = scatter(o1, d1, t1, id1, 1)
    r2 = orders(o2, d2, None, "bounce2")
    t2, id2, _ = trace(o2, d2, reps=1)
    o3, d3, _ = scatter(o2, d2, t2, id2, 1)
    r3 = orders(o3, d3, None, "bounce3")
    json.dump({"bounce1": r1, "bounce2": r2, "bounce3": r3}, open(os.path.join(ROOT, "gpurun_out", "exp_coherence.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
