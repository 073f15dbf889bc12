"""Parity at BASELINE config 4's FULL size (10 M spheres + box walls, 3840x2160) through size-independent
properties (the oracle cannot render this scene in test time, but it can brute-force a few hundred
rays against all 10 M primitives):
  * closest hit of sampled primary and scattered rays == the oracle's brute force, ids and t bit-exact;
  * the 8-wide tree and the binary LBVH give the same ids and the same radiance bits for the whole frame;
  * the samples-per-pixel partition does not change a single bit of the frame."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu
N, RADIUS, W, H = 10_000_000, 1.3e-3, 3840, 2160
MISS = 0xFFFFFFFF


@pytest.fixture(scope="module")
def c4(pkg, synth):
    centers, radii = synth.uniform_spheres(N, RADIUS, 0x5EED0001)
    ramp = pkg.bake_color_ramp("terrain", N)
    roughs = np.concatenate([np.float32([[0.73, 0.73, 0.73]]), ramp]).astype(np.float32)
    d = scenes.SceneDesc(roughs, np.float32([[0.8, 0.85, 0.88, 0.0]]))
    g = d.add_spheres(centers, radii)
    d.add_instance(g, None, 0, 1, 1)
    v, nn = synth.box_walls(True)
    g = d.add_triangles(v, nn)
    d.add_instance(g, None, 0, 0, 0)
    cam = pkg.camera_configure(**synth.c4_camera())
    return d, cam


def _primary_rays(cam, px):
    """Shader.cu:249-261 in fp32 for pixel indices px (the oracle and the kernel normalise themselves:
    rays handed to rt_trace_rays keep the un-normalised direction, t is in its units)."""
    x, y = (px % W).astype(np.float32), (px // W).astype(np.float32)
    nx = ((x + np.float32(0.5)) / np.float32(W)) * np.float32(2) - np.float32(1)
    ny = ((y + np.float32(0.5)) / np.float32(H)) * np.float32(2) - np.float32(1)
    asp = np.float32(W) / np.float32(H)
    U, V, Wv, C = (np.float32(list(v)) for v in (cam.U, cam.V, cam.W, cam.center))
    d = (nx * asp)[:, None] * U + ny[:, None] * V + Wv
    o = np.broadcast_to(C, d.shape)
    n = len(px)
    return np.concatenate([o, np.full((n, 1), 1e-6, np.float32), d.astype(np.float32), np.full((n, 1), 1e16, np.float32)], 1).astype(np.float32)


def test_full_size_hits_equal_oracle_brute_force(pkg, oracle, c4):
    d, cam = c4
    world = scenes.flatten_world(d)
    gs = scenes.GpuScene(pkg, d)
    rng = np.random.default_rng(5)
    # pixels of the central region, where the unit cube projects to (it covers ~24 % of the 16:9 frame)
    px = rng.integers(int(0.2 * H), int(0.8 * H), 160) * W + rng.integers(int(0.34 * W), int(0.66 * W), 160)
    rays = _primary_rays(cam, px)
    t0, i0 = oracle.trace(world, rays, accel=0)   # brute force over all 10 000 010 primitives
    t1, i1 = gs.trace(rays)
    assert (i0 != MISS).sum() > 100
    assert np.array_equal(i0, i1)
    assert np.array_equal(t0.view(np.uint32), t1.view(np.uint32))
    hit = i0 != MISS
    o2 = (rays[hit, 0:3] + t0[hit, None] * rays[hit, 4:7]).astype(np.float32)
    d2 = rng.normal(size=o2.shape).astype(np.float32)
    rays2 = np.concatenate([o2, np.full((len(o2), 1), 1e-6, np.float32), d2, np.full((len(o2), 1), 1e16, np.float32)], 1)
    t2, i2 = oracle.trace(world, rays2, skip=i0[hit], accel=0)
    t3, i3 = gs.trace(rays2, skip=i0[hit])
    assert np.array_equal(i2, i3) and np.array_equal(t2.view(np.uint32), t3.view(np.uint32))
    gs.close()


def test_full_size_frame_is_independent_of_the_tree_and_of_the_sample_partition(pkg, c4):
    d, cam = c4
    gs = scenes.GpuScene(pkg, d)
    a = gs.render(cam, W, H, 3, 5, 7, want=("accum", "prim_id"))
    assert int(a["stats"].n_prims) == N + 10
    covered = (a["prim_id"] != MISS).mean()
    assert 0.15 < covered <= 1.0
    # samples [0,3) in one call == [0,1) then [1,3) accumulated
    b1 = gs.render(cam, W, H, 1, 5, 7, want=("accum",))
    b2 = gs.render(cam, W, H, 2, 5, 7, sample_offset=1, flags=pkg.RT_FLAG_ACCUMULATE, want=("accum",), accum_init=b1["accum"])
    assert np.array_equal(a["accum"].view(np.uint32), b2["accum"].view(np.uint32))
    assert a["segments"] == b1["segments"] + b2["segments"] - W * H  # primary rays are traced once per call
    gs.close()
    # binary LBVH + the other traversal kernel: same ids, same radiance bits
    gb = scenes.GpuScene(pkg, d, options={"wide": 0})
    c = gb.render(cam, W, H, 3, 5, 7, want=("accum", "prim_id"))
    assert np.array_equal(a["prim_id"], c["prim_id"])
    assert np.array_equal(a["accum"].view(np.uint32), c["accum"].view(np.uint32))
    assert a["segments"] == c["segments"]
    gb.close()
