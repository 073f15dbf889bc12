"""rt_spheres_advect (device-side Mesh-mode particle update, src/Global/RendererMesh.cu:378-391) against
the host path it replaces: rt_transform_mesh per particle + instance flattening in rt_accel_build."""
import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu
W, H = 200, 120


def _particles(n, seed):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-0.4, 0.4, (n, 3)).astype(np.float32)
    r = rng.uniform(0.01, 0.03, n).astype(np.float32)
    v = rng.uniform(-3.0, 3.0, (n, 3)).astype(np.float32)
    return c, r, v


@pytest.mark.parametrize("pscale", [1.0, 0.75])
def test_advect_matches_host_transform_bit_for_bit(pkg, pscale):
    import torch
    n = 5003  # not a multiple of 4: the vector body and the scalar tail both run
    c, r, v = _particles(n, 3)
    duration, frames, frame = np.float32(0.01), 9, 4
    pshift, ps = [0.02, -0.01, 0.03], [pscale] * 3
    ctx = pkg.Context(0)
    dc, dr, dv = (torch.from_numpy(x).cuda() for x in (c, r, v))
    oc, orr = torch.empty_like(dc), torch.empty_like(dr)
    ctx.spheres_advect(dc.data_ptr(), dr.data_ptr(), dv.data_ptr(), n, float(duration), frames, frame, pshift, ps,
                       oc.data_ptr(), orr.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got_c, got_r = oc.cpu().numpy(), orr.cpu().numpy()
    want_c = np.empty_like(c)
    want_r = np.empty_like(r)
    for i in range(n):
        m = np.float32(pkg.transform_mesh(v[i], float(duration), frames, frame, pshift, ps))
        want_c[i] = scenes._xform_points(m, c[i:i + 1])[0]
        want_r[i] = r[i] * np.sqrt((m[0] * m[0] + m[4] * m[4]) + m[8] * m[8], dtype=np.float32)
    assert np.array_equal(got_c.view(np.uint32), want_c.view(np.uint32))
    assert np.array_equal(got_r.view(np.uint32), want_r.view(np.uint32))
    ctx.close()


def test_advected_array_renders_like_per_particle_instances(pkg):
    """One sphere group advected on the device == one instance per particle with the host transform."""
    import torch
    n = 60
    c, r, v = _particles(n, 5)
    r = (r * np.float32(3.0)).astype(np.float32)
    duration, frames, frame = 0.01, 9, 7
    pshift, ps = [0.0, 0.01, -0.02], [1.0, 1.0, 1.0]
    roughs = np.random.default_rng(1).uniform(0.2, 0.9, (n, 3)).astype(np.float32)
    cam = pkg.camera_configure([2.0, 0.2, 0.1], [0, 0, 0], [0, 0, 1], True)
    # host path: particle i = its own group + instance transform (the reference's Mesh-mode layout)
    d1 = scenes.SceneDesc(roughs, np.zeros((0, 4), np.float32))
    for i in range(n):
        g = d1.add_spheres(c[i:i + 1], r[i:i + 1])
        d1.add_instance(g, pkg.transform_mesh(v[i], duration, frames, frame, pshift, ps), 0, i, 0)
    a = scenes.GpuScene(pkg, d1).render(cam, W, H, 3, 5, 11)
    # device path
    ctx = pkg.Context(0)
    dc, dr, dv = (torch.from_numpy(x).cuda() for x in (c, r, v))
    oc, orr = torch.empty_like(dc), torch.empty_like(dr)
    ctx.spheres_advect(dc.data_ptr(), dr.data_ptr(), dv.data_ptr(), n, duration, frames, frame, pshift, ps, oc.data_ptr(),
                       orr.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    d2 = scenes.SceneDesc(roughs, np.zeros((0, 4), np.float32))
    g = d2.add_spheres(oc.cpu().numpy(), orr.cpu().numpy())
    d2.add_instance(g, None, 0, 0, 1)
    b = scenes.GpuScene(pkg, d2).render(cam, W, H, 3, 5, 11)
    assert (a["prim_id"] != 0xFFFFFFFF).sum() > 500
    assert np.array_equal(a["prim_id"], b["prim_id"])
    assert np.array_equal(a["accum"].view(np.uint32), b["accum"].view(np.uint32))
    ctx.close()


def _time_particles(n, seed):
    rng = np.random.default_rng(seed)
    pos = rng.uniform(-0.4, 0.4, (n, 3)).astype(np.float32)
    vel = rng.uniform(-3.0, 3.0, (n, 3)).astype(np.float32)
    qa = rng.normal(size=(n, 4)).astype(np.float32)
    qa /= np.linalg.norm(qa, axis=1, keepdims=True)
    qb = rng.normal(size=(n, 4)).astype(np.float32)
    qb /= np.linalg.norm(qb, axis=1, keepdims=True)
    k = n // 3
    qb[:k] = (qa[:k] + rng.normal(scale=0.01, size=(k, 4))).astype(np.float32)   # near-equal key frames: the lerp branch (dot > 0.9995)
    qb[k:k + 50] = -qa[k:k + 50]                                                  # negative dot: the sign flip
    qa[k + 50:k + 60] = np.float32([0, 0.70710678, 0, 0.70710678])                # pitch at the asin clamp
    qb[k + 50:k + 60] = qa[k + 50:k + 60]
    return pos, vel, qa, qb


# Tolerance of the device Time-mode chain (include/rtb200.h rt_transforms_time): the translation column is plain
# fp32 arithmetic in the reference's order -> bit-exact; the 3x3 part goes through CUDA's sinf/cosf/acosf/asinf/atan2f
# (<= 2 ulp each) instead of glibc's -> absolute difference <= 4e-6 on entries of magnitude <= |scale|.
TIME_ROT_ATOL = 4e-6


@pytest.mark.parametrize("frame,frames", [(0, 9), (4, 9), (8, 9), (0, 1)])
def test_transforms_time_device_matches_host_chain(pkg, frame, frames):
    import torch
    n = 3001
    pos, vel, qa, qb = _time_particles(n, 7 + frame)
    duration, pshift, ps = 0.01, [0.02, -0.01, 0.03], [0.5, 0.5, 0.5]
    ctx = pkg.Context(0)
    dp, dv, da, db = (torch.from_numpy(x).cuda() for x in (pos, vel, qa, qb))
    out = torch.empty((n, 12), dtype=torch.float32, device="cuda")
    ctx.transforms_time(dp.data_ptr(), dv.data_ptr(), da.data_ptr(), db.data_ptr(), n, duration, frames, frame, pshift, ps,
                        out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    want = np.stack([np.float32(pkg.transform_time(pos[i], vel[i], qa[i], qb[i], duration, frames, frame, pshift, ps)) for i in range(n)])
    assert np.array_equal(got[:, [3, 7, 11]].view(np.uint32), want[:, [3, 7, 11]].view(np.uint32)), "translation column must be bit-exact"
    rot = [0, 1, 2, 4, 5, 6, 8, 9, 10]
    assert np.isfinite(got).all()
    err = np.abs(got[:, rot].astype(np.float64) - want[:, rot].astype(np.float64)).max()
    assert err <= TIME_ROT_ATOL, err
    ctx.close()


def test_device_resident_transforms_render_like_host_transforms(pkg):
    """rt_scene_set_transforms_device (no host round trip) == rt_scene_set_transforms with the same values, and a
    later host-side call for the same instances takes over again."""
    import torch
    n = 40
    pos, vel, qa, qb = _time_particles(n, 21)
    rng = np.random.default_rng(2)
    roughs = rng.uniform(0.2, 0.9, (n, 3)).astype(np.float32)
    cam = pkg.camera_configure([2.0, 0.2, 0.1], [0, 0, 0], [0, 0, 1], True)
    tri_v = np.float32([[-0.06, -0.05, -0.03], [0.07, -0.04, 0.0], [0.0, 0.06, 0.04]] * 1).reshape(3, 3)
    tri_n = np.float32([[0, 0, 1]] * 3)
    d = scenes.SceneDesc(roughs, np.zeros((0, 4), np.float32))
    g = d.add_triangles(tri_v, tri_n)
    tf = [pkg.transform_time(pos[i], vel[i], qa[i], qb[i], 0.01, 9, 3, [0, 0, 0], [1.5, 1.5, 1.5]) for i in range(n)]
    for i in range(n):
        d.add_instance(g, tf[i], 0, i, 0)
    a = scenes.GpuScene(pkg, d).render(cam, W, H, 3, 5, 11)
    assert (a["prim_id"] != 0xFFFFFFFF).sum() > 100
    # same scene built with identity transforms, then the transforms arrive as a device array
    d2 = scenes.SceneDesc(roughs, np.zeros((0, 4), np.float32))
    g2 = d2.add_triangles(tri_v, tri_n)
    for i in range(n):
        d2.add_instance(g2, None, 0, i, 0)
    gs = scenes.GpuScene(pkg, d2)
    dev_tf = torch.from_numpy(np.float32(tf).reshape(n, 12)).cuda()
    gs.scene.set_transforms_device(dev_tf.data_ptr(), 0, n)
    gs.scene.build(torch.cuda.current_stream().cuda_stream)
    b = gs.render(cam, W, H, 3, 5, 11)
    assert np.array_equal(a["prim_id"], b["prim_id"])
    assert np.array_equal(a["accum"].view(np.uint32), b["accum"].view(np.uint32))
    # the array is read again by the next build: new values, no new call
    tf2 = [pkg.transform_time(pos[i], vel[i], qa[i], qb[i], 0.01, 9, 6, [0, 0, 0], [1.5, 1.5, 1.5]) for i in range(n)]
    dev_tf.copy_(torch.from_numpy(np.float32(tf2).reshape(n, 12)).cuda())
    gs.scene.build(torch.cuda.current_stream().cuda_stream)
    c = gs.render(cam, W, H, 3, 5, 11)
    d3 = scenes.SceneDesc(roughs, np.zeros((0, 4), np.float32))
    g3 = d3.add_triangles(tri_v, tri_n)
    for i in range(n):
        d3.add_instance(g3, tf2[i], 0, i, 0)
    e = scenes.GpuScene(pkg, d3).render(cam, W, H, 3, 5, 11)
    assert np.array_equal(c["accum"].view(np.uint32), e["accum"].view(np.uint32))
    assert not np.array_equal(c["prim_id"], a["prim_id"])
    # host-side values for the same range replace the device array
    host_tf = np.float32(tf).reshape(n, 12)
    gs.scene.set_transforms(host_tf.ctypes.data, 0, n)
    gs.scene.build(torch.cuda.current_stream().cuda_stream)
    f = gs.render(cam, W, H, 3, 5, 11)
    assert np.array_equal(f["accum"].view(np.uint32), a["accum"].view(np.uint32))
    with pytest.raises(pkg.RtError):
        gs.scene.set_transforms_device(dev_tf.data_ptr(), n - 1, 2)
    gs.close()
