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
