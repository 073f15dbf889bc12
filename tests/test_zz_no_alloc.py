"""rt_trace_rays_static: closest hits for callers that are not allowed to allocate device memory.

Some users of this library reach a point where cudaMalloc is forbidden or would fail: a frame loop with a memory plan
fixed at start-up, a plugin whose host application owns the allocator, a process that shares the GPU with a job that
has taken the rest of the memory.  For them the library carries one fixed scratch array per device, and
rt_trace_rays_static runs the traversal kernel on it instead of on buffers that rt_trace_rays grows on demand.

This file is collected last on purpose (pytest -x stops at the first failure; a failure here must not hide the other
tests) and passes both alone in a fresh process — where its first call is the one that looks up the array's address —
and after the rest of the suite."""
import numpy as np
import pytest

import scenes

MISS = 0xFFFFFFFF


def _rays(rng, n):
    o = rng.uniform(-0.55, 0.55, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    return np.concatenate([o, np.full((n, 1), 1e-6, np.float32), d, np.full((n, 1), 1e16, np.float32)], 1)


def _trace(gs, rays, skip, static):
    """Both entry points through the C ABI on the same device buffers; outputs pre-filled with a pattern neither writes."""
    torch = gs.torch
    n = len(rays)
    r = torch.from_numpy(np.ascontiguousarray(rays, np.float32)).to(gs.dev)
    t = torch.full((max(n, 1),), -7.0, dtype=torch.float32, device=gs.dev)
    ids = torch.full((max(n, 1),), 123456789, dtype=torch.int32, device=gs.dev)
    sk = torch.from_numpy(np.ascontiguousarray(skip, np.uint32).view(np.int32)).to(gs.dev) if skip is not None else None
    call = gs.scene.trace_rays_static if static else gs.scene.trace_rays
    call(r.data_ptr(), sk.data_ptr() if sk is not None else None, n, t.data_ptr(), ids.data_ptr(),
         stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return t.cpu().numpy()[:n], ids.cpu().numpy().view(np.uint32)[:n]


@pytest.mark.gpu
@pytest.mark.parametrize("opts", [{"wide": 1}, {"wide": 0}])
def test_static_scratch_trace_equals_allocating_trace(pkg, synth, opts):
    """Two calls in a row on the shared array, the second smaller and with self-exclusion, so that it runs over what the
    first one left behind.  The bar is the one tests/test_gpu_parity.py sets for traversal: ids and t bit for bit."""
    d = scenes.random_scene(pkg, synth, 30000, n_tris=3000, seed=33, with_ground=True)
    gs = scenes.GpuScene(pkg, d, options=opts)
    rng = np.random.default_rng(17)
    cap = pkg.RT_TRACE_STATIC_MAX_RAYS

    # call 1: the largest batch the array takes
    rays1 = _rays(rng, cap)
    t_ref, i_ref = _trace(gs, rays1, None, static=False)
    t_new, i_new = _trace(gs, rays1, None, static=True)
    assert (i_ref != MISS).sum() > cap // 10, "scene / rays do not exercise the hit path"
    assert np.array_equal(i_ref, i_new)
    assert np.array_equal(t_ref.view(np.uint32), t_new.view(np.uint32))

    # call 2, straight after: fewer rays, started on the primitives call 1 hit and excluding them
    hit = np.flatnonzero(i_ref != MISS)[:20001]
    o2 = (rays1[hit, 0:3] + t_ref[hit, None] * rays1[hit, 4:7]).astype(np.float32)
    rays2 = _rays(rng, len(hit))
    rays2[:, 0:3] = o2
    t_ref2, i_ref2 = _trace(gs, rays2, i_ref[hit], static=False)
    t_new2, i_new2 = _trace(gs, rays2, i_ref[hit], static=True)
    assert np.array_equal(i_ref2, i_new2)
    assert np.array_equal(t_ref2.view(np.uint32), t_new2.view(np.uint32))
    assert (i_new2 != MISS).any() and not np.array_equal(i_new2, i_new[: len(hit)])  # really call 2's answer

    # call 3: call 1 again gives call 1's answer (nothing of call 2 leaks)
    t_again, i_again = _trace(gs, rays1, None, static=True)
    assert np.array_equal(i_again, i_ref) and np.array_equal(t_again.view(np.uint32), t_ref.view(np.uint32))
    gs.close()


@pytest.mark.gpu
def test_static_scratch_trace_rejects_oversize_batches(pkg, synth):
    """One ray more than the array holds is an error that names the limit; nothing is launched and the outputs stay as
    they were.  n = 0 is accepted (the documented way to do the per-device set-up ahead of time)."""
    d = scenes.random_scene(pkg, synth, 2000, seed=5)
    gs = scenes.GpuScene(pkg, d)
    rng = np.random.default_rng(3)
    _trace(gs, _rays(rng, 0).reshape(0, 8), None, static=True)
    with pytest.raises(pkg.RtError) as e:
        _trace(gs, _rays(rng, pkg.RT_TRACE_STATIC_MAX_RAYS + 1), None, static=True)
    assert "RT_TRACE_STATIC_MAX_RAYS" in str(e.value) and str(pkg.RT_TRACE_STATIC_MAX_RAYS) in str(e.value)
    # the scene is still usable afterwards
    rays = _rays(rng, 1000)
    t0, i0 = _trace(gs, rays, None, static=False)
    t1, i1 = _trace(gs, rays, None, static=True)
    assert np.array_equal(i0, i1) and np.array_equal(t0.view(np.uint32), t1.view(np.uint32))
    gs.close()
