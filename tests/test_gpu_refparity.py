"""CUDA path vs the REFERENCE'S OWN shader programs (golden vectors produced by shader/Shader.cu compiled as host
code: tests/golden/ref_shader.npz, see tests/test_ref_shader.py) and vs the oracle in compat mode.

With RT_FLAG_REF_NO_SELF_SKIP | RT_FLAG_REF_MUL_ON_RETURN (and scene option compat_shading = 3 on the scenes that
carry the reference's object-space shading data) rt_render reproduces the linear radiance of the reference's
closest-hit / miss recursion bit for bit."""
import os

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "ref_shader.npz"))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("idx", range(6))
@pytest.mark.parametrize("wide", [1, 0])
def test_cuda_reproduces_reference_shader_golden(pkg, synth, idx, wide):
    name, desc, cam, compat = scenes.ref_pin_cases(pkg, synth)[idx]
    W, H, spp, seed = scenes.REF_PIN_SIZE
    gs = scenes.GpuScene(pkg, desc, options={"compat_shading": 3 if compat else 0, "wide": wide})
    g = gs.render(cam, W, H, spp, 5, seed, flags=pkg.RT_FLAG_REF_NO_SELF_SKIP | pkg.RT_FLAG_REF_MUL_ON_RETURN, want=("accum",))
    gs.close()
    assert np.array_equal(bits(g["accum"][:, :3]), bits(GOLD[name + "_accum"])), name


@pytest.mark.parametrize("flags", [0, 2, 4, 6])
def test_cuda_compat_and_flags_match_oracle(pkg, synth, oracle, flags):
    """Every combination of the two transport flags, on the instanced scene with object-space shading data, equals
    the oracle with the same switches (radiance, primary ids, first-hit AOVs)."""
    d = scenes.instanced_scene(pkg, synth)
    cam = pkg.camera_configure([2.0, 0.5, 0.6], [0, 0, 0], [0, 0, 1], False)
    gs = scenes.GpuScene(pkg, d, options={"compat_shading": 3})
    g = gs.render(cam, 120, 80, 5, 5, 77, flags=flags)
    gs.close()
    o = oracle.render(scenes.flatten_world(d, compat=True), cam, 120, 80, 5, 5, 77, flags=flags >> 1)
    assert np.array_equal(g["prim_id"], o["prim_id"])
    assert np.array_equal(bits(g["accum"]), bits(o["accum"]))
    assert np.array_equal(bits(g["albedo"]), bits(o["albedo"]))
    assert np.array_equal(bits(g["normal"]), bits(o["normal"]))


def test_compat_bits_are_independent(pkg, synth, oracle):
    """compat_shading = 1 (object-space sphere centre only) and = 2 (unrotated triangle normals only)."""
    d = scenes.instanced_scene(pkg, synth)
    cam = pkg.camera_configure([2.0, 0.5, 0.6], [0, 0, 0], [0, 0, 1], False)
    full, none = scenes.flatten_world(d, compat=True), scenes.flatten_world(d)
    for mode in (1, 2):
        world = dict(none)
        if mode == 1:
            world["sphere_shade"] = full["sphere_shade"]
        else:
            world["tri_normals"] = full["tri_normals"]
        gs = scenes.GpuScene(pkg, d, options={"compat_shading": mode})
        g = gs.render(cam, 96, 64, 3, 5, 5)
        gs.close()
        o = oracle.render(world, cam, 96, 64, 3, 5, 5)
        assert np.array_equal(bits(g["accum"]), bits(o["accum"])), mode
        assert np.array_equal(bits(g["normal"]), bits(o["normal"])), mode


def test_ref_mul_flag_rejects_deep_paths(pkg, synth):
    d = scenes.instanced_scene(pkg, synth)
    cam = pkg.camera_configure([2.0, 0.5, 0.6], [0, 0, 0], [0, 0, 1], False)
    gs = scenes.GpuScene(pkg, d)
    with pytest.raises(pkg.RtError):
        gs.render(cam, 16, 16, 1, 6, 1, flags=pkg.RT_FLAG_REF_MUL_ON_RETURN, want=("accum",))
    gs.close()


def test_material_index_out_of_range_is_rejected(pkg, synth):
    """scatter() reads roughs[] / metals[] unchecked, so rt_accel_build validates every instance's
    material range (index + (count-1)*stride) against the table of its type."""
    import torch
    ctx = pkg.Context(0)
    sc = pkg.Scene(ctx)
    c = torch.zeros((10, 3), device="cuda")
    r = torch.full((10,), 0.1, device="cuda")
    roughs = np.zeros((5, 3), np.float32)
    metals = np.zeros((1, 4), np.float32)
    sc.set_materials(roughs.ctypes.data, 5, metals.ctypes.data, 1)
    g = sc.add_spheres(c.data_ptr(), r.data_ptr(), 10)
    sc.add_instance(g, None, pkg.RT_MAT_ROUGH, 0, 1)      # indices 0..9 > 5 entries
    with pytest.raises(pkg.RtError, match="material index"):
        sc.build()
    sc.clear_geometry()
    g = sc.add_spheres(c.data_ptr(), r.data_ptr(), 10)
    sc.add_instance(g, None, pkg.RT_MAT_METAL, 1, 0)      # metal 1 of 1
    with pytest.raises(pkg.RtError, match="material index"):
        sc.build()
    sc.clear_geometry()
    g = sc.add_spheres(c.data_ptr(), r.data_ptr(), 5)
    sc.add_instance(g, None, pkg.RT_MAT_ROUGH, 0, 1)      # 0..4: fits
    sc.build()
    sc.close()
    ctx.close()


@pytest.mark.gpu
def test_materials_after_build_and_shrinking_tables(pkg, synth):
    """the range check lives in rt_accel_build, so (a) a build before any rt_scene_set_materials fails against the
    empty tables (documented order), (b) equal-size tables set after a build need no rebuild, (c) a table that shrinks
    after a build invalidates it: render is refused until the build has re-checked, and the re-check fails."""
    import torch
    ctx = pkg.Context(0)
    sc = pkg.Scene(ctx)
    c = torch.zeros((8, 3), device="cuda")
    c[:, 0] = torch.arange(8, device="cuda") * 0.3
    r = torch.full((8,), 0.1, device="cuda")
    g = sc.add_spheres(c.data_ptr(), r.data_ptr(), 8)
    sc.add_instance(g, None, pkg.RT_MAT_ROUGH, 0, 1)      # indices 0..7
    with pytest.raises(pkg.RtError, match="table of 0 entries"):
        sc.build()
    roughs = np.full((8, 3), 0.5, np.float32)
    sc.set_materials(roughs.ctypes.data, 8, None, 0)
    sc.build()
    cam = pkg.camera_configure([1.0, -3.0, 0.0], [1.0, 0, 0], [0, 0, 1], False)
    acc = torch.zeros((32 * 16, 4), device="cuda")

    def render():
        sc.render(cam, 32, 16, 1, 5, 7, (0.7, 0.8, 0.9), 0, 0, acc.data_ptr(), None, None, None, None)
        torch.cuda.synchronize()

    render()
    roughs2 = np.full((8, 3), 0.25, np.float32)
    sc.set_materials(roughs2.ctypes.data, 8, None, 0)     # same size: still built
    render()
    sc.set_materials(roughs2.ctypes.data, 5, None, 0)     # smaller: indices 5..7 would read stale rows
    with pytest.raises(pkg.RtError, match="before rt_accel_build"):
        render()
    with pytest.raises(pkg.RtError, match="material index"):
        sc.build()
    sc.set_materials(roughs2.ctypes.data, 8, None, 0)
    sc.build()
    render()
    sc.close()
    ctx.close()
