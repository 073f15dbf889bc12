"""The fused form of rt_render (scene option "fused" = 1: one persistent kernel follows every path from its first
scatter to its end, no ray queues) must give the wavefront form's frame, and the oracle's, bit for bit."""
import os

import numpy as np
import pytest

import scenes

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
IMG = np.load(os.path.join(HERE, "golden", "oracle_images.npz"))
GOLD = np.load(os.path.join(HERE, "golden", "ref_shader.npz"))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("n,ntri,spp,depth,res", [
    (0, 0, 2, 5, (64, 48)),
    (1, 0, 3, 5, (64, 48)),
    (50, 10, 4, 5, (160, 90)),
    (2000, 300, 4, 8, (200, 120)),
    (20000, 0, 19, 5, (256, 144)),   # 19 samples = one 16-sample unit + a ragged one per pixel
    (3000, 0, 3, 1, (96, 64)),       # depth 1: no path is ever started
])
@pytest.mark.parametrize("opts", [{}, {"chunk_spp": 1}, {"chunk_spp": 5, "batch_spp": 7, "fetch_min": 1}])
def test_fused_image_parity_vs_oracle(pkg, synth, oracle, n, ntri, spp, depth, res, opts):
    d = scenes.random_scene(pkg, synth, n, n_tris=ntri, seed=n + 1, metal_every=7 if n > 10 else 0, fuzz=0.25,
                            with_walls=n != 0, with_ground=n >= 50)
    cam = pkg.camera_configure([2.2, 0.2, 0.3], [0, 0, 0], [0, 0, 1], False)
    gs = scenes.GpuScene(pkg, d, options=dict(opts, fused=1))
    g = gs.render(cam, res[0], res[1], spp, depth, 7 + n)
    gs.close()
    o = oracle.render(scenes.flatten_world(d), cam, res[0], res[1], spp, depth, 7 + n, accel=1 if n > 500 else 0)
    assert np.array_equal(g["prim_id"], o["prim_id"])
    assert np.array_equal(bits(g["accum"]), bits(o["accum"]))
    assert np.array_equal(bits(g["albedo"]), bits(o["albedo"])) and np.array_equal(bits(g["normal"]), bits(o["normal"]))
    assert g["segments"] == o["segments"]


def test_fused_golden_fixture_and_config1(pkg, synth):
    d = scenes.random_scene(pkg, synth, 300, n_tris=40, seed=3, metal_every=5, fuzz=0.2, with_ground=True)
    cam = pkg.camera_configure([2.2, 0.3, 0.4], [0, 0, 0], [0, 0, 1], False)
    gs = scenes.GpuScene(pkg, d, options={"fused": 1})
    g = gs.render(cam, 96, 64, 3, 6, 99)
    gs.close()
    assert np.array_equal(bits(g["accum"]), bits(IMG["rnd_accum"])) and g["segments"] == int(IMG["rnd_segments"][0])
    gs = scenes.GpuScene(pkg, scenes.config1_scene(pkg), options={"fused": 1})   # outlier primitive (r = 1000 ground)
    g = gs.render(scenes.config_camera(pkg, "VK"), 240, 135, 2, 5, 1234)
    gs.close()
    assert np.array_equal(g["prim_id"], IMG["c1_prim_id"]) and np.array_equal(bits(g["accum"]), bits(IMG["c1_accum"]))


def test_fused_sample_partitions_streams_and_form_switches(pkg, synth):
    """Per-pixel XORWOW streams are created lazily in the fused form (a pixel's state is stepped only while the
    pixel is hit): contiguous blocks, blocks with gaps (the multi-GPU spp split), going back to an earlier offset
    and switching between the two forms on one scene all give the wavefront form's samples."""
    d = scenes.random_scene(pkg, synth, 3000, seed=4, metal_every=5)
    cam = pkg.camera_configure([2.2, 0, 0.1], [0, 0, 0], [0, 0, 1], False)
    ref = scenes.GpuScene(pkg, d)
    gs = scenes.GpuScene(pkg, d, options={"fused": 1})
    W, H = 160, 96
    full = ref.render(cam, W, H, 8, 5, 11, want=("accum",))["accum"]
    assert np.array_equal(bits(gs.render(cam, W, H, 8, 5, 11, want=("accum",))["accum"]), bits(full))
    a = gs.render(cam, W, H, 3, 5, 11, sample_offset=0, want=("accum",))            # back to offset 0
    b = gs.render(cam, W, H, 5, 5, 11, sample_offset=3, flags=pkg.RT_FLAG_ACCUMULATE, accum_init=a["accum"], want=("accum",))
    assert np.array_equal(bits(b["accum"]), bits(full))
    for off, cnt in ((40, 2), (100, 3), (7000, 2), (5, 4)):                          # gaps, a far jump, backwards
        want = ref.render(cam, W, H, cnt, 5, 11, sample_offset=off, want=("accum",))["accum"]
        got = gs.render(cam, W, H, cnt, 5, 11, sample_offset=off, want=("accum",))["accum"]
        assert np.array_equal(bits(got), bits(want)), (off, cnt)
    gs.scene.set_option("fused", 0)                                                  # same scene, other form, and back
    assert np.array_equal(bits(gs.render(cam, W, H, 8, 5, 11, want=("accum",))["accum"]), bits(full))
    gs.scene.set_option("fused", 1)
    assert np.array_equal(bits(gs.render(cam, W, H, 8, 5, 12, want=("accum",))["accum"]),
                          bits(ref.render(cam, W, H, 8, 5, 12, want=("accum",))["accum"]))  # new seed
    gs.close()
    ref.close()


@pytest.mark.parametrize("idx", [0, 2, 4])
def test_fused_reproduces_reference_shader_golden(pkg, synth, idx):
    name, desc, cam, compat = scenes.ref_pin_cases(pkg, synth)[idx]
    W, H, spp, seed = scenes.REF_PIN_SIZE
    gs = scenes.GpuScene(pkg, desc, options={"compat_shading": 3 if compat else 0, "fused": 1})
    g = gs.render(cam, W, H, spp, 5, seed, flags=pkg.RT_FLAG_REF_NO_SELF_SKIP | pkg.RT_FLAG_REF_MUL_ON_RETURN, want=("accum",))
    gs.close()
    assert np.array_equal(bits(g["accum"][:, :3]), bits(GOLD[name + "_accum"])), name


def test_fused_uses_a_fraction_of_the_wavefront_memory(pkg, synth):
    d = scenes.random_scene(pkg, synth, 20000, seed=2)
    cam = pkg.camera_configure([2.2, 0.2, 0.3], [0, 0, 0], [0, 0, 1], False)
    a = scenes.GpuScene(pkg, d, options={"compact_queues": 0})   # queues sized for every pixel, rt_render never blocks
    b = scenes.GpuScene(pkg, d, options={"fused": 1})
    c = scenes.GpuScene(pkg, d)                                   # default: queues sized from the frame's primary-hit count
    ra = a.render(cam, 640, 360, 16, 5, 3, want=("accum",))
    rb = b.render(cam, 640, 360, 16, 5, 3, want=("accum",))
    rc = c.render(cam, 640, 360, 16, 5, 3, want=("accum",))
    assert np.array_equal(bits(ra["accum"]), bits(rb["accum"])) and np.array_equal(bits(ra["accum"]), bits(rc["accum"]))
    assert ra["segments"] == rb["segments"] == rc["segments"]
    assert rb["stats"].device_bytes < 0.4 * ra["stats"].device_bytes
    assert rc["stats"].device_bytes < ra["stats"].device_bytes
    assert rb["stats"].kernel_launches < ra["stats"].kernel_launches
    # a frame with MORE hit pixels than the one the compact queues were sized for: they grow, same result as the full ones
    cam2 = pkg.camera_configure([1.2, 0.1, 0.2], [0, 0, 0], [0, 0, 1], False)
    ra2 = a.render(cam2, 640, 360, 16, 5, 3, want=("accum",))
    rc2 = c.render(cam2, 640, 360, 16, 5, 3, want=("accum",))
    assert np.array_equal(bits(ra2["accum"]), bits(rc2["accum"])) and ra2["segments"] == rc2["segments"]
    a.close()
    b.close()
    c.close()
