"""The oracle's shading / transport restatement pinned against the REFERENCE'S OWN device programs.

oracle/_ref/librt_ref_shader.so is shader/Shader.cu (+ Shader.cuh, DeviceFunctions.cuh, ColorRamp.cuh), unmodified,
compiled as host code with a host emulation of the OptiX device API (oracle/ref_shader.cpp, oracle/ref_stub/);
only the closest-hit query behind optixTrace comes from the oracle (OptiX's traversal is closed source).
tests/golden/ref_shader.npz and ref_color_ramp.npz hold outputs of that library (tests/golden/make_golden.py), so the
pin holds on machines without /root/reference; where the library exists the same checks also run live.

Bar: BIT-EXACT.  With ORC_FLAG_NO_SELF_SKIP | ORC_FLAG_MUL_ON_RETURN the oracle reproduces the linear radiance the
reference's raygen receives, bit for bit, on every scene (rough / metal / fuzz, spheres / triangles, translated,
rotated and scaled instances with the reference's object-space shading data).  The default oracle deviates in
exactly those two documented ways; test_default_deviation_* bound their size.
"""
import os

import numpy as np
import pytest

import oracle_lib
import scenes

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "ref_shader.npz"))
RAMPS = np.load(os.path.join(HERE, "golden", "ref_color_ramp.npz"))
REF_FLAGS = oracle_lib.FLAG_NO_SELF_SKIP | oracle_lib.FLAG_MUL_ON_RETURN


@pytest.fixture(scope="module")
def ref_shader():
    return oracle_lib.load_ref_shader()


@pytest.fixture(scope="module")
def cases(pkg, synth):
    return scenes.ref_pin_cases(pkg, synth)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_golden_depth_is_the_reference_constant():
    assert int(GOLD["depth"][0]) == 5  # include/Global/Shader.cuh:8


@pytest.mark.parametrize("idx", range(6))
def test_oracle_reproduces_reference_shader_golden(oracle, cases, idx):
    name, desc, cam, compat = cases[idx]
    W, H, spp, seed = scenes.REF_PIN_SIZE
    o = oracle.render(scenes.flatten_world(desc, compat=compat), cam, W, H, spp, 5, seed, flags=REF_FLAGS)
    assert np.array_equal(bits(o["accum"][:, :3]), bits(GOLD[name + "_accum"])), name
    # the reference traces the primary ray of every launch; the oracle counts it once per pixel
    assert int(GOLD[name + "_segments"][0]) == o["segments"] + (spp - 1) * W * H


def test_reference_colour_buffer_is_srgb_of_the_linear_result(oracle, cases):
    """raygen stores colorToFloat4(result) (Shader.cu:270): encode the oracle's last sample and compare bit for bit."""
    name, desc, cam, compat = cases[2]
    W, H, spp, seed = scenes.REF_PIN_SIZE
    o = oracle.render(scenes.flatten_world(desc, compat=compat), cam, W, H, 1, 5, seed, sample_offset=spp - 1, flags=REF_FLAGS)
    enc = np.array([oracle.color_to_float4(list(px[:3])) for px in o["accum"]], np.float32)
    assert np.array_equal(bits(enc), bits(GOLD[name + "_color"]))


def test_each_deviation_alone_changes_pixels(oracle, cases):
    """Both flags are needed: dropping either one leaves pixels that differ from the reference (so the pin really
    exercises self-intersection and the product order)."""
    name, desc, cam, compat = cases[2]
    W, H, spp, seed = scenes.REF_PIN_SIZE
    world = scenes.flatten_world(desc, compat=compat)
    for flags in (0, oracle_lib.FLAG_NO_SELF_SKIP, oracle_lib.FLAG_MUL_ON_RETURN):
        o = oracle.render(world, cam, W, H, spp, 5, seed, flags=flags)
        assert not np.array_equal(bits(o["accum"][:, :3]), bits(GOLD[name + "_accum"])), flags


def test_default_deviation_multiply_order_is_rounding_only(oracle, cases):
    """Throughput multiplied on the way down instead of on the way back (Shader.cu:236-238): <= 4 products of
    factors <= 1, i.e. a few ulp of a value <= 0.9 per sample.  Stated bound: 4 spp * 3e-7."""
    W, H, spp, seed = scenes.REF_PIN_SIZE
    for name, desc, cam, compat in cases:
        o = oracle.render(scenes.flatten_world(desc, compat=compat), cam, W, H, spp, 5, seed, flags=oracle_lib.FLAG_NO_SELF_SKIP)
        assert np.abs(o["accum"][:, :3] - GOLD[name + "_accum"]).max() <= spp * 3e-7, name


def test_default_deviation_self_exclusion_psnr(oracle, cases):
    """Self-exclusion (a scattered ray never re-hits the primitive it leaves) against the reference's tmin = 1e-6.
    The size of this deviation is set by how often a scattered ray re-hits its own primitive within (1e-6, inf) in
    fp32, i.e. by primitive size: measured at 4 spp 45 dB on the random scene (r = 0.05 spheres: 9 of 2304 pixels
    differ), 30-35 dB on the instanced scene (r = 0.3 sphere, scaled walls: 57-76 pixels), 16-24 dB on the shipped
    config with its r = 1000 ground sphere: 1e-6 is below the fp32 resolution of a hit point at |P| ~ 1000 (6e-5), so
    with OUR sphere intersector the reference's rule re-hits the ground at once (shadow acne, paths die at the depth
    cut-off) — what OptiX's closed-source intersector does there is unknown.  Stated bounds: 40 / 28 / 12 dB."""
    W, H, spp, seed = scenes.REF_PIN_SIZE
    for name, desc, cam, compat in cases:
        o = oracle.render(scenes.flatten_world(desc, compat=compat), cam, W, H, spp, 5, seed)
        mse = float(np.mean(((o["accum"][:, :3] - GOLD[name + "_accum"]) / spp).astype(np.float64) ** 2))
        psnr = 99.0 if mse == 0 else 10 * np.log10(1.0 / mse)
        assert psnr >= (12.0 if name.startswith("c1") else 28.0 if name.startswith("inst") else 40.0), (name, psnr)


# ---------------------------------------------------------------------------------- colour ramps (ColorRamp.cuh)
@pytest.mark.parametrize("key", sorted(RAMPS.files))
def test_color_ramp_bit_exact_vs_reference(pkg, key):
    preset, n = key.rsplit("_", 1)
    ours = pkg.bake_color_ramp(preset, int(n))
    assert np.array_equal(bits(ours), bits(RAMPS[key])), key


# ---------------------------------------------------------------------------------- live (library present)
def test_live_reference_matches_golden_and_oracle(oracle, ref_shader, cases, pkg, synth):
    if ref_shader is None:
        pytest.skip("oracle/_ref/librt_ref_shader.so absent (no /root/reference on this machine)")
    W, H, spp, seed = scenes.REF_PIN_SIZE
    name, desc, cam, compat = cases[4]
    r = ref_shader.render(scenes.flatten_world(desc, compat=compat), cam, W, H, spp, seed)
    assert np.array_equal(bits(r["accum"][:, :3]), bits(GOLD[name + "_accum"]))
    assert not r["albedo"].any() and not r["normal"].any()  # the AOV buffers end up zero
    # other seeds / sizes / sample offsets, BVH-accelerated closest hit: live reference == oracle, bit for bit
    d = scenes.random_scene(pkg, synth, 1500, n_tris=60, seed=9, metal_every=3, fuzz=0.35, with_ground=True)
    world = scenes.flatten_world(d, compat=True)
    cam2 = pkg.camera_configure([1.9, -0.6, 0.7], [0, 0, 0], [0, 0, 1], True)
    r = ref_shader.render(world, cam2, 80, 45, 3, 31337, sample_offset=5, accel=1)
    o = oracle.render(world, cam2, 80, 45, 3, 5, 31337, sample_offset=5, accel=1, flags=REF_FLAGS)
    assert np.array_equal(bits(r["accum"]), bits(o["accum"]))


def test_live_rsqrt_deviation_is_noise_level(oracle, ref_shader, cases):
    """normalize(): IEEE 1/sqrt (ours, needed for CPU/GPU bit parity) vs the device's <= 2 ulp rsqrtf, emulated by a
    correctly rounded double-precision reciprocal square root.  Paths diverge chaotically after a 1-ulp change, so
    this is a per-sample noise term, not a bias: stated bound >= 30 dB at 4 spp (46-59 dB at 64 spp)."""
    if ref_shader is None:
        pytest.skip("oracle/_ref/librt_ref_shader.so absent")
    W, H, spp, seed = scenes.REF_PIN_SIZE
    name, desc, cam, compat = cases[2]
    r = ref_shader.render(scenes.flatten_world(desc, compat=compat), cam, W, H, spp, seed, rsqrt_mode=1)
    mse = float(np.mean(((r["accum"][:, :3] - GOLD[name + "_accum"]) / spp).astype(np.float64) ** 2))
    assert 10 * np.log10(1.0 / max(mse, 1e-30)) >= 30.0
    assert abs(float(r["accum"][:, :3].mean()) - float(GOLD[name + "_accum"].mean())) < 0.02 * spp


def test_live_color_ramp_and_preset_names(pkg, ref_shader):
    if ref_shader is None:
        pytest.skip("oracle/_ref/librt_ref_shader.so absent")
    for preset in ("Viridis", "PLASMA", "spectral", "terrain", "heatmap", "grayscale", "unknown-preset"):
        for n in (1, 3, 25, 1000):
            assert np.array_equal(bits(pkg.bake_color_ramp(preset, n)), bits(ref_shader.bake_color_ramp(preset, n))), (preset, n)
    assert ref_shader.preset_name("unknown-preset") == "viridis"  # ColorRamp.cuh:146-150
