"""Host-side maths of the path (camera basis, per-frame transforms, colour ramps): the product's
C ABI (rt_*) and the oracle (orc_*) against values produced by the REFERENCE's own code
(tests/golden/ref_host_math.json, generated from oracle/_ref by tests/golden/make_golden.py) and,
when oracle/_ref is present, against the live reference functions.  Bit-exact (fp32)."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
G = json.load(open(os.path.join(HERE, "golden", "ref_host_math.json")))


def f32list(x):
    return [float(np.float32(v)) for v in x]


def bits(x):
    return np.asarray(x, np.float32).view(np.uint32).tolist()


def test_constants():
    assert bits(G["constants"]) == bits([1e-6, 1e16, 3.1415926])


@pytest.mark.parametrize("case", G["transform"])
def test_construct_transform_matches_reference(pkg, oracle, built_lib, case):
    want = bits(case["out"])
    assert bits(pkg.construct_transform(case["shift"], case["rotate"], case["scale"])) == want
    assert bits(oracle.construct_transform(case["shift"], case["rotate"], case["scale"])) == want


def test_known_transform_value(pkg, built_lib):
    # constructTransformMatrix({0,0,-1000.5}, 0, 1) -> identity + tz
    assert pkg.construct_transform([0, 0, -1000.5], [0, 0, 0], [1, 1, 1]) == [1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, -1000.5]


@pytest.mark.parametrize("case", G["camera"])
def test_camera_matches_reference(pkg, oracle, built_lib, case):
    want = bits(case["out"])
    cam = pkg.camera_configure(case["center"], case["target"], case["up"], not case["opengl"])
    assert bits(list(cam.center) + list(cam.U) + list(cam.V) + list(cam.W)) == want
    oc = oracle.camera_configure(case["center"], case["target"], case["up"], not case["opengl"])
    assert bits(list(oc.center) + list(oc.U) + list(oc.V) + list(oc.W)) == want


@pytest.mark.parametrize("case", G["slerp"])
def test_slerp_matches_reference(oracle, case):
    assert bits(oracle.slerp(case["q1"], case["q2"], case["t"])) == bits(case["out"])


@pytest.mark.parametrize("case", G["euler"])
def test_quat_to_euler_matches_reference(oracle, case):
    assert bits(oracle.quat_to_euler(case["q"])) == bits(case["out"])


@pytest.mark.parametrize("case", G["color"])
def test_color_matches_reference(oracle, case):
    assert bits(oracle.color_to_float4(case["rgb"])) == bits(case["f4"])
    assert oracle.color_to_uchar4(case["rgb"]) == case["u4"]


def test_known_color_value(oracle):
    # colorToFloat4(0.7,0.8,0.9)
    got = oracle.color_to_float4([0.7, 0.8, 0.9])
    assert ["%.9f" % v for v in got[:3]] == ["0.854305804", "0.906331718", "0.954687059"]


def test_transform_time_chain(pkg, oracle, built_lib):
    """rt_transform_time == oracle chain == slerp -> quatToEuler -> constructTransformMatrix built from
    the pinned pieces (RendererTime.cu:436-472)."""
    fr = json.load(open(os.path.join(HERE, "golden", "particle_frame0.json")))
    q = np.asarray(fr["quat"], np.float32)
    p = np.asarray(fr["points"], np.float32)
    v = np.asarray(fr["vel"], np.float32)
    frames = int(0.01 * 240 * 4)  # (size_t)(duration * fps * ratio), RendererTime.cu:427-428
    for i in range(len(q)):
        for frame in (0, 1, 4, frames - 1):
            qn = q[(i + 1) % len(q)]
            a = pkg.transform_time(p[i], v[i], q[i], qn, 0.01, frames, frame, [0.1, 0, -0.2], [1, 1, 1])
            b = oracle.transform_time(p[i], v[i], q[i], qn, 0.01, frames, frame, [0.1, 0, -0.2], [1, 1, 1])
            assert bits(a) == bits(b)
            # recompose from pinned pieces
            factor = np.float32(frame) / np.float32(frames - 1)
            e = oracle.quat_to_euler(oracle.slerp(q[i], qn, factor))
            shift = np.float32([0.1, 0, -0.2]) + (p[i] + ((v[i] * np.float32(0.01)) / np.float32(frames)) * np.float32(frame))
            c = oracle.construct_transform(shift, e, [1, 1, 1])
            assert bits(a) == bits(c)


def test_transform_mesh(pkg, oracle, built_lib):
    for frame in (0, 3, 8):
        a = pkg.transform_mesh([0.5, -1, -2], 0.01, 9, frame, [0.1, 0.2, 0.3], [1, 1, 1])
        b = oracle.transform_mesh([0.5, -1, -2], 0.01, 9, frame, [0.1, 0.2, 0.3], [1, 1, 1])
        assert bits(a) == bits(b)
        assert a[0] == 1 and a[5] == 1 and a[10] == 1


def test_live_reference_if_present(pkg, oracle, ref, built_lib):
    if ref is None:
        pytest.skip("oracle/_ref not built (reference checkout absent)")
    rng = np.random.default_rng(5)
    for _ in range(50):
        s, r, sc = rng.uniform(-3, 3, 3), rng.uniform(-360, 360, 3), rng.uniform(0.1, 4, 3)
        want = bits(ref.construct_transform(s, r, sc))
        assert bits(pkg.construct_transform(s, r, sc)) == want
        assert bits(oracle.construct_transform(s, r, sc)) == want
        q1, q2 = rng.normal(size=4), rng.normal(size=4)
        q1, q2 = q1 / np.linalg.norm(q1), q2 / np.linalg.norm(q2)
        t = rng.uniform()
        assert bits(oracle.slerp(q1, q2, t)) == bits(ref.slerp(q1, q2, t))
        assert bits(oracle.quat_to_euler(q1)) == bits(ref.quat_to_euler(q1))
        c = rng.uniform(-0.2, 1.2, 3)
        assert bits(oracle.color_to_float4(c)) == bits(ref.color_to_float4(c))
        ce, ta, up = rng.uniform(-5, 5, 3), rng.uniform(-1, 1, 3), rng.normal(size=3)
        cam = pkg.camera_configure(ce, ta, up, True)
        assert bits(list(cam.center) + list(cam.U) + list(cam.V) + list(cam.W)) == bits(ref.camera_configure(ce, ta, up, 0))


def test_color_ramp(pkg, built_lib):
    # include/Util/ColorRamp.cuh:83-112: endpoints are the first/last stop, count==1 -> last stop
    r = pkg.bake_color_ramp("terrain", 5)
    assert np.allclose(r[0], [0.149, 0.149, 0.149]) and np.allclose(r[-1], [0.996, 0.922, 0.545])
    assert np.allclose(r[2], [0.639, 0.784, 0.325])
    assert np.allclose(pkg.bake_color_ramp("TeRrAiN", 5), r)
    assert np.allclose(pkg.bake_color_ramp("nonsense", 3), pkg.bake_color_ramp("viridis", 3))
    assert np.allclose(pkg.bake_color_ramp("grayscale", 1), [[0.95, 0.95, 0.95]])
    g = pkg.bake_color_ramp("grayscale", 11)
    assert np.allclose(g[:, 0], 0.05 + 0.9 * np.arange(11) / 10, atol=1e-6)
    # numpy restatement of the piecewise-linear bake for a long ramp
    n = 1000
    stops = np.array([[0.0, 0.267, 0.004, 0.329], [0.25, 0.283, 0.141, 0.458], [0.5, 0.254, 0.265, 0.530],
                      [0.75, 0.196, 0.509, 0.364], [1.0, 0.993, 0.906, 0.144]], np.float32)
    u = np.arange(n, dtype=np.float32) / np.float32(n - 1)
    want = np.stack([np.interp(u, stops[:, 0], stops[:, k]) for k in (1, 2, 3)], 1)
    assert np.allclose(pkg.bake_color_ramp("viridis", n), want, atol=2e-6)
