"""What the parity contract rests on but no standard pins, checked against the oracle binary and the device:

* the engine of oracle/rng_shim.h (raw draws), libstdc++'s float and piecewise draws from it — numpy restatement
  vs `ref_harness pinning` (CPU) and vs `b2pt_debug_draws` (GPU);
* the evaluation order of `vec2(x + g.sample(), y + g.sample())` (Technique.cpp:322) and
  `vec2(g.sample(), g.sample())` (AreaLights.cpp:224) in THIS build of the reference: the FIRST draw lands in y
  (g++ evaluates the constructor arguments right to left) — and the device mirrors it (`b2pt_debug_primary`,
  `b2pt_debug_light_samples`);
* the known answers the reference keeps in its never-compiled unittest() blocks (Cameras.cpp:164-189,
  Technique.cpp:118-152) and its stale gtest file (unit_tests/Cameras.test.cpp:38-47), evaluated by the reference's
  own ray_direction / pixel_position (`ref_harness camera`) and by the device (`b2pt_debug_camera`).
"""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import REF_HARNESS, golden_path

needs_harness = pytest.mark.skipif(not os.path.exists(REF_HARNESS), reason="oracle/_ref/ref_harness not built")
f32 = np.float32


def harness(*args):
    out = subprocess.run([REF_HARNESS] + [str(a) for a in args], check=True, capture_output=True, text=True).stdout
    return json.loads(out.strip().splitlines()[-1])


# ---------------------------------------------------------------- numpy restatement of the stream definition (rng_shim.h)
def mix32(x):
    x = np.uint32(x)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint32(16); x = np.uint32(x * np.uint32(0x7feb352d))
        x ^= x >> np.uint32(15); x = np.uint32(x * np.uint32(0x846ca68b))
        x ^= x >> np.uint32(16)
    return np.uint32(x)


def stream_key(seed, domain, index, sample):
    with np.errstate(over="ignore"):
        k = mix32(np.uint32(seed) + np.uint32(0x9E3779B9) * np.uint32(domain + 1))
        k = mix32(k ^ np.uint32(index))
        k = mix32(k + np.uint32(0x85EBCA6B) * np.uint32(sample))
    return k


def draw(key, ctr):
    with np.errstate(over="ignore"):
        return mix32(np.uint32(key) ^ mix32(np.uint32(ctr) * np.uint32(0x9E3779B9) + np.uint32(0x7F4A7C15)))


def draw_float(key, ctr):
    """std::uniform_real_distribution<float> over a 32-bit URBG: generate_canonical<float, 24> = float(u) * 2^-32, 1.0 -> nextafter."""
    f = f32(f32(draw(key, ctr)) * f32(2.0 ** -32))
    return f32(0.99999994) if f >= f32(1.0) else f


def draw_double(key, ctr):
    """generate_canonical<double, 53>: two engine calls."""
    r = (float(draw(key, ctr)) + float(draw(key, ctr + 1)) * 4294967296.0) * 2.0 ** -64
    return np.nextafter(1.0, 0.0) if r >= 1.0 else r


def piecewise(key, ctr, weights):
    """std::piecewise_constant_distribution<float>(n, 0, 1, fw) as libstdc++ initialises and samples it (bits/random.tcc)."""
    n = len(weights)
    delta = f32(1.0) / f32(n)
    ints = [f32(f32(k) * delta) for k in range(n + 1)]
    # fw is evaluated at the interval mid points: weights[min(size_t(x * n), n - 1)] (Sample.inl:237-243)
    den = [float(weights[min(int(f32(f32(ints[k] + f32(0.5) * delta)) * n), n - 1)]) for k in range(n)]
    s = sum(den)
    den = [d / s for d in den]
    cp, acc = [], 0.0
    for d in den:
        acc += d
        cp.append(acc)
    cp[-1] = 1.0
    den = [den[k] / float(ints[k + 1] - ints[k]) for k in range(n)]
    p = draw_double(key, ctr)
    i = int(np.searchsorted(np.array(cp), p, side="left"))
    pref = cp[i - 1] if i > 0 else 0.0
    return f32(float(ints[i]) + (p - pref) / den[i])


def ray_direction(px, py, width, height, flen):
    """Cameras.cpp:120-127 in float32."""
    inv = f32(1.0) / f32(height)
    x = f32(f32(f32(px) * inv) * f32(2.0)) - f32(f32(width) * inv)
    y = f32(f32(f32(py) * inv) * f32(2.0)) - f32(1.0)
    v = np.array([x, y, -f32(flen)], dtype=f32)
    d = f32(f32(v[0] * v[0]) + f32(v[1] * v[1])) + f32(v[2] * v[2])  # glm::dot: tmp = a * b; tmp.x + tmp.y + tmp.z
    return v * f32(f32(1.0) / np.sqrt(d, dtype=f32))  # glm::normalize = v * inversesqrt(dot)


def to_world(v2w, d):
    """glm mat3 * vec3: m[0] * v.x + m[1] * v.y + m[2] * v.z, component-wise, left to right (float32)."""
    m = v2w.reshape(3, 3).astype(f32)
    return np.array([f32(f32(f32(m[0][k] * d[0]) + f32(m[1][k] * d[1])) + f32(m[2][k] * d[2])) for k in range(3)], dtype=f32)


SCENE = "TestCaseFurnace"  # six area lights: the light pick is a real piecewise draw


@pytest.fixture(scope="module")
def pin():
    return harness("pinning", golden_path("scenes", SCENE + ".b2s"), "count=24", "width=64", "height=48")


@needs_harness
def test_engine_and_libstdcxx_draws_match_the_restatement(pin):
    key = np.uint32(pin["key"])
    assert key == stream_key(pin["seed"], 0, 7, 3)
    assert [int(draw(key, i)) for i in range(8)] == pin["raw"]
    assert [float(draw_float(key, i)) for i in range(8)] == [float(f32(x)) for x in pin["floats"]]
    assert pin["piecewise_draws_used"] == 16  # one double = two engine calls per pick
    ours = [float(piecewise(key, 2 * i, pin["piecewise_weights"])) for i in range(8)]
    assert ours == [float(f32(x)) for x in pin["piecewise"]]


@needs_harness
def test_first_jitter_draw_lands_in_y(pin, flat_scenes):
    """Technique.cpp:322 — `vec2(x + g.sample(), y + g.sample())`: this build draws y's jitter first."""
    flat = flat_scenes(SCENE)
    W, H = pin["width"], pin["height"]
    ctx = flat.camera_context(0, W, H)
    assert float(ctx["focal_length_y"]) == float(f32(pin["focal_length_y"]))
    y_first = x_first = 0
    for p in pin["primary"]:
        key = stream_key(pin["seed"], 0, p["pixel"], p["sample"])
        d0, d1 = draw_float(key, 0), draw_float(key, 1)
        px, py = p["pixel"] % W, p["pixel"] // W
        ref = np.array(p["direction"], dtype=f32)
        a = to_world(ctx["view_to_world"], ray_direction(f32(px) + d1, f32(py) + d0, W, H, ctx["focal_length_y"]))
        b = to_world(ctx["view_to_world"], ray_direction(f32(px) + d0, f32(py) + d1, W, H, ctx["focal_length_y"]))
        y_first += int(np.array_equal(a, ref))
        x_first += int(np.array_equal(b, ref))
        assert np.allclose(np.array(p["origin"], dtype=f32), ctx["position"])
    assert y_first == len(pin["primary"]) and x_first == 0


@needs_harness
def test_first_light_position_draw_lands_in_y(pin, flat_scenes):
    """AreaLights.cpp:224 — `vec2(g.sample(), g.sample())`: the draw after the light pick is the `up` coordinate."""
    flat = flat_scenes(SCENE)
    L = flat.lights
    weights = L["weight"]
    hits = 0
    for s in pin["lights"]:
        key = np.uint32(s["key"])
        assert s["draws"] == 4  # one double for the pick + two floats
        pick = piecewise(key, 0, weights)
        lid = min(int(f32(pick) * f32(len(L))), len(L) - 1)
        assert int(L["material_id"][lid]) == s["material_id"]
        s1, s0 = draw_float(key, 2), draw_float(key, 3)  # first draw -> y
        tan = L["tangent"][lid].reshape(3, 3)
        ux = f32(f32(s0 - f32(0.5)) * L["size"][lid][0])
        uy = f32(f32(s1 - f32(0.5)) * L["size"][lid][1])
        pos = (L["position"][lid] + ux * tan[0]).astype(f32)
        pos = (pos + uy * tan[2]).astype(f32)
        hits += int(np.array_equal(pos, np.array(s["position"], dtype=f32)))
    assert hits == len(pin["lights"])


@needs_harness
def test_reference_known_answers_hold():
    """The assertions of the reference's own unittest() blocks, on the reference's own functions (assert_almost_eq = 64 ulp or FLT_EPSILON)."""
    c = harness("camera")
    close = lambda a, b, tol=1e-5: np.allclose(np.array(a, dtype=f32), np.array(b, dtype=f32), atol=tol, rtol=64 * 1.2e-7)
    assert close(c["roundtrip_pixel"], [123.4, 345.0], tol=1e-4)                 # Cameras.cpp:164-173
    assert close(c["dir_400_300"], [0, 0, -1])                                   # unit_tests/Cameras.test.cpp:38-47
    assert close(c["dir_0_0"], [-0.685994, -0.514496, -0.514496], tol=1e-6)
    assert close(c["dir_800_600"], [0.685994, 0.514496, -0.514496], tol=1e-4)
    assert close(c["lookat_position"], [1, 3, 2])                                # Cameras.cpp:175-189
    assert close(c["lookat_tangent0"], [0, 1, 0]) and close(c["lookat_tangent1"], [0, 0, 1]) and close(c["lookat_tangent2"], [1, 0, 0])
    assert close(c["frame_tangent0"], [0, 1, 0])                                 # Technique.cpp:118-138
    assert close(c["frame_tangent1"], c["frame_expected1"]) and close(c["frame_tangent2"], c["frame_expected2"])
    assert close(c["frame2_col2"], c["frame2_expected"])                         # Technique.cpp:144-152
    # and the numpy restatement used by the tests above agrees with the reference's ray_direction to the bit
    assert np.array_equal(ray_direction(123.4, 345.0, 800, 600, c["flength_90"]), np.array(c["roundtrip_direction"], dtype=f32))
    assert np.array_equal(ray_direction(0.0, 0.0, 800, 600, c["flength_90"]), np.array(c["dir_0_0"], dtype=f32))


# ---------------------------------------------------------------- the device against the oracle binary
@pytest.mark.gpu
@needs_harness
def test_device_engine_matches(pin):
    from master_b200 import api
    raw, fl, db = api.debug_draws(int(pin["key"]), 8)
    assert raw.tolist() == pin["raw"]
    assert fl.tolist() == [float(f32(x)) for x in pin["floats"]]
    assert db.tolist() == [draw_double(np.uint32(pin["key"]), 2 * i) for i in range(8)]


@pytest.mark.gpu
@needs_harness
def test_device_primary_rays_and_light_samples_match_the_reference(pin, flat_scenes, device_scenes):
    from master_b200 import api
    flat = flat_scenes(SCENE)
    p = api.make_params(flat, technique="PT", width=pin["width"], height=pin["height"], seed=pin["seed"])
    pixels = [q["pixel"] for q in pin["primary"]]
    samples = [q["sample"] for q in pin["primary"]]
    dirs = api.debug_primary(p, pixels, samples)
    ref = np.array([q["direction"] for q in pin["primary"]], dtype=f32)
    assert np.array_equal(dirs.view(np.uint32), ref.view(np.uint32))  # bit for bit
    pos, mat, draws = device_scenes(SCENE).debug_light_samples([s["key"] for s in pin["lights"]])
    assert np.array_equal(pos.view(np.uint32), np.array([s["position"] for s in pin["lights"]], dtype=f32).view(np.uint32))
    assert mat.tolist() == [s["material_id"] for s in pin["lights"]]
    assert draws.tolist() == [s["draws"] for s in pin["lights"]]


@pytest.mark.gpu
@needs_harness
def test_device_camera_known_answers(flat_scenes):
    """ray_direction / pixel_position on the device at the reference's unittest() inputs: 800 x 600, fov 90 degrees, identity camera."""
    from master_b200 import api
    c = harness("camera")
    flat = flat_scenes("CornellBoxDiffuse")
    p = api.make_params(flat, technique="PT", width=800, height=600, focal_length_y=c["flength_90"])
    eye = [1, 0, 0, 0, 1, 0, 0, 0, 1]
    p.camera.view_to_world[:] = eye
    p.camera.world_to_view[:] = eye
    dirs, pix = api.debug_camera(p, [[123.4, 345.0], [400.0, 300.0], [0.0, 0.0], [800.0, 600.0]])
    ref = np.array([c["roundtrip_direction"], c["dir_400_300"], c["dir_0_0"], c["dir_800_600"]], dtype=f32)
    assert np.array_equal(dirs.view(np.uint32), ref.view(np.uint32))
    assert np.array_equal(pix[0].view(np.uint32), np.array(c["roundtrip_pixel"], dtype=f32).view(np.uint32))
    assert np.allclose(pix[0], [123.4, 345.0], atol=1e-4)
