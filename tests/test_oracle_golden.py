"""The CPU oracle (oracle/spt_oracle.c) against golden vectors frozen from the UNMODIFIED reference
(tests/golden/make_golden.py).  Bit-exact everywhere: this is what pins the oracle."""
import os

import numpy as np
import pytest

from conftest import golden_path


def load(name):
    return np.load(golden_path(name))


def scene_from(z):
    from oracle import Scene

    return Scene.of(type("S", (), {k: z[k] for k in ("centers", "radii", "materials", "colors", "fuzz")}))


def cam_from(z):
    from oracle import Camera

    return Camera(z["view"], z["eye"], z["sky"])


def test_rng_front_end(oracle):
    g = load("rng.npz")
    assert np.array_equal(oracle.uniform_seq(77, -1, 1, 4096), g["uniform_m1_1_seed77"])
    assert np.array_equal(oracle.uniform_seq(78, -0.5, 0.5, 4096), g["uniform_half_seed78"])
    assert np.array_equal(oracle.uniform_seq(79, 0, 255, 1024), g["uniform_0_255_seed79"])
    assert np.array_equal(oracle.shell_seq(5, 2048), g["shell_seed5"])


def test_shell_sampler_distribution(oracle):
    """Random.hpp:115-127 keeps the OUTSIDE of the radius-0.5 ball: |v| in [0.5, sqrt(0.75)] (quirk Q5)."""
    v = oracle.shell_seq(11, 20000)
    ln = np.linalg.norm(v.astype(np.float64), axis=1)
    assert ln.min() >= 0.5 - 1e-6 and ln.max() <= np.sqrt(0.75) + 1e-6
    u = oracle.uniform_seq(3, -1, 1, 50000)
    assert u.min() >= -1 and u.max() < 1 and abs(u.mean()) < 0.02


def test_camera_basis(oracle):
    g = load("camera_default.npz")
    c = oracle.camera()
    assert np.array_equal(c.view, g["view"]) and np.array_equal(c.eye, g["eye"]) and np.array_equal(c.sky, g["sky"])
    g = load("camera_offaxis.npz")  # exercises the reference's Cross() z-lane bug (Math.hpp:113-120)
    c = oracle.camera(tuple(g["eye_in"]), tuple(g["look_in"]))
    assert np.array_equal(c.view, g["view"])


@pytest.mark.parametrize("seed", [1, 2])
def test_scene_generators(oracle, seed):
    for name, gen in (("init", oracle.init_spheres), ("random", oracle.generate_spheres)):
        g = load(f"scene_{name}_seed{seed}.npz")
        s = gen(seed)
        for k in ("centers", "radii", "materials", "colors", "fuzz"):
            assert np.array_equal(getattr(s, k), g[k]), (name, k)


@pytest.mark.parametrize("name", ["init", "random"])
@pytest.mark.parametrize("size", [(160, 90), (96, 96)])
def test_first_hit(oracle, name, size):
    g = load("first_hit.npz")
    scene = scene_from(load(f"scene_{name}_seed1.npz"))
    cam = cam_from(load("camera_default.npz"))
    w, h = size
    idx, t, nrm, d = oracle.first_hit(scene, cam, w, h)
    key = f"{name}_{w}x{h}"
    assert np.array_equal(idx, g[key + "_idx"].astype(np.uint32))
    assert np.array_equal(t, g[key + "_t"])
    assert np.array_equal(nrm, g[key + "_normal"])
    assert np.array_equal(d, g[key + "_dir"])


@pytest.mark.parametrize("name", ["init", "random"])
@pytest.mark.parametrize("tracer", [0, 1])
def test_render_replay_of_reference_sequence(oracle, name, tracer):
    """RenderSegment / RenderSegmentTask replayed with the reference's own RNG sequence: identical floats and bytes."""
    g = load("renders_refseq.npz")
    w, h, spp, depth, seed = [int(v) for v in g["config"]]
    scene = scene_from(load(f"scene_{name}_seed1.npz"))
    cam = cam_from(load("camera_default.npz"))
    mean, rgb, ctr = oracle.render_refseq(scene, cam, w, h, spp, depth, tracer, seed)
    assert np.array_equal(mean, g[f"{name}_tracer{tracer}_mean"], equal_nan=True)
    assert np.array_equal(rgb, g[f"{name}_tracer{tracer}_rgb"])
    assert ctr["samples"] == w * h * spp and ctr["tests"] == ctr["segments"] * scene.n


@pytest.mark.parametrize("name", ["init", "random"])
def test_render_replay_segment(oracle, name):
    g = load("renders_refseq.npz")
    w, h, spp, depth, _ = [int(v) for v in g["config"]]
    scene = scene_from(load(f"scene_{name}_seed1.npz"))
    cam = cam_from(load("camera_default.npz"))
    mean, rgb, _ = oracle.render_refseq(scene, cam, w, h, spp, depth, 0, 99, seg=(5, 30, 7, 47))
    assert np.array_equal(mean, g[f"{name}_seg_mean"]) and np.array_equal(rgb, g[f"{name}_seg_rgb"])


def test_write_pixel(oracle):
    g = load("write_pixel.npz")
    got = np.array([oracle.write_pixel((v, v * 0.5, 255 - v)) for v in g["values"]], np.uint8)
    assert np.array_equal(got, g["bytes"])


def test_pow5_modes(oracle):
    """The device-matching x^5 (binary64 chain, one rounding) stays within 1 ulp of the libm powf the
    compiled reference calls, and pow(q, 2.f) is exactly one rounded product."""
    import math

    xs = np.concatenate([np.linspace(0, 1, 20001), np.linspace(-1e-6, 1e-6, 101)]).astype(np.float32)
    a = np.array([oracle.pow5(float(x), 0) for x in xs], np.float32)
    b = np.array([oracle.pow5(float(x), 1) for x in xs], np.float32)
    ulp = np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1
    exact = np.array([np.float32(float(x) ** 5) for x in xs], np.float32)
    assert np.array_equal(a, exact)
    q = np.float32(-0.5) / np.float32(2.5)
    assert np.float32(q * q) == np.float32(math.pow(float(q), 2.0))


def test_counter_mode_matches_converged_reference(oracle):
    """Counter-RNG mode is statistically the same estimator: a moderate-spp render is within Monte-Carlo
    noise of the reference's converged render (tight check lives in the GPU tests)."""
    g = load("converged_task.npz")
    w, h, _, depth = [int(v) for v in g["init_config"]]
    scene = scene_from(load("scene_init_seed1.npz"))
    cam = cam_from(load("camera_default.npz"))
    spp = 64
    acc, _ = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=10, seed=5)
    mean = acc[..., :3] / acc[..., 3:4]
    ref = g["init_mean"][::-1]  # fixture is at the reference's flipped slots
    rmse = float(np.sqrt(np.mean(((mean - ref) / 255.0) ** 2)))
    assert rmse < 0.03
    bias = float(np.abs(np.mean((mean - ref) / 255.0)))
    assert bias < 2e-3


def test_restatement_independent_of_pass_cap_when_paths_are_short(oracle):
    scene = scene_from(load("scene_init_seed1.npz"))
    cam = cam_from(load("camera_default.npz"))
    a, ca = oracle.render_accum(scene, cam, 32, 32, 4, 8, pass_cap=0, seed=2)
    b, cb = oracle.render_accum(scene, cam, 32, 32, 4, 8, pass_cap=1000, seed=2)
    assert np.array_equal(a, b) and cb["dropped"] == 0


def test_shell_acceptance_needs_no_square_root():
    """The CUDA path tests lensq < 0.25f where Random.hpp:124 tests sqrtf(lensq) < 0.5f.  They agree for
    EVERY float: sqrt is monotone, so it is enough to look at both sides of the boundary."""
    q = np.float32(0.25)
    around = q.view(np.uint32).astype(np.int64) + np.arange(-100000, 100001)
    xs = around.astype(np.uint32).view(np.float32)
    assert np.array_equal(np.sqrt(xs) < np.float32(0.5), xs < q)
    below, at = np.nextafter(q, np.float32(0)), q
    assert np.sqrt(below) < np.float32(0.5) and not (np.sqrt(at) < np.float32(0.5))
    rng = np.random.default_rng(0)
    xs = rng.uniform(0, 0.75, 2_000_000).astype(np.float32)
    assert np.array_equal(np.sqrt(xs) < np.float32(0.5), xs < q)


def test_shell_fixed_point_fields_through_the_mantissa():
    """spt_device.cuh's shell_attempt builds k * 2^-21 - 0.5 (2^-22 for z) by placing k in the mantissa of a float in [1, 2) and
    subtracting 1.5 instead of convert * scale + min.  Both are exact, so they agree for EVERY k, the +0 at the midpoint
    included -- the definition the oracle implements (canonical * (max - min) + min) is unchanged."""
    for bits, shift in ((21, 2), (22, 1)):
        k = np.arange(1 << bits, dtype=np.uint32)
        want = (k.astype(np.float32) * np.float32(2.0 ** -bits) + np.float32(-0.5)).astype(np.float32)
        got = ((np.uint32(0x3F800000) | (k << np.uint32(shift))).view(np.float32) + np.float32(-1.5)).astype(np.float32)
        assert np.array_equal(want.view(np.uint32), got.view(np.uint32))
