"""The CPU oracle against the UNMODIFIED reference compiled here (oracle/_ref/libspt_ref*.so), live.
Skipped when the reference build is absent; the frozen equivalents are in test_oracle_golden.py."""
import numpy as np
import pytest


def same_scene(a, b):
    return a.n == b.n and all(np.array_equal(getattr(a, k), getattr(b, k)) for k in ("centers", "radii", "materials", "colors", "fuzz"))


def test_rng(oracle, reference):
    assert np.array_equal(oracle.uniform_seq(123, -1, 1, 5000), reference.uniform_seq(123, -1, 1, 5000))
    assert np.array_equal(oracle.shell_seq(9, 3000), reference.shell_seq(9, 3000))


@pytest.mark.parametrize("seed", [3, 17, 99])
def test_generators(oracle, reference, seed):
    assert same_scene(oracle.init_spheres(seed), reference.init_spheres(seed))
    assert same_scene(oracle.generate_spheres(seed), reference.generate_spheres(seed))


def test_first_hit_full_config1(oracle, reference):
    cam = reference.set_camera()
    scene = reference.generate_spheres(1)
    reference.set_config(640, 360, 1, 8)
    r = reference.first_hit()
    o = oracle.first_hit(scene, cam, 640, 360)
    for a, b in zip(r, o):
        assert np.array_equal(a, b)


def test_closest_hit_secondary_rays(oracle, reference):
    reference.set_camera()
    scene = reference.generate_spheres(2)
    rng = np.random.default_rng(1)
    n = 50000
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    pick = rng.integers(0, scene.n, n)
    surf = rng.normal(size=(n, 3)).astype(np.float32)
    surf /= np.linalg.norm(surf, axis=1, keepdims=True).astype(np.float32)
    o = (scene.centers[pick, :3] + surf * scene.radii[pick, None]).astype(np.float32)
    assert np.array_equal(reference.closest_hit_many(d, o), oracle.closest_hit_many(scene, d, o))


@pytest.mark.parametrize("tracer", [0, 1])
@pytest.mark.parametrize("which", ["init", "random"])
def test_render_bit_exact(oracle, reference, tracer, which):
    cam = reference.set_camera()
    scene = reference.init_spheres(5) if which == "init" else reference.generate_spheres(5)
    w, h, spp, depth = 64, 64, 3, 6
    reference.set_config(w, h, spp, depth)
    mr, rr = reference.render(tracer, 777)
    mo, ro, _ = oracle.render_refseq(scene, cam, w, h, spp, depth, tracer, 777)
    assert np.array_equal(mr, mo, equal_nan=True) and np.array_equal(rr, ro)


def test_wide_index_build_is_identical_below_256(oracle, reference, reference_wide):
    """The uint8_t -> uint32_t sphere index (needed for the >255-sphere configs) changes nothing for n <= 255."""
    cam = reference.set_camera()
    reference_wide.set_camera()
    scene = reference.generate_spheres(1)
    reference_wide.set_scene(scene)
    for r in (reference, reference_wide):
        r.set_config(80, 80, 2, 8)
    a = reference.render(1, 31)
    b = reference_wide.render(1, 31)
    assert np.array_equal(a[0], b[0], equal_nan=True) and np.array_equal(a[1], b[1])


def test_wide_index_above_255(oracle, reference_wide):
    cam = reference_wide.set_camera()
    rng = np.random.default_rng(3)
    n = 600
    from oracle import Scene

    scene = Scene(
        centers=np.concatenate([rng.uniform(-15, 15, (n, 1)), rng.uniform(0.3, 4, (n, 1)), rng.uniform(0, 25, (n, 1)), np.zeros((n, 1))], 1).astype(np.float32),
        radii=rng.uniform(0.3, 0.5, n).astype(np.float32),
        materials=rng.integers(1, 4, n).astype(np.uint8),
        colors=np.concatenate([rng.uniform(0, 255, (n, 3)), np.zeros((n, 1))], 1).astype(np.float32),
        fuzz=rng.uniform(0, 1, n).astype(np.float32),
    )
    reference_wide.set_scene(scene)
    w, h, spp, depth = 64, 48, 2, 8
    reference_wide.set_config(w, h, spp, depth)
    ri = reference_wide.first_hit()
    oi = oracle.first_hit(scene, cam, w, h)
    for a, b in zip(ri, oi):
        assert np.array_equal(a, b)
    assert ri[0].max() > 255
    for tracer in (0, 1):
        mr, rr = reference_wide.render(tracer, 5)
        mo, ro, _ = oracle.render_refseq(scene, cam, w, h, spp, depth, tracer, 5)
        assert np.array_equal(mr, mo, equal_nan=True) and np.array_equal(rr, ro)
