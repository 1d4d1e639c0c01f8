"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same inputs.

Bars (BASELINE.json north_star): first-hit sphere index bit-exact, t / normal within 4 ulp
(we require and get 0 ulp); renders within RMSE <= 1e-3 / PSNR >= 45 dB of the reference.
Because the CUDA path spells every rounded operation in the reference's order and shares the
counter-based RNG definition with the oracle, the float accumulators are compared for EQUALITY.
"""
import numpy as np
import pytest

import simplepathtracer_b200 as spt

pytestmark = pytest.mark.gpu


def ulp_diff(a, b):
    a = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    a = np.where(a < 0, -(a & 0x7FFFFFFF), a)
    b = np.where(b < 0, -(b & 0x7FFFFFFF), b)
    return np.abs(a - b)


def assert_accum_equal(gpu, cpu):
    assert gpu.shape == cpu.shape
    same = (gpu == cpu) | (np.isnan(gpu) & np.isnan(cpu))
    if not same.all():
        bad = np.argwhere(~same)
        raise AssertionError(f"{len(bad)} accumulator words differ; first at {bad[0]}: gpu={gpu[tuple(bad[0])]} cpu={cpu[tuple(bad[0])]}")


# ----------------------------------------------------------------------------- first hit
@pytest.mark.parametrize("which", ["init", "random"])
@pytest.mark.parametrize("size", [(640, 360), (256, 256)])
@pytest.mark.parametrize("flags", [0, spt.FLAG_STRICT_LOOP])
def test_first_hit_bit_exact(tracer, oracle, cam, scene_init, scene_random, which, size, flags):
    scene = scene_init if which == "init" else scene_random
    w, h = size
    tracer.set_scene(scene)
    gi, gt, gn = tracer.first_hit(cam, w, h, flags=flags)
    oi, ot, on, _ = oracle.first_hit(scene, cam, w, h)
    assert np.array_equal(gi, oi)  # index: bit-exact
    assert ulp_diff(gt, ot).max() == 0  # t: north star allows 4 ulp
    assert ulp_diff(gn, on).max() == 0  # normal: north star allows 4 ulp
    assert len(np.unique(oi)) > 3


def test_first_hit_config2_full_resolution(tracer, oracle, cam):
    """BASELINE config 2: single diffuse sphere + ground sphere, 1920x1080, all 2.07 M pinhole rays."""
    scene = spt.Scene(
        centers=np.array([[0, -1000.5, 0, 0], [0, 1, 0, 0]], np.float32),
        radii=np.array([1000.0, 0.5], np.float32),
        materials=np.array([spt.DIFFUSE, spt.DIFFUSE], np.uint8),
        colors=np.array([[30, 144, 255, 0], [255, 100, 230, 0]], np.float32),
        fuzz=np.zeros(2, np.float32),
    )
    tracer.set_scene(scene)
    gi, gt, gn = tracer.first_hit(cam, 1920, 1080)
    oi, ot, on, _ = oracle.first_hit(scene, cam, 1920, 1080)
    assert np.array_equal(gi, oi)
    assert ulp_diff(gt, ot).max() == 0
    assert ulp_diff(gn, on).max() == 0
    assert set(np.unique(oi)) == {0, 1, 2}


def test_first_hit_segment(tracer, oracle, cam, scene_random):
    tracer.set_scene(scene_random)
    seg = (40, 120, 100, 300)  # yBegin, yEnd, xBegin, xEnd
    gi, gt, gn = tracer.first_hit(cam, 640, 360, seg=seg)
    oi, ot, on, _ = oracle.first_hit(scene_random, cam, 640, 360, seg=seg)
    assert np.array_equal(gi, oi) and np.array_equal(gt, ot) and np.array_equal(gn, on)


def secondary_like_rays(scene, n, seed):
    """Origins on sphere surfaces (and some anywhere, some far away), arbitrary unit directions,
    a share of them aimed to graze a sphere's silhouette (the pre-filter's hard case)."""
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    pick = rng.integers(0, scene.n, n)
    surf = rng.normal(size=(n, 3)).astype(np.float32)
    surf /= np.linalg.norm(surf, axis=1, keepdims=True).astype(np.float32)
    o = (scene.centers[pick, :3].astype(np.float64) + surf * scene.radii[pick, None].astype(np.float64)).astype(np.float32)
    q = n // 8
    o[:q] = rng.uniform(-20, 20, size=(q, 3)).astype(np.float32)
    o[q : q + q // 4] = rng.uniform(-3000, 3000, size=(q // 4, 3)).astype(np.float32)  # far origins: widened tC test
    # grazing rays: aim at a point on the silhouette circle of another sphere, +- a hair
    k = slice(2 * q, 5 * q)
    m = k.stop - k.start
    tgt = rng.integers(0, scene.n, m)
    c = scene.centers[tgt, :3].astype(np.float64)
    r = scene.radii[tgt].astype(np.float64)
    to = c - o[k].astype(np.float64)
    dist = np.linalg.norm(to, axis=1, keepdims=True)
    axis = to / np.maximum(dist, 1e-9)
    perp = np.cross(axis, rng.normal(size=(m, 3)))
    perp /= np.maximum(np.linalg.norm(perp, axis=1, keepdims=True), 1e-9)
    scale = (1.0 + rng.normal(scale=3e-5, size=(m, 1))) * r[:, None]
    aim = c + perp * scale - o[k]
    aim /= np.maximum(np.linalg.norm(aim, axis=1, keepdims=True), 1e-9)
    d[k] = aim.astype(np.float32)
    return d, o


@pytest.mark.parametrize("which", ["init", "random"])
def test_closest_hit_random_rays(tracer, oracle, scene_init, scene_random, which):
    scene = scene_init if which == "init" else scene_random
    d, o = secondary_like_rays(scene, 200_000, 7)
    tracer.set_scene(scene)
    want = oracle.closest_hit_many(scene, d, o)
    got, evals = tracer.closest_hit(d, o, want_evals=True)  # product path: pre-filter + exact candidates
    assert np.array_equal(got, want)
    assert np.array_equal(tracer.closest_hit(d, o, flags=spt.FLAG_STRICT_LOOP), want)
    assert (want < scene.n).mean() > 0.3
    assert evals.mean() < 0.25 * scene.n + 4  # the pre-filter really filters


def test_prefilter_never_rejects_a_true_hit(tracer, scene_random, scene_init):
    """Property at scale: pre-filter + exact candidates == strict loop on millions of rays (device-side
    cross-check; 0xFFFFFFFF marks a disagreement)."""
    for scene, seed in ((scene_random, 11), (scene_init, 12)):
        tracer.set_scene(scene)
        for rep in range(3):
            d, o = secondary_like_rays(scene, 1_000_000, seed + 100 * rep)
            got = tracer.closest_hit(d, o, flags=spt.FLAG_CROSS_CHECK)
            assert not (got == 0xFFFFFFFF).any()


# ----------------------------------------------------------------------------- renders
@pytest.mark.parametrize("which", ["init", "random"])
@pytest.mark.parametrize("pass_cap", [0, 10])
@pytest.mark.parametrize("flags", [0, spt.FLAG_SINGLE_KERNEL, spt.FLAG_STRICT_LOOP])
def test_render_accumulators_bit_exact(tracer, oracle, cam, scene_init, scene_random, which, pass_cap, flags):
    scene = scene_init if which == "init" else scene_random
    w, h, spp, depth = 96, 64, 8, 8
    tracer.set_scene(scene)
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=42, want_accum=True, flags=flags)
    oacc, octr = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=pass_cap, seed=42)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))
    assert acc[..., 3].max() == spp


def test_render_pass_cap_drops_samples(tracer, oracle, cam, scene_init):
    """With a tiny pass cap the task tracer's drop rule shows up as count < spp, identically on both sides."""
    w, h, spp, depth = 64, 64, 8, 8
    tracer.set_scene(scene_init)
    _, acc = tracer.render(cam, w, h, spp, depth, pass_cap=1, seed=3, want_accum=True)
    oacc, octr = oracle.render_accum(scene_init, cam, w, h, spp, depth, pass_cap=1, seed=3)
    assert_accum_equal(acc, oacc)
    assert octr["dropped"] > 0 and acc[..., 3].min() < spp


def test_render_counters_match_oracle(tracer, oracle, cam, scene_random):
    w, h, spp, depth = 64, 48, 4, 8
    tracer.set_scene(scene_random)
    tracer.reset_counters()
    tracer.render(cam, w, h, spp, depth, seed=9, flags=spt.FLAG_COUNTERS)
    c = tracer.counters()
    _, o = oracle.render_accum(scene_random, cam, w, h, spp, depth, seed=9)
    assert c["segments"] == o["segments"]
    assert c["tests"] == o["tests"] == o["segments"] * scene_random.n
    assert c["samples"] == o["samples"] == w * h * spp


def test_render_segment_and_ragged_sizes(tracer, oracle, cam, scene_random):
    """Segments that do not align with the 32x8 block tile, odd image sizes, 1-pixel segments, empty segments."""
    w, h, spp, depth = 70, 45, 3, 5
    tracer.set_scene(scene_random)
    for seg in [(0, 45, 0, 70), (3, 29, 11, 58), (44, 45, 69, 70), (10, 10, 5, 20)]:
        rgb0 = np.full((h, w, 3), 7, np.uint8)
        rgb, acc = tracer.render(cam, w, h, spp, depth, seed=5, seg=seg, rgb8=rgb0.copy(), want_accum=True)
        oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, seed=5, seg=seg)
        assert_accum_equal(acc, oacc)
        y0, y1, x0, x1 = seg
        want = rgb0.copy()
        if y1 > y0 and x1 > x0:
            full = oracle.resolve(oacc, seg=seg)
            want[h - y1 : h - y0, x0:x1] = full[h - y1 : h - y0, x0:x1]
        assert np.array_equal(rgb, want)  # bytes outside the segment are untouched


def test_sample_ranges_compose(tracer, oracle, cam, scene_init):
    """sample_begin/sample_end shards (the multi-GPU split) reproduce the same per-sample radiance."""
    import torch

    w, h, depth = 48, 32, 6
    tracer.set_scene(scene_init)
    tracer.set_stream(torch.cuda.current_stream().cuda_stream)  # same stream as the torch fills below
    acc = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    parts = []
    for s0, s1 in [(0, 5), (5, 6), (6, 16)]:
        part = torch.zeros_like(acc)
        tracer.accumulate(cam, w, h, s0, s1, depth, part, seed=11)
        tracer.synchronize()
        o, _ = oracle.render_accum(scene_init, cam, w, h, s1 - s0, depth, seed=11, sample_begin=s0)
        assert_accum_equal(part.cpu().numpy(), o)
        parts.append(o)
    # accumulating ranges one after another into ONE buffer == oracle doing the same
    for s0, s1 in [(0, 5), (5, 6), (6, 16)]:
        tracer.accumulate(cam, w, h, s0, s1, depth, acc, seed=11)
    tracer.synchronize()
    want = np.zeros((h, w, 4), np.float32)
    for s0, s1 in [(0, 5), (5, 6), (6, 16)]:
        oracle.render_accum(scene_init, cam, w, h, s1 - s0, depth, seed=11, sample_begin=s0, accum=want)
    assert_accum_equal(acc.cpu().numpy(), want)
    tracer.set_stream(None)


def test_resolve_device(tracer, oracle):
    import torch

    rng = np.random.default_rng(0)
    w, h = 37, 19
    acc = rng.uniform(0, 255 * 9, size=(h, w, 4)).astype(np.float32)
    acc[..., 3] = rng.integers(1, 12, size=(h, w)).astype(np.float32)
    acc[0, 0] = [0, 0, 0, 0]  # count 0 -> NaN colour -> byte 0 (TaskBasedPathTracer.hpp:198)
    tracer.set_stream(torch.cuda.current_stream().cuda_stream)
    d_acc = torch.from_numpy(acc).cuda()
    d_rgb = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda")
    tracer.resolve(d_acc, w, h, d_rgb)
    tracer.synchronize()
    tracer.set_stream(None)
    assert np.array_equal(d_rgb.cpu().numpy(), oracle.resolve(acc))


def test_config1_full_size(tracer, oracle, cam, scene_random):
    """BASELINE config 1 at full size: random-spheres scene, 640x360, 16 spp, depth 8, task-tracer semantics."""
    w, h, spp, depth = 640, 360, 16, 8
    tracer.set_scene(scene_random)
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=1, want_accum=True)
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=1)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))


def test_determinism(tracer, cam, scene_random):
    tracer.set_scene(scene_random)
    a = tracer.render(cam, 128, 72, 8, 8, seed=77, want_accum=True)[1]
    b = tracer.render(cam, 128, 72, 8, 8, seed=77, want_accum=True)[1]
    c = tracer.render(cam, 128, 72, 8, 8, seed=78, want_accum=True)[1]
    assert np.array_equal(a, b) and not np.array_equal(a, c)


def test_errors(tracer, cam, scene_init):
    tracer.set_scene(scene_init)
    with pytest.raises(spt.SptError):
        tracer.render(cam, 64, 64, 1, 4, seg=(0, 65, 0, 64))  # segment outside the image
    bad = spt.Camera.of(cam)
    bad.eye = np.array([0, 1, -3, 1], np.float32)
    with pytest.raises(spt.SptError):
        tracer.render(bad, 64, 64, 1, 4)


# ----------------------------------------------------------------------------- statistical parity with the reference
def _psnr_rmse(mean, ref):
    err = (mean.astype(np.float64) - ref.astype(np.float64)) / 255.0  # radiance is on a 0..255 scale (Globals.hpp:29)
    mse = float(np.mean(err ** 2))
    return 10.0 * np.log10(1.0 / mse), float(np.sqrt(mse))


def _converged_case(which, scene_init, scene_random):
    """(scene, w, h, ref_spp, depth, reference mean at image rows) of one converged REFERENCE render frozen under tests/golden/."""
    from conftest import golden_path

    if which in ("init", "random"):
        g = np.load(golden_path("converged_task.npz"))
        scene = scene_init if which == "init" else scene_random
    else:  # BASELINE config 2 (its "convergence check") and a config-4 style metal/glass scene (wide-index reference build)
        g = np.load(golden_path("converged_task2.npz"))
        scene = spt.Scene(**{k: g[f"{which}_scene_{k}"] for k in ("centers", "radii", "materials", "colors", "fuzz")})
    w, h, ref_spp, depth = [int(v) for v in g[f"{which}_config"]]
    return scene, w, h, ref_spp, depth, g[f"{which}_mean"][::-1]  # fixture rows are at the reference's flipped slots


@pytest.mark.parametrize("which", ["init", "random", "c2", "c4"])
def test_converged_render_matches_reference_own_render(tracer, cam, scene_init, scene_random, which):
    """North-star gate: a converged CUDA render against the REFERENCE'S OWN high-spp CPU render
    (RenderSegmentTask over square tiles, frozen in tests/golden/converged_task*.npz by make_golden.py):
    PSNR >= 45 dB in linear radiance, with the RMSE recorded.  The two use unrelated random numbers, so this
    is the statistical check that the counter-based RNG and the restated shading are the same estimator.
    Cases: the 10-sphere scene, the random scene (config 1/3's), BASELINE config 2 at depth 4, and a 501-sphere
    mirror/glass scene at depth 50 (config 4's, where a path draws the most random numbers)."""
    scene, w, h, ref_spp, depth, ref = _converged_case(which, scene_init, scene_random)
    tracer.set_scene(scene)
    spp = 65536 if which != "c4" else 32768
    _, acc = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=2024, want_accum=True)
    mean = acc[..., :3] / acc[..., 3:4]
    psnr, rmse = _psnr_rmse(mean, ref)
    print(f"converged {which}: {w}x{h} depth {depth}, CUDA {spp} spp vs reference {ref_spp} spp: PSNR {psnr:.2f} dB, RMSE {rmse:.3e}")
    # The reference render carries the Monte-Carlo noise of its own ref_spp samples, which alone is an RMSE of a few 1e-3:
    # the north star's "RMSE <= 1e-3 OR PSNR >= 45 dB" is asserted on its PSNR arm, and the RMSE must stay at the level that
    # noise explains (sigma / sqrt(ref_spp) with sigma <= 0.5 on the 0..1 scale).
    assert psnr >= 45.0, (psnr, rmse)
    assert rmse <= 0.5 / np.sqrt(ref_spp) + 1e-3, (psnr, rmse)
    # and no systematic offset: mean signed error well below the noise floor, per channel
    bias = np.mean((mean.astype(np.float64) - ref) / 255.0, axis=(0, 1))
    assert np.abs(bias).max() < 5e-4, bias


# ----------------------------------------------------------------------------- wavefront variant
@pytest.mark.parametrize("which", ["init", "random"])
@pytest.mark.parametrize("pass_cap", [0, 10, 1])
@pytest.mark.parametrize("feed", [0, spt.FLAG_STREAMED])
def test_wavefront_variant_bit_exact(tracer, oracle, cam, scene_init, scene_random, which, pass_cap, feed):
    """The wavefront kernel (persistent CTAs, per-material queues) computes the same per-sample radiance
    and the same in-order per-pixel sums as the megakernel and the oracle -- with the list resident in shared memory and
    with its intersect stage fed by the TMA chunk ring (what config 5's list takes)."""
    scene = scene_init if which == "init" else scene_random
    w, h, spp, depth = 100, 45, 6, 8  # not a multiple of the 32x8 tile
    tracer.set_scene(scene)
    tracer.reset_counters()
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=21, want_accum=True, variant=spt.WAVEFRONT,
                             flags=spt.FLAG_COUNTERS | feed)
    c = tracer.counters()
    oacc, octr = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=pass_cap, seed=21)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))
    assert c["segments"] == octr["segments"] and c["samples"] == octr["samples"] and c["dropped"] == octr["dropped"]


def test_wavefront_segment(tracer, oracle, cam, scene_random):
    w, h, spp, depth = 70, 45, 3, 5
    tracer.set_scene(scene_random)
    seg = (3, 29, 11, 58)
    _, acc = tracer.render(cam, w, h, spp, depth, seed=5, seg=seg, want_accum=True, variant=spt.WAVEFRONT)
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, seed=5, seg=seg)
    assert_accum_equal(acc, oacc)


# ----------------------------------------------------------------------------- BASELINE configs 4 and 5
@pytest.fixture(scope="module")
def scene_metal_glass():
    return spt.generate_metal_glass(4, 500)  # config 4: ground + 500 mirror / glass spheres (N > 255: wide index)


@pytest.fixture(scope="module")
def scene_lattice():
    return spt.generate_lattice(5, 128)  # config 5: ground + 16 383 spheres (list larger than shared memory)


@pytest.mark.parametrize("variant", [spt.MEGAKERNEL, spt.WAVEFRONT])
@pytest.mark.parametrize("pass_cap", [0, 10])
def test_config4_metal_glass_bit_exact(tracer, oracle, cam, scene_metal_glass, variant, pass_cap):
    """Divergence stress: every small sphere is specular, paths are long.  Accumulators equal the oracle's."""
    w, h, spp, depth = 96, 56, 4, 50
    tracer.set_scene(scene_metal_glass)
    tracer.reset_counters()
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=4, want_accum=True, variant=variant,
                             flags=spt.FLAG_COUNTERS)
    c = tracer.counters()
    oacc, octr = oracle.render_accum(scene_metal_glass, cam, w, h, spp, depth, pass_cap=pass_cap, seed=4)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))
    assert c["segments"] == octr["segments"] and c["dropped"] == octr["dropped"]
    assert c["segments"] > 2.0 * w * h * spp


def test_config4_first_hit_wide_index(tracer, oracle, cam, scene_metal_glass):
    tracer.set_scene(scene_metal_glass)
    gi, gt, gn = tracer.first_hit(cam, 480, 270)
    oi, ot, on, _ = oracle.first_hit(scene_metal_glass, cam, 480, 270)
    assert np.array_equal(gi, oi) and np.array_equal(gt, ot) and np.array_equal(gn, on)
    assert gi[gi < scene_metal_glass.n].max() > 255  # indices the reference's uint8_t could not hold


@pytest.mark.parametrize("variant,flags", [(spt.MEGAKERNEL, 0), (spt.MEGAKERNEL, spt.FLAG_SINGLE_KERNEL), (spt.MEGAKERNEL, spt.FLAG_RESIDENT),
                                           (spt.MEGAKERNEL, spt.FLAG_STRICT_LOOP), (spt.WAVEFRONT, 0)])
def test_config5_lattice_bit_exact(tracer, oracle, cam, scene_lattice, variant, flags):
    """16 384 spheres: neither the geometry nor the pre-filter records fit in shared memory."""
    w, h, spp, depth = 64, 40, 2, 16
    tracer.set_scene(scene_lattice)
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=5, want_accum=True, variant=variant, flags=flags)
    oacc, _ = oracle.render_accum(scene_lattice, cam, w, h, spp, depth, pass_cap=10, seed=5)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))


def test_config5_first_hit_and_prefilter(tracer, oracle, cam, scene_lattice):
    tracer.set_scene(scene_lattice)
    gi, gt, gn = tracer.first_hit(cam, 320, 180)
    oi, ot, on, _ = oracle.first_hit(scene_lattice, cam, 320, 180)
    assert np.array_equal(gi, oi) and np.array_equal(gt, ot) and np.array_equal(gn, on)
    rng = np.random.default_rng(55)
    d = rng.normal(size=(20000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    o = (rng.uniform(-1, 1, size=(20000, 3)) * np.array([80, 0, 80]) + np.array([0, 1.0, 80])).astype(np.float32)
    got = tracer.closest_hit(d, o, flags=spt.FLAG_CROSS_CHECK)
    assert not (got == 0xFFFFFFFF).any()
    assert np.array_equal(got, oracle.closest_hit_many(scene_lattice, d, o))


def _full_size_case(which, scene_random, scene_metal_glass, scene_lattice):
    scene, depth, spp = {"c3": (scene_random, 50, 4), "c4": (scene_metal_glass, 50, 4), "c5": (scene_lattice, 16, 2)}[which]
    w, h = (3840, 2160) if which == "c3" else (1920, 1080)
    return scene, w, h, depth, spp


def _pick_crops(oracle, scene, cam, w, h, cw=64, ch=32):
    """Crops (yBegin, yEnd, xBegin, xEnd) of the full-size image chosen from a coarse first-hit map of the ORACLE (the ray
    generator is scale-consistent: pixel (x, y) of a w/8 x h/8 image looks where (8x, 8y) of the full one does): the largest
    glass sphere, the largest mirror, the horizon line, the busiest far field of small spheres, two image corners, and a
    crop that is not aligned with any CTA tile."""
    f = 8
    idx = oracle.first_hit(scene, cam, w // f, h // f)[0].reshape(h // f, w // f)
    crops = {}

    def around(name, y, x):
        y0 = int(np.clip(y * f - ch // 2, 0, h - ch))
        x0 = int(np.clip(x * f - cw // 2, 0, w - cw))
        crops[name] = (y0, y0 + ch, x0, x0 + cw)

    hit = idx < scene.n
    for name, mat in (("glass", spt.REFRACTIVE), ("mirror", spt.REFLECTIVE)):
        cand = [i for i in np.unique(idx[hit]) if scene.materials[i] == mat]
        if cand:
            best = max(cand, key=lambda i: (idx == i).sum())
            ys, xs = np.nonzero(idx == best)
            around(name, ys.mean(), xs.mean())
    # horizon: a pixel whose vertical neighbour is sky while it is not
    edge = np.argwhere(hit[1:, :] != hit[:-1, :])
    if len(edge):
        y, x = edge[len(edge) // 2]
        around("horizon", y + 0.5, x)
    # far field: the coarse 8x4 window with the most distinct spheres
    best, where = -1, (0, 0)
    for y in range(0, h // f - 4, 2):
        for x in range(0, w // f - 8, 4):
            k = len(np.unique(idx[y:y + 4, x:x + 8]))
            if k > best:
                best, where = k, (y + 2, x + 4)
    around("far_field", *where)
    crops["corner00"] = (0, ch, 0, cw)
    crops["corner11"] = (h - ch, h, w - cw, w)
    crops["ragged"] = (h // 2 + 3, h // 2 + 3 + ch - 5, w // 3 + 5, w // 3 + 5 + cw - 7)
    return crops


@pytest.mark.parametrize("which", ["c3", "c4", "c5"])
def test_full_size_render_equals_oracle_on_crops(tracer, oracle, cam, scene_random, scene_metal_glass, scene_lattice, which):
    """BASELINE configs 3, 4, 5 at their FULL image size, depth and tracer semantics (pass_cap 10), rendered whole by the
    DEFAULT path (split kernels; streamed list for c5): seven crops of the accumulator -- glass sphere, mirror, horizon, far
    field, two corners, a ragged rectangle -- equal the oracle's render of the same crop bit for bit
    (RenderSegmentTask, TaskBasedPathTracer.hpp:54-206; Collision.hpp:87-109)."""
    import torch

    scene, w, h, depth, spp = _full_size_case(which, scene_random, scene_metal_glass, scene_lattice)
    tracer.set_scene(scene)
    tracer.set_stream(torch.cuda.current_stream().cuda_stream)
    acc = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    tracer.accumulate(cam, w, h, 0, spp, depth, acc, pass_cap=10, seed=8)
    torch.cuda.synchronize()
    tracer.set_stream(None)
    crops = _pick_crops(oracle, scene, cam, w, h)
    assert len(crops) >= 6, crops
    for name, (y0, y1, x0, x1) in crops.items():
        oacc, _ = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=10, seed=8, seg=(y0, y1, x0, x1))
        got = acc[y0:y1, x0:x1].cpu().numpy()
        try:
            assert_accum_equal(got, oacc[y0:y1, x0:x1])
        except AssertionError as e:
            raise AssertionError(f"{which} crop {name} {(y0, y1, x0, x1)}: {e}") from None
    # and the same crops through the host-buffer entry point (spt_cuda_render with a segment = a RenderSegmentData tile)
    name, seg = next(iter(crops.items()))
    rgb, part = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=8, seg=seg, want_accum=True)
    y0, y1, x0, x1 = seg
    assert_accum_equal(part[y0:y1, x0:x1], acc[y0:y1, x0:x1].cpu().numpy())
    full = oracle.resolve(part, seg=seg)
    assert np.array_equal(rgb[h - y1:h - y0, x0:x1], full[h - y1:h - y0, x0:x1])


@pytest.mark.parametrize("which", ["c3", "c4", "c5"])
def test_full_size_properties(tracer, cam, scene_random, scene_metal_glass, scene_lattice, which):
    """Size-independent properties at BASELINE's full image sizes (the oracle comparison at these sizes is
    test_full_size_render_equals_oracle_on_crops): every code path (split, single kernel, streamed, resident, the opt-in grid,
    strict loop, wavefront) produces identical accumulators over the WHOLE image, sample ranges compose, and count + dropped == spp."""
    import torch

    scene, depth, spp = {"c3": (scene_random, 50, 4), "c4": (scene_metal_glass, 50, 4), "c5": (scene_lattice, 16, 1)}[which]
    w, h = (3840, 2160) if which == "c3" else (1920, 1080)
    tracer.set_scene(scene)
    outs = []
    for variant, flags in ((spt.MEGAKERNEL, 0), (spt.MEGAKERNEL, spt.FLAG_SINGLE_KERNEL), (spt.MEGAKERNEL, spt.FLAG_STREAMED),
                           (spt.MEGAKERNEL, spt.FLAG_RESIDENT), (spt.MEGAKERNEL, spt.FLAG_GRID),
                           (spt.MEGAKERNEL, spt.FLAG_STRICT_LOOP), (spt.WAVEFRONT, 0)):
        if which == "c5" and flags == spt.FLAG_STRICT_LOOP:  # the strict loop over 16k spheres at full HD is minutes: use a band of it
            continue
        acc = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
        tracer.set_stream(torch.cuda.current_stream().cuda_stream)
        tracer.accumulate(cam, w, h, 0, spp, depth, acc, pass_cap=10, seed=8, variant=variant, flags=flags)
        torch.cuda.synchronize()
        outs.append(acc)
    for other in outs[1:]:
        assert torch.equal(outs[0], other)
    cnt = outs[0][..., 3]
    assert float(cnt.max()) == spp and float(cnt.min()) >= 0
    if spp >= 2:  # sample ranges [0, k) + [k, spp): the running sums continue from the accumulator, so the bits are the same
        acc = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
        tracer.accumulate(cam, w, h, 0, 1, depth, acc, pass_cap=10, seed=8)
        tracer.accumulate(cam, w, h, 1, spp, depth, acc, pass_cap=10, seed=8)
        torch.cuda.synchronize()
        assert torch.equal(acc, outs[0])
    tracer.set_stream(None)


# ----------------------------------------------------------------------------- streamed sphere list (TMA chunk ring)
@pytest.mark.parametrize("which", ["init", "random", "metal_glass", "lattice40"])
@pytest.mark.parametrize("pass_cap", [0, 10])
@pytest.mark.parametrize("single", [0, spt.FLAG_SINGLE_KERNEL])
def test_streamed_megakernel_bit_exact(tracer, oracle, cam, scene_init, scene_random, scene_metal_glass, which, pass_cap, single, monkeypatch):
    """SPT_FLAG_STREAMED forces the chunk-ring kernel on lists of every length: a fraction of one chunk (10, 149),
    one almost full chunk (501) and several chunks with a ragged tail (1600 spheres = 3 full chunks + 63 slots)."""
    scene = {"init": scene_init, "random": scene_random, "metal_glass": scene_metal_glass}.get(which) or spt.generate_lattice(7, 40)
    w, h, spp, depth = 100, 45, 5, 12  # ragged tiles: CTAs with dead warps and dead lanes must keep the ring moving
    monkeypatch.setattr(tracer, "split_chunk", 2)  # SptParams.split_chunk: 3 chunks of samples through the two staging buffers
    tracer.set_scene(scene)
    tracer.reset_counters()
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=33, want_accum=True,
                             flags=spt.FLAG_STREAMED | spt.FLAG_COUNTERS | single)
    c = tracer.counters()
    oacc, octr = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=pass_cap, seed=33)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))
    assert c["segments"] == octr["segments"] and c["samples"] == octr["samples"] and c["dropped"] == octr["dropped"]


def test_streamed_equals_resident_on_segments(tracer, cam, scene_random):
    w, h, spp, depth = 200, 120, 8, 8
    tracer.set_scene(scene_random)
    seg = (5, 101, 17, 190)
    _, a = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=2, seg=seg, want_accum=True)
    _, b = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=2, seg=seg, want_accum=True, flags=spt.FLAG_STREAMED)
    _, c = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=2, seg=seg, want_accum=True, flags=spt.FLAG_SINGLE_KERNEL)
    assert_accum_equal(a, b)
    assert_accum_equal(a, c)


# ----------------------------------------------------------------------------- progressive rendering (context-owned accumulator)
@pytest.mark.parametrize("variant", [spt.MEGAKERNEL, spt.WAVEFRONT])
def test_progressive_passes_equal_one_render(tracer, oracle, cam, scene_random, variant):
    """P passes over the samples, and per-tile calls in the reference scheduler's grid (Renderer.hpp:264-274), leave
    the bits of one render: the per-pixel running sums continue from the accumulator."""
    w, h, spp, depth = 96, 64, 12, 8
    tracer.set_scene(scene_random)
    _, one = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=6, want_accum=True, variant=variant)
    tracer.progressive_begin(w, h)
    for s0, s1 in [(0, 1), (1, 6), (6, 12)]:
        for seg in [(0, 32, 0, 48), (0, 32, 48, 96), (32, 64, 0, 48), (32, 64, 48, 96)]:
            tracer.progressive_add(cam, s0, s1, depth, pass_cap=10, seed=6, seg=seg, variant=variant)
        rgb, acc = tracer.progressive_image(want_accum=True)
        assert acc[..., 3].max() == s1
    assert_accum_equal(acc, one)
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=6)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))
    # resume: load a checkpoint of the first 6 samples into a fresh accumulator and finish
    tracer.progressive_begin(w, h)
    tracer.progressive_add(cam, 0, 6, depth, pass_cap=10, seed=6, variant=variant)
    _, half = tracer.progressive_image(want_accum=True)
    tracer.progressive_begin(w, h)
    tracer.progressive_load(half)
    tracer.progressive_add(cam, 6, 12, depth, pass_cap=10, seed=6, variant=variant)
    _, acc2 = tracer.progressive_image(want_accum=True)
    assert_accum_equal(acc2, one)


# ----------------------------------------------------------------------------- split rendering (primary pass + secondary bounce loop)
@pytest.mark.parametrize("which", ["init", "random", "metal_glass"])
@pytest.mark.parametrize("pass_cap", [0, 10, 1])
def test_split_render_bit_exact(tracer, oracle, cam, scene_init, scene_random, scene_metal_glass, which, pass_cap, monkeypatch):
    """SPT_FLAG_SPLIT: camera segments traced by primary_hits_kernel with the camera form of the pre-filter, bounce loop
    on secondary rays only.  Same bits as the oracle, same counters; several chunks of samples on purpose."""
    scene = {"init": scene_init, "random": scene_random, "metal_glass": scene_metal_glass}[which]
    monkeypatch.setattr(tracer, "split_chunk", 3)
    w, h, spp, depth = 100, 45, 7, 12
    tracer.set_scene(scene)
    tracer.reset_counters()
    rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=17, want_accum=True, flags=spt.FLAG_SPLIT | spt.FLAG_COUNTERS)
    c = tracer.counters()
    oacc, octr = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=pass_cap, seed=17)
    assert_accum_equal(acc, oacc)
    assert np.array_equal(rgb, oracle.resolve(oacc))
    assert c["segments"] == octr["segments"] and c["samples"] == octr["samples"] and c["dropped"] == octr["dropped"]


def test_split_render_segment_and_other_camera(tracer, oracle, scene_random):
    """An off-axis, far-away eye (the camera records are rebuilt per eye; epsT > 0 beyond |E| + Amax > 1024) and a segment."""
    tracer.set_scene(scene_random)
    for eye in ((3.0, 2.5, -6.0), (0.0, 400.0, -1500.0)):
        cam = oracle.camera(eye=eye, look=(0, 1, 5))
        w, h, spp, depth = 90, 50, 4, 8
        seg = (4, 41, 9, 77)
        _, acc = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=3, seg=seg, want_accum=True, flags=spt.FLAG_SPLIT)
        oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=3, seg=seg)
        assert_accum_equal(acc, oacc)


@pytest.mark.parametrize("feed", [0, spt.FLAG_STREAMED])
@pytest.mark.parametrize("pass_cap", [0, 10, 2])
def test_samples_shared_inside_the_warp_keep_the_bits(tracer, oracle, cam, scene_random, scene_metal_glass, feed, pass_cap, monkeypatch):
    """The bounce kernel's default shares the samples of a warp's 8x4 tile among its lanes (a finished sample is deposited into
    its staging record, every pixel's deposits are added in sample order afterwards); SPT_FLAG_OWN_PIXELS keeps one lane = one
    pixel.  Both leave the oracle's accumulators, drops and counts included, on ragged images (lanes without a pixel work for
    their neighbours), over several chunks, resident and ring feed; sharing makes fewer closest-hit trips per warp."""
    for scene, (w, h, spp, depth) in ((scene_random, (107, 61, 9, 12)), (scene_metal_glass, (50, 29, 6, 50))):
        tracer.set_scene(scene)
        oacc, octr = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=pass_cap, seed=23)
        trips = {}
        for chunk in (0, 4):
            monkeypatch.setattr(tracer, "split_chunk", chunk)
            for own in (0, spt.FLAG_OWN_PIXELS):
                tracer.reset_counters()
                rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=23, want_accum=True, flags=feed | own | spt.FLAG_COUNTERS)
                c = tracer.counters()
                assert_accum_equal(acc, oacc)
                assert np.array_equal(rgb, oracle.resolve(oacc))
                assert c["segments"] == octr["segments"] and c["dropped"] == octr["dropped"]
                trips[(chunk, own)] = c["warp_iterations"]
        if not feed:  # (the ring's loop is CTA-uniform: its trip count is the CTA's)
            assert trips[(0, 0)] < trips[(0, spt.FLAG_OWN_PIXELS)]


# ----------------------------------------------------------------------------- interleaved row bands (multi-GPU pixel partition)
@pytest.mark.parametrize("variant,flags", [(spt.MEGAKERNEL, 0), (spt.MEGAKERNEL, spt.FLAG_SINGLE_KERNEL), (spt.MEGAKERNEL, spt.FLAG_STREAMED),
                                           (spt.MEGAKERNEL, spt.FLAG_STRICT_LOOP), (spt.WAVEFRONT, 0)])
def test_row_bands_compose_to_the_full_render(tracer, oracle, cam, scene_random, variant, flags, monkeypatch):
    """SptParams.band_count / band_index: the calls of G 'ranks' touch disjoint 8-row bands and together leave the bits of
    one full render (what the multi-GPU partition relies on), on a ragged segment, for every kernel family."""
    import torch

    monkeypatch.setattr(tracer, "split_chunk", 3)  # several chunks of samples per call
    tracer.set_scene(scene_random)
    tracer.set_stream(torch.cuda.current_stream().cuda_stream)  # torch's fills and our kernels in one stream order
    w, h, spp, depth = 150, 93, 5, 8
    seg = (3, 90, 7, 141)  # 87 rows: 11 bands, the last one 7 rows
    full = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    tracer.accumulate(cam, w, h, 0, spp, depth, full, pass_cap=10, seed=12, seg=seg, variant=variant, flags=flags)
    for world in (2, 3, 16):  # 16 > 11 bands: some ranks own nothing
        total = torch.zeros_like(full)
        for rank in range(world):
            part = torch.zeros_like(full)
            tracer.accumulate(cam, w, h, 0, spp, depth, part, pass_cap=10, seed=12, seg=seg, variant=variant, flags=flags,
                              bands=(world, rank))
            torch.cuda.synchronize()
            rows = torch.nonzero(part[..., 3].sum(dim=1)).flatten().tolist()
            assert all(((y - seg[0]) // 8) % world == rank for y in rows)
            assert not torch.any((part[..., 3] > 0) & (total[..., 3] > 0))
            total += part
        assert torch.equal(total, full)
    torch.cuda.synchronize()
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=12, seg=seg)
    assert_accum_equal(full.cpu().numpy(), oacc)
    with pytest.raises(spt.SptError):
        tracer.accumulate(cam, w, h, 0, spp, depth, full, bands=(2, 2))
    tracer.set_stream(None)


# ----------------------------------------------------------------------------- staging limit: strips, chunks, auxiliary streams
def test_staging_limit_strips_and_chunks_keep_the_bits(tracer, oracle, cam, scene_random, monkeypatch):
    """spt_cuda_set_staging_limit / SptParams.strip_rows / split_chunk: a call that does not fit the budget is cut into strips
    of tile rows and, below that, into sample chunks; work items alternate between two staging buffers, camera kernels run on
    an auxiliary stream, the bounce kernels of neighbouring strips on two streams side by side.  Whatever the cut, the per-pixel
    running sums continue across the chunks of a strip and strips touch disjoint pixels: the accumulators are the oracle's."""
    w, h, spp, depth = 100, 45, 11, 12  # 6 tile rows of 8 (the last one 5 rows), 7 CTA columns
    tracer.set_scene(scene_random)
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=5)
    try:
        cases = [  # (staging limit, strip_rows, split_chunk, kernels launched = 2 per work item + resolve)
            (0, 0, 0, 2 * 1 + 1),        # everything fits: one strip, one chunk
            (0, 2, 0, 2 * 3 + 1),        # three strips, all samples each
            (0, 4, 3, 2 * 2 * 4 + 1),    # two strips (4 + 2 tile rows) x four chunks (3 + 3 + 3 + 2)
            (0, 1, 1, 2 * 6 * 11 + 1),   # one tile row, one sample per launch
            (4 * w * 8 * 16 * 6, 0, 0, None),  # a budget of four samples of the image: the library cuts by itself
            (1, 0, 0, None),             # absurdly small: one sample per launch, buffers as small as they can be
        ]
        for limit, strips, chunk, launches in cases:
            tracer.set_staging_limit(limit)
            monkeypatch.setattr(tracer, "strip_rows", strips)
            monkeypatch.setattr(tracer, "split_chunk", chunk)
            for flags in (0, spt.FLAG_SERIAL_CHUNKS):
                _, acc = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=5, want_accum=True, flags=flags)
                assert_accum_equal(acc, oacc)
                got = tracer.timing()["launches"]
                assert got == launches if launches else got > 3, (limit, strips, chunk, got)
        # a ragged segment, bands of another "rank" and strips together
        seg = (3, 43, 7, 91)
        monkeypatch.setattr(tracer, "strip_rows", 2)
        monkeypatch.setattr(tracer, "split_chunk", 4)
        tracer.set_staging_limit(0)
        import torch

        tracer.set_stream(torch.cuda.current_stream().cuda_stream)
        total = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
        for rank in range(2):
            tracer.accumulate(cam, w, h, 0, spp, depth, total, pass_cap=10, seed=5, seg=seg, bands=(2, rank))
        torch.cuda.synchronize()
        tracer.set_stream(None)
        oseg, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=5, seg=seg)
        assert_accum_equal(total.cpu().numpy(), oseg)
    finally:
        tracer.set_staging_limit(0)


# ----------------------------------------------------------------------------- several GPUs behind ONE context (spt_cuda_create_multi)
def _device_list(n):
    """n CUDA ordinals for a multi-device context: distinct GPUs when the box has them, otherwise the same GPU listed
    several times (several internal contexts on one device: every code path but the NVLink hop)."""
    import torch

    have = torch.cuda.device_count()
    return [k % have for k in range(n)]


@pytest.mark.parametrize("ndev", [2, 3])
@pytest.mark.parametrize("variant", [spt.MEGAKERNEL, spt.WAVEFRONT])
def test_multi_device_context_bands_leave_the_single_gpu_bits(tracer, oracle, cam, scene_random, ndev, variant):
    """One spt_cuda_render call on a multi-device context, band partition: the devices trace disjoint 8-row bands, the fused
    reduce-and-resolve kernel adds zeros, and both the accumulators and the RGB8 bytes equal the oracle's (ragged segment,
    bytes outside it untouched).  Renderer.hpp:257-302: one call spreads one render over all workers."""
    w, h, spp, depth = 150, 93, 5, 8
    seg = (3, 90, 7, 141)
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=12, seg=seg)
    with spt.PathTracer(_device_list(ndev)) as multi:
        assert multi.device_count() == ndev
        multi.set_scene(scene_random)
        multi.set_partition(spt.PARTITION_BANDS)
        multi.split_chunk = 2
        rgb0 = np.full((h, w, 3), 9, np.uint8)
        multi.reset_counters()
        rgb, acc = multi.render(cam, w, h, spp, depth, pass_cap=10, seed=12, seg=seg, want_accum=True, rgb8=rgb0.copy(), variant=variant,
                                flags=spt.FLAG_COUNTERS)
        assert_accum_equal(acc, oacc)
        want = rgb0.copy()
        y0, y1, x0, x1 = seg
        want[h - y1:h - y0, x0:x1] = oracle.resolve(oacc, seg=seg)[h - y1:h - y0, x0:x1]
        assert np.array_equal(rgb, want)
        c = multi.counters()
        assert c["samples"] == (y1 - y0) * (x1 - x0) * spp
        with pytest.raises(spt.SptError):  # the library partitions a multi-device context itself
            multi.progressive_begin(w, h) or multi.progressive_add(cam, 0, 1, depth, bands=(2, 0))


@pytest.mark.parametrize("ndev", [2, 4])
def test_multi_device_context_sample_ranges(tracer, oracle, cam, scene_random, ndev):
    """Sample-range partition (the default when a call has at least as many samples as devices): device k traces samples
    [k S / n, (k+1) S / n) of every pixel; the per-device sums are the oracle's for those ranges and the fused epilogue adds
    them in device order -- so the total is reproduced EXACTLY by adding the oracle's partial sums in the same order."""
    w, h, spp, depth = 96, 64, 10, 8
    parts = []
    for k in range(ndev):
        s0, s1 = spp * k // ndev, spp * (k + 1) // ndev
        o, _ = oracle.render_accum(scene_random, cam, w, h, s1 - s0, depth, pass_cap=10, seed=6, sample_begin=s0)
        parts.append(o)
    want = parts[0].copy()
    for o in parts[1:]:
        want = want + o  # float32 adds in device order, as reduce_resolve_kernel does
    with spt.PathTracer(_device_list(ndev)) as multi:
        multi.set_scene(scene_random)
        rgb, acc = multi.render(cam, w, h, spp, depth, pass_cap=10, seed=6, want_accum=True)
        assert_accum_equal(acc, want)
        assert np.array_equal(rgb, oracle.resolve(want))
        assert multi.timing()["launches"] == ndev * 3  # camera + bounce + epilogue slice on every device
        # against the single-device render: same samples, sums re-associated
        _, one = tracer_render_single(tracer, scene_random, cam, w, h, spp, depth)
        assert np.array_equal(acc[..., 3], one[..., 3])
        np.testing.assert_allclose(acc[..., :3], one[..., :3], rtol=2e-6, atol=0)
        # the opt-in grid on every device of the context: the same sums
        _, gacc = multi.render(cam, w, h, spp, depth, pass_cap=10, seed=6, want_accum=True, flags=spt.FLAG_GRID)
        assert_accum_equal(gacc, want)
        # fewer samples than devices: AUTO falls back to bands, which are bit-identical to one GPU
        _, few = multi.render(cam, w, h, 1, depth, pass_cap=10, seed=6, want_accum=True)
        _, few1 = tracer.render(cam, w, h, 1, depth, pass_cap=10, seed=6, want_accum=True)
        assert_accum_equal(few, few1)


def tracer_render_single(tracer, scene, cam, w, h, spp, depth):
    tracer.set_scene(scene)
    return tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=6, want_accum=True)


def test_multi_device_progressive(tracer, oracle, cam, scene_random):
    """The progressive entry points on a multi-device context: passes and tiles accumulate per device, every image is the
    fused reduction of the per-device sums, and with the band partition it is bit-identical to one render."""
    w, h, spp, depth = 96, 64, 6, 8
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=6)
    with spt.PathTracer(_device_list(2)) as multi:
        multi.set_scene(scene_random)
        multi.set_partition(spt.PARTITION_BANDS)
        multi.progressive_begin(w, h)
        for s0, s1 in [(0, 1), (1, 4), (4, 6)]:
            for seg in [(0, 32, 0, 48), (0, 32, 48, 96), (32, 64, 0, 48), (32, 64, 48, 96)]:
                multi.progressive_add(cam, s0, s1, depth, pass_cap=10, seed=6, seg=seg)
            rgb, acc = multi.progressive_image(want_accum=True)
            assert acc[..., 3].max() == s1
        assert_accum_equal(acc, oacc)
        assert np.array_equal(rgb, oracle.resolve(oacc))
        # resume from a checkpoint of the totals
        part, _ = oracle.render_accum(scene_random, cam, w, h, 4, depth, pass_cap=10, seed=6)
        multi.progressive_begin(w, h)
        multi.progressive_load(part)
        multi.progressive_add(cam, 4, 6, depth, pass_cap=10, seed=6)
        _, acc2 = multi.progressive_image(want_accum=True)
        assert_accum_equal(acc2, oacc)


# ----------------------------------------------------------------------------- shared-reciprocal division
def test_shared_reciprocal_division_equals_fdiv_rn(tracer):
    """Normalize and the camera ray divide by a shared, once-refined reciprocal (csrc/spt_device.cuh: the inline fast path of
    __fdiv_rn with its first three instructions hoisted).  On the device, 4e9 operand sets -- random mantissas, exponents
    inside, at the edges of and beyond the fast path's domain, zeros, denormals, infinities, NaNs -- give the bits of
    __fdiv_rn itself, for single quotients and for whole Normalize calls (Math.hpp:140-154)."""
    total = 0
    for seed in (1, 2, 3, 4):
        tested, bad = tracer.self_check_division(1 << 30, seed=seed)
        assert bad == 0, (seed, tested, bad)
        total += tested
    assert total >= 4 << 30


# ----------------------------------------------------------------------------- pixel-cone candidate lists of the camera kernel
@pytest.mark.parametrize("size", [(16, 900), (900, 16), (3, 3), (257, 131), (1, 40)])
@pytest.mark.parametrize("eye,look", [((0, 1, -3), (0, 1, 0)), ((0.3, 0.45, 8.0), (4, 0.3, 12)), ((40, 25, -60), (0, 0, 10)), ((0, 500, -2000), (0, 1, 5))])
def test_pixel_cone_lists_never_lose_a_hit(tracer, oracle, scene_random, size, eye, look):
    """The camera kernel tests every sphere ONCE per pixel and launch against the pixel's central ray with records fattened by
    the pixel's angular footprint, and runs only the exact sequence on that list for every sample.  The list must contain every
    sphere any jittered sample can hit: renders equal the oracle's bit for bit for extreme aspect ratios (quirk Q1 stretches the
    ray fan: px reaches 2 W / H - 1), tiny images (footprints of tens of degrees: lists overflow, the per-sample filter takes
    over), an eye inside the sphere field, and far / off-axis eyes."""
    w, h = size
    cam = oracle.camera(eye=eye, look=look)
    spp, depth = 6, 4
    tracer.set_scene(scene_random)
    _, acc = tracer.render(cam, w, h, spp, depth, pass_cap=10, seed=9, want_accum=True)
    oacc, _ = oracle.render_accum(scene_random, cam, w, h, spp, depth, pass_cap=10, seed=9)
    assert_accum_equal(acc, oacc)


def test_pixel_cone_lists_streamed_and_dense(tracer, oracle, cam):
    """The same on the streamed list (one pass of the blob through the TMA ring per launch) for a lattice whose far rows put
    more spheres into one pixel footprint than a list holds (those CTAs stream per sample), at two image sizes."""
    scene = spt.generate_lattice(7, 48)
    tracer.set_scene(scene)
    for (w, h) in ((120, 64), (40, 300)):
        _, acc = tracer.render(cam, w, h, 3, 6, pass_cap=10, seed=2, want_accum=True, flags=spt.FLAG_STREAMED)
        oacc, _ = oracle.render_accum(scene, cam, w, h, 3, 6, pass_cap=10, seed=2)
        assert_accum_equal(acc, oacc)


# ----------------------------------------------------------------------------- SPT_FLAG_GRID (opt-in acceleration structure)
def _probe_rays(rng, count, centre, spread, far):
    d = rng.normal(size=(count, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    o = (rng.uniform(-1, 1, size=(count, 3)) * np.array(spread) + np.array(centre)).astype(np.float32)
    # a share of axis-parallel directions (the walk's degenerate axes), of origins on the ground far outside the grid's
    # reach (they take the fattened-box test or the full list) aimed back at the scene, and of grazing low rays
    k = count // 8
    d[:k] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, k)] * rng.choice([-1, 1], size=(k, 1)).astype(np.float32)
    o[k:2 * k] = (rng.uniform(-1, 1, size=(k, 3)) * np.array([far, 0, far]) + np.array([0, 0.05, 0])).astype(np.float32)
    aim = (np.array(centre) + rng.uniform(-1, 1, size=(k, 3)) * np.array(spread) - o[k:2 * k]).astype(np.float32)
    d[k:2 * k] = aim / np.linalg.norm(aim, axis=1, keepdims=True).astype(np.float32)
    d[2 * k:3 * k, 1] *= np.float32(0.02)
    d[2 * k:3 * k] /= np.linalg.norm(d[2 * k:3 * k], axis=1, keepdims=True).astype(np.float32)
    return d, o


@pytest.mark.parametrize("which", ["init", "random", "metal_glass", "lattice"])
def test_grid_closest_hit_equals_the_linear_loop(tracer, oracle, scene_init, scene_random, scene_metal_glass, scene_lattice, which):
    """SPT_FLAG_GRID: the closest hit found by walking the uniform grid is FindClosestIntersectionSphere's for every ray --
    random, axis-parallel, grazing, starting inside spheres, starting thousands of units out on the ground -- on the device
    against the strict loop (cross-check) and against the oracle."""
    scene, centre, spread, far = {"init": (scene_init, (0, 1, 0), (3, 2, 3), 400.0), "random": (scene_random, (0, 2, 8), (25, 3, 14), 3000.0),
                                  "metal_glass": (scene_metal_glass, (0, 1, 12), (30, 2, 20), 3000.0),
                                  "lattice": (scene_lattice, (0, 1, 80), (85, 1.5, 85), 5000.0)}[which]
    tracer.set_scene(scene)
    rng = np.random.default_rng(77)
    d, o = _probe_rays(rng, 40000 if which != "lattice" else 24000, centre, spread, far)
    got, evals = tracer.closest_hit(d, o, flags=spt.FLAG_GRID | spt.FLAG_CROSS_CHECK, want_evals=True)
    assert not (got == 0xFFFFFFFF).any()
    assert np.array_equal(got, oracle.closest_hit_many(scene, d, o))
    if which == "lattice":  # the point of it: a few dozen exact evaluations per ray, not 16 384 (far origins aimed at the scene take them all)
        assert np.median(evals) < 64 and (evals >= 16383).mean() < 0.2


def test_grid_far_from_the_world_origin_and_degenerate_scenes(tracer, oracle, scene_random):
    """The grid's paddings grow with the coordinates (the float test is sloppier far from the world origin): the random scene
    moved 2 500 units away -- float spacing 2.4e-4 there, every discriminant a difference of numbers around 1e7 -- still gives
    the strict loop's and the oracle's answers; so do a scene of one small sphere and one of coincident spheres (ties)."""
    rng = np.random.default_rng(91)
    moved = spt.Scene.of(scene_random)
    moved = spt.Scene(moved.centers.copy(), moved.radii.copy(), moved.materials.copy(), moved.colors.copy(), moved.fuzz.copy())
    small = moved.radii < 64
    moved.centers[small, 0] += np.float32(1800.0)
    moved.centers[small, 2] += np.float32(1700.0)
    tracer.set_scene(moved)
    d, o = _probe_rays(rng, 40000, (1800, 2, 1708), (25, 3, 14), 6000.0)
    got = tracer.closest_hit(d, o, flags=spt.FLAG_GRID | spt.FLAG_CROSS_CHECK)
    assert not (got == 0xFFFFFFFF).any()
    assert np.array_equal(got, oracle.closest_hit_many(moved, d, o))
    assert (got < moved.n).mean() > 0.3
    # one filterable sphere; then eight coincident ones (the lowest index must win every tie)
    for count in (1, 8):
        centers = np.zeros((count + 1, 4), np.float32)
        centers[0] = (0, -1000.5, 0, 0)
        centers[1:] = (0.25, 1.0, 3.0, 0)
        radii = np.array([1000.0] + [0.5] * count, np.float32)
        sc = spt.Scene(centers, radii, np.full(count + 1, 3, np.uint8), np.full((count + 1, 4), 100, np.float32), np.zeros(count + 1, np.float32))
        tracer.set_scene(sc)
        d, o = _probe_rays(rng, 8000, (0.25, 1.0, 3.0), (3, 2, 3), 500.0)
        got = tracer.closest_hit(d, o, flags=spt.FLAG_GRID | spt.FLAG_CROSS_CHECK)
        assert not (got == 0xFFFFFFFF).any()
        assert np.array_equal(got, oracle.closest_hit_many(sc, d, o))
        assert set(np.unique(got)) <= {0, 1, count + 1}


@pytest.mark.parametrize("pass_cap", [0, 10])
def test_grid_render_keeps_the_bits(tracer, oracle, cam, scene_random, scene_metal_glass, scene_lattice, pass_cap, monkeypatch):
    """Renders with the bounce kernel on the grid: the oracle's accumulators and counters (segments and tests still count the
    reference's work), resident and ring-fed camera kernels, both sample policies, several chunks, another eye."""
    for scene, (w, h, spp, depth) in ((scene_random, (107, 61, 6, 12)), (scene_metal_glass, (50, 29, 4, 50)), (scene_lattice, (64, 40, 2, 16))):
        tracer.set_scene(scene)
        oacc, octr = oracle.render_accum(scene, cam, w, h, spp, depth, pass_cap=pass_cap, seed=41)
        for chunk, own in ((0, 0), (3, 0), (0, spt.FLAG_OWN_PIXELS)):
            monkeypatch.setattr(tracer, "split_chunk", chunk)
            tracer.reset_counters()
            rgb, acc = tracer.render(cam, w, h, spp, depth, pass_cap=pass_cap, seed=41, want_accum=True, flags=spt.FLAG_GRID | own | spt.FLAG_COUNTERS)
            c = tracer.counters()
            assert_accum_equal(acc, oacc)
            assert np.array_equal(rgb, oracle.resolve(oacc))
            assert c["segments"] == octr["segments"] and c["dropped"] == octr["dropped"]
    monkeypatch.setattr(tracer, "split_chunk", 0)
    tracer.set_scene(scene_random)
    for eye in ((3.0, 2.5, -6.0), (0.0, 400.0, -1500.0)):
        cam2 = oracle.camera(eye=eye, look=(0, 1, 5))
        _, acc = tracer.render(cam2, 90, 50, 4, 8, pass_cap=pass_cap, seed=3, want_accum=True, flags=spt.FLAG_GRID)
        oacc, _ = oracle.render_accum(scene_random, cam2, 90, 50, 4, 8, pass_cap=pass_cap, seed=3)
        assert_accum_equal(acc, oacc)
