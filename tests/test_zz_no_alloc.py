"""spt_cuda_closest_hit_static -- the closest hit through the library's file-scope static device scratch, for callers that
cannot call cudaMalloc (no counterpart in the reference).

Collected LAST on purpose (file name): a failure here hides nothing after it.  Passes
alone in a fresh process and after the other tests: it uses a context of its own and uploads its own scene.
"""
import numpy as np
import pytest

import simplepathtracer_b200 as spt

pytestmark = pytest.mark.gpu


def _rays(scene, n, seed):
    """Origins on sphere surfaces and in free space, arbitrary unit directions."""
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    pick = rng.integers(0, scene.n, n)
    surf = rng.normal(size=(n, 3)).astype(np.float32)
    surf /= np.linalg.norm(surf, axis=1, keepdims=True).astype(np.float32)
    o = (scene.centers[pick, :3].astype(np.float64) + surf * scene.radii[pick, None].astype(np.float64)).astype(np.float32)
    o[: n // 4] = rng.uniform(-20, 20, size=(n // 4, 3)).astype(np.float32)
    return d, o


def test_closest_hit_static_equals_closest_hit_on_two_calls_in_a_row():
    """Through the C ABI: the no-allocation entry point returns what spt_cuda_closest_hit returns -- sphere indices and exact
    evaluation counts are integers, so the tolerance is equality, as everywhere in this suite -- on two calls in a row; the
    second call is SHORTER than the first, so the shared array still holds the first call's rays and answers behind it, and
    a third call runs the strict loop.  Sizes beyond the stated maximum are refused with a message that names the limit, and
    the calls change nothing about the device's free memory."""
    import torch

    scene = spt.generate_spheres(1)  # the random-spheres scene of the benchmark configurations (host-side generator)
    with spt.PathTracer(0) as pt:
        pt.set_scene(scene)
        d1, o1 = _rays(scene, spt.STATIC_SCRATCH_RAYS, 101)  # exactly the maximum
        d2, o2 = _rays(scene, 20_001, 202)
        want1, wev1 = pt.closest_hit(d1, o1, want_evals=True)
        want2, wev2 = pt.closest_hit(d2, o2, want_evals=True)
        assert (want1 < scene.n).mean() > 0.3 and (want2 < scene.n).mean() > 0.3  # the rays do hit things

        got1, gev1 = pt.closest_hit_static(d1, o1, want_evals=True)
        free_after_first = torch.cuda.mem_get_info(0)[0]
        got2, gev2 = pt.closest_hit_static(d2, o2, want_evals=True)
        got3 = pt.closest_hit_static(d2, o2, flags=spt.FLAG_STRICT_LOOP)
        free_after_third = torch.cuda.mem_get_info(0)[0]
        assert np.array_equal(got1, want1) and np.array_equal(gev1, wev1)
        assert np.array_equal(got2, want2) and np.array_equal(gev2, wev2)
        assert np.array_equal(got3, want2)
        assert free_after_third == free_after_first  # nothing was allocated or freed on the device

        # beyond the maximum: a clear error, and the context stays usable
        d4, o4 = _rays(scene, spt.STATIC_SCRATCH_RAYS + 1, 303)
        with pytest.raises(spt.SptError, match="SPT_STATIC_SCRATCH_RAYS"):
            pt.closest_hit_static(d4, o4)
        assert np.array_equal(pt.closest_hit_static(d2, o2), want2)
        assert pt.closest_hit_static(d2[:0], o2[:0]).shape == (0,)
