"""Host-side sharding logic (simplepathtracer_b200/sharding.py) on CPU: partitions, and the N > 1 path with
two gloo processes.  The oracle stands in for the GPU inside the worker (it is the checker of the test: a
fake tracer object with the PathTracer.accumulate signature), so what is verified is the partition + reduce
plumbing: the reduced accumulator equals the sum of the per-rank accumulators, the band partition reproduces
the single-process render bit for bit, and sample ranges reproduce it up to float re-association."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from simplepathtracer_b200 import sharding  # noqa: E402

W, H, SPP, DEPTH, SEED = 48, 40, 6, 4, 9


def test_sample_ranges_tile_the_samples():
    for world in (1, 2, 3, 4, 8):
        for spp in (0, 1, 5, 8, 1024, 1000):
            r = [sharding.sample_range(k, world, spp) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.sample_range(2, 2, 8)


def test_row_bands_partition_the_image():
    for world in (1, 2, 4, 8):
        for height in (1, 8, 37, 360):
            seen = np.zeros(height, np.int32)
            for k in range(world):
                for y0, y1, x0, x1 in sharding.row_bands(k, world, height, 64):
                    assert (x0, x1) == (0, 64) and 0 <= y0 < y1 <= height
                    seen[y0:y1] += 1
            assert np.all(seen == 1)


def test_choose_partition():
    assert sharding.choose_partition(8, 1024) == "samples"
    assert sharding.choose_partition(8, 4) == "bands"
    assert sharding.band_params(3, 8) == (8, 3)
    with pytest.raises(ValueError):
        sharding.band_params(8, 8)


class OracleTracer:
    """PathTracer.accumulate's signature over the CPU oracle (test double)."""

    def __init__(self, scene):
        from oracle import Oracle

        self.o, self.scene = Oracle(), scene

    def accumulate(self, cam, width, height, s0, s1, depth, accum, pass_cap=0, seed=1, seg=None, variant=0, flags=0, bands=None):
        import torch

        # SptParams.band_count / band_index (include/spt_cuda.h): the 8-row bands b of the segment with b % count == index
        y0, y1, x0, x1 = seg if seg is not None else (0, height, 0, width)
        count, index = bands if bands and bands[0] > 1 else (1, 0)
        for b in range(index, (y1 - y0 + sharding.BAND_ROWS - 1) // sharding.BAND_ROWS, count):
            band = (y0 + b * sharding.BAND_ROWS, min(y0 + (b + 1) * sharding.BAND_ROWS, y1), x0, x1)
            part, _ = self.o.render_accum(self.scene, cam, width, height, s1 - s0, depth, pass_cap=pass_cap, seed=seed, seg=band,
                                          sample_begin=s0)
            accum += torch.from_numpy(np.ascontiguousarray(part).reshape(height, width, 4))


def _worker(rank, world, port, partition, out_dir):
    import torch
    import torch.distributed as dist

    from oracle import Oracle

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        o = Oracle()
        scene, cam = o.init_spheres(1), o.camera()
        accum = torch.zeros((H, W, 4), dtype=torch.float32)
        n = sharding.accumulate_sharded(OracleTracer(scene), cam, W, H, SPP, DEPTH, accum, rank=rank, world=world, pass_cap=10,
                                        seed=SEED, partition=partition)
        assert n >= 1
        np.save(os.path.join(out_dir, f"part{rank}.npy"), accum.numpy())
        sharding.reduce_accumulators(accum, dst=0)
        if rank == 0:
            np.save(os.path.join(out_dir, "reduced.npy"), accum.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("partition", ["samples", "bands"])
def test_two_ranks_gloo(tmp_path, partition, oracle):
    import torch.multiprocessing as mp

    world = 2
    mp.spawn(_worker, args=(world, _free_port(), partition, str(tmp_path)), nprocs=world, join=True)
    parts = [np.load(tmp_path / f"part{r}.npy") for r in range(world)]
    reduced = np.load(tmp_path / "reduced.npy")
    assert np.array_equal(reduced, parts[0] + parts[1])  # the collective is a plain sum in rank order

    scene, cam = oracle.init_spheres(1), oracle.camera()
    full, _ = oracle.render_accum(scene, cam, W, H, SPP, DEPTH, pass_cap=10, seed=SEED)
    full = np.asarray(full).reshape(H, W, 4)
    assert np.array_equal(reduced[..., 3], full[..., 3])  # sample counts are small integers: exact either way
    if partition == "bands":
        # disjoint regions: the reduce adds zeros, so the sharded render IS the single-process render
        assert np.array_equal(reduced, full)
        assert np.count_nonzero(parts[0][..., 3]) and np.count_nonzero(parts[1][..., 3])
        assert not np.any((parts[0][..., 3] > 0) & (parts[1][..., 3] > 0))
    else:
        # (a0+a1+a2) + (a3+a4+a5) vs ((((a0+a1)+a2)+a3)+a4)+a5: equal up to float re-association
        np.testing.assert_allclose(reduced, full, rtol=1e-6, atol=1e-3)
        a, b = oracle.resolve(reduced).astype(int), oracle.resolve(full).astype(int)
        assert np.max(np.abs(a - b)) <= 1  # a rounding boundary may be crossed by the last ulp of the sum
