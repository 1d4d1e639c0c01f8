"""Sharding one render over the ranks of a single node.  One process per GPU.

Every (pixel, sample) is independent under the counter-based RNG, and the scene (<= 672 KB) is replicated,
so the path shards with no exchange inside the tracing kernel.  Two partitions are offered:

* ``sample_range``  -- rank r traces samples [r*spp/G, (r+1)*spp/G) of EVERY pixel.  Balanced by construction (sky
  and sphere-dense image regions cost differently) and every rank launches the full grid of CTAs; the per-rank float4
  accumulators (sum r, sum g, sum b, count) are then combined by ONE sum-reduce to rank 0
  (``reduce_accumulators``: NCCL over NVLink on the GPUs, gloo in the CPU tests).  The default.
* ``row_bands``     -- rank r owns the 8-row bands b with b % G == r (interleaved) and traces ALL samples of its pixels
  in ONE call (``SptParams.band_count`` / ``band_index``).  The regions are disjoint, so the same sum-reduce of
  zero-initialised accumulators is a gather with no arithmetic and the result is bit-identical to the single-GPU
  render.  Used when spp < G; measured on 8 B200s it is slower than sample ranges at 1024 spp (a rank's grid shrinks to
  1/G of the CTAs, 6.9 instead of 55 waves per launch, and the tail of each launch weighs accordingly: 40.5 against
  51.4 Gsamples/s).

The reference's analogue is the static threadCount x threadCount tile grid of RenderImageParallelMain
(/root/reference/include/Renderer.hpp:264-274).  This module is host-side plumbing: it never computes a pixel.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

BAND_ROWS = 8  # one CTA tile row of the tracing kernels (kTileH, csrc/spt_kernels.cuh)


def sample_range(rank: int, world: int, spp: int) -> Tuple[int, int]:
    """[begin, end) of the samples rank `rank` of `world` traces; the ranges tile [0, spp) in rank order."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    if spp < 0:
        raise ValueError("spp must be >= 0")
    return rank * spp // world, (rank + 1) * spp // world


def row_bands(rank: int, world: int, height: int, width: int, band_rows: int = BAND_ROWS) -> List[Tuple[int, int, int, int]]:
    """Segments (yBegin, yEnd, xBegin, xEnd) of the interleaved row bands owned by `rank`."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    segs = []
    nbands = (height + band_rows - 1) // band_rows
    for b in range(rank, nbands, world):
        segs.append((b * band_rows, min((b + 1) * band_rows, height), 0, width))
    return segs


def band_params(rank: int, world: int) -> Tuple[int, int]:
    """(band_count, band_index) of ``SptParams`` for rank `rank` of `world`."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return world, rank


def choose_partition(world: int, spp: int) -> str:
    """Sample ranges whenever every rank gets at least one sample, otherwise interleaved row bands."""
    return "samples" if spp >= world else "bands"


def reduce_accumulators(accum, dst: int = 0, group=None) -> None:
    """Sum the per-rank float4 accumulators into rank `dst`'s tensor, in place (async on the current stream
    for NCCL).  A no-op for a single process."""
    import torch.distributed as dist

    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)


def accumulate_sharded(tracer, cam, width: int, height: int, spp: int, depth: int, accum, *, rank: int, world: int,
                       pass_cap: int = 0, seed: int = 1, variant: int = 0, flags: int = 0, partition: Optional[str] = None) -> int:
    """This rank's share of one render into its (zeroed) device accumulator.  Returns the kernels launched."""
    partition = partition or choose_partition(world, spp)
    if partition == "samples":
        s0, s1 = sample_range(rank, world, spp)
        if s1 > s0:
            tracer.accumulate(cam, width, height, s0, s1, depth, accum, pass_cap=pass_cap, seed=seed, variant=variant, flags=flags)
            return 1
        return 0
    if partition == "bands":
        # one call: the library's grid covers only this rank's bands (SptParams.band_count / band_index)
        if rank * BAND_ROWS >= height or spp <= 0:
            return 0
        tracer.accumulate(cam, width, height, 0, spp, depth, accum, pass_cap=pass_cap, seed=seed, variant=variant, flags=flags,
                          bands=band_params(rank, world))
        return 1
    raise ValueError(f"unknown partition {partition!r}")
