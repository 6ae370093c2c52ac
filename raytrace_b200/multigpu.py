"""Sample-sharded rendering across GPUs: one process per GPU, scene replicated read-only.

Every (pixel, sample) path is independent (Ray.hs:228-232), and the device RNG is a Philox counter keyed by
(seed; pixel, GLOBAL sample index, bounce), so rank r of G renders global sample indices
[first_r, first_r + count_r) of every pixel into its own fp32 sum buffer and ONE ``reduce(SUM)`` to rank 0 over
NCCL/NVLink assembles the image; the union of samples — hence the image up to fp32 summation order — does not
depend on G.  No other collective exists on this path.

``torch`` / ``torch.distributed`` are plumbing only (device buffers, the process group); rendering goes through
librt_b200.so (``CompiledScene.render_device``).
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple


def split_samples(spp: int, world: int, rank: int) -> Tuple[int, int]:
    """(first, count) of rank's global sample range; a remainder is spread over the first ranks.
    The ranges of ranks 0..world-1 tile [0, spp) exactly, in rank order."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank / world size")
    base, rem = divmod(int(spp), world)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def reduce_image(acc, spp: int, group=None, dst: int = 0):
    """Sum the per-rank un-normalised radiance buffers on ``dst`` and divide by spp there (Ray.hs:232).
    ``acc`` is a torch tensor (cuda → NCCL, cpu → gloo); returns it (meaningful on ``dst`` only)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM, group=group)
        if dist.get_rank(group) != dst:
            return acc
    acc.mul_(1.0 / float(spp))
    return acc


def render_sharded(render_range: Callable[[int, int, "object"], None], acc, spp: int, group=None, dst: int = 0):
    """Generic driver: ``render_range(first, count, acc)`` must ADD the un-normalised sums of global samples
    [first, first + count) into ``acc``; the result lands on rank ``dst`` divided by spp."""
    import torch.distributed as dist
    world = rank = None
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    first, count = split_samples(spp, world, rank)
    acc.zero_()
    if count:
        render_range(first, count, acc)
    return reduce_image(acc, spp, group=group, dst=dst)


def raytrace_distributed(settings, world_geometry, seed, compiled=None, device: Optional[int] = None, group=None):
    """``raytrace`` across the ranks of the current process group.  Returns the H×W×3 float32 image as a numpy array on
    rank 0 and None elsewhere.  Call from every rank (torchrun, one rank per GPU)."""
    import torch
    from .ray import CompiledScene, imageHeight
    if device is None:
        device = torch.cuda.current_device()
    cs = compiled if compiled is not None else CompiledScene(world_geometry, device=device)
    h, w = imageHeight(settings), settings.cs_imageWidth
    dev = torch.device("cuda", device)
    acc = torch.zeros(h * w * 3, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev)

    def render_range(first, count, buf):
        cs.render_device(settings, seed, first, count, buf.data_ptr(), 0, stream.cuda_stream)

    render_sharded(render_range, acc, settings.cs_samplesPerPixel, group=group)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_rank(group) != 0:
        return None
    return acc.view(h, w, 3).cpu().numpy()
