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


class ImageReducer:
    """The image reduce of the sample-sharded path over NVLink peer memory (``rt_reduce_image``).

    Every rank accumulates into a symmetric-memory buffer (``torch.distributed._symmetric_memory``: plumbing that maps
    each rank's buffer into all peers and, on NVSwitch systems, creates the multicast mapping); ``reduce`` is a
    device-side barrier across the ranks followed by ONE kernel on the root that reads the image through the multicast
    address (``multimem.ld_reduce``: the switch adds the G copies in the fabric) - or through plain peer loads when the
    fabric has no multicast - and applies 1/spp.  Buffers are double-buffered so that a rank may start zeroing and
    filling the next image while the root still reads the previous one (the next barrier orders the reuse).

    ``mode``: "nvls" (multicast; falls back to "p2p" when unsupported), "p2p", or "nccl" (``dist.reduce`` + scale, the
    SURVEY 8e baseline; also what CPU/gloo runs use).  ``RT_B200_REDUCE`` overrides."""

    def __init__(self, n_floats: int, device, group=None, dst: int = 0, mode: Optional[str] = None):
        import os
        import torch
        import torch.distributed as dist
        self.n = int(n_floats)
        self.n_pad = (self.n + 3) // 4 * 4
        self.dst = dst
        self.group = group
        self.device = torch.device(device)
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        mode = os.environ.get("RT_B200_REDUCE", mode or "nvls")
        self.mode = "nccl"
        self.why = ""
        self.k = 0
        self.handles = None
        if mode in ("nvls", "p2p") and self.world > 1 and self.device.type == "cuda":
            try:
                import torch.distributed._symmetric_memory as symm
                g = group if group is not None else dist.group.WORLD
                self.bufs = [symm.empty(self.n_pad, dtype=torch.float32, device=self.device) for _ in range(2)]
                self.handles = [symm.rendezvous(b, g.group_name) for b in self.bufs]
                mc = all(int(h.multicast_ptr) != 0 for h in self.handles)
                self.mode = "nvls" if (mode == "nvls" and mc) else "p2p"
                if mode == "nvls" and not mc:
                    self.why = "no multicast support on this fabric: peer loads"
                self.out = torch.empty(self.n_pad, dtype=torch.float32, device=self.device)
            except Exception as e:      # symmetric memory unavailable: NCCL reduce (still on the GPUs)
                self.handles = None
                self.why = f"symmetric memory unavailable ({type(e).__name__}: {e}): ncclReduce"
        if self.handles is None:
            self.bufs = [torch.zeros(self.n_pad, dtype=torch.float32, device=self.device)]

    def acc(self):
        """This rank's accumulation buffer for the next image (zero it, render into it, then call ``reduce``)."""
        return self.bufs[self.k]

    def reduce(self, scale: float):
        """Stream-ordered on the current stream.  Returns the normalised image (n floats) on ``dst``, None elsewhere."""
        import torch
        import torch.distributed as dist
        buf = self.bufs[self.k]
        if self.handles is None:
            if self.world > 1:
                dist.reduce(buf, dst=self.dst, op=dist.ReduceOp.SUM, group=self.group)
            if self.rank != self.dst:
                return None
            buf.mul_(scale)
            return buf[: self.n]
        import ctypes as C
        from . import _ffi
        h = self.handles[self.k]
        self.k ^= 1
        h.barrier(channel=0)                     # every rank's render has completed and is visible
        if self.rank != self.dst:
            return None
        stream = torch.cuda.current_stream(self.device).cuda_stream
        if self.mode == "nvls":
            _ffi.check(_ffi.lib().rt_reduce_image(self.device.index, C.c_void_p(int(h.multicast_ptr)), None, 0,
                                                  C.c_void_p(self.out.data_ptr()), self.n_pad, float(scale), C.c_void_p(stream)))
        else:
            ptrs = (C.c_void_p * self.world)(*[int(p) for p in h.buffer_ptrs])
            _ffi.check(_ffi.lib().rt_reduce_image(self.device.index, None, ptrs, self.world,
                                                  C.c_void_p(self.out.data_ptr()), self.n_pad, float(scale), C.c_void_p(stream)))
        return self.out[: self.n]


_REDUCERS: dict = {}


def cached_reducer(n_floats: int, device, group=None) -> ImageReducer:
    """One ImageReducer per (size, device, group): its symmetric-memory buffers are a collective allocation, and a rank
    must not free them while the root may still be reading them through the multicast / peer mapping - so they live as
    long as the process (what bench.py does by hand)."""
    import torch
    key = (int(n_floats), str(torch.device(device)), id(group))
    if key not in _REDUCERS:
        _REDUCERS[key] = ImageReducer(n_floats, device, group=group)
    return _REDUCERS[key]


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
    stream = torch.cuda.current_stream(dev)
    import torch.distributed as dist
    spp = settings.cs_samplesPerPixel
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    red = cached_reducer(h * w * 3, dev, group=group)
    acc = red.acc()
    acc.zero_()
    first, count = split_samples(spp, world, rank)
    if count:
        cs.render_device(settings, seed, first, count, acc.data_ptr(), 0, stream.cuda_stream, want_stats=False)
    img = red.reduce(1.0 / float(spp))
    if img is None:
        return None
    return img.view(h, w, 3).cpu().numpy()
