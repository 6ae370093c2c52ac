"""N > 1 path on CPU: world-size-2 gloo process group exercising the sample-range partition and the one reduce
(raytrace_b200/multigpu.py).  The renderer itself needs a GPU, so a deterministic stand-in adds per-sample
contributions keyed by (pixel, GLOBAL sample index) — exactly the property the device RNG counter has — and the
test checks that the sharded sum equals the single-rank sum and that only the destination rank normalises."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytrace_b200.multigpu import reduce_image, render_sharded, split_samples


def test_split_samples_tiles_the_range():
    for spp in (1, 7, 50, 200, 1024, 1025):
        for world in (1, 2, 3, 4, 8):
            nxt = 0
            for r in range(world):
                first, count = split_samples(spp, world, r)
                assert first == nxt and count >= 0
                nxt += count
            assert nxt == spp
            counts = [split_samples(spp, world, r)[1] for r in range(world)]
            assert max(counts) - min(counts) <= 1
    with pytest.raises(ValueError):
        split_samples(10, 2, 2)


def _fake_range(n_px):
    def render_range(first, count, acc):
        px = torch.arange(n_px * 3, dtype=torch.float64)
        for s in range(first, first + count):                      # contribution depends on (pixel, global sample) only
            acc += (torch.sin(px * 0.37 + s * 1.7) ** 2).to(torch.float32)
    return render_range


def _worker(rank, world, port, spp, n_px, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    acc = torch.zeros(n_px * 3, dtype=torch.float32)
    render_sharded(_fake_range(n_px), acc, spp)
    np.save(os.path.join(out_dir, f"r{rank}.npy"), acc.numpy())
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("spp", [8, 13])
def test_sharded_sum_equals_single_rank(tmp_path, spp):
    n_px, world = 257, 2
    mp.spawn(_worker, args=(world, _free_port(), spp, n_px, str(tmp_path)), nprocs=world, join=True)
    single = torch.zeros(n_px * 3, dtype=torch.float32)
    render_sharded(_fake_range(n_px), single, spp)               # no process group: world = 1
    r0 = np.load(tmp_path / "r0.npy")
    r1 = np.load(tmp_path / "r1.npy")
    assert np.allclose(r0, single.numpy(), rtol=1e-5, atol=1e-6)
    # rank 1 keeps its own un-normalised partial sum (gloo reduce leaves non-root buffers unspecified but finite)
    assert np.isfinite(r1).all()


def test_reduce_image_without_group_just_normalises():
    a = torch.full((12,), 6.0)
    reduce_image(a, 3)
    assert torch.allclose(a, torch.full((12,), 2.0))


def _reducer_worker(rank, world, port, spp, n_px, out_dir):
    from raytrace_b200.multigpu import ImageReducer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    red = ImageReducer(n_px * 3, "cpu")               # no CUDA: the reducer must pick the dist.reduce path
    assert red.mode == "nccl" and red.handles is None
    for step in range(3):                             # reused across images like bench.py does
        acc = red.acc()
        acc.zero_()
        first, count = split_samples(spp, world, rank)
        _fake_range(n_px)(first + step, count, acc[: n_px * 3])
        img = red.reduce(1.0 / spp)
        assert (img is None) == (rank != 0)
        if rank == 0:
            np.save(os.path.join(out_dir, f"img{step}.npy"), img.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_image_reducer_gloo_world_size_2(tmp_path):
    """ImageReducer without CUDA / symmetric memory: dist.reduce + scale, image padded to a multiple of 4 floats
    (n_px * 3 = 771 is not one) and the padding never leaks into the result."""
    n_px, world, spp = 257, 2, 10
    mp.spawn(_reducer_worker, args=(world, _free_port(), spp, n_px, str(tmp_path)), nprocs=world, join=True)
    for step in range(3):
        single = torch.zeros(n_px * 3, dtype=torch.float32)
        _fake_range(n_px)(step, spp, single)
        got = np.load(tmp_path / f"img{step}.npy")
        assert got.shape == (n_px * 3,)
        assert np.allclose(got, single.numpy() / spp, rtol=1e-5, atol=1e-6)
