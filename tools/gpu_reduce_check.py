"""Multi-GPU check of the image reduce (run under torchrun, one rank per GPU): rt_reduce_image through the NVSwitch
multicast mapping (multimem.ld_reduce) and through peer loads, against ncclReduce on the same random buffers, with the
device time of each (median of 20, max over ranks).
usage: python -m torch.distributed.run --nproc-per-node N tools/gpu_reduce_check.py [n_floats]"""
import os, sys, json, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from raytrace_b200.multigpu import ImageReducer

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 600 * 600 * 3
stream = torch.cuda.current_stream(dev)
res = {"world": world, "n_floats": n}
g = torch.Generator(device=dev).manual_seed(1234 + rank)
data = [torch.rand(n, generator=g, device=dev) for _ in range(2)]
want = []
for d in data:                                  # reference: ncclReduce of the same data
    t = d.clone()
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)
    want.append(t * 0.25)
for mode in ("nvls", "p2p", "nccl"):
    red = ImageReducer(n, dev, mode=mode)
    ok = True
    for step in range(4):                       # alternates the two symmetric buffers
        a = red.acc()
        a.zero_()
        a[:n].copy_(data[step % 2])
        img = red.reduce(0.25)
        if rank == 0:
            err = float((img - want[step % 2]).abs().max())
            ok = ok and err <= 1e-5
    ts = []
    for k in range(22):
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        red.reduce(0.25)
        e1.record(stream)
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    t = torch.tensor([statistics.median(ts[2:])], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res[mode] = {"mode_used": red.mode, "why": red.why, "ok": ok, "ms": float(t.item())}
if rank == 0:
    print(json.dumps(res), flush=True)
dist.barrier()
dist.destroy_process_group()
