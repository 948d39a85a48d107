"""Aggregate PCIe bandwidth of one box with N ranks copying at once (torchrun, one rank per GPU): device-to-host alone,
host-to-device alone and both together, pinned host memory, 80 MB per copy, CUDA events + barriers, max time over ranks.
Answers whether the host-buffer step of `bench.py --gpus 8` (88 GB/s of results for all eight ranks) is at the box's limit.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/probes/pcie_aggregate.py"""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if os.environ.get("ACB200_PROBE_BIND", "0") == "1":
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
    import acb200
    acb200.sharding.bind_host_to_gpu(local)

SIZE = 80 << 20
dev = torch.empty(SIZE, dtype=torch.uint8, device="cuda")
host = torch.empty(SIZE, dtype=torch.uint8, pin_memory=True)
dev2 = torch.empty(16 << 20, dtype=torch.uint8, device="cuda")
host2 = torch.empty(16 << 20, dtype=torch.uint8, pin_memory=True)
side = torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def timed(fn, reps=20):
    fn()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1) / reps * 1e-3], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def both():
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        dev2.copy_(host2, non_blocking=True)
    host.copy_(dev, non_blocking=True)
    torch.cuda.current_stream().wait_stream(side)


out = {"ranks": world, "bound": os.environ.get("ACB200_PROBE_BIND", "0")}
out["d2h_GBs_all_ranks"] = world * SIZE / timed(lambda: host.copy_(dev, non_blocking=True)) / 1e9
out["h2d_GBs_all_ranks"] = world * SIZE / timed(lambda: dev.copy_(host, non_blocking=True)) / 1e9
out["d2h_80MB_with_h2d_16MB_GBs_all_ranks"] = world * SIZE / timed(both) / 1e9
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
