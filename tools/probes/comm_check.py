"""torchrun --nproc-per-node N tools/probes/comm_check.py: the C-ABI statistics all-reduce over N GPUs of one box.
Every rank steps its shard, the library's own communicator (acb200_comm_init, id carried by torch.distributed) sums the
statistics (acb200_episode_stats); checked against torch.distributed's all_reduce of the same vectors and timed."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import acb200  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 65536
env = acb200.BatchedRateControlEnv(n, controller="rate_pid", precision=32, device=local, seed=5, id_offset=rank * n, episode_length=0.4)
env.reset()
for _ in range(40):
    env.step()
assert acb200.sharding.init_comm(env.ctx) is True
info = acb200.sharding.comm_info(env.ctx)
ours = acb200.sharding.allreduce_episode_stats(env.stats, ctx=env.ctx)
ref = env.stats.clone(); dist.all_reduce(ref)
torch.cuda.synchronize()
ok = bool(torch.allclose(ours, ref, rtol=1e-13, atol=0)) and float(ours[15]) == world * n * 40
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
for a, b in ev:
    a.record(); acb200.sharding.allreduce_episode_stats(env.stats, ctx=env.ctx); b.record()
torch.cuda.synchronize()
us = sorted(1e3 * a.elapsed_time(b) for a, b in ev)
tv = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(200)]
for a, b in tv:
    a.record(); acb200.sharding.allreduce_episode_stats(env.stats); b.record()
torch.cuda.synchronize()
ust = sorted(1e3 * a.elapsed_time(b) for a, b in tv)
flags = torch.tensor([int(ok)], device="cuda"); dist.all_reduce(flags, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"world": world, "all_ranks_ok": bool(flags.item()), "comm": info, "aircraft_steps_all_ranks": float(ours[15]),
                      "c_abi_allreduce_us": {"median": us[100], "p90": us[180]}, "torch_distributed_allreduce_us": {"median": ust[100], "p90": ust[180]}}))
env.ctx.check(env.ctx.lib.acb200_comm_destroy(env.ctx.handle))
dist.destroy_process_group()
sys.exit(0 if ok else 1)
