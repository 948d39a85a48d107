import sys, torch, numpy as np
sys.path.insert(0, ".")
import acb200, bench
n = 1 << 20
env = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=32, seed=42, command_type="step")
difficulty, att, phase = bench.make_inputs(n, 0)
env.set_curriculum(torch.as_tensor(difficulty), "step"); env.att_cmd.copy_(torch.as_tensor(att).float()); env.reset()
pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
att_host = pin((4, n), torch.float32); att_host.copy_(torch.as_tensor(att).float())
obs_h, rew_h = pin((n, 18), torch.float32), pin((n,), torch.float32)
term_h, trunc_h = pin((n,), torch.uint8), pin((n,), torch.uint8)
for chunks in (0, 8, 0, 8, 0, 8):
    env.ctx.check(env.ctx.lib.acb200_set_host_chunks(env.ctx.handle, chunks))
    for _ in range(3):
        env.step_host(None, obs_h, rew_h, term_h, trunc_h, att_cmd_host=att_host)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(30):
        env.step_host(None, obs_h, rew_h, term_h, trunc_h, att_cmd_host=att_host)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 30
    print(f"chunks={chunks}: {ms:.3f} ms/step, {n/ms*1e3:.3e} env-steps/s, D2H {81.79/ms:.1f} GB/s", flush=True)
