"""bench.py --workload c4: BASELINE.json configs[3] -- 262 144 aircraft per GPU flown by the learned LSTM rate policy
(random init, torch seed 0), medium difficulty, dt 0.02 (20 RK4 substeps).  One step = policy forward
(policy_forward_kernel, tcgen05) + env step (env_step_pair_kernel, external action) for every aircraft, issued as ONE
acb200_policy_env_rollout call for all the timed steps (two halves of the fleet on their own streams)."""
from __future__ import annotations

import json
import os
import time

METRIC = "aircraft env-steps/sec (learned LSTM policy forward + env step, 20 RK4 substeps per env step)"
UNIT = "env-steps/s"


def workload_name(n):
    return (f"C4: {n} aircraft/GPU, LSTMFeaturesExtractor + stateful actor LSTM + pi-MLP policy (random init, seed 0) "
            f"-> RateControlEnv (external action, medium, step commands), dt=0.02 K=20 RK4 substeps")


def measure(args, rt, steps, warmup, e2e_steps, cpu_baseline, clocks=None):
    """Every rank runs it; returns the result object (bench.py prints rank 0's)."""
    import acb200
    from acb200.policy import random_policy_state_dict
    torch, dist, dev = rt.torch, rt.dist, rt.dev
    rank, world, local = rt.rank, rt.world, rt.local
    barrier, max_over_ranks = rt.barrier, rt.max_over_ranks
    n = args.envs
    env = acb200.BatchedRateControlEnv(n, difficulty="medium", controller="external", precision=32, device=local, seed=42,
                                       id_offset=rank * n, command_type="step")
    pol = acb200.LearnedPolicyB200(random_policy_state_dict("lstm", 0), kind="lstm", n=n, device=local, ctx=env.ctx)
    env.reset()
    flags = (env.terminated, env.truncated)

    def step():
        env.step(pol(env.obs, flags))

    for _ in range(warmup):
        step()
    barrier()
    if clocks is not None and rank == 0:
        clocks.start()
    l0 = env.ctx.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    ev0.record()
    for a, b in pev:                       # events bracket the policy kernel on its own (the launching) stream
        a.record()
        act = pol(env.obs, flags)
        b.record()
        env.step(act)
    ev1.record()
    barrier()
    seq_ms = max_over_ranks(ev0.elapsed_time(ev1))
    policy_ms = sum(a.elapsed_time(b) for a, b in pev) / steps
    # ---- the same closed loop as ONE call (acb200_policy_env_rollout): two halves of the fleet on their own streams, one
    # half's env step beside the other half's policy forward.  This is the C4 step that `value` reports; the sequence above
    # stays for the per-kernel times (the policy kernel alone, on the launching stream) behind the roofline.
    parts = int(os.environ.get("ACB200_C4_PARTS", "2"))
    pol.rollout(env, 4, parts=parts)
    barrier()
    l0 = env.ctx.launch_count
    ev0.record()
    pol.rollout(env, steps, parts=parts)
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = env.ctx.launch_count - l0
    # and replayed from a CUDA graph (one graph launch per 20 steps)
    graph_ms = None
    try:
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                pol.rollout(env, 20, parts=parts)
        barrier()
        reps = max(1, steps // 20)
        ev0.record()
        for _ in range(reps):
            g.replay()
        ev1.record()
        barrier()
        graph_ms = max_over_ranks(ev0.elapsed_time(ev1)) / (20 * reps)
    except Exception as exc:                      # a failed capture must not cost the measurement
        graph_ms = f"capture failed: {exc}"
    stats = acb200.sharding.allreduce_episode_stats(env.stats)
    value = n * world * steps / (ms * 1e-3)

    # ---- end to end: the observation comes from host memory every step, results go back to host memory
    pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
    obs_h, rew_h = pin((n, 18), torch.float32), pin((n,), torch.float32)
    term_h, trunc_h = pin((n,), torch.uint8), pin((n,), torch.uint8)
    obs_h.copy_(env.obs); term_h.copy_(env.terminated); trunc_h.copy_(env.truncated)
    obs_d = torch.empty_like(env.obs)
    from acb200.policy import step_host_pipelined
    e2e_chunks = int(os.environ.get("ACB200_C4_E2E_CHUNKS", "4"))

    def step_host():
        if e2e_chunks > 1:      # the public host-buffer step: H2D / policy + env / D2H pipelined over parts of the fleet
            step_host_pipelined(env, pol, obs_h, rew_h, term_h, trunc_h, obs_d, chunks=e2e_chunks)
            return
        obs_d.copy_(obs_h, non_blocking=True)
        env.step(pol(obs_d, flags))
        obs_h.copy_(env.obs, non_blocking=True); rew_h.copy_(env.reward_f32 if hasattr(env, "reward_f32") else env.reward, non_blocking=True)
        term_h.copy_(env.terminated, non_blocking=True); trunc_h.copy_(env.truncated, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(2):
        step_host()
    barrier()
    ev0.record()
    for _ in range(e2e_steps):
        step_host()
    ev1.record()
    barrier()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1))
    e2e_value = n * world * e2e_steps / (e2e_ms * 1e-3)
    clk = clocks.stop() if (clocks is not None and rank == 0) else None
    h2d = obs_h.numel() * 4
    d2h = sum(t.numel() * t.element_size() for t in (obs_h, rew_h, term_h, trunc_h))
    peaks = rt.peaks
    info = pol.info()
    flops = info["flops_per_aircraft"] * n
    achieved = flops / (policy_ms * 1e-3) / 1e12
    peak = peaks.get("bf16_tflops_sustained", 1401.4)     # a kernel timed inside a long step: the sustained figure
    traffic = rt.traffic.get(f"policy_forward_{n}")
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 operands / f32 accumulate and gates (policy), f32 (env)", "data": "synthetic",
        "config": {"workload": workload_name(n), "envs_per_gpu": n, "policy": "lstm", "substeps": 20,
                   "l2": "per-step working set (recurrent state 805 MB + env state) > 126 MB L2 (inputs larger than L2)",
                   "sharding": f"dp{world}: aircraft sharded by global id, no data-path collective"},
        "roofline": {"bound": "tensor", "kernel": "policy_forward_kernel", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "kernel_ms": policy_ms,
                     "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if "bf16_tflops_sustained" in peaks else "fallback",
                     "flop_per_aircraft": info["flops_per_aircraft"], "algorithmic_flop_per_launch": flops,
                     "weight_bytes_per_tile": info["weight_bytes"], "traffic": traffic, "traffic_source": rt.traffic.get("_source")},
        "kernels_ms": {"policy_forward_kernel": policy_ms, "env_step_kernel": seq_ms / steps - policy_ms,
                       "sequential_step": seq_ms / steps, "rollout_call_step": ms / steps, "rollout_graph_step": graph_ms,
                       "rollout_parts": parts},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "ms_per_step": e2e_ms / e2e_steps},
        "gpu_launches": int(launches),
        "clocks": clk,
        "episode_stats": dict(zip(acb200.envs.STATS_NAMES, stats.cpu().tolist())),
    }
    if world == 1 and cpu_baseline and rank == 0:
        v, done, el, cores, sample = time_cpu(args.ref_envs_c4, budget_s=args.cpu_seconds)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    del env, pol
    torch.cuda.empty_cache()
    return out


def time_cpu(n, budget_s=None, steps=None, warmup=1):
    """CPU arm: torch fp32 restatement of the policy (all cores, torch's own threading) + the C oracle env with an
    external action (OpenMP over aircraft)."""
    import numpy as np
    import torch
    from oracle import oracle as orc
    from oracle import policy_oracle as po
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    net = po.make_policy("lstm", seed=0)
    env = orc.Env(n, controller=orc.CTRL_EXTERNAL, seed=42, difficulty=np.full(n, 0.5, np.float32), command_type=orc.CMD_STEP,
                  nthreads=cores)
    env.reset()
    state = net.initial_state(n)
    done_flags = torch.zeros(n, dtype=torch.bool)

    def step():
        nonlocal state, done_flags
        act, state = net(torch.from_numpy(np.ascontiguousarray(env.obs, dtype=np.float32)), state, done_flags)
        env.step(act.numpy())
        done_flags = torch.from_numpy((np.asarray(env.terminated) | np.asarray(env.truncated)).astype(bool))

    for _ in range(warmup):
        step()
    t0 = time.perf_counter(); done = 0
    while True:
        step(); done += 1
        if steps is not None and done >= steps:
            break
        if budget_s is not None and done >= 3 and time.perf_counter() - t0 >= budget_s:
            break
    el = time.perf_counter() - t0
    sample = (f"{n} aircraft x {done} steps ({el:.1f} s): torch fp32 policy restatement ({cores} threads) + C oracle env "
              f"(OpenMP), same workload")
    return n * done / el, done, el, cores, sample


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    v, done, el, cores, sample = time_cpu(args.ref_envs_c4, steps=args.steps, warmup=args.warmup)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (policy), f64 (env)", "data": "synthetic",
        "config": {"workload": workload_name(262144), "sample_envs": args.ref_envs_c4},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))
