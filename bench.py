#!/usr/bin/env python
"""bench.py -- env-steps/s of the batched flight-control path on B200 (BASELINE.json metric).

Workload (BASELINE.json configs[2], "C3"): 2^20 aircraft per GPU, each a RateControlEnv (dt 0.02 =
20 RK4 substeps of 1 ms) flown by the full cascaded attitude -> rate -> surface PID, "step" rate
commands, per-aircraft difficulty drawn from {easy, medium, hard}, in-kernel Philox curriculum
resets.  One "step" = one control step of every aircraft = ONE launch of env_step_pair_kernel (fp32: two aircraft per thread on packed FP32
instructions; env_step_kernel<T>, one aircraft per thread, for the fp64 parity build and ACB200_ENV_PAIR=0).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
    python bench.py --impl reference [...]                          # the CPU arm (oracle port, all cores)

For N > 1 launch under torchrun (one rank per GPU); aircraft are sharded by global id with no
data-path collective (weak scaling), the only NCCL traffic is the episode-statistics all-reduce.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "aircraft env-steps/sec (full cascaded PID, 20 RK4 substeps per env step)"
UNIT = "env-steps/s"
K_SUB = 20
# algorithmic work per env step, A(K) = 1014 K + 320 FLOP; HBM bytes at the control-step boundary
FLOP_PER_ENV_STEP = 1014 * K_SUB + 320
FP32_PEAK_DERIVED_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12     # 74.4: 148 SM x 128 lanes x FMA x max clock


def workload_name(n):
    return (f"C3: {n} aircraft/GPU, attitude->rate->surface PID cascade + RateControlEnv reward/termination/"
            f"Philox curriculum reset (easy/medium/hard mix), dt=0.02 K=20 RK4 substeps")


def env_bytes_per_step(precision):
    """Algorithmic HBM bytes per aircraft per env step of env_step_kernel (attitude-PID controller): every buffer
    the kernel loads (kernels.cuh phases A and C) and stores (scatter + outputs), once each."""
    r = precision // 8
    rd = (13 + 4 + 3 + 7 + 7 + 1 + 1) * r + 18 * 4 + 4 * 4 + 4 + 4 + 1 + 1 + 4      # state, att_cmd, rate_cmd, rew_mem, sched, env_time, ep_return | pid, prev_action, counters, flags, difficulty
    wr = (13 + 4 + 4 + 3 + 7 + 1 + 1 + 1) * r + 18 * 4 + 4 * 4 + 4 + 4 + 1 + 72 + 1 + 1   # state, derived, controls, rate_cmd, rew_mem, env_time, ep_return, reward | pid, prev_action, counters, settled, obs, done flags
    return rd + wr


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1])); pw.append(float(p[2]))
            except ValueError:
                continue
            for nm, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def make_inputs(n, id_offset, seed=42):
    """Synthetic C3 inputs by GLOBAL aircraft id, so results do not depend on the sharding."""
    import numpy as np
    ids = np.arange(id_offset, id_offset + n, dtype=np.uint64)
    # splitmix-style hash -> three uniform streams (host side, cheap, shard-independent)
    def u(stream):
        x = (ids + np.uint64(seed * 0x9E3779B97F4A7C15 & 0xFFFFFFFFFFFFFFFF) + np.uint64(stream) * np.uint64(0xBF58476D1CE4E5B9))
        x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
        return (x >> np.uint64(11)).astype(np.float64) / float(1 << 53)
    with np.errstate(over="ignore"):
        difficulty = np.array([0.3, 0.5, 0.7], np.float32)[np.minimum((u(1) * 3).astype(np.int64), 2)]
        att = np.zeros((4, n), np.float64)
        att[0] = np.radians(-20 + 40 * u(2)); att[1] = np.radians(-20 + 40 * u(3))
        att[2] = np.inf                                   # hold current yaw (what HSAAgent commands, hsa_agent.py:214-216)
        att[3] = 0.6
        phase = (u(4) * 500).astype(np.int32)             # de-synchronise the episodes: resets are spread over steps
    return difficulty, att, phase


def oracle_env(n, nthreads, id_offset=0, dt=0.02):
    import numpy as np
    from oracle import oracle as orc
    difficulty, att, phase = make_inputs(n, id_offset)
    env = orc.Env(n, controller=orc.CTRL_ATTITUDE_PID, seed=42, id_offset=id_offset, difficulty=difficulty,
                  command_type=orc.CMD_STEP, nthreads=nthreads, dt=dt)
    env.att_cmd[:] = att
    env.reset()
    env.step_count[:] = phase
    return env


def time_oracle(n, nthreads, warmup, steps=None, budget_s=None, dt=0.02):
    env = oracle_env(n, nthreads, dt=dt)
    for _ in range(warmup):
        env.step()
    t0 = time.perf_counter()
    done = 0
    while True:
        env.step(); done += 1
        if steps is not None and done >= steps:
            break
        if budget_s is not None and done >= 3 and time.perf_counter() - t0 >= budget_s:
            break
    el = time.perf_counter() - t0
    return n * done / el, done, el


def run_reference(args):
    """CPU arm: the reference algorithm (oracle port, C, OpenMP over aircraft) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n = args.ref_envs
    k1 = args.workload == "k1"
    value, done, el = time_oracle(n, cores, args.warmup, steps=args.steps, dt=0.001 if k1 else 0.02)
    sample = f"{n} aircraft x {done} env steps of the same C3 workload (bounded sample), OpenMP over aircraft"
    print(json.dumps({
        "impl": "reference", "metric": METRIC.replace("20 RK4 substeps", "1 RK4 substep") if k1 else METRIC, "value": value,
        "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(1 << 20) + (" [at K=1: dt=0.001]" if k1 else ""), "sample_envs": n},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


class Runtime:
    """One process per GPU: rank / device / process group, the barrier and the max-over-ranks the contract asks for."""

    def __init__(self):
        import torch
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0")); self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                self.peaks = json.load(f)
        except Exception:
            pass
        self.traffic = {}
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                self.traffic = json.load(f)
        except Exception:
            pass

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.dist is None:
            return ms
        t = self.torch.tensor([ms], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def measure_c3(args, rt, k_sub, steps, warmup, e2e_steps, cpu_baseline, clocks=None):
    """C3 (k_sub = 20) or its K = 1 end (k_sub = 1) on every rank; returns the result object (meaningful on rank 0)."""
    import numpy as np
    import acb200
    torch, dist, dev = rt.torch, rt.dist, rt.dev
    rank, world, local = rt.rank, rt.world, rt.local
    barrier, max_over_ranks = rt.barrier, rt.max_over_ranks
    n = args.envs
    precision = args.precision
    flop_per_env_step = 1014 * k_sub + 320
    env = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=precision, device=local, seed=42,
                                       id_offset=rank * n, command_type="step", dt=0.001 * k_sub)
    difficulty, att, phase = make_inputs(n, rank * n)
    env.set_curriculum(torch.as_tensor(difficulty), "step")
    env.att_cmd.copy_(torch.as_tensor(att).to(env.dtype))
    env.reset()
    env.step_count.copy_(torch.as_tensor(phase))
    torch.cuda.synchronize()

    # ---- device-resident arm: `value` -------------------------------------------------------------
    for _ in range(warmup):
        env.step()
    own_comm = acb200.sharding.init_comm(env.ctx)                    # the library's own NCCL communicator (C ABI) when world > 1
    stats = acb200.sharding.allreduce_episode_stats(env.stats, ctx=env.ctx)       # warm it outside the timed region
    barrier()
    if clocks is not None and rank == 0:
        clocks.start()
    l0 = env.ctx.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cev = []
    barrier()
    ev0.record()
    for s_ in range(steps):
        env.step()
        # the one collective of the path -- episode statistics summed over ranks, once per report interval (100 steps)
        # and at the end of the run -- INSIDE the timed region, bracketed by its own events
        if (s_ + 1) % 100 == 0 or s_ + 1 == steps:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            stats = acb200.sharding.allreduce_episode_stats(env.stats, ctx=env.ctx)
            b.record()
            cev.append((a, b))
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = env.ctx.launch_count - l0
    coll_us = [1e3 * a.elapsed_time(b) for a, b in cev]
    collective = {"op": "all_reduce(sum) of the episode statistics, float64[16]" + ("" if world > 1 else " (one rank: a device copy, no NCCL call)"),
                  "bytes": 128, "calls_in_timed_region": len(cev), "us_mean": sum(coll_us) / len(coll_us), "us_max": max(coll_us),
                  "backend": ("nccl, issued by the library (acb200_episode_stats, C ABI)" if own_comm else "nccl (torch.distributed)") if world > 1 else None,
                  "interval_steps": 100}
    value = n * world * steps / (ms * 1e-3)

    # ---- end-to-end arm: host buffers, H2D of the step's commands + D2H of obs/reward/done every step
    numa_cores = acb200.sharding.bind_host_to_gpu(local) if world > 1 else []   # staging buffers on the GPU's own NUMA node
    pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
    att_host = pin((4, n), env.dtype); att_host.copy_(torch.as_tensor(att).to(env.dtype))
    obs_h, rew_h = pin((n, 18), torch.float32), pin((n,), torch.float32)
    term_h, trunc_h = pin((n,), torch.uint8), pin((n,), torch.uint8)
    for _ in range(2):
        env.step_host(None, obs_h, rew_h, term_h, trunc_h, att_cmd_host=att_host)
    barrier()
    ev0.record()
    for _ in range(e2e_steps):
        env.step_host(None, obs_h, rew_h, term_h, trunc_h, att_cmd_host=att_host)
    ev1.record()
    barrier()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1))
    e2e_value = n * world * e2e_steps / (e2e_ms * 1e-3)
    clk = clocks.stop() if (clocks is not None and rank == 0) else None           # sampled over both timed regions
    h2d = att_host.numel() * att_host.element_size()
    d2h = sum(t.numel() * t.element_size() for t in (obs_h, rew_h, term_h, trunc_h))

    # ---- the same call with fewer bytes: the controller runs in the kernel, so a trainer needs reward + flags (6 B per
    # aircraft) and the observations of the aircraft that FINISHED (acb200_env_gather_done), not the (n, 18) matrix.
    # Reported beside the full-copy number, which stays the headline (it is the reference-shaped call).
    done_rows = 0
    for _ in range(2):
        env.step_host(None, None, rew_h, term_h, trunc_h, att_cmd_host=att_host); env.gather_done()
    barrier()
    ev0.record()
    for _ in range(e2e_steps):
        env.step_host(None, None, rew_h, term_h, trunc_h, att_cmd_host=att_host)
        done_rows += len(env.gather_done()[0])
    ev1.record()
    barrier()
    red_ms = max_over_ranks(ev0.elapsed_time(ev1))
    e2e_reduced = {"value": n * world * e2e_steps / (red_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": 6 * n + int(done_rows / e2e_steps * (8 + 2 * 72)) + 8, "steps": e2e_steps,
                   "ms_per_step": red_ms / e2e_steps,
                   "note": "obs output NULL: reward + flags for every aircraft, ids + terminal/reset observations of the finished ones only"}

    peaks = rt.peaks
    per_gpu = value / world
    bytes_step = env_bytes_per_step(precision)
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    # non-tensor FP32 peak: measured on this device by the library's FMA-chain kernel (MEASURED_PEAKS.json has no
    # non-tensor compute figure); fp64 denominator: half of it (B200: 64 FP64 lanes per SM against 128 FP32)
    torch.cuda.synchronize()
    fp32_measured = env.ctx.measure_fp32_peak()
    fp_peak = fp32_measured if precision == 32 else fp32_measured / 2
    traffic = None
    # fp32 fleets of even size run the two-aircraft-per-thread kernel on packed FP32 (csrc/kernels_pair_f32.cuh)
    pair = precision == 32 and n % 2 == 0 and os.environ.get("ACB200_ENV_PAIR", "1")[:1] != "0"
    kernel_name = "env_step_pair_kernel" if pair else "env_step_kernel"
    traffic = rt.traffic.get(f"env_step{'_pair' if pair else ''}{'_k1' if k_sub == 1 else ''}_f{precision}_{n}")
    achieved_tflops = per_gpu * flop_per_env_step / 1e12
    achieved_gbs = per_gpu * bytes_step / 1e9
    roof_fp = {"bound": "fp32", "note": "non-tensor FP32 issue-bound; the path is neither HBM- nor tensor-bound (DESIGN.md 4); HBM view in roofline_hbm",
               "achieved": achieved_tflops, "peak": fp_peak, "unit": "TFLOP/s", "frac": achieved_tflops / fp_peak,
               "peak_source": "measured here: FMA-chain kernel, all SMs (derived nominal: %.1f)" % FP32_PEAK_DERIVED_TFLOPS,
               "flop_per_env_step": flop_per_env_step, "kernel": kernel_name, "kernel_ms": ms / steps,
               "traffic": traffic, "traffic_source": rt.traffic.get("_source")}
    roof_hbm = {"bound": "hbm", "achieved": achieved_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": achieved_gbs / hbm_peak,
                "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback",
                "bytes_per_env_step": bytes_step, "algorithmic_bytes_per_launch": bytes_step * n, "kernel": kernel_name,
                "kernel_ms": ms / steps, "traffic": traffic, "traffic_source": rt.traffic.get("_source")}
    k1 = k_sub == 1
    out = {
        "metric": METRIC.replace("20 RK4 substeps", "1 RK4 substep") if k1 else METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if precision == 32 else "f64", "data": "synthetic",
        "config": {"workload": workload_name(n).replace("C3:", "C3 at K=1 (HBM-bound end of the path):").replace("dt=0.02 K=20 RK4 substeps", "dt=0.001 K=1 RK4 substep") if k1 else workload_name(n),
                   "envs_per_gpu": n, "controller": "attitude_pid", "substeps": k_sub,
                   "l2": f"working set {n * bytes_step / 2 / 1e6:.0f} MB per step > 126 MB L2 (inputs larger than L2)",
                   "sharding": f"dp{world}: aircraft sharded by global id, no data-path collective"},
        "roofline": roof_hbm if k1 else roof_fp,
        ("roofline_fp32" if k1 else "roofline_hbm"): roof_fp if k1 else roof_hbm,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "ms_per_step": e2e_ms / e2e_steps, "host_cores_bound": len(numa_cores)},
        "e2e_reduced": e2e_reduced,
        "collective": collective,
        "gpu_launches": int(launches),
        "clocks": clk,
        "episode_stats": dict(zip(acb200.envs.STATS_NAMES, stats.cpu().tolist())),
    }
    if world == 1 and cpu_baseline and rank == 0:
        cores = os.cpu_count() or 1
        v, done, el = time_oracle(args.ref_envs, cores, 1, budget_s=args.cpu_seconds, dt=0.001 * k_sub)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                               "sample": f"{args.ref_envs} aircraft x {done} env steps ({el:.1f} s) of the same workload, C oracle with OpenMP"}
    del env
    torch.cuda.empty_cache()
    return out


# what a sub-workload contributes to the headline line's "workloads" object
_BRIEF = ("metric", "value", "unit", "steps", "warmup", "ms_per_step", "dtype", "config", "roofline", "e2e", "gpu_launches",
          "kernels_ms", "mission", "collective")


def run_gpu(args):
    """The workload --workload names on every rank; with the default (c3, the BASELINE metric) the other three are
    measured right after it in the same process and reported in "workloads", so that one driver run of
    `bench.py --gpus N` puts k1, C4 and C5 under the same clock (about a second each)."""
    from tools import bench_c4, bench_c5
    rt = Runtime()
    clocks = ClockSampler(rt.local)
    e2e_steps = max(3, min(args.steps, args.e2e_steps))
    cpu = not args.no_cpu_baseline
    if args.workload in ("c3", "k1"):
        out = measure_c3(args, rt, 1 if args.workload == "k1" else K_SUB, args.steps, args.warmup, e2e_steps, cpu, clocks)
    elif args.workload == "c4":
        out = bench_c4.measure(args, rt, args.steps, args.warmup, e2e_steps, cpu, clocks)
    else:
        out = bench_c5.measure(args, rt, args.steps, args.warmup, e2e_steps, cpu, clocks)
    if args.workload == "c3" and not args.no_extra_workloads:
        import copy
        extra = {}
        sub = copy.copy(args)
        sub.envs = args.envs
        extra["k1"] = measure_c3(sub, rt, 1, 300, 5, 20, False)
        sub = copy.copy(args); sub.envs = 262144 if args.envs == 1 << 20 else max(256, args.envs // 4 // 256 * 256)
        extra["c4"] = bench_c4.measure(sub, rt, 100, 5, 20, False)
        sub = copy.copy(args)
        extra["c5"] = bench_c5.measure(sub, rt, 300, 5, 20, False)
        out["workloads"] = {k: {f: v[f] for f in _BRIEF if f in v} for k, v in extra.items()}
    if rt.rank == 0:
        print(json.dumps(out))
    rt.close()


def main():
    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version banner under torchrun)
    # goes to stderr instead
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=1 << 20, help="aircraft per GPU")
    ap.add_argument("--precision", type=int, default=32, choices=[32, 64])
    ap.add_argument("--ref-envs", type=int, default=16384, help="aircraft in the bounded CPU sample")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--e2e-steps", type=int, default=50)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra-workloads", action="store_true", help="c3 only: skip the k1 / c4 / c5 lines of the \"workloads\" object")
    ap.add_argument("--workload", default="c3", choices=["c3", "c4", "c5", "k1"],
                    help="c3 (default, BASELINE metric): 2^20 aircraft, cascaded PID, K=20; c4: 262144 aircraft, learned LSTM policy; "
                         "k1: c3 with one RK4 substep per env step (the HBM-bound K=1 path); c5: 2^20 aircraft, waypoint mission, 5-level cascade, K=1")
    ap.add_argument("--ref-envs-c4", type=int, default=4096, help="aircraft in the bounded CPU sample of --workload c4")
    args = ap.parse_args()
    if args.workload == "c5":
        from tools import bench_c5
        args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
        if args.steps == 1000 and args.impl == "reference":
            args.steps = 20
        return bench_c5.run_reference(args) if args.impl == "reference" else run_gpu(args)
    if args.workload == "c4":
        from tools import bench_c4
        if args.envs == 1 << 20:
            args.envs = 262144
        args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
        if args.steps == 1000:
            args.steps = 200 if args.impl == "b200" else 5
        return bench_c4.run_reference(args) if args.impl == "reference" else run_gpu(args)
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
