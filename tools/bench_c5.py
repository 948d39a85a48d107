"""bench.py --workload c5: BASELINE.json configs[4] -- 2^20 aircraft per GPU flying a looping square mission
(300 m side at 100 m altitude, 15 m/s, pure-pursuit guidance) through the full 5-level cascade
(waypoint -> HSA/TECS -> attitude -> rate -> surface), dt = dt_physics = 0.01 (K = 1: ONE RK4 substep per control step),
initial states jittered +-20 m / +-10 deg heading by global id.  One step = one launch of step_guided_staged_kernel<1>
(at K = 1 the one-aircraft code with its arrays staged through shared memory as 2-D TMA boxes; the pair kernel takes over from
K = 4; ACB200_GUIDED_STAGED=0 selects the direct kernel step_guided_kernel<float, 1>)."""
from __future__ import annotations

import json
import os
import time

METRIC = "aircraft control-steps/sec (waypoint->HSA->attitude->rate->surface cascade, 1 RK4 substep per step)"
UNIT = "env-steps/s"
DT = 0.01
SQUARE = [[300.0, 0.0, -100.0, 15.0], [300.0, 300.0, -100.0, 15.0], [0.0, 300.0, -100.0, 15.0], [0.0, 0.0, -100.0, 15.0]]
# (+ 10 warp-reduced mission statistics per block: 80 B per 128 aircraft, not counted)
# step_guided_staged_kernel<1> / step_guided_kernel<T,1>: loads state 13 + derived 4 + pid 18 + pid_outer 9 + wp_index 1, stores state 13 + controls 4 +
# derived 4 + pid 18 + pid_outer 9 + wp_index 1 (4-byte words, fp32 build)
BYTES_PER_STEP = (13 + 4 + 18 + 9 + 1 + 13 + 4 + 4 + 18 + 9 + 1) * 4


def workload_name(n):
    return (f"C5: {n} aircraft/GPU, looping square mission 300 m @ 100 m / 15 m/s, pure-pursuit guidance + HSA (TECS) + "
            f"attitude->rate->surface PID cascade, dt=0.01 K=1 RK4 substep")


def initial_states(n, id_offset, seed=7):
    import numpy as np
    ids = np.arange(id_offset, id_offset + n, dtype=np.uint64)

    def u(stream):
        with np.errstate(over="ignore"):
            x = ids + np.uint64(seed * 0x9E3779B97F4A7C15 & 0xFFFFFFFFFFFFFFFF) + np.uint64(stream) * np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(30); x *= np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(27); x *= np.uint64(0x94D049BB133111EB)
            x ^= x >> np.uint64(31)
        return (x >> np.uint64(11)).astype(np.float64) / float(1 << 53)

    # The fleet is spread over the whole circuit: arc position by global id, +-20 m across the leg, +-10 deg of heading
    # (BASELINE config 5's jitter), active waypoint = the end of the leg the aircraft is on.  In any window of the benchmark
    # some aircraft are inside an acceptance radius, so MissionPlanner's index-advance branch and the corner guidance run
    # (r01 started everyone at the first corner: nobody reached a waypoint in the timed 10 s).
    arc = 1200.0 * u(4)
    leg = np.minimum((arc // 300.0).astype(np.int64), 3)
    along, across = arc - 300.0 * leg, -20.0 + 40.0 * u(1)
    corner = np.array([[0.0, 0.0], [300.0, 0.0], [300.0, 300.0], [0.0, 300.0]])[leg]
    direc = np.array([[1.0, 0.0], [0.0, 1.0], [-1.0, 0.0], [0.0, -1.0]])[leg]
    s = np.zeros((13, n), np.float64)
    s[0] = corner[:, 0] + along * direc[:, 0] - across * direc[:, 1]
    s[1] = corner[:, 1] + along * direc[:, 1] + across * direc[:, 0]
    s[2] = -100.0
    s[3] = 15.0
    yaw = np.array([0.0, np.pi / 2, np.pi, -np.pi / 2])[leg] + np.radians(-10.0 + 20.0 * u(3))
    s[8] = np.arctan2(np.sin(yaw), np.cos(yaw))
    return s, leg.astype(np.int32)


def measure(args, rt, steps, warmup, e2e_steps, cpu_baseline, clocks=None):
    """Every rank runs it; returns the result object (bench.py prints rank 0's)."""
    import acb200
    torch, dist, dev = rt.torch, rt.dist, rt.dev
    rank, world, local = rt.rank, rt.world, rt.local
    barrier, max_over_ranks = rt.barrier, rt.max_over_ranks
    n = args.envs
    fleet = acb200.AircraftFleet(n, precision=32, device=local, config={"dt_physics": DT})
    s0, idx0 = initial_states(n, rank * n)
    fleet.reset(torch.as_tensor(s0))
    fleet.set_mission(SQUARE, acceptance_radius=40.0, guidance="PP", loop=True, stats=True)
    if os.environ.get("ACB200_C5_STATS", "1") == "0":        # A/B aid: the kernel without the statistics accumulation
        keep = fleet.mission_stats_buf
        fleet.set_mission(SQUARE, acceptance_radius=40.0, guidance="PP", loop=True, stats=False)
        fleet.mission_stats_buf = keep
    fleet.wp_index.copy_(torch.as_tensor(idx0))

    for _ in range(warmup):
        fleet.step_waypoint(DT)
    own_comm = acb200.sharding.init_comm(fleet.ctx)
    acb200.sharding.allreduce_episode_stats(fleet.mission_stats_buf, ctx=fleet.ctx)        # warm the communicator outside the timed region
    fleet.mission_stats_buf.zero_()
    barrier()
    if clocks is not None and rank == 0:
        clocks.start()
    l0 = fleet.ctx.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    cev, mstats = [], None
    for s_ in range(steps):
        fleet.step_waypoint(DT)
        # the one collective of the path: mission statistics summed over ranks once per report interval (100 steps) and at the
        # end of the run, inside the timed region, bracketed by its own events
        if (s_ + 1) % 100 == 0 or s_ + 1 == steps:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            mstats = acb200.sharding.allreduce_episode_stats(fleet.mission_stats_buf, ctx=fleet.ctx)
            b.record()
            cev.append((a, b))
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = fleet.ctx.launch_count - l0
    coll_us = [1e3 * a.elapsed_time(b) for a, b in cev]
    mvals = dict(zip(acb200._cabi.MISSION_STATS_NAMES, mstats.cpu().tolist()))
    value = n * world * steps / (ms * 1e-3)

    # end to end: the caller reads every aircraft's state back each step (AircraftInterface.step returns the state);
    # the mission is autonomous, so nothing travels host -> device per step
    state_h = torch.empty((13, n), dtype=torch.float32, pin_memory=True)
    derived_h = torch.empty((4, n), dtype=torch.float32, pin_memory=True)
    def step_host():
        fleet.step_waypoint(DT)
        state_h.copy_(fleet.state, non_blocking=True); derived_h.copy_(fleet.derived, non_blocking=True)
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
    wp = fleet.wp_index.cpu()
    alt = fleet.derived[1].float().cpu()
    peaks = rt.peaks
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    achieved = value / world * BYTES_PER_STEP / 1e9
    traffic = rt.traffic.get(f"step_waypoint_f32_{n}")
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(n), "envs_per_gpu": n, "substeps": 1,
                   "l2": f"working set {n * BYTES_PER_STEP / 2 / 1e6:.0f} MB per step > 126 MB L2 (inputs larger than L2)",
                   "sharding": f"dp{world}: aircraft sharded by global id, no data-path collective"},
        "roofline": {"bound": "hbm", "kernel": "step_guided_kernel<float,1>" if os.environ.get("ACB200_GUIDED_STAGED", "1")[:1] == "0" else "step_guided_staged_kernel<1>", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                     "frac": achieved / hbm_peak, "kernel_ms": ms / steps, "bytes_per_env_step": BYTES_PER_STEP,
                     "algorithmic_bytes_per_launch": BYTES_PER_STEP * n,
                     "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback", "traffic": traffic},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": (13 + 4) * 4 * n,
                "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                "note": "autonomous mission: no per-step host input; state + derived outputs read back every step"},
        "gpu_launches": int(launches),
        "clocks": clk,
        "collective": {"op": "all_reduce(sum) of the mission statistics, float64[16]" + ("" if world > 1 else " (one rank: a device copy, no NCCL call)"),
                       "bytes": 128, "calls_in_timed_region": len(cev), "us_mean": sum(coll_us) / len(coll_us), "us_max": max(coll_us),
                       "backend": ("nccl, issued by the library (acb200_episode_stats, C ABI)" if own_comm else "nccl (torch.distributed)") if world > 1 else None,
                       "interval_steps": 100},
        "mission": {"stats_all_ranks_timed_region": mvals,
                    "waypoints_reached_per_1000_aircraft_steps": 1e3 * mvals["waypoints_reached"] / max(mvals["aircraft_steps"], 1.0),
                    "cross_track_mean_m": mvals["cross_track_abs_sum"] / max(mvals["aircraft_steps"], 1.0),
                    "altitude_error_mean_m": mvals["altitude_error_abs_sum"] / max(mvals["aircraft_steps"], 1.0),
                    "airspeed_mean": mvals["airspeed_sum"] / max(mvals["aircraft_steps"], 1.0),
                    "altitude_mean_rank0": float(alt.mean()), "altitude_min_rank0": float(alt.min()),
                    "layout": "fleet spread over the circuit by global id (arc position, +-20 m across, +-10 deg heading)",
                    "sim_seconds": (warmup + steps + e2e_steps + 2) * DT},
    }
    if world == 1 and cpu_baseline and rank == 0:
        v, done, el, cores, sample = time_cpu(args.ref_envs, budget_s=args.cpu_seconds)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    del fleet
    torch.cuda.empty_cache()
    return out


def time_cpu(n, budget_s=None, steps=None, warmup=1):
    import numpy as np
    from oracle import oracle as orc
    cores = os.cpu_count() or 1
    P, CP = orc.default_aircraft_params(), orc.default_controller_params()
    state, idx = initial_states(n, 0)
    state = np.ascontiguousarray(state); idx = np.ascontiguousarray(idx)
    wp = np.asarray(SQUARE, np.float64)
    pid = np.zeros((18, n), np.float32); pid_outer = np.zeros((9, n), np.float32)

    def step():
        orc.step_waypoint(P, CP, state, wp, idx, pid, pid_outer, DT, guidance=orc.GUIDANCE_PP, acceptance_radius=40.0, loop=True,
                          dt_physics=DT, nthreads=cores)

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
    return n * done / el, done, el, cores, f"{n} aircraft x {done} steps ({el:.1f} s) of the same mission, C oracle with OpenMP"


def run_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    v, done, el, cores, sample = time_cpu(args.ref_envs, steps=args.steps, warmup=args.warmup)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * el / done, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(1 << 20), "sample_envs": args.ref_envs},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))
