"""The reference's OWN Python env timed as a baseline (north_star: "next to the reference's Python env, SubprocVecEnv across all
host cores, core count stated").  TEST / MEASUREMENT INFRASTRUCTURE: imports the unmodified reference from /root/reference
(with oracle/_ref's compiled PID module and gymnasium stand-in), so it runs in the build container only -- the GPU box has
no reference checkout; bench.py's reference arm there times the C port (a far stronger baseline).

stable-baselines3 is not installed, so SubprocVecEnv is reproduced by its worker protocol: one process per core, one Pipe
round trip (action -> obs, reward, done, info) per env step, auto-reset in the worker.  Two variants, both reported:
  piped    -- the trainer-facing rate: the parent sends every worker an action and waits for all results, step by step;
  free_run -- each worker steps its env alone for the whole interval (no pipes): the ceiling of the reference on these cores.
Actions: the reference's own PID composition (utils/pid_demonstrations.py:55-78) or zeros.

    make -C oracle ref && PYTHONPATH=oracle/_ref:/root/reference python tools/ref_python_env_baseline.py [--seconds 20]
"""
import argparse
import json
import logging
import multiprocessing as mp
import os
import platform
import time

import numpy as np


def _make(seed, difficulty):
    logging.disable(logging.CRITICAL)          # the reference logs every ground contact (simplified_6dof.py:278-282)
    from learned_controllers.envs.rate_env import RateControlEnv
    env = RateControlEnv(difficulty=difficulty, command_type="step", rng_seed=seed)
    env.reset(seed=seed)
    return env


def _pid_action(env, agent):
    from controllers.types import ControlCommand, ControlMode
    st = env.sim.get_state()
    p, q, r = env.rate_command
    sf = agent.compute_action(ControlCommand(mode=ControlMode.RATE, roll_rate=p, pitch_rate=q, yaw_rate=r, throttle=0.6), st, dt=0.02)
    return np.array([sf.aileron, sf.elevator, sf.rudder, sf.throttle], dtype=np.float32)


def _worker_piped(conn, seed, difficulty):
    env = _make(seed, difficulty)
    while True:
        cmd, a = conn.recv()
        if cmd == "close":
            break
        obs, rew, term, trunc, info = env.step(a)
        if term or trunc:                          # SB3's worker: keep the terminal observation, reset, return the reset one
            info["terminal_observation"] = obs
            obs, _ = env.reset()
        conn.send((obs, rew, term or trunc, info))


def _worker_free(seed, difficulty, seconds, policy, out):
    env = _make(seed, difficulty)
    agent = None
    if policy == "pid":
        from controllers.rate_agent import RateAgent
        from controllers.types import ControllerConfig
        agent = RateAgent(ControllerConfig())
    zero = np.array([0, 0, 0, 0.6], np.float32)
    for _ in range(20):
        env.step(zero)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        a = _pid_action(env, agent) if agent is not None else zero
        _, _, term, trunc, _ = env.step(a)
        n += 1
        if term or trunc:
            env.reset()
            if agent is not None:
                agent.reset()
    out.put((n, time.perf_counter() - t0))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=20.0)
    ap.add_argument("--workers", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--difficulty", default="medium")
    args = ap.parse_args()
    w = args.workers
    res = {"what": "reference RateControlEnv (unmodified Python, numpy physics + its C++ PID), dt=0.02 = 20 RK4 substeps per env step",
           "host": platform.processor() or platform.machine(), "cores": os.cpu_count(), "workers": w, "seconds_per_variant": args.seconds,
           "numpy": np.__version__}
    # ---- piped (SubprocVecEnv protocol), zero action from the parent
    pipes, procs = [], []
    for k in range(w):
        a, b = mp.Pipe()
        p = mp.Process(target=_worker_piped, args=(b, 100 + k, args.difficulty), daemon=True)
        p.start(); pipes.append(a); procs.append(p)
    zero = np.array([0, 0, 0, 0.6], np.float32)
    def vec_step():
        for c in pipes:
            c.send(("step", zero))
        return [c.recv() for c in pipes]
    for _ in range(20):
        vec_step()
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < args.seconds:
        vec_step(); n += 1
    el = time.perf_counter() - t0
    for c in pipes:
        c.send(("close", None))
    for p in procs:
        p.join(timeout=5)
    res["piped_zero_action"] = {"env_steps_per_s": w * n / el, "vec_steps": n, "seconds": el}
    # ---- free-running workers
    for policy in ("zero", "pid"):
        q = mp.Queue()
        procs = [mp.Process(target=_worker_free, args=(200 + k, args.difficulty, args.seconds, policy, q), daemon=True) for k in range(w)]
        for p in procs:
            p.start()
        got = [q.get() for _ in procs]
        for p in procs:
            p.join(timeout=5)
        res[f"free_run_{policy}_action"] = {"env_steps_per_s": sum(n_ / t for n_, t in got), "per_worker": sum(n_ / t for n_, t in got) / w}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
