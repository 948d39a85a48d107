"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

    make -C oracle ref
    PYTHONPATH=oracle/_ref:/root/reference python tests/golden/make_golden.py

The reference (numpy physics, Python agents/env, its own C++ PID compiled into oracle/_ref) is imported
from /root/reference; nothing is copied.  Each fixture stores the inputs (initial states, command /
action sequences) next to the reference's outputs, so the tests need neither the reference nor this
script at run time.  Conventions: SoA [field][n]; state rows = N,E,D,u,v,w,phi,theta,psi,p,q,r,time;
controls rows = elevator, aileron, rudder, throttle (ControlSurfaces.to_array order);
env actions = aileron, elevator, rudder, throttle (rate_env.py:228-233).
"""
import logging
import multiprocessing as mp
import os
import sys

import numpy as np

logging.disable(logging.CRITICAL)
HERE = os.path.dirname(os.path.abspath(__file__))
# The fixtures pin the reference AS RUN UNDER NumPy >= 2 (NEP 50 weak scalars): with float32 actions the env's smoothness
# term and the residual bonus are float32 arithmetic (rate_env.py:260-265,282; residual_rate_env.py:153-155).  The reference's
# requirements.txt pins numpy<2, where the same expressions promote to float64 -- a difference of <= ~1e-8 in the reward,
# inside the 1e-7 reward tolerance of the tests but outside bit parity.  DESIGN.md section 3 states this assumption.
assert int(np.__version__.split(".")[0]) >= 2, f"fixtures are defined under NumPy >= 2, found {np.__version__}"

import aircraft_controls_bindings as acb  # noqa: E402  (reference C++ PID, oracle/_ref)
from controllers.attitude_agent import AttitudeAgent  # noqa: E402
from controllers.rate_agent import RateAgent  # noqa: E402
from controllers.types import (AircraftState, ControlCommand, ControlMode, ControllerConfig,  # noqa: E402
                               ControlSurfaces)
from simulation.simulation_backend import SimulationAircraftBackend  # noqa: E402


def state_vec(backend):
    ph = backend._physics
    return np.concatenate([ph._state, [ph._time]])


def derived_vec(st):
    return np.array([st.airspeed, st.altitude, st.ground_speed, st.heading])


def make_state(v):
    return AircraftState(time=float(v[12]), position=v[0:3].copy(), velocity=v[3:6].copy(),
                         attitude=v[6:9].copy(), angular_rate=v[9:12].copy())


def sample_initial_states(rng, n):
    """SURVEY.md 8(d) C2 inputs: V~U(15,30), h~U(50,200), phi,theta~U(+-15deg), psi~U(0,2pi), pqr~U(+-0.1)."""
    s = np.zeros((13, n))
    s[2] = -rng.uniform(50, 200, n)
    s[3] = rng.uniform(15, 30, n)
    s[6] = rng.uniform(-np.radians(15), np.radians(15), n)
    s[7] = rng.uniform(-np.radians(15), np.radians(15), n)
    s[8] = rng.uniform(0, 2 * np.pi, n)
    s[9:12] = rng.uniform(-0.1, 0.1, (3, n))
    return s


# ------------------------------------------------------------------------------------------------
def gen_pid_kat():
    """Known-answer vectors for cpp/src/pid_controller.cpp through its own pybind module."""
    rng = np.random.default_rng(7)
    cases = []
    cfgs = [
        dict(kp=1.3, ki=0.4, kd=0.012, omin=-1.0, omax=1.0, imin=-25.0, imax=25.0, alpha=0.1),
        dict(kp=8.0, ki=2.0, kd=0.3, omin=-np.pi, omax=np.pi, imin=-25.0, imax=25.0, alpha=0.1),
        dict(kp=0.12, ki=0.03, kd=0.03, omin=-0.5, omax=0.5, imin=-10.0, imax=10.0, alpha=0.1),
        dict(kp=2.0, ki=5.0, kd=0.5, omin=-0.5, omax=0.5, imin=-0.2, imax=0.2, alpha=1.0),
        dict(kp=0.0, ki=1.0, kd=0.0, omin=-10.0, omax=10.0, imin=-5.0, imax=5.0, alpha=0.3),
    ]
    T = 400
    cfg_arr = np.zeros((len(cfgs), 8), np.float32)
    sp = np.zeros((len(cfgs), T), np.float32); meas = np.zeros_like(sp); dts = np.zeros_like(sp)
    out = np.zeros((len(cfgs), T, 4), np.float32)  # output, error, integral, derivative_filtered
    for ci, c in enumerate(cfgs):
        pc = acb.PIDConfig()
        pc.gains = acb.PIDGains(c["kp"], c["ki"], c["kd"])
        pc.output_min, pc.output_max = c["omin"], c["omax"]
        pc.integral_min, pc.integral_max = c["imin"], c["imax"]
        pc.derivative_filter_alpha = c["alpha"]
        pid = acb.PIDController(pc)
        cfg_arr[ci] = [c["kp"], c["ki"], c["kd"], c["omin"], c["omax"], c["imin"], c["imax"], c["alpha"]]
        for t in range(T):
            s_ = np.float32(rng.normal(0, 1.0) if t % 50 else rng.normal(0, 5.0))
            m_ = np.float32(rng.normal(0, 1.0))
            d_ = np.float32([0.01, 0.001, 0.02, 1e-7][t % 4] if t % 37 else 0.0)
            if t == 123:
                m_ = np.float32(np.nan)   # NaN path through std::min/std::max
            if t == 130:
                pid.reset(); out[ci, t] = np.nan; sp[ci, t] = np.nan; continue  # marker: reset here
            o = pid.compute(float(s_), float(m_), float(d_))
            sp[ci, t], meas[ci, t], dts[ci, t] = s_, m_, d_
            out[ci, t] = [o, pid.get_error(), pid.get_integral(), pid.get_derivative()]
    np.savez(os.path.join(HERE, "pid_kat.npz"), cfg=cfg_arr, sp=sp, meas=meas, dt=dts, out=out)
    print("pid_kat", out.shape)


def gen_dynamics_kat():
    """Simplified6DOF._dynamics on random (incl. degenerate) states."""
    from simulation.simplified_6dof import Simplified6DOF
    rng = np.random.default_rng(5)
    sim = Simplified6DOF()
    n = 3000
    S = np.zeros((n, 12)); Cc = np.zeros((n, 4)); O = np.zeros((n, 12))
    for i in range(n):
        s = S[i]
        s[0:3] = rng.uniform(-500, 500, 3); s[3] = rng.uniform(-5, 40); s[4:6] = rng.uniform(-8, 8, 2)
        s[6] = rng.uniform(-3.2, 3.2); s[7] = rng.uniform(-1.6, 1.6); s[8] = rng.uniform(-3.2, 3.2)
        s[9:12] = rng.uniform(-4, 4, 3)
        if i % 50 == 0: s[3] = rng.uniform(-1e-6, 1e-6)       # u_safe branch
        if i % 77 == 0: s[3:6] = 0.0                          # zero airspeed
        if i % 91 == 0: s[3:6] = rng.uniform(-90, 90, 3)      # derivative clamps
        c = Cc[i]; c[0:3] = rng.uniform(-1, 1, 3); c[3] = rng.uniform(0, 1)
        ctl = ControlSurfaces(elevator=np.float64(c[0]), aileron=np.float64(c[1]), rudder=np.float64(c[2]),
                              throttle=np.float64(c[3]))
        O[i] = sim._dynamics(s, ctl)
    np.savez(os.path.join(HERE, "dynamics_kat.npz"), s=S, ctrl=Cc, out=O)
    print("dynamics_kat", O.shape)


# ------------------------------------------------------------------------------------------------
def _open_loop_worker(args):
    s0, ctrl_seq, hold, steps, dt, dt_physics, every = args
    b = SimulationAircraftBackend({"aircraft_type": "rc_plane", "dt_physics": dt_physics})
    b.reset(make_state(s0))
    snaps, ders = [], []
    for t in range(steps):
        c = ctrl_seq[t // hold]
        b.set_controls(ControlSurfaces(elevator=c[0], aileron=c[1], rudder=c[2], throttle=c[3]))
        st = b.step(dt)
        if (t + 1) % every == 0:
            snaps.append(state_vec(b)); ders.append(derived_vec(st))
    return np.array(snaps), np.array(ders)


def gen_open_loop(pool):
    """Config 2: open-loop surface-level stepping, piecewise-constant random surfaces."""
    for name, n, steps, dt, dtp, every in [("open_loop_k1", 16, 10000, 0.01, 0.01, 25),
                                           ("open_loop_k10", 8, 1000, 0.01, 0.001, 5),
                                           ("open_loop_k20", 8, 300, 0.02, 0.001, 1)]:
        rng = np.random.default_rng(1234)
        s0 = sample_initial_states(rng, n)
        hold = 50
        nseg = (steps + hold - 1) // hold
        ctrl = np.zeros((nseg, 4, n))
        # surface amplitude per aircraft: U(+-0.3) tumbles this airframe out of the env's envelope
        # within ~0.5 s, so most aircraft get gentler inputs that keep them flying for the whole run
        amp = np.array([0.01, 0.01, 0.03, 0.01, 0.1, 0.01, 0.03, 0.3] * 2)[:n]
        ctrl[:, 0:3] = rng.uniform(-1.0, 1.0, (nseg, 3, n)) * amp
        ctrl[:, 3] = rng.uniform(0.3, 0.9, (nseg, n))
        # a few out-of-range commands to exercise set_controls clipping
        ctrl[1, 0, 0] = 1.7; ctrl[1, 3, 1] = -0.2; ctrl[2, 1, 2] = -3.0
        res = pool.map(_open_loop_worker, [(s0[:, i], ctrl[:, :, i], hold, steps, dt, dtp, every) for i in range(n)])
        states = np.stack([r[0] for r in res], axis=2)   # [snap][13][n]
        ders = np.stack([r[1] for r in res], axis=2)     # [snap][4][n]
        np.savez(os.path.join(HERE, name + ".npz"), s0=s0, ctrl=ctrl, hold=hold, steps=steps, dt=dt,
                 dt_physics=dtp, every=every, states=states, derived=ders)
        print(name, states.shape)


# ------------------------------------------------------------------------------------------------
def _cascade_worker(args):
    level, s0, cmd_seq, hold, steps, dt, dt_ctrl, every = args
    b = SimulationAircraftBackend({"aircraft_type": "rc_plane"})
    agent = RateAgent(ControllerConfig()) if level == 4 else AttitudeAgent(ControllerConfig())
    state = b.reset(make_state(s0))
    S, U, PI = [], [], []
    for t in range(steps):
        c = cmd_seq[t // hold]
        if level == 4:
            cmd = ControlCommand(mode=ControlMode.RATE, roll_rate=c[0], pitch_rate=c[1], yaw_rate=c[2], throttle=c[3])
        else:
            # NaN = yaw_angle None; +inf = hold the current yaw, the command HSAAgent issues (hsa_agent.py:214-216)
            yaw = None if np.isnan(c[2]) else (state.yaw if np.isinf(c[2]) else c[2])
            cmd = ControlCommand(mode=ControlMode.ATTITUDE, roll_angle=c[0], pitch_angle=c[1],
                                 yaw_angle=yaw, throttle=c[3])
        surf = agent.compute_action(cmd, state, dt=dt_ctrl)
        b.set_controls(surf)
        state = b.step(dt)
        if (t + 1) % every:
            continue
        S.append(state_vec(b)); U.append(surf.to_array())
        rc = agent.rate_controller if level == 4 else agent.rate_agent.rate_controller
        e = rc.get_error(); i_ = rc.get_integral()
        row = [e.x, e.y, e.z, i_.x, i_.y, i_.z]
        if level == 3:
            e2 = agent.angle_controller.get_error(); i2 = agent.angle_controller.get_integral()
            row += [e2.x, e2.y, e2.z, i2.x, i2.y, i2.z]
        PI.append(row)
    return np.array(S), np.array(U), np.array(PI)


def gen_cascade(pool):
    """Config 1 (RateAgent, 30 deg/s roll step) + random rate / attitude command sequences."""
    for name, level, n, steps, dt, dt_ctrl, every in [("rate_agent", 4, 8, 1000, 0.01, 0.01, 4),
                                                      ("attitude_agent", 3, 8, 2000, 0.01, 0.01, 8),
                                                      ("attitude_agent_dt02", 3, 4, 500, 0.02, 0.02, 2),
                                                      ("attitude_agent_10k", 3, 8, 10000, 0.01, 0.01, 50)]:
        rng = np.random.default_rng(99 + level)
        s0 = sample_initial_states(rng, n)
        # aircraft 0 is the reference's hello-controls case: level 100 m, 20 m/s
        s0[:, 0] = 0.0; s0[2, 0] = -100.0; s0[3, 0] = 20.0
        hold = 250
        nseg = (steps + hold - 1) // hold
        cmd = np.zeros((nseg, 4, n))
        if level == 4:
            cmd[:, 0:3] = rng.uniform(-0.6, 0.6, (nseg, 3, n))
            cmd[:, :, 0] = 0.0; cmd[:, 0, 0] = np.radians(30.0)          # examples/01_hello_controls.py
            cmd[1, 0, 1] = 5.0                                            # beyond max_roll_rate -> clipped
        else:
            # gentle attitude hold: the default gains are only well-conditioned near level flight; aggressive
            # commands tumble the aircraft and 1-ulp differences then grow by 10x per 0.1 s (measured)
            cmd[:, 0] = rng.uniform(-0.35, 0.35, (nseg, n))
            cmd[:, 1] = rng.uniform(-0.03, 0.08, (nseg, n))
            cmd[:, 2] = np.inf                                             # hold current yaw
            cmd[1, 0, 1] = 1.0                                            # beyond max_roll -> clipped
            if name == "attitude_agent_dt02":
                s0[8] = rng.uniform(-0.3, 0.3, n)
                cmd[:, 0] = rng.uniform(-0.05, 0.05, (nseg, n))
                cmd[:, 2, 0:2] = np.nan                                   # yaw_angle=None -> 0.0
                cmd[:, 2, 2:4] = s0[8, 2:4] + rng.uniform(-0.2, 0.2, (nseg, 2)) + 2 * np.pi   # wrapped setpoint
        cmd[:, 3] = rng.uniform(0.4, 0.8, (nseg, n)); cmd[:, 3, 0] = 0.7
        res = pool.map(_cascade_worker, [(level, s0[:, i], cmd[:, :, i], hold, steps, dt, dt_ctrl, every) for i in range(n)])
        np.savez(os.path.join(HERE, name + ".npz"), level=level, s0=s0, cmd=cmd, hold=hold, steps=steps, dt=dt,
                 dt_ctrl=dt_ctrl, every=every,
                 states=np.stack([r[0] for r in res], axis=2),      # [snap][13][n]
                 surfaces=np.stack([r[1] for r in res], axis=2),    # [snap][4][n]
                 pid=np.stack([r[2] for r in res], axis=2))         # [snap][6 or 12][n]
        print(name, res[0][0].shape)


# ------------------------------------------------------------------------------------------------
def _env_worker(args):
    from learned_controllers.envs.rate_env import RateControlEnv
    from learned_controllers.envs.residual_rate_env import ResidualRateControlEnv
    seed, difficulty, command_type, controller, steps = args
    rng = np.random.default_rng(seed)
    if controller == 3:
        env = ResidualRateControlEnv(difficulty=difficulty, command_type=command_type, rng_seed=seed)
        base = env.base_env
    else:
        env = RateControlEnv(difficulty=difficulty, command_type=command_type, rng_seed=seed)
        base = env
    obs0, info = env.reset(seed=seed)
    obs0 = obs0.copy()
    s0 = state_vec(base.sim)
    cmd0 = base.rate_command.copy()
    sched = np.zeros(7)
    if command_type == "ramp":
        sched[0:3] = base.command_schedule["end"]
    elif command_type == "sine":
        sched[0:3] = base.sine_params["amplitudes"]; sched[3] = base.sine_params["frequency"]
    agent = RateAgent(ControllerConfig()) if controller == 1 else None
    A, O, R, TE, TR, CMP, S, CMD, SET = [], [], [], [], [], [], [], [], []
    for t in range(steps):
        if controller == 0:      # external float32 policy actions (random, some out of the Box)
            a = rng.uniform(-1.2, 1.2, 4).astype(np.float32)
            if t < 150:          # start gently so the episode survives long enough to be interesting
                a = (a * 0.1).astype(np.float32); a[3] = np.float32(0.6)
        elif controller == 3:    # residual policy output
            a = rng.uniform(-1.0, 1.0, 4).astype(np.float32)
        else:                    # pid_demonstrations.py:55-78
            st = base.sim.get_state()
            p, q, r = base.rate_command
            sf = agent.compute_action(ControlCommand(mode=ControlMode.RATE, roll_rate=p, pitch_rate=q, yaw_rate=r,
                                                     throttle=0.6), st, dt=0.02)
            a = np.array([sf.aileron, sf.elevator, sf.rudder, sf.throttle], dtype=np.float32)
        obs, rew, term, trunc, info = env.step(a)
        rc = info["reward_components"]
        comp = [rc["tracking"], rc["smoothness"], rc["stability"], rc["oscillation"], rc["survival"],
                0.0, rc.get("crash_penalty", 0.0), 0.0]
        comp[5] = float(rew) - float(rc["total"]) - comp[6]      # settle bonus (+ residual bonus for controller 3)
        A.append(a); O.append(obs.copy()); R.append(float(rew)); TE.append(term); TR.append(trunc)
        CMP.append(comp); S.append(state_vec(base.sim)); CMD.append(base.rate_command.copy())
        SET.append(base.settle_bonus.is_settled)
        if term or trunc:
            break
    return dict(s0=s0, cmd0=cmd0, sched=sched, obs0=obs0, actions=np.array(A), obs=np.array(O), reward=np.array(R),
                terminated=np.array(TE), truncated=np.array(TR), components=np.array(CMP), states=np.array(S),
                rate_cmd=np.array(CMD), is_settled=np.array(SET))


def gen_env(pool):
    """RateControlEnv / ResidualRateControlEnv episodes with the reference's own initial draws recorded."""
    scale = {"easy": 0.3, "medium": 0.5, "hard": 0.7}
    cases = []
    for ctrl in (0, 1, 3):
        for diff, ctype in [("easy", "step"), ("medium", "step"), ("hard", "step"), ("medium", "ramp"),
                            ("medium", "sine"), ("hard", "sine")]:
            cases.append((1000 + len(cases), diff, ctype, ctrl, 500))
    res = pool.map(_env_worker, cases)
    out = {"n_cases": len(cases)}
    for k, (c, r) in enumerate(zip(cases, res)):
        out[f"c{k}_meta"] = np.array([c[0], scale[c[1]], {"step": 0, "ramp": 1, "sine": 2}[c[2]], c[3]], dtype=np.float64)
        for key, v in r.items():
            out[f"c{k}_{key}"] = v
        print("env case", k, c, "steps", len(r["reward"]), "return", r["reward"].sum())
    np.savez_compressed(os.path.join(HERE, "rate_env.npz"), **out)


class _RecordingRandomState:
    """Stands in for RateCommandGenerator.rng (a plain attribute, generators.py:35): forwards every call to the real
    RandomState and keeps what randn() returned, so that the fixture can carry the reference's own normals."""

    def __init__(self, rs):
        self._rs = rs
        self.normals = []

    def randn(self, *shape):
        z = self._rs.randn(*shape)
        self.normals.append(np.array(z, dtype=np.float64).copy())
        return z

    def __getattr__(self, name):
        return getattr(self._rs, name)


def _env_random_worker(args):
    """command_type="random" (rate_env.py:333-336,358-365; generators.py:145-164): the command is a clipped random walk,
    one randn(3) per env step.  The fixture records those normals next to everything _env_worker records."""
    from learned_controllers.envs.rate_env import RateControlEnv
    from learned_controllers.envs.residual_rate_env import ResidualRateControlEnv
    seed, difficulty, controller, steps, boost = args
    rng = np.random.default_rng(seed)
    if controller == 3:
        env = ResidualRateControlEnv(difficulty=difficulty, command_type="random", rng_seed=seed)
        base = env.base_env
    else:
        env = RateControlEnv(difficulty=difficulty, command_type="random", rng_seed=seed)
        base = env
    rec = _RecordingRandomState(base.cmd_generator.rng)
    base.cmd_generator.rng = rec
    obs0, info = env.reset(seed=seed)
    obs0 = obs0.copy()
    s0 = state_vec(base.sim)
    assert not rec.normals and np.all(base.rate_command == 0.0)        # the walk starts from zero, no draw at reset
    agent = RateAgent(ControllerConfig()) if controller == 1 else None
    A, O, R, TE, TR, CMP, S, CMD, SET, Z = [], [], [], [], [], [], [], [], [], []
    for t in range(steps):
        if boost and t == boost:
            # push the walk to the command limit once so that the fixture exercises the clip (rate_env.py:362-365):
            # the command is the env's public attribute, callers set it (eval_rate.py:104-108)
            base.rate_command[:] = np.array([0.999, -0.999, 0.999]) * base._max_rates
        if controller == 0:
            a = (rng.uniform(-1.2, 1.2, 4) * 0.1).astype(np.float32); a[3] = np.float32(0.6)
        elif controller == 3:
            a = rng.uniform(-1.0, 1.0, 4).astype(np.float32)
        else:
            st = base.sim.get_state()
            p, q, r = base.rate_command
            sf = agent.compute_action(ControlCommand(mode=ControlMode.RATE, roll_rate=p, pitch_rate=q, yaw_rate=r,
                                                     throttle=0.6), st, dt=0.02)
            a = np.array([sf.aileron, sf.elevator, sf.rudder, sf.throttle], dtype=np.float32)
        obs, rew, term, trunc, info = env.step(a)
        rc = info["reward_components"]
        comp = [rc["tracking"], rc["smoothness"], rc["stability"], rc["oscillation"], rc["survival"],
                0.0, rc.get("crash_penalty", 0.0), 0.0]
        comp[5] = float(rew) - float(rc["total"]) - comp[6]
        A.append(a); O.append(obs.copy()); R.append(float(rew)); TE.append(term); TR.append(trunc)
        CMP.append(comp); S.append(state_vec(base.sim)); CMD.append(base.rate_command.copy())
        SET.append(base.settle_bonus.is_settled)
        assert len(rec.normals) == t + 1
        Z.append(rec.normals[-1])
        if term or trunc:
            break
    return dict(s0=s0, obs0=obs0, actions=np.array(A), obs=np.array(O), reward=np.array(R), terminated=np.array(TE),
                truncated=np.array(TR), components=np.array(CMP), states=np.array(S), rate_cmd=np.array(CMD),
                is_settled=np.array(SET), walk_z=np.array(Z), boost=np.array(boost))


def gen_env_random(pool):
    """Random-walk command episodes: external / rate-PID / residual control, all three difficulties."""
    scale = {"easy": 0.3, "medium": 0.5, "hard": 0.7}
    cases = []
    for ctrl in (0, 1, 3):
        for diff in ("easy", "medium", "hard"):
            cases.append((2000 + len(cases), diff, ctrl, 500, 40 if diff == "hard" else 0))
    res = pool.map(_env_random_worker, cases)
    out = {"n_cases": len(cases)}
    for k, (c, r) in enumerate(zip(cases, res)):
        out[f"c{k}_meta"] = np.array([c[0], scale[c[1]], 3, c[2]], dtype=np.float64)
        for key, v in r.items():
            out[f"c{k}_{key}"] = v
        hit = np.abs(r["rate_cmd"]) >= np.radians([180, 180, 160]) - 1e-15
        print("env random-walk case", k, c, "steps", len(r["reward"]), "return", r["reward"].sum(), "clip hits", int(hit.sum()))
    np.savez_compressed(os.path.join(HERE, "rate_env_random.npz"), **out)


# ------------------------------------------------------------------------------------------------
ENVELOPE_SUPPORT = [(15.0, 30.0), (50.0, 200.0), (-np.radians(15), np.radians(15)), (-np.radians(15), np.radians(15)),
                    (0.0, 2 * np.pi), (-0.1, 0.1), (-0.1, 0.1), (-0.1, 0.1)]     # generators.py:173-176,199-209
ENV_BINS, MAG_BINS, Z_BINS, Z_RANGE = 4096, 1024, 4096, 6.0


def _reset_distribution_worker(args):
    """R resets of ONE unmodified RateControlEnv (the three MT19937 streams run on, rate_env.py:64-77): what the envelope
    sampler and the command generator drew, reduced to counts -- exact supports and fine histograms (= the empirical CDF on
    a grid), so that the fixture stays small whatever R is."""
    from learned_controllers.envs.rate_env import RateControlEnv
    difficulty, ctype, seed, R = args
    scale = {"easy": 0.3, "medium": 0.5, "hard": 0.7}[difficulty]
    env = RateControlEnv(difficulty=difficulty, command_type=ctype, rng_seed=seed)
    max_rates = env._max_rates
    env_q = np.zeros((R, 8)); cmd = np.zeros((R, 3)); freq = np.zeros(R)
    for i in range(R):
        env.reset()
        st = env.sim.get_state()
        env_q[i] = [st.airspeed, st.altitude, st.roll, st.pitch, st.yaw, st.p, st.q, st.r]
        assert st.position[0] == 0.0 and st.position[1] == 0.0 and st.velocity[1] == 0.0 and st.velocity[2] == 0.0
        assert st.velocity[0] == st.airspeed and st.position[2] == -st.altitude                  # rate_env.py:174-183
        if ctype == "step":
            cmd[i] = env.rate_command
        elif ctype == "ramp":
            cmd[i] = env.command_schedule["end"]
            assert np.all(env.rate_command == 0.0) and env.command_schedule["duration"] == 3.0
        else:
            cmd[i] = env.sine_params["amplitudes"]; freq[i] = env.sine_params["frequency"]
            assert np.all(env.rate_command == 0.0)
    env_hist = np.stack([np.histogram(env_q[:, j], bins=ENV_BINS, range=ENVELOPE_SUPPORT[j])[0] for j in range(8)])
    assert env_hist.sum() == 8 * R                                     # nothing outside the stated support
    active = cmd != 0.0
    subset = np.bincount((active * np.array([1, 2, 4])).sum(axis=1), minlength=8)
    norm = np.abs(cmd) / ((0.5 if ctype == "sine" else 1.0) * scale * max_rates)        # U(0.3, 1) where active
    mag_hist = np.stack([np.histogram(norm[active[:, a], a], bins=MAG_BINS, range=(0.3, 1.0))[0] for a in range(3)])
    assert mag_hist.sum() == active.sum()
    out = dict(n=np.array(R), env_hist=env_hist, env_corr=np.corrcoef(env_q.T), subset=subset,
               active=active.sum(axis=0), positive=(cmd > 0).sum(axis=0), mag_hist=mag_hist,
               mag_mean=np.array([norm[active[:, a], a].mean() for a in range(3)]))
    if ctype == "sine":
        out["freq_hist"] = np.histogram(freq, bins=MAG_BINS, range=(0.1, 2.0))[0]
        assert out["freq_hist"].sum() == R
    return out


def _walk_distribution_worker(args):
    """RateCommandGenerator.generate_random_walk(dt=0.02) (generators.py:145-164), R calls: the delta per axis, normalised by
    diffusion * sqrt(dt) * difficulty_scale * max_rate, is a standard normal."""
    from learned_controllers.data.generators import RateCommandGenerator
    difficulty, seed, R = args
    scale = {"easy": 0.3, "medium": 0.5, "hard": 0.7}[difficulty]
    gen = RateCommandGenerator(difficulty=difficulty, rng_seed=seed)
    d = np.stack([gen.generate_random_walk(dt=0.02)[0] for _ in range(R)])
    z = d / (0.1 * np.sqrt(0.02) * scale * gen._max_rates)
    return dict(n=np.array(R), delta_std=d.std(axis=0), delta_mean=d.mean(axis=0),
                z_hist=np.stack([np.histogram(z[:, a], bins=Z_BINS, range=(-Z_RANGE, Z_RANGE))[0] for a in range(3)]),
                z_outside=np.array(int((np.abs(z) >= Z_RANGE).sum())), z_corr=np.corrcoef(z.T))


def gen_reset_distributions(pool):
    """What the reference's FlightEnvelopeSampler / RateCommandGenerator / RateControlEnv._generate_new_command DRAW
    (generators.py:47-164,190-214; rate_env.py:302-340), 100 000 resets per (difficulty, command type): the pin for the
    Philox reset sampler, whose stream cannot replay MT19937 (RNG row, item ii)."""
    R = 100_000
    combos = [(d, c) for d in ("easy", "medium", "hard") for c in ("step", "ramp", "sine")]
    res = pool.map(_reset_distribution_worker, [(d, c, 7000 + k, R) for k, (d, c) in enumerate(combos)])
    out = {"combos": np.array([f"{d}/{c}" for d, c in combos]), "env_support": np.array(ENVELOPE_SUPPORT),
           "env_hist": sum(r["env_hist"] for r in res), "env_n": np.array(R * len(combos)),
           "env_corr_max_offdiag": np.array(max(float(np.max(np.abs(r["env_corr"] - np.eye(8)))) for r in res))}
    for (d, c), r in zip(combos, res):
        for key, v in r.items():
            if key not in ("env_hist", "env_corr"):
                out[f"{d}/{c}/{key}"] = v
        print("reset distribution", d, c, "subsets", r["subset"].tolist(), "positive/active", r["positive"].tolist(), r["active"].tolist())
    walk = pool.map(_walk_distribution_worker, [(d, 7100 + k, 400_000) for k, d in enumerate(("easy", "medium", "hard"))])
    for d, r in zip(("easy", "medium", "hard"), walk):
        for key, v in r.items():
            out[f"{d}/walk/{key}"] = v
        print("random walk", d, "delta std", r["delta_std"], "outside", int(r["z_outside"]))
    np.savez_compressed(os.path.join(HERE, "reset_distributions.npz"), **out)


# ------------------------------------------------------------------------------------------------
def _hsa_worker(args):
    """HSAAgent closed loop through the backend (tests/test_control_integration.py:435-540 shape)."""
    from controllers.hsa_agent import HSAAgent
    s0, cmd_seq, hold, steps, dt, every = args
    b = SimulationAircraftBackend({"aircraft_type": "rc_plane"})
    agent = HSAAgent(ControllerConfig())
    state = b.reset(make_state(s0))
    S, U = [], []
    for t in range(steps):
        c = cmd_seq[t // hold]
        cmd = ControlCommand(mode=ControlMode.HSA, heading=c[0], speed=c[1], altitude=c[2])
        surf = agent.compute_action(cmd, state, dt)
        b.set_controls(surf)
        state = b.step(dt)
        if (t + 1) % every == 0:
            S.append(state_vec(b)); U.append(surf.to_array())
    return np.array(S), np.array(U)


def _mission_worker(args):
    """examples/03_waypoint_square_demo.py:147-209: MissionPlanner.update -> WaypointAgent -> Simplified6DOF.step(DT)
    directly (one RK4 step per control step), square mission, LOS or pure-pursuit guidance."""
    from controllers.mission_planner import MissionPlanner
    from controllers.types import Waypoint
    from controllers.waypoint_agent import WaypointAgent
    from simulation.simplified_6dof import AircraftParams, Simplified6DOF
    s0, wps, guidance, radius, steps, dt, every = args
    sim = Simplified6DOF(AircraftParams())
    sim.reset(make_state(s0))
    agent = WaypointAgent(ControllerConfig(), guidance_type=guidance)
    mission = MissionPlanner([Waypoint(north=w[0], east=w[1], down=w[2], speed=None if np.isnan(w[3]) else float(w[3]))
                              for w in wps], acceptance_radius=radius)
    mission.start()
    S, U, IDX, HC = [], [], [], []
    surf = ControlSurfaces()
    hc = np.zeros(3)
    for t in range(steps):
        state = sim.get_state()
        mission.update(state)
        if not mission.is_complete():
            cmd = mission.get_waypoint_command()
            surf = agent.compute_action(cmd, state, dt)
            sim.set_controls(surf)
            sim.step(dt)
        if (t + 1) % every == 0:
            S.append(np.concatenate([sim._state, [sim._time]])); U.append(surf.to_array())
            IDX.append(mission.current_waypoint_index)
            HC.append(agent.last_heading_cmd if agent.last_heading_cmd is not None else 0.0)
    return np.array(S), np.array(U), np.array(IDX), np.array(HC)


def gen_guidance(pool):
    """Levels 2 (HSA) and 1 (waypoint mission) -- BASELINE config C5's controller."""
    rng = np.random.default_rng(2024)
    n, steps, dt, every, hold = 8, 3000, 0.01, 10, 1000
    s0 = np.zeros((13, n)); s0[2] = -rng.uniform(80, 120, n); s0[3] = rng.uniform(14, 20, n); s0[8] = rng.uniform(-0.5, 0.5, n)
    s0[:, 0] = 0.0; s0[2, 0] = -100.0; s0[3, 0] = 20.0                    # level 100 m / 20 m/s (conftest level_flight_state)
    nseg = steps // hold
    cmd = np.zeros((nseg, 3, n))
    cmd[:, 0] = rng.uniform(-np.pi, np.pi, (nseg, n))                      # heading
    cmd[:, 1] = rng.uniform(14, 20, (nseg, n))                             # speed
    cmd[:, 2] = rng.uniform(80, 130, (nseg, n))                            # altitude
    cmd[:, :, 0] = [[np.radians(45.0)], [15.0], [120.0]]                   # test_control_integration.py: heading 45 deg, climb to 120 m
    res = pool.map(_hsa_worker, [(s0[:, i], cmd[:, :, i], hold, steps, dt, every) for i in range(n)])
    np.savez(os.path.join(HERE, "hsa_agent.npz"), s0=s0, cmd=cmd, hold=hold, steps=steps, dt=dt, every=every,
             states=np.stack([r[0] for r in res], axis=2), surfaces=np.stack([r[1] for r in res], axis=2))
    print("hsa_agent", res[0][0].shape)

    # square mission of the demo (configs/missions/square_pattern.yaml: 300 m, 100 m altitude, 15 m/s, radius 40 m),
    # one waypoint without a speed (None -> current airspeed)
    size, alt, spd, radius = 300.0, 100.0, 15.0, 40.0
    wps = np.array([[0, 0, -alt, spd], [size, 0, -alt, spd], [size, size, -alt, np.nan], [0, size, -alt, spd], [0, 0, -alt, spd]])
    n, steps, every = 8, 9000, 25
    s0 = np.zeros((13, n)); s0[2] = -alt; s0[3] = spd
    s0[0, 1:] = rng.uniform(-20, 20, n - 1); s0[1, 1:] = rng.uniform(-20, 20, n - 1)       # C5: +-20 m, +-10 deg jitter
    s0[8, 1:] = rng.uniform(-np.radians(10), np.radians(10), n - 1)
    for guidance in ("PP", "LOS"):
        res = pool.map(_mission_worker, [(s0[:, i], wps, guidance, radius, steps, dt, every) for i in range(n)])
        np.savez(os.path.join(HERE, f"waypoint_{guidance.lower()}.npz"), s0=s0, waypoints=wps, guidance=0 if guidance == "LOS" else 1,
                 acceptance_radius=radius, steps=steps, dt=dt, every=every,
                 states=np.stack([r[0] for r in res], axis=2), surfaces=np.stack([r[1] for r in res], axis=2),
                 wp_index=np.stack([r[2] for r in res], axis=1), heading_cmd=np.stack([r[3] for r in res], axis=1))
        print("waypoint", guidance, "final waypoint index per aircraft", res[0][2][-1], [int(r[2][-1]) for r in res])


# ------------------------------------------------------------------------------------------------
def gen_sensor():
    """NoisySensorInterface (interfaces/sensor.py:137-263): per case, T updates on a moving truth.  The fixture keeps the
    truth, the reference's measured states and bias history, and the 20 standard normals per update the reference's PCG64
    Generator consumed (replayed here from the same seed: Generator.normal(0, s, k) = s * standard_normal(k), in
    order), so the test can push the SAME normals through the restatement and compare bit for bit."""
    from interfaces.sensor import NoisySensorInterface
    cases = {
        "default_seed42": {"enabled": True, "seed": 42},
        "custom_seed123": {"enabled": True, "imu_gyro_stddev": 0.1, "gps_position_stddev": 5.0, "attitude_stddev": 0.05,
                           "airspeed_stddev": 1.0, "seed": 123},              # tests/test_interfaces.py:361-369
        "gps_seed999": {"enabled": True, "gps_position_stddev": 1.0, "seed": 999},   # tests/test_interfaces.py:406-411
        "disabled": {"enabled": False, "imu_gyro_stddev": 100.0, "seed": 5},         # tests/test_interfaces.py:391-395
    }
    T = 120
    out = {}
    for name, nc in cases.items():
        rng = np.random.default_rng(1000 + len(name))
        truth = sample_initial_states(rng, T)                 # [13][T]: a different truth per update
        truth[0:2] = rng.uniform(-500, 500, (2, T)); truth[4:6] = rng.uniform(-2, 2, (2, T)); truth[12] = 0.01 * np.arange(T)
        der = np.stack([rng.uniform(12, 30, T), -truth[2], rng.uniform(10, 30, T), rng.uniform(0, 2 * np.pi, T)])
        sensor = NoisySensorInterface(dict(nc))
        meas = np.zeros((T, 17)); bias = np.zeros((T, 6))
        for t in range(T):
            st = AircraftState(time=float(truth[12, t]), position=truth[0:3, t].copy(), velocity=truth[3:6, t].copy(),
                               attitude=truth[6:9, t].copy(), angular_rate=truth[9:12, t].copy(), airspeed=float(der[0, t]),
                               altitude=float(der[1, t]), ground_speed=float(der[2, t]), heading=float(der[3, t]))
            sensor.update(st)
            m = sensor.get_state()
            meas[t] = np.concatenate([m.position, m.velocity, m.attitude, m.angular_rate, [m.time, m.airspeed, m.altitude,
                                                                                           m.ground_speed, m.heading]])
            bias[t] = np.concatenate([sensor._gyro_bias, sensor._accel_bias])
        z = np.random.default_rng(nc["seed"]).standard_normal((T, 20)) if nc["enabled"] else np.zeros((T, 20))
        out[name + "/truth"] = np.concatenate([truth, der]).T.copy()      # [T][17]
        out[name + "/measured"] = meas
        out[name + "/bias"] = bias
        out[name + "/z"] = z
        out[name + "/params"] = np.array([float(nc["enabled"])] + [sensor.get_noise_parameters()[k] for k in (
            "imu_gyro_stddev", "imu_accel_stddev", "gps_position_stddev", "gps_velocity_stddev", "airspeed_stddev",
            "altitude_stddev", "attitude_stddev")] + [float(nc["seed"])])
    # distribution summary of the reference's own noise: 4000 sensors x 64 updates at the default config
    R, U = 4000, 64
    noise = np.zeros((R, 14)); walk = np.zeros((R, 6))
    zero = AircraftState()
    for r in range(R):
        sensor = NoisySensorInterface({"enabled": True, "seed": 10_000 + r})
        for _ in range(U):
            sensor.update(zero)
        m = sensor.get_state()
        noise[r] = np.concatenate([m.position, m.velocity, m.attitude, m.angular_rate - 0.0, [m.airspeed, m.altitude]])
        walk[r] = np.concatenate([sensor._gyro_bias, sensor._accel_bias])
    out["stats/updates"] = np.array([U]); out["stats/sensors"] = np.array([R])
    out["stats/noise_std"] = noise.std(axis=0); out["stats/noise_mean"] = noise.mean(axis=0)
    out["stats/walk_std"] = walk.std(axis=0); out["stats/walk_mean"] = walk.mean(axis=0)
    np.savez_compressed(os.path.join(HERE, "sensor.npz"), **out)
    print("sensor.npz", {k: v.shape for k, v in out.items() if k.startswith("default")}, out["stats/noise_std"], out["stats/walk_std"])


# ------------------------------------------------------------------------------------------------
def gen_telemetry():
    """TelemetryLogger's JSON layout (visualization/logger.py:72-147) for 3 aircraft x 6 steps: the fixture keeps the
    records the states were built from ([T][21][k]: state 13 | derived 4 | controls 4 in ControlSurfaces.to_array order)
    and the file the unmodified logger wrote, byte for byte."""
    import tempfile
    import importlib.util          # visualization/__init__.py pulls in matplotlib (absent here): load logger.py by path
    ref = os.environ.get("REFERENCE", "/root/reference")
    spec = importlib.util.spec_from_file_location("ref_telemetry_logger", os.path.join(ref, "visualization", "logger.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    TelemetryLogger = mod.TelemetryLogger
    rng = np.random.default_rng(77)
    T, k = 6, 3
    rec = rng.normal(size=(T, 21, k))
    rec[:, 12, :] = (0.01 * np.arange(1, T + 1))[:, None]
    ids = ["lead", "wing_left", "wing_right"]
    meta = {"lead": {"type": "rc_plane", "formation_index": 0}, "wing_left": {"type": "rc_plane", "formation_index": 1},
            "wing_right": {}}
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "log.json")
        lg = TelemetryLogger(path)
        for a in ids:
            lg.register_aircraft(a, metadata=meta[a] or None)
        for t in range(T):
            for j, a in enumerate(ids):
                r = rec[t, :, j]
                st = AircraftState(time=float(r[12]), position=r[0:3].copy(), velocity=r[3:6].copy(), attitude=r[6:9].copy(),
                                   angular_rate=r[9:12].copy(), airspeed=float(r[13]), altitude=float(r[14]),
                                   ground_speed=float(r[15]), heading=float(r[16]))
                lg.log_state(a, st)
                lg.log_command(a, ControlCommand(mode=ControlMode.WAYPOINT), float(r[12]))
                lg.log_surfaces(a, ControlSurfaces(elevator=float(r[17]), aileron=float(r[18]), rudder=float(r[19]),
                                                   throttle=float(r[20])), float(r[12]))
        lg.close()
        text = open(path).read()
    open(os.path.join(HERE, "telemetry.json"), "w").write(text)
    np.save(os.path.join(HERE, "telemetry_records.npy"), rec)
    print("telemetry.json", len(text), "bytes")


def gen_policy_features():
    """The half of the learned policy that lives in the reference: LSTMFeaturesExtractor (networks/lstm_policy.py:13-97)
    run as it is (its SB3 base class comes from oracle/sb3_stub.py: an nn.Module that stores features_dim), plus what
    LSTMPolicy / SimpleMLPPolicy.get_policy_kwargs (:107-164) hand to sb3-contrib.  The fixture keeps the extractor's
    state_dict (torch default init under manual_seed(0)), 512 observations, the features of the single-step call and of
    a 3-step sequence call (the [:, -1, :] branch, :88-91)."""
    import torch
    from gymnasium import spaces
    from learned_controllers.networks.lstm_policy import LSTMFeaturesExtractor, LSTMPolicy, SimpleMLPPolicy
    sys.path.insert(0, os.path.join(HERE, "..", ".."))
    from oracle.policy_oracle import synthetic_obs
    torch.manual_seed(0)
    space = spaces.Box(low=-np.inf, high=np.inf, shape=(18,), dtype=np.float32)
    kw = LSTMPolicy.get_policy_kwargs()
    assert kw["features_extractor_class"] is LSTMFeaturesExtractor
    fe = kw["features_extractor_class"](space, **kw["features_extractor_kwargs"]).eval()
    obs = synthetic_obs(512, seed=123)
    seq = torch.stack([synthetic_obs(64, seed=200 + t) for t in range(3)], dim=1)          # [64, 3, 18]
    # the sequence branch needs weight_hh (h is non-zero from the second step on): pinned on a SMALL instance of the same
    # class (hidden 32, features 16) whose whole state_dict is a few kB
    small = LSTMFeaturesExtractor(space, features_dim=16, lstm_hidden_size=32, n_lstm_layers=2).eval()
    with torch.no_grad():
        feats = fe(obs); feats_seq = small(seq); feats_small = small(obs[:64])
    out = {"obs": obs.numpy(), "features": feats.numpy(), "obs_seq": seq.numpy(), "features_seq": feats_seq.numpy(), "features_small": feats_small.numpy(),
           "features_dim": np.array(fe.features_dim),
           "lstm_kwargs": np.array([kw["features_extractor_kwargs"][k] for k in ("features_dim", "lstm_hidden_size", "n_lstm_layers")]),
           "lstm_pi": np.array(kw["net_arch"]["pi"]), "lstm_vf": np.array(kw["net_arch"]["vf"]),
           "lstm_activation": np.array(kw["activation_fn"].__name__),
           "mlp_pi": np.array(SimpleMLPPolicy.get_policy_kwargs()["net_arch"]["pi"]),
           "mlp_activation": np.array(SimpleMLPPolicy.get_policy_kwargs()["activation_fn"].__name__),
           "torch_version": np.array(torch.__version__)}
    for k, v in fe.state_dict().items():
        if "weight_hh" in k:        # 2 MB that only ever multiply h0 = 0 (no state is passed, :86): the test loads random ones
            continue
        out["sd/" + k] = v.numpy()
    for k, v in small.state_dict().items():
        out["small/" + k] = v.numpy()
    np.savez_compressed(os.path.join(HERE, "policy_features.npz"), **out)
    print("policy_features", feats.shape, "state_dict", [k for k in fe.state_dict()], "|features| max", float(feats.abs().max()))


def gen_flight_episode():
    """TensorBoard flight plugin: one episode logged by the UNMODIFIED FlightLoggerCallback._extract_flight_data
    (tensorboard_flight/callbacks.py:152-263) and FlightLogger (logger.py:93-293) from a recorded RateControlEnv episode
    (first 40 steps of case 0 of rate_env.npz, info dicts rebuilt from the recorded states).  Run with
    PYTHONPATH=oracle/_ref:/root/reference:/root/reference/tensorboard_flight_plugin/src.  The fixture keeps the inputs,
    the reference's step records and the episode JSON read back from the event file it wrote."""
    import json
    import tempfile
    from tensorboard.backend.event_processing.event_file_loader import EventFileLoader
    from tensorboard_flight.callbacks import FlightLoggerCallback
    from tensorboard_flight.logger import FlightLogger
    g = np.load(os.path.join(HERE, "rate_env.npz"))
    T = 40
    names = ("tracking", "smoothness", "stability", "oscillation", "survival")
    with tempfile.TemporaryDirectory() as d:
        logger = FlightLogger(d)
        cb = FlightLoggerCallback(logger, agent_id="golden_agent")
        inputs, records = [], []
        for t in range(T):
            obs, act, rew, st = g["c0_obs"][t], g["c0_actions"][t], float(g["c0_reward"][t]), g["c0_states"][t]
            comp = {n: float(g["c0_components"][t][j]) for j, n in enumerate(names)}
            info = {"time": 0.02 * (t + 1), "step": t + 1, "position": st[0:3].copy(), "reward_components": comp}
            if t == T - 1:
                info.update(termination_reason="crash_altitude", dones=True)
            rec = cb._extract_flight_data(obs, act, rew, info, 1000 + t)
            logger.log_flight_data(**rec)
            inputs.append({"obs": obs.tolist(), "action": act.tolist(), "reward": rew, "step": 1000 + t,
                           "info": {k: (v.tolist() if isinstance(v, np.ndarray) else v) for k, v in info.items()}})
            records.append(json.loads(json.dumps(rec, default=lambda o: o.tolist())))
        logger.end_episode(success=False, termination_reason="crash_altitude", config={"episode_num": 7}, tags=["training_step_1040"])
        logger.close()
        episodes = []
        for fn in sorted(os.listdir(d)):
            for ev in EventFileLoader(os.path.join(d, fn)).Load():
                for v in ev.summary.value:
                    if v.metadata.plugin_data.plugin_name == "flight":
                        episodes.append({"tag": v.tag, "step": ev.step, "content": json.loads(v.metadata.plugin_data.content.decode())})
    assert len(episodes) == 1
    with open(os.path.join(HERE, "flight_episode.json"), "w") as f:
        json.dump({"inputs": inputs, "records": records, "episode": episodes[0]}, f)
    print("flight_episode.json", len(inputs), "steps; tag", episodes[0]["tag"], "keys", sorted(episodes[0]["content"]))


if __name__ == "__main__":
    if "flight" in sys.argv[1:]:
        gen_flight_episode()
        sys.exit(0)
    if "policy" in sys.argv[1:]:
        gen_policy_features()
        sys.exit(0)
    which = sys.argv[1:] or ["pid", "dyn", "open", "cascade", "env", "env_random", "reset_dist", "guidance", "sensor", "telemetry"]
    with mp.Pool(8) as pool:
        if "pid" in which: gen_pid_kat()
        if "dyn" in which: gen_dynamics_kat()
        if "open" in which: gen_open_loop(pool)
        if "cascade" in which: gen_cascade(pool)
        if "env" in which: gen_env(pool)
        if "env_random" in which: gen_env_random(pool)
        if "reset_dist" in which: gen_reset_distributions(pool)
        if "guidance" in which: gen_guidance(pool)
        if "sensor" in which: gen_sensor()
        if "telemetry" in which: gen_telemetry()
