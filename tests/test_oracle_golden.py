"""Pin the CPU restatement (oracle/) against fixtures produced by the unmodified reference.

tests/golden/*.npz come from tests/golden/make_golden.py (reference Python + reference C++ PID).
Tolerances: PID bit-exact; fp64 physics 1e-9 * max(1,|x|) (BASELINE.json north_star), which leaves
room for the 1-ulp libm differences between numpy's SIMD tan/arcsin/arctan2/exp and glibc.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err
from oracle import oracle as orc

TOL = 1e-9


def _load(name):
    return np.load(os.path.join(GOLDEN, name))


def test_pid_known_answers_bit_exact():
    g = _load("pid_kat.npz")
    for ci in range(g["cfg"].shape[0]):
        kp, ki, kd, omin, omax, imin, imax, alpha = [float(x) for x in g["cfg"][ci]]
        pid = orc.Pid(orc.PidConfig(kp=kp, ki=ki, kd=kd, output_min=omin, output_max=omax, integral_min=imin,
                                    integral_max=imax, derivative_filter_alpha=alpha))
        for t in range(g["sp"].shape[1]):
            if np.isnan(g["sp"][ci, t]):            # reset marker
                pid = orc.Pid(pid.cfg)
                continue
            o = pid.compute(g["sp"][ci, t], g["meas"][ci, t], g["dt"][ci, t])
            got = np.array([o, pid.st.error, pid.st.integral, pid.st.derivative_filtered], np.float32)
            exp = g["out"][ci, t]
            assert got.tobytes() == exp.tobytes() or (np.isnan(got) == np.isnan(exp)).all() and \
                np.array_equal(got[~np.isnan(got)], exp[~np.isnan(exp)]), (ci, t, got, exp)


def test_pid_reference_analytic_cases():
    """The analytic cases of the reference's tests/test_pid_bindings.py:89-209."""
    mk = lambda **kw: orc.Pid(orc.PidConfig(**{**dict(kp=0, ki=0, kd=0, output_min=-1, output_max=1, integral_min=-10,
                                                      integral_max=10, derivative_filter_alpha=0.1), **kw}))
    assert mk(kp=2.0).compute(1.0, 0.0, 0.01) == 1.0                       # P-only saturates   :89-102
    pid = mk(ki=1.0)
    for _ in range(10):
        out = pid.compute(1.0, 0.0, 0.1)
    assert abs(out - 1.0) < 0.01                                           # I accumulates      :104-123
    pid = mk(kp=10.0, output_min=-0.5, output_max=0.5)
    assert pid.compute(1.0, 0.0, 0.01) == 0.5 and pid.compute(-1.0, 0.0, 0.01) == -0.5   # :160-173
    pid = mk(ki=1.0, integral_min=-5, integral_max=5, output_min=-100, output_max=100)
    for _ in range(100):
        pid.compute(10.0, 0.0, 0.1)
    assert abs(pid.st.integral - 5.0) < 0.01                               # anti-windup        :175-190
    pid = mk(kd=1.0, derivative_filter_alpha=1.0, output_min=-100, output_max=100)
    assert pid.compute(0.0, 0.0, 0.01) == 0.0
    assert abs(pid.compute(1.0, 0.5, 0.01) - 50.0) < 1.0                   # derivative         :192-209


def _terminated(states, derived):
    """RateControlEnv._check_termination (rate_env.py:437-460) on golden snapshots [13][n], [4][n]."""
    return ((derived[1] < 5.0) | (np.abs(states[6]) > np.radians(120)) | (np.abs(states[7]) > np.radians(80))
            | (derived[0] < 8.0))


@pytest.mark.parametrize("name", ["open_loop_k1", "open_loop_k10", "open_loop_k20"])
def test_open_loop_trajectories(name):
    """Config 2.  Each aircraft is compared until it first meets the env's termination condition
    (altitude < 5 m etc.): beyond that it is bouncing along the ground clamp, which is chaotic --
    a 1-ulp libm difference grows past any tolerance (measured: 1e-13 before contact, O(1) by 8k steps)."""
    g = _load(name + ".npz")
    P = orc.default_aircraft_params()
    state = np.ascontiguousarray(g["s0"]); n = state.shape[1]
    hold, steps, every = int(g["hold"]), int(g["steps"]), int(g["every"])
    dt, dtp = float(g["dt"]), float(g["dt_physics"])
    der = np.zeros((4, n))
    alive = np.ones(n, bool)
    worst, compared = 0.0, 0
    for t in range(steps):
        ctrl = np.ascontiguousarray(g["ctrl"][t // hold])
        orc.step_surface(P, state, ctrl, dt, dtp, derived_out=der)
        if (t + 1) % every == 0:
            k = (t + 1) // every - 1
            alive &= ~_terminated(g["states"][k], g["derived"][k])
            if not alive.any():
                break
            e = max(rel_err(state[:, alive], g["states"][k][:, alive]),
                    rel_err(der[:3, alive], g["derived"][k][:3, alive]))
            dh = np.angle(np.exp(1j * (der[3, alive] - g["derived"][k][3, alive])))   # heading lives on a circle
            e = max(e, float(np.max(np.abs(dh))))
            worst = max(worst, e); compared += int(alive.sum())
            assert e <= TOL, f"{name}: step {t + 1}: rel err {e:.3e}"
    assert compared > 0
    print(f"{name}: worst rel err {worst:.3e} over {compared} aircraft-snapshots")


@pytest.mark.parametrize("name", ["rate_agent", "attitude_agent", "attitude_agent_dt02", "attitude_agent_10k"])
def test_cascade_trajectories(name):
    """Config 1 and the attitude cascade, incl. 10 000 control steps (100 000 RK4 substeps) closed loop."""
    g = _load(name + ".npz")
    P, Cp = orc.default_aircraft_params(), orc.default_controller_params()
    level = int(g["level"]); state = np.ascontiguousarray(g["s0"]); n = state.shape[1]
    hold, steps, dt, dt_ctrl, every = int(g["hold"]), int(g["steps"]), float(g["dt"]), float(g["dt_ctrl"]), int(g["every"])
    pid = np.zeros((18, n), np.float32); surf = np.zeros((4, n)); der = np.zeros((4, n))
    alive = np.ones(n, bool)
    worst = 0.0
    for t in range(steps):
        cmd = np.ascontiguousarray(g["cmd"][t // hold])
        orc.step_cascade(P, Cp, level, state, cmd, pid, dt, dt_ctrl=dt_ctrl, surfaces_out=surf, derived_out=der)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        gs = g["states"][k]
        alive &= ~((-gs[2] < 5.0) | (np.abs(gs[6]) > np.radians(120)) | (np.abs(gs[7]) > np.radians(80)))
        if not alive.any():
            break
        e = max(rel_err(state[:, alive], gs[:, alive]), rel_err(surf[:, alive], g["surfaces"][k][:, alive]))
        worst = max(worst, e)
        assert e <= TOL, f"{name}: step {t + 1} rel err {e:.3e}"
        # PID integrals are float: bit-identical wherever the fp64 states are bit-identical
        same = np.all(state == gs, axis=0) & alive
        integ = np.stack([pid[(0 * 3 + a) * 3 + 0] for a in range(3)])
        assert np.array_equal(integ[:, same], g["pid"][k][3:6].astype(np.float32)[:, same])
    assert alive.any(), "every aircraft left the envelope; fixture is useless"
    print(f"{name}: worst rel err over {steps} steps = {worst:.3e}, {int(alive.sum())}/{n} aircraft in envelope at the end")


def test_dynamics_known_answers():
    g = _load("dynamics_kat.npz")
    P = orc.default_aircraft_params()
    exact = 0
    for s, c, ref in zip(g["s"], g["ctrl"], g["out"]):
        got = orc.dynamics(P, np.ascontiguousarray(s), np.ascontiguousarray(c))
        exact += int(np.array_equal(got, ref))
        assert np.all(np.abs(got - ref) <= 1e-12 * np.maximum(np.abs(ref), 1.0)), (s, c, got - ref)
    assert exact >= 0.9 * len(g["s"]), exact     # the rest differ by libm ulps only


def _run_env_case(g, k):
    seed, scale, ctype, ctrl = g[f"c{k}_meta"]
    env = orc.Env(1, controller=int(ctrl), difficulty=float(scale), command_type=int(ctype), auto_reset=False)
    # inject the reference's own initial draws (RNG streams cannot be replayed)
    env.state[:, 0] = g[f"c{k}_s0"]
    env.derived[:, 0] = orc.derived(np.ascontiguousarray(g[f"c{k}_s0"][:12]))
    env.rate_cmd[:, 0] = g[f"c{k}_cmd0"]
    env.sched[:, 0] = g[f"c{k}_sched"]
    env.prev_action[:, 0] = [0, 0, 0, 0.5]
    return env


def test_rate_env_episodes():
    g = _load("rate_env.npz")
    for k in range(int(g["n_cases"])):
        env = _run_env_case(g, k)
        ctrl = int(g[f"c{k}_meta"][3])
        acts, T = g[f"c{k}_actions"], len(g[f"c{k}_reward"])
        for t in range(T):
            a = acts[t][None, :] if ctrl in (0, 3) else None
            obs, rew, term, trunc = env.step(a)
            ctx = f"case {k} step {t + 1}"
            assert rel_err(env.state[:, 0], g[f"c{k}_states"][t]) <= TOL, ctx
            assert rel_err(obs[0], g[f"c{k}_obs"][t]) <= 2e-7, ctx          # float32 observation
            assert abs(rew[0] - g[f"c{k}_reward"][t]) <= 1e-7, (ctx, rew[0], g[f"c{k}_reward"][t])
            assert bool(term[0]) == bool(g[f"c{k}_terminated"][t]), ctx
            assert bool(trunc[0]) == bool(g[f"c{k}_truncated"][t]), ctx
            assert bool(env.is_settled[0]) == bool(g[f"c{k}_is_settled"][t]), ctx
            assert rel_err(env.rate_cmd[:, 0], g[f"c{k}_rate_cmd"][t]) <= 1e-12, ctx
            comp = env.components[:, 0]
            exp = g[f"c{k}_components"][t]
            assert np.allclose(comp[:5], exp[:5], rtol=0, atol=1e-7), ctx
            assert abs((comp[5] + comp[7]) - exp[5]) <= 1e-7 and comp[6] == exp[6], ctx
            if ctrl == 1:   # the PID-composed action must be the reference's float32 action bit for bit
                assert np.array_equal(env.prev_action[:, 0], acts[t]), ctx
        assert bool(term[0]) or bool(trunc[0]) or T == 500


def test_rate_env_random_walk_episodes():
    """command_type="random" (rate_env.py:333-336,358-365; generators.py:145-164): the reference's own MT19937 randn(3) values,
    recorded per step, are replayed through the restatement's hook -- the command after the clip must match to 1e-12, the
    rest of the step to the usual bars.  Cases: external / rate-PID / residual control x easy / medium / hard; the hard cases
    are pushed to the command limit at step `boost` so that the clip is exercised."""
    g = _load("rate_env_random.npz")
    max_rates = np.radians([180.0, 180.0, 160.0])
    clip_hits = 0
    for k in range(int(g["n_cases"])):
        seed, scale, ctype, ctrl = g[f"c{k}_meta"]
        assert int(ctype) == orc.CMD_RANDOM
        env = orc.Env(1, controller=int(ctrl), difficulty=float(scale), command_type=orc.CMD_RANDOM, auto_reset=False)
        env.state[:, 0] = g[f"c{k}_s0"]
        env.derived[:, 0] = orc.derived(np.ascontiguousarray(g[f"c{k}_s0"][:12]))
        env.prev_action[:, 0] = [0, 0, 0, 0.5]
        acts, T, boost = g[f"c{k}_actions"], len(g[f"c{k}_reward"]), int(g[f"c{k}_boost"])
        for t in range(T):
            if boost and t == boost:
                env.rate_cmd[:, 0] = np.array([0.999, -0.999, 0.999]) * max_rates
            a = acts[t][None, :] if int(ctrl) in (0, 3) else None
            obs, rew, term, trunc = env.step(a, walk_z=g[f"c{k}_walk_z"][t][:, None])
            ctx = f"random-walk case {k} step {t + 1}"
            assert rel_err(env.rate_cmd[:, 0], g[f"c{k}_rate_cmd"][t]) <= 1e-12, ctx
            assert rel_err(env.state[:, 0], g[f"c{k}_states"][t]) <= TOL, ctx
            assert rel_err(obs[0], g[f"c{k}_obs"][t]) <= 2e-7, ctx
            assert abs(rew[0] - g[f"c{k}_reward"][t]) <= 1e-7, (ctx, rew[0], g[f"c{k}_reward"][t])
            assert bool(term[0]) == bool(g[f"c{k}_terminated"][t]) and bool(trunc[0]) == bool(g[f"c{k}_truncated"][t]), ctx
            assert bool(env.is_settled[0]) == bool(g[f"c{k}_is_settled"][t]), ctx
            comp, exp = env.components[:, 0], g[f"c{k}_components"][t]
            assert np.allclose(comp[:5], exp[:5], rtol=0, atol=1e-7), ctx
            assert abs((comp[5] + comp[7]) - exp[5]) <= 1e-7 and comp[6] == exp[6], ctx
            if int(ctrl) == 1:
                assert np.array_equal(env.prev_action[:, 0], acts[t]), ctx
            clip_hits += int(np.sum(np.abs(env.rate_cmd[:, 0]) == max_rates))
        assert bool(term[0]) or bool(trunc[0]) or T == 500
    assert clip_hits > 50, "the fixture must exercise the command clip"


def test_difficulty_scale_is_the_reference_python_float():
    """generators.py:41-45: 0.3 / 0.5 / 0.7 as Python floats; the fleet stores float32 ((double)0.3f is 4e-8 off)."""
    for v in (0.3, 0.5, 0.7):
        assert orc.difficulty_scale(np.float32(v)) == v


def test_hsa_agent_trajectories():
    """Level 2: HSAAgent (heading PID -> bank, TECS energy/balance PIDs -> throttle/pitch) closed loop, 3000 control steps."""
    g = _load("hsa_agent.npz")
    P, Cp = orc.default_aircraft_params(), orc.default_controller_params()
    state = np.ascontiguousarray(g["s0"]); n = state.shape[1]
    hold, steps, dt, every = int(g["hold"]), int(g["steps"]), float(g["dt"]), int(g["every"])
    pid = np.zeros((18, n), np.float32); outer = np.zeros((9, n), np.float32); surf = np.zeros((4, n))
    worst = 0.0
    alive = np.ones(n, bool)
    for t in range(steps):
        orc.step_hsa(P, Cp, state, np.ascontiguousarray(g["cmd"][t // hold]), pid, outer, dt, surfaces_out=surf)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        gs = g["states"][k]
        alive &= ~((-gs[2] < 5.0) | (np.abs(gs[6]) > np.radians(120)) | (np.abs(gs[7]) > np.radians(80)))
        e = max(rel_err(state[:, alive], gs[:, alive]), rel_err(surf[:, alive], g["surfaces"][k][:, alive]))
        worst = max(worst, e)
        assert e <= TOL, f"hsa: step {t + 1} rel err {e:.3e}"
    assert alive.sum() >= n // 2
    print(f"hsa_agent: worst rel err over {steps} steps = {worst:.3e}, {int(alive.sum())}/{n} in envelope")


@pytest.mark.parametrize("name", ["waypoint_pp", "waypoint_los"])
def test_waypoint_mission_trajectories(name):
    """Level 1: MissionPlanner + WaypointAgent + direct physics step (examples/03_waypoint_square_demo.py), 9000 steps."""
    g = _load(name + ".npz")
    P, Cp = orc.default_aircraft_params(), orc.default_controller_params()
    state = np.ascontiguousarray(g["s0"]); n = state.shape[1]
    steps, dt, every = int(g["steps"]), float(g["dt"]), int(g["every"])
    wps = np.ascontiguousarray(g["waypoints"]); idx = np.zeros(n, np.int32)
    pid = np.zeros((18, n), np.float32); outer = np.zeros((9, n), np.float32); surf = np.zeros((4, n)); hc = np.zeros((3, n))
    worst = 0.0
    for t in range(steps):
        orc.step_waypoint(P, Cp, state, wps, idx, pid, outer, dt, guidance=int(g["guidance"]),
                          acceptance_radius=float(g["acceptance_radius"]), dt_physics=dt, surfaces_out=surf, hsa_cmd_out=hc)
        if (t + 1) % every:
            continue
        k = (t + 1) // every - 1
        assert np.array_equal(idx, g["wp_index"][k]), f"{name}: waypoint index at step {t + 1}"
        e = max(rel_err(state, g["states"][k]), rel_err(surf, g["surfaces"][k]))
        worst = max(worst, e)
        assert e <= TOL, f"{name}: step {t + 1} rel err {e:.3e}"
        assert rel_err(hc[0], g["heading_cmd"][k]) <= 1e-9
    assert idx.max() >= 3
    print(f"{name}: worst rel err over {steps} steps = {worst:.3e}; waypoint indices {idx.tolist()}")
