"""Measures the error distribution of the fp32 throughput build against the fp64 oracle (GPU box).
Output feeds the tolerances stated in tests/test_gpu_parity_f32.py and DESIGN.md."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import acb200
from oracle import oracle as orc
from helpers import in_envelope, rel_err_cols, sample_initial_states, surface_sequence

Q = [50, 90, 99, 99.9, 99.99, 100]
out = {}

def quant(x):
    return {str(q): float(np.percentile(x, q)) for q in Q}

# per-step, re-synchronised
for dt, dtp in ((0.01, 0.01), (0.01, 0.001), (0.02, 0.001)):
    n, steps, hold = 4096, 120, 30
    rng = np.random.default_rng(99)
    s0 = sample_initial_states(rng, n); ctrl = surface_sequence(rng, n, steps // hold)
    P = orc.default_aircraft_params()
    fleet = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp})
    ref = s0.copy(); rder = np.zeros((4, n)); errs = []; field_worst = np.zeros(12); der_worst = np.zeros(4)
    for t in range(steps):
        c = np.ascontiguousarray(ctrl[t // hold])
        start32 = ref.astype(np.float32)
        fleet.reset(torch.as_tensor(start32, device="cuda"))
        fleet.step_surface(torch.as_tensor(c, device="cuda", dtype=torch.float32), dt)
        ref = start32.astype(np.float64)
        orc.step_surface(P, ref, c.astype(np.float32).astype(np.float64), dt, dtp, derived_out=rder)
        alive = in_envelope(ref, rder)
        got = fleet.state.double().cpu().numpy()
        ef = np.abs(got[:12] - ref[:12]) / np.maximum(np.abs(ref[:12]), 1.0)
        field_worst = np.maximum(field_worst, ef[:, alive].max(axis=1))
        errs.append(ef.max(axis=0)[alive])
        gder = fleet.derived.double().cpu().numpy()
        ed = np.abs(gder[:3] - rder[:3]) / np.maximum(np.abs(rder[:3]), 1.0)
        dh = np.abs(np.angle(np.exp(1j * (gder[3] - rder[3]))))
        der_worst = np.maximum(der_worst, np.concatenate([ed[:, alive].max(axis=1), [dh[alive].max()]]))
    out[f"per_step dt={dt} dtp={dtp}"] = {"q": quant(np.concatenate(errs)), "field_worst": field_worst.tolist(),
                                          "derived_worst(airspeed,altitude,ground_speed,heading_abs)": der_worst.tolist()}

# closed loop trajectories
for controller in ("rate_pid", "attitude_pid"):
    n, steps = 8192, 100
    rng = np.random.default_rng(5)
    diff = np.array([0.3, 0.5, 0.7], np.float32)[rng.integers(0, 3, n)]
    env = acb200.BatchedRateControlEnv(n, controller=controller, precision=32, seed=11, auto_reset=False)
    env.set_curriculum(torch.as_tensor(diff), "step")
    ref = orc.Env(n, controller=acb200.envs.CONTROLLERS[controller], seed=11, difficulty=diff, auto_reset=False)
    att = np.zeros((4, n)); att[0] = rng.uniform(-0.3, 0.3, n); att[1] = rng.uniform(-0.03, 0.08, n); att[2] = np.inf; att[3] = 0.6
    env.att_cmd.copy_(torch.as_tensor(att, dtype=torch.float32)); ref.att_cmd[:] = att.astype(np.float32)
    env.reset(); ref.reset()
    # sensitivity probe: the oracle itself from a start one float32-ulp away in every state
    pert = orc.Env(n, controller=acb200.envs.CONTROLLERS[controller], seed=11, difficulty=diff, auto_reset=False)
    pert.att_cmd[:] = att.astype(np.float32); pert.reset()
    s32 = pert.state[:12].astype(np.float32)
    pert.state[:12] = np.nextafter(s32, np.float32(np.inf)).astype(np.float64)
    alive = np.ones(n, bool); traj = {}
    for t in range(steps):
        obs, rew, term, trunc = env.step(); ref.step(); pert.step()
        alive &= in_envelope(ref.state, ref.derived) & ~ref.terminated.astype(bool)
        if (t + 1) in (1, 5, 10, 25, 50, 100):
            e = rel_err_cols(env.state.double().cpu().numpy()[:12], ref.state[:12])
            pid_same = np.all(env.pid.cpu().numpy().view(np.uint32) == ref.pid.view(np.uint32), axis=0)
            re = np.abs(rew.double().cpu().numpy() - ref.reward)
            ep = rel_err_cols(pert.state[:12], ref.state[:12])
            traj[t + 1] = {"alive": float(alive.mean()), "state": quant(e[alive]), "oracle_ulp32_sensitivity": quant(ep[alive]), "reward_abs": quant(re[alive]),
                           "pid_bit_identical": float(pid_same[alive].mean()),
                           "worst_by_difficulty": {str(d): float(e[alive & (diff == np.float32(d))].max()) for d in (0.3, 0.5, 0.7)}}
    out[f"traj {controller}"] = traj
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "f32_tolerance_probe.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
