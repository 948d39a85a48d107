import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import acb200
from oracle import oracle as orc
from tools.bench_c5 import SQUARE, initial_states
n, dt = 1024, 0.01
for K, loop, steps in ((1, True, 1), (1, True, 50), (10, False, 50), (10, False, 400)):
    dtp = dt / K
    s0, idx0 = initial_states(n, 0)
    wps = np.asarray(SQUARE, np.float64)
    P, CP = orc.default_aircraft_params(), orc.default_controller_params()
    state = s0.copy(); idx = idx0.copy()        # copies: the oracle advances them in place
    pid = np.zeros((18, n), np.float32); outer = np.zeros((9, n), np.float32); ref = np.zeros(16)
    for _ in range(steps):
        orc.step_waypoint(P, CP, state, wps, idx, pid, outer, dt, guidance=orc.GUIDANCE_PP, acceptance_radius=40.0, loop=loop, dt_physics=dtp, mission_stats=ref)
    fleet = acb200.AircraftFleet(n, precision=64, config={"dt_physics": dtp})
    fleet.reset(torch.as_tensor(s0)); fleet.set_mission(SQUARE, acceptance_radius=40.0, guidance="PP", loop=loop); fleet.wp_index.copy_(torch.as_tensor(idx0))
    for _ in range(steps):
        fleet.step_waypoint(dt)
    got = fleet.mission_stats_buf.cpu().numpy()
    err = np.abs(fleet.state.cpu().numpy() - state)
    gi = fleet.wp_index.cpu().numpy()
    print(K, loop, steps, "idx equal", np.array_equal(gi, idx), "mismatching idx", int((gi != idx).sum()), "per-field max err", err.max(axis=1).round(6).tolist())
    bad = np.flatnonzero(gi != idx)[:5]
    print("  first mismatches: ids", bad.tolist(), "idx0", idx0[bad].tolist(), "gpu", gi[bad].tolist(), "ref", idx[bad].tolist(), "pos0", s0[:2, bad].round(1).tolist())
    print("  got", got[:10].round(3)); print("  ref", ref[:10].round(3))
