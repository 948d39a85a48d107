import sys, os
import numpy as np
ROOT="/root/repo"; sys.path.insert(0, ROOT); sys.path.insert(0, ROOT+"/tests")
import torch, acb200
from oracle import oracle as orc
from helpers import in_envelope, sample_initial_states, surface_sequence
np.set_printoptions(precision=6, linewidth=200)
dt,dtp=0.02,0.001
n, steps, hold = 4096, 120, 30
rng = np.random.default_rng(99)
s0 = sample_initial_states(rng, n); ctrl = surface_sequence(rng, n, steps // hold)
P = orc.default_aircraft_params()
fleet = acb200.AircraftFleet(n, precision=32, config={"dt_physics": dtp})
ref = s0.copy(); rder = np.zeros((4, n))
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
    e = ef.max(axis=0) * alive
    i = int(np.argmax(e))
    if e[i] > 2.5e-6:
        print("step", t+1, "aircraft", i, "err", e[i])
        print(" start", start32[:12, i]); print(" ctrl ", c[:, i]); print(" ref  ", ref[:12, i]); print(" got  ", got[:12, i]); print(" relerr", ef[:, i])
        V=np.linalg.norm(start32[3:6,i]); print(" V", V, "alpha", np.degrees(np.arctan2(start32[5,i], start32[3,i])), "beta", np.degrees(np.arcsin(start32[4,i]/max(V,10))))
