"""Where do the one-aircraft and the pair guided kernels part? (per array, after each of the first steps)"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import acb200

n, dt = 2000, 0.01
g = torch.Generator().manual_seed(8)
init = torch.zeros(13, n)
init[0].uniform_(-20, 20, generator=g); init[1].uniform_(-20, 20, generator=g); init[2] = -100.0
init[3] = 15.0; init[8].uniform_(-0.2, 0.2, generator=g)
wps = [[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0]]
fleets = []
for pair in (False, True):
    f = acb200.AircraftFleet(n, precision=32)
    f.reset(init.cuda())
    f.set_mission(wps, acceptance_radius=12.0, guidance="PP", loop=False)
    fleets.append(f)
names = ("state", "pid", "pid_outer", "controls", "derived", "wp_index")
for step in range(int(sys.argv[1]) if len(sys.argv) > 1 else 12):
    for pair, f in zip((False, True), fleets):
        os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
        f.step_waypoint(dt)
    torch.cuda.synchronize()
    out = []
    for nm in names:
        a, b = getattr(fleets[0], nm).cpu().numpy(), getattr(fleets[1], nm).cpu().numpy()
        bad = (a.view(np.uint32) != b.view(np.uint32)) if a.dtype == np.float32 else (a != b)
        out.append(f"{nm}:{int(bad.sum())}" + (f"(rows {sorted(set(np.nonzero(bad)[0].tolist()))[:8]})" if bad.any() and bad.ndim == 2 else ""))
    print(step, " ".join(out), "finished:", int((fleets[0].wp_index >= 2).sum()))
