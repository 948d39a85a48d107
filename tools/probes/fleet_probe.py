"""Fleet stepping (AircraftFleet.step_surface / step_attitude, fp32) on the pair kernels against the one-aircraft kernels:
ms per step of a 2^20-aircraft fleet at dt = 0.01 (K = 10) and dt = 0.001 (K = 1).  Prints one JSON line."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import acb200

n, steps = 1 << 20, 200
out = {}
for pair in (0, 1):
    os.environ["ACB200_ENV_PAIR"] = str(pair)
    fleet = acb200.AircraftFleet(n, precision=32)
    g = torch.Generator(device="cuda").manual_seed(1)
    surf = torch.empty(4, n, device="cuda").uniform_(-0.2, 0.2, generator=g); surf[3] = 0.6
    att = torch.zeros(4, n, device="cuda"); att[0].uniform_(-0.3, 0.3, generator=g); att[2] = float("inf"); att[3] = 0.6
    hsa = torch.zeros(3, n, device="cuda"); hsa[0].uniform_(-1, 1, generator=g); hsa[1] = 15.0; hsa[2] = 100.0
    for name, fn in (("hsa_K10", lambda: fleet.step_hsa(hsa, 0.01)), ("hsa_K1", lambda: fleet.step_hsa(hsa, 0.001)), ("surface_K10", lambda: fleet.step_surface(surf, 0.01)), ("attitude_K10", lambda: fleet.step_attitude(att, 0.01)),
                     ("surface_K1", lambda: fleet.step_surface(surf, 0.001)), ("attitude_K1", lambda: fleet.step_attitude(att, 0.001))):
        fleet.reset()
        for _ in range(5):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(steps):
            fn()
        e1.record(); torch.cuda.synchronize()
        out[f"{name}_{'pair' if pair else 'one'}_ms"] = e0.elapsed_time(e1) / steps
os.environ.pop("ACB200_ENV_PAIR", None)
out["aircraft"] = n
print(json.dumps(out))
