"""Times acb200_sensor_update on 2^20 aircraft (fp32 and fp64 builds) against the measured HBM peak, and the CPU
restatement on a bounded sample.  Prints one JSON line per build."""
import json, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import acb200
from oracle import oracle as orc

n = 1 << 20
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6552.0)
for precision in (32, 64):
    f = acb200.AircraftFleet(n, precision=precision)
    sensor = acb200.NoisySensorB200({"seed": 3}, n, precision=precision)
    for _ in range(5):
        sensor.update(f)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        sensor.update(f)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    word = precision // 8
    bytes_alg = n * (13 + 4 + 6) * 2 * word            # truth + bias in, measured + bias out
    ns = 1 << 16
    cfg = orc.sensor_config({"seed": 3})
    s = np.zeros((13, ns)); d = np.zeros((4, ns)); b = np.zeros((6, ns))
    t0 = time.perf_counter(); reps = 0
    while time.perf_counter() - t0 < 3.0:
        orc.sensor_update(cfg, s, d, b, reps); reps += 1
    cpu = ns * reps / (time.perf_counter() - t0)
    print(json.dumps({"kernel": f"sensor_update_kernel<{'float' if precision == 32 else 'double'}>", "aircraft": n, "ms": ms,
                      "updates_per_s": n / (ms * 1e-3),
                      "roofline": {"bound": "hbm", "achieved": bytes_alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                   "frac": bytes_alg / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": bytes_alg},
                      "cpu_baseline": {"value": cpu, "unit": "updates/s", "cores": 1, "kind": "port",
                                       "sample": f"{ns} aircraft x {reps} updates, C restatement, one thread"}}))
