"""Times acb200_gae on a C4-sized rollout (T=128 steps x 262144 environments) against the measured HBM peak, and the numpy
host implementation it replaces on a bounded sample.  Prints one JSON line."""
import json, os, sys, time
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import acb200
from oracle import rollout_oracle as ro

T, n = 128, 262144
buf = acb200.RolloutBufferB200(T, n)
buf.rewards.normal_(); buf.values.normal_(); buf.episode_starts.copy_((torch.rand(T, n, device="cuda") < 0.02).to(torch.uint8))
last = torch.randn(n, device="cuda"); dones = torch.zeros(n, dtype=torch.uint8, device="cuda")
for _ in range(3):
    buf.compute_returns_and_advantage(last, dones)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    buf.compute_returns_and_advantage(last, dones)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
bytes_alg = T * n * (4 + 4 + 1 + 4 + 4) + n * 5
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6552.0)
ns = 4096
r = buf.rewards[:, :ns].cpu().numpy(); v = buf.values[:, :ns].cpu().numpy(); s = buf.episode_starts[:, :ns].float().cpu().numpy()
t0 = time.perf_counter(); reps = 0
while time.perf_counter() - t0 < 5.0:
    ro.compute_returns_and_advantage(r, v, s, last[:ns].cpu().numpy(), np.zeros(ns, bool)); reps += 1
cpu = T * ns * reps / (time.perf_counter() - t0)
print(json.dumps({"kernel": "gae_kernel", "T": T, "n_envs": n, "ms": ms, "samples_per_s": T * n / (ms * 1e-3),
                  "roofline": {"bound": "hbm", "achieved": bytes_alg / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                               "frac": bytes_alg / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_launch": bytes_alg},
                  "cpu_baseline": {"value": cpu, "unit": "samples/s", "cores": 1, "kind": "port",
                                   "sample": f"numpy float32 loop, {T} x {ns} samples x {reps} repetitions"}}))
