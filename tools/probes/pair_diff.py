"""Debug aid: first step at which the pair kernel and the one-aircraft kernel disagree, and the state before it."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import acb200
from test_gpu_pair_kernel import _make

n, seed = 2048 + 74, 5
envs = []
for pair in (False, True):
    e = _make(acb200, n, "attitude_pid", seed, True)
    envs.append(e)
def step(e, pair):
    os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
    e.step()
for pair, e in zip((False, True), envs):
    os.environ["ACB200_ENV_PAIR"] = "1" if pair else "0"
    e.reset()
np.set_printoptions(precision=9, linewidth=200)
for t in range(130):
    before = envs[0].state.cpu().numpy().copy()
    ctrl_before = envs[0].controls.cpu().numpy().copy()
    step(envs[0], False); step(envs[1], True)
    a, b = envs[0].state.cpu().numpy(), envs[1].state.cpu().numpy()
    bad = np.unique(np.nonzero(a.view(np.uint32) != b.view(np.uint32))[1])
    if len(bad):
        print("step", t, "aircraft", bad[:10], "count", len(bad))
        for i in bad[:3]:
            print(" before ", before[:, i]); print(" scalar ", a[:, i]); print(" pair   ", b[:, i])
            print(" ulps   ", (a[:, i].view(np.int32) - b[:, i].view(np.int32)))
            print(" controls(after)", envs[0].controls.cpu().numpy()[:, i], envs[1].controls.cpu().numpy()[:, i], "step_count", envs[0].step_count[i].item())
        break
else:
    print("no difference in 130 steps")
