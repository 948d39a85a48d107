"""C4 rollout: timing per `parts`, and the CUDA-graph replay checked stage by stage."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import acb200
from acb200.policy import random_policy_state_dict

n = int(os.environ.get("N", "262144"))
sd = random_policy_state_dict("lstm", 0)
def make():
    env = acb200.BatchedRateControlEnv(n, difficulty="medium", controller="external", precision=32, seed=42)
    pol = acb200.LearnedPolicyB200(sd, kind="lstm", n=n, ctx=env.ctx)
    env.reset()
    return env, pol
env, pol = make()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for parts in (1, 2, 3, 2):
    pol.rollout(env, 5, parts=parts); torch.cuda.synchronize()
    ev0.record(); pol.rollout(env, 100, parts=parts); ev1.record(); torch.cuda.synchronize()
    print(f"parts={parts}: {ev0.elapsed_time(ev1) / 100:.4f} ms/step", flush=True)
# graph replay
ea, pa = make(); ec, pc = make()
print("ec state nonzero after reset:", bool(ec.state.abs().sum() > 0))
for _ in range(30):
    ea.step(pa(ea.obs, (ea.terminated, ea.truncated)))
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    pc.rollout(ec, 10, parts=2)
s.synchronize()
print("after warm-up on side stream: nonzero", bool(ec.state.abs().sum() > 0), "step_count", int(ec.step_count[0]))
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=s):
    pc.rollout(ec, 10, parts=2)
print("after capture: nonzero", bool(ec.state.abs().sum() > 0), "step_count", int(ec.step_count[0]))
g.replay(); g.replay(); torch.cuda.synchronize()
print("after 2 replays: nonzero", bool(ec.state.abs().sum() > 0), "step_count", int(ec.step_count[0]), "equal to loop:", bool(torch.equal(ea.state, ec.state)))
