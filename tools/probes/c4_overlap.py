"""C4 device step with the env step of one part of the fleet running beside the policy forward of the next part (two streams).
Needs the policy built with a 2-stage weight ring (ACB_POLICY_STAGES=2) so that one env block fits beside a policy block."""
import ctypes as C, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import acb200
from acb200.policy import random_policy_state_dict

n, steps = 262144, 200
parts = int(sys.argv[1]) if len(sys.argv) > 1 else 2
env = acb200.BatchedRateControlEnv(n, difficulty="medium", controller="external", precision=32, seed=42)
pol = acb200.LearnedPolicyB200(random_policy_state_dict("lstm", 0), kind="lstm", n=n, ctx=env.ctx)
env.reset()
flags = (env.terminated, env.truncated)
per = n // parts
s_pol, s_env = torch.cuda.Stream(), torch.cuda.Stream()
env_done = [None] * parts

def step_seq():
    env.step(pol(env.obs, flags))

def step_overlap():
    for k in range(parts):
        lo = k * per
        with torch.cuda.stream(s_pol):
            if env_done[k] is not None:
                s_pol.wait_event(env_done[k])          # this part's previous env step
            pol.forward_range(env.obs, flags, lo, per)
            ready = s_pol.record_event()
        with torch.cuda.stream(s_env):
            s_env.wait_event(ready)
            env._bind_stream()
            env.ctx.check(env.ctx.lib.acb200_env_step_range(env.ctx.handle, C.byref(env.c), C.byref(env.cfg),
                                                            C.c_void_p(pol.action.data_ptr()), lo, per))
            env_done[k] = s_env.record_event()

out = {}
for name, fn in (("sequential", step_seq), ("overlap", step_overlap)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    if name == "overlap":
        torch.cuda.current_stream().wait_stream(s_pol); torch.cuda.current_stream().wait_stream(s_env)
    e1.record(); torch.cuda.synchronize()
    out[name + "_ms"] = e0.elapsed_time(e1) / steps
out["parts"] = parts
print(json.dumps(out))
