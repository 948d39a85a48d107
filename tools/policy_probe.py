"""GPU probe for the policy-forward kernel: prints error against the bf16-emulated and fp32 torch oracles and a
first timing.  Development aid (tests/test_gpu_policy.py holds the bars)."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from oracle import policy_oracle as po
import acb200

def run(kind, n, scale=1.0, steps=3):
    net = po.make_policy(kind, scale=scale)
    pol = acb200.LearnedPolicyB200(net.state_dict(), kind=kind, n=n)
    st = net.initial_state(n)
    st_e = net.initial_state(n)
    for t in range(steps):
        obs = po.synthetic_obs(n, seed=t)
        es = torch.zeros(n, dtype=torch.bool) if t else torch.ones(n, dtype=torch.bool)
        act = pol(obs.cuda(), es.cuda()).cpu()
        torch.cuda.synchronize()
        ref, st = net(obs, st, es)
        emu, st_e = po.forward_bf16_emulated(net, obs, st_e, es)
        msg = f"{kind} n={n} scale={scale} t={t}: |act-emu|={float((act-emu).abs().max()):.3e} |act-f32|={float((act-ref).abs().max()):.3e} |act|={float(act.abs().mean()):.3f}"
        if kind == "lstm":
            h, c = pol.get_lstm_states()
            msg += f" |h-emu|={np.abs(h[0]-st_e[0].numpy()).max():.3e} |c-emu|={np.abs(c[0]-st_e[1].numpy()).max():.3e}"
        print(msg, flush=True)
    if not torch.isfinite(act).all():
        print("NON-FINITE", act[:4])
    return pol

def bench(kind, n, iters=20):
    net = po.make_policy(kind)
    pol = acb200.LearnedPolicyB200(net.state_dict(), kind=kind, n=n)
    obs = po.synthetic_obs(4096, 0).cuda().repeat(n // 4096, 1).contiguous()
    for _ in range(3):
        pol(obs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        pol(obs)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = pol.info()["flops_per_aircraft"] * n
    print(f"bench {kind} n={n}: {ms:.3f} ms/forward, {n/ms*1e3:.3e} aircraft/s, {fl/ms/1e9:.1f} TFLOP/s", flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "bench":
        bench("lstm", 262144, iters=int(sys.argv[2]) if len(sys.argv) > 2 else 5)
        sys.exit(0)
    run("mlp", 128); run("mlp", 300)
    run("lstm", 128); run("lstm", 300, steps=5); run("lstm", 640, scale=2.5, steps=5)
    bench("mlp", 262144); bench("lstm", 262144)
