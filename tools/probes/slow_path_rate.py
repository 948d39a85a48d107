"""How often would a pair substep that TESTS the model's rarely-binding clamps fall back to the one-aircraft code?
Belongs to profiles/experiments/r02_pair_substep_tests_instead_of_clamps.diff (not part of the shipped kernel): apply that
diff, build with `make EXTRA=-DACB_SLOW_DBG` -- every fallback then adds 1 to the nonfinite_repairs statistic (field 14).
Measured on the C3 workload: 0 in the first 200 steps, 4.3e-4 per pair-substep once episodes start to end in crashes."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import acb200, bench
n, steps, K = 1 << 18, 300, 20
env = acb200.BatchedRateControlEnv(n, controller="attitude_pid", precision=32, seed=42, command_type="step", dt=0.001 * K)
difficulty, att, phase = bench.make_inputs(n, 0)
env.set_curriculum(torch.as_tensor(difficulty), "step")
env.att_cmd.copy_(torch.as_tensor(att).to(env.dtype))
env.reset(); env.step_count.copy_(torch.as_tensor(phase))
for blk in range(3):
    env.stats.zero_()
    for _ in range(steps // 3):
        env.step()
    torch.cuda.synchronize()
    st = env.stats.cpu().numpy()
    print(f"steps {blk * 100}..: fallbacks per pair-substep {st[14] / (n / 2 * (steps // 3) * K):.4e}  episodes {st[0]:.0f} crashes {st[4]:.0f}")
    s = env.state.cpu().numpy()
    print("   u<0.1:", (s[3] < 0.1).mean(), " |theta|>84deg:", (np.abs(s[7]) > np.radians(84)).mean(), " altitude<1:", (-s[2] < 1).mean(),
          " |v|max", np.abs(s[3:6]).max(), " rate max", np.abs(s[9:12]).max())
