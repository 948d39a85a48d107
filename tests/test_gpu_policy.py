"""Row a20: acb200_policy_forward (csrc/policy.cu, tcgen05 + tensor memory) against the torch restatement of the
reference's networks (oracle/policy_oracle.py; learned_controllers/networks/lstm_policy.py:13-164).

Two bars, both stated here:
  * vs the bf16-EMULATED oracle (same operand rounding as the kernel): EMU_TOL.  What is left is fp32 accumulation
    order and tanh.approx (2^-11 relative) in the gates -- a wrong descriptor, gate order or state layout fails this.
  * vs the plain fp32 oracle: F32_TOL on the clipped action, the cost of bf16 operands.
"""
import numpy as np
import pytest
import torch

from oracle import policy_oracle as po

pytestmark = pytest.mark.gpu

EMU_TOL = 4e-3      # LSTM policy: max |action - emulated| and max |h,c - emulated|   (measured 1e-4 / 5e-4, r01)
F32_TOL = 5e-3      # LSTM policy: max |action - fp32 oracle|                         (measured 1.5e-4, r01)
# The MLP policy has no gates: the raw observation (altitude ~100 m, airspeed ~20 m/s) drives hidden activations to
# O(30), and ONE bf16 rounding of such an activation is 0.06 -- an fp32-accumulation-order difference of one ulp that
# flips a rounding already moves the action by ~3e-3.  Its bars are therefore wider (measured 2.6e-3 / 1.0e-2, r01).
# r02: both bars narrowed to what SURVEY.md 8c asks for bf16 ("<= 2e-2 abs on actions").  A CPU study with the emulated
# forward (hi+lo split of the first hidden layer's activations and of its weights as well) showed the 1e-2 is spread over
# EVERY layer of the un-normalised net (relative bf16 precision 2^-9 of pre-clip actions of a few units), not the first one:
# the split moved 1.1e-2 to 1.0e-2, so it was not built.
EMU_TOL_MLP = 8e-3
F32_TOL_MLP = 2e-2


def _policy(kind, n, seed=0, scale=1.0):
    import acb200
    net = po.make_policy(kind, seed=seed, scale=scale)
    return net, acb200.LearnedPolicyB200(net.state_dict(), kind=kind, n=n)


@pytest.mark.parametrize("n", [1, 128, 300, 4096])
def test_mlp_policy_vs_oracle(n):
    net, pol = _policy("mlp", n)
    obs = po.synthetic_obs(n, seed=n)
    act = pol(obs.cuda()).cpu()
    torch.cuda.synchronize()
    emu, _ = po.forward_bf16_emulated(net, obs)
    ref, _ = net(obs)
    assert torch.isfinite(act).all()
    assert (act - emu).abs().max().item() < EMU_TOL_MLP
    assert (act - ref).abs().max().item() < F32_TOL_MLP


@pytest.mark.parametrize("n,scale", [(1, 1.0), (300, 1.0), (2048, 1.0), (640, 2.5)])
def test_lstm_policy_sequence_vs_oracle(n, scale):
    """20 steps with carried state; episode_start raised at random (and for everyone at step 0, as SB3 does)."""
    net, pol = _policy("lstm", n, scale=scale)
    rng = np.random.default_rng(7)
    st_ref = net.initial_state(n)
    st_emu = net.initial_state(n)
    worst_emu = worst_ref = worst_state = 0.0
    for t in range(20):
        obs = po.synthetic_obs(n, seed=1000 + t)
        es = torch.from_numpy(rng.random(n) < (1.0 if t == 0 else 0.1))
        act = pol(obs.cuda(), es.cuda()).cpu()
        emu, st_emu = po.forward_bf16_emulated(net, obs, st_emu, es)
        ref, st_ref = net(obs, st_ref, es)
        h, c = pol.get_lstm_states()
        worst_emu = max(worst_emu, (act - emu).abs().max().item())
        worst_ref = max(worst_ref, (act - ref).abs().max().item())
        worst_state = max(worst_state, np.abs(h[0] - st_emu[0].numpy()).max(), np.abs(c[0] - st_emu[1].numpy()).max())
        # keep the emulation on the device's own state so that one step's rounding does not compound into the bar
        st_emu = (torch.from_numpy(h[0].copy()), torch.from_numpy(c[0].copy()))
    assert worst_emu < EMU_TOL * max(1.0, scale), worst_emu
    assert worst_state < 2 * EMU_TOL * max(1.0, scale), worst_state
    if scale == 1.0:
        assert worst_ref < F32_TOL, worst_ref


def test_sm_pair_kernel_matches_single_block_kernel(monkeypatch):
    """The cta_group::2 variant (ACB200_POLICY_CTA=2: one MMA stream for an SM pair, weight halves loaded through tensor
    maps onto the leader's barrier) computes the same sums in the same order: actions and state bit for bit."""
    import acb200
    n = 900                                            # 8 tiles: the last pair is half padding
    net = po.make_policy("lstm", seed=1)
    out = {}
    for cta in ("1", "2"):
        monkeypatch.setenv("ACB200_POLICY_CTA", cta)
        pol = acb200.LearnedPolicyB200(net.state_dict(), kind="lstm", n=n)
        acts = []
        for t in range(4):
            obs = po.synthetic_obs(n, seed=50 + t).cuda()
            es = torch.full((n,), t == 0, dtype=torch.bool, device="cuda")
            acts.append(pol(obs, es).clone())
        out[cta] = (torch.stack(acts), *pol.get_lstm_states())
    assert torch.equal(out["1"][0], out["2"][0])
    assert np.array_equal(out["1"][1], out["2"][1]) and np.array_equal(out["1"][2], out["2"][2])
    emu, _ = po.forward_bf16_emulated(net, po.synthetic_obs(n, seed=50), net.initial_state(n), torch.ones(n, dtype=torch.bool))
    assert (out["2"][0][0].cpu() - emu).abs().max().item() < EMU_TOL


def test_state_import_export_roundtrip():
    n = 300
    net, pol = _policy("lstm", n)
    rng = np.random.default_rng(3)
    h = rng.uniform(-1, 1, (1, n, 256)).astype(np.float32)
    c = rng.uniform(-3, 3, (1, n, 256)).astype(np.float32)
    pol.set_lstm_states(h, c)
    h2, c2 = pol.get_lstm_states()
    assert np.array_equal(c2, c)                                                     # c travels as fp32
    assert np.array_equal(h2, torch.from_numpy(h).bfloat16().float().numpy())        # h as bf16


def test_predict_signature_matches_sb3():
    net, pol = _policy("lstm", 4)
    obs = po.synthetic_obs(4, seed=5).numpy()
    a1, st = pol.predict(obs, state=None, episode_start=np.ones(4, bool), deterministic=True)
    assert a1.shape == (4, 4) and st[0].shape == (1, 4, 256) and st[1].shape == (1, 4, 256)
    a2, _ = pol.predict(obs, state=st, episode_start=np.zeros(4, bool))
    ref1, s = net(torch.from_numpy(obs), net.initial_state(4), torch.ones(4, dtype=torch.bool))
    ref2, _ = net(torch.from_numpy(obs), s, torch.zeros(4, dtype=torch.bool))
    assert np.abs(a1 - ref1.numpy()).max() < F32_TOL and np.abs(a2 - ref2.numpy()).max() < F32_TOL
    assert (a1[:, :3] >= -1).all() and (a1 <= 1).all() and (a1[:, 3] >= 0).all()


def test_full_size_tile_invariance():
    """BASELINE C4 size (262 144 aircraft): every aircraft's action depends on its own row only -- the first 1024
    rows repeated through the fleet must reproduce the 1024-aircraft run bit for bit, on every tile and SM."""
    n, base = 262144, 1024
    net, small = _policy("lstm", base)
    obs = po.synthetic_obs(base, seed=11).cuda()
    es = torch.ones(base, dtype=torch.uint8, device="cuda")
    a_small = [small(obs, es).clone(), small(obs).clone()]
    import acb200
    big = acb200.LearnedPolicyB200(net.state_dict(), kind="lstm", n=n)
    obs_big = obs.repeat(n // base, 1).contiguous()
    a_big = [big(obs_big, es.repeat(n // base)).clone(), big(obs_big).clone()]
    torch.cuda.synchronize()
    for s, b in zip(a_small, a_big):
        assert torch.equal(b.view(n // base, base, 4), s.unsqueeze(0).expand(n // base, base, 4))


def test_argument_errors():
    net, pol = _policy("lstm", 8)
    with pytest.raises(ValueError):
        pol(torch.zeros(8, 17, device="cuda"))
    with pytest.raises(ValueError):
        pol(torch.zeros(8, 18))
    import acb200
    with pytest.raises(KeyError):
        acb200.LearnedPolicyB200({}, kind="lstm", n=8)
    with pytest.raises(ValueError):
        acb200.LearnedPolicyB200(net.state_dict(), kind="gru", n=8)


def test_policy_env_closed_loop_vs_oracle():
    """C4 in small: the device policy flies the device env; the torch oracle policy flies the C oracle env, both
    from the same Philox resets.  The loop is a feedback system, so the bar is per step on re-synchronised inputs:
    every step the device action is compared with the oracle policy's action ON THE DEVICE'S OWN observation and
    state (LSTM bar), and the fleet keeps flying (finite observations, episodes end and restart)."""
    import acb200
    n = 512
    net = po.make_policy("lstm", seed=0)
    env = acb200.BatchedRateControlEnv(n, difficulty="medium", controller="external", precision=64, seed=42, episode_length=1.0)
    pol = acb200.LearnedPolicyB200(net.state_dict(), kind="lstm", n=n, ctx=env.ctx)
    env.reset()
    worst, restarts = 0.0, 0
    for t in range(120):
        h, c = pol.get_lstm_states()
        obs = env.obs.clone()
        flags = (env.terminated.clone(), env.truncated.clone())
        act = pol(obs, (env.terminated, env.truncated)).clone()
        es = (flags[0] | flags[1]).bool().cpu()
        ref, _ = net(obs.cpu(), (torch.from_numpy(h[0].copy()), torch.from_numpy(c[0].copy())), es)
        worst = max(worst, (act.cpu() - ref).abs().max().item())
        env.step(act)
        restarts += int((env.terminated | env.truncated).sum().item())
        assert torch.isfinite(env.obs).all()
    assert worst < F32_TOL, worst
    assert restarts >= n          # 1 s episodes: everyone restarted at least once, the recurrent state was re-zeroed


def test_pipelined_host_step_equals_the_plain_closed_loop():
    """policy.step_host_pipelined (H2D / policy + env / D2H over parts of the fleet on three streams) against the plain
    sequence on one stream: observations, rewards, flags and the recurrent state bit for bit over 12 closed-loop steps."""
    import acb200
    from acb200.policy import random_policy_state_dict, step_host_pipelined
    n = 3000                                     # ragged: chunks of 768 aircraft, a 696-aircraft tail
    outs = []
    for piped in (False, True):
        env = acb200.BatchedRateControlEnv(n, difficulty="medium", controller="external", precision=32, seed=7,
                                           episode_length=0.2)
        pol = acb200.LearnedPolicyB200(random_policy_state_dict("lstm", 0), kind="lstm", n=n, ctx=env.ctx)
        env.reset()
        pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
        obs_h, rew_h = pin((n, 18), torch.float32), pin((n,), torch.float32)
        term_h, trunc_h = pin((n,), torch.uint8), pin((n,), torch.uint8)
        obs_h.copy_(env.obs); obs_d = torch.empty_like(env.obs)
        hist = []
        for t in range(12):
            if piped:
                step_host_pipelined(env, pol, obs_h, rew_h, term_h, trunc_h, obs_d, chunks=4)
            else:
                obs_d.copy_(obs_h, non_blocking=True)
                env.step(pol(obs_d, (env.terminated, env.truncated)))
                obs_h.copy_(env.obs); rew_h.copy_(env.reward); term_h.copy_(env.terminated); trunc_h.copy_(env.truncated)
                torch.cuda.synchronize()
            hist.append([x.clone() for x in (obs_h, rew_h, term_h, trunc_h)])
        torch.cuda.synchronize()
        outs.append((hist, pol.h_state.cpu().clone(), pol.c_state.cpu().clone(), env.state.cpu().clone()))
    for a, b in zip(outs[0][0], outs[1][0]):
        for x, y in zip(a, b):
            assert torch.equal(x.view(torch.uint8), y.view(torch.uint8))
    assert torch.equal(outs[0][1], outs[1][1]) and torch.equal(outs[0][2].view(torch.int32), outs[1][2].view(torch.int32))
    assert torch.equal(outs[0][3].view(torch.int32), outs[1][3].view(torch.int32))
    assert sum(int(h[2].sum() + h[3].sum()) for h in outs[1][0]) >= n          # episodes ended and restarted on the way


@pytest.mark.parametrize("kind,n,parts", [("lstm", 1024, 2), ("lstm", 1300, 3), ("mlp", 2048, 2), ("lstm", 512, 1)])
def test_rollout_call_equals_the_step_by_step_closed_loop(kind, n, parts):
    """acb200_policy_env_rollout (K closed-loop steps in one call, the fleet in `parts` ranges on their own streams, the
    2-slot-ring policy kernel beside the env kernel) against forward + env.step in a Python loop: every buffer bit-identical,
    through auto-resets; and the same rollout replayed from a CUDA graph."""
    import acb200
    from acb200.policy import random_policy_state_dict
    sd = random_policy_state_dict(kind, 0)
    envs, pols = [], []
    for _ in range(3):
        env = acb200.BatchedRateControlEnv(n, controller="external", precision=32, seed=13, episode_length=0.3)     # 15-step episodes
        pol = acb200.LearnedPolicyB200(sd, kind=kind, n=n, ctx=env.ctx)
        env.reset()
        envs.append(env); pols.append(pol)
    (ea, pa), (eb, pb), (ec, pc) = zip(envs, pols)
    steps = 40
    for _ in range(steps):
        ea.step(pa(ea.obs, (ea.terminated, ea.truncated)))
    pb.rollout(eb, steps, parts=parts)
    torch.cuda.synchronize()
    names = ("state", "obs", "reward", "terminated", "truncated", "prev_action", "rew_mem", "step_count", "episode_idx")
    for name in names:
        assert torch.equal(getattr(ea, name), getattr(eb, name)), name
    assert torch.equal(pa.action, pb.action)
    if kind == "lstm":
        assert torch.equal(pa.h_state, pb.h_state) and torch.equal(pa.c_state, pb.c_state)
    assert float(ea.stats[0]) > 0                                     # episodes ended on the way
    # CUDA graph: 10 steps in a warm-up call, 10 steps captured (a capture records, it does not run), replayed 3 times = 40
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        pc.rollout(ec, 10, parts=parts)
    s.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        pc.rollout(ec, 10, parts=parts)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for name in names:
        assert torch.equal(getattr(ea, name), getattr(ec, name)), "graph: " + name


def test_rollout_argument_errors():
    import acb200
    from acb200.policy import random_policy_state_dict
    env = acb200.BatchedRateControlEnv(512, controller="rate_pid", precision=32)
    pol = acb200.LearnedPolicyB200(random_policy_state_dict("mlp", 0), kind="mlp", n=512, ctx=env.ctx)
    env.reset()
    with pytest.raises(ValueError, match="external or residual"):
        pol.rollout(env, 2)
    other = acb200.BatchedRateControlEnv(512, controller="external", precision=32)
    with pytest.raises(ValueError, match="share one context"):
        pol.rollout(other, 2)
    ext = acb200.BatchedRateControlEnv(512, controller="external", precision=32, ctx=env.ctx)
    ext.reset()
    with pytest.raises(ValueError):
        pol.rollout(ext, 2, parts=7)
    pol.rollout(ext, 0)                                                # zero steps: a no-op
