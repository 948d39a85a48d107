"""acb200_env_step_host (host buffers, chunked three-stream pipeline) must give exactly what acb200_env_step gives, with
pinned and with pageable result buffers."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision,controller,n,chunks,pinned", [(32, "attitude_pid", 100_003, 0, True), (64, "external", 5_001, 7, True),
                                                                 (32, "residual", 777, 3, True), (32, "rate_pid", 64, 0, True),
                                                                 (32, "attitude_pid", 40_000, 0, False)])
def test_host_step_equals_device_step(precision, controller, n, chunks, pinned):
    import torch
    import acb200
    rng = np.random.default_rng(1)
    diff = np.array([0.3, 0.5, 0.7], np.float32)[rng.integers(0, 3, n)]
    dt = torch.float64 if precision == 64 else torch.float32
    att = np.zeros((4, n)); att[0] = rng.uniform(-0.3, 0.3, n); att[1] = rng.uniform(-0.03, 0.08, n); att[2] = np.inf; att[3] = 0.6
    envs = []
    for _ in range(2):
        e = acb200.BatchedRateControlEnv(n, controller=controller, precision=precision, seed=9, episode_length=0.2)
        e.set_curriculum(torch.as_tensor(diff), "step")
        e.reset()
        envs.append(e)
    a, b = envs
    b.ctx.check(b.ctx.lib.acb200_set_host_chunks(b.ctx.handle, chunks))
    pin = lambda shape, d: torch.empty(shape, dtype=d, pin_memory=True)
    res = pin if pinned else (lambda shape, d: torch.full(shape, 77, dtype=d))
    obs_h, rew_h, term_h, trunc_h = res((n, 18), torch.float32), res((n,), torch.float32), res((n,), torch.uint8), res((n,), torch.uint8)
    att_h = pin((4, n), dt)
    act_h = pin((n, 4), torch.float32)
    needs_action = controller in ("external", "residual")
    for t in range(25):          # episodes of 10 steps: the in-kernel auto-reset is exercised twice
        att_t = att.copy(); att_t[0] += 0.01 * t
        att_h.copy_(torch.as_tensor(att_t).to(dt))
        act = (rng.uniform(-1, 1, (n, 4)) * 0.1).astype(np.float32); act[:, 3] = 0.6
        act_h.copy_(torch.as_tensor(act))
        a.att_cmd.copy_(att_h)
        o, r, te, tr = a.step(act_h.cuda() if needs_action else None)
        b.step_host(act_h if needs_action else None, obs_h, rew_h, term_h, trunc_h,
                    att_cmd_host=att_h if controller == "attitude_pid" else None)
        assert torch.equal(o.cpu(), obs_h), t
        assert torch.equal(r.float().cpu(), rew_h), t
        assert torch.equal(te.cpu(), term_h) and torch.equal(tr.cpu(), trunc_h), t
        assert torch.equal(a.state, b.state) and torch.equal(a.pid, b.pid), t
    sa, sb = a.episode_stats(), b.episode_stats()
    assert sa["episodes"] == sb["episodes"] >= 2 * n
    assert abs(sa["return_sum"] - sb["return_sum"]) <= 1e-6 * abs(sa["return_sum"]) + 1e-6    # atomics: order differs
