"""BASELINE.json's full size (C3: 2^20 aircraft on one GPU) through size-independent properties.  The oracle cannot run a
million aircraft in a test, but aircraft are independent and every random draw is keyed by the GLOBAL aircraft id, so any slice
of the big fleet must equal, bit for bit, a small fleet built for the same ids -- and the small fleets are what the oracle
parity tests (test_gpu_parity_f32.py / _f64.py, golden fixtures) cover.  Plus bookkeeping identities over the whole fleet:
episode counters against the flags the step returned, no non-finite repair, statistics against a recount on the host."""
import pytest
import torch

pytestmark = pytest.mark.gpu

N = 1 << 20
STEPS = 45                       # 20-step episodes: every aircraft resets twice inside the kernel
TABLE = torch.tensor([0.3, 0.5, 0.7])


def _build(acb200, n, id_offset, controller, precision=32):
    env = acb200.BatchedRateControlEnv(n, controller=controller, precision=precision, seed=42, episode_length=0.4, id_offset=id_offset)
    ids = torch.arange(id_offset, id_offset + n)
    env.set_curriculum(TABLE[ids % 3], (ids % 4).to(torch.uint8))
    att = torch.zeros(4, n)
    att[0] = ((ids % 41).float() / 40.0 - 0.5) * 0.7          # roll command  +-20 deg, by global id
    att[1] = ((ids % 29).float() / 28.0 - 0.5) * 0.3
    att[2] = float("inf")
    att[3] = 0.5 + (ids % 7).float() * 0.05
    env.att_cmd.copy_(att.to(env.dtype))
    env.reset()
    return env


@pytest.mark.parametrize("controller,precision,N", [("attitude_pid", 32, N), ("rate_pid", 32, N), ("attitude_pid", 64, N // 4)])
def test_slices_of_the_full_fleet_equal_small_fleets_with_the_same_ids(controller, precision, N):
    """fp32 throughput build at C3's 2^20 aircraft; fp64 parity build at the 2^18 it is benchmarked at."""
    import acb200
    big = _build(acb200, N, 0, controller, precision)
    slices = [(0, 4096), (N // 2 - 128, 2050), (N - 1024, 1024)]          # first block, across the middle, the last blocks
    small = [_build(acb200, cnt, lo, controller, precision) for lo, cnt in slices]
    flags = torch.zeros((), dtype=torch.int64, device="cuda")
    crashes = torch.zeros((), dtype=torch.int64, device="cuda")
    for _ in range(STEPS):
        _, _, term, trunc = big.step()
        flags += ((term | trunc) != 0).sum()
        crashes += ((term != 0) & (trunc == 0)).sum()
        for e in small:
            e.step()
    for (lo, cnt), e in zip(slices, small):
        for name in ("state", "pid", "rate_cmd", "rew_mem", "sched", "controls", "derived"):
            assert torch.equal(getattr(e, name), getattr(big, name)[:, lo:lo + cnt]), (name, lo)
        for name in ("obs", "reward", "terminated", "truncated", "episode_idx", "step_count", "ep_return", "env_time"):
            assert torch.equal(getattr(e, name), getattr(big, name)[lo:lo + cnt]), (name, lo)
    stats = big.episode_stats()
    assert stats["episodes"] == int(flags) >= 2 * N                       # every aircraft finished at least two episodes
    assert stats["crashes"] == int(crashes)
    assert big.stats[14].item() == 0                                       # no non-finite state was repaired
    assert big.stats[15].item() == STEPS * N                               # every aircraft stepped every time
    assert torch.isfinite(big.state).all() and torch.isfinite(big.obs).all() and torch.isfinite(big.reward).all()
    # the physical envelope the reference clamps to (simplified_6dof.py:258-291)
    assert float(big.state[3:6].abs().max()) <= 100.0 and float(big.state[7].abs().max()) <= 1.4836 and float(big.state[2].max()) <= 0.0
    assert int(big.step_count.max()) < 20 and int(big.episode_idx.min()) >= 2


def test_full_fleet_pair_and_one_aircraft_kernels_agree(monkeypatch):
    """The packed two-aircraft kernel against the one-aircraft kernel at the full size (the small-size bit-identity is
    test_gpu_pair_kernel.py): checksums of every state word after 25 steps."""
    import acb200
    sums = []
    for pair in ("1", "0"):
        monkeypatch.setenv("ACB200_ENV_PAIR", pair)
        env = _build(acb200, N, 0, "attitude_pid")
        for _ in range(25):
            env.step()
        words = torch.cat([env.state.view(torch.int32).flatten(), env.pid.view(torch.int32).flatten(),
                           env.obs.view(torch.int32).flatten(), env.reward.view(torch.int32).flatten()]).to(torch.int64)
        weights = torch.arange(words.numel(), device="cuda", dtype=torch.int64) % 8191 + 1
        sums.append((int(words.sum()), int((words * weights).sum())))
        del env
    assert sums[0] == sums[1]


def test_field_offsets_beyond_4_GiB_are_addressed_correctly():
    """64 M aircraft: fields 8..12 of the state array start more than 2^31 bytes into it and fields 16..17 of the PID memory
    (the yaw angle loop) more than 2^32 bytes, so 32-bit index arithmetic anywhere in the load/store path would alias
    aircraft.  Slices at the start, in the middle and at the ragged end must equal small fleets with the same initial
    states (attitude cascade, K = 1 and 10)."""
    import acb200
    n = (1 << 26) + 130
    big = acb200.AircraftFleet(n, precision=32)

    def perturb(fleet, lo):
        ids = torch.arange(lo, lo + fleet.n, device="cuda")
        fleet.state[3] = 15.0 + (ids % 100).float() * 0.1
        fleet.state[6] = ((ids % 37).float() - 18.0) * 0.01
        fleet.state[2] = -100.0 - (ids % 50).float()

    def command(cnt, lo):
        ids = torch.arange(lo, lo + cnt, device="cuda")
        cmd = torch.zeros(4, cnt, device="cuda")
        cmd[0] = ((ids % 41).float() / 40.0 - 0.5) * 0.6
        cmd[3] = 0.6
        return cmd

    perturb(big, 0)
    cmd = command(n, 0)
    slices = [(0, 1024), (n // 2 // 128 * 128, 2048), (n - 130 - 1024, 1154)]
    small = []
    for lo, cnt in slices:
        f = acb200.AircraftFleet(cnt, precision=32)
        perturb(f, lo)
        small.append((f, command(cnt, lo)))
    for dt in (0.001, 0.01, 0.001):
        big.step_attitude(cmd, dt)
        for f, c in small:
            f.step_attitude(c, dt)
    for (lo, cnt), (f, _) in zip(slices, small):
        for name in ("state", "pid", "controls", "derived"):
            assert torch.equal(getattr(f, name), getattr(big, name)[:, lo:lo + cnt]), (name, lo)
    assert torch.isfinite(big.state).all()


@pytest.mark.parametrize("dt,steps", [(0.01, 260), (0.001, 200)])
def test_slices_of_the_full_mission_fleet_equal_small_fleets(dt, steps):
    """C5's 2^20 aircraft flying the waypoint mission through the five-level cascade (K = 10: pair kernel, K = 1: one-aircraft
    kernel): slices against small fleets with the same initial states, waypoint indices included."""
    import acb200
    wps = [[25.0, 0.0, -100.0, 15.0], [50.0, 20.0, -100.0, 15.0], [20.0, 40.0, -100.0, 15.0]]

    def build(cnt, lo):
        f = acb200.AircraftFleet(cnt, precision=32)
        ids = torch.arange(lo, lo + cnt, device="cuda")
        init = torch.zeros(13, cnt, device="cuda")
        init[0] = ((ids % 41).float() - 20.0)
        init[1] = ((ids % 37).float() - 18.0)
        init[2] = -100.0
        init[3] = 15.0
        init[8] = ((ids % 21).float() - 10.0) * 0.0175
        f.reset(init)
        f.set_mission(wps, acceptance_radius=12.0, guidance="PP", loop=True)
        return f

    big = build(N, 0)
    slices = [(0, 2048), (N // 2 - 128, 2050), (N - 1024, 1024)]
    small = [build(cnt, lo) for lo, cnt in slices]
    for _ in range(steps):
        big.step_waypoint(dt)
        for f in small:
            f.step_waypoint(dt)
    for (lo, cnt), f in zip(slices, small):
        for name in ("state", "pid", "pid_outer", "controls", "derived"):
            assert torch.equal(getattr(f, name), getattr(big, name)[:, lo:lo + cnt]), (name, lo)
        assert torch.equal(f.wp_index, big.wp_index[lo:lo + cnt]), lo
    assert torch.isfinite(big.state).all()
    if dt == 0.01:
        assert int(big.wp_index.max()) >= 1 and len(torch.unique(big.wp_index)) >= 2        # the fleet is spread over the legs
